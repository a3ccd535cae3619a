/*
 * oic_build.cuh -- K0 (2-bit pack, both strands) and K1 (k-mer index of a hash block).
 *
 * Replaces Build_Hash_Index / Put_String_In_Hash / Hash_Insert / Mark_Skip_Kmers / Hash_Mark_Empty and the
 * chain-coalescing sweep (OIC/overlapInCore-Build_Hash_Index.C:75-675).  The reference's 21-way bucket table,
 * check bytes and Hash_Check_Array are invisible to the output; what is reproduced is
 * (a) the set of k-mers, (b) each k-mer's occurrence list in the reference's chain order, (c) the Empty bit and
 * the screened-end flags, (d) the hash-load cut-off (which read is the last one loaded).
 *
 *   k_pack_reads   ASCII -> 2-bit words, forward and reverse-complement streams
 *   k_insert       one warp per 32-base word of the block: 32 windows, one per lane; claim/find the slot
 *                  (atomicCAS on the key), count the occurrence, remember the slot of every window
 *   k_scan_*       exclusive scan of the per-slot counts -> CSR starts (3-phase, in place in the table)
 *   k_fill         scatter every window's (string, offset, previous base) into its list
 *   k_sort_small   lists of 2..32 entries: one warp per list, bitonic network in registers (descending)
 *   k_sort_big     longer lists: one block per list, LSD radix sort (descending) through a ping-pong buffer
 *   k_first_occ    histogram of "read in which a k-mer first appears" -> hash-load cut-off on the host
 *   k_mark_skip    Empty bit + screened ends for the -k skip list (and absent k-mers when the hopeless check is on)
 */
#ifndef OIC_BUILD_CUH
#define OIC_BUILD_CUH
#include "oic_dev_common.cuh"

namespace oic {

// ------------------------------------------------------------------------------------------------ K0
// one thread per 64-bit word of the padded stream (padding words are written as zero), both strands.
__global__ void k_pack_reads(const char *__restrict__ ascii, const uint64_t *__restrict__ asc_off,
                             const uint32_t *__restrict__ rd_len, const int64_t *__restrict__ rd_pos, uint32_t n_reads,
                             uint64_t n_words_total, uint64_t *__restrict__ fwd, uint64_t *__restrict__ rc, DevCounters *ctr) {
  uint64_t gw = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gw >= n_words_total) return;
  int64_t p0 = (int64_t)gw * 32;
  uint64_t wf = 0, wr = 0;
  unsigned bad = 0;
  if (n_reads && p0 >= rd_pos[0]) {
    uint32_t lo = 0, hi = n_reads;   // largest r with rd_pos[r] <= p0
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (rd_pos[mid] <= p0) lo = mid; else hi = mid;
    }
    uint32_t r = lo;
    uint32_t len = rd_len[r];
    int64_t i0 = p0 - rd_pos[r];
    if (i0 < (int64_t)len) {
      const char *s = ascii + asc_off[r];
#pragma unroll 4
      for (int k = 0; k < 32; k++) {
        uint32_t i = (uint32_t)i0 + k;
        if (i < len) {
          char c = s[i];
          uint32_t code;
          switch (c) {
            case 'a': case 'A': code = 0; break;
            case 'c': case 'C': code = 1; break;
            case 'g': case 'G': code = 2; break;
            case 't': case 'T': code = 3; break;
            default: code = 0; bad = 1; break;
          }
          wf |= (uint64_t)code << (2 * k);
          // reverse-complement base i = complement of forward base len-1-i  (AS_UTL_reverseComplement.C:114-143)
          char d = s[len - 1 - i];
          uint32_t rcode;
          switch (d) {
            case 'a': case 'A': rcode = 3; break;
            case 'c': case 'C': rcode = 2; break;
            case 'g': case 'G': rcode = 1; break;
            case 't': case 'T': rcode = 0; break;
            default: rcode = 0; bad = 1; break;
          }
          wr |= (uint64_t)rcode << (2 * k);
        }
      }
    }
  }
  fwd[gw] = wf;
  rc[gw] = wr;
  if (bad) atomicOr(&ctr->bad_base, 1u);
}

// ------------------------------------------------------------------------------------------------ K1
struct BuildArgs {
  const uint64_t *fwd;
  const int64_t *rd_pos;     // per read index (0-based within the loaded set)
  const uint32_t *rd_len;
  const uint8_t *str_use;    // per string slot of this block: 1 = read is loaded into the hash
  uint32_t ri_bgn;           // read index of string 0
  uint32_t n_strings;
  int64_t region_pos;        // stream position of string 0
  uint64_t region_words;     // words from region_pos to the end of the last string
  int K;
  Slot *table;
  uint64_t slot_mask;
  uint32_t *win_slot;        // per stream position of the region: slot of the window starting there
  uint64_t *occ;
  DevCounters *ctr;
};

__device__ __forceinline__ uint32_t find_string(const BuildArgs &a, int64_t p0) {
  uint32_t lo = 0, hi = a.n_strings;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (a.rd_pos[a.ri_bgn + mid] <= p0) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ uint32_t table_insert(Slot *table, uint64_t mask, unsigned long long key, bool &claimed) {
  uint64_t h = hash_key(key) & mask;
  claimed = false;
  while (true) {
    unsigned long long cur = table[h].key;
    if (cur == key) return (uint32_t)h;
    if (cur == EMPTY_KEY) {
      unsigned long long old = atomicCAS(&table[h].key, EMPTY_KEY, key);
      if (old == EMPTY_KEY) { claimed = true; return (uint32_t)h; }
      if (old == key) return (uint32_t)h;
    }
    h = (h + 1) & mask;
  }
}

__device__ __forceinline__ bool table_find(const Slot *__restrict__ table, uint64_t mask, unsigned long long key, uint32_t &slot,
                                           uint32_t &start, uint32_t &lenflags) {
  uint64_t h = hash_key(key) & mask;
  while (true) {
    uint4 v = __ldg(reinterpret_cast<const uint4 *>(table + h));
    unsigned long long cur = ((unsigned long long)v.y << 32) | v.x;
    if (cur == key) { slot = (uint32_t)h; start = v.z; lenflags = v.w; return true; }
    if (cur == EMPTY_KEY) return false;
    h = (h + 1) & mask;
  }
}

// one warp per 32-base word of the hash block's stream region
__global__ void __launch_bounds__(256) k_insert(BuildArgs a) {
  uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= a.region_words) return;
  int l = lane_id();
  int64_t p0 = a.region_pos + (int64_t)gw * 32;
  uint32_t s = find_string(a, p0);
  int64_t spos = a.rd_pos[a.ri_bgn + s];
  uint32_t len = a.rd_len[a.ri_bgn + s];
  int64_t o = p0 - spos + l;
  bool valid = a.str_use[s] && o >= 0 && o + a.K <= (int64_t)len;
  uint64_t W0 = a.fwd[p0 >> 5], W1 = a.fwd[(p0 >> 5) + 1];
  uint64_t key = l ? ((W0 >> (2 * l)) | (W1 << (64 - 2 * l))) : W0;
  key &= (a.K == 32) ? ~0ULL : ((1ULL << (2 * a.K)) - 1);
  bool claimed = false;
  if (valid) {
    uint32_t slot = table_insert(a.table, a.slot_mask, key, claimed);
    atomicAdd(&a.table[slot].lenflags, 1u);
    a.win_slot[(p0 - a.region_pos) + l] = slot;
  }
  uint32_t nv = __popc(__ballot_sync(FULL, valid)), nc = __popc(__ballot_sync(FULL, claimed));
  if (l == 0) {
    if (nv) atomicAdd(&a.ctr->n_inserts, (unsigned long long)nv);
    if (nc) atomicAdd(&a.ctr->n_distinct, (unsigned long long)nc);
  }
}

// 3-phase exclusive scan over table[i].lenflags -> table[i].start   (SCAN_ITEMS slots per block)
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_PER_THREAD = 16;
constexpr int SCAN_ITEMS = SCAN_THREADS * SCAN_PER_THREAD;

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *total) {
  __shared__ uint32_t wsum[SCAN_THREADS / 32];
  __shared__ uint32_t tot;
  uint32_t incl = warp_incl_scan(v);
  int w = threadIdx.x >> 5, l = lane_id();
  if (l == 31) wsum[w] = incl;
  __syncthreads();
  if (w == 0) {
    uint32_t x = (l < SCAN_THREADS / 32) ? wsum[l] : 0;
    uint32_t xi = warp_incl_scan(x);
    if (l < SCAN_THREADS / 32) wsum[l] = xi - x;
    if (l == SCAN_THREADS / 32 - 1) tot = xi;
  }
  __syncthreads();
  uint32_t r = incl - v + wsum[w];
  if (total) *total = tot;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_slots_reduce(const Slot *__restrict__ table, uint64_t n, uint32_t *blk) {
  uint64_t base = (uint64_t)blockIdx.x * SCAN_ITEMS + (uint64_t)threadIdx.x * SCAN_PER_THREAD;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++)
    if (base + i < n) s += table[base + i].lenflags & LEN_MASK;
  uint32_t tot;
  block_excl_scan(s, &tot);
  if (threadIdx.x == 0) blk[blockIdx.x] = tot;
}
// single block: exclusive scan of blk[0..nb) in place; total written to blk[nb]
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_u32_single(uint32_t *blk, uint32_t nb) {
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < nb; base += SCAN_THREADS) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = (i < nb) ? blk[i] : 0;
    uint32_t tot;
    uint32_t ex = block_excl_scan(v, &tot);
    uint32_t c = carry;
    if (i < nb) blk[i] = ex + c;
    __syncthreads();
    if (threadIdx.x == 0) carry = c + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) blk[nb] = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_slots_apply(Slot *table, uint64_t n, const uint32_t *__restrict__ blk) {
  uint64_t base = (uint64_t)blockIdx.x * SCAN_ITEMS + (uint64_t)threadIdx.x * SCAN_PER_THREAD;
  uint32_t c[SCAN_PER_THREAD];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++) {
    c[i] = (base + i < n) ? (table[base + i].lenflags & LEN_MASK) : 0;
    s += c[i];
  }
  uint32_t ex = block_excl_scan(s, nullptr) + blk[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++) {
    if (base + i < n && c[i]) table[base + i].start = ex;
    ex += c[i];
  }
}

// generic exclusive scan of a uint32 array (used for per-query-strand run counts), same 3 phases
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_u32_reduce(const uint32_t *__restrict__ in, uint64_t n, uint32_t *blk) {
  uint64_t base = (uint64_t)blockIdx.x * SCAN_ITEMS + (uint64_t)threadIdx.x * SCAN_PER_THREAD;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++)
    if (base + i < n) s += in[base + i];
  uint32_t tot;
  block_excl_scan(s, &tot);
  if (threadIdx.x == 0) blk[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_u32_apply(const uint32_t *__restrict__ in, uint32_t *out, uint64_t n,
                                                                 const uint32_t *__restrict__ blk) {
  uint64_t base = (uint64_t)blockIdx.x * SCAN_ITEMS + (uint64_t)threadIdx.x * SCAN_PER_THREAD;
  uint32_t c[SCAN_PER_THREAD];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++) {
    c[i] = (base + i < n) ? in[base + i] : 0;
    s += c[i];
  }
  uint32_t ex = block_excl_scan(s, nullptr) + blk[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_PER_THREAD; i++) {
    if (base + i < n) out[base + i] = ex;
    ex += c[i];
  }
}

// same mapping as k_insert
__global__ void __launch_bounds__(256) k_fill(BuildArgs a) {
  uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= a.region_words) return;
  int l = lane_id();
  int64_t p0 = a.region_pos + (int64_t)gw * 32;
  uint32_t slot = a.win_slot[(p0 - a.region_pos) + l];
  if (slot == NO_SLOT) return;
  uint32_t s = find_string(a, p0);   // (uniform over the lanes that reach here: a word never spans two strings)
  int64_t spos = a.rd_pos[a.ri_bgn + s];
  uint32_t o = (uint32_t)(p0 - spos + l);
  uint32_t pv = o > 0, pb = 0;
  if (pv) pb = base_at(a.fwd, p0 + l - 1);
  uint32_t idx = atomicAdd(&a.table[slot].start, 1u);   // start doubles as the fill cursor; k_sort_small rewinds it
  a.occ[idx] = occ_pack(s, o, pv, pb);
}

// descending bitonic sort of up to 32 values held one per lane (0 pads sort to the end)
__device__ __forceinline__ uint64_t warp_sort_desc(uint64_t v) {
  int l = lane_id();
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      uint64_t o = __shfl_xor_sync(FULL, v, j);
      bool upper = (l & j) != 0;          // this lane holds the second element of the compared couple
      bool desc = (l & k) == 0;           // direction of this sub-sequence
      bool take_max = (desc != upper);    // first element of a descending couple keeps the max
      uint64_t mx = v > o ? v : o, mn = v > o ? o : v;
      v = take_max ? mx : mn;
    }
  }
  return v;
}

// one warp per 32 consecutive slots; lists of 2..32 are sorted here, longer ones are queued
__global__ void __launch_bounds__(256) k_sort_small(Slot *table, uint64_t n_slots, uint64_t *occ, uint32_t *big_list, DevCounters *ctr) {
  uint64_t gwarp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint64_t s0 = gwarp * 32;
  if (s0 >= n_slots) return;
  int l = lane_id();
  uint32_t len = 0, start = 0;
  if (s0 + l < n_slots) {
    Slot sl = table[s0 + l];
    if (sl.key != EMPTY_KEY) {
      len = sl.lenflags & LEN_MASK;
      start = sl.start - len;              // rewind the fill cursor
      if (len) table[s0 + l].start = start;
    }
  }
  unsigned big = __ballot_sync(FULL, len > 32);
  if (big) {
    uint32_t base = 0;
    if (l == 0) base = atomicAdd(&ctr->n_big_lists, (unsigned)__popc(big));
    base = __shfl_sync(FULL, base, 0);
    if (len > 32) big_list[base + __popc(big & ((1u << l) - 1))] = (uint32_t)(s0 + l);
  }
  unsigned todo = __ballot_sync(FULL, len >= 2 && len <= 32);
  while (todo) {
    int src = __ffs(todo) - 1;
    todo &= todo - 1;
    uint32_t n = __shfl_sync(FULL, len, src), st = __shfl_sync(FULL, start, src);
    uint64_t v = (l < (int)n) ? occ[st + l] : 0;
    v = warp_sort_desc(v);
    if (l < (int)n) occ[st + l] = v;
  }
}

// one block per long list: LSD radix sort, 8-bit digits over bits [12, 64), descending, stable
constexpr int BIG_THREADS = 256;
__global__ void __launch_bounds__(BIG_THREADS) k_sort_big(const Slot *__restrict__ table, const uint32_t *__restrict__ big_list,
                                                          uint32_t n_big, uint64_t *occ, uint64_t *tmp) {
  __shared__ uint32_t bins[256];
  __shared__ uint32_t wcnt[BIG_THREADS / 32][256];
  __shared__ int skip;
  for (uint32_t li = blockIdx.x; li < n_big; li += gridDim.x) {
    Slot sl = table[big_list[li]];
    uint32_t n = sl.lenflags & LEN_MASK;
    uint64_t *src = occ + sl.start, *dst = tmp + sl.start;
    for (int shift = 12; shift < 64; shift += 8) {
      bins[threadIdx.x] = 0;
      __syncthreads();
      for (uint32_t i = threadIdx.x; i < n; i += BIG_THREADS) atomicAdd(&bins[255 - (int)((src[i] >> shift) & 255)], 1u);
      __syncthreads();
      if (threadIdx.x == 0) {
        skip = 0;
        uint32_t run = 0;
        for (int b = 0; b < 256; b++) {
          uint32_t c = bins[b];
          if (c == n) skip = 1;
          bins[b] = run;
          run += c;
        }
      }
      __syncthreads();
      if (skip) { __syncthreads(); continue; }
      int w = threadIdx.x >> 5, l = lane_id();
      for (uint32_t base = 0; base < n; base += BIG_THREADS) {
        for (int b = threadIdx.x; b < (BIG_THREADS / 32) * 256; b += BIG_THREADS) (&wcnt[0][0])[b] = 0;
        __syncthreads();
        uint32_t i = base + threadIdx.x;
        bool act = i < n;
        uint64_t v = act ? src[i] : 0;
        int dg = act ? 255 - (int)((v >> shift) & 255) : 256 + l;   // inactive lanes never match anybody
        unsigned m = __match_any_sync(FULL, dg);
        int rank = __popc(m & ((1u << l) - 1));
        if (act && rank == 0) wcnt[w][dg] = __popc(m);
        __syncthreads();
        if (act) {
          uint32_t pos = bins[dg] + rank;
          for (int ww = 0; ww < w; ww++) pos += wcnt[ww][dg];
          dst[pos] = v;
        }
        __syncthreads();
        {
          uint32_t add = 0;
          for (int ww = 0; ww < BIG_THREADS / 32; ww++) add += wcnt[ww][threadIdx.x];
          bins[threadIdx.x] += add;
        }
        __syncthreads();
      }
      uint64_t *t = src; src = dst; dst = t;
    }
    // result must end up in occ
    if (src != occ + sl.start)
      for (uint32_t i = threadIdx.x; i < n; i += BIG_THREADS) occ[sl.start + i] = src[i];
    __syncthreads();
  }
}

// read in which every distinct k-mer first appears (the LAST entry of its descending list)
__global__ void k_first_occ(const Slot *__restrict__ table, uint64_t n_slots, const uint64_t *__restrict__ occ, uint32_t *hist) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_slots) return;
  Slot sl = table[i];
  uint32_t n = sl.lenflags & LEN_MASK;
  if (sl.key == EMPTY_KEY || n == 0) return;
  atomicAdd(&hist[occ_str(occ[sl.start + n - 1])], 1u);
}

// Mark_Skip_Kmers: one thread per (k-mer, orientation) key.
// str_flags: bit0 = left end screened, bit1 = right end screened  (Hash_Frag_Info_t, OIC/overlapInCore.H)
__global__ void k_mark_skip(Slot *table, uint64_t mask, const uint64_t *__restrict__ keys, uint64_t n_keys,
                            const uint64_t *__restrict__ occ, const uint32_t *__restrict__ rd_len, uint32_t ri_bgn,
                            uint32_t n_strings, uint8_t *str_flags8, int K, int add_absent, unsigned int *has_empty) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_keys) return;
  unsigned long long key = keys[i];
  uint64_t h = hash_key(key) & mask;
  while (true) {
    unsigned long long cur = table[h].key;
    if (cur == key) break;
    if (cur == EMPTY_KEY) {
      if (!add_absent) return;
      unsigned long long old = atomicCAS(&table[h].key, EMPTY_KEY, key);
      if (old == EMPTY_KEY || old == key) break;
    }
    h = (h + 1) & mask;
  }
  uint32_t old = atomicOr(&table[h].lenflags, EMPTY_FLAG);
  *has_empty = 1;
  uint32_t n = old & LEN_MASK;
  uint32_t st = table[h].start;
  unsigned int *flags32 = reinterpret_cast<unsigned int *>(str_flags8);
  for (uint32_t j = 0; j < n; j++) {
    uint64_t oc = occ[st + j];
    uint32_t s = occ_str(oc), off = occ_off(oc);
    if (s >= n_strings) continue;
    uint32_t f = 0;
    if ((int)off < HOPELESS_MATCH) f |= 1;
    if ((int64_t)rd_len[ri_bgn + s] - (int64_t)off - K + 1 < HOPELESS_MATCH) f |= 2;
    if (f) atomicOr(&flags32[s >> 2], f << ((s & 3) * 8));
  }
}

}  // namespace oic
#endif
