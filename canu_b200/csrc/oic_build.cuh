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
 *   k_insert       one window per lane: one CAS on the key per probe, the occurrence ORed into the slot (k-mers seen once stay
 *                  inline in the slot); repeats go to an overflow list; a sharded build first packs the rank's windows
 *                  (k_compact_windows) so that every lane of an insert warp has work
 *   k_filter_build Bloom filter of the block's k-mers (L2-resident while filled), tested by the seed scan before the table
 *   k_multi_prep / _count / _alloc / _fill   per repeated k-mer: count its overflow entries, reserve its run of the
 *                  occurrence array (bump allocator), scatter the entries into it
 *   k_multi_sort   runs of 2..6 entries: one thread each (sorting network in registers, descending);
 *   k_sort_mid     7..32 entries: one warp per run (bitonic);  k_sort_big: longer runs, one block each (LSD radix)
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
                             uint64_t n_words_total, uint64_t *__restrict__ fwd, uint64_t *__restrict__ rc, DevCounters *ctr, uint64_t n_off) {
  // 'N' (the only non-ACGT letter a gkpStore holds: gatekeeperCreate.C:107-136 rewrites every other one) packs as code 0 plus
  // 0b11 in the N-mask stream that follows the base stream at word offset n_off (same layout, so every extraction helper
  // works on it).  n_off == 0: detection pass only (ctr->bad_base bit 0 = an N was seen, bit 1 = some other letter).
  uint64_t gw = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gw >= n_words_total) return;
  int64_t p0 = (int64_t)gw * 32;
  uint64_t wf = 0, wr = 0, mf = 0, mr = 0;
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
            case 'n': case 'N': code = 0; bad |= 1; mf |= 3ULL << (2 * k); break;
            default: code = 0; bad |= 2; break;
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
            case 'n': case 'N': rcode = 0; bad |= 1; mr |= 3ULL << (2 * k); break;   // NUL in the reference's reverse strand
            default: rcode = 0; bad |= 2; break;
          }
          wr |= (uint64_t)rcode << (2 * k);
        }
      }
    }
  }
  fwd[gw] = wf;
  rc[gw] = wr;
  if (n_off) { fwd[gw + n_off] = mf; rc[gw + n_off] = mr; }
  if (bad) atomicOr(&ctr->bad_base, bad);
}

// Same streams from the gkpStore's own 2-bit encoding (ST/gkStoreEncode.C:81-114: 4 bases per byte, FIRST base in bits 7-6,
// a=0 c=1 g=2 t=3, the last byte left-aligned; every read starts on a byte boundary): a quarter of the PCIe bytes of the
// ASCII path and no decode on the host.  One thread per 64-bit word of the padded stream.
__global__ void k_pack_2bit(const uint8_t *__restrict__ packed, const uint64_t *__restrict__ byte_off,
                            const uint32_t *__restrict__ rd_len, const int64_t *__restrict__ rd_pos, uint32_t n_reads,
                            uint64_t n_words_total, uint64_t *__restrict__ fwd, uint64_t *__restrict__ rc) {
  uint64_t gw = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gw >= n_words_total) return;
  int64_t p0 = (int64_t)gw * 32;
  uint64_t wf = 0, wr = 0;
  if (n_reads && p0 >= rd_pos[0]) {
    uint32_t lo = 0, hi = n_reads;   // largest r with rd_pos[r] <= p0
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (rd_pos[mid] <= p0) lo = mid; else hi = mid;
    }
    const uint32_t len = rd_len[lo];
    const int64_t i0 = p0 - rd_pos[lo];
    if (i0 < (int64_t)len) {
      const uint8_t *s = packed + byte_off[lo];
      const uint32_t nb = (len - (uint32_t)i0 < 32u) ? len - (uint32_t)i0 : 32u;   // bases of this word
      // forward: 8 bytes, the four 2-bit groups of every byte in reverse order (first base to the low bits)
#pragma unroll
      for (int j = 0; j < 8; j++) {
        if (4u * j < nb) {
          const uint32_t b = s[(i0 >> 2) + j];
          const uint32_t r = ((b & 3u) << 6) | ((b & 0xcu) << 2) | ((b & 0x30u) >> 2) | ((b & 0xc0u) >> 6);
          wf |= (uint64_t)r << (8 * j);
        }
      }
      if (nb < 32) wf &= (1ULL << (2 * nb)) - 1;
      // reverse complement: base i of the rc strand = 3 - base (len-1-i) of the read (AS_UTL_reverseComplement.C:114-143)
      for (uint32_t k = 0; k < nb; k++) {
        const uint32_t j = len - 1 - ((uint32_t)i0 + k);
        const uint32_t x = (s[j >> 2] >> (6 - 2 * (j & 3))) & 3u;
        wr |= (uint64_t)(x ^ 3u) << (2 * k);
      }
    }
  }
  fwd[gw] = wf;
  rc[gw] = wr;
}

// ------------------------------------------------------------------------------------------------ K1
struct BuildArgs {
  const uint64_t *fwd;
  uint64_t n_off;            // word offset of the N-mask stream behind the base stream (0: no read holds an N)
  const int64_t *rd_pos;     // per read index (0-based within the loaded set)
  const uint32_t *rd_len;
  const uint32_t *word_read; // per stream word: index of the read it belongs to (k_word_read)
  const uint8_t *str_use;    // per string slot of this block: 1 = read is loaded into the hash
  uint32_t ri_bgn;           // read index of string 0
  uint32_t n_strings;
  int64_t region_pos;        // stream position of string 0
  uint64_t region_words;     // words from region_pos to the end of the last string
  int K;
  Slot *table;
  TableGeom geom;
  uint32_t slice_lo, slice_hi;   // only k-mers whose home slice is in [slice_lo, slice_hi) are inserted (sharded build: this rank's slices)
  uint32_t *filter;          // Bloom filter words (zeroed before the insert)
  uint32_t *ovf_slot;        // overflow list: every occurrence of a k-mer that occurs more than once ...
  uint64_t *ovf_occ;         // ... {slot, occurrence word}
  uint32_t *multi_list;      // slots of those k-mers
  uint64_t *first_occ;       // per list: the occurrence that claimed the slot
  uint32_t *cnt;             // per list: occurrences on the overflow list, then the fill cursor
  uint32_t *mid_list, *big_list;
  uint64_t *occ;
  DevCounters *ctr;
};

// read that owns every 64-bit word of the stream (the last read that starts at or before it), filled once per load: the build
// kernels look the string of a word up instead of searching the read table (14 dependent loads per warp on a 10 k-read block)
__global__ void __launch_bounds__(256) k_word_read(const int64_t *__restrict__ rd_pos, uint32_t n_reads, uint64_t n_words, uint32_t *word_read) {
  const uint64_t gw = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gw >= n_words) return;
  const int64_t p0 = (int64_t)gw * 32;
  uint32_t lo = 0, hi = n_reads;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (rd_pos[mid] <= p0) lo = mid; else hi = mid;
  }
  word_read[gw] = lo;
}
__device__ __forceinline__ uint32_t find_string(const BuildArgs &a, int64_t p0) {
  const uint32_t r = a.word_read[p0 >> 5];
  return r > a.ri_bgn ? (r - a.ri_bgn < a.n_strings ? r - a.ri_bgn : a.n_strings - 1) : 0;
}

/* 16-byte compare-and-swap on a slot: returns the slot's previous content */
__device__ __forceinline__ void slot_cas(Slot *sl, unsigned long long ck, unsigned long long cp, unsigned long long nk, unsigned long long np,
                                         unsigned long long &ok, unsigned long long &op) {
  asm volatile("{\n .reg .b128 c, n, o;\n mov.b128 c, {%3, %4};\n mov.b128 n, {%5, %6};\n"
               " atom.global.relaxed.gpu.cas.b128 o, [%2], c, n;\n mov.b128 {%0, %1}, o;\n}"
               : "=l"(ok), "=l"(op) : "l"(sl), "l"(ck), "l"(cp), "l"(nk), "l"(np) : "memory");
}

/* load that bypasses L1 (the table is modified by atomics of other SMs while it is read) */
__device__ __forceinline__ unsigned long long ld_key(const Slot *sl) {
  unsigned long long v;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(&sl->key) : "memory");
  return v;
}
#ifndef OIC_CAS128
#define OIC_CAS128 1   /* claim a free slot with ONE 16-byte compare-and-swap {EMPTY, 0} -> {key, occurrence} instead of a CAS on the key + an OR into the payload */
#endif
// returns the key the slot held before (EMPTY_KEY: the claim succeeded and the payload is in place)
__device__ __forceinline__ unsigned long long slot_claim(Slot *sl, unsigned long long key, unsigned long long occ) {
#if OIC_CAS128
  unsigned long long ok, op;
  asm volatile("{\n\t.reg .b128 c, n, d;\n\tmov.b128 c, {%3, %4};\n\tmov.b128 n, {%5, %6};\n\t"
               "atom.global.relaxed.gpu.cas.b128 d, [%2], c, n;\n\tmov.b128 {%0, %1}, d;\n\t}"
               : "=l"(ok), "=l"(op) : "l"(sl), "l"(EMPTY_KEY), "l"(0ULL), "l"(key), "l"(occ) : "memory");
  return ok;
#else
  const unsigned long long cur = atomicCAS(&sl->key, EMPTY_KEY, key);
  if (cur == EMPTY_KEY && occ) atomicOr(&sl->pay, occ);
  return cur;
#endif
}
__device__ __forceinline__ bool table_find(const Slot *__restrict__ table, const TableGeom &g, uint64_t hk, unsigned long long key, uint32_t &slot,
                                           unsigned long long &pay) {
  uint64_t h = geom_home(g, hk);
  while (true) {
    uint4 v = __ldg(reinterpret_cast<const uint4 *>(table + h));
    unsigned long long cur = ((unsigned long long)v.y << 32) | v.x;
    if (cur == key) { slot = (uint32_t)h; pay = ((unsigned long long)v.w << 32) | v.z; return true; }
    if (cur == EMPTY_KEY) return false;
    h = geom_next(g, h);
  }
}

/* The k-mer window of lane l of stream word gw of the block: its key, its occurrence word (string, offset, previous base) and
 * the hash of the key.  false: no window starts there (read end, string not loaded, a window with an N: Char_Is_Bad,
 * Build_Hash_Index.C:376-393). */
__device__ __forceinline__ bool block_window(const BuildArgs &a, uint64_t gw, int l, unsigned long long &key, unsigned long long &occ, uint64_t &hk) {
  const int64_t p0 = a.region_pos + (int64_t)gw * 32;
  const uint32_t s = find_string(a, p0);
  const int64_t spos = a.rd_pos[a.ri_bgn + s];
  const uint32_t len = a.rd_len[a.ri_bgn + s];
  const int64_t o = p0 - spos + l;
  bool valid_n = true;
  const uint64_t W0 = a.fwd[p0 >> 5], W1 = a.fwd[(p0 >> 5) + 1];
  key = l ? ((W0 >> (2 * l)) | (W1 << (64 - 2 * l))) : W0;
  key &= (a.K == 32) ? ~0ULL : ((1ULL << (2 * a.K)) - 1);
  bool prev_n = false;
  if (a.n_off) {
    const uint64_t M0 = a.fwd[(p0 >> 5) + a.n_off], M1 = a.fwd[(p0 >> 5) + 1 + a.n_off];
    uint64_t mk = l ? ((M0 >> (2 * l)) | (M1 << (64 - 2 * l))) : M0;
    mk &= (a.K == 32) ? ~0ULL : ((1ULL << (2 * a.K)) - 1);
    if (mk) valid_n = false;
    prev_n = l ? ((M0 >> (2 * (l - 1))) & 3u) != 0 : (o > 0 && ((a.fwd[(p0 >> 5) - 1 + a.n_off] >> 62) & 3u) != 0);
  }
  const bool valid = valid_n && a.str_use[s] && o >= 0 && o + a.K <= (int64_t)len;
  if (!valid) return false;
  const uint32_t pv = o > 0 && !prev_n;   // the window one base earlier holds an N: it is not in the table, a run starts here
  const uint32_t pb = l ? (uint32_t)(W0 >> (2 * (l - 1))) & 3u : (pv ? (uint32_t)(a.fwd[(p0 >> 5) - 1] >> 62) & 3u : 0u);
  occ = occ_pack(s, (uint32_t)o, pv, pb);
  hk = hash_key(key);
  const uint32_t slice = geom_slice(a.geom, hk);
  return slice >= a.slice_lo && slice < a.slice_hi;   // (a sharded build only takes the k-mers of this rank's slices)
}

// block-wide exclusive scan helper + exclusive scan of a uint32 array (per-query-strand run counts), 3 phases
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

/*
 * The insert (Hash_Insert's chained bucket, Build_Hash_Index.C:296-350, appears as: slot = k-mer, list = chain).  Per probe: a
 * plain load of the key (a load that misses is much cheaper than an atomic that misses), then one 16-byte CAS;
 *   - the slot was free: the window owns the k-mer, its occurrence word went in with the same CAS (first occurrence of a k-mer:
 *     ~90 % of the windows of a 1x block never need anything else);
 *   - the key is already there: the window ORs the list mark into the payload and appends {slot, occurrence} to the overflow
 *     list; the arrival that finds the mark clear also queues the slot for the list kernels.
 * Both payload updates are ORs into disjoint bits, so they commute: no arrival waits for another, and the result does not
 * depend on the order of the windows.
 *
 * INS_R windows per thread, INS_THREADS threads per block: the block's appends to the overflow list / the list-of-lists and its
 * contributions to the two counters are aggregated in shared memory, so a block does FOUR atomics on the context's counters
 * (one atomic per warp on those few addresses -- 10 M same-line atomics for a 100 Mbp block -- serialise in one L2 slice).
 *
 * FROM_RECS = false: the windows are taken straight from the stream (one warp per 32-base word).  FROM_RECS = true: from a
 * compacted record array (k_compact_windows) -- the sharded build of a GPU group, where a rank owns 1/N of the windows and
 * would otherwise run warps with 1/N of their lanes busy: the insert is a chain of dependent memory operations per warp, so
 * its time follows the number of WARPS, not of inserts (measured: half the inserts in 81 % of the time).
 *
 * Measured and dropped in round 2: ordering the records by table region so that the lines being filled stay in L2 (256
 * partitions, contiguous or hashed by line).  DRAM traffic fell as planned, the insert time did not (7.1 ms either way): what
 * bounds it on this GPU is the ~200 M 64-bit atomics themselves, not where their lines come from.  Hence slot_claim: a free slot is
 * taken with ONE 16-byte compare-and-swap {EMPTY, 0} -> {key, occurrence} (atom.cas.b128) instead of a CAS on the key followed
 * by an OR into the payload -- 90 % of the windows of a 1x block are first occurrences: 190 M atomics become 100 M, the build
 * 10.0 -> 8.9 ms.
 */
struct PartRec { unsigned long long key, occ; };
constexpr int INS_THREADS = 256, INS_R = 4, INS_TILE = INS_THREADS * INS_R;

// the windows a rank owns, packed: block-aggregated append of {key, occurrence} records (one atomic per block)
__global__ void __launch_bounds__(INS_THREADS) k_compact_windows(BuildArgs a, PartRec *recs, uint32_t *n_recs) {
  __shared__ uint32_t s_cnt[INS_THREADS / 32], s_base;
  const int l = lane_id(), w = threadIdx.x >> 5;
  unsigned long long key[INS_R], occ[INS_R];
  uint32_t have = 0, n = 0;
#pragma unroll
  for (int r = 0; r < INS_R; r++) {
    const uint64_t gw = ((uint64_t)blockIdx.x * INS_TILE + (uint64_t)r * INS_THREADS + threadIdx.x) >> 5;
    uint64_t hk;
    if (gw < a.region_words && block_window(a, gw, l, key[r], occ[r], hk)) { have |= 1u << r; n++; }
  }
  const uint32_t incl = warp_incl_scan(n);
  if (l == 31) s_cnt[w] = incl;
  __syncthreads();
  uint32_t off = incl - n;
  for (int ww = 0; ww < w; ww++) off += s_cnt[ww];
  if (threadIdx.x == INS_THREADS - 1) s_base = (off + n) ? atomicAdd(n_recs, off + n) : 0;
  __syncthreads();
  uint32_t wpos = s_base + off;
#pragma unroll
  for (int r = 0; r < INS_R; r++)
    if ((have >> r) & 1) recs[wpos++] = PartRec{key[r], occ[r]};
}

template <bool FROM_RECS>
__global__ void __launch_bounds__(INS_THREADS) k_insert(BuildArgs a, const PartRec *__restrict__ recs, const uint32_t *__restrict__ n_recs_p) {
  __shared__ uint32_t s_later[INS_THREADS / 32], s_conv[INS_THREADS / 32], s_base[2], s_tot[4];
  const uint32_t n_recs = FROM_RECS ? *n_recs_p : 0u;
  const int l = lane_id(), w = threadIdx.x >> 5;
  unsigned long long occ_l[INS_R];
  uint32_t slot_l[INS_R];
  uint32_t n_later = 0, n_conv = 0, conv_mask = 0, later_mask = 0, n_valid = 0, n_claimed = 0;
#pragma unroll
  for (int r = 0; r < INS_R; r++) {
    const uint64_t i = (uint64_t)blockIdx.x * INS_TILE + (uint64_t)r * INS_THREADS + threadIdx.x;
    unsigned long long key, my_occ;
    uint64_t hk;
    if constexpr (FROM_RECS) {
      if (i >= n_recs) continue;
      const uint4 rv = __ldcs(reinterpret_cast<const uint4 *>(recs + i));   // read once
      key = ((unsigned long long)rv.y << 32) | rv.x;
      my_occ = ((unsigned long long)rv.w << 32) | rv.z;
      hk = hash_key(key);
    } else {
      if ((i >> 5) >= a.region_words || !block_window(a, i >> 5, l, key, my_occ, hk)) continue;
    }
    n_valid++;
    uint64_t h = geom_home(a.geom, hk);
    unsigned long long cur = ld_key(a.table + h);
    while (true) {
      if (cur == EMPTY_KEY) cur = slot_claim(a.table + h, key, my_occ);
      if (cur == EMPTY_KEY) {
        n_claimed++;
        break;
      }
      if (cur == key) {
        if (!pay_is_list(atomicOr(&a.table[h].pay, PAY_MULTI))) { conv_mask |= 1u << r; n_conv++; }
        slot_l[r] = (uint32_t)h;
        occ_l[r] = my_occ;
        later_mask |= 1u << r;
        n_later++;
        break;
      }
      h = geom_next(a.geom, h);
      cur = ld_key(a.table + h);
    }
  }
  // block-wide exclusive ranks of this thread's list entries: warp scan, then the warp totals through shared memory
  const uint32_t il = warp_incl_scan(n_later), ic = warp_incl_scan(n_conv);
  if (l == 31) { s_later[w] = il; s_conv[w] = ic; }
  const uint32_t bv = warp_sum(n_valid), bc = warp_sum(n_claimed);
  if (threadIdx.x < 4) s_tot[threadIdx.x] = 0;
  __syncthreads();
  if (l == 0) { if (bv) atomicAdd(&s_tot[0], bv); if (bc) atomicAdd(&s_tot[1], bc); }
  uint32_t off_l = il - n_later, off_c = ic - n_conv;
  for (int ww = 0; ww < w; ww++) { off_l += s_later[ww]; off_c += s_conv[ww]; }
  __syncthreads();
  if (threadIdx.x == INS_THREADS - 1) {
    const uint32_t tl = off_l + n_later, tc = off_c + n_conv;
    s_base[0] = tl ? atomicAdd(&a.ctr->n_ovf, tl) : 0;
    s_base[1] = tc ? atomicAdd(&a.ctr->n_multi, tc) : 0;
    if (s_tot[0]) atomicAdd(&a.ctr->n_inserts, (unsigned long long)s_tot[0]);
    if (s_tot[1]) atomicAdd(&a.ctr->n_distinct, (unsigned long long)s_tot[1]);
  }
  __syncthreads();
  uint32_t wl = s_base[0] + off_l, wc = s_base[1] + off_c;
#pragma unroll
  for (int r = 0; r < INS_R; r++) {
    if ((later_mask >> r) & 1) {
      a.ovf_slot[wl] = slot_l[r];
      a.ovf_occ[wl] = occ_l[r];
      wl++;
      if ((conv_mask >> r) & 1) a.multi_list[wc++] = slot_l[r];
    }
  }
}

constexpr int MULTI_GRID = 148 * 16;   // grid-stride kernels over device-side counts (no host round trip inside the build)
constexpr int SMALL_LIST = 6;           // lists up to this length are sorted by one thread

// The list kernels.  After k_insert a repeated k-mer's slot holds {first occurrence | list mark}; its other occurrences are
// on the overflow list.  i = position of the slot in multi_list indexes the per-list arrays first_occ[] / cnt[].
// (1) one thread per list: take the first occurrence out of the slot, leave the list index there
__global__ void __launch_bounds__(256) k_multi_prep(BuildArgs a) {
  const uint32_t n = a.ctr->n_multi;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t h = a.multi_list[i];
    a.first_occ[i] = a.table[h].pay & ~PAY_MULTI;
    a.table[h].pay = PAY_MULTI | i;
    a.cnt[i] = 0;
  }
}
// (2) one thread per overflow entry: count it; from here on the entry names its list, not its slot
__global__ void __launch_bounds__(256) k_multi_count(BuildArgs a) {
  const uint32_t n = a.ctr->n_ovf;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const uint32_t i = (uint32_t)a.table[a.ovf_slot[j]].pay;
    atomicAdd(&a.cnt[i], 1u);
    a.ovf_slot[j] = i;
  }
}
// (3) one thread per list: reserve its run in `occ`, place the first occurrence, publish {length, start} in the slot
__global__ void __launch_bounds__(256) k_multi_alloc(BuildArgs a) {
  const uint32_t n = a.ctr->n_multi;
  const int l = lane_id();
  for (uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; i0 < n; i0 += gridDim.x * blockDim.x) {
    const uint32_t i = i0 + l;
    const uint32_t len = (i < n) ? a.cnt[i] + 1 : 0;
    const uint32_t incl = warp_incl_scan(len);
    uint32_t base = 0;
    if (l == 31) base = atomicAdd(&a.ctr->occ_cursor, incl);
    base = __shfl_sync(FULL, base, 31);
    if (i < n) {
      const uint32_t start = base + incl - len;
      a.occ[start] = a.first_occ[i];
      a.cnt[i] = start + 1;   // fill cursor
      a.table[a.multi_list[i]].pay = PAY_MULTI | ((unsigned long long)len << 32) | start;
    }
  }
}
// (4) one thread per overflow entry
__global__ void __launch_bounds__(256) k_multi_fill(BuildArgs a) {
  const uint32_t n = a.ctr->n_ovf;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x)
    a.occ[atomicAdd(&a.cnt[a.ovf_slot[j]], 1u)] = a.ovf_occ[j];
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

// (5) one thread per list: short lists are sorted here (descending (string, offset): the order of
// the reference's head-pushed chain, Build_Hash_Index.C:317-323), longer ones are queued for a warp / a block
__global__ void __launch_bounds__(256) k_multi_sort(BuildArgs a) {
  const uint32_t n = a.ctr->n_multi;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t h = a.multi_list[i];
    const unsigned long long p = a.table[h].pay;
    const uint32_t len = (uint32_t)(p >> 32) & PAY_LEN_MASK, start = (uint32_t)p;
    if (len > 32) { a.big_list[atomicAdd(&a.ctr->n_big_lists, 1u)] = h; continue; }
    if (len > SMALL_LIST) { a.mid_list[atomicAdd(&a.ctr->n_mid_lists, 1u)] = h; continue; }
    uint64_t v[SMALL_LIST];
#pragma unroll
    for (int k = 0; k < SMALL_LIST; k++) v[k] = (k < (int)len) ? a.occ[start + k] : 0ULL;
#pragma unroll
    for (int r = 0; r < SMALL_LIST; r++) {
#pragma unroll
      for (int k = (r & 1); k + 1 < SMALL_LIST; k += 2) {
        const uint64_t x = v[k], y = v[k + 1];
        v[k] = x > y ? x : y;
        v[k + 1] = x > y ? y : x;
      }
    }
#pragma unroll
    for (int k = 0; k < SMALL_LIST; k++) if (k < (int)len) a.occ[start + k] = v[k];
  }
}

// one warp per list of SMALL_LIST+1 .. 32 entries
__global__ void __launch_bounds__(256) k_sort_mid(BuildArgs a) {
  const uint32_t n = a.ctr->n_mid_lists;
  const int l = lane_id();
  for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += (gridDim.x * blockDim.x) >> 5) {
    const unsigned long long p = a.table[a.mid_list[i]].pay;
    const uint32_t len = (uint32_t)(p >> 32) & PAY_LEN_MASK, st = (uint32_t)p;
    uint64_t v = (l < (int)len) ? a.occ[st + l] : 0;
    v = warp_sort_desc(v);
    if (l < (int)len) a.occ[st + l] = v;
  }
}

// one block per long list: LSD radix sort, 8-bit digits over bits [12, 64), descending, stable
constexpr int BIG_THREADS = 256;
__global__ void __launch_bounds__(BIG_THREADS) k_sort_big(const Slot *__restrict__ table, const uint32_t *__restrict__ big_list,
                                                          const DevCounters *ctr, uint64_t *occ, uint64_t *tmp) {
  __shared__ uint32_t bins[256];
  __shared__ uint32_t wcnt[BIG_THREADS / 32][256];
  __shared__ int skip;
  const uint32_t n_big = ctr->n_big_lists;
  for (uint32_t li = blockIdx.x; li < n_big; li += gridDim.x) {
    Slot sl = table[big_list[li]];
    const uint32_t n = pay_len(sl.pay), st0 = pay_start(sl.pay);
    uint64_t *src = occ + st0, *dst = tmp + st0;
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
    if (src != occ + st0)
      for (uint32_t i = threadIdx.x; i < n; i += BIG_THREADS) occ[st0 + i] = src[i];
    __syncthreads();
  }
}

// read in which every distinct k-mer first appears (the LAST entry of its descending list)
__global__ void k_first_occ(const Slot *__restrict__ table, uint64_t n_slots, const uint64_t *__restrict__ occ, uint32_t *hist) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_slots) return;
  Slot sl = table[i];
  if (sl.key == EMPTY_KEY) return;
  if (!pay_is_list(sl.pay)) { atomicAdd(&hist[occ_str(sl.pay)], 1u); return; }
  const uint32_t n = pay_len(sl.pay);
  if (n) atomicAdd(&hist[occ_str(occ[pay_start(sl.pay) + n - 1])], 1u);
}

// Bloom filter of the finished slices [slot_lo, slot_hi): one thread per slot, coalesced read of the table, one reduction per
// k-mer into the filter.  On its own (not from inside the insert, where the filter competes with the table lines for L2) the
// 64 MB filter stays L2-resident while it is filled.
__global__ void __launch_bounds__(256) k_filter_build(const Slot *__restrict__ table, uint64_t slot_lo, uint64_t slot_hi, TableGeom geom,
                                                      uint32_t *filter) {
  const uint64_t i = slot_lo + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= slot_hi) return;
  const unsigned long long key = __ldg(&table[i].key);
  if (key == EMPTY_KEY) return;
  uint32_t fw, fb;
  geom_filter(geom, hash_key(key), fw, fb);
  atomicOr(filter + fw, fb);
}

// Mark_Skip_Kmers: one thread per (k-mer, orientation) key.
// str_flags: bit0 = left end screened, bit1 = right end screened  (Hash_Frag_Info_t, OIC/overlapInCore.H)
__global__ void k_mark_skip(Slot *table, TableGeom geom, uint32_t *filter, const uint64_t *__restrict__ keys, uint64_t n_keys,
                            const uint64_t *__restrict__ occ, const uint32_t *__restrict__ rd_len, uint32_t ri_bgn,
                            uint32_t n_strings, uint8_t *str_flags8, int K, int add_absent, unsigned int *has_empty) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_keys) return;
  const unsigned long long key = keys[i];
  const uint64_t hk = hash_key(key);
  uint64_t h = geom_home(geom, hk);
  unsigned long long pay = 0;
  while (true) {
    Slot cur = table[h];
    if (cur.key == EMPTY_KEY) {
      if (!add_absent) return;
      // absent k-mer, hopeless check on: present as an Empty entry without occurrences (Build_Hash_Index.C:207-219)
      unsigned long long ok, op;
      slot_cas(table + h, EMPTY_KEY, 0ULL, key, PAY_MULTI | PAY_MULTI_EMPTY, ok, op);
      if (ok == EMPTY_KEY) {
        uint32_t fw, fb;   // the scan must reach this entry: it screens the query's ends (Find_Overlaps.C:323-325)
        geom_filter(geom, hk, fw, fb);
        atomicOr(filter + fw, fb);
        *has_empty = 1;
        return;
      }
      cur.key = ok; cur.pay = op;
    }
    if (cur.key == key) { pay = cur.pay; break; }
    h = geom_next(geom, h);
  }
  atomicOr(&table[h].pay, pay_is_list(pay) ? PAY_MULTI_EMPTY : PAY_SINGLE_EMPTY);
  *has_empty = 1;
  const uint32_t n = pay_len(pay), st = pay_start(pay);
  unsigned int *flags32 = reinterpret_cast<unsigned int *>(str_flags8);
  for (uint32_t j = 0; j < n; j++) {
    const uint64_t oc = pay_is_list(pay) ? occ[st + j] : pay;
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
