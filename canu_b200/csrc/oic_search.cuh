/*
 * oic_search.cuh -- K2: seed scan of a batch of query reads against the current hash block.
 *
 * Replaces Find_Overlaps / Hash_Find / Add_Ref / Add_Match (OIC/overlapInCore-Find_Overlaps.C:77-385) and the
 * String_Olap accumulator (OIC/overlapInCore.H:245-255).  The reference pushes every k-mer hit through a
 * per-pair linked list; here only the FIRST hit of every maximal exact-match run leaves the scan kernel
 * (a hit continues a run iff the previous query base equals the previous hash base, which is stored in the
 * occurrence entry), so the traffic after the probe is one 16-byte record per run instead of one per hit:
 *
 *   k_seed_scan    one warp per (query, strand); 32 windows per step from two 64-bit packed words; one 16-byte
 *                  probe per window; occurrence lists walked warp-cooperatively (balanced over the lanes);
 *                  run starts appended, in arrival order, to the strand's chunk chain
 *   k_run_length   one thread per run: length by 32-base word compares (plus the Empty-window clip)
 *   k_group_runs   one warp per (query, strand): gather the chain, stable LSD radix sort by hash string
 *                  (shared-memory digit bins), emit one pair descriptor per hash read
 *   k_replay       one thread per pair: Add_Match replayed on runs (oic_core.cuh) -> seed list in list order
 */
#ifndef OIC_SEARCH_CUH
#define OIC_SEARCH_CUH
#include "oic_build.cuh"

namespace oic {

constexpr int CHUNK = 128;             // run records per chunk
constexpr uint32_t NO_CHUNK = 0xffffffffu;

struct Pair {
  uint32_t qs;      // (query index in batch) * 2 + strand
  uint32_t str;     // hash-block string index
  uint32_t first;   // first seed in the seed arrays
  uint32_t n;       // seeds
};
struct PairInfo {
  int32_t consistent, diag_ct, diag_bgn, diag_end;
};

struct ScanArgs {
  const uint64_t *fwd, *rc;
  uint64_t n_off;                 // word offset of the N-mask streams (0: none)
  const int64_t *rd_pos;
  const uint32_t *rd_len;
  const uint32_t *qs_read;        // per query in batch: read index, or 0xffffffff if the read is skipped
  const uint32_t *q_order;        // queries of the batch, longest first: the order the warps take them in (the scan of a strand is one
                                  // serial chain of dependent loads, so a long strand started last would be the tail of the kernel)
  uint32_t n_queries;             // reads in the batch (strands = 2x)
  uint32_t first_id;              // read ID of read index 0
  uint32_t hash_bgn_id;           // read ID of string 0
  int K;
  const Slot *table;
  TableGeom geom;
  const uint64_t *occ;
  const uint32_t *filter;         // Bloom filter of the block's k-mers (nullptr = none)
  int has_empty;
  // outputs
  Seed *recs;                     // [max_chunks * CHUNK]
  uint32_t max_chunks;
  uint32_t *chunk_next, *chunk_cnt, *chunk_qs;
  uint32_t *qs_first_chunk, *qs_nrec;   // [2 * n_queries]
  uint8_t *qs_screen;             // bit0 left end screened, bit1 right end screened
  const uint32_t *qs_tile_base;   // [2 * n_queries] index into tile_empty (only if has_empty)
  uint32_t *tile_empty;           // ballot of Empty windows per 32-window tile
  DevCounters *ctr;
};

constexpr int SCAN_WARPS = 4;

// resident blocks per SM the register allocation aims at.  The scan is a chain of dependent loads per warp (filter word -> table
// line -> occurrence list), so throughput comes from resident warps: measured on B200 (c3 job) 1 (69 registers, 28 warps per SM)
// 5.47 ms, 8 (56 registers) 4.52 ms, 10 (48) 4.28 ms, 12 (40 registers, 16 bytes of spills, 48 warps) 4.02 ms.
#ifndef OIC_SCAN_BLOCKS
#define OIC_SCAN_BLOCKS 12
#endif
__global__ void __launch_bounds__(SCAN_WARPS * 32, OIC_SCAN_BLOCKS) k_seed_scan(ScanArgs a) {
  __shared__ uint32_t s_pref[SCAN_WARPS][33];
  __shared__ unsigned long long s_pay[SCAN_WARPS][32];
  const int wib = threadIdx.x >> 5, l = lane_id();
  const uint32_t slot = blockIdx.x * SCAN_WARPS + wib;
  if (slot >= 2 * a.n_queries) return;
  const uint32_t q = a.q_order[slot >> 1], strand = slot & 1;
  const uint32_t qs = 2 * q + strand;
  const uint32_t ri = a.qs_read[q];
  if (l == 0) { a.qs_first_chunk[qs] = NO_CHUNK; a.qs_nrec[qs] = 0; a.qs_screen[qs] = 0; }
  if (ri == 0xffffffffu) return;
  const uint32_t len = a.rd_len[ri];
  const int K = a.K;
  if ((int)len < K) return;
  const uint64_t *W = strand ? a.rc : a.fwd;
  const int64_t wbase = a.rd_pos[ri] >> 5;
  const uint32_t q_id = a.first_id + ri;
  const uint64_t kmask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
  uint32_t n_win = len - K + 1, eff_len = len;
  if (a.n_off && strand) {
    // The reference's reverse strand holds NUL where the read has N (AS_UTL_reverseComplement.C:39-72) and its scan is
    // NUL-driven: the first K characters are taken as they are, then it stops at the first NUL (Find_Overlaps.C:309-341).
    for (uint32_t b0 = 0; b0 < len && eff_len == len; b0 += 1024) {
      const uint32_t wi = (b0 >> 5) + l;
      uint64_t mw = (wi * 32u < len) ? W[wbase + wi + a.n_off] : 0;
      if (wi * 32u < (uint32_t)K) mw &= ((uint32_t)K - wi * 32u >= 32u) ? 0ULL : (~0ULL << (2 * ((uint32_t)K - wi * 32u)));
      const unsigned hit = __ballot_sync(FULL, mw != 0);
      if (hit) {
        const int fl = __ffs(hit) - 1;
        const uint64_t fw = __shfl_sync(FULL, mw, fl);
        const uint32_t pos = b0 + 32u * fl + ((__ffsll((long long)fw) - 1) >> 1);
        if (pos < len) eff_len = pos;
      }
    }
    n_win = eff_len - K + 1;   // eff_len >= K by construction
  }

  // per-warp output cursor
  uint32_t cur_chunk = NO_CHUNK, cur_fill = 0, n_rec = 0;
  unsigned screen = 0;
  unsigned long long c_found = 0, c_refs = 0, c_add = 0;
  uint32_t prev_tile_last_empty = 0;   // Empty flag of the window just before this tile
  uint32_t prev_tile_last_base = 0;

  for (uint32_t t0 = 0; t0 < n_win; t0 += 32) {
    uint64_t W0 = W[wbase + (t0 >> 5)], W1 = W[wbase + (t0 >> 5) + 1];
    uint32_t off = t0 + l;
    bool valid = off < n_win;
    uint64_t key = (l ? ((W0 >> (2 * l)) | (W1 << (64 - 2 * l))) : W0) & kmask;
    uint64_t M0 = 0;
    if (a.n_off) {   // a window with an N (or NUL) never verifies against the all-acgt table strings (Hash_Find's strncmp)
      M0 = W[wbase + (t0 >> 5) + a.n_off];
      const uint64_t M1 = W[wbase + (t0 >> 5) + 1 + a.n_off];
      if ((l ? ((M0 >> (2 * l)) | (M1 << (64 - 2 * l))) : M0) & kmask) valid = false;
    }
    // Bloom filter first (an L2 hit): most windows of a query are absent from the block and never touch the table
    const uint64_t hk = hash_key(key);
    bool maybe = valid;
    if (a.filter) {
      uint32_t fw, fb;
      geom_filter(a.geom, hk, fw, fb);
      maybe = valid && ((__ldg(a.filter + fw) & fb) == fb);
    }
    uint32_t slot;
    unsigned long long pay = 0;
    bool found = maybe && table_find(a.table, a.geom, hk, key, slot, pay);
    bool empty = found && pay_empty(pay);
    if (found) c_found++;
    if (empty) {   // hi_hits: OIC/overlapInCore-Find_Overlaps.C:323-325, 359-366
      if (off == 0) screen |= 1;
      else {
        if ((int)off < HOPELESS_MATCH) screen |= 1;
        if ((int)len - (int)off - K + 1 < HOPELESS_MATCH) screen |= 2;
      }
    }
    unsigned empty_mask = __ballot_sync(FULL, empty);
    if (a.has_empty && l == 0) a.tile_empty[a.qs_tile_base[qs] + (t0 >> 5)] = empty_mask;
    uint32_t cnt = (found && !empty) ? pay_len(pay) : 0;
    uint32_t incl = warp_incl_scan(cnt);
    uint32_t total = __shfl_sync(FULL, incl, 31);
    // previous query base / previous window Empty for every lane's window
    uint32_t my_prev_base = l ? (uint32_t)(W0 >> (2 * (l - 1))) & 3u : prev_tile_last_base;
    if (a.n_off && l && ((M0 >> (2 * (l - 1))) & 3u)) my_prev_base = 4;   // previous query base is N: that window was no hit
    uint32_t my_prev_empty = l ? ((empty_mask >> (l - 1)) & 1u) : prev_tile_last_empty;
    prev_tile_last_base = ((M0 >> 62) & 3u) ? 4u : (uint32_t)(W0 >> 62) & 3u;
    prev_tile_last_empty = (empty_mask >> 31) & 1u;
    if (total == 0) continue;
    __syncwarp();
    s_pref[wib][l] = incl - cnt;
    s_pay[wib][l] = pay;
    if (l == 31) s_pref[wib][32] = total;
    __syncwarp();

    for (uint32_t r0 = 0; r0 < total; r0 += 32) {
      uint32_t r = r0 + l;
      bool act = r < total;
      int owner = 0;
      uint64_t oc = 0;
      if (act) {
        // owner = last lane whose exclusive prefix is <= r
        int lo = 0, hi = 32;
        while (hi - lo > 1) {
          int mid = (lo + hi) >> 1;
          if (s_pref[wib][mid] <= r) lo = mid; else hi = mid;
        }
        owner = lo;
        const unsigned long long op = s_pay[wib][owner];   // a k-mer seen once carries its occurrence in the slot
        oc = pay_is_list(op) ? __ldg(a.occ + pay_start(op) + (r - s_pref[wib][owner])) : op;
      }
      uint32_t pbase = __shfl_sync(FULL, my_prev_base, owner);
      uint32_t pempty = __shfl_sync(FULL, my_prev_empty, owner);
      bool is_start = false;
      Seed rec;
      if (act) {
        uint32_t str = occ_str(oc), hoff = occ_off(oc);
        c_refs++;
        if (q_id < a.hash_bgn_id + str) {   // only pairs with queryID < hashID (Find_Overlaps.C:328,369)
          c_add++;
          uint32_t qoff = t0 + owner;
          // the hit continues a run iff the window one base earlier was an Add_Ref'd hit to (string, offset-1):
          // previous bases equal (the hash side's is stored in the entry) and that window was not screened
          bool cont = qoff > 0 && ((oc >> 2) & 1) && ((uint32_t)(oc & 3) == pbase) && !pempty;
          is_start = !cont;
          rec.start = (int32_t)qoff; rec.offset = (int32_t)hoff; rec.len = 0; rec.str = (int32_t)str;
        }
      }
      unsigned smask = __ballot_sync(FULL, is_start);
      if (smask) {
        uint32_t ns = __popc(smask);
        if (cur_chunk == NO_CHUNK || cur_fill + ns > CHUNK) {
          uint32_t nc = NO_CHUNK;
          if (l == 0) {
            nc = atomicAdd(&a.ctr->rec_chunks, 1u);
            if (nc >= a.max_chunks) { atomicOr(&a.ctr->rec_overflow, 1u); nc = NO_CHUNK; }
            else {
              a.chunk_next[nc] = NO_CHUNK;
              a.chunk_qs[nc] = qs;
              if (cur_chunk == NO_CHUNK) a.qs_first_chunk[qs] = nc;
              else { a.chunk_next[cur_chunk] = nc; a.chunk_cnt[cur_chunk] = cur_fill; }
            }
          }
          nc = __shfl_sync(FULL, nc, 0);
          if (nc == NO_CHUNK) {   // out of run space: the host retries the batch with fewer queries
            if (l == 0 && cur_chunk != NO_CHUNK) a.chunk_cnt[cur_chunk] = cur_fill;
            return;
          }
          cur_chunk = nc;
          cur_fill = 0;
        }
        if (is_start) a.recs[(uint64_t)cur_chunk * CHUNK + cur_fill + __popc(smask & ((1u << l) - 1))] = rec;
        cur_fill += ns;
        n_rec += ns;
      }
    }
    __syncwarp();
  }
  if (l == 0) {
    if (cur_chunk != NO_CHUNK) a.chunk_cnt[cur_chunk] = cur_fill;
    a.qs_nrec[qs] = n_rec;
  }
  // reductions
  screen = __reduce_or_sync(FULL, screen);
  c_found = warp_sum64(c_found); c_refs = warp_sum64(c_refs); c_add = warp_sum64(c_add);
  if (l == 0) {
    a.qs_screen[qs] = (uint8_t)screen;
    atomicAdd(&a.ctr->n_query_bases, (unsigned long long)eff_len);
    atomicAdd(&a.ctr->n_windows, (unsigned long long)n_win);
    if (c_found) atomicAdd(&a.ctr->n_found, c_found);
    if (c_refs) atomicAdd(&a.ctr->n_refs, c_refs);
    if (c_add) atomicAdd(&a.ctr->n_add_ref, c_add);
    if (n_rec) atomicAdd(&a.ctr->n_runs, (unsigned long long)n_rec);
  }
}

// ------------------------------------------------------------------------------------------------
struct RunLenArgs {
  const uint64_t *fwd, *rc;
  uint64_t n_off;                // word offset of the N-mask streams (0: none)
  const int64_t *rd_pos;
  const uint32_t *rd_len;
  const uint32_t *qs_read;
  uint32_t hash_ri_bgn;          // read index of string 0
  int K;
  Seed *recs;
  uint32_t n_chunks;
  const uint32_t *chunk_cnt, *chunk_qs;
  int has_empty;
  const uint32_t *qs_tile_base, *tile_empty;
};

__global__ void __launch_bounds__(256) k_run_length(RunLenArgs a) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t chunk = (uint32_t)(i / CHUNK), j = (uint32_t)(i % CHUNK);
  if (chunk >= a.n_chunks || j >= a.chunk_cnt[chunk]) return;
  uint32_t qs = a.chunk_qs[chunk];
  uint32_t ri = a.qs_read[qs >> 1];
  Seed rec = a.recs[i];
  const uint64_t *S = (qs & 1) ? a.rc : a.fwd;
  int64_t sp = a.rd_pos[ri];
  int s_len = (int)a.rd_len[ri];
  uint32_t hri = a.hash_ri_bgn + (uint32_t)rec.str;
  int64_t tp = a.rd_pos[hri];
  int t_len = (int)a.rd_len[hri];
  int K = a.K;
  int limit = imin(s_len - (rec.start + K), t_len - (rec.offset + K));
  int m = 0;
  if (limit > 0) {
    if (a.n_off == 0) m = match_fwd(S, sp + rec.start + K, a.fwd, tp + rec.offset + K, limit);
    else {   // a window that holds an N on either side is no hit: the run of hits ends in front of it
      const int64_t pa = sp + rec.start + K, pt = tp + rec.offset + K;
      while (m < limit) {
        const uint64_t x = (ext32(S, pa + m) ^ ext32(a.fwd, pt + m)) | ext32(S + a.n_off, pa + m) | ext32(a.fwd + a.n_off, pt + m);
        if (x) { m += (__ffsll((long long)x) - 1) >> 1; break; }
        m += 32;
      }
      if (m > limit) m = limit;
    }
  }
  int hits = 1 + m;
  if (a.has_empty && hits > 1) {
    // the run stops in front of the first screened (Empty) window: that window produces no Add_Ref
    const uint32_t *te = a.tile_empty + a.qs_tile_base[qs];
    int w = rec.start + 1, wend = rec.start + hits;   // windows (w .. wend-1) must be clear
    while (w < wend) {
      uint32_t bits = te[w >> 5] >> (w & 31);
      if (bits) {
        int f = w + __ffs(bits) - 1;
        if (f < wend) hits = f - rec.start;
        break;
      }
      w = (w | 31) + 1;
    }
  }
  a.recs[i].len = K + hits - 1;
}

// ------------------------------------------------------------------------------------------------
struct GroupArgs {
  const Seed *recs;
  const uint32_t *chunk_next, *chunk_cnt;
  const uint32_t *qs_first_chunk, *qs_nrec, *qs_base;   // qs_base = exclusive scan of qs_nrec
  uint32_t n_qs;
  Seed *bufA, *bufB;             // dense seed arrays (ping-pong)
  int passes;                    // 8-bit digits needed to cover the string index
  Pair *pairs;
  DevCounters *ctr;
};

constexpr int GROUP_WARPS = 4;

// result of the sort lives in bufA if passes is even, else in bufB
__global__ void __launch_bounds__(GROUP_WARPS * 32) k_group_runs(GroupArgs a) {
  __shared__ uint32_t s_bins[GROUP_WARPS][256];
  const int wib = threadIdx.x >> 5, l = lane_id();
  const uint32_t qs = blockIdx.x * GROUP_WARPS + wib;
  if (qs >= a.n_qs) return;
  const uint32_t n = a.qs_nrec[qs];
  if (n == 0) return;
  const uint32_t base = a.qs_base[qs];
  // gather the chunk chain, in order
  {
    uint32_t c = a.qs_first_chunk[qs], w = 0;
    while (c != NO_CHUNK) {
      uint32_t cnt = a.chunk_cnt[c];
      for (uint32_t j = l; j < cnt; j += 32) a.bufA[base + w + j] = a.recs[(uint64_t)c * CHUNK + j];
      w += cnt;
      c = a.chunk_next[c];
    }
  }
  __syncwarp();
  Seed *src = a.bufA + base, *dst = a.bufB + base;
  uint32_t *bins = s_bins[wib];
  for (int p = 0; p < a.passes; p++) {
    int shift = 8 * p;
    for (int b = l; b < 256; b += 32) bins[b] = 0;
    __syncwarp();
    for (uint32_t i = l; i < n; i += 32) atomicAdd(&bins[((uint32_t)src[i].str >> shift) & 255], 1u);
    __syncwarp();
    // exclusive scan of 256 bins: 8 per lane
    {
      uint32_t v[8], s = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) { v[k] = bins[l * 8 + k]; s += v[k]; }
      uint32_t ex = warp_incl_scan(s) - s;
#pragma unroll
      for (int k = 0; k < 8; k++) { bins[l * 8 + k] = ex; ex += v[k]; }
    }
    __syncwarp();
    for (uint32_t i0 = 0; i0 < n; i0 += 32) {
      uint32_t i = i0 + l;
      bool act = i < n;
      Seed rec;
      int dg = 256 + l;
      if (act) { rec = src[i]; dg = ((uint32_t)rec.str >> shift) & 255; }
      unsigned m = __match_any_sync(FULL, dg);
      int rank = __popc(m & ((1u << l) - 1));
      uint32_t pos = 0;
      if (act) pos = bins[dg] + rank;
      __syncwarp();
      if (act && rank == 0) bins[dg] += __popc(m);
      if (act) dst[pos] = rec;
      __syncwarp();
    }
    Seed *t = src; src = dst; dst = t;
  }
  // pair descriptors: one per distinct string of the (now grouped) segment
  uint32_t n_pairs = 0;
  for (uint32_t i0 = 0; i0 < n; i0 += 32) {
    uint32_t i = i0 + l;
    bool head = i < n && (i == 0 || src[i].str != src[i - 1].str);
    n_pairs += __popc(__ballot_sync(FULL, head));
  }
  uint32_t pbase = 0;
  if (l == 0) pbase = (uint32_t)atomicAdd(&a.ctr->n_pairs, (unsigned long long)n_pairs);
  pbase = __shfl_sync(FULL, pbase, 0);
  uint32_t seen = 0;
  for (uint32_t i0 = 0; i0 < n; i0 += 32) {
    uint32_t i = i0 + l;
    bool head = i < n && (i == 0 || src[i].str != src[i - 1].str);
    unsigned hm = __ballot_sync(FULL, head);
    if (head) {
      uint32_t pi = pbase + seen + __popc(hm & ((1u << l) - 1));
      Pair pr;
      pr.qs = qs; pr.str = (uint32_t)src[i].str; pr.first = base + i; pr.n = 0;
      a.pairs[pi] = pr;
    }
    seen += __popc(hm);
  }
}

// pair sizes: n = next pair's first - first (pairs of one strand are contiguous and ordered) -- computed by the
// replay kernel from the seed array itself: walk until the string changes or the strand's segment ends.
struct ReplayArgs {
  const Seed *src;      // grouped runs
  Seed *dst;            // seeds in list order
  int32_t *scratch;     // 4 ints per seed
  Pair *pairs;
  PairInfo *info;
  uint32_t n_pairs;
  const uint32_t *qs_base, *qs_nrec;
  int K;
};

__global__ void __launch_bounds__(128) k_replay(ReplayArgs a) {
  uint32_t pi = blockIdx.x * blockDim.x + threadIdx.x;
  if (pi >= a.n_pairs) return;
  Pair pr = a.pairs[pi];
  uint32_t seg_end = a.qs_base[pr.qs] + a.qs_nrec[pr.qs];
  uint32_t n = 1;
  while (pr.first + n < seg_end && (uint32_t)a.src[pr.first + n].str == pr.str) n++;
  a.pairs[pi].n = n;
  PairInfo inf;
  replay_add_match(a.src + pr.first, (int)n, a.K, a.dst + pr.first, a.scratch + 4ull * pr.first, inf.consistent, inf.diag_ct,
                   inf.diag_bgn, inf.diag_end);
  a.info[pi] = inf;
}

// ---- longest-processing-time-first order of the pairs: pairs with many seeds need many sequential extensions,
// so they enter the work queue first (counting sort by seed count, descending)
constexpr int ORDER_BINS = 1024;
// cost proxy of a pair: many seeds = many sequential extensions; one or two seeds may be a spurious k-mer hit whose
// band is never trimmed (up to error_limit^2 cells), so those go first as well
__device__ __forceinline__ uint32_t pair_bin(uint32_t n_seeds) {
  uint32_t k = n_seeds <= 2 ? ORDER_BINS - 1 : (n_seeds < ORDER_BINS - 2 ? n_seeds : ORDER_BINS - 2);
  return ORDER_BINS - 1 - k;
}
__global__ void k_pair_hist(const Pair *__restrict__ pairs, uint32_t n, uint32_t *hist) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  atomicAdd(&hist[pair_bin(pairs[i].n)], 1u);
}
__global__ void __launch_bounds__(ORDER_BINS) k_pair_bins_scan(uint32_t *hist) {
  __shared__ uint32_t s[ORDER_BINS];
  s[threadIdx.x] = hist[threadIdx.x];
  __syncthreads();
  for (int o = 1; o < ORDER_BINS; o <<= 1) {
    uint32_t v = threadIdx.x >= (unsigned)o ? s[threadIdx.x - o] : 0;
    __syncthreads();
    s[threadIdx.x] += v;
    __syncthreads();
  }
  hist[threadIdx.x] = s[threadIdx.x] - hist[threadIdx.x];   // exclusive
}
__global__ void k_pair_scatter(const Pair *__restrict__ pairs, uint32_t n, uint32_t *cursor, uint32_t *order) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  order[atomicAdd(&cursor[pair_bin(pairs[i].n)], 1u)] = i;
}

}  // namespace oic
#endif
