/*
 * oic_count.cuh -- canonical K-mer counts of the loaded reads: the step immediately upstream of overlapInCore in the
 * pipeline.  The reference gets its `-k` skip list from meryl: `meryl -B -C -m K -s gkpStore` counts the
 * CANONICAL K-mers of every read (the lexicographically smaller of a K-mer and its reverse complement; windows with a
 * non-ACGT base are skipped; src/meryl/meryl-build.C) and `meryl -Dt -n T` lists those seen >= T times in ascending K-mer
 * order (src/meryl/meryl-dump.C, Meryl.pm:634).  Here the reads are already 2-bit packed in HBM, so a count is one pass of
 * the insert machinery of the hash build:
 *
 *   k_count_kmers    one warp per 32-base word of the read stream, one window per lane; canonical key with the FIRST base in
 *                    the HIGH bits (ascending keys = meryl's order); plain load of the slot first, CAS on the key, atomicAdd
 *                    on the count in the same 16-byte slot (key and count in separate arrays cost two lines per window:
 *                    measured 53 GB of DRAM traffic for 200 M windows).  With n_pass > 1 a pass only takes the keys of its hash partition (reads larger than
 *                    the table budget are counted in several passes over the resident stream).
 *   k_count_collect  one thread per slot: keys with count >= T appended to the output (or only counted: cap == 0)
 *   k_count_hist     histogram of the counts (`meryl -Dh`)
 */
#ifndef OIC_COUNT_CUH
#define OIC_COUNT_CUH
#include "oic_build.cuh"

namespace oic {

struct CountArgs {
  const uint64_t *fwd;
  uint64_t n_off;            // word offset of the N-mask stream (0: none)
  const int64_t *rd_pos;
  const uint32_t *rd_len;
  uint32_t n_reads;
  uint64_t n_words;          // words of the read stream
  int K;
  Slot *table;               // {key (EMPTY_KEY = free), count}: one 16-byte slot, so a window touches ONE line of the table
  uint64_t slot_mask;
  uint32_t pass, n_pass;
};

/* K-mer `key` (first base in the LOW bits, as extracted from the stream) -> canonical key with the first base in the HIGH bits */
__device__ __forceinline__ uint64_t canonical_msb(uint64_t key, int K, uint64_t kmask) {
  uint64_t x = __brevll(key);                                                    // bit-reversed: groups reversed, pairs swapped
  x = ((x >> 1) & 0x5555555555555555ULL) | ((x & 0x5555555555555555ULL) << 1);   // swap the two bits of every base back
  const uint64_t f = x >> (64 - 2 * K);                                          // the K-mer itself, first base high
  const uint64_t r = ~key & kmask;                                               // its reverse complement, first base high
  return f < r ? f : r;
}

__global__ void __launch_bounds__(256) k_count_kmers(CountArgs a) {
  const uint64_t gw = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (gw >= a.n_words) return;
  const int l = lane_id();
  const int64_t p0 = (int64_t)gw * 32;
  if (a.n_reads == 0 || p0 < a.rd_pos[0]) return;
  uint32_t lo = 0, hi = a.n_reads;   // the read this word belongs to (reads start on word boundaries)
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (a.rd_pos[mid] <= p0) lo = mid; else hi = mid;
  }
  const int64_t o = p0 - a.rd_pos[lo] + l;
  bool valid = o + a.K <= (int64_t)a.rd_len[lo];
  if (!__any_sync(FULL, valid)) return;
  const uint64_t kmask = (a.K == 32) ? ~0ULL : ((1ULL << (2 * a.K)) - 1);
  const uint64_t W0 = a.fwd[gw], W1 = a.fwd[gw + 1];
  const uint64_t key = (l ? ((W0 >> (2 * l)) | (W1 << (64 - 2 * l))) : W0) & kmask;
  if (a.n_off) {
    const uint64_t M0 = a.fwd[gw + a.n_off], M1 = a.fwd[gw + 1 + a.n_off];
    if ((l ? ((M0 >> (2 * l)) | (M1 << (64 - 2 * l))) : M0) & kmask) valid = false;
  }
  if (!valid) return;
  const uint64_t ck = canonical_msb(key, a.K, kmask);
  const uint64_t hk = hash_key(ck);
  if (a.n_pass > 1 && (uint32_t)((hk >> 40) % a.n_pass) != a.pass) return;
  uint64_t h = hk & a.slot_mask;
  unsigned long long cur;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(cur) : "l"(&a.table[h].key) : "memory");
  while (true) {
    if (cur == EMPTY_KEY) cur = atomicCAS(&a.table[h].key, EMPTY_KEY, (unsigned long long)ck);
    if (cur == EMPTY_KEY || cur == ck) { atomicAdd(&a.table[h].pay, 1ull); break; }
    h = (h + 1) & a.slot_mask;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(cur) : "l"(&a.table[h].key) : "memory");
  }
}

__global__ void __launch_bounds__(256) k_count_collect(const Slot *__restrict__ table,
                                                       uint64_t n_slots, uint32_t min_count, unsigned long long *out_keys,
                                                       uint32_t *out_cnt, unsigned long long cap, unsigned long long *n_out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (i < n_slots) v = __ldg(reinterpret_cast<const uint4 *>(table + i));
  const uint32_t c = v.w ? 0xffffffffu : v.z;   // counts are far below 2^32 (at most one per window of the read set)
  const bool take = c != 0 && c >= min_count;
  const unsigned m = __ballot_sync(FULL, take);
  if (m == 0) return;
  const int l = lane_id();
  unsigned long long base = 0;
  if (l == 0) base = atomicAdd(n_out, (unsigned long long)__popc(m));
  base = __shfl_sync(FULL, base, 0);
  if (take) {
    const unsigned long long w = base + __popc(m & ((1u << l) - 1));
    if (w < cap) { out_keys[w] = ((unsigned long long)v.y << 32) | v.x; out_cnt[w] = c; }
  }
}

/* Histogram of the counts of one pass (`meryl -Dh`, src/meryl/meryl-dump.C:120-152: how many distinct K-mers were seen exactly
 * i times): counts below HIST_SMEM are binned in shared memory per block (most K-mers of a read set are seen once or
 * coverage-many times: a global atomic per slot would serialise on two or three addresses), the rest go to the global bins,
 * counts >= n_bins - 1 to the last bin's overflow list. */
constexpr int HIST_SMEM = 1024;
__global__ void __launch_bounds__(256) k_count_hist(const Slot *__restrict__ table, uint64_t n_slots, unsigned long long *hist, uint32_t n_bins,
                                                    uint32_t *big, uint32_t big_cap, unsigned int *n_big) {
  __shared__ unsigned int sh[HIST_SMEM];
  for (int i = threadIdx.x; i < HIST_SMEM; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x) {
    const unsigned long long c64 = table[i].pay;
    if (c64 == 0) continue;
    const uint32_t c = c64 > 0xffffffffull ? 0xffffffffu : (uint32_t)c64;
    if (c < HIST_SMEM) atomicAdd(&sh[c], 1u);
    else if (c < n_bins - 1) atomicAdd(&hist[c], 1ull);
    else {
      const unsigned int w = atomicAdd(n_big, 1u);
      if (w < big_cap) big[w] = c;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < HIST_SMEM; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

}  // namespace oic
#endif
