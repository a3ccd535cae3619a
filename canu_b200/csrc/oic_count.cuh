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
 *                    on the count.  With n_pass > 1 a pass only takes the keys of its hash partition (reads larger than
 *                    the table budget are counted in several passes over the resident stream).
 *   k_count_collect  one thread per slot: keys with count >= T appended to the output (or only counted: cap == 0)
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
  unsigned long long *keys;  // EMPTY_KEY = free
  uint32_t *cnt;
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
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(cur) : "l"(a.keys + h) : "memory");
  while (true) {
    if (cur == EMPTY_KEY) cur = atomicCAS(a.keys + h, EMPTY_KEY, (unsigned long long)ck);
    if (cur == EMPTY_KEY || cur == ck) { atomicAdd(a.cnt + h, 1u); break; }
    h = (h + 1) & a.slot_mask;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(cur) : "l"(a.keys + h) : "memory");
  }
}

__global__ void __launch_bounds__(256) k_count_collect(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ cnt,
                                                       uint64_t n_slots, uint32_t min_count, unsigned long long *out_keys,
                                                       uint32_t *out_cnt, unsigned long long cap, unsigned long long *n_out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t c = i < n_slots ? cnt[i] : 0;
  const bool take = c != 0 && c >= min_count;
  const unsigned m = __ballot_sync(FULL, take);
  if (m == 0) return;
  const int l = lane_id();
  unsigned long long base = 0;
  if (l == 0) base = atomicAdd(n_out, (unsigned long long)__popc(m));
  base = __shfl_sync(FULL, base, 0);
  if (take) {
    const unsigned long long w = base + __popc(m & ((1u << l) - 1));
    if (w < cap) { out_keys[w] = keys[i]; out_cnt[w] = c; }
  }
}

}  // namespace oic
#endif
