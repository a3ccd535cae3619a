/*
 * oic_store.cuh -- the step downstream of overlapInCore: overlap-store construction.
 *
 * Replaces the compute of ovStoreBuild's in-core path (ST/ovStoreBuild.C:500-645): every overlap is filtered and mirrored
 * (ovStoreFilter::filterOverlap ST/ovStoreFilter.C:118-260, ovOverlap::swapIDs ST/ovOverlap.C:114-144), the 2N records are
 * sorted by ovOverlap::operator< (ST/ovOverlap.H:327-339: a_iid, b_iid, then the two 64-bit data words) and handed to the store
 * writer.  The reference buckets the records into files and std::sorts bucket by bucket on one core; here the records of a job
 * (or of a whole sweep) fit HBM, so it is one pass:
 *
 *   k_store_mirror   one thread per overlap: decode, filter, emit the forward and the mirrored record that still carry a flag
 *   k_radix_hist / k_scan_u32_* / k_radix_scatter   LSD radix sort of (a_iid << 32 | b_iid, record index), 8-bit digits, only
 *                    over the bits the read IDs use; stable (block-local ranks by __match_any_sync)
 *   k_store_tiefix   runs of equal (a_iid, b_iid) ordered by the data words (one thread per run: they hold 1-6 records)
 *   k_store_gather   records permuted into the sorted order
 *
 * Record layout = the .ovb file words (ST/ovStoreFile.C:196-221): a_iid, b_iid, dat0.hi, dat0.lo, dat1.hi, dat1.lo with
 * dat0 = ahg5:21 ahg3:21 evalue:12 flipped forOBT forDUP forUTG extra:6, dat1 = bhg5:21 bhg3:21 span:21 extra:1 (LSB first).
 */
#ifndef OIC_STORE_CUH
#define OIC_STORE_CUH
#include "oic_build.cuh"

namespace oic {

constexpr uint64_t OV_M21 = (1ULL << 21) - 1;
constexpr int OV_FLIPPED = 54, OV_OBT = 55, OV_DUP = 56, OV_UTG = 57;
constexpr uint64_t OV_FLAGS = (1ULL << OV_OBT) | (1ULL << OV_DUP) | (1ULL << OV_UTG);
constexpr int OBT_FAR5PRIME = 29, OBT_MIN_LENGTH = 75;   // ST/ovStoreFilter.C:29-30

struct StoreRec { uint32_t w[6]; };

struct MirrorArgs {
  const StoreRec *in;
  uint64_t n;
  uint32_t max_id;              // numReads + 1: IDs must be 1 .. max_id - 1
  const uint32_t *rd_len;       // [max_id] sequence length per read ID
  const uint8_t *skip_dup;      // [max_id] library does not ask for deduplication
  const uint8_t *lib;           // [max_id] library of the read
  uint32_t max_evalue;
  StoreRec *out;                // capacity 2 n
  unsigned long long *n_out;
  unsigned long long *counts;   // [0] bad IDs, [1] skipERATE, [2] skipOBTbad, [3] skipOBTshort, [4] skipDUP, [5] skipDUPlib, [6] saveUTG, [7] saveOBT, [8] saveDUP
};

__global__ void __launch_bounds__(256) k_store_mirror(MirrorArgs a) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int l = lane_id();
  bool wf = false, wr = false;
  uint32_t fa = 0, fb = 0;
  uint64_t f0 = 0, f1 = 0, r0 = 0, r1 = 0;
  if (i < a.n) {
    const StoreRec s = a.in[i];
    fa = s.w[0]; fb = s.w[1];
    f0 = ((uint64_t)s.w[2] << 32) | s.w[3];
    f1 = ((uint64_t)s.w[4] << 32) | s.w[5];
    if (fa == 0 || fb == 0 || fa >= a.max_id || fb >= a.max_id) atomicAdd(&a.counts[0], 1ULL);   // the reference exits: the host reports it
    else {
      const uint64_t ahg5 = f0 & OV_M21, ahg3 = (f0 >> 21) & OV_M21, bhg5 = f1 & OV_M21, bhg3 = (f1 >> 21) & OV_M21;
      const bool flipped = (f0 >> OV_FLIPPED) & 1;
      // swapIDs: the hangs trade places; a flipped overlap also trades 5' and 3' so that the new A read is forward
      const uint64_t rest0 = f0 & ~((1ULL << 42) - 1), rest1 = f1 & ~((1ULL << 42) - 1);
      if (!flipped) { r0 = rest0 | bhg5 | (bhg3 << 21); r1 = rest1 | ahg5 | (ahg3 << 21); }
      else          { r0 = rest0 | bhg3 | (bhg5 << 21); r1 = rest1 | ahg3 | (ahg5 << 21); }
      if (((f0 >> 42) & 4095) > a.max_evalue) {
        f0 &= ~OV_FLAGS; r0 &= ~OV_FLAGS;
        atomicAdd(&a.counts[1], 2ULL);
      }
      // (the two "don't OBT if not requested" tests of the reference are no-ops as written, ST/ovStoreFilter.C:152-160)
      const int32_t alen = (int32_t)a.rd_len[fa], blen = (int32_t)a.rd_len[fb];
      const int32_t ab = (int32_t)ahg5, ae = alen - (int32_t)ahg3;
      const int32_t bb = flipped ? blen - (int32_t)bhg5 : (int32_t)bhg5, be = flipped ? (int32_t)bhg3 : blen - (int32_t)bhg3;
      bool is_diff = true;                                   // 5' ends far apart? (a flipped overlap always is)
      if (!flipped) {
        const uint32_t ua = (uint32_t)ab, ub = (uint32_t)bb;
        is_diff = (ua > ub ? ua - ub : ub - ua) > (uint32_t)OBT_FAR5PRIME;
      }
      int32_t blength = be - bb;
      if (be < bb) blength = bb - be;
      const bool is_long = (ae - ab > OBT_MIN_LENGTH) && (blength > OBT_MIN_LENGTH);
      unsigned long long c_bad = 0, c_short = 0, c_dup = 0, c_lib = 0;
      if (!is_diff) { if ((f0 >> OV_OBT) & 1) { f0 &= ~(1ULL << OV_OBT); c_bad++; } if ((r0 >> OV_OBT) & 1) { r0 &= ~(1ULL << OV_OBT); c_bad++; } }
      if (!is_long) { if ((f0 >> OV_OBT) & 1) { f0 &= ~(1ULL << OV_OBT); c_short++; } if ((r0 >> OV_OBT) & 1) { r0 &= ~(1ULL << OV_OBT); c_short++; } }
      // don't dedupe if the library does not ask for it (both tests look at the forward A read: roverlap.b_iid == foverlap.a_iid)
      if (((f0 >> OV_DUP) & 1) && a.skip_dup[fa]) { f0 &= ~(1ULL << OV_DUP); c_dup++; }
      if (((r0 >> OV_DUP) & 1) && a.skip_dup[fa]) { r0 &= ~(1ULL << OV_DUP); c_dup++; }
      if ((((f0 >> OV_DUP) & 1) || ((r0 >> OV_DUP) & 1)) && a.lib[fa] != a.lib[fb]) {   // no duplicates between libraries
        if ((f0 >> OV_DUP) & 1) { f0 &= ~(1ULL << OV_DUP); c_lib++; }
        if ((r0 >> OV_DUP) & 1) { r0 &= ~(1ULL << OV_DUP); c_lib++; }
      }
      if (c_bad) atomicAdd(&a.counts[2], c_bad);
      if (c_short) atomicAdd(&a.counts[3], c_short);
      if (c_dup) atomicAdd(&a.counts[4], c_dup);
      if (c_lib) atomicAdd(&a.counts[5], c_lib);
      // (the reference counts the saved flags of the forward record only, ST/ovStoreFilter.C:254-258)
      if ((f0 >> OV_UTG) & 1) atomicAdd(&a.counts[6], 1ULL);
      if ((f0 >> OV_OBT) & 1) atomicAdd(&a.counts[7], 1ULL);
      if ((f0 >> OV_DUP) & 1) atomicAdd(&a.counts[8], 1ULL);
      wf = (f0 & OV_FLAGS) != 0;
      wr = (r0 & OV_FLAGS) != 0;
    }
  }
  const unsigned mf = __ballot_sync(FULL, wf), mr = __ballot_sync(FULL, wr);
  if ((mf | mr) == 0) return;
  unsigned long long base = 0;
  if (l == 0) base = atomicAdd(a.n_out, (unsigned long long)(__popc(mf) + __popc(mr)));
  base = __shfl_sync(FULL, base, 0);
  const unsigned below = (1u << l) - 1;
  if (wf) {
    StoreRec o;
    o.w[0] = fa; o.w[1] = fb; o.w[2] = (uint32_t)(f0 >> 32); o.w[3] = (uint32_t)f0; o.w[4] = (uint32_t)(f1 >> 32); o.w[5] = (uint32_t)f1;
    a.out[base + __popc(mf & below)] = o;
  }
  if (wr) {
    StoreRec o;
    o.w[0] = fb; o.w[1] = fa; o.w[2] = (uint32_t)(r0 >> 32); o.w[3] = (uint32_t)r0; o.w[4] = (uint32_t)(r1 >> 32); o.w[5] = (uint32_t)r1;
    a.out[base + __popc(mf) + __popc(mr & below)] = o;
  }
}

__global__ void __launch_bounds__(256) k_store_keys(const StoreRec *__restrict__ rec, uint64_t n, unsigned long long *keys, uint32_t *idx) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = ((unsigned long long)rec[i].w[0] << 32) | rec[i].w[1];
  idx[i] = (uint32_t)i;
}

// ---- LSD radix sort pass over (key, idx): RADIX_TILE consecutive items per block, 8-bit digit at `shift`
constexpr int RADIX_THREADS = 256, RADIX_ITEMS = 16, RADIX_TILE = RADIX_THREADS * RADIX_ITEMS;

// hist[digit * n_blocks + block] = items of the block's tile with that digit (digit-major: its exclusive scan is the output offset)
__global__ void __launch_bounds__(RADIX_THREADS) k_radix_hist(const unsigned long long *__restrict__ keys, uint64_t n, int shift, uint32_t *hist, uint32_t n_blocks) {
  __shared__ uint32_t bins[256];
  bins[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t base = (uint64_t)blockIdx.x * RADIX_TILE;
#pragma unroll
  for (int k = 0; k < RADIX_ITEMS; k++) {
    const uint64_t i = base + (uint64_t)k * RADIX_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&bins[(keys[i] >> shift) & 255], 1u);
  }
  __syncthreads();
  hist[(uint64_t)threadIdx.x * n_blocks + blockIdx.x] = bins[threadIdx.x];
}

__global__ void __launch_bounds__(RADIX_THREADS) k_radix_scatter(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ idx, uint64_t n,
                                                                 int shift, const uint32_t *__restrict__ offs, uint32_t n_blocks,
                                                                 unsigned long long *keys_out, uint32_t *idx_out) {
  __shared__ uint32_t bins[256];                          // next output position per digit
  __shared__ uint32_t wcnt[RADIX_THREADS / 32][256];
  bins[threadIdx.x] = offs[(uint64_t)threadIdx.x * n_blocks + blockIdx.x];
  const int w = threadIdx.x >> 5, l = lane_id();
  const uint64_t base = (uint64_t)blockIdx.x * RADIX_TILE;
  for (int k = 0; k < RADIX_ITEMS; k++) {                 // chunks of 256 consecutive items, in order: the pass is stable
    for (int b = threadIdx.x; b < (RADIX_THREADS / 32) * 256; b += RADIX_THREADS) (&wcnt[0][0])[b] = 0;
    __syncthreads();
    const uint64_t i = base + (uint64_t)k * RADIX_THREADS + threadIdx.x;
    const bool act = i < n;
    const unsigned long long key = act ? keys[i] : 0;
    const int dg = act ? (int)((key >> shift) & 255) : 256 + l;   // inactive lanes never match anybody
    const unsigned m = __match_any_sync(FULL, dg);
    const int rank = __popc(m & ((1u << l) - 1));
    if (act && rank == 0) wcnt[w][dg] = __popc(m);
    __syncthreads();
    if (act) {
      uint32_t pos = bins[dg] + rank;
      for (int ww = 0; ww < w; ww++) pos += wcnt[ww][dg];
      keys_out[pos] = key;
      idx_out[pos] = idx[i];
    }
    __syncthreads();
    uint32_t add = 0;
    for (int ww = 0; ww < RADIX_THREADS / 32; ww++) add += wcnt[ww][threadIdx.x];
    bins[threadIdx.x] += add;
    __syncthreads();
  }
}

// runs of equal (a_iid, b_iid): order their records by the two data words (ovOverlap::operator<); one thread per run
__global__ void __launch_bounds__(256) k_store_tiefix(const unsigned long long *__restrict__ keys, uint32_t *idx, const StoreRec *__restrict__ rec, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || (i > 0 && keys[i - 1] == keys[i])) return;
  uint64_t j = i + 1;
  while (j < n && keys[j] == keys[i]) j++;
  if (j - i < 2) return;
  auto less = [&](uint32_t x, uint32_t y) {
    const StoreRec &p = rec[x], &q = rec[y];
    const uint64_t p0 = ((uint64_t)p.w[2] << 32) | p.w[3], q0 = ((uint64_t)q.w[2] << 32) | q.w[3];
    if (p0 != q0) return p0 < q0;
    const uint64_t p1 = ((uint64_t)p.w[4] << 32) | p.w[5], q1 = ((uint64_t)q.w[4] << 32) | q.w[5];
    return p1 < q1;
  };
  for (uint64_t x = i + 1; x < j; x++) {   // insertion sort
    const uint32_t v = idx[x];
    uint64_t y = x;
    while (y > i && less(v, idx[y - 1])) { idx[y] = idx[y - 1]; y--; }
    idx[y] = v;
  }
}

__global__ void __launch_bounds__(256) k_store_gather(const StoreRec *__restrict__ rec, const uint32_t *__restrict__ idx, uint64_t n, StoreRec *out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = rec[idx[i]];
}

}  // namespace oic
#endif
