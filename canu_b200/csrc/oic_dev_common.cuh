/*
 * oic_dev_common.cuh -- device-side data layout shared by the build / scan / extend kernels.
 *
 * HBM layout:
 *   reads     2-bit packed, 32 bases per uint64 word, base i of the stream at bits [2i, 2i+2) of word i/32
 *             (first base in the LOW bits, the same convention as the reference's rolling k-mer key,
 *             OIC/overlapInCore-Build_Hash_Index.C:374-379).  Every read starts on a word boundary and is
 *             followed by >= 64 zero bases, so unaligned 32-base extractions never leave the buffer.
 *             Two streams with identical read positions: forward strand and reverse complement.
 *   table     open-addressing k-mer table, 16-byte slots {key, start, len|Empty}; one 16-byte load answers a probe.
 *   occ       occurrence array (CSR payload), 8 bytes per k-mer occurrence, DESCENDING (string, offset) per key:
 *             the order the reference's chain has after pushing at the head (Build_Hash_Index.C:317-323).
 */
#ifndef OIC_DEV_COMMON_CUH
#define OIC_DEV_COMMON_CUH
#include <cuda_runtime.h>
#include <stdint.h>

#include "oic_core.cuh"

namespace oic {

struct __align__(16) Slot {
  unsigned long long key;
  unsigned long long pay;   // one occurrence inline, or {Empty, length, start} of a list in `occ` (see below)
};
constexpr unsigned long long EMPTY_KEY = ~0ULL;
constexpr uint32_t NO_SLOT = 0xffffffffu;

/* Slot payload.  A k-mer that occurs once in the block (most of them) keeps its occurrence word in the slot, so a probe
 * that hits needs no second memory access:
 *   single : bit 63 = 0, the occurrence word itself (oic_core.cuh occ_pack: [62:33] string, [32:12] offset, [2:0] previous
 *            base), bit 4 = Empty (screened k-mer)
 *   list   : bit 63 = 1, bit 62 = Empty, [61:32] length, [31:0] first entry in `occ` (length 0: a screened k-mer that is
 *            absent from the block, Build_Hash_Index.C:207-219)
 * While the block is being built the low word of a list payload is first the arrival count, then the fill cursor. */
constexpr unsigned long long PAY_MULTI = 1ULL << 63;
constexpr unsigned long long PAY_MULTI_EMPTY = 1ULL << 62;
constexpr unsigned long long PAY_SINGLE_EMPTY = 1ULL << 4;
constexpr uint32_t PAY_LEN_MASK = 0x3fffffffu;
OIC_HD bool pay_is_list(unsigned long long p) { return (p >> 63) != 0; }
OIC_HD uint32_t pay_len(unsigned long long p) { return pay_is_list(p) ? (uint32_t)(p >> 32) & PAY_LEN_MASK : 1u; }
OIC_HD uint32_t pay_start(unsigned long long p) { return (uint32_t)p; }
OIC_HD bool pay_empty(unsigned long long p) { return pay_is_list(p) ? ((p >> 62) & 1) != 0 : ((p >> 4) & 1) != 0; }

/* Geometry of the k-mer table.  The table is cut into n_slices slices of 2^slice_bits slots; a k-mer's home slot is
 * (slice picked by the HIGH hash bits, offset by the LOW hash bits) and linear probing wraps INSIDE the slice, so slices are
 * independent hash tables: the table size follows the load factor in steps of one slice instead of doubling, a slice can be
 * built by one GPU of a group and exchanged as a plain byte range, and a slice-ordered insert keeps its lines in L2.
 * filter: a blocked Bloom filter over the k-mers of the block (two bits in one 32-bit word per k-mer), tested by the seed
 * scan before it touches the table -- the role of the reference's Hash_Check_Array (OIC/overlapInCore.H:329).  Like it, it
 * never changes a result: a false positive costs one probe.  The filter is cut into FILTER_REGIONS regions that follow the
 * slices (region = slice * FILTER_REGIONS / n_slices; n_slices is a multiple of FILTER_REGIONS), so the GPU that owns a
 * range of slices also owns whole regions of the filter. */
constexpr uint32_t FILTER_REGIONS = 8;
struct TableGeom {
  uint32_t n_slices;                // multiple of FILTER_REGIONS
  uint32_t slice_bits;
  unsigned long long slice_mask;    // 2^slice_bits - 1
  uint32_t region_bits;             // log2(filter words per region)
};
__host__ __device__ __forceinline__ uint64_t geom_slots(const TableGeom &g) { return (uint64_t)g.n_slices << g.slice_bits; }
__device__ __forceinline__ uint32_t geom_slice(const TableGeom &g, uint64_t h) { return __umulhi((uint32_t)(h >> 32), g.n_slices); }
__device__ __forceinline__ uint64_t geom_home(const TableGeom &g, uint64_t h) {
  return ((uint64_t)geom_slice(g, h) << g.slice_bits) | (h & g.slice_mask);
}
__device__ __forceinline__ uint64_t geom_next(const TableGeom &g, uint64_t s) { return (s & ~g.slice_mask) | ((s + 1) & g.slice_mask); }
/* filter word and the two bits of a k-mer (from a second mix of the hash, so the bits are independent of the slot) */
__host__ __device__ __forceinline__ uint64_t geom_filter_words(const TableGeom &g) { return (uint64_t)FILTER_REGIONS << g.region_bits; }
__device__ __forceinline__ void geom_filter(const TableGeom &g, uint64_t h, uint32_t &word, uint32_t &bits) {
  const uint64_t h2 = h * 0x9e3779b97f4a7c15ULL;
  const uint32_t region = (uint32_t)(h >> 61);   // = slice * FILTER_REGIONS / n_slices, because n_slices is a multiple of FILTER_REGIONS = 8
  word = (region << g.region_bits) | ((uint32_t)(h2 >> 32) & ((1u << g.region_bits) - 1u));
  bits = (1u << ((uint32_t)h2 & 31u)) | (1u << (((uint32_t)h2 >> 5) & 31u));
}

constexpr int READ_PAD_BASES = 64;   // zero bases after every read
constexpr int READ_LEAD_BASES = 64;  // zero bases before the first read

constexpr unsigned FULL = 0xffffffffu;

// phase clocks of the pair kernel: compile with -DOIC_PHASE_CLOCKS to enable (perturbs scheduling)
#ifdef OIC_PHASE_CLOCKS
#define OIC_CLOCK() clock64()
#else
#define OIC_CLOCK() 0LL
#endif

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

/* 32 bases starting at stream position p (p >= 0) */
__device__ __forceinline__ uint64_t ext32(const uint64_t *__restrict__ W, int64_t p) {
  int64_t w = p >> 5;
  int s = (int)(p & 31) * 2;
  uint64_t lo = W[w];
  if (s == 0) return lo;
  uint64_t hi = W[w + 1];
  return (lo >> s) | (hi << (64 - s));
}

/* number of equal bases comparing A[pa + i] with T[pt + i], i = 0.. , at most `limit` */
__device__ __forceinline__ int match_fwd(const uint64_t *__restrict__ A, int64_t pa, const uint64_t *__restrict__ T,
                                         int64_t pt, int limit) {
  int r = 0;
  while (r < limit) {
    uint64_t x = ext32(A, pa + r) ^ ext32(T, pt + r);
    if (x) { r += (__ffsll((long long)x) - 1) >> 1; break; }
    r += 32;
  }
  return r < limit ? r : limit;
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
/* inclusive scan */
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
  int l = lane_id();
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(FULL, v, o);
    if (l >= o) v += t;
  }
  return v;
}

/* counters shared by all kernels of a context (one cudaMemcpy to read them) */
struct DevCounters {
  unsigned long long n_inserts, n_distinct;
  unsigned long long n_query_bases, n_windows, n_found, n_refs, n_add_ref;
  unsigned long long n_runs, n_pairs;
  unsigned long long ext_calls, ext_cells, ext_words;
  unsigned long long kmer_hits_without_olap, kmer_hits_with_olap, kmer_hits_skipped, multi_overlap;
  unsigned long long total_overlaps, contained, dovetail;
  // transient
  unsigned int rec_chunks;     // bump allocator of seed-run chunks
  unsigned int rec_overflow;   // set when the run buffer was too small
  unsigned int n_big_lists;
  unsigned int pair_cursor;    // work queue of the pair kernel
  unsigned int n_retry;        // pairs whose edit rows did not fit the per-warp scratch
  unsigned int n_out;          // overlap records written
  unsigned int bad_base;       // a non-ACGT base was seen while packing
  unsigned int out_overflow;
  // speculative second round of the pair stage (oic_extend.cuh)
  unsigned int n_pend, pend_cursor;        // pairs handed over after their first extension
  unsigned int n_spec_tasks, task_cursor;  // (pair, seed) extension tasks of those pairs
  unsigned int n_spec_used;                // tasks whose result the commit pass consumed
  unsigned int n_wide_tasks, wide_cursor;  // tasks rerun with the generic row loop
  unsigned int n_ovf;          // hash build: occurrences of k-mers that occur more than once (entries of the overflow list)
  unsigned int n_multi;        // hash build: k-mers that occur more than once
  unsigned int occ_cursor;     // hash build: bump allocator of `occ`
  unsigned int n_mid_lists;    // hash build: lists of 7..32 occurrences (sorted by a warp each)
  unsigned int n_strands;      // -l path: query strands that have partners
  unsigned long long bad_short_window, bad_long_window;   // -w: overlaps rejected by the 50- / the 100-base window
  unsigned long long dpool_used;           // ints of the delta pool handed out
  // phase clocks of the pair kernel (SM cycles summed over warps; t_warp_max = busiest warp)
  unsigned long long t_dp, t_tb, t_prune, t_warp_sum, t_warp_max, t_pair_max;
};

}  // namespace oic
#endif
