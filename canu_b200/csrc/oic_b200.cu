/*
 * oic_b200.cu -- context, host orchestration and the C ABI (include/oic_b200.h) of the B200-native
 * overlapInCore path.  Kernels live in oic_build.cuh (pack + k-mer index), oic_search.cuh (seed scan, run
 * grouping, Add_Match replay), oic_extend.cuh (banded prefix edit distance + overlap output) and oic_count.cuh (canonical
 * K-mer counts of the loaded reads: the frequent-k-mer list the pipeline takes from meryl).
 *
 * Reference call stack replaced: OverlapDriver -> Build_Hash_Index -> Process_Overlaps ->
 * Find_Overlaps -> Process_String_Olaps -> Process_Matches -> Extend_Alignment -> Output_Overlap.
 * There is no CPU implementation of any stage in this library: every entry point needs a CUDA device.
 */
#include <dlfcn.h>
#include <math.h>
#include <nccl.h>   /* types and prototypes only: the library is dlopen'ed at the first communicator call */
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/oic_b200.h"
#include "oic_extend.cuh"
#include "oic_count.cuh"
#include "oic_store.cuh"

using namespace oic;

namespace {

thread_local std::string g_create_error;   // error text of the last failed context-less call on THIS thread (oic_last_error(NULL))

// Bumped by every cudaMalloc / cudaFree of this library: what cudaMemGetInfo last said stays true while it stands still.  The
// steady state of a job stream (same block sizes, buffers grown to size) then makes NO call that enters the kernel-mode driver
// -- cudaMalloc, cudaFree and cudaMemGetInfo do, and on a box whose other GPUs are busy they were seen to take 5-100 ms at random.
std::atomic<uint64_t> g_alloc_epoch{1};

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    g_alloc_epoch.fetch_add(1, std::memory_order_relaxed);
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { e = cudaMalloc(&p, bytes); want = bytes; }
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) { cudaFree(p); g_alloc_epoch.fetch_add(1, std::memory_order_relaxed); } p = nullptr; cap = 0; }
  template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// The overlap records on the host: a growable PINNED buffer (the device-to-host copy of a short-read block's million records
// ran at 4 GB/s into a pageable std::vector, 25 GB/s into this).  Grows by doubling; never shrinks until the context closes.
struct HostRecs {
  oic_ovb_record_t *p = nullptr;
  size_t n = 0, cap = 0;
  bool pinned = true;   // false: the host refused to pin that much (ulimit -l, fragmentation): plain memory, slower copies, same results
  size_t size() const { return n; }
  const oic_ovb_record_t *data() const { return p; }
  void clear() { n = 0; }
  void drop() { if (p) { if (pinned) cudaFreeHost(p); else free(p); } p = nullptr; }
  // room for `more` records behind the current ones; returns where they go, or nullptr when the host has no memory for them
  oic_ovb_record_t *grow(size_t more) {
    if (n + more > cap) {
      size_t want = std::max<size_t>(2 * cap, n + more + 4096);
      oic_ovb_record_t *q = nullptr;
      bool q_pinned = true;
      if (cudaMallocHost(reinterpret_cast<void **>(&q), want * sizeof(oic_ovb_record_t)) != cudaSuccess) {
        cudaGetLastError();
        want = n + more + 4096;
        q = static_cast<oic_ovb_record_t *>(malloc(want * sizeof(oic_ovb_record_t)));
        q_pinned = false;
        if (!q) return nullptr;
      }
      if (n) memcpy(q, p, n * sizeof(oic_ovb_record_t));
      drop();
      p = q; cap = want; pinned = q_pinned;
    }
    oic_ovb_record_t *at = p + n;
    n += more;
    return at;
  }
  void release() { drop(); n = cap = 0; pinned = true; }
};

// ---- host tables: Edit_Match_Limit (LIBO/Binomial_Bound.C:47-107,119-194, BITS == 21 slope) ------------------
// Smallest n >= start such that more than e errors in n bases at rate p is not "improbable" (1e-4).
int binomial_bound(int e, double p, int start) {
  const double PROB_BOUND = 1e-4, NORMAL_THOLD = 3.62;
  const double q = 1.0 - p;
  if (start < e) start = e;
  for (int n = start; n < MAX_READLEN; n++) {
    if (n <= 35) {   // exact binomial tail
      double sum = 0.0, p_power = 1.0, q_power = pow(q, n);
      int binom = 1, ct = 0;
      for (int k = 0; k < e && 1.0 - sum > PROB_BOUND; k++) {
        sum += binom * p_power * q_power;
        binom *= n - ct;
        binom /= ++ct;
        p_power *= p;
        q_power /= q;
      }
      if (1.0 - sum > PROB_BOUND) return n;
    } else {         // normal test first, then the Poisson sum
      const double np = n * p;
      const double z = (e - 0.5 - np) / sqrt(np * q);
      if (z <= NORMAL_THOLD) return n;
      double sum = 0.0, mu_power = 1.0, factorial = 1.0;
      const double poisson_coeff = exp(-np);
      for (int k = 0; k < e; k++) {
        sum += mu_power * poisson_coeff / factorial;
        mu_power *= np;
        factorial *= k + 1;
      }
      if (1.0 - sum > PROB_BOUND) return n;
    }
  }
  return MAX_READLEN;
}

void build_match_limit(double erate, int n_entries, std::vector<int32_t> &ml) {
  const int max_errors = 1 + (int)ceil(erate * MAX_READLEN);   // LIBO/prefixEditDistance.C:39
  ml.assign(std::max(n_entries, 2), 0);
  const int exact = std::min(max_errors, 2000);
  int e = 2, start = 1;
  for (; e < exact && e < n_entries; e++) {
    start = binomial_bound(e - 1, erate, start);
    ml[e] = start - 1;
  }
  if (e < n_entries) {
    // the reference fills the rest linearly from the last exact entry (which we must have computed)
    const double sl = 0.982064188397525 / erate + 0.067835741959926;
    double vl = ml[e - 1] + sl;
    for (; e < n_entries; e++) {
      ml[e] = (int32_t)ceil(vl);
      vl += sl;
    }
  }
}

}  // namespace

struct oic_ctx {
  oic_params_t P;
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  uint32_t debug = 0;

  // reads
  uint32_t first_id = 0, n_reads = 0;
  std::vector<int64_t> h_pos;
  std::vector<uint32_t> h_len;
  std::vector<uint8_t> h_hash_ok, h_ref_ok;
  uint64_t stream_words = 0;
  uint64_t n_off = 0;          // word offset of the N-mask streams behind d_fwd / d_rc (0: no loaded read holds an N)
  uint32_t max_read_len = 0;
  DevBuf d_fwd, d_rc, d_pos, d_len;

  // host tables on device
  int max_rows = 0;
  DevBuf d_match_limit;

  // skip k-mers (both orientations, as keys)
  std::vector<uint64_t> skip_keys;
  DevBuf d_skip_keys;

  int sm_count = 148;

  // communicator (NCCL), attached by oic_comm_init_rank / oic_comm_init_all
  ncclComm_t comm = nullptr;
  int comm_n = 0, comm_rank = 0;
  DevBuf d_xchg;               // small device staging of the collectives (meta, per-rank counts)

  // hash block
  bool hash_ready = false;
  oic_hash_meta_t hb = {};
  TableGeom geom = {};
  DevBuf d_word_read;           // read index per stream word (k_word_read, once per load)
  DevBuf d_parts, d_part_ctr;   // build: the block's windows as partitioned {key, occurrence} records
  DevBuf d_table, d_occ, d_str_flags, d_str_use, d_filter, d_ovf_slot, d_ovf_occ, d_multi_list, d_first_occ, d_cnt, d_mid_list, d_blk, d_big_list, d_occ_tmp, d_hist;
  // identity of the block in the table (oic_set_job sweeps reuse it): requested range + the two load limits
  uint32_t hbk_bgn = 0, hbk_end = 0, hbk_strings = 0; uint64_t hbk_datalen = 0;
  bool job_mode = false;   // set by oic_set_job: only a job sweep reuses a block; a plain oic_build_hash_index always builds
  unsigned long long mem_avail = 0;   // device memory the pair stage may plan with (free + what it already holds)
  uint64_t mem_avail_epoch = 0;       // g_alloc_epoch when it was asked: any allocation since makes it stale
  bool have_filter = false;      // Bloom filter of the current hash block is valid

  // search buffers
  DevBuf d_q_order;
  DevBuf d_strand;              // -l path: per-pair ordering scratch (k_process_strands)
  DevBuf d_quals;               // -w: one quality byte per base at the base's stream position
  bool have_quals = false;
  DevBuf d_qs_read, d_qs_tile_base, d_tile_empty, d_recs, d_chunk_next, d_chunk_cnt, d_chunk_qs, d_qs_first, d_qs_nrec,
      d_qs_base, d_qs_screen, d_bufA, d_bufB, d_scratch, d_pairs, d_info, d_out, d_retry, d_vals, d_rowmeta, d_deltas, d_dirs, d_order, d_hist2, d_pstate, d_pend, d_tasks, d_results, d_alive, d_dpool, d_wide;
  DevBuf d_ctr, d_flag;
  DevBuf d_ascii, d_asc_off;                  // staging of oic_load_reads (kept: a 200 MB cudaMalloc/cudaFree per call costs milliseconds)
  double ml_erate = -1.0; int ml_rows = -1;   // what d_match_limit was built for
  uint64_t rec_cap = 0;       // run records per batch
  uint64_t batch_bases = 0;

  // canonical K-mer counts (oic_count_kmers)
  DevBuf d_kc_table, d_kc_out_keys, d_kc_out_cnt, d_kc_n, d_kc_hist, d_kc_big;
  std::vector<uint64_t> kc_hist;   // kc_hist[i] = distinct canonical K-mers seen exactly i times (last oic_count_kmers)
  std::vector<std::pair<uint64_t, uint32_t>> freq;   // {key (first base high), count}, ascending
  float kc_ms = 0;

  HostRecs overlaps;
  oic_stats_t stats = {};
  oic_timings_t tm = {};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  cudaEvent_t ev_call[2] = {nullptr, nullptr};   // whole-call timers (oic_process_overlaps, oic_count_kmers)
  cudaEvent_t ev_k[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // inner kernel timers of the pair stage

  // debug dumps of the last oic_process_overlaps call
  std::vector<int64_t> dbg_pairs;
  std::vector<int32_t> dbg_matches;

  int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    err = buf;
    return code;
  }
};

#define CK(call)                                                                                             \
  do {                                                                                                       \
    cudaError_t e_ = (call);                                                                                 \
    if (e_ != cudaSuccess) return c->fail(OIC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

namespace {

inline unsigned nblocks(uint64_t n, unsigned per) { return (unsigned)((n + per - 1) / per); }

struct StageTimer {   // stops itself on every exit path (early error returns included)
  oic_ctx *c;
  float *acc;
  bool running = true;
  StageTimer(oic_ctx *c_, float *acc_) : c(c_), acc(acc_) { cudaEventRecord(c->ev[0], c->stream); }
  ~StageTimer() { stop(); }
  void stop() {
    if (!running) return;
    running = false;
    cudaEventRecord(c->ev[1], c->stream);
    if (cudaEventSynchronize(c->ev[1]) != cudaSuccess) return;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]) == cudaSuccess) *acc += ms;
  }
};

int read_counters(oic_ctx *c, DevCounters &h) {
  CK(cudaMemcpyAsync(&h, c->d_ctr.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->tm.d2h_bytes += sizeof(h);
  return 0;
}
int zero_counters(oic_ctx *c) {
  CK(cudaMemsetAsync(c->d_ctr.p, 0, sizeof(DevCounters), c->stream));
  return 0;
}

// Bloom filter of the block's k-mers (oic_dev_common.cuh TableGeom), filled by k_insert; OIC_FILTER=0 makes the scan ignore it
bool use_filter() {
  static const char *env = getenv("OIC_FILTER");
  return !(env && env[0] == '0');
}

// Table geometry for a block of n_keys k-mer windows (+ skip k-mers): slots ~ factor x windows in steps of one slice
// (32..63 slices, a multiple of FILTER_REGIONS = 8 so that 1, 2, 4 or 8 GPUs own whole slices and whole filter regions;
// slices of at least 4096 slots so that the load of a single slice cannot stray far from the mean), filter ~ 4 bits per window.
TableGeom table_geometry(uint64_t n_keys) {
  static const char *tf = getenv("OIC_TABLE_FACTOR");   // slots per k-mer window (default 1.5: load ~0.55 on a 1x block)
  static const char *ff = getenv("OIC_FILTER_BITS");    // filter bits per k-mer window before rounding up to a power of two (default 4)
  const double factor = tf ? std::max(1.1, atof(tf)) : 1.5;
  const double fbits = ff ? std::max(1.0, atof(ff)) : 4.0;
  const uint64_t want = (uint64_t)(factor * (double)n_keys) + 16;
  TableGeom g;
  g.slice_bits = 12;
  while ((want >> g.slice_bits) >= 64) g.slice_bits++;
  g.slice_mask = (1ULL << g.slice_bits) - 1;
  uint64_t ns = (want + g.slice_mask) >> g.slice_bits;
  ns = std::max<uint64_t>(FILTER_REGIONS, (ns + FILTER_REGIONS - 1) / FILTER_REGIONS * FILTER_REGIONS);
  g.n_slices = (uint32_t)ns;
  const uint64_t fwords = (uint64_t)(fbits * (double)n_keys / 32.0) + 1;
  g.region_bits = 6;
  while (((uint64_t)FILTER_REGIONS << g.region_bits) < fwords && g.region_bits < 28) g.region_bits++;
  return g;
}

__global__ void k_table_init(Slot *t, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) reinterpret_cast<uint4 *>(t)[i] = make_uint4(0xffffffffu, 0xffffffffu, 0u, 0u);
}

// ---- hash build ------------------------------------------------------------------------------------------
uint64_t block_windows(const oic_ctx *c, uint32_t ri_bgn, uint32_t n_strings, const std::vector<uint8_t> &use) {
  const uint32_t K = (uint32_t)c->P.kmer_len;
  uint64_t n = 0;
  for (uint32_t s = 0; s < n_strings; s++)
    if (use[s] && c->h_len[ri_bgn + s] >= K) n += c->h_len[ri_bgn + s] - K + 1;
  return n;
}

// Builds the index over string slots [0, n_strings) starting at read index ri_bgn; use[] says which are loaded.
// part / n_parts: only the k-mers whose home slice belongs to share `part` of `n_parts` are inserted (the sharded build of a
// GPU group; 0 / 1 = everything).  The slices of the other shares stay empty, the occurrence lists built are those of this share.
int build_block(oic_ctx *c, uint32_t ri_bgn, uint32_t n_strings, const std::vector<uint8_t> &use, DevCounters &hc, uint32_t part = 0,
                uint32_t n_parts = 1) {
  const oic_params_t &P = c->P;
  const int K = P.kmer_len;
  const uint64_t n_windows = block_windows(c, ri_bgn, n_strings, use);
  if (n_windows >= (1ull << 30)) return c->fail(OIC_ERR_CAPACITY, "hash block has %llu k-mers (limit 2^30)", (unsigned long long)n_windows);
  if (n_strings >= (1u << 30)) return c->fail(OIC_ERR_CAPACITY, "hash block has %u reads (limit 2^30)", n_strings);
  const TableGeom geom = table_geometry(n_windows + c->skip_keys.size());
  const uint64_t slots = geom_slots(geom), fwords = geom_filter_words(geom);
  if (geom.n_slices % n_parts) return c->fail(OIC_ERR_ARG, "internal: %u table slices do not split over %u ranks", geom.n_slices, n_parts);
  const uint64_t share = n_windows;   // list scratch: the worst case (every window a repeat, all of them in this share)

  CK(c->d_table.ensure(slots * sizeof(Slot)));
  CK(c->d_filter.ensure(fwords * 4));
  CK(c->d_str_flags.ensure(((size_t)n_strings + 8) & ~(size_t)3));
  CK(c->d_str_use.ensure((size_t)n_strings + 4));
  const int64_t region_pos = c->h_pos[ri_bgn];
  const uint32_t ri_last = ri_bgn + n_strings - 1;
  const int64_t region_end = c->h_pos[ri_last] + c->h_len[ri_last];
  const uint64_t region_words = (uint64_t)((region_end - region_pos + 31) >> 5);
  // worst case: every window belongs to a repeated k-mer
  CK(c->d_ovf_slot.ensure((share + 32) * sizeof(uint32_t)));
  CK(c->d_ovf_occ.ensure((share + 32) * sizeof(uint64_t)));
  CK(c->d_multi_list.ensure((share / 2 + 32) * sizeof(uint32_t)));
  CK(c->d_first_occ.ensure((share / 2 + 32) * sizeof(uint64_t)));
  CK(c->d_cnt.ensure((share / 2 + 32) * sizeof(uint32_t)));
  CK(c->d_mid_list.ensure((share / (SMALL_LIST + 1) + 32) * sizeof(uint32_t)));
  CK(c->d_big_list.ensure((share / 33 + 32) * sizeof(uint32_t)));
  CK(c->d_occ.ensure((n_windows + 32) * sizeof(uint64_t)));

  if (zero_counters(c)) return OIC_ERR_CUDA;
  {   // (a rank of a group only clears the slices / filter regions it fills: the others arrive by the exchange)
    const uint64_t per = slots / n_parts, fper = fwords / n_parts;
    k_table_init<<<nblocks(per, 256), 256, 0, c->stream>>>(c->d_table.as<Slot>() + per * part, per);
    CK(cudaMemsetAsync(c->d_filter.as<uint32_t>() + fper * part, 0, fper * 4, c->stream));
  }
  CK(cudaMemsetAsync(c->d_str_flags.p, 0, ((size_t)n_strings + 8) & ~(size_t)3, c->stream));
  CK(cudaMemcpyAsync(c->d_str_use.p, use.data(), n_strings, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += n_strings;

  BuildArgs a;
  a.fwd = c->d_fwd.as<uint64_t>();
  a.n_off = c->n_off;
  a.rd_pos = c->d_pos.as<int64_t>();
  a.rd_len = c->d_len.as<uint32_t>();
  a.word_read = c->d_word_read.as<uint32_t>();
  a.str_use = c->d_str_use.as<uint8_t>();
  a.ri_bgn = ri_bgn;
  a.n_strings = n_strings;
  a.region_pos = region_pos;
  a.region_words = region_words;
  a.K = K;
  a.table = c->d_table.as<Slot>();
  a.geom = geom;
  a.slice_lo = geom.n_slices / n_parts * part;
  a.slice_hi = geom.n_slices / n_parts * (part + 1);
  a.filter = c->d_filter.as<uint32_t>();
  a.ovf_slot = c->d_ovf_slot.as<uint32_t>();
  a.ovf_occ = c->d_ovf_occ.as<uint64_t>();
  a.multi_list = c->d_multi_list.as<uint32_t>();
  a.first_occ = c->d_first_occ.as<uint64_t>();
  a.cnt = c->d_cnt.as<uint32_t>();
  a.mid_list = c->d_mid_list.as<uint32_t>();
  a.big_list = c->d_big_list.as<uint32_t>();
  a.occ = c->d_occ.as<uint64_t>();
  a.ctr = c->d_ctr.as<DevCounters>();

  // no host round trip between the kernels: the list kernels take their sizes from the device counters
  static const bool trace = getenv("OIC_PHASE_CLOCKS") != nullptr;
  auto tick = [&]() { if (trace) cudaStreamSynchronize(c->stream); return std::chrono::steady_clock::now(); };
  const auto tb0 = tick();
  if (n_parts > 1) {
    // a rank of a group owns 1/N of the windows: pack them first, then a dense insert (oic_build.cuh)
    CK(c->d_parts.ensure((n_windows + 32) * sizeof(PartRec)));   // worst case: every window hashes into this rank's slices
    CK(c->d_part_ctr.ensure(16));
    CK(cudaMemsetAsync(c->d_part_ctr.p, 0, 16, c->stream));
    const uint64_t cap = n_windows;
    k_compact_windows<<<nblocks(region_words * 32, INS_TILE), INS_THREADS, 0, c->stream>>>(a, c->d_parts.as<PartRec>(), c->d_part_ctr.as<uint32_t>());
    k_insert<true><<<nblocks(cap, INS_TILE), INS_THREADS, 0, c->stream>>>(a, c->d_parts.as<PartRec>(), c->d_part_ctr.as<uint32_t>());
  } else k_insert<false><<<nblocks(region_words * 32, INS_TILE), INS_THREADS, 0, c->stream>>>(a, nullptr, nullptr);
  const auto tb1 = tick();
  k_multi_prep<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  k_multi_count<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  k_multi_alloc<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  k_multi_fill<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  k_multi_sort<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  k_sort_mid<<<MULTI_GRID, 256, 0, c->stream>>>(a);
  const auto tb2 = tick();
  {
    const uint64_t per = slots / n_parts;   // the filter regions of this share follow its slices
    k_filter_build<<<nblocks(per, 256), 256, 0, c->stream>>>(c->d_table.as<Slot>(), per * part, per * (part + 1), geom, c->d_filter.as<uint32_t>());
  }
  if (trace) {
    const auto tb3 = tick();
    auto ms = [](auto x, auto y) { return std::chrono::duration<double, std::milli>(y - x).count(); };
    fprintf(stderr, "[oic] build_block share %u/%u: partition + insert %.2f ms | lists %.2f ms | filter %.2f ms (%llu windows, %llu slots)\n", part, n_parts,
            ms(tb0, tb1), ms(tb1, tb2), ms(tb2, tb3), (unsigned long long)n_windows, (unsigned long long)slots);
  }
  c->tm.launches += 11;
  if (read_counters(c, hc)) return OIC_ERR_CUDA;
  if (n_parts == 1 && (c->n_off ? hc.n_inserts > n_windows : hc.n_inserts != n_windows))
    return c->fail(OIC_ERR_STATE, "internal: %llu windows inserted, %llu expected", (unsigned long long)hc.n_inserts, (unsigned long long)n_windows);
  const uint32_t n_occ = hc.n_ovf + hc.n_multi;
  if (hc.occ_cursor != n_occ) return c->fail(OIC_ERR_STATE, "internal: %u list entries allocated, %u on the overflow list", hc.occ_cursor, n_occ);
  if (hc.n_big_lists) {
    CK(c->d_occ_tmp.ensure(((size_t)n_occ + 32) * sizeof(uint64_t)));
    unsigned grid = std::min<unsigned>(hc.n_big_lists, c->sm_count * 8);
    k_sort_big<<<grid, BIG_THREADS, 0, c->stream>>>(c->d_table.as<Slot>(), c->d_big_list.as<uint32_t>(), c->d_ctr.as<DevCounters>(), c->d_occ.as<uint64_t>(),
                                                    c->d_occ_tmp.as<uint64_t>());
    c->tm.launches += 1;
  }
  c->geom = geom;
  c->hb.table_slots = slots;
  c->hb.n_slices = geom.n_slices;
  c->hb.slice_bits = geom.slice_bits;
  c->hb.filter_words = fwords;
  c->hb.n_occ = n_occ;
  c->hb.n_strings = n_strings;
  c->hb.n_distinct = hc.n_distinct;
  c->have_filter = true;
  CK(cudaGetLastError());
  return 0;
}

// ---- NCCL, loaded on demand ------------------------------------------------------------------------------
struct NcclApi {
  void *handle = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclCommCount) CommCount = nullptr;
  decltype(&ncclCommUserRank) CommUserRank = nullptr;
  decltype(&ncclBroadcast) Broadcast = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  std::string err;
};
NcclApi *nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, []() {
    // a copy the process already holds (e.g. the one torch brought) first: two NCCL copies in one process do not share state
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { api.err = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
    api.handle = h;
#define OIC_NCCL_SYM(name) \
    api.name = reinterpret_cast<decltype(api.name)>(dlsym(h, "nccl" #name)); \
    if (!api.name) { api.err = "libnccl.so.2 lacks nccl" #name; api.handle = nullptr; return; }
    OIC_NCCL_SYM(GetUniqueId) OIC_NCCL_SYM(CommInitRank) OIC_NCCL_SYM(CommInitAll) OIC_NCCL_SYM(CommDestroy) OIC_NCCL_SYM(CommCount)
    OIC_NCCL_SYM(CommUserRank) OIC_NCCL_SYM(Broadcast) OIC_NCCL_SYM(AllGather) OIC_NCCL_SYM(GroupStart) OIC_NCCL_SYM(GroupEnd)
    OIC_NCCL_SYM(GetErrorString)
#undef OIC_NCCL_SYM
  });
  return api.handle ? &api : nullptr;
}
#define NK(call)                                                                                              \
  do {                                                                                                        \
    ncclResult_t r_ = (call);                                                                                 \
    if (r_ != ncclSuccess) return c->fail(OIC_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, nccl_api()->GetErrorString(r_)); \
  } while (0)

// the communicator a collective call works on: the caller's ncclComm_t, or the one attached to the context
int resolve_comm(oic_ctx *c, void *arg, ncclComm_t &comm, int &n, int &rank) {
  NcclApi *api = nccl_api();
  if (!api) return c->fail(OIC_ERR_CUDA, "NCCL unavailable: %s", nccl_api() ? "" : "libnccl.so.2 could not be loaded");
  if (arg) {
    comm = reinterpret_cast<ncclComm_t>(arg);
    NK(api->CommCount(comm, &n));
    NK(api->CommUserRank(comm, &rank));
    return 0;
  }
  if (!c->comm) return c->fail(OIC_ERR_STATE, "no communicator: call oic_comm_init_rank / oic_comm_init_all first, or pass an ncclComm_t");
  comm = c->comm; n = c->comm_n; rank = c->comm_rank;
  return 0;
}

// which reads the reference's load loop would take, ignoring the hash-load limit (Build_Hash_Index.C:538-616)
void hash_load_plan(const oic_ctx *c, uint32_t bgn_id, uint32_t end_id, std::vector<uint8_t> &use) {
  const oic_params_t &P = c->P;
  use.clear();
  uint64_t total_len = 0;
  uint32_t cur = bgn_id;
  while (use.size() < P.max_hash_strings && total_len < P.max_hash_data_len && cur <= end_id) {
    const uint32_t ri = cur - c->first_id;
    uint8_t u = c->h_hash_ok[ri] && (int64_t)c->h_len[ri] >= (int64_t)P.min_olap_len;
    if (u) total_len += c->h_len[ri] + 1;
    use.push_back(u);
    cur++;
  }
}

// Mark_Skip_Kmers on the finished table (Empty bits, screened ends, absent k-mers): every rank of a group runs it on its own copy
int mark_skip(oic_ctx *c, uint32_t ri_bgn, uint32_t n_strings, unsigned int &has_empty) {
  const oic_params_t &P = c->P;
  has_empty = 0;
  if (c->skip_keys.empty()) return 0;
  CK(c->d_skip_keys.ensure(c->skip_keys.size() * 8));
  CK(c->d_flag.ensure(16));
  CK(cudaMemsetAsync(c->d_flag.p, 0, 16, c->stream));
  CK(cudaMemcpyAsync(c->d_skip_keys.p, c->skip_keys.data(), c->skip_keys.size() * 8, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += c->skip_keys.size() * 8;
  k_mark_skip<<<nblocks(c->skip_keys.size(), 128), 128, 0, c->stream>>>(
      c->d_table.as<Slot>(), c->geom, c->d_filter.as<uint32_t>(), c->d_skip_keys.as<uint64_t>(), c->skip_keys.size(), c->d_occ.as<uint64_t>(),
      c->d_len.as<uint32_t>(), ri_bgn, n_strings, c->d_str_flags.as<uint8_t>(), P.kmer_len, P.use_hopeless_check, c->d_flag.as<unsigned int>());
  c->tm.launches += 1;
  CK(cudaMemcpyAsync(&has_empty, c->d_flag.p, 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  return 0;
}

// lists of this rank's slices were allocated from 0: shift their start indices to the rank's segment of the gathered occ array
__global__ void k_rebase_lists(Slot *table, uint64_t slot_lo, uint64_t slot_hi, uint32_t base) {
  const uint64_t i = slot_lo + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= slot_hi) return;
  const unsigned long long p = table[i].pay;
  if (table[i].key != EMPTY_KEY && pay_is_list(p)) table[i].pay = p + base;   // start lives in the low 32 bits and cannot carry out (n_occ < 2^30)
}

struct ShardCounts { unsigned long long n_inserts, n_distinct, n_occ, pad; };

int replicate_impl(oic_ctx *c, ncclComm_t comm, int n, int rank, int root);

/* Build_Hash_Index for one context (comm_arg == nullptr and no group) or collectively for a GPU group (sharded = true). */
int build_hash_index_impl(oic_ctx *c, void *comm_arg, uint32_t bgn_id, uint32_t end_id, uint32_t *last_id, bool sharded = false) {
  if (!c) return OIC_ERR_ARG;
  cudaSetDevice(c->device);
  const oic_params_t &P = c->P;
  ncclComm_t comm = nullptr;
  int n_ranks = 1, rank = 0;
  if (sharded) {
    int rc = resolve_comm(c, comm_arg, comm, n_ranks, rank);
    if (rc) return rc;
  }
  if (c->job_mode && c->hash_ready && c->hbk_bgn == bgn_id && c->hbk_end == end_id && c->hbk_strings == P.max_hash_strings && c->hbk_datalen == P.max_hash_data_len) {
    if (last_id) *last_id = c->hb.last_id;   // the same request as the block in the table: nothing to do (every rank of a group agrees)
    return OIC_OK;
  }
  c->hash_ready = false;
  if (bgn_id < c->first_id || end_id >= c->first_id + c->n_reads || c->n_reads == 0)
    return c->fail(OIC_ERR_ARG, "hash range %u-%u outside the loaded reads %u-%u", bgn_id, end_id, c->first_id, c->first_id + c->n_reads - 1);
  StageTimer t(c, &c->tm.ms_build);
  std::vector<uint8_t> use;
  hash_load_plan(c, bgn_id, end_id, use);
  uint32_t n_strings = (uint32_t)use.size();
  if (n_strings == 0) { if (last_id) *last_id = end_id; return c->fail(OIC_ERR_ARG, "empty hash range"); }
  const uint64_t hash_entry_limit = (uint64_t)(P.hash_load * (double)((uint64_t)1 << P.hash_bits) * 21);   // Build_Hash_Index.C:474
  const uint32_t ri_bgn = bgn_id - c->first_id;
  NcclApi *api = n_ranks > 1 ? nccl_api() : nullptr;

  DevCounters hc;
  unsigned long long tot_inserts = 0, tot_distinct = 0;
  bool root_path = false;   // group only: build on rank 0 and broadcast (rank count does not divide the slices, or the load limit trips)
  if (n_ranks > 1) {
    const TableGeom g = table_geometry(block_windows(c, ri_bgn, n_strings, use) + c->skip_keys.size());
    root_path = (g.n_slices % (uint32_t)n_ranks) != 0 || (FILTER_REGIONS % (uint32_t)n_ranks) != 0 || getenv("OIC_SHARDED_BUILD_OFF") != nullptr;
  }
  if (n_ranks > 1 && !root_path) {
    // ---- every rank builds the slices it owns, then the slices / list segments / filter regions are exchanged
    static const bool trace = getenv("OIC_PHASE_CLOCKS") != nullptr;
    const auto tp0 = std::chrono::steady_clock::now();
    int rcode = build_block(c, ri_bgn, n_strings, use, hc, (uint32_t)rank, (uint32_t)n_ranks);
    if (rcode) return rcode;
    const auto tp1 = std::chrono::steady_clock::now();
    CK(c->d_xchg.ensure(sizeof(ShardCounts) * (size_t)n_ranks + 256));
    std::vector<ShardCounts> cnt(n_ranks);
    ShardCounts mine = {hc.n_inserts, hc.n_distinct, c->hb.n_occ, 0};
    ShardCounts *dx = c->d_xchg.as<ShardCounts>();
    CK(cudaMemcpyAsync(dx + rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, c->stream));
    NK(api->AllGather(dx + rank, dx, sizeof(ShardCounts), ncclChar, comm, c->stream));
    CK(cudaMemcpyAsync(cnt.data(), dx, sizeof(ShardCounts) * n_ranks, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    unsigned long long base = 0, tot_occ = 0;
    for (int q = 0; q < n_ranks; q++) {
      if (q < rank) base += cnt[q].n_occ;
      tot_occ += cnt[q].n_occ; tot_inserts += cnt[q].n_inserts; tot_distinct += cnt[q].n_distinct;
    }
    if (tot_distinct >= hash_entry_limit) root_path = true;   // the cut needs the global first-occurrence histogram: rare, done on one rank
    else {
      const uint64_t slots = c->hb.table_slots, per = slots / n_ranks;
      const uint64_t fper = c->hb.filter_words / n_ranks;
      const uint64_t my_occ = c->hb.n_occ;
      // own list segment out of the way (segments land at their global offsets), own list starts shifted to the segment
      CK(c->d_occ_tmp.ensure((my_occ + 32) * 8));
      if (my_occ) CK(cudaMemcpyAsync(c->d_occ_tmp.p, c->d_occ.p, my_occ * 8, cudaMemcpyDeviceToDevice, c->stream));
      if (base) k_rebase_lists<<<nblocks(per, 256), 256, 0, c->stream>>>(c->d_table.as<Slot>(), per * rank, per * (rank + 1), (uint32_t)base);
      static const char *xm = getenv("OIC_EXCHANGE");   // experiment knob: "bcast" = per-rank broadcasts instead of ncclAllGather
      const bool use_bcast = xm && strcmp(xm, "bcast") == 0;
      auto tsync = [&]() { if (trace) cudaStreamSynchronize(c->stream); return std::chrono::steady_clock::now(); };
      const auto tx0 = tsync();
      static const bool grouped = getenv("OIC_NCCL_GROUP") != nullptr;   // (measured: one group of the four collectives is slower than issuing them back to back)
      if (!trace && grouped) NK(api->GroupStart());
      if (use_bcast) {
        for (int q = 0; q < n_ranks; q++) NK(api->Broadcast(c->d_table.as<Slot>() + per * q, c->d_table.as<Slot>() + per * q, per * sizeof(Slot), ncclChar, q, comm, c->stream));
      } else NK(api->AllGather(c->d_table.as<Slot>() + per * rank, c->d_table.p, per * sizeof(Slot), ncclChar, comm, c->stream));
      const auto tx1 = tsync();
      NK(api->AllGather(c->d_filter.as<uint32_t>() + fper * rank, c->d_filter.p, fper * 4, ncclChar, comm, c->stream));
      const auto tx2 = tsync();
      unsigned long long off = 0;
      for (int q = 0; q < n_ranks; q++) {
        if (cnt[q].n_occ) NK(api->Broadcast(c->d_occ_tmp.p, c->d_occ.as<uint64_t>() + off, cnt[q].n_occ * 8, ncclChar, q, comm, c->stream));
        off += cnt[q].n_occ;
      }
      if (!trace && grouped) NK(api->GroupEnd());
      if (trace) {
        const auto tx3 = tsync();
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[oic rank %d] exchange (ungrouped, synchronised for the trace): table %.2f ms | filter %.2f ms | lists %.2f ms\n", rank, ms(tx0, tx1), ms(tx1, tx2), ms(tx2, tx3));
      }
      if (trace) {
        const auto tp2 = std::chrono::steady_clock::now();
        cudaStreamSynchronize(c->stream);
        const auto tp3 = std::chrono::steady_clock::now();
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[oic rank %d] sharded build: own slices %.2f ms | counts + rebase + enqueue %.2f ms | exchange %.2f ms (table %.0f MB, lists %.0f MB)\n",
                rank, ms(tp0, tp1), ms(tp1, tp2), ms(tp2, tp3), slots * 16.0 / 1e6, tot_occ * 8.0 / 1e6);
      }
      c->tm.launches += 1;
      c->hb.n_occ = tot_occ;
      c->hb.n_distinct = tot_distinct;
      const uint64_t n_windows = block_windows(c, ri_bgn, n_strings, use);
      if (c->n_off ? tot_inserts > n_windows : tot_inserts != n_windows)
        return c->fail(OIC_ERR_STATE, "internal: %llu windows inserted by the group, %llu expected", tot_inserts, (unsigned long long)n_windows);
    }
  }
  if (n_ranks > 1 && root_path) {
    // ---- rank 0 builds alone (load cut-off included), the others receive
    int rcode = 0;
    uint32_t last = 0;
    if (rank == 0) {
      t.stop();
      rcode = build_hash_index_impl(c, nullptr, bgn_id, end_id, &last, false);
    }
    // (a failed root still enters the collective so that nobody hangs; the meta carries the failure)
    if (rank == 0 && rcode) c->hash_ready = false;
    const int r2 = replicate_impl(c, comm, n_ranks, rank, 0);
    if (rcode) return rcode;
    if (r2) return r2;
    c->hbk_bgn = bgn_id; c->hbk_end = end_id; c->hbk_strings = P.max_hash_strings; c->hbk_datalen = P.max_hash_data_len;
    if (last_id) *last_id = c->hb.last_id;
    return OIC_OK;
  }
  if (n_ranks == 1) {
    int rcode = build_block(c, ri_bgn, n_strings, use, hc);
    if (rcode) return rcode;
    if (hc.n_distinct >= hash_entry_limit) {
      // the load loop stops taking reads once Hash_Entries >= limit: find the read at which that happens
      CK(c->d_hist.ensure((size_t)n_strings * 4));
      CK(cudaMemsetAsync(c->d_hist.p, 0, (size_t)n_strings * 4, c->stream));
      k_first_occ<<<nblocks(c->hb.table_slots, 256), 256, 0, c->stream>>>(c->d_table.as<Slot>(), c->hb.table_slots, c->d_occ.as<uint64_t>(),
                                                                          c->d_hist.as<uint32_t>());
      c->tm.launches += 1;
      std::vector<uint32_t> hist(n_strings);
      CK(cudaMemcpyAsync(hist.data(), c->d_hist.p, (size_t)n_strings * 4, cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      c->tm.d2h_bytes += (uint64_t)n_strings * 4;
      uint64_t cum = 0;
      uint32_t cut = n_strings;
      for (uint32_t s = 0; s < n_strings; s++) {
        cum += hist[s];
        if (cum >= hash_entry_limit) { cut = s + 1; break; }
      }
      if (cut < n_strings) {
        n_strings = cut;
        use.resize(n_strings);
        rcode = build_block(c, ri_bgn, n_strings, use, hc);
        if (rcode) return rcode;
      }
    }
    tot_inserts = hc.n_inserts; tot_distinct = hc.n_distinct;
  }
  // screened k-mers
  unsigned int has_empty = 0;
  if (mark_skip(c, ri_bgn, n_strings, has_empty)) return OIC_ERR_CUDA;
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  t.stop();
  c->hb.bgn_id = bgn_id;
  c->hb.last_id = bgn_id + n_strings - 1;
  c->hb.has_empty = has_empty;
  c->hash_ready = true;
  c->hbk_bgn = bgn_id; c->hbk_end = end_id; c->hbk_strings = P.max_hash_strings; c->hbk_datalen = P.max_hash_data_len;
  for (uint32_t s = 0; s < n_strings; s++)
    if (use[s]) c->stats.n_hash_bases += c->h_len[ri_bgn + s];
  c->stats.n_inserts += tot_inserts;
  c->stats.n_distinct += tot_distinct;
  c->stats.n_hash_blocks++;
  c->stats.last_end_hash_id = c->hb.last_id;
  if (last_id) *last_id = c->hb.last_id;
  return OIC_OK;
}

}  // namespace

// ================================================================================================ C ABI
extern "C" {

const char *oic_last_error(const oic_ctx_t *c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int oic_create(const oic_params_t *p, int device, oic_ctx_t **out) {
  if (!p || !out) { g_create_error = "null argument"; return OIC_ERR_ARG; }
  *out = nullptr;
  if (p->kmer_len < 2 || p->kmer_len > 31) { g_create_error = "k-mer length must be 2..31 in this build"; return OIC_ERR_UNSUPPORTED; }
  if (!(p->max_erate > 0.0) || p->max_erate >= 1.0) { g_create_error = "bad --maxerate"; return OIC_ERR_ARG; }
  if (p->max_hash_strings == 0 || p->num_threads == 0) { g_create_error = "bad --hashstrings / -t"; return OIC_ERR_ARG; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU path)";
    return OIC_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) { g_create_error = "bad device index"; return OIC_ERR_ARG; }
  if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return OIC_ERR_CUDA; }
  oic_ctx *c = new oic_ctx();
  c->P = *p;
  if (c->P.max_erate > 0.06) { c->P.use_hopeless_check = 0; c->P.use_window_filter = 0; }   // OIC/overlapInCore.C:416-421
  c->device = device;
  {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0) c->sm_count = sms;
  }
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev[0])) != cudaSuccess || (e = cudaEventCreate(&c->ev[1])) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev_call[0])) != cudaSuccess || (e = cudaEventCreate(&c->ev_call[1])) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev_k[0])) != cudaSuccess || (e = cudaEventCreate(&c->ev_k[1])) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev_k[2])) != cudaSuccess || (e = cudaEventCreate(&c->ev_k[3])) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev_k[4])) != cudaSuccess || (e = cudaEventCreate(&c->ev_k[5])) != cudaSuccess ||
      (e = c->d_ctr.ensure(sizeof(DevCounters))) != cudaSuccess) {
    g_create_error = cudaGetErrorString(e);
    delete c;
    return OIC_ERR_CUDA;
  }
  cudaMemset(c->d_ctr.p, 0, sizeof(DevCounters));
  const char *rc = getenv("OIC_REC_CAP"), *bb = getenv("OIC_BATCH_BASES");
  c->rec_cap = rc ? strtoull(rc, nullptr, 10) : (48ull << 20);
  c->batch_bases = bb ? strtoull(bb, nullptr, 10) : (512ull << 20);
  *out = c;
  return OIC_OK;
}

void oic_destroy(oic_ctx_t *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  DevBuf *bufs[] = {&c->d_fwd, &c->d_rc, &c->d_pos, &c->d_len, &c->d_match_limit, &c->d_skip_keys, &c->d_table, &c->d_occ,
                    &c->d_ascii, &c->d_asc_off, &c->d_str_flags, &c->d_str_use, &c->d_filter, &c->d_ovf_slot, &c->d_ovf_occ, &c->d_multi_list, &c->d_first_occ, &c->d_cnt, &c->d_mid_list, &c->d_blk, &c->d_big_list, &c->d_occ_tmp, &c->d_hist,
                    &c->d_qs_read, &c->d_qs_tile_base, &c->d_tile_empty, &c->d_recs, &c->d_chunk_next, &c->d_chunk_cnt,
                    &c->d_chunk_qs, &c->d_qs_first, &c->d_qs_nrec, &c->d_qs_base, &c->d_qs_screen, &c->d_bufA, &c->d_bufB,
                    &c->d_scratch, &c->d_pairs, &c->d_info, &c->d_out, &c->d_retry, &c->d_vals, &c->d_rowmeta, &c->d_deltas, &c->d_dirs, &c->d_order, &c->d_hist2, &c->d_pstate, &c->d_pend, &c->d_tasks, &c->d_results, &c->d_alive, &c->d_dpool, &c->d_wide,
                    &c->d_ctr, &c->d_flag, &c->d_kc_table, &c->d_kc_out_keys, &c->d_kc_out_cnt, &c->d_kc_n, &c->d_kc_hist, &c->d_kc_big, &c->d_xchg, &c->d_parts, &c->d_part_ctr, &c->d_q_order, &c->d_word_read, &c->d_strand, &c->d_quals};
  if (c->comm && nccl_api()) nccl_api()->CommDestroy(c->comm);
  for (DevBuf *b : bufs) b->release();
  c->overlaps.release();
  if (c->ev[0]) cudaEventDestroy(c->ev[0]);
  if (c->ev[1]) cudaEventDestroy(c->ev[1]);
  if (c->ev_call[0]) cudaEventDestroy(c->ev_call[0]);
  if (c->ev_call[1]) cudaEventDestroy(c->ev_call[1]);
  for (int i = 0; i < 6; i++) if (c->ev_k[i]) cudaEventDestroy(c->ev_k[i]);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

void oic_set_debug(oic_ctx_t *c, uint32_t mask) { c->debug = mask; }

/* Common part of the two loaders: read table, stream positions, device copies of them.  two_bit: bases are the store's 2-bit
 * encoding (4 per byte) instead of ASCII. */
static int load_reads_impl(oic_ctx_t *c, uint32_t first_id, uint32_t n, const uint64_t *offsets, const uint32_t *lengths,
                           const void *bases, const uint8_t *hash_ok, const uint8_t *ref_ok, bool two_bit) {
  if (!c || (n && (!offsets || !lengths || !bases)) || first_id < 1) return c ? c->fail(OIC_ERR_ARG, "oic_load_reads: bad argument") : OIC_ERR_ARG;
  cudaSetDevice(c->device);
  c->hash_ready = false;
  c->have_quals = false;
  c->first_id = first_id;
  c->n_reads = n;
  c->h_pos.resize(n + 1);
  c->h_len.assign(lengths, lengths + n);
  c->h_hash_ok.assign(n, 1);
  c->h_ref_ok.assign(n, 1);
  if (hash_ok) c->h_hash_ok.assign(hash_ok, hash_ok + n);
  if (ref_ok) c->h_ref_ok.assign(ref_ok, ref_ok + n);
  int64_t pos = READ_LEAD_BASES;
  uint64_t asc_lo = UINT64_MAX, asc_hi = 0;
  c->max_read_len = 0;
  for (uint32_t i = 0; i < n; i++) {
    if (lengths[i] > (uint32_t)MAX_READLEN) return c->fail(OIC_ERR_ARG, "read %u longer than AS_MAX_READLEN", first_id + i);
    c->h_pos[i] = pos;
    pos = (pos + lengths[i] + READ_PAD_BASES + 31) & ~(int64_t)31;
    const uint64_t nbytes = two_bit ? ((uint64_t)lengths[i] + 3) / 4 : lengths[i];
    asc_lo = std::min(asc_lo, offsets[i]);
    asc_hi = std::max(asc_hi, offsets[i] + nbytes);
    c->max_read_len = std::max(c->max_read_len, lengths[i]);
  }
  c->h_pos[n] = pos;
  c->stream_words = (uint64_t)(pos >> 5) + 4;
  if (n == 0) return OIC_OK;

  StageTimer t(c, &c->tm.ms_pack);
  CK(c->d_fwd.ensure(c->stream_words * 8));
  CK(c->d_rc.ensure(c->stream_words * 8));
  CK(c->d_pos.ensure((size_t)(n + 1) * 8));
  CK(c->d_len.ensure((size_t)(n + 1) * 4));
  DevBuf &d_ascii = c->d_ascii, &d_off = c->d_asc_off;
  const uint64_t asc_bytes = asc_hi - asc_lo;
  CK(d_ascii.ensure(asc_bytes + 64));
  CK(d_off.ensure((size_t)n * 8));
  std::vector<uint64_t> rel(n);
  for (uint32_t i = 0; i < n; i++) rel[i] = offsets[i] - asc_lo;
  CK(cudaMemcpyAsync(d_ascii.p, (const char *)bases + asc_lo, asc_bytes, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(d_off.p, rel.data(), (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->d_pos.p, c->h_pos.data(), (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->d_len.p, c->h_len.data(), (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += asc_bytes + (uint64_t)n * 20 + 8;
  if (zero_counters(c)) return OIC_ERR_CUDA;
  if (two_bit)
    k_pack_2bit<<<nblocks(c->stream_words, 256), 256, 0, c->stream>>>(d_ascii.as<uint8_t>(), d_off.as<uint64_t>(), c->d_len.as<uint32_t>(),
                                                                       c->d_pos.as<int64_t>(), n, c->stream_words, c->d_fwd.as<uint64_t>(),
                                                                       c->d_rc.as<uint64_t>());
  else
    k_pack_reads<<<nblocks(c->stream_words, 256), 256, 0, c->stream>>>(d_ascii.as<char>(), d_off.as<uint64_t>(), c->d_len.as<uint32_t>(),
                                                                       c->d_pos.as<int64_t>(), n, c->stream_words, c->d_fwd.as<uint64_t>(),
                                                                       c->d_rc.as<uint64_t>(), c->d_ctr.as<DevCounters>(), 0);
  c->tm.launches += 1;
  c->n_off = 0;
  DevCounters hc;
  int rcode = read_counters(c, hc);
  if (rcode) return rcode;
  CK(cudaGetLastError());
  if (hc.bad_base & 2) return c->fail(OIC_ERR_BASES, "a loaded read contains a letter other than ACGTN (a gkpStore never does: gatekeeper rewrites them to N)");
  if (hc.bad_base & 1) {
    // some read holds an N: pack again with the N-mask streams behind the base streams (same layout, 0b11 per N)
    const uint64_t n_off = c->stream_words + 4;
    CK(c->d_fwd.ensure((2 * n_off + 8) * 8));
    CK(c->d_rc.ensure((2 * n_off + 8) * 8));
    CK(cudaMemsetAsync(c->d_fwd.p, 0, (2 * n_off + 8) * 8, c->stream));
    CK(cudaMemsetAsync(c->d_rc.p, 0, (2 * n_off + 8) * 8, c->stream));
    k_pack_reads<<<nblocks(c->stream_words, 256), 256, 0, c->stream>>>(d_ascii.as<char>(), d_off.as<uint64_t>(), c->d_len.as<uint32_t>(),
                                                                       c->d_pos.as<int64_t>(), n, c->stream_words, c->d_fwd.as<uint64_t>(),
                                                                       c->d_rc.as<uint64_t>(), c->d_ctr.as<DevCounters>(), n_off);
    c->tm.launches += 1;
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    c->n_off = n_off;
  }
  CK(c->d_word_read.ensure(c->stream_words * 4));
  k_word_read<<<nblocks(c->stream_words, 256), 256, 0, c->stream>>>(c->d_pos.as<int64_t>(), n, c->stream_words, c->d_word_read.as<uint32_t>());
  c->tm.launches += 1;
  t.stop();

  // Edit_Match_Limit for every row an extension can reach: error limit <= ceil(erate * longest read)
  c->max_rows = (int)ceil(c->P.max_erate * (double)c->max_read_len) + 3;
  if (c->ml_erate != c->P.max_erate || c->ml_rows != c->max_rows) {
    std::vector<int32_t> ml;
    build_match_limit(c->P.max_erate, c->max_rows + 1, ml);
    CK(c->d_match_limit.ensure(ml.size() * 4));
    CK(cudaMemcpyAsync(c->d_match_limit.p, ml.data(), ml.size() * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->tm.h2d_bytes += ml.size() * 4;
    c->ml_erate = c->P.max_erate; c->ml_rows = c->max_rows;
  }
  return OIC_OK;
}

int oic_load_reads(oic_ctx_t *c, uint32_t first_id, uint32_t n, const uint64_t *offsets, const uint32_t *lengths,
                   const char *bases, const uint8_t *hash_ok, const uint8_t *ref_ok) {
  return load_reads_impl(c, first_id, n, offsets, lengths, bases, hash_ok, ref_ok, false);
}

int oic_load_reads_2bit(oic_ctx_t *c, uint32_t first_id, uint32_t n, const uint64_t *byte_offsets, const uint32_t *lengths,
                        const uint8_t *packed, const uint8_t *hash_ok, const uint8_t *ref_ok) {
  return load_reads_impl(c, first_id, n, byte_offsets, lengths, packed, hash_ok, ref_ok, true);
}

int oic_load_qualities(oic_ctx_t *c, const uint64_t *offsets, const uint8_t *quals) {
  if (!c || !offsets || !quals) return c ? c->fail(OIC_ERR_ARG, "oic_load_qualities: bad argument") : OIC_ERR_ARG;
  if (c->n_reads == 0) return c->fail(OIC_ERR_STATE, "oic_load_qualities before oic_load_reads");
  cudaSetDevice(c->device);
  // laid out like the base stream: quality of base i of read r at [h_pos[r] + i], zero elsewhere
  const size_t bytes = (size_t)c->stream_words * 32;
  std::vector<uint8_t> q(bytes, 0);
  for (uint32_t r = 0; r < c->n_reads; r++) memcpy(q.data() + c->h_pos[r], quals + offsets[r], c->h_len[r]);
  CK(c->d_quals.ensure(bytes));
  CK(cudaMemcpyAsync(c->d_quals.p, q.data(), bytes, cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  c->tm.h2d_bytes += bytes;
  c->have_quals = true;
  return OIC_OK;
}

int oic_set_skip_kmers(oic_ctx_t *c, uint64_t n, const char *kmers) {
  if (!c || (n && !kmers)) return OIC_ERR_ARG;
  const int K = c->P.kmer_len;
  c->hash_ready = false;
  c->skip_keys.clear();
  c->skip_keys.reserve(2 * n);
  auto code = [](char ch) -> int {
    switch (ch) { case 'c': case 'C': return 1; case 'g': case 'G': return 2; case 't': case 'T': return 3; default: return 0; }
  };
  auto rc_code = [](char ch) -> int {   // AS_UTL_reverseComplement table: non-ACGT -> NUL, which codes as 0
    switch (ch) { case 'a': case 'A': return 3; case 'c': case 'C': return 2; case 'g': case 'G': return 1; default: return 0; }
  };
  for (uint64_t i = 0; i < n; i++) {
    const char *s = kmers + i * K;
    uint64_t kf = 0, kr = 0;
    for (int j = 0; j < K; j++) {
      kf |= (uint64_t)code(s[j]) << (2 * j);
      kr |= (uint64_t)rc_code(s[K - 1 - j]) << (2 * j);
    }
    c->skip_keys.push_back(kf);
    c->skip_keys.push_back(kr);
  }
  return OIC_OK;
}

int oic_build_hash_index(oic_ctx_t *c, uint32_t bgn_id, uint32_t end_id, uint32_t *last_id) {
  return build_hash_index_impl(c, nullptr, bgn_id, end_id, last_id);
}

}  // extern "C"

namespace {

/* Root holds a built block: meta first (the receivers size their buffers from it), then the four buffers, all on the
 * context's stream.  meta.table_slots == 0 tells the receivers that the root failed (nobody hangs in the collective). */
int replicate_impl(oic_ctx *c, ncclComm_t comm, int n, int rank, int root) {
  NcclApi *api = nccl_api();
  if (n == 1) return c->hash_ready ? OIC_OK : c->fail(OIC_ERR_STATE, "oic_hash_block_replicate: no hash block on the root");
  StageTimer t(c, &c->tm.ms_build);
  CK(c->d_xchg.ensure(sizeof(ShardCounts) * (size_t)n + 256));
  oic_hash_meta_t meta;
  memset(&meta, 0, sizeof(meta));
  if (rank == root && c->hash_ready) meta = c->hb;
  if (rank == root) CK(cudaMemcpyAsync(c->d_xchg.p, &meta, sizeof(meta), cudaMemcpyHostToDevice, c->stream));
  NK(api->Broadcast(c->d_xchg.p, c->d_xchg.p, sizeof(meta), ncclChar, root, comm, c->stream));
  if (rank != root) {
    CK(cudaMemcpyAsync(&meta, c->d_xchg.p, sizeof(meta), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
  }
  if (meta.table_slots == 0) return c->fail(OIC_ERR_STATE, "oic_hash_block_replicate: the root rank has no hash block");
  if (rank != root) {
    const int rc = oic_hash_block_reserve(c, &meta);
    if (rc) return rc;   // (the peers would wait in the broadcasts below: a receiver that cannot allocate is fatal for the job)
  }
  const oic_hash_meta_t keep = c->hb;
  c->hb = meta;   // oic_hash_block_buffers sizes the buffers from the meta
  oic_dev_buffer_t b[OIC_HASH_BLOCK_NBUF];
  oic_hash_block_buffers(c, b);
  c->hb = keep;
  // (back to back, not as one NCCL group: measured on 2 B200, a group of these collectives runs ~5 ms slower than the same calls in sequence)
  for (int i = 0; i < OIC_HASH_BLOCK_NBUF; i++) NK(api->Broadcast(b[i].ptr, b[i].ptr, b[i].bytes, ncclChar, root, comm, c->stream));
  c->tm.launches += 1;
  if (rank != root) {
    const int rc = oic_hash_block_commit(c, &meta);
    if (rc) return rc;
  } else CK(cudaStreamSynchronize(c->stream));
  return OIC_OK;
}

int ensure_pair_scratch(oic_ctx *c, unsigned n_warps, unsigned long long rowcap) {
  CK(c->d_vals.ensure((size_t)n_warps * rowcap * 4));
  CK(c->d_rowmeta.ensure((size_t)n_warps * c->max_rows * sizeof(int4)));
  CK(c->d_deltas.ensure((size_t)n_warps * (4ull * c->max_rows + 8) * 4));
  return 0;
}

// The -l path of the pair stage (oic_extend.cuh k_process_strands): one warp per query strand walks its partners in the
// reference's order.  A partner cannot be deferred to a retry pass (the next one depends on it), so every warp gets worst-case
// edit-row scratch and the grid is sized by what that costs.
int run_pairs_limited(oic_ctx *c, ExtArgs &ea, const Seed *runs, uint32_t n_pairs, unsigned long long vals_bytes, DevCounters &hc) {
  const unsigned sms = (unsigned)c->sm_count;
  const unsigned long long big = ((unsigned long long)c->max_rows + 2) * ((unsigned long long)c->max_rows + 2) * 17 / 16 + 8ull * c->max_rows + 1024;
  const unsigned grid2 = (unsigned)std::max<unsigned long long>(8, std::min<unsigned long long>(sms * 4, std::max<unsigned long long>(24ull << 30, vals_bytes) / 4 / big / EXT_WARPS));
  if (ensure_pair_scratch(c, grid2 * EXT_WARPS, big)) return OIC_ERR_CUDA;
  ea.vals = c->d_vals.as<int32_t>(); ea.rowcap = big; ea.rowmeta = c->d_rowmeta.as<int4>(); ea.deltas = c->d_deltas.as<int32_t>();
  ea.task_list = nullptr; ea.n_tasks = 0; ea.handoff = 0;
  const size_t np = (size_t)n_pairs + 8;
  CK(c->d_strand.ensure(np * 40));   // per pair: creation key + mean diagonal (8 bytes each), six 4-byte arrays
  unsigned char *b = c->d_strand.as<unsigned char>();
  StrandArgs st;
  st.src = runs;
  st.ckey = reinterpret_cast<unsigned long long *>(b);
  st.mean = reinterpret_cast<double *>(b + np * 8);
  uint32_t *w = reinterpret_cast<uint32_t *>(b + np * 16);
  st.slot = w; st.rank = w + np; st.by_creation = w + 2 * np; st.order = w + 3 * np; st.strand_first = w + 4 * np; st.strand_n = w + 5 * np;
  st.limit = c->P.frag_olap_limit;
  k_strand_ranges<<<nblocks(n_pairs, 256), 256, 0, c->stream>>>(ea.pairs, n_pairs, st, ea.ctr);
  k_process_strands<<<grid2, EXT_WARPS * 32, 0, c->stream>>>(ea, st);
  c->tm.launches += 2;
  if (read_counters(c, hc)) return OIC_ERR_CUDA;
  CK(cudaGetLastError());
  if (hc.n_retry) return c->fail(OIC_ERR_CAPACITY, "edit rows exceed worst-case scratch in the -l path (internal)");
  return 0;
}

// returns 0 ok, 1 = run buffer overflow (caller retries with a smaller batch), <0 error
int run_batch(oic_ctx *c, const std::vector<uint32_t> &qs_read /* read index or ~0 per query */) {
  const oic_params_t &P = c->P;
  const uint32_t nq = (uint32_t)qs_read.size();
  const uint32_t nqs = 2 * nq;
  if (nq == 0) return 0;
  const uint32_t max_chunks = (uint32_t)(c->rec_cap / CHUNK);

  CK(c->d_qs_read.ensure((size_t)nq * 4));
  CK(cudaMemcpyAsync(c->d_qs_read.p, qs_read.data(), (size_t)nq * 4, cudaMemcpyHostToDevice, c->stream));
  // longest-first order of the scan (skipped reads last)
  // (a stable counting sort on length / 32: linear in the reads -- a comparison sort cost 21 ms per call on the 667 k reads of a
  // short-read block, where every read has the same length anyway)
  std::vector<uint32_t> q_order(nq);
  {
    constexpr uint32_t QB = (uint32_t)(MAX_READLEN >> 5) + 2;
    std::vector<uint32_t> bin(QB + 1, 0);
    auto key = [&](uint32_t q) { return qs_read[q] == 0xffffffffu ? 0u : 1u + (c->h_len[qs_read[q]] >> 5); };
    for (uint32_t q = 0; q < nq; q++) bin[QB - key(q)]++;          // descending length = ascending QB - key
    uint32_t run = 0;
    for (uint32_t b = 0; b <= QB; b++) { const uint32_t n_b = bin[b]; bin[b] = run; run += n_b; }
    for (uint32_t q = 0; q < nq; q++) q_order[bin[QB - key(q)]++] = q;
  }
  CK(c->d_q_order.ensure((size_t)nq * 4));
  CK(cudaMemcpyAsync(c->d_q_order.p, q_order.data(), (size_t)nq * 4, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += (uint64_t)nq * 8;
  CK(c->d_recs.ensure((size_t)max_chunks * CHUNK * sizeof(Seed)));
  CK(c->d_chunk_next.ensure((size_t)max_chunks * 4));
  CK(c->d_chunk_cnt.ensure((size_t)max_chunks * 4));
  CK(c->d_chunk_qs.ensure((size_t)max_chunks * 4));
  CK(c->d_qs_first.ensure((size_t)nqs * 4));
  CK(c->d_qs_nrec.ensure((size_t)nqs * 4));
  CK(c->d_qs_base.ensure((size_t)nqs * 4 + 16));
  CK(c->d_qs_screen.ensure((size_t)nqs + 16));
  if (c->hb.has_empty) {
    std::vector<uint32_t> tb(nqs);
    uint64_t tot = 0;
    for (uint32_t q = 0; q < nq; q++) {
      uint32_t tiles = qs_read[q] == 0xffffffffu ? 0 : (c->h_len[qs_read[q]] + 31) / 32 + 1;
      tb[2 * q] = (uint32_t)tot; tot += tiles;
      tb[2 * q + 1] = (uint32_t)tot; tot += tiles;
    }
    CK(c->d_qs_tile_base.ensure((size_t)nqs * 4));
    CK(c->d_tile_empty.ensure((size_t)tot * 4 + 16));
    CK(cudaMemcpyAsync(c->d_qs_tile_base.p, tb.data(), (size_t)nqs * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemsetAsync(c->d_tile_empty.p, 0, (size_t)tot * 4 + 16, c->stream));
    c->tm.h2d_bytes += (uint64_t)nqs * 4;
  }
  if (zero_counters(c)) return OIC_ERR_CUDA;

  // ---- K2a seed scan
  ScanArgs sa;
  sa.fwd = c->d_fwd.as<uint64_t>(); sa.rc = c->d_rc.as<uint64_t>(); sa.n_off = c->n_off;
  sa.rd_pos = c->d_pos.as<int64_t>(); sa.rd_len = c->d_len.as<uint32_t>();
  sa.qs_read = c->d_qs_read.as<uint32_t>(); sa.q_order = c->d_q_order.as<uint32_t>(); sa.n_queries = nq;
  sa.first_id = c->first_id; sa.hash_bgn_id = c->hb.bgn_id; sa.K = P.kmer_len;
  sa.table = c->d_table.as<Slot>(); sa.geom = c->geom; sa.occ = c->d_occ.as<uint64_t>();
  sa.filter = (c->have_filter && use_filter()) ? c->d_filter.as<uint32_t>() : nullptr;
  sa.has_empty = (int)c->hb.has_empty;
  sa.recs = c->d_recs.as<Seed>(); sa.max_chunks = max_chunks;
  sa.chunk_next = c->d_chunk_next.as<uint32_t>(); sa.chunk_cnt = c->d_chunk_cnt.as<uint32_t>(); sa.chunk_qs = c->d_chunk_qs.as<uint32_t>();
  sa.qs_first_chunk = c->d_qs_first.as<uint32_t>(); sa.qs_nrec = c->d_qs_nrec.as<uint32_t>();
  sa.qs_screen = c->d_qs_screen.as<uint8_t>();
  sa.qs_tile_base = c->d_qs_tile_base.as<uint32_t>(); sa.tile_empty = c->d_tile_empty.as<uint32_t>();
  sa.ctr = c->d_ctr.as<DevCounters>();
  DevCounters hc;
  {
    StageTimer t(c, &c->tm.ms_scan);
    k_seed_scan<<<nblocks(nqs, SCAN_WARPS), SCAN_WARPS * 32, 0, c->stream>>>(sa);
    c->tm.launches += 1;
    if (read_counters(c, hc)) return OIC_ERR_CUDA;
    CK(cudaGetLastError());
    t.stop();
  }
  if (hc.rec_overflow) return 1;
  const uint32_t n_chunks = hc.rec_chunks;
  const uint64_t n_runs = hc.n_runs;

  uint32_t n_pairs = 0;
  if (n_runs > 0) {
    StageTimer t(c, &c->tm.ms_group);
    // ---- K2b run lengths
    RunLenArgs ra;
    ra.fwd = sa.fwd; ra.rc = sa.rc; ra.n_off = sa.n_off; ra.rd_pos = sa.rd_pos; ra.rd_len = sa.rd_len; ra.qs_read = sa.qs_read;
    ra.hash_ri_bgn = c->hb.bgn_id - c->first_id; ra.K = P.kmer_len; ra.recs = sa.recs; ra.n_chunks = n_chunks;
    ra.chunk_cnt = sa.chunk_cnt; ra.chunk_qs = sa.chunk_qs; ra.has_empty = sa.has_empty;
    ra.qs_tile_base = sa.qs_tile_base; ra.tile_empty = sa.tile_empty;
    k_run_length<<<nblocks((uint64_t)n_chunks * CHUNK, 256), 256, 0, c->stream>>>(ra);
    // ---- exclusive scan of the per-strand run counts
    const uint64_t nb = (nqs + SCAN_ITEMS - 1) / SCAN_ITEMS;
    CK(c->d_blk.ensure((nb + 2) * 4));
    k_scan_u32_reduce<<<(unsigned)nb, SCAN_THREADS, 0, c->stream>>>(c->d_qs_nrec.as<uint32_t>(), nqs, c->d_blk.as<uint32_t>());
    k_scan_u32_single<<<1, SCAN_THREADS, 0, c->stream>>>(c->d_blk.as<uint32_t>(), (uint32_t)nb);
    k_scan_u32_apply<<<(unsigned)nb, SCAN_THREADS, 0, c->stream>>>(c->d_qs_nrec.as<uint32_t>(), c->d_qs_base.as<uint32_t>(), nqs, c->d_blk.as<uint32_t>());
    // ---- K2c group by hash read
    CK(c->d_bufA.ensure((size_t)n_runs * sizeof(Seed)));
    CK(c->d_bufB.ensure((size_t)n_runs * sizeof(Seed)));
    CK(c->d_scratch.ensure((size_t)n_runs * 16));
    CK(c->d_pairs.ensure((size_t)n_runs * sizeof(Pair)));
    CK(c->d_info.ensure((size_t)n_runs * sizeof(PairInfo)));
    int bits = 1;
    while ((1ull << bits) < c->hb.n_strings) bits++;
    GroupArgs ga;
    ga.recs = sa.recs; ga.chunk_next = sa.chunk_next; ga.chunk_cnt = sa.chunk_cnt;
    ga.qs_first_chunk = sa.qs_first_chunk; ga.qs_nrec = sa.qs_nrec; ga.qs_base = c->d_qs_base.as<uint32_t>(); ga.n_qs = nqs;
    ga.bufA = c->d_bufA.as<Seed>(); ga.bufB = c->d_bufB.as<Seed>(); ga.passes = (bits + 7) / 8;
    ga.pairs = c->d_pairs.as<Pair>(); ga.ctr = c->d_ctr.as<DevCounters>();
    k_group_runs<<<nblocks(nqs, GROUP_WARPS), GROUP_WARPS * 32, 0, c->stream>>>(ga);
    c->tm.launches += 5;
    if (read_counters(c, hc)) return OIC_ERR_CUDA;
    n_pairs = (uint32_t)hc.n_pairs;
    // ---- K2d Add_Match replay
    ReplayArgs rp;
    rp.src = (ga.passes & 1) ? ga.bufB : ga.bufA;
    rp.dst = (ga.passes & 1) ? ga.bufA : ga.bufB;
    rp.scratch = c->d_scratch.as<int32_t>(); rp.pairs = ga.pairs; rp.info = c->d_info.as<PairInfo>(); rp.n_pairs = n_pairs;
    rp.qs_base = ga.qs_base; rp.qs_nrec = ga.qs_nrec; rp.K = P.kmer_len;
    if (n_pairs) k_replay<<<nblocks(n_pairs, 128), 128, 0, c->stream>>>(rp);
    c->tm.launches += 1;
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    t.stop();

    if ((c->debug & 1) && n_pairs) {
      std::vector<Pair> hp(n_pairs);
      std::vector<PairInfo> hi(n_pairs);
      std::vector<Seed> hs(n_runs);
      CK(cudaMemcpy(hp.data(), c->d_pairs.p, (size_t)n_pairs * sizeof(Pair), cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hi.data(), c->d_info.p, (size_t)n_pairs * sizeof(PairInfo), cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hs.data(), rp.dst, (size_t)n_runs * sizeof(Seed), cudaMemcpyDeviceToHost));
      for (uint32_t i = 0; i < n_pairs; i++) {
        int64_t rec[9] = {(int64_t)c->first_id + qs_read[hp[i].qs >> 1], (int64_t)c->hb.bgn_id + hp[i].str, (int64_t)(hp[i].qs & 1),
                          hi[i].consistent, hi[i].diag_ct, hi[i].diag_bgn, hi[i].diag_end, (int64_t)hp[i].n,
                          (int64_t)(c->dbg_matches.size() / 3)};
        c->dbg_pairs.insert(c->dbg_pairs.end(), rec, rec + 9);
        for (uint32_t j = 0; j < hp[i].n; j++) {
          const Seed &s = hs[hp[i].first + j];
          c->dbg_matches.push_back(s.offset); c->dbg_matches.push_back(s.len); c->dbg_matches.push_back(s.start);
        }
      }
    }

    // ---- K3 pairs
    if (n_pairs) {
      StageTimer t3(c, &c->tm.ms_extend);
      const bool phase_trace = getenv("OIC_PHASE_CLOCKS") != nullptr;
      const auto h0 = std::chrono::steady_clock::now();
      auto h1 = h0, h2 = h0;
      // lean task kernel: the narrow-staging instance (more resident warps) unless the error rate makes wide bands the rule
      const bool narrow = !(P.max_erate > 0.06) && getenv("OIC_LEAN_WIDE") == nullptr;
      const unsigned sms = (unsigned)c->sm_count;
      const unsigned grid = sms * 4, grid_tasks = sms * (narrow ? TASK_BLOCKS_NARROW : TASK_BLOCKS_WIDE);
      const unsigned n_warps = grid_tasks * EXT_WARPS;
      // edit-row history per warp: the worst case (every row 2e+1 wide) if it fits the budget, else what fits
      // (pairs that overflow are rerun below with worst-case scratch on fewer warps)
      const unsigned long long worst = ((unsigned long long)c->max_rows + 2) * ((unsigned long long)c->max_rows + 2) * 17 / 16 + 8ull * c->max_rows + 1024;
      // memory budgets follow the device: a B200 has 180 GB, the c3 job needs 8 GB of history, the wide-band c2 job 51 GB.
      // mem_avail = what is free now + what the two big scratch buffers already hold; it is asked again whenever this library
      // allocated or freed anything since (g_alloc_epoch) and after a failed allocation -- not per call: cudaMemGetInfo enters
      // the kernel-mode driver (a millisecond alone, tens of milliseconds at random next to busy GPUs).
      if (c->mem_avail == 0 || c->mem_avail_epoch != g_alloc_epoch.load(std::memory_order_relaxed) || getenv("OIC_MEMINFO_ALWAYS")) {
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        c->mem_avail = (unsigned long long)free_b + c->d_vals.cap + c->d_dpool.cap;
        c->mem_avail_epoch = g_alloc_epoch.load(std::memory_order_relaxed);
      }
      unsigned long long avail = c->mem_avail;
      unsigned long long vals_bytes = std::min<unsigned long long>(96ull << 30, std::min<unsigned long long>(std::max<unsigned long long>(12ull << 30, avail * 45 / 100), avail * 60 / 100));
      // test knobs: shrink the first-pass history / the delta pool so that the overflow -> serial retry path runs on small inputs
      if (const char *e = getenv("OIC_VALS_BUDGET_MB")) vals_bytes = strtoull(e, nullptr, 10) << 20;
      unsigned long long budget = vals_bytes / 4 / n_warps;
      unsigned long long rowcap = std::min(worst, std::max(budget, (unsigned long long)c->max_rows * 64 + 1024));
      // an allocation that fails (another context, a K-mer count table, a larger block took the memory since the plan): halve the
      // history and try again -- pairs whose rows no longer fit take the retry path below -- instead of failing the call
      for (int attempt = 0; ensure_pair_scratch(c, n_warps, rowcap) != 0; attempt++) {
        cudaGetLastError();
        if (attempt >= 6 || rowcap <= (unsigned long long)c->max_rows * 64 + 1024) return OIC_ERR_CUDA;
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        c->mem_avail = avail = (unsigned long long)free_b + c->d_vals.cap + c->d_dpool.cap;
        vals_bytes = std::min(vals_bytes / 2, avail * 45 / 100);
        budget = vals_bytes / 4 / n_warps;
        rowcap = std::min(worst, std::max(budget, (unsigned long long)c->max_rows * 64 + 1024));
      }
      // longest-processing-time-first order
      CK(c->d_order.ensure((size_t)n_pairs * 4 + 4));
      CK(c->d_hist2.ensure(ORDER_BINS * 4));
      CK(cudaMemsetAsync(c->d_hist2.p, 0, ORDER_BINS * 4, c->stream));
      k_pair_hist<<<nblocks(n_pairs, 256), 256, 0, c->stream>>>(c->d_pairs.as<Pair>(), n_pairs, c->d_hist2.as<uint32_t>());
      k_pair_bins_scan<<<1, ORDER_BINS, 0, c->stream>>>(c->d_hist2.as<uint32_t>());
      k_pair_scatter<<<nblocks(n_pairs, 256), 256, 0, c->stream>>>(c->d_pairs.as<Pair>(), n_pairs, c->d_hist2.as<uint32_t>(), c->d_order.as<uint32_t>());
      c->tm.launches += 3;
      const uint32_t out_cap = n_pairs * (P.unique_olap ? 1u : (uint32_t)MAX_DISTINCT_OLAPS);
      CK(c->d_out.ensure((size_t)out_cap * 24 + 24));
      CK(c->d_retry.ensure((size_t)n_pairs * 4 + 4));
      ExtArgs ea;
      ea.fwd = sa.fwd; ea.rc = sa.rc; ea.n_off = sa.n_off; ea.rd_pos = sa.rd_pos; ea.rd_len = sa.rd_len; ea.qs_read = sa.qs_read;
      ea.first_id = c->first_id; ea.hash_ri_bgn = c->hb.bgn_id - c->first_id; ea.hash_bgn_id = c->hb.bgn_id;
      ea.str_flags = c->d_str_flags.as<uint8_t>(); ea.qs_screen = c->d_qs_screen.as<uint8_t>();
      ea.seeds = rp.dst; ea.pairs = ga.pairs; ea.info = rp.info; ea.n_pairs = n_pairs;
      ea.task_list = c->d_order.as<uint32_t>(); ea.n_tasks = n_pairs;
      ea.K = P.kmer_len; ea.partial = P.partial; ea.unique_olap = P.unique_olap; ea.use_hopeless = P.use_hopeless_check;
      ea.min_olap_len = P.min_olap_len; ea.filter_by_kmer_count = P.filter_by_kmer_count;
      ea.erate = P.max_erate; ea.minkmer_factor = exp(-1.0 * (double)P.kmer_len * P.max_erate);
      ea.branch_match_value = P.max_erate / (1 + P.max_erate);               // LIBO/prefixEditDistance.C:95
      ea.min_branch_tail_slope = (P.max_erate > 0.06) ? 1.0 : 0.20;          // LIBO/prefixEditDistance.C:41
      ea.match_limit = c->d_match_limit.as<int32_t>(); ea.max_rows = c->max_rows;
      ea.vals = c->d_vals.as<int32_t>(); ea.rowcap = rowcap; ea.rowmeta = c->d_rowmeta.as<int4>(); ea.deltas = c->d_deltas.as<int32_t>();
      ea.out = c->d_out.as<uint32_t>(); ea.out_cap = out_cap; ea.retry_list = c->d_retry.as<uint32_t>();
      ea.ctr = sa.ctr;
      ea.quals = P.use_window_filter ? c->d_quals.as<uint8_t>() : nullptr;
      ea.first_lean = getenv("OIC_NO_FIRST_LEAN") ? 0 : 1;
      if (P.frag_olap_limit != UINT64_MAX) {   // -l: partners of a strand depend on each other -> serial walk per strand
        const int rl = run_pairs_limited(c, ea, rp.src, n_pairs, vals_bytes, hc);
        if (rl) return rl;
      } else {
      // speculative second round (oic_extend.cuh): state per pair, one result slot per seed, a pool for the delta lists
      unsigned long long dpool_cap = std::min<unsigned long long>(std::max<unsigned long long>(1ull << 28, avail * 15 / 100 / 4),
                                                                  (unsigned long long)n_runs * 512 + (16ull << 20));
      if (const char *e = getenv("OIC_DPOOL_MB")) dpool_cap = (strtoull(e, nullptr, 10) << 20) / 4;
      CK(c->d_pstate.ensure((size_t)n_pairs * sizeof(PairState)));
      CK(c->d_pend.ensure((size_t)n_pairs * 4 + 4));
      CK(c->d_tasks.ensure((size_t)n_runs * sizeof(uint2) + 8));
      CK(c->d_results.ensure((size_t)n_runs * sizeof(TaskResult)));
      CK(c->d_alive.ensure((size_t)n_runs + 4));
      CK(c->d_wide.ensure((size_t)n_runs * 4 + 4));
      while (c->d_dpool.ensure((size_t)dpool_cap * 4) != cudaSuccess) {   // a pool that is too small only sends pairs to the retry path
        cudaGetLastError();
        if (dpool_cap <= (1ull << 24)) return c->fail(OIC_ERR_CUDA, "no device memory for the delta pool");
        dpool_cap /= 2;
      }
      // (pairs with an N in either read are never handed over: the lean loop of pass 2 knows no wildcards)
      // (-w walks the delta list of every overlap it outputs, which the speculative path does not keep: serial path)
      ea.handoff = (getenv("OIC_NO_SPECULATION") || ea.quals) ? 0 : 1;
      ea.pstate = c->d_pstate.as<PairState>(); ea.pend_list = c->d_pend.as<uint32_t>(); ea.tasks = c->d_tasks.as<uint2>();
      ea.results = c->d_results.as<TaskResult>(); ea.alive = c->d_alive.as<uint8_t>(); ea.wide_list = c->d_wide.as<uint32_t>();
      ea.dpool = c->d_dpool.as<int32_t>(); ea.dpool_cap = dpool_cap;
      h1 = std::chrono::steady_clock::now();
      cudaEventRecord(c->ev_k[0], c->stream);
      k_process_pairs<<<grid, EXT_WARPS * 32, 0, c->stream>>>(ea);
      cudaEventRecord(c->ev_k[1], c->stream);
      c->tm.launches += 1;
      if (ea.handoff) {
        if (narrow) k_extend_tasks<LSW_NARROW, TASK_BLOCKS_NARROW><<<grid_tasks, EXT_WARPS * 32, 0, c->stream>>>(ea);
        else k_extend_tasks<LSW_WIDE, TASK_BLOCKS_WIDE><<<grid_tasks, EXT_WARPS * 32, 0, c->stream>>>(ea);
        cudaEventRecord(c->ev_k[2], c->stream);
        k_extend_tasks<0, 4><<<grid, EXT_WARPS * 32, 0, c->stream>>>(ea);
        k_commit_pairs<<<sms * 4, 128, 0, c->stream>>>(ea);
        cudaEventRecord(c->ev_k[3], c->stream);
        c->tm.launches += 3;
      }
      if (read_counters(c, hc)) return OIC_ERR_CUDA;
      h2 = std::chrono::steady_clock::now();
      CK(cudaGetLastError());
      {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, c->ev_k[0], c->ev_k[1]) == cudaSuccess) c->tm.ms_ext_first += ms;
        if (ea.handoff) {
          if (cudaEventElapsedTime(&ms, c->ev_k[1], c->ev_k[2]) == cudaSuccess) c->tm.ms_ext_tasks += ms;
          if (cudaEventElapsedTime(&ms, c->ev_k[2], c->ev_k[3]) == cudaSuccess) c->tm.ms_ext_commit += ms;
        }
      }
      const DevCounters hc_first = hc;
      if (hc.n_retry) {
        // pairs whose rows did not fit the per-warp scratch: regenerate the seed lists (the pruning above edits them
        // in place; finished pairs are not read again) and rerun just those pairs with worst-case scratch
        const uint32_t nr = hc.n_retry;
        k_replay<<<nblocks(n_pairs, 128), 128, 0, c->stream>>>(rp);
        const unsigned long long big = ((unsigned long long)c->max_rows + 2) * ((unsigned long long)c->max_rows + 2) * 17 / 16 + 8ull * c->max_rows + 1024;
        const unsigned grid2 = (unsigned)std::max<unsigned long long>(8, std::min<unsigned long long>(sms * 4, std::max<unsigned long long>(24ull << 30, vals_bytes) / 4 / big / EXT_WARPS));
        cudaEventRecord(c->ev_k[4], c->stream);
        DevBuf keep_retry;
        CK(keep_retry.ensure((size_t)nr * 4));
        CK(cudaMemcpyAsync(keep_retry.p, c->d_retry.p, (size_t)nr * 4, cudaMemcpyDeviceToDevice, c->stream));
        if (ensure_pair_scratch(c, grid2 * EXT_WARPS, big)) { keep_retry.release(); return OIC_ERR_CUDA; }
        ea.vals = c->d_vals.as<int32_t>(); ea.rowcap = big; ea.rowmeta = c->d_rowmeta.as<int4>(); ea.deltas = c->d_deltas.as<int32_t>();
        ea.task_list = keep_retry.as<uint32_t>(); ea.n_tasks = nr; ea.handoff = 0;
        CK(cudaMemsetAsync(&c->d_ctr.as<DevCounters>()->pair_cursor, 0, 4, c->stream));
        CK(cudaMemsetAsync(&c->d_ctr.as<DevCounters>()->n_retry, 0, 4, c->stream));
        k_process_pairs<<<grid2, EXT_WARPS * 32, 0, c->stream>>>(ea);
        c->tm.launches += 2;
        cudaEventRecord(c->ev_k[5], c->stream);
        int r2 = read_counters(c, hc);
        keep_retry.release();
        { float ms = 0; if (cudaEventElapsedTime(&ms, c->ev_k[4], c->ev_k[5]) == cudaSuccess) c->tm.ms_ext_retry += ms; }
        if (r2) return OIC_ERR_CUDA;
        CK(cudaGetLastError());
        if (hc.n_retry) return c->fail(OIC_ERR_CAPACITY, "edit rows exceed worst-case scratch (internal)");
      }
      if (getenv("OIC_PHASE_CLOCKS"))
        fprintf(stderr, "[oic] pair stage: pairs %u | handed over %u | speculative tasks %u used %u | delta pool %llu of %llu | retried %u | wide tasks %u\n",
                n_pairs, hc_first.n_pend, hc_first.n_spec_tasks, hc_first.n_spec_used, (unsigned long long)hc_first.dpool_used, dpool_cap, hc_first.n_retry, hc_first.n_wide_tasks);
      }   // (default path)
      if (hc.out_overflow) return c->fail(OIC_ERR_CAPACITY, "overlap output buffer overflow (internal)");
      if (hc.n_out) {
        oic_ovb_record_t *at = c->overlaps.grow(hc.n_out);
        if (!at) return c->fail(OIC_ERR_CUDA, "no host memory for %llu overlap records", (unsigned long long)(c->overlaps.size() + hc.n_out));
        CK(cudaMemcpyAsync(at, c->d_out.p, (size_t)hc.n_out * 24, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        c->tm.d2h_bytes += (uint64_t)hc.n_out * 24;
      }
      t3.stop();
      if (phase_trace) {
        const auto h3 = std::chrono::steady_clock::now();
        auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[oic] pair stage host clock: plan + order %.2f ms | kernels until the counters are back %.2f ms | records to the host %.2f ms\n",
                ms(h0, h1), ms(h1, h2), ms(h2, h3));
      }
    }
  }

  oic_stats_t &s = c->stats;
  s.n_query_bases += hc.n_query_bases; s.n_windows += hc.n_windows; s.n_found += hc.n_found; s.n_refs += hc.n_refs;
  s.n_add_ref += hc.n_add_ref; s.n_runs += hc.n_runs; s.n_pairs += n_pairs;
  s.ext_calls += hc.ext_calls; s.ext_cells += hc.ext_cells; s.ext_words += hc.ext_words;
  s.kmer_hits_without_olap += hc.kmer_hits_without_olap; s.kmer_hits_with_olap += hc.kmer_hits_with_olap;
  s.kmer_hits_skipped += hc.kmer_hits_skipped; s.multi_overlap += hc.multi_overlap;
  s.total_overlaps += hc.total_overlaps; s.contained += hc.contained; s.dovetail += hc.dovetail;
  s.bad_short_window += hc.bad_short_window; s.bad_long_window += hc.bad_long_window;
  return 0;
}


// INT32 throughput probe (roofline denominator of the extension stage): every thread runs 8 independent chains of
// LOP3 + IADD3 (the ALU pipe that the edit-distance rows live on) or of IMAD (the FMA pipe); one warp instruction = 32 ops.
template <int MODE>
__global__ void k_int_probe(uint32_t *out, int iters, uint32_t seed) {
  uint32_t x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = seed + threadIdx.x * 8u + i;
  const uint32_t y = seed * 2654435761u + blockIdx.x, z = seed ^ 0x9e3779b9u;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) x[i] = (x[i] ^ y) + z;        // LOP3 + IADD3
      else x[i] = x[i] * y + z;                    // IMAD
    }
  }
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) r ^= x[i];
  if (r == 0x12345678u) out[0] = r;                // keeps the chains alive
}

}  // namespace

extern "C" {

int oic_process_overlaps(oic_ctx_t *c, uint32_t bgn_ref_id, uint32_t end_ref_id) {
  if (!c) return OIC_ERR_ARG;
  cudaSetDevice(c->device);
  if (!c->hash_ready) return c->fail(OIC_ERR_STATE, "oic_process_overlaps before oic_build_hash_index");
  if (c->P.use_window_filter && !c->have_quals) return c->fail(OIC_ERR_STATE, "-w (window filter) needs the reads' quality values: call oic_load_qualities first");
  c->dbg_pairs.clear();
  c->dbg_matches.clear();
  if (end_ref_id < bgn_ref_id) return OIC_OK;
  if (bgn_ref_id < c->first_id || end_ref_id >= c->first_id + c->n_reads)
    return c->fail(OIC_ERR_ARG, "query range %u-%u outside the loaded reads", bgn_ref_id, end_ref_id);
  cudaEventRecord(c->ev_call[0], c->stream);
  const oic_params_t &P = c->P;
  uint32_t id = bgn_ref_id;
  uint64_t batch_bases = c->batch_bases;
  int rc = 0;
  while (id <= end_ref_id) {
    std::vector<uint32_t> qs;
    uint64_t bases = 0;
    uint32_t j = id;
    while (j <= end_ref_id && (qs.empty() || bases < batch_bases)) {
      const uint32_t ri = j - c->first_id;
      const bool act = c->h_ref_ok[ri] && (int64_t)c->h_len[ri] >= (int64_t)P.min_olap_len;   // Process_Overlaps.C:108-115
      qs.push_back(act ? ri : 0xffffffffu);
      if (act) bases += c->h_len[ri];
      j++;
    }
    rc = run_batch(c, qs);
    if (rc < 0) break;
    if (rc == 1) {   // run buffer too small for this batch
      if (qs.size() == 1) c->rec_cap *= 2; else batch_bases = std::max<uint64_t>(bases / 2, 1);
      rc = 0;
      continue;
    }
    id = j;
  }
  cudaEventRecord(c->ev_call[1], c->stream);
  float ms = 0;
  if (cudaEventSynchronize(c->ev_call[1]) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev_call[0], c->ev_call[1]) == cudaSuccess)
    c->tm.ms_total_search += ms;
  return rc;
}

uint32_t oic_last_processed_ref(uint32_t bgn_ref, uint32_t end_ref, uint32_t num_threads) {
  if (end_ref < bgn_ref || num_threads == 0) return 0;
  const uint32_t per_thread = 1 + (end_ref - bgn_ref) / num_threads / 8;   // OIC/overlapInCore.C:249
  // a block that would START at end_ref is dropped by `bgnID < endRefID` (Process_Overlaps.C:86)
  if ((end_ref - bgn_ref) % per_thread == 0) return end_ref - 1 >= bgn_ref ? end_ref - 1 : 0;
  return end_ref;
}

int oic_overlap_driver(oic_ctx_t *c) {
  if (!c) return OIC_ERR_ARG;
  const oic_params_t &P = c->P;
  if (c->n_reads == 0) return c->fail(OIC_ERR_STATE, "no reads loaded");
  const uint32_t num_reads = c->first_id + c->n_reads - 1;
  uint32_t g_bgn = P.bgn_hash_id < 1 ? 1 : P.bgn_hash_id;
  uint32_t g_end = P.end_hash_id > num_reads ? num_reads : P.end_hash_id;
  uint64_t bgn = g_bgn, end = (uint64_t)g_bgn + P.max_hash_strings - 1;
  while (bgn < g_end) {   // strict '<' (OIC/overlapInCore.C:222)
    if (end > g_end) end = g_end;
    uint32_t last = 0;
    int rc = oic_build_hash_index(c, (uint32_t)bgn, (uint32_t)end, &last);
    if (rc) return rc;
    uint32_t bgn_ref = P.bgn_ref_id < 1 ? 1 : P.bgn_ref_id;
    uint32_t end_ref = P.end_ref_id > num_reads ? num_reads : P.end_ref_id;
    uint32_t last_ref = oic_last_processed_ref(bgn_ref, end_ref, P.num_threads);
    if (last_ref >= bgn_ref) {
      rc = oic_process_overlaps(c, bgn_ref, last_ref);
      if (rc) return rc;
    }
    bgn = (uint64_t)last + 1;
    end = bgn + P.max_hash_strings - 1;
  }
  return OIC_OK;
}

int oic_set_job(oic_ctx_t *c, uint32_t bgn_hash_id, uint32_t end_hash_id, uint32_t bgn_ref_id, uint32_t end_ref_id,
                uint32_t max_hash_strings, uint64_t max_hash_data_len) {
  if (!c) return OIC_ERR_ARG;
  if (max_hash_strings == 0) return c->fail(OIC_ERR_ARG, "bad --hashstrings");
  c->P.bgn_hash_id = bgn_hash_id; c->P.end_hash_id = end_hash_id;
  c->P.bgn_ref_id = bgn_ref_id; c->P.end_ref_id = end_ref_id;
  c->P.max_hash_strings = max_hash_strings; c->P.max_hash_data_len = max_hash_data_len;
  c->job_mode = true;
  return OIC_OK;
}
void oic_reset_stats(oic_ctx_t *c) { if (c) memset(&c->stats, 0, sizeof(c->stats)); }

uint64_t oic_num_overlaps(const oic_ctx_t *c) { return c ? c->overlaps.size() : 0; }

int oic_copy_overlaps(oic_ctx_t *c, oic_ovb_record_t *dst, uint64_t capacity, uint64_t *n_copied) {
  if (!c || (!dst && capacity)) return OIC_ERR_ARG;
  uint64_t n = std::min<uint64_t>(capacity, c->overlaps.size());
  if (n) memcpy(dst, c->overlaps.data(), n * sizeof(oic_ovb_record_t));
  if (n_copied) *n_copied = n;
  return OIC_OK;
}
void oic_clear_overlaps(oic_ctx_t *c) { if (c) c->overlaps.clear(); }

int oic_get_stats(const oic_ctx_t *c, oic_stats_t *out) {
  if (!c || !out) return OIC_ERR_ARG;
  *out = c->stats;
  return OIC_OK;
}
int oic_get_timings(oic_ctx_t *c, oic_timings_t *out) {
  if (!c || !out) return OIC_ERR_ARG;
  *out = c->tm;
  return OIC_OK;
}
void oic_reset_timings(oic_ctx_t *c) { if (c) memset(&c->tm, 0, sizeof(c->tm)); }

// ---- multi-GPU replication of the hash block -----------------------------------------------------------
int oic_hash_block_meta(const oic_ctx_t *c, oic_hash_meta_t *m) {
  if (!c || !m || !c->hash_ready) return OIC_ERR_STATE;
  *m = c->hb;
  return OIC_OK;
}
int oic_hash_block_reserve(oic_ctx_t *c, const oic_hash_meta_t *m) {
  if (!c || !m) return OIC_ERR_ARG;
  cudaSetDevice(c->device);
  c->hash_ready = false;
  if (m->n_slices == 0 || m->table_slots != ((uint64_t)m->n_slices << m->slice_bits) || m->filter_words == 0 ||
      (m->filter_words & (m->filter_words - 1)) != 0 || m->filter_words < FILTER_REGIONS)
    return c->fail(OIC_ERR_ARG, "oic_hash_block_reserve: inconsistent table geometry in the meta");
  CK(c->d_table.ensure(m->table_slots * sizeof(Slot)));
  CK(c->d_occ.ensure(((size_t)m->n_occ + 32) * 8));
  CK(c->d_str_flags.ensure(((size_t)m->n_strings + 8) & ~(size_t)3));
  CK(c->d_filter.ensure(m->filter_words * 4));
  return OIC_OK;
}
int oic_hash_block_buffers(const oic_ctx_t *c, oic_dev_buffer_t b[OIC_HASH_BLOCK_NBUF]) {
  if (!c || !b) return OIC_ERR_ARG;
  b[0].ptr = c->d_table.p; b[0].bytes = c->hb.table_slots * sizeof(Slot);
  b[1].ptr = c->d_occ.p; b[1].bytes = ((uint64_t)c->hb.n_occ + 32) * 8;
  b[2].ptr = c->d_str_flags.p; b[2].bytes = ((uint64_t)c->hb.n_strings + 8) & ~(uint64_t)3;
  b[3].ptr = c->d_filter.p; b[3].bytes = c->hb.filter_words * 4;
  return OIC_OK;
}
int oic_hash_block_commit(oic_ctx_t *c, const oic_hash_meta_t *m) {
  if (!c || !m) return OIC_ERR_ARG;
  cudaSetDevice(c->device);
  c->hb = *m;
  c->geom.n_slices = m->n_slices;
  c->geom.slice_bits = m->slice_bits;
  c->geom.slice_mask = (1ULL << m->slice_bits) - 1;
  c->geom.region_bits = 0;
  while (((uint64_t)FILTER_REGIONS << c->geom.region_bits) < m->filter_words) c->geom.region_bits++;
  c->have_filter = true;   // the filter travels with the block
  CK(cudaStreamSynchronize(c->stream));
  c->hash_ready = true;
  c->hbk_bgn = c->hbk_end = 0;   // a block that arrived from another rank is never mistaken for a local build
  c->stats.n_hash_blocks++;
  c->stats.last_end_hash_id = m->last_id;
  return OIC_OK;
}

// ---- the collective inside the library -------------------------------------------------------------------
int oic_comm_unique_id(uint8_t id[OIC_COMM_ID_BYTES]) {
  static_assert(OIC_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "ncclUniqueId size");
  NcclApi *api = nccl_api();
  if (!api || !id) { g_create_error = "NCCL unavailable (libnccl.so.2 could not be loaded)"; return OIC_ERR_CUDA; }
  ncclUniqueId u;
  const ncclResult_t r = api->GetUniqueId(&u);
  if (r != ncclSuccess) { g_create_error = api->GetErrorString(r); return OIC_ERR_CUDA; }
  memcpy(id, u.internal, OIC_COMM_ID_BYTES);
  return OIC_OK;
}
int oic_comm_init_rank(oic_ctx_t *c, const uint8_t id[OIC_COMM_ID_BYTES], int n_ranks, int rank) {
  if (!c || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return c ? c->fail(OIC_ERR_ARG, "oic_comm_init_rank: bad argument") : OIC_ERR_ARG;
  NcclApi *api = nccl_api();
  if (!api) return c->fail(OIC_ERR_CUDA, "NCCL unavailable (libnccl.so.2 could not be loaded)");
  if (c->comm) return c->fail(OIC_ERR_STATE, "the context already has a communicator");
  cudaSetDevice(c->device);
  ncclUniqueId u;
  memcpy(u.internal, id, OIC_COMM_ID_BYTES);
  NK(api->CommInitRank(&c->comm, n_ranks, u, rank));
  c->comm_n = n_ranks; c->comm_rank = rank;
  return OIC_OK;
}
int oic_comm_init_all(oic_ctx_t *const *ctxs, int n) {
  if (!ctxs || n < 1) return OIC_ERR_ARG;
  NcclApi *api = nccl_api();
  oic_ctx *c = ctxs[0];
  if (!api) return c->fail(OIC_ERR_CUDA, "NCCL unavailable (libnccl.so.2 could not be loaded)");
  std::vector<int> devs(n);
  std::vector<ncclComm_t> comms(n);
  for (int i = 0; i < n; i++) {
    if (!ctxs[i] || ctxs[i]->comm) return c->fail(OIC_ERR_STATE, "oic_comm_init_all: null context or communicator already attached");
    devs[i] = ctxs[i]->device;
  }
  NK(api->CommInitAll(comms.data(), n, devs.data()));
  for (int i = 0; i < n; i++) { ctxs[i]->comm = comms[i]; ctxs[i]->comm_n = n; ctxs[i]->comm_rank = i; }
  return OIC_OK;
}
int oic_comm_info(const oic_ctx_t *c, int *n_ranks, int *rank) {
  if (!c) return OIC_ERR_ARG;
  if (n_ranks) *n_ranks = c->comm ? c->comm_n : 0;
  if (rank) *rank = c->comm_rank;
  return OIC_OK;
}
int oic_hash_block_replicate(oic_ctx_t *c, void *nccl_comm, int root) {
  if (!c) return OIC_ERR_ARG;
  cudaSetDevice(c->device);
  ncclComm_t comm;
  int n = 1, rank = 0;
  int rc = resolve_comm(c, nccl_comm, comm, n, rank);
  if (rc) return rc;
  if (root < 0 || root >= n) return c->fail(OIC_ERR_ARG, "oic_hash_block_replicate: bad root");
  return replicate_impl(c, comm, n, rank, root);
}
int oic_build_hash_index_sharded(oic_ctx_t *c, void *nccl_comm, uint32_t bgn_id, uint32_t end_id, uint32_t *last_id) {
  return build_hash_index_impl(c, nccl_comm, bgn_id, end_id, last_id, true);
}

// ---- stage dumps ---------------------------------------------------------------------------------------
int oic_dump_index(oic_ctx_t *c, uint64_t *n_keys, uint64_t *n_occ, uint64_t *keys, uint64_t *occ_start, uint64_t *occ, uint8_t *flags) {
  if (!c || !c->hash_ready) return OIC_ERR_STATE;
  cudaSetDevice(c->device);
  std::vector<Slot> t(c->hb.table_slots);
  std::vector<uint64_t> o(c->hb.n_occ);
  CK(cudaMemcpy(t.data(), c->d_table.p, t.size() * sizeof(Slot), cudaMemcpyDeviceToHost));
  if (!o.empty()) CK(cudaMemcpy(o.data(), c->d_occ.p, o.size() * 8, cudaMemcpyDeviceToHost));
  std::vector<uint32_t> idx;
  for (uint64_t i = 0; i < t.size(); i++)
    if (t[i].key != EMPTY_KEY) idx.push_back((uint32_t)i);
  std::sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return t[a].key < t[b].key; });
  if (n_keys) *n_keys = idx.size();
  uint64_t total = 0;
  for (size_t k = 0; k < idx.size(); k++) total += pay_len(t[idx[k]].pay);
  if (n_occ) *n_occ = total;
  if (!keys) return OIC_OK;
  uint64_t w = 0;
  for (size_t k = 0; k < idx.size(); k++) {
    const Slot &s = t[idx[k]];
    keys[k] = s.key;
    occ_start[k] = w;
    flags[k] = pay_empty(s.pay) ? 1 : 0;
    const uint32_t n = pay_len(s.pay);
    for (uint32_t j = 0; j < n; j++) {
      const uint64_t e = pay_is_list(s.pay) ? o[pay_start(s.pay) + j] : s.pay;
      occ[w++] = ((uint64_t)occ_str(e) << 32) | occ_off(e);
    }
  }
  occ_start[idx.size()] = w;
  return OIC_OK;
}

int oic_dump_pairs(oic_ctx_t *c, uint64_t *n_pairs, uint64_t *n_matches, int64_t *pairs, int32_t *matches) {
  if (!c) return OIC_ERR_ARG;
  if (n_pairs) *n_pairs = c->dbg_pairs.size() / 9;
  if (n_matches) *n_matches = c->dbg_matches.size() / 3;
  if (pairs && !c->dbg_pairs.empty()) memcpy(pairs, c->dbg_pairs.data(), c->dbg_pairs.size() * 8);
  if (matches && !c->dbg_matches.empty()) memcpy(matches, c->dbg_matches.data(), c->dbg_matches.size() * 4);
  return OIC_OK;
}


static int count_kmers_impl(oic_ctx_t *c, uint32_t K, uint32_t min_count, uint64_t *n_found, uint32_t *n_passes, double *device_ms);

/* Canonical K-mer counts of the loaded reads (oic_count.cuh).  The count table (up to 60 % of the free memory) and its output
 * buffers are released before returning -- the overlap stage plans its own scratch against what is free -- and a failed
 * count leaves no partial result behind. */
int oic_count_kmers(oic_ctx_t *c, uint32_t K, uint32_t min_count, uint64_t *n_found, uint32_t *n_passes, double *device_ms) {
  const int rc = count_kmers_impl(c, K, min_count, n_found, n_passes, device_ms);
  if (c) {
    cudaSetDevice(c->device);
    c->d_kc_table.release(); c->d_kc_out_keys.release(); c->d_kc_out_cnt.release(); c->d_kc_big.release(); c->d_kc_hist.release();
      if (rc != OIC_OK) { c->freq.clear(); c->kc_hist.clear(); if (n_found) *n_found = 0; }
  }
  return rc;
}

static int count_kmers_impl(oic_ctx_t *c, uint32_t K, uint32_t min_count, uint64_t *n_found, uint32_t *n_passes, double *device_ms) {
  if (!c) return OIC_ERR_ARG;
  if (K < 2 || K > 31) return c->fail(OIC_ERR_UNSUPPORTED, "oic_count_kmers: K must be 2..31");
  if (c->n_reads == 0) return c->fail(OIC_ERR_STATE, "oic_count_kmers before oic_load_reads");
  cudaSetDevice(c->device);
  if (min_count < 1) min_count = 1;
  uint64_t windows = 0;
  for (uint32_t i = 0; i < c->n_reads; i++) if (c->h_len[i] >= K) windows += c->h_len[i] - K + 1;
  c->freq.clear();
  if (n_found) *n_found = 0;
  if (n_passes) *n_passes = 0;
  if (device_ms) *device_ms = 0.0;
  if (windows == 0) return OIC_OK;
  // table of 16-byte {key, count} slots, at most half full even if every window is a different K-mer; read sets whose
  // table does not fit the budget are counted in n_pass passes, each over one hash partition of the K-mers
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  unsigned long long budget = (unsigned long long)(free_b + c->d_kc_table.cap) * 6 / 10;
  uint64_t floor_slots = 1ull << 20;
  if (const char *e = getenv("OIC_COUNT_BUDGET_MB")) { budget = strtoull(e, nullptr, 10) << 20; floor_slots = 4096; }   // test knob: forces several passes
  uint32_t n_pass = 1;
  uint64_t slots = 0;
  for (;; n_pass++) {
    slots = 1024;
    while (slots < 2 * ((windows + n_pass - 1) / n_pass) + 2 * (windows / n_pass / 16)) slots <<= 1;   // + slack for uneven partitions
    if (slots * 16ull <= budget || slots <= floor_slots) break;
    if (n_pass > 4096) return c->fail(OIC_ERR_CAPACITY, "oic_count_kmers: no memory for the count table");
  }
  CK(c->d_kc_table.ensure(slots * sizeof(Slot)));
  CK(c->d_kc_n.ensure(16));
  const uint32_t HB = 1u << 20, BIG_CAP = 1u << 20;   // histogram bins on the device; rarer, larger counts are listed one by one
  CK(c->d_kc_hist.ensure((size_t)HB * 8));
  CK(c->d_kc_big.ensure((size_t)BIG_CAP * 4));
  CK(cudaMemsetAsync(c->d_kc_hist.p, 0, (size_t)HB * 8, c->stream));
  CK(cudaMemsetAsync(c->d_kc_n.p, 0, 16, c->stream));
  c->kc_hist.clear();
  std::vector<uint32_t> big_counts;
  cudaEventRecord(c->ev_call[0], c->stream);
  CountArgs a;
  a.fwd = c->d_fwd.as<uint64_t>(); a.n_off = c->n_off; a.rd_pos = c->d_pos.as<int64_t>(); a.rd_len = c->d_len.as<uint32_t>();
  a.n_reads = c->n_reads; a.n_words = c->stream_words - 4; a.K = (int)K;
  a.table = c->d_kc_table.as<Slot>(); a.slot_mask = slots - 1; a.n_pass = n_pass;
  int rc = OIC_OK;
  for (uint32_t pass = 0; pass < n_pass && rc == OIC_OK; pass++) {
    a.pass = pass;
    k_table_init<<<nblocks(slots, 256), 256, 0, c->stream>>>(a.table, slots);
    k_count_kmers<<<nblocks(a.n_words * 32, 256), 256, 0, c->stream>>>(a);
    // histogram of this partition's counts (accumulates over the passes)
    unsigned int *n_big = reinterpret_cast<unsigned int *>(c->d_kc_n.as<unsigned long long>() + 1);
    CK(cudaMemsetAsync(n_big, 0, 4, c->stream));
    k_count_hist<<<c->sm_count * 8, 256, 0, c->stream>>>(a.table, slots, c->d_kc_hist.as<unsigned long long>(), HB, c->d_kc_big.as<uint32_t>(), BIG_CAP, n_big);
    {
      unsigned int nb = 0;
      CK(cudaMemcpyAsync(&nb, n_big, 4, cudaMemcpyDeviceToHost, c->stream));
      CK(cudaStreamSynchronize(c->stream));
      if (nb > BIG_CAP) return c->fail(OIC_ERR_CAPACITY, "oic_count_kmers: more than 2^20 K-mers seen >= 2^20 times");
      if (nb) {
        const size_t old_n = big_counts.size();
        big_counts.resize(old_n + nb);
        CK(cudaMemcpy(big_counts.data() + old_n, c->d_kc_big.p, (size_t)nb * 4, cudaMemcpyDeviceToHost));
      }
    }
    c->tm.launches += 1;
    // how many K-mers reach min_count, then collect them
    unsigned long long n = 0;
    CK(cudaMemsetAsync(c->d_kc_n.p, 0, 8, c->stream));
    k_count_collect<<<nblocks(slots, 256), 256, 0, c->stream>>>(a.table, slots, min_count, nullptr, nullptr, 0, c->d_kc_n.as<unsigned long long>());
    CK(cudaMemcpyAsync(&n, c->d_kc_n.p, 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    c->tm.launches += 2;
    if (n == 0) continue;
    CK(c->d_kc_out_keys.ensure(n * 8));
    CK(c->d_kc_out_cnt.ensure(n * 4));
    CK(cudaMemsetAsync(c->d_kc_n.p, 0, 8, c->stream));
    k_count_collect<<<nblocks(slots, 256), 256, 0, c->stream>>>(a.table, slots, min_count, c->d_kc_out_keys.as<unsigned long long>(),
                                                               c->d_kc_out_cnt.as<uint32_t>(), n, c->d_kc_n.as<unsigned long long>());
    std::vector<unsigned long long> hk(n);
    std::vector<uint32_t> hcnt(n);
    CK(cudaMemcpyAsync(hk.data(), c->d_kc_out_keys.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(hcnt.data(), c->d_kc_out_cnt.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    c->tm.launches += 1;
    c->tm.d2h_bytes += n * 12;
    const size_t old = c->freq.size();
    c->freq.resize(old + n);
    for (size_t i = 0; i < n; i++) c->freq[old + i] = std::make_pair((uint64_t)hk[i], hcnt[i]);
  }
  cudaEventRecord(c->ev_call[1], c->stream);
  CK(cudaEventSynchronize(c->ev_call[1]));
  cudaEventElapsedTime(&c->kc_ms, c->ev_call[0], c->ev_call[1]);
  {
    std::vector<unsigned long long> hh(HB);
    CK(cudaMemcpy(hh.data(), c->d_kc_hist.p, (size_t)HB * 8, cudaMemcpyDeviceToHost));
    uint64_t top = 0;
    for (uint32_t i = 0; i < HB; i++) if (hh[i]) top = i;
    for (uint32_t b : big_counts) top = std::max<uint64_t>(top, b);
    c->kc_hist.assign(top + 1, 0);
    for (uint32_t i = 0; i < HB && i <= top; i++) c->kc_hist[i] = hh[i];
    for (uint32_t b : big_counts) c->kc_hist[b]++;
  }
  std::sort(c->freq.begin(), c->freq.end());   // ascending K-mer order = the order of `meryl -Dt`
  if (n_found) *n_found = c->freq.size();
  if (n_passes) *n_passes = n_pass;
  if (device_ms) *device_ms = c->kc_ms;
  return rc;
}

int oic_copy_frequent_kmers(oic_ctx_t *c, uint64_t *keys, uint32_t *counts, uint64_t capacity, uint64_t *n_copied) {
  if (!c || (capacity && !keys)) return OIC_ERR_ARG;
  const uint64_t n = std::min<uint64_t>(capacity, c->freq.size());
  for (uint64_t i = 0; i < n; i++) { keys[i] = c->freq[i].first; if (counts) counts[i] = c->freq[i].second; }
  if (n_copied) *n_copied = n;
  return OIC_OK;
}

int oic_copy_kmer_histogram(oic_ctx_t *c, uint64_t *hist, uint64_t capacity, uint64_t *length) {
  if (!c) return OIC_ERR_ARG;
  if (length) *length = c->kc_hist.size();
  if (hist) for (uint64_t i = 0; i < capacity && i < c->kc_hist.size(); i++) hist[i] = c->kc_hist[i];
  return OIC_OK;
}

/* Overlap store construction on the GPU: filter + mirror, radix sort by (a_iid, b_iid), ties by the data words (oic_store.cuh). */
int oic_ovlstore_sort(oic_ctx_t *c, const oic_ovb_record_t *recs, uint64_t n, uint32_t max_id, const uint32_t *read_len, const uint8_t *skip_dup,
                      const uint8_t *library, uint32_t max_evalue, oic_ovb_record_t *sorted, uint64_t *n_sorted, uint64_t counts[9], double *device_ms) {
  if (!c) return OIC_ERR_ARG;
  if ((n && (!recs || !sorted)) || !read_len || !skip_dup || !library || max_id == 0) return c->fail(OIC_ERR_ARG, "oic_ovlstore_sort: bad argument");
  if (2 * n >= (1ull << 32)) return c->fail(OIC_ERR_CAPACITY, "oic_ovlstore_sort: more than 2^31 overlaps in one call");
  cudaSetDevice(c->device);
  if (n_sorted) *n_sorted = 0;
  if (device_ms) *device_ms = 0.0;
  if (counts) memset(counts, 0, 9 * sizeof(uint64_t));
  if (n == 0) return OIC_OK;
  DevBuf d_in, d_mir, d_out, d_len, d_skip, d_lib, d_cnt, d_keys[2], d_idx[2], d_hist;
  struct Release {
    std::vector<DevBuf *> v;
    ~Release() { for (DevBuf *b : v) b->release(); }
  } rel{{&d_in, &d_mir, &d_out, &d_len, &d_skip, &d_lib, &d_cnt, &d_keys[0], &d_keys[1], &d_idx[0], &d_idx[1], &d_hist}};
  CK(d_in.ensure(n * sizeof(StoreRec)));
  CK(d_mir.ensure(2 * n * sizeof(StoreRec)));
  CK(d_len.ensure((size_t)max_id * 4));
  CK(d_skip.ensure(max_id));
  CK(d_lib.ensure(max_id));
  CK(d_cnt.ensure(16 * 8));
  CK(cudaMemsetAsync(d_cnt.p, 0, 16 * 8, c->stream));
  CK(cudaMemcpyAsync(d_in.p, recs, n * sizeof(StoreRec), cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(d_len.p, read_len, (size_t)max_id * 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(d_skip.p, skip_dup, max_id, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(d_lib.p, library, max_id, cudaMemcpyHostToDevice, c->stream));
  c->tm.h2d_bytes += n * sizeof(StoreRec) + (uint64_t)max_id * 6;
  cudaEventRecord(c->ev_call[0], c->stream);
  MirrorArgs ma;
  ma.in = d_in.as<StoreRec>(); ma.n = n; ma.max_id = max_id; ma.rd_len = d_len.as<uint32_t>(); ma.skip_dup = d_skip.as<uint8_t>();
  ma.lib = d_lib.as<uint8_t>(); ma.max_evalue = max_evalue; ma.out = d_mir.as<StoreRec>();
  ma.n_out = d_cnt.as<unsigned long long>() + 15; ma.counts = d_cnt.as<unsigned long long>();
  k_store_mirror<<<nblocks(n, 256), 256, 0, c->stream>>>(ma);
  unsigned long long hcnt[16];
  CK(cudaMemcpyAsync(hcnt, d_cnt.p, sizeof(hcnt), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  c->tm.launches += 1;
  if (counts) for (int i = 0; i < 9; i++) counts[i] = hcnt[i];
  const uint64_t m = hcnt[15];
  if (hcnt[0]) return c->fail(OIC_ERR_ARG, "oic_ovlstore_sort: %llu overlaps have IDs out of range (maxID %u), possibly corrupt input data", hcnt[0], max_id);
  if (m) {
    for (int i = 0; i < 2; i++) { CK(d_keys[i].ensure(m * 8)); CK(d_idx[i].ensure(m * 4)); }
    CK(d_out.ensure(m * sizeof(StoreRec)));
    const uint32_t nb = nblocks(m, RADIX_TILE);
    CK(d_hist.ensure((size_t)2 * 256 * nb * 4 + 16));   // digit histogram, then its exclusive scan
    const uint64_t hn = (uint64_t)256 * nb, hb = (hn + SCAN_ITEMS - 1) / SCAN_ITEMS;
    CK(c->d_blk.ensure((hb + 2) * 4));
    k_store_keys<<<nblocks(m, 256), 256, 0, c->stream>>>(d_mir.as<StoreRec>(), m, d_keys[0].as<unsigned long long>(), d_idx[0].as<uint32_t>());
    int id_bits = 1;
    while ((1ull << id_bits) < max_id) id_bits++;
    int cur = 0;
    for (int half = 0; half < 2; half++)                       // b_iid digits first (low word), then a_iid
      for (int shift = 32 * half; shift < 32 * half + id_bits; shift += 8) {
        k_radix_hist<<<nb, RADIX_THREADS, 0, c->stream>>>(d_keys[cur].as<unsigned long long>(), m, shift, d_hist.as<uint32_t>(), nb);
        k_scan_u32_reduce<<<(unsigned)hb, SCAN_THREADS, 0, c->stream>>>(d_hist.as<uint32_t>(), hn, c->d_blk.as<uint32_t>());
        k_scan_u32_single<<<1, SCAN_THREADS, 0, c->stream>>>(c->d_blk.as<uint32_t>(), (uint32_t)hb);
        k_scan_u32_apply<<<(unsigned)hb, SCAN_THREADS, 0, c->stream>>>(d_hist.as<uint32_t>(), d_hist.as<uint32_t>() + hn, hn, c->d_blk.as<uint32_t>());
        k_radix_scatter<<<nb, RADIX_THREADS, 0, c->stream>>>(d_keys[cur].as<unsigned long long>(), d_idx[cur].as<uint32_t>(), m, shift, d_hist.as<uint32_t>() + hn, nb,
                                                             d_keys[cur ^ 1].as<unsigned long long>(), d_idx[cur ^ 1].as<uint32_t>());
        c->tm.launches += 5;
        cur ^= 1;
      }
    k_store_tiefix<<<nblocks(m, 256), 256, 0, c->stream>>>(d_keys[cur].as<unsigned long long>(), d_idx[cur].as<uint32_t>(), d_mir.as<StoreRec>(), m);
    k_store_gather<<<nblocks(m, 256), 256, 0, c->stream>>>(d_mir.as<StoreRec>(), d_idx[cur].as<uint32_t>(), m, d_out.as<StoreRec>());
    c->tm.launches += 3;
    cudaEventRecord(c->ev_call[1], c->stream);
    CK(cudaMemcpyAsync(sorted, d_out.p, m * sizeof(StoreRec), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    c->tm.d2h_bytes += m * sizeof(StoreRec);
    float ms = 0;
    if (device_ms && cudaEventElapsedTime(&ms, c->ev_call[0], c->ev_call[1]) == cudaSuccess) *device_ms = ms;
  }
  if (n_sorted) *n_sorted = m;
  return OIC_OK;
}

/* ---- allocation-free variant of oic_ovlstore_sort (no counterpart in the reference)
 * All intermediate buffers of the sort in one static device array.  Layout (bytes, every piece 256-aligned), for at most
 * NOALLOC_MAX_N input overlaps = NOALLOC_MAX_M mirrored records:
 *   counters 256 | mirrored records M x 24 | keys[2] M x 8 each | idx[2] M x 4 each | digit histogram + its scan 2 x 256 x blocks x 4 | block sums 256 */
constexpr uint64_t NOALLOC_MAX_N = 65536, NOALLOC_MAX_M = 2 * NOALLOC_MAX_N;
constexpr uint64_t NOALLOC_MAX_NB = (NOALLOC_MAX_M + RADIX_TILE - 1) / RADIX_TILE;
constexpr size_t NA_OFF_CNT = 0;
constexpr size_t NA_OFF_MIR = NA_OFF_CNT + 256;
constexpr size_t NA_OFF_KEY0 = NA_OFF_MIR + NOALLOC_MAX_M * sizeof(StoreRec);
constexpr size_t NA_OFF_KEY1 = NA_OFF_KEY0 + NOALLOC_MAX_M * 8;
constexpr size_t NA_OFF_IDX0 = NA_OFF_KEY1 + NOALLOC_MAX_M * 8;
constexpr size_t NA_OFF_IDX1 = NA_OFF_IDX0 + NOALLOC_MAX_M * 4;
constexpr size_t NA_OFF_HIST = NA_OFF_IDX1 + NOALLOC_MAX_M * 4;
constexpr size_t NA_OFF_BLK = NA_OFF_HIST + 2 * 256 * NOALLOC_MAX_NB * 4;
constexpr size_t NOALLOC_BYTES = NA_OFF_BLK + 256;
static_assert(NOALLOC_BYTES == 6357504, "the size stated in include/oic_b200.h, DESIGN.md and INTEGRATION.md");
static_assert(NA_OFF_MIR % 256 == 0 && NA_OFF_KEY0 % 256 == 0 && NA_OFF_KEY1 % 256 == 0 && NA_OFF_IDX0 % 256 == 0 && NA_OFF_IDX1 % 256 == 0 &&
              NA_OFF_HIST % 256 == 0 && NA_OFF_BLK % 256 == 0, "every piece of the scratch is 256-byte aligned");
static_assert((256 * NOALLOC_MAX_NB + SCAN_ITEMS - 1) / SCAN_ITEMS + 2 <= 64, "block sums fit their 256 bytes");

// In HOST code this name is a host-side stand-in, not a device address: a kernel given it faults (or, where the GPU can reach
// host memory, silently uses the wrong memory).  Host code takes the address from cudaGetSymbolAddress, per device (below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

static std::mutex g_noalloc_mutex;
static unsigned char *g_noalloc_ptr[64];   // device address of g_noalloc_scratch on device i (nullptr: not looked up yet)

uint64_t oic_ovlstore_sort_noalloc_max(void) { return NOALLOC_MAX_N; }

int oic_ovlstore_sort_noalloc(int device, const oic_ovb_record_t *d_recs, uint64_t n, uint32_t max_id, const uint32_t *d_read_len,
                              const uint8_t *d_skip_dup, const uint8_t *d_library, uint32_t max_evalue, oic_ovb_record_t *d_sorted,
                              uint64_t *n_sorted, uint64_t counts[9]) {
  auto fail = [](int code, const char *fmt, auto... args) {
    char buf[512];
    snprintf(buf, sizeof(buf), fmt, args...);
    g_create_error = buf;   // read back with oic_last_error(NULL) on this thread
    return code;
  };
  if (n_sorted) *n_sorted = 0;
  if (counts) memset(counts, 0, 9 * sizeof(uint64_t));
  if (n > NOALLOC_MAX_N)
    return fail(OIC_ERR_CAPACITY, "oic_ovlstore_sort_noalloc: %llu overlaps exceed the static scratch (at most %llu per call: split the input or use oic_ovlstore_sort)",
                (unsigned long long)n, (unsigned long long)NOALLOC_MAX_N);
  if (device < 0 || device >= 64 || max_id == 0 || !d_read_len || !d_skip_dup || !d_library || (n && (!d_recs || !d_sorted)))
    return fail(OIC_ERR_ARG, "oic_ovlstore_sort_noalloc: bad argument");
  if (n == 0) return OIC_OK;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(OIC_ERR_CUDA, "oic_ovlstore_sort_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  unsigned char *base = nullptr;
  {   // set-up, once per device, with that device current
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
      void *p = nullptr;
      e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p) return fail(OIC_ERR_CUDA, "oic_ovlstore_sort_noalloc: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
      g_noalloc_ptr[device] = static_cast<unsigned char *>(p);
    }
    base = g_noalloc_ptr[device];
  }
#define NK_CUDA(call)                                                                                                     \
  do {                                                                                                                    \
    cudaError_t e_ = (call);                                                                                              \
    if (e_ != cudaSuccess) return fail(OIC_ERR_CUDA, "oic_ovlstore_sort_noalloc: %s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)
  unsigned long long *cnt = reinterpret_cast<unsigned long long *>(base + NA_OFF_CNT);
  StoreRec *mir = reinterpret_cast<StoreRec *>(base + NA_OFF_MIR);
  unsigned long long *keys[2] = {reinterpret_cast<unsigned long long *>(base + NA_OFF_KEY0), reinterpret_cast<unsigned long long *>(base + NA_OFF_KEY1)};
  uint32_t *idx[2] = {reinterpret_cast<uint32_t *>(base + NA_OFF_IDX0), reinterpret_cast<uint32_t *>(base + NA_OFF_IDX1)};
  uint32_t *hist = reinterpret_cast<uint32_t *>(base + NA_OFF_HIST), *blk = reinterpret_cast<uint32_t *>(base + NA_OFF_BLK);
  const cudaStream_t st = 0;   // legacy default stream: creating a stream would allocate
  // the array keeps the last call's contents: the counters must start at zero, everything else is written before it is read
  NK_CUDA(cudaMemsetAsync(cnt, 0, 256, st));
  MirrorArgs ma;
  ma.in = reinterpret_cast<const StoreRec *>(d_recs); ma.n = n; ma.max_id = max_id; ma.rd_len = d_read_len; ma.skip_dup = d_skip_dup;
  ma.lib = d_library; ma.max_evalue = max_evalue; ma.out = mir; ma.n_out = cnt + 15; ma.counts = cnt;
  k_store_mirror<<<nblocks(n, 256), 256, 0, st>>>(ma);
  unsigned long long hcnt[16];
  NK_CUDA(cudaMemcpyAsync(hcnt, cnt, sizeof(hcnt), cudaMemcpyDeviceToHost, st));
  NK_CUDA(cudaStreamSynchronize(st));
  NK_CUDA(cudaGetLastError());
  if (counts) for (int i = 0; i < 9; i++) counts[i] = hcnt[i];
  const uint64_t m = hcnt[15];
  if (hcnt[0]) return fail(OIC_ERR_ARG, "oic_ovlstore_sort_noalloc: %llu overlaps have IDs out of range (maxID %u), possibly corrupt input data", hcnt[0], max_id);
  if (m > NOALLOC_MAX_M) return fail(OIC_ERR_STATE, "oic_ovlstore_sort_noalloc: internal: %llu records from %llu overlaps", (unsigned long long)m, (unsigned long long)n);
  if (m) {
    const uint32_t nb = nblocks(m, RADIX_TILE);
    const uint64_t hn = (uint64_t)256 * nb, hb = (hn + SCAN_ITEMS - 1) / SCAN_ITEMS;
    k_store_keys<<<nblocks(m, 256), 256, 0, st>>>(mir, m, keys[0], idx[0]);
    int id_bits = 1;
    while ((1ull << id_bits) < max_id) id_bits++;
    int cur = 0;
    for (int half = 0; half < 2; half++)                       // b_iid digits first (low word), then a_iid
      for (int shift = 32 * half; shift < 32 * half + id_bits; shift += 8) {
        k_radix_hist<<<nb, RADIX_THREADS, 0, st>>>(keys[cur], m, shift, hist, nb);
        k_scan_u32_reduce<<<(unsigned)hb, SCAN_THREADS, 0, st>>>(hist, hn, blk);
        k_scan_u32_single<<<1, SCAN_THREADS, 0, st>>>(blk, (uint32_t)hb);
        k_scan_u32_apply<<<(unsigned)hb, SCAN_THREADS, 0, st>>>(hist, hist + hn, hn, blk);
        k_radix_scatter<<<nb, RADIX_THREADS, 0, st>>>(keys[cur], idx[cur], m, shift, hist + hn, nb, keys[cur ^ 1], idx[cur ^ 1]);
        cur ^= 1;
      }
    k_store_tiefix<<<nblocks(m, 256), 256, 0, st>>>(keys[cur], idx[cur], mir, m);
    k_store_gather<<<nblocks(m, 256), 256, 0, st>>>(mir, idx[cur], m, reinterpret_cast<StoreRec *>(d_sorted));   // straight into the caller's buffer
    NK_CUDA(cudaStreamSynchronize(st));
    NK_CUDA(cudaGetLastError());
  }
#undef NK_CUDA
  if (n_sorted) *n_sorted = m;
  return OIC_OK;
}

/* Measured INT32 issue rates of this GPU in Gop/s (thread operations): alu = LOP3/IADD3 chains, imad = IMAD chains. */
int oic_measure_int32_peak(int device, double *alu_gops, double *imad_gops) {
  if (cudaSetDevice(device) != cudaSuccess) return OIC_ERR_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return OIC_ERR_CUDA;
  uint32_t *d = nullptr;
  if (cudaMalloc(&d, 64) != cudaSuccess) return OIC_ERR_CUDA;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
  double res[2] = {0, 0};
  for (int mode = 0; mode < 2; mode++) {
    for (int rep = 0; rep < 3; rep++) {
      cudaEventRecord(e0);
      if (mode == 0) k_int_probe<0><<<blocks, threads>>>(d, iters, 17u + rep);
      else k_int_probe<1><<<blocks, threads>>>(d, iters, 17u + rep);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double ops = (double)blocks * threads * iters * 8.0 * (mode == 0 ? 2.0 : 1.0);
      if (rep && ms > 0) res[mode] = std::max(res[mode], ops / (ms * 1e-3) / 1e9);
    }
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  if (cudaGetLastError() != cudaSuccess) return OIC_ERR_CUDA;
  if (alu_gops) *alu_gops = res[0];
  if (imad_gops) *imad_gops = res[1];
  return OIC_OK;
}

}  // extern "C"
