/*
 * overlapInCore.cpp -- drop-in replacement of the reference's overlapInCore executable (OIC/overlapInCore.C:306-595):
 * same command line, gkpStore input, <-o>.ovb / <base>.counts / <-s> stats outputs.  All overlap computation is done
 * by liboic_b200.so (CUDA, include/oic_b200.h); this file is argv, file formats and the OverlapDriver loop.
 *
 * Additive options (ignored by nothing in the pipeline because the reference does not have them):
 *   --gpus N      shard the query range of every hash sub-block over N GPUs (one host thread + one context per GPU).  The hash
 *                 block is built ONCE by the group: every GPU inserts the k-mers of the table slices it owns and the slices are
 *                 exchanged with NCCL over NVLink (oic_comm_init_all + oic_build_hash_index_sharded); overlap records are
 *                 gathered on the host (one writer).  --replicate rebuild: every GPU builds the whole block itself, no collective.
 *   --device D    first CUDA device to use (default 0)
 *   --sweep PFX   run EVERY job of an overlapInCorePartition job list (PFX.ovlbat / .ovljob / .ovlopt) in this one process:
 *                 the store is opened and its reads sent to the GPU once, consecutive query blocks of a hash block reuse
 *                 the built index, and every job writes the files the pipeline expects from its own process
 *                 (OverlapInCore.pm:184-233): DIR/<batch>/<job>.ovb, .counts, .stats with DIR from --sweep-out (default ".").
 *                 -h / -r / --hashstrings / --hashdatalen come from the job list, every other option applies to all jobs.
 */
#include <ctype.h>
#include <errno.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "../../include/oic_b200.h"
#include "partition.h"
#include "stores.h"

using namespace oichost;

namespace {

// AS_UTL_decodeRange (UTL/AS_UTL_decodeRange.C:103-152): "a-b" or "a"
void decode_range(const char *s, uint32_t &lo, uint32_t &hi) {
  char *end = nullptr;
  unsigned long long a = strtoull(s, &end, 10);
  if (end == s) { fprintf(stderr, "ERROR: invalid range '%s'\n", s); exit(1); }
  unsigned long long b = a;
  if (*end == '-') b = strtoull(end + 1, nullptr, 10);
  lo = (uint32_t)a;
  hi = (uint32_t)b;
}

void usage(const char *argv0) {
  fprintf(stderr, "USAGE:  %s [options] <gkpStorePath>\n\n", argv0);
  fprintf(stderr, "-G          do partial overlaps\n");
  fprintf(stderr, "-h <range>  to specify fragments to put in hash table\n");
  fprintf(stderr, "-H <range>  libraries to put in the hash table\n");
  fprintf(stderr, "-l <n>      at most <n> overlaps off each end of a read (default: no limit)\n");
  fprintf(stderr, "-k          if one or two digits, the length of a kmer, otherwise\n");
  fprintf(stderr, "            the filename containing a list of kmers to ignore in the hash table\n");
  fprintf(stderr, "-m          allow multiple overlaps per oriented fragment pair\n");
  fprintf(stderr, "-o          specify output file name\n");
  fprintf(stderr, "-r <range>  specify old fragments to overlap\n");
  fprintf(stderr, "-R <range>  libraries to overlap\n");
  fprintf(stderr, "-s          specify statistics file name\n");
  fprintf(stderr, "-t <n>      use <n> parallel threads (only decides which reads the reference's scheduler would process)\n");
  fprintf(stderr, "-u          allow only 1 overlap per oriented fragment pair\n");
  fprintf(stderr, "-w          filter out overlaps with too many errors in a window\n");
  fprintf(stderr, "-z          skip the hopeless check\n\n");
  fprintf(stderr, "--maxerate <n>     only output overlaps with fraction <n> or less error (e.g., 0.06 == 6%%)\n");
  fprintf(stderr, "--minlength <n>    only output overlaps of <n> or more bases\n");
  fprintf(stderr, "--minkmers         filter pairs by expected k-mer count\n\n");
  fprintf(stderr, "--hashbits n       n bits for the hash mask (here: only sets the load limit).\n");
  fprintf(stderr, "--hashstrings n    Load at most n strings into the hash table at one time.\n");
  fprintf(stderr, "--hashdatalen n    Load at most n bytes into the hash table at one time.\n");
  fprintf(stderr, "--hashload f       Load to at most 0.0 < f < 1.0 capacity (default 0.7).\n");
  fprintf(stderr, "--maxreadlen n     accepted, no effect on results\n\n");
  fprintf(stderr, "--gpus n           B200s to shard the query reads over (default 1)\n");
  fprintf(stderr, "--device d         first CUDA device (default 0)\n");
  fprintf(stderr, "--replicate how    --gpus > 1: 'sharded' (default: the group builds the hash block once, NCCL exchange) or 'rebuild'\n");
  fprintf(stderr, "--sweep prefix     run all jobs of prefix.ovlbat/.ovljob/.ovlopt (overlapInCorePartition) in this process\n");
  fprintf(stderr, "--sweep-out dir    directory of the per-job outputs <dir>/<batch>/<job>.ovb/.counts/.stats (default .)\n");
}

double elapsed() {   // seconds since the first call (progress lines on stderr)
  static const auto t0 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

struct Shard {
  oic_ctx_t *ctx = nullptr;
  int rc = 0;
  std::string err;
};

}  // namespace

int main(int argc, char **argv) {
  oic_params_t P;
  memset(&P, 0, sizeof(P));
  // oicParameters::initialize() (OIC/overlapInCore.H:425-464)
  P.max_erate = 0.06;
  P.bgn_hash_id = 1; P.end_hash_id = UINT32_MAX;
  P.bgn_ref_id = 1; P.end_ref_id = UINT32_MAX;
  P.frag_olap_limit = UINT64_MAX;
  P.unique_olap = 1;
  P.hash_bits = 22; P.hash_load = 0.6; P.max_hash_strings = 10000; P.max_hash_data_len = 100000000;
  P.num_threads = 1;
  P.use_hopeless_check = 1;
  uint32_t min_lib_hash = 0, max_lib_hash = UINT32_MAX, min_lib_ref = 0, max_lib_ref = UINT32_MAX;
  const char *skip_file = nullptr, *out_name = nullptr, *stat_name = nullptr, *store_path = nullptr;
  bool window_filter = false;
  int n_gpus = 1, device0 = 0;
  bool sharded_build = true;
  const char *sweep_prefix = nullptr, *sweep_out = ".";
  uint64_t max_string_num = (1ull << 31) - 1;

  if (argc == 2 && strcmp(argv[1], "--version") == 0) { fprintf(stdout, "overlapInCore (B200-native, Canu 1.6 compatible)\n"); return 0; }

  int err = 0;
  for (int arg = 1; arg < argc; arg++) {
    const char *a = argv[arg];
    auto need = [&](void) -> const char * {
      if (arg + 1 >= argc) { fprintf(stderr, "ERROR: option '%s' needs a value\n", a); exit(1); }
      return argv[++arg];
    };
    if (strcmp(a, "-G") == 0) P.partial = 1;
    else if (strcmp(a, "-h") == 0) decode_range(need(), P.bgn_hash_id, P.end_hash_id);
    else if (strcmp(a, "-H") == 0) decode_range(need(), min_lib_hash, max_lib_hash);
    else if (strcmp(a, "-r") == 0) decode_range(need(), P.bgn_ref_id, P.end_ref_id);
    else if (strcmp(a, "-R") == 0) decode_range(need(), min_lib_ref, max_lib_ref);
    else if (strcmp(a, "-k") == 0) {
      const char *v = need();   // one or two digits = k-mer length, anything else = skip file (OIC/overlapInCore.C:332-342)
      if ((isdigit(v[0]) && v[1] == 0) || (isdigit(v[0]) && isdigit(v[1]) && v[2] == 0)) P.kmer_len = (int32_t)strtoull(v, nullptr, 10);
      else {
        FILE *t = fopen(v, "r");
        if (!t) { fprintf(stderr, "ERROR: Failed to open -k '%s': %s\n", v, strerror(errno)); exit(1); }
        fclose(t);
        skip_file = v;
      }
    } else if (strcmp(a, "-l") == 0) {
      long v = strtol(need(), nullptr, 10);
      P.frag_olap_limit = v < 1 ? UINT64_MAX : (uint64_t)v;
    } else if (strcmp(a, "-m") == 0) P.unique_olap = 0;
    else if (strcmp(a, "-u") == 0) P.unique_olap = 1;
    else if (strcmp(a, "--hashbits") == 0) P.hash_bits = (uint32_t)strtoull(need(), nullptr, 10);
    else if (strcmp(a, "--hashstrings") == 0) P.max_hash_strings = (uint32_t)strtoull(need(), nullptr, 10);
    else if (strcmp(a, "--hashdatalen") == 0) P.max_hash_data_len = strtoull(need(), nullptr, 10);
    else if (strcmp(a, "--hashload") == 0) P.hash_load = atof(need());
    else if (strcmp(a, "--maxreadlen") == 0) {
      // re-splits the reference's 30-bit string reference (OIC/overlapInCore.C:366-378): only the string limit matters here
      uint32_t desired = (uint32_t)strtoul(need(), nullptr, 10), bits = 1;
      while ((1u << bits) < desired) bits++;
      max_string_num = (1ull << (30 - bits)) - 1;
    } else if (strcmp(a, "-o") == 0) out_name = need();
    else if (strcmp(a, "-s") == 0) stat_name = need();
    else if (strcmp(a, "-t") == 0) P.num_threads = (uint32_t)strtoull(need(), nullptr, 10);
    else if (strcmp(a, "--minlength") == 0) P.min_olap_len = (int32_t)strtol(need(), nullptr, 10);
    else if (strcmp(a, "--minkmers") == 0)   // evaluated with the K / erate / minlength seen SO FAR (OIC/overlapInCore.C:392-393)
      P.filter_by_kmer_count = (uint64_t)(int)floor(exp(-1.0 * (double)P.kmer_len * P.max_erate) * (P.min_olap_len - P.kmer_len + 1));
    else if (strcmp(a, "--maxerate") == 0) P.max_erate = strtof(need(), nullptr);   // float precision, widened (:395)
    else if (strcmp(a, "-w") == 0) window_filter = true;
    else if (strcmp(a, "-z") == 0) P.use_hopeless_check = 0;
    else if (strcmp(a, "--gpus") == 0) n_gpus = atoi(need());
    else if (strcmp(a, "--device") == 0) device0 = atoi(need());
    else if (strcmp(a, "--replicate") == 0) {
      const char *v = need();
      if (strcmp(v, "sharded") == 0) sharded_build = true;
      else if (strcmp(v, "rebuild") == 0) sharded_build = false;
      else { fprintf(stderr, "ERROR: --replicate takes 'sharded' or 'rebuild'\n"); err++; }
    }
    else if (strcmp(a, "--sweep") == 0) sweep_prefix = need();
    else if (strcmp(a, "--sweep-out") == 0) sweep_out = need();
    else if (store_path == nullptr) store_path = a;
    else { fprintf(stderr, "Unknown option '%s'\n", a); err++; }
  }
  if (P.max_erate > 0.06) {   // OIC/overlapInCore.C:416-421
    if (window_filter) fprintf(stderr, "High error rates requested -- window-filter turned off despite -w flag!\n");
    window_filter = false;
    P.use_hopeless_check = 0;
  }
  if (P.max_hash_strings == 0) fprintf(stderr, "* No memory model supplied; -M needed!\n"), err++;
  if (P.kmer_len == 0) fprintf(stderr, "* No kmer length supplied; -k needed!\n"), err++;
  if (P.max_hash_strings > max_string_num) fprintf(stderr, "Too many strings (--hashstrings), must be less than %llu\n", (unsigned long long)max_string_num), err++;
  if (out_name == nullptr && sweep_prefix == nullptr) fprintf(stderr, "ERROR:  No output file name specified\n"), err++;
  if (err || store_path == nullptr) { usage(argv[0]); return 1; }
  P.use_window_filter = window_filter ? 1 : 0;
  if (P.num_threads == 0) P.num_threads = 1;
  if (n_gpus < 1) n_gpus = 1;

  // ---- input
  elapsed();
  GkStore store(store_path);
  const uint32_t num_reads = store.num_reads();
  std::vector<OvlJob> jobs;   // --sweep: the job list; the reads loaded are the union of all ranges
  if (sweep_prefix) {
    if (!read_job_files(sweep_prefix, jobs) || jobs.empty()) { fprintf(stderr, "ERROR: cannot read job list '%s.ovlbat/.ovljob/.ovlopt'\n", sweep_prefix); return 1; }
    P.bgn_hash_id = P.bgn_ref_id = UINT32_MAX;
    P.end_hash_id = P.end_ref_id = 0;
    for (const OvlJob &j : jobs) {
      P.bgn_hash_id = std::min(P.bgn_hash_id, j.hash_bgn); P.end_hash_id = std::max(P.end_hash_id, j.hash_end);
      P.bgn_ref_id = std::min(P.bgn_ref_id, j.ref_bgn); P.end_ref_id = std::max(P.end_ref_id, j.ref_end);
    }
  }
  if (P.bgn_hash_id < 1) P.bgn_hash_id = 1;
  if (P.end_hash_id > num_reads) P.end_hash_id = num_reads;
  if (P.bgn_ref_id < 1) P.bgn_ref_id = 1;
  if (P.end_ref_id > num_reads) P.end_ref_id = num_reads;
  const uint32_t first_id = std::min(P.bgn_hash_id, P.bgn_ref_id);
  const uint32_t last_id = std::max(P.end_hash_id, P.end_ref_id);
  const uint32_t n = last_id >= first_id ? last_id - first_id + 1 : 0;
  std::vector<uint64_t> offsets(n);
  std::vector<uint32_t> lengths(n);
  std::vector<uint8_t> hash_ok(n), ref_ok(n);
  uint64_t total = 0;
  for (uint32_t i = 0; i < n; i++) {
    const GkRead &r = store.read(first_id + i);
    offsets[i] = total;
    lengths[i] = r.seq_len();
    total += r.seq_len();
    const uint32_t lib = r.library_id();
    hash_ok[i] = !(lib < min_lib_hash || lib > max_lib_hash);   // Build_Hash_Index.C:555-557
    ref_ok[i] = !(lib < min_lib_ref || lib > max_lib_ref);      // Process_Overlaps.C:108-110
  }
  // The store keeps ACGT-only reads 2-bit encoded (`2SEQ`); those bytes go to the GPU as they are (oic_load_reads_2bit).  If any
  // read of the range is stored unencoded (`USEQ`), the whole range takes the ASCII path (gkRead_loadData semantics).
  bool two_bit = getenv("OIC_ASCII_LOAD") == nullptr;
  for (uint32_t i = 0; i < n && two_bit; i++)
    if (lengths[i] && !store.sequence_2bit(first_id + i)) two_bit = false;
  std::vector<char> bases;
  const unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  if (two_bit) {
    uint64_t tb = 0;
    for (uint32_t i = 0; i < n; i++) { offsets[i] = tb; tb += ((uint64_t)lengths[i] + 3) / 4; }
    bases.resize(tb + 8);
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
      th.emplace_back([&, t]() {
        for (uint64_t i = (uint64_t)n * t / nt, e = (uint64_t)n * (t + 1) / nt; i < e; i++)
          if (lengths[i]) memcpy(bases.data() + offsets[i], store.sequence_2bit(first_id + (uint32_t)i), ((uint64_t)lengths[i] + 3) / 4);
      });
    for (auto &t : th) t.join();
  } else {
    bases.resize(total + 1);
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
      th.emplace_back([&, t]() {
        for (uint64_t i = (uint64_t)n * t / nt, e = (uint64_t)n * (t + 1) / nt; i < e; i++)
          store.load_sequence(first_id + (uint32_t)i, bases.data() + offsets[i]);
      });
    for (auto &t : th) t.join();
  }
  fprintf(stderr, "Loaded reads %u-%u (%llu bases) from '%s'  [%.2f s]\n", first_id, last_id, (unsigned long long)total, store_path, elapsed());

  // -w: the reads' quality values (QVAL constants / UQLT bytes), one byte per base, in read order
  std::vector<uint8_t> quals;
  std::vector<uint64_t> qual_off;
  if (window_filter) {
    qual_off.resize(n);
    uint64_t at = 0;
    for (uint32_t i = 0; i < n; i++) { qual_off[i] = at; at += lengths[i]; }
    quals.resize(at + 1);
    for (uint32_t i = 0; i < n; i++) store.load_qualities(first_id + i, quals.data() + qual_off[i]);
  }

  std::vector<char> skip;   // Mark_Skip_Kmers input: '>' header line, then exactly K bases (Build_Hash_Index.C:233-286)
  uint64_t n_skip = 0;
  if (skip_file) {
    FILE *f = fopen(skip_file, "r");
    char line[1000];
    int ct = 0;
    while (fgets(line, sizeof(line), f)) {
      ct++;
      int len = (int)strlen(line) - 1;
      if (line[0] != '>' || len < 0 || line[len] != '\n') { fprintf(stderr, "ERROR:  Bad line %d in kmer skip file\n%s", ct, line); return 1; }
      if (!fgets(line, sizeof(line), f)) { fprintf(stderr, "ERROR:  Bad line after %d in kmer skip file\n", ct); return 1; }
      ct++;
      len = (int)strlen(line) - 1;
      if (len != P.kmer_len || line[len] != '\n') { fprintf(stderr, "ERROR:  Bad line %d in kmer skip file\n%s", ct, line); return 1; }
      skip.insert(skip.end(), line, line + len);
      n_skip++;
    }
    fclose(f);
    fprintf(stderr, "Read %llu kmers to mark to skip\n", (unsigned long long)n_skip);
  }

  // ---- contexts, one per GPU
  std::vector<Shard> sh(n_gpus);
  // run fn(g) on the contexts g0 .. g0+ng-1, one host thread each (the calling thread when there is only one)
  auto on_range = [&](int g0, int ng, auto fn) -> bool {
    auto body = [&](int g) { sh[g].rc = fn(g); if (sh[g].rc) sh[g].err = oic_last_error(sh[g].ctx); };
    if (ng == 1) body(g0);
    else {
      std::vector<std::thread> th;
      for (int g = g0; g < g0 + ng; g++) th.emplace_back(body, g);
      for (auto &t : th) t.join();
    }
    for (int g = g0; g < g0 + ng; g++)
      if (sh[g].rc) { fprintf(stderr, "ERROR (GPU %d): %s\n", device0 + g, sh[g].err.c_str()); return false; }
    return true;
  };
  if (n == 0) { fprintf(stderr, "ERROR: empty read range\n"); return 1; }
  // CUDA context + the reads of the whole range resident on GPU g
  auto bring_up = [&](int g) -> int {
    int rc = oic_create(&P, device0 + g, &sh[g].ctx);
    if (rc) { sh[g].err = oic_last_error(nullptr); return rc; }
    rc = two_bit ? oic_load_reads_2bit(sh[g].ctx, first_id, n, offsets.data(), lengths.data(), (const uint8_t *)bases.data(), hash_ok.data(), ref_ok.data())
                 : oic_load_reads(sh[g].ctx, first_id, n, offsets.data(), lengths.data(), bases.data(), hash_ok.data(), ref_ok.data());
    if (!rc && window_filter) rc = oic_load_qualities(sh[g].ctx, qual_off.data(), quals.data());
    if (!rc && n_skip) rc = oic_set_skip_kmers(sh[g].ctx, n_skip, skip.data());
    if (rc) sh[g].err = oic_last_error(sh[g].ctx);
    fprintf(stderr, "GPU %d: context up, reads resident  [%.2f s]\n", device0 + g, elapsed());
    return rc;
  };
  const bool lazy_bring_up = sweep_prefix != nullptr && n_gpus > 1;   // sweep workers bring their own GPU up (see below)
  if (!lazy_bring_up) {
    std::vector<std::thread> th;
    for (int g = 0; g < n_gpus; g++) th.emplace_back([&, g]() { sh[g].rc = bring_up(g); });
    for (auto &t : th) t.join();
    for (int g = 0; g < n_gpus; g++)
      if (sh[g].rc) { fprintf(stderr, "ERROR (GPU %d): %s\n", device0 + g, sh[g].err.c_str()); return 1; }
  }
  // a plain job on several GPUs: one NCCL communicator over the contexts (a sweep hands WHOLE jobs to the GPUs: no group there)
  bool have_comm = false;
  if (n_gpus > 1 && !sweep_prefix && sharded_build) {
    std::vector<oic_ctx_t *> ctxs(n_gpus);
    for (int g = 0; g < n_gpus; g++) ctxs[g] = sh[g].ctx;
    if (oic_comm_init_all(ctxs.data(), n_gpus) != 0) {
      fprintf(stderr, "ERROR: NCCL communicator over %d GPUs: %s (use --replicate rebuild to run without one)\n", n_gpus, oic_last_error(sh[0].ctx));
      return 1;
    }
    have_comm = true;
    fprintf(stderr, "NCCL communicator over %d GPUs  [%.2f s]\n", n_gpus, elapsed());
  }

  // One job = what one process of the reference does with its -h / -r / --hashstrings / --hashdatalen.
  // It runs on the contexts g0 .. g0+ng-1: the query range of every hash sub-block is split over them.
  auto run_job = [&](const oic_params_t &J, const char *job_out, const char *job_stat, int g0, int ng) -> int {
    auto on_all = [&](auto fn) -> bool { return on_range(g0, ng, fn); };
    const int n_gpus = ng;
    if (!on_all([&](int g) {
          oic_reset_stats(sh[g].ctx);
          return oic_set_job(sh[g].ctx, J.bgn_hash_id, J.end_hash_id, J.bgn_ref_id, J.end_ref_id, J.max_hash_strings, J.max_hash_data_len);
        })) return 1;
    // ---- OverlapDriver (OIC/overlapInCore.C:190-300)
    OvbWriter out(job_out);
    std::vector<uint32_t> rec;
    uint64_t bgn = J.bgn_hash_id, end = (uint64_t)J.bgn_hash_id + J.max_hash_strings - 1;
    while (bgn < J.end_hash_id) {
      if (end > J.end_hash_id) end = J.end_hash_id;
      std::vector<uint32_t> last(n_gpus, 0);
      if (!on_all([&](int g) {
            return (have_comm && ng > 1) ? oic_build_hash_index_sharded(sh[g].ctx, nullptr, (uint32_t)bgn, (uint32_t)end, &last[g - g0])
                                         : oic_build_hash_index(sh[g].ctx, (uint32_t)bgn, (uint32_t)end, &last[g - g0]);
          })) return 1;
      fprintf(stderr, "Hash block %llu-%u built\n", (unsigned long long)bgn, last[0]);
      const uint32_t last_ref = oic_last_processed_ref(J.bgn_ref_id, J.end_ref_id, J.num_threads);
      if (last_ref >= J.bgn_ref_id) {
        // contiguous shards balanced by bases x expected partners: a query only pairs with hash reads of higher ID
        std::vector<double> w(last_ref - J.bgn_ref_id + 2, 0.0);
        for (uint32_t id = J.bgn_ref_id; id <= last_ref; id++) {
          double partners = id < bgn ? (double)(last[0] - bgn + 1) : (id < last[0] ? (double)(last[0] - id) : 0.0);
          w[id - J.bgn_ref_id + 1] = w[id - J.bgn_ref_id] + (double)lengths[id - first_id] * partners + 1.0;
        }
        std::vector<uint32_t> cut(n_gpus + 1, J.bgn_ref_id);
        for (int g = 1; g <= n_gpus; g++) {
          double target = w.back() * g / n_gpus;
          cut[g] = J.bgn_ref_id + (uint32_t)(std::lower_bound(w.begin(), w.end(), target) - w.begin());
          if (cut[g] > last_ref + 1 || g == n_gpus) cut[g] = last_ref + 1;
        }
        if (!on_all([&](int g) { const int i = g - g0; return cut[i + 1] > cut[i] ? oic_process_overlaps(sh[g].ctx, cut[i], cut[i + 1] - 1) : 0; })) return 1;
        for (int g = g0; g < g0 + ng; g++) {
          uint64_t k = oic_num_overlaps(sh[g].ctx), got = 0;
          rec.resize(6 * k);
          oic_copy_overlaps(sh[g].ctx, reinterpret_cast<oic_ovb_record_t *>(rec.data()), k, &got);
          out.write(rec.data(), got);
          oic_clear_overlaps(sh[g].ctx);
        }
      }
      bgn = (uint64_t)last[0] + 1;
      end = bgn + J.max_hash_strings - 1;
    }
    out.close();

    // ---- -s stats (OIC/overlapInCore.C:569-588)
    oic_stats_t tot, s;
    memset(&tot, 0, sizeof(tot));
    for (int g = g0; g < g0 + ng; g++) {
      oic_get_stats(sh[g].ctx, &s);
      tot.kmer_hits_without_olap += s.kmer_hits_without_olap; tot.kmer_hits_with_olap += s.kmer_hits_with_olap;
      tot.multi_overlap += s.multi_overlap; tot.total_overlaps += s.total_overlaps;
      tot.contained += s.contained; tot.dovetail += s.dovetail;
      tot.bad_short_window += s.bad_short_window; tot.bad_long_window += s.bad_long_window;
    }
    FILE *stats = stderr;
    if (job_stat) {
      stats = fopen(job_stat, "w");
      if (!stats) { fprintf(stderr, "WARNING: failed to open '%s' for writing: %s\n", job_stat, strerror(errno)); stats = stderr; }
    }
    fprintf(stats, " Kmer hits without olaps = %lld\n", (long long)tot.kmer_hits_without_olap);
    fprintf(stats, "    Kmer hits with olaps = %lld\n", (long long)tot.kmer_hits_with_olap);
    fprintf(stats, "  Multiple overlaps/pair = %lld\n", (long long)tot.multi_overlap);
    fprintf(stats, " Total overlaps produced = %lld\n", (long long)tot.total_overlaps);
    fprintf(stats, "      Contained overlaps = %lld\n", (long long)tot.contained);
    fprintf(stats, "       Dovetail overlaps = %lld\n", (long long)tot.dovetail);
    fprintf(stats, "Rejected by short window = %lld\n", (long long)tot.bad_short_window);
    fprintf(stats, " Rejected by long window = %lld\n", (long long)tot.bad_long_window);
    if (stats != stderr) fclose(stats);
    return 0;
  };

  int rcode = 0;
  if (!sweep_prefix) rcode = run_job(P, out_name, stat_name, 0, n_gpus);
  else {
    // --sweep: jobs are independent (own hash range, own output files), so with several GPUs every GPU takes WHOLE jobs
    // from a shared list, heaviest first (a job's work ~ bases of its query range): no cross-GPU barrier per job, a hash
    // block is built only by the GPU that needs it, and every GPU thread frames / writes its own .ovb files.
    std::vector<uint64_t> cum(n + 1, 0);
    for (uint32_t i = 0; i < n; i++) cum[i + 1] = cum[i] + lengths[i];
    std::vector<size_t> order(jobs.size());
    std::vector<uint64_t> work(jobs.size(), 0);
    for (size_t k = 0; k < jobs.size(); k++) {
      order[k] = k;
      const uint32_t a = std::max(first_id, std::max<uint32_t>(1, jobs[k].ref_bgn)), b = std::min(last_id, std::min(num_reads, jobs[k].ref_end));
      if (b >= a) work[k] = cum[b - first_id + 1] - cum[a - first_id];
    }
    if (n_gpus > 1) std::stable_sort(order.begin(), order.end(), [&](size_t x, size_t y) { return work[x] > work[y]; });
    std::atomic<size_t> next(0);
    std::atomic<int> failed(0);
    mkdir(sweep_out, 0777);
    // CUDA start-up of the devices of one box is serialised by the driver (measured on 8 B200: ~1 s per device, in threads or in
    // separate processes alike), so every worker brings up its own GPU and starts on the job list as soon as it is ready
    // instead of waiting for the last one.
    auto worker = [&](int g) {
      if (lazy_bring_up && (sh[g].rc = bring_up(g)) != 0) {
        fprintf(stderr, "ERROR (GPU %d): %s\n", device0 + g, sh[g].err.c_str());
        failed.store(sh[g].rc);
        return;
      }
      for (;;) {
        const size_t k = next.fetch_add(1);
        if (k >= order.size() || failed.load()) return;
        const OvlJob &j = jobs[order[k]];
        oic_params_t J = P;
        J.bgn_hash_id = std::max<uint32_t>(1, j.hash_bgn); J.end_hash_id = std::min(num_reads, j.hash_end);
        J.bgn_ref_id = std::max<uint32_t>(1, j.ref_bgn); J.end_ref_id = std::min(num_reads, j.ref_end);
        if (j.max_num_reads) { J.max_hash_strings = j.max_num_reads; J.max_hash_data_len = j.max_length; }
        char bat[16], job[16];
        snprintf(bat, sizeof(bat), "%03u", j.batch);
        snprintf(job, sizeof(job), "%06u", j.job);
        const std::string dir = std::string(sweep_out) + "/" + bat, base = dir + "/" + job;
        mkdir(dir.c_str(), 0777);
        const std::string working = base + ".ovb.WORKING", final_name = base + ".ovb", st = base + ".stats";
        fprintf(stderr, "Job %06u (GPU %d): %s  [%.2f s]\n", j.job, device0 + g, j.options().c_str(), elapsed());
        const int rc = run_job(J, working.c_str(), st.c_str(), g, 1);
        if (rc) { failed.store(rc); return; }
        rename(working.c_str(), final_name.c_str());   // the pipeline's success test is "<job>.ovb exists" (OverlapInCore.pm:220-225,405-429)
      }
    };
    if (n_gpus == 1) worker(0);
    else {
      std::vector<std::thread> th;
      for (int g = 0; g < n_gpus; g++) th.emplace_back(worker, g);
      for (auto &t : th) t.join();
    }
    rcode = failed.load();
  }
  for (int g = 0; g < n_gpus; g++) if (sh[g].ctx) oic_destroy(sh[g].ctx);
  if (rcode) return rcode;
  fprintf(stderr, "Bye.  [%.2f s]\n", elapsed());
  return 0;
}
