/*
 * ovStoreBuild.cpp -- drop-in for the reference's ovStoreBuild (ST/ovStoreBuild.C:348-660), the step that follows the
 * overlapInCore jobs in the pipeline (src/pipelines/canu/OverlapStore.pm): `ovStoreBuild -O asm.ovlStore -G asm.gkpStore
 * [-e erate] [-L fileList | *.ovb]`.  The reference buckets the overlaps into temporary files so that each bucket fits its
 * memory limit (-M / -F) and std::sorts bucket by bucket on one core; here the filter, the mirroring and the sort run on the
 * GPU in one piece (oic_ovlstore_sort, canu_b200/csrc/oic_store.cuh) and this file is argv, the .ovb reader and the store
 * writer.  -M / -F are accepted and ignored (no temporary buckets exist); -evalues and -config are not part of this build.
 */
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <chrono>
#include <string>
#include <atomic>
#include <thread>
#include <vector>

#include "../../include/oic_b200.h"
#include "ovstore.h"
#include "stores.h"

using namespace oichost;

static bool file_exists(const char *p) {
  struct stat st;
  return stat(p, &st) == 0 && !S_ISDIR(st.st_mode);
}

int main(int argc, char **argv) {
  const char *ovl_name = nullptr, *gkp_name = nullptr;
  double max_error = 1.0;
  int device = 0;
  std::vector<std::string> files;
  int err = 0;
  for (int arg = 1; arg < argc; arg++) {
    const char *a = argv[arg];
    auto need = [&]() -> const char * {
      if (arg + 1 >= argc) { fprintf(stderr, "%s: option '%s' needs a value\n", argv[0], a); exit(1); }
      return argv[++arg];
    };
    if (strcmp(a, "-O") == 0) ovl_name = need();
    else if (strcmp(a, "-G") == 0) gkp_name = need();
    else if (strcmp(a, "-F") == 0 || strcmp(a, "-M") == 0) need();   // memory / file limits of the bucketed CPU build: nothing to limit here
    else if (strcmp(a, "-e") == 0) max_error = atof(need());
    else if (strcmp(a, "-L") == 0) {
      const char *ln = need();
      FILE *f = fopen(ln, "r");
      if (!f) { fprintf(stderr, "%s: failed to open file list '%s': %s\n", argv[0], ln, strerror(errno)); exit(1); }
      char line[4096];
      while (fgets(line, sizeof(line), f)) {
        size_t n = strlen(line);
        while (n && (line[n - 1] == '\n' || line[n - 1] == '\r' || line[n - 1] == ' ')) line[--n] = 0;
        if (n) files.push_back(line);
      }
      fclose(f);
    } else if (strcmp(a, "--device") == 0) device = atoi(need());
    else if (strcmp(a, "-evalues") == 0 || strcmp(a, "-config") == 0) {
      fprintf(stderr, "%s: option '%s' is not supported by this build\n", argv[0], a);
      err++;
    } else if (file_exists(a)) files.push_back(a);
    else { fprintf(stderr, "%s: unknown option '%s'.\n", argv[0], a); err++; }
  }
  if (!ovl_name || !gkp_name || files.empty() || err) {
    fprintf(stderr, "usage: %s -O asm.ovlStore -G asm.gkpStore [opts] [-L fileList | *.ovb]\n", argv[0]);
    fprintf(stderr, "  -O asm.ovlStore       path to store to create\n");
    fprintf(stderr, "  -G asm.gkpStore       path to gkpStore for this assembly\n");
    fprintf(stderr, "  -L fileList           read input filenames from 'fileList'\n");
    fprintf(stderr, "  -e e                  filter overlaps above e fraction error\n");
    fprintf(stderr, "  -F f / -M g           accepted for compatibility (the sort runs on the GPU in one piece)\n");
    fprintf(stderr, "  --device d            CUDA device (default 0)\n");
    if (!ovl_name) fprintf(stderr, "ERROR: No overlap store (-O) supplied.\n");
    if (!gkp_name) fprintf(stderr, "ERROR: No gatekeeper store (-G) supplied.\n");
    if (files.empty()) fprintf(stderr, "ERROR: No input overlap files (-L or last on the command line) supplied.\n");
    return 1;
  }
  const auto t0 = std::chrono::steady_clock::now();
  auto secs = [&]() { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };

  // the GPU first: there is no CPU path, so fail before any file is touched
  oic_params_t P;
  memset(&P, 0, sizeof(P));
  P.kmer_len = 22; P.max_erate = 0.06; P.unique_olap = 1; P.hash_bits = 22; P.hash_load = 0.6; P.max_hash_strings = 1; P.max_hash_data_len = 1;
  P.frag_olap_limit = UINT64_MAX; P.num_threads = 1; P.bgn_hash_id = P.bgn_ref_id = 1; P.end_hash_id = P.end_ref_id = UINT32_MAX;
  oic_ctx_t *ctx = nullptr;
  if (oic_create(&P, device, &ctx) != 0) { fprintf(stderr, "ERROR: %s\n", oic_last_error(nullptr)); return 1; }

  GkStore gkp(gkp_name);
  const StoreReadInfo ri = store_read_info(gkp, gkp_name);
  const uint32_t max_id = gkp.num_reads() + 1;
  const uint32_t max_evalue = (max_error < 4095 / 10000.0) ? (uint32_t)(int)(10000.0 * max_error + 0.5) : 4095;   // AS_OVS_encodeEvalue

  OvlStoreWriter store(ovl_name, gkp);
  std::vector<uint32_t> rec;
  fprintf(stderr, "\n-- LOADING --\n\n");
  {   // the files are independent (framed, snappy blocks): decompress them on all host cores, then concatenate in command-line order
    std::vector<std::vector<uint32_t>> part(files.size());
    std::atomic<size_t> next{0};
    const unsigned nt = std::max(1u, std::min<unsigned>((unsigned)files.size(), std::thread::hardware_concurrency()));
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
      th.emplace_back([&]() { for (size_t i; (i = next.fetch_add(1)) < files.size();) read_ovb_file(files[i], part[i]); });
    for (std::thread &t : th) t.join();
    size_t total = 0;
    for (const auto &v : part) total += v.size();
    rec.resize(total);
    size_t at = 0;
    for (size_t i = 0; i < files.size(); i++) {
      fprintf(stderr, "-  Loading '%s'\n", files[i].c_str());
      if (!part[i].empty()) memcpy(rec.data() + at, part[i].data(), part[i].size() * 4);
      at += part[i].size();
      std::vector<uint32_t>().swap(part[i]);
    }
  }
  const uint64_t n = rec.size() / 6;
  fprintf(stderr, "-  %llu overlaps loaded  [%.2f s]\n\n-- FILTERING AND SORTING (GPU) --\n\n", (unsigned long long)n, secs());
  std::vector<uint32_t> sorted(12 * n + 6);
  uint64_t m = 0, counts[9];
  double ms = 0;
  if (oic_ovlstore_sort(ctx, reinterpret_cast<const oic_ovb_record_t *>(rec.data()), n, max_id, ri.length.data(), ri.skip_dup.data(), ri.library.data(),
                        max_evalue, reinterpret_cast<oic_ovb_record_t *>(sorted.data()), &m, counts, &ms) != 0) {
    fprintf(stderr, "ERROR: %s\n", oic_last_error(ctx));
    return 1;
  }
  if (counts[8] > 0) fprintf(stderr, "-- Saved      %llu dedupe overlaps\n", (unsigned long long)counts[8]);
  if (counts[7] > 0) fprintf(stderr, "-- Saved      %llu trimming overlaps\n-- Discarded  %llu too similar %llu too short\n", (unsigned long long)counts[7],
                             (unsigned long long)counts[2], (unsigned long long)counts[3]);
  if (counts[6] > 0) fprintf(stderr, "-- Saved      %llu unitigging overlaps\n", (unsigned long long)counts[6]);
  if (counts[1] > 0) fprintf(stderr, "-- Discarded  %llu low quality, more than %.4f fraction error\n", (unsigned long long)counts[1], max_error);
  fprintf(stderr, "-  %llu store records sorted in %.2f ms on the GPU  [%.2f s]\n\n-- WRITING --\n\n", (unsigned long long)m, ms, secs());
  store.write(sorted.data(), m);
  store.finish();
  oic_destroy(ctx);
  fprintf(stderr, "-  done  [%.2f s]\n", secs());
  return 0;
}
