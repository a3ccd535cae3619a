/*
 * frequentMers.cpp -- the frequent K-mer list of a gkpStore, counted on the GPU (liboic_b200.so: oic_count_kmers).
 *
 * In the pipeline the `-k <file>` skip list of overlapInCore is the output of two meryl runs (src/pipelines/canu/Meryl.pm:
 * 434-450 `meryl -B -C -L 2 -v -m K -threads t -memory m -s <gkpStore> -o <db>`, then :634 `meryl -Dt -n T -s <db> > file`).
 * This tool writes the same bytes as that second command -- ">count\nKMER\n" for every canonical K-mer seen >= T times in
 * the reads of the store, upper case, ascending -- straight from the store:
 *
 *     frequentMers -m K -n T -s <gkpStore> [-o <file>] [--histogram <file> [--info <file>]] [--device d]
 *
 * (-o defaults to stdout, like the shell redirection of the pipeline.)  --histogram writes what `meryl -Dh -s <db>` prints on
 * stdout (count, distinct K-mers with that count, cumulative fraction of distinct, cumulative fraction of all; meryl-dump.C:
 * 120-152 -- the input of estimate-mer-threshold, Meryl.pm:449-470) and --info its four stderr lines; with --histogram, -n may be
 * omitted (no list is written).  No CPU path: without a CUDA device it exits 1.
 */
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/oic_b200.h"
#include "stores.h"

using namespace oichost;

int main(int argc, char **argv) {
  uint32_t K = 0, T = 0;
  const char *store_path = nullptr, *out_name = nullptr, *hist_name = nullptr, *info_name = nullptr;
  int device = 0;
  for (int i = 1; i < argc; i++) {
    const char *a = argv[i];
    auto need = [&]() -> const char * {
      if (i + 1 >= argc) { fprintf(stderr, "ERROR: option '%s' needs a value\n", a); exit(1); }
      return argv[++i];
    };
    if (!strcmp(a, "-m")) K = (uint32_t)strtoul(need(), nullptr, 10);
    else if (!strcmp(a, "-n")) T = (uint32_t)strtoul(need(), nullptr, 10);
    else if (!strcmp(a, "-s")) store_path = need();
    else if (!strcmp(a, "-o")) out_name = need();
    else if (!strcmp(a, "--histogram")) hist_name = need();
    else if (!strcmp(a, "--info")) info_name = need();
    else if (!strcmp(a, "--device")) device = atoi(need());
    else { fprintf(stderr, "Unknown option '%s'\n", a); return 1; }
  }
  if (K < 2 || K > 31 || (T < 1 && !hist_name) || !store_path) {
    fprintf(stderr, "usage: %s -m <mersize 2..31> -n <min count> -s <gkpStore> [-o file] [--histogram file [--info file]] [--device d]\n", argv[0]);
    return 1;
  }

  GkStore store(store_path);
  const uint32_t n = store.num_reads();
  if (n == 0) { fprintf(stderr, "ERROR: empty store\n"); return 1; }
  std::vector<uint64_t> offsets(n);
  std::vector<uint32_t> lengths(n);
  bool two_bit = true;
  uint64_t total = 0;
  for (uint32_t i = 0; i < n; i++) {
    lengths[i] = store.read(i + 1).seq_len();
    total += lengths[i];
    if (lengths[i] && !store.sequence_2bit(i + 1)) two_bit = false;
  }
  std::vector<char> bases;
  uint64_t at = 0;
  for (uint32_t i = 0; i < n; i++) { offsets[i] = at; at += two_bit ? ((uint64_t)lengths[i] + 3) / 4 : lengths[i]; }
  bases.resize(at + 8);
  const unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  {
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
      th.emplace_back([&, t]() {
        for (uint64_t i = (uint64_t)n * t / nt, e = (uint64_t)n * (t + 1) / nt; i < e; i++) {
          if (!lengths[i]) continue;
          if (two_bit) memcpy(bases.data() + offsets[i], store.sequence_2bit((uint32_t)i + 1), ((uint64_t)lengths[i] + 3) / 4);
          else store.load_sequence((uint32_t)i + 1, bases.data() + offsets[i]);
        }
      });
    for (auto &t : th) t.join();
  }

  oic_params_t P;
  memset(&P, 0, sizeof(P));
  P.kmer_len = (int32_t)K; P.max_erate = 0.06; P.bgn_hash_id = P.bgn_ref_id = 1; P.end_hash_id = P.end_ref_id = n;
  P.frag_olap_limit = UINT64_MAX; P.unique_olap = 1; P.hash_bits = 22; P.hash_load = 0.6; P.max_hash_strings = n;
  P.max_hash_data_len = total + 1; P.num_threads = 1; P.use_hopeless_check = 1;
  oic_ctx_t *ctx = nullptr;
  if (oic_create(&P, device, &ctx)) { fprintf(stderr, "ERROR: %s\n", oic_last_error(nullptr)); return 1; }
  int rc = two_bit ? oic_load_reads_2bit(ctx, 1, n, offsets.data(), lengths.data(), (const uint8_t *)bases.data(), nullptr, nullptr)
                   : oic_load_reads(ctx, 1, n, offsets.data(), lengths.data(), bases.data(), nullptr, nullptr);
  uint64_t found = 0;
  uint32_t passes = 0;
  double ms = 0;
  const bool want_list = T >= 1;
  if (!rc) rc = oic_count_kmers(ctx, K, want_list ? T : UINT32_MAX, &found, &passes, &ms);
  if (rc) { fprintf(stderr, "ERROR: %s\n", oic_last_error(ctx)); return 1; }
  std::vector<uint64_t> keys(found);
  std::vector<uint32_t> counts(found);
  uint64_t got = 0;
  oic_copy_frequent_kmers(ctx, keys.data(), counts.data(), found, &got);
  fprintf(stderr, "%llu reads, %llu bases: %llu canonical %u-mers seen >= %u times (%u pass%s, %.1f ms on the GPU)\n", (unsigned long long)n,
          (unsigned long long)total, (unsigned long long)got, K, T, passes, passes == 1 ? "" : "es", ms);

  if (hist_name) {
    uint64_t hl = 0;
    oic_copy_kmer_histogram(ctx, nullptr, 0, &hl);
    std::vector<uint64_t> hist(hl);
    oic_copy_kmer_histogram(ctx, hist.data(), hl, &hl);
    uint64_t n_distinct = 0, n_total = 0;
    for (uint64_t i = 1; i < hl; i++) { n_distinct += hist[i]; n_total += hist[i] * i; }
    FILE *hf = fopen(hist_name, "w");
    if (!hf) { fprintf(stderr, "ERROR: cannot write '%s': %s\n", hist_name, strerror(errno)); return 1; }
    uint64_t distinct = 0, tot = 0;
    for (uint64_t i = 1; i < hl; i++) {
      if (hist[i] == 0) continue;
      distinct += hist[i];
      tot += hist[i] * i;
      fprintf(hf, "%u\t%llu\t%.4f\t%.4f\n", (unsigned)i, (unsigned long long)hist[i], distinct / (double)n_distinct, tot / (double)n_total);
    }
    fclose(hf);
    FILE *inf = info_name ? fopen(info_name, "w") : stderr;
    if (!inf) { fprintf(stderr, "ERROR: cannot write '%s': %s\n", info_name, strerror(errno)); return 1; }
    fprintf(inf, "Found %llu mers.\nFound %llu distinct mers.\nFound %llu unique mers.\nLargest mercount is %llu.\n", (unsigned long long)n_total,
            (unsigned long long)n_distinct, (unsigned long long)(hl > 1 ? hist[1] : 0), (unsigned long long)(hl ? hl - 1 : 0));
    if (inf != stderr) fclose(inf);
  }
  if (!want_list) { oic_destroy(ctx); return 0; }

  FILE *out = stdout;
  if (out_name) {
    out = fopen(out_name, "w");
    if (!out) { fprintf(stderr, "ERROR: cannot write '%s': %s\n", out_name, strerror(errno)); return 1; }
  }
  std::string line;
  char mer[40];
  mer[K] = 0;
  for (uint64_t i = 0; i < got; i++) {
    for (uint32_t j = 0; j < K; j++) mer[j] = "ACGT"[(keys[i] >> (2 * (K - 1 - j))) & 3];
    fprintf(out, ">%u\n%s\n", counts[i], mer);
  }
  if (out != stdout) fclose(out);
  oic_destroy(ctx);
  return 0;
}
