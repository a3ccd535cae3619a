/*
 * overlapInCorePartition.cpp -- drop-in for the reference's overlapInCorePartition (OIC/overlapInCorePartition.C:380-511):
 * same options, same three output files (<prefix>.ovlbat / .ovljob / .ovlopt).  The partitioning itself is partition.cpp;
 * the same function drives the in-process sweep of canu_b200/bin/overlapInCore (--sweep).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "partition.h"
#include "stores.h"

using namespace oichost;

// AS_UTL_decodeRange into a set (UTL/AS_UTL_decodeRange.C:40-100): "a", "a-b", comma separated
static void decode_set(const char *s, std::set<uint32_t> &out) {
  while (*s) {
    char *e;
    unsigned long a = strtoul(s, &e, 10), b = a;
    if (e == s) { fprintf(stderr, "ERROR: invalid range '%s'\n", s); exit(1); }
    if (*e == '-') { s = e + 1; b = strtoul(s, &e, 10); }
    for (unsigned long v = a; v <= b; v++) out.insert((uint32_t)v);
    s = (*e == ',') ? e + 1 : e;
    if (*e && *e != ',') break;
  }
}

int main(int argc, char **argv) {
  const char *store_name = nullptr, *prefix = nullptr;
  PartitionParams p;
  bool check_all_lib_used = true;
  int err = 0;
  for (int arg = 1; arg < argc; arg++) {
    const char *a = argv[arg];
    auto need = [&]() -> const char * { if (arg + 1 >= argc) { fprintf(stderr, "ERROR: option '%s' needs a value\n", a); exit(1); } return argv[++arg]; };
    if (strcmp(a, "-g") == 0) store_name = need();
    else if (strcmp(a, "-bl") == 0) p.hash_block_length = strtoull(need(), nullptr, 10);
    else if (strcmp(a, "-bs") == 0) p.hash_block_size = strtoull(need(), nullptr, 10);
    else if (strcmp(a, "-rl") == 0) p.ref_block_length = strtoull(need(), nullptr, 10);
    else if (strcmp(a, "-rs") == 0) p.ref_block_size = strtoull(need(), nullptr, 10);
    else if (strcmp(a, "-ol") == 0) p.min_overlap_length = (uint32_t)strtoull(need(), nullptr, 10);
    else if (strcmp(a, "-H") == 0) decode_set(need(), p.lib_to_hash);
    else if (strcmp(a, "-R") == 0) decode_set(need(), p.lib_to_ref);
    else if (strcmp(a, "-C") == 0) check_all_lib_used = false;
    else if (strcmp(a, "-o") == 0) prefix = need();
    else { fprintf(stderr, "ERROR:  Unknown option '%s'\n", a); err++; }
  }
  if (err || !store_name || !prefix) { fprintf(stderr, "usage: %s [opts]\n", argv[0]); return 1; }
  if (p.hash_block_length > 0 && p.hash_block_size > 0) { fprintf(stderr, "ERROR:  At most one of -bl and -bs can be non-zero.\n"); return 1; }
  if (p.ref_block_length > 0 && p.ref_block_size > 0) { fprintf(stderr, "ERROR:  At most one of -rl and -rs can be non-zero.\n"); return 1; }
  fprintf(stderr, "HASH: %llu reads or %llu length.\n", (unsigned long long)p.hash_block_size, (unsigned long long)p.hash_block_length);
  fprintf(stderr, "REF:  %llu reads or %llu length.\n", (unsigned long long)p.ref_block_size, (unsigned long long)p.ref_block_length);

  GkStore store(store_name);
  const uint32_t num_libs = store.num_libraries();
  uint32_t invalid = 0;
  for (uint32_t l : p.lib_to_hash) if (num_libs < l) { fprintf(stderr, "ERROR: -H %u is invalid; only %u libraries in '%s'\n", l, num_libs, store_name); invalid++; }
  for (uint32_t l : p.lib_to_ref) if (num_libs < l) { fprintf(stderr, "ERROR: -R %u is invalid; only %u libraries in '%s'\n", l, num_libs, store_name); invalid++; }
  if (!p.lib_to_hash.empty() && !p.lib_to_ref.empty())
    for (uint32_t lib = 1; lib <= num_libs; lib++)
      if (!p.lib_to_hash.count(lib) && !p.lib_to_ref.count(lib)) {
        if (check_all_lib_used) { fprintf(stderr, "ERROR: library %u is not mentioned in either -H or -R.\n", lib); invalid++; }
        else fprintf(stderr, "Warning: library %u is not mentioned in either -H or -R.\n", lib);
      }
  if (invalid) { fprintf(stderr, "ERROR: one of -H and/or -R are invalid.\n"); return 1; }

  const uint32_t n = store.num_reads();
  std::vector<uint32_t> len(n + 2, 0), lib(n + 2, 0);
  for (uint32_t i = 1; i <= n; i++) { len[i] = store.read(i).seq_len(); lib[i] = store.read(i).library_id(); }
  const std::vector<OvlJob> jobs = partition_jobs(p, n, len, lib);
  for (const OvlJob &j : jobs) {
    if (j.max_num_reads == 0) fprintf(stderr, "HASH %10d-%10d  REFR %10d-%10d JOB %d\n", j.hash_bgn, j.hash_end, j.ref_bgn, j.ref_end, j.job);
    else fprintf(stderr, "HASH %10d-%10d  REFR %10d-%10d  STRINGS %10d  BASES %10d JOB %d\n", j.hash_bgn, j.hash_end, j.ref_bgn, j.ref_end, j.max_num_reads, j.max_length, j.job);
  }
  return write_job_files(prefix, jobs) ? 0 : 1;
}
