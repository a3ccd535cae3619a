// partition.cpp -- see partition.h for the reference citations.
#include "partition.h"

#include <stdio.h>
#include <string.h>

namespace oichost {

namespace {
const uint32_t kBatchMax = 1000;   // jobs per batch directory (OIC/overlapInCorePartition.C:53)

struct Names {
  uint32_t batch_size = 0, batch = 1, job = 1;
  void next() {   // outputJob's bookkeeping (:84-92)
    batch_size++;
    if (batch_size >= kBatchMax) { batch_size = 0; batch++; }
    job++;
  }
};

// The query blocks of one hash block (the inner while of both partitioners, :205-238 / :311-345).
void ref_blocks(const PartitionParams &p, const std::vector<uint32_t> &read_len, uint32_t hash_bgn, uint32_t hash_end,
                uint32_t ref_min, uint32_t ref_max, uint32_t max_num_reads, uint32_t max_length, Names &nm, std::vector<OvlJob> &out) {
  const bool same_libs = !p.lib_to_hash.empty() && p.lib_to_hash == p.lib_to_ref;
  const bool clip_to_hash = p.lib_to_hash.empty() || p.lib_to_hash != p.lib_to_ref;
  uint32_t ref_bgn = ref_min, ref_end = 0;
  while (ref_bgn < ref_max && (ref_bgn < hash_end || same_libs)) {
    if (p.ref_block_length > 0) {
      uint64_t ref_len = 0;
      do {
        ref_end++;
        if (read_len[ref_end] < p.min_overlap_length) continue;
        ref_len += read_len[ref_end];
      } while (ref_len < p.ref_block_length && ref_end < ref_max);
    } else {
      ref_end = ref_bgn + (uint32_t)p.ref_block_size - 1;
    }
    if (ref_end > ref_max) ref_end = ref_max;
    if (ref_end > hash_end && clip_to_hash) ref_end = hash_end;
    out.push_back(OvlJob{hash_bgn, hash_end, ref_bgn, ref_end, max_num_reads, max_length, nm.batch, nm.job});
    nm.next();
    ref_bgn = ref_end + 1;
  }
}
}  // namespace

std::string OvlJob::options() const {
  char b[256];
  if (max_num_reads == 0) snprintf(b, sizeof(b), "-h %u-%u -r %u-%u", hash_bgn, hash_end, ref_bgn, ref_end);
  else snprintf(b, sizeof(b), "-h %u-%u -r %u-%u --hashstrings %u --hashdatalen %u", hash_bgn, hash_end, ref_bgn, ref_end, max_num_reads, max_length);
  return b;
}

std::vector<OvlJob> partition_jobs(const PartitionParams &p, uint32_t num_reads, const std::vector<uint32_t> &read_len,
                                   const std::vector<uint32_t> &read_lib) {
  std::vector<OvlJob> jobs;
  uint32_t hash_min = 1, hash_max = UINT32_MAX, ref_min = 1, ref_max = UINT32_MAX;
  // loadReadLengths (:95-165): the ID range of the reads whose library is listed in -H / -R
  if (!p.lib_to_hash.empty()) {
    hash_min = UINT32_MAX; hash_max = 0;
    for (uint32_t i = 1; i <= num_reads; i++)
      if (p.lib_to_hash.count(read_lib[i])) { if (i < hash_min) hash_min = i; if (hash_max < i) hash_max = i; }
  }
  if (!p.lib_to_ref.empty()) {
    ref_min = UINT32_MAX; ref_max = 0;
    for (uint32_t i = 1; i <= num_reads; i++)
      if (p.lib_to_ref.count(read_lib[i])) { if (i < ref_min) ref_min = i; if (ref_max < i) ref_max = i; }
  }
  if (hash_max > num_reads) hash_max = num_reads;
  if (ref_max > num_reads) ref_max = num_reads;

  Names nm;
  uint32_t hash_bgn = hash_min;
  if (p.hash_block_length == 0) {            // partitionFrags: hash blocks of a fixed number of reads (:169-246)
    while (hash_bgn < hash_max) {
      uint32_t hash_end = hash_bgn + (uint32_t)p.hash_block_size - 1;
      if (hash_end > hash_max) hash_end = hash_max;
      ref_blocks(p, read_len, hash_bgn, hash_end, ref_min, ref_max, 0, 0, nm, jobs);
      hash_bgn = hash_end + 1;
    }
  } else {                                   // partitionLength: hash blocks of a fixed number of bases (:252-352)
    uint32_t hash_end = hash_min - 1;
    while (hash_bgn < hash_max) {
      uint64_t hash_len = 0;
      do {   // a loaded read costs its bases plus the terminating zero (:297-306)
        hash_end++;
        if (read_len[hash_end] < p.min_overlap_length) continue;
        hash_len += (uint64_t)read_len[hash_end] + 1;
      } while (hash_len < p.hash_block_length && hash_end < hash_max);
      ref_blocks(p, read_len, hash_bgn, hash_end, ref_min, ref_max, hash_end - hash_bgn + 1, (uint32_t)hash_len, nm, jobs);
      hash_bgn = hash_end + 1;
    }
  }
  return jobs;
}

bool write_job_files(const std::string &prefix, const std::vector<OvlJob> &jobs) {
  const char *types[3] = {"ovlbat", "ovljob", "ovlopt"};
  FILE *f[3];
  for (int i = 0; i < 3; i++) {
    f[i] = fopen((prefix + "." + types[i] + ".WORKING").c_str(), "w");
    if (!f[i]) { fprintf(stderr, "Failed to open '%s.%s.WORKING'\n", prefix.c_str(), types[i]); return false; }
  }
  for (const OvlJob &j : jobs) {
    fprintf(f[0], "%03u\n", j.batch);
    fprintf(f[1], "%06u\n", j.job);
    fprintf(f[2], "%s\n", j.options().c_str());
  }
  for (int i = 0; i < 3; i++) {
    fclose(f[i]);
    rename((prefix + "." + types[i] + ".WORKING").c_str(), (prefix + "." + types[i]).c_str());
  }
  return true;
}

bool read_job_files(const std::string &prefix, std::vector<OvlJob> &jobs) {
  FILE *b = fopen((prefix + ".ovlbat").c_str(), "r"), *j = fopen((prefix + ".ovljob").c_str(), "r"), *o = fopen((prefix + ".ovlopt").c_str(), "r");
  if (!b || !j || !o) { if (b) fclose(b); if (j) fclose(j); if (o) fclose(o); return false; }
  char line[512];
  jobs.clear();
  bool ok = true;
  while (fgets(line, sizeof(line), o)) {
    OvlJob x = {};
    int n = sscanf(line, "-h %u-%u -r %u-%u --hashstrings %u --hashdatalen %u", &x.hash_bgn, &x.hash_end, &x.ref_bgn, &x.ref_end, &x.max_num_reads, &x.max_length);
    if (n != 4 && n != 6) { ok = false; break; }
    if (fscanf(b, "%u", &x.batch) != 1 || fscanf(j, "%u", &x.job) != 1) { ok = false; break; }
    jobs.push_back(x);
  }
  fclose(b); fclose(j); fclose(o);
  return ok;
}

}  // namespace oichost
