/*
 * partition.h -- the job list of the overlap stage: which (hash ID range, query ID range) pairs one invocation of
 * overlapInCore handles.  Restates the reference's overlapInCorePartition (OIC/overlapInCorePartition.C:55-93 outputJob,
 * :169-246 partitionFrags, :252-352 partitionLength): overlapInCore only computes overlaps for queryID < hashID, so for every
 * hash block the query blocks cover IDs 1 .. end of the hash block (the triangular sweep).  Host-only C++.
 */
#ifndef OIC_HOST_PARTITION_H
#define OIC_HOST_PARTITION_H
#include <stdint.h>

#include <set>
#include <string>
#include <vector>

namespace oichost {

struct OvlJob {
  uint32_t hash_bgn, hash_end, ref_bgn, ref_end;
  uint32_t max_num_reads, max_length;   // --hashstrings / --hashdatalen of the job (0, 0: not passed)
  uint32_t batch, job;                  // names: %03u / %06u
  std::string options() const;          // the line of the .ovlopt file
};

struct PartitionParams {
  uint32_t min_overlap_length = 0;      // -ol
  uint64_t hash_block_length = 0;       // -bl  (bases per hash block; 0: use hash_block_size)
  uint64_t hash_block_size = 0;         // -bs  (reads per hash block)
  uint64_t ref_block_length = 0;        // -rl
  uint64_t ref_block_size = 0;          // -rs
  std::set<uint32_t> lib_to_hash, lib_to_ref;   // -H / -R
};

/* read_len[1 .. num_reads], read_lib[1 .. num_reads] (index 0 unused) */
std::vector<OvlJob> partition_jobs(const PartitionParams &p, uint32_t num_reads, const std::vector<uint32_t> &read_len,
                                   const std::vector<uint32_t> &read_lib);

/* <prefix>.ovlbat / .ovljob / .ovlopt, written as .WORKING and renamed (OIC/overlapInCorePartition.C:355-377,491-506) */
bool write_job_files(const std::string &prefix, const std::vector<OvlJob> &jobs);
/* the inverse: parses the three files */
bool read_job_files(const std::string &prefix, std::vector<OvlJob> &jobs);

}  // namespace oichost
#endif
