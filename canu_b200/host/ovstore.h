/*
 * ovstore.h -- the overlap store, the format downstream of the overlapInCore stage.
 *   reader of .ovb files (ovFileFull)                  ST/ovStoreFile.C:267-330
 *   writer of <asm>.ovlStore/{info,index,NNNN,evalueLen,overlapScores}
 *                                                      ST/ovStoreWriter.C:48-230 (sequential store), ST/ovStore.H:53-205,
 *                                                      ST/ovStoreHistogram.C:130-345
 * Host-only C++; the sort that feeds the writer is oic_ovlstore_sort (CUDA).
 */
#ifndef OIC_HOST_OVSTORE_H
#define OIC_HOST_OVSTORE_H
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "stores.h"

namespace oichost {

/* appends the 6-word records of an .ovb file (blocks of size_t length + snappy raw data); exits on a malformed file */
void read_ovb_file(const std::string &name, std::vector<uint32_t> &records6);

/* per-read inputs of the store filter (ST/ovStoreFilter.C:36-75), indexed by read ID, numReads + 1 entries */
struct StoreReadInfo {
  std::vector<uint32_t> length;
  std::vector<uint8_t> skip_dup;   // library does not remove duplicate reads
  std::vector<uint8_t> library;
};
StoreReadInfo store_read_info(const GkStore &gkp, const std::string &gkp_path);

/* Sequential overlap store writer: feed it the records in ovOverlap::operator< order. */
class OvlStoreWriter {
 public:
  OvlStoreWriter(const std::string &path, const GkStore &gkp);   // exits if `path` already is a complete store
  void write(const uint32_t *records6, uint64_t n);
  void finish();                                                  // index tail, info with the real magic, evalueLen, overlapScores
  uint64_t num_overlaps() const { return info_[5]; }
  ~OvlStoreWriter();

 private:
  struct Offt { uint32_t a_iid, fileno, offset, num_olaps; uint64_t overlap_id; };   // ovStoreOfft, 24 bytes on disk
  void save_info(bool complete);
  void process_scores(uint32_t aid);
  std::string path_;
  const GkStore &gkp_;
  uint64_t info_[8];               // magic, version, unused, smallestIID, largestIID, numOverlapsTotal, highestFileIndex, maxReadLenInBits
  FILE *index_ = nullptr, *data_ = nullptr;
  std::vector<uint32_t> dbuf_;     // 20-byte records of the current data file not yet written (one fwrite per 4 MB, not per record)
  void flush_data() { if (data_ && !dbuf_.empty()) fwrite(dbuf_.data(), 4, dbuf_.size(), data_); dbuf_.clear(); }
  Offt offt_, offm_;
  uint64_t this_file_ = 0, this_file_max_ = 0;
  uint32_t file_index_ = 0;
  bool finished_ = false;
  // overlaps per (evalue, length / 250)
  uint32_t opel_len_ = 0, max_olength_ = 0, max_evalue_ = 0;
  std::vector<std::vector<uint32_t>> opel_;
  // overlap scores per read
  struct Sco { uint16_t points[16], scores[16]; };
  std::vector<uint16_t> score_list_;
  uint32_t score_aid_ = 0, scores_bgn_ = 0, scores_len_ = 0;
  std::vector<Sco> scores_;
};

}  // namespace oichost
#endif
