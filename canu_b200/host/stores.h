/*
 * stores.h -- the two on-disk formats either side of the overlapInCore stage:
 *   input : gkpStore directory (reader only)        ST/gkStore.H:333-401, ST/gkStore.C:49-156,592-727, ST/gkStoreEncode.C:81-114
 *   output: .ovb overlap file, .counts, -s stats    ST/ovStoreFile.C:67-76,137-224, ST/ovStoreHistogram.C:226-241,322-339,
 *                                                   AS_UTL/AS_UTL_fileIO.C:56-68, OIC/overlapInCore.C:569-588
 * Host-only C++ (no CUDA); written from the format description, not from the reference sources.
 */
#ifndef OIC_HOST_STORES_H
#define OIC_HOST_STORES_H
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

namespace oichost {

struct GkRead {           // 16 bytes: word0 = readID:37 | libID:6 | seqLen:21 (from the LSB); word1 = mPtr:48 | pID:16
  uint64_t w0, w1;
  uint64_t read_id() const { return w0 & ((1ull << 37) - 1); }
  uint32_t library_id() const { return (uint32_t)((w0 >> 37) & 63); }
  uint32_t seq_len() const { return (uint32_t)((w0 >> 43) & ((1u << 21) - 1)); }
  uint64_t m_ptr() const { return w1 & ((1ull << 48) - 1); }
};

class GkStore {
 public:
  // exits the process with a message on any format error, like the reference (ST/gkStore.C:592-727)
  explicit GkStore(const std::string &path);
  ~GkStore();
  uint32_t num_reads() const { return num_reads_; }
  uint32_t num_libraries() const { return num_libraries_; }
  const GkRead &read(uint32_t id) const { return reads_[id]; }
  // sequence of read `id` as upper-case ASCII into out[0 .. seq_len) (gkRead_loadData: 2SEQ / USEQ chunks)
  void load_sequence(uint32_t id, char *out) const;
  // the read's `2SEQ` chunk as stored (4 bases per byte, first base in bits 7-6; ceil(seq_len/4) bytes), or nullptr when the
  // read is kept unencoded (`USEQ`: it contains non-ACGT bases) -- fed to oic_load_reads_2bit without decoding
  const uint8_t *sequence_2bit(uint32_t id) const;
  // quality values of read `id` into out[0 .. seq_len), as gkReadData_getQualities returns them (ST/gkStore.C:117-145): the `QVAL`
  // constant for every base, or the `UQLT` bytes; a read with neither gets zeros (what the reference's untouched buffer holds)
  void load_qualities(uint32_t id, uint8_t *out) const;

 private:
  std::string path_;
  uint32_t num_reads_ = 0, num_libraries_ = 0;
  const GkRead *reads_ = nullptr;
  size_t reads_bytes_ = 0;
  const uint8_t *blobs_ = nullptr;
  size_t blobs_bytes_ = 0;
};

/* ovFile in ovFileFullWrite mode: blocks of <= 262080 uint32 (whole 6-word records), each written as
 * size_t compressedLength + a snappy raw block (snappy_lite.h: a greedy LZ77 compressor of the raw format; ovStoreBuild /
 * overlapConvert read the blocks with the stock snappy decoder). */
class OvbWriter {
 public:
  explicit OvbWriter(const std::string &name);   // exits on failure
  void write(const uint32_t *records6, uint64_t n_records);
  void close();                                   // flushes and writes <base>.counts
  ~OvbWriter();

 private:
  void flush();
  std::string name_;
  FILE *f_ = nullptr;
  std::vector<uint32_t> buf_;
  std::vector<uint32_t> opr_;   // overlaps per read
  uint32_t opr_len_ = 0;
  static const size_t kBufferMax = 262080;
};

std::string counts_name_for(const std::string &ovb_name);

}  // namespace oichost
#endif
