// Test harness for canu_b200/host/stores.cpp (no CUDA): exported with C linkage for ctypes.
#include <string.h>

#include "../canu_b200/host/ovstore.h"
#include "../canu_b200/host/snappy_lite.h"
#include "../canu_b200/host/stores.h"

using namespace oichost;

extern "C" {

// number of reads; lengths[] / libs[] (size num_reads, may be null) are filled for IDs 1..num_reads
uint32_t sh_store_info(const char *path, uint32_t *lengths, uint32_t *libs, uint32_t cap) {
  GkStore s(path);
  for (uint32_t id = 1; id <= s.num_reads() && id <= cap; id++) {
    if (lengths) lengths[id - 1] = s.read(id).seq_len();
    if (libs) libs[id - 1] = s.read(id).library_id();
  }
  return s.num_reads();
}

// concatenated ASCII sequences of reads 1..num_reads into out (caller sized it from the lengths)
void sh_store_bases(const char *path, char *out) {
  GkStore s(path);
  uint64_t at = 0;
  for (uint32_t id = 1; id <= s.num_reads(); id++) {
    s.load_sequence(id, out + at);
    at += s.read(id).seq_len();
  }
}

// the stored 2-bit chunks of reads 1..num_reads, concatenated (ceil(len/4) bytes each); returns 0 if a read has none
int sh_store_2bit(const char *path, uint8_t *out) {
  GkStore s(path);
  uint64_t at = 0;
  for (uint32_t id = 1; id <= s.num_reads(); id++) {
    const uint64_t nb = ((uint64_t)s.read(id).seq_len() + 3) / 4;
    const uint8_t *p = s.sequence_2bit(id);
    if (!p && nb) return 0;
    memcpy(out + at, p, nb);
    at += nb;
  }
  return 1;
}

// write n 24-byte records in `pieces` write() calls (exercises the block boundary logic)
void sh_write_ovb(const char *name, const uint32_t *rec, uint64_t n, uint32_t pieces) {
  OvbWriter w(name);
  if (pieces == 0) pieces = 1;
  uint64_t done = 0;
  for (uint32_t p = 0; p < pieces; p++) {
    uint64_t upto = n * (p + 1) / pieces;
    w.write(rec + 6 * done, upto - done);
    done = upto;
  }
  w.close();
}

void sh_counts_name(const char *name, char *out, uint32_t cap) {
  std::string s = counts_name_for(name);
  strncpy(out, s.c_str(), cap - 1);
  out[cap - 1] = 0;
}

// snappy raw format: compress into out (cap bytes), returns the compressed size; uncompress returns the size or -1
uint64_t sh_snappy_compress(const uint8_t *in, uint64_t n, uint8_t *out, uint64_t cap) {
  std::vector<uint8_t> tmp(snappy_max_compressed(n));
  const size_t bl = snappy_compress(in, n, tmp.data());
  if (bl <= cap) memcpy(out, tmp.data(), bl);
  return bl;
}
int64_t sh_snappy_uncompress(const uint8_t *in, uint64_t n, uint8_t *out, uint64_t cap) {
  std::vector<uint8_t> tmp;
  if (!snappy_uncompress(in, n, tmp) || tmp.size() > cap) return -1;
  memcpy(out, tmp.data(), tmp.size());
  return (int64_t)tmp.size();
}

// records of an .ovb file (any snappy-framed ovFileFull file, the reference's included); returns the record count
uint64_t sh_read_ovb(const char *name, uint32_t *out, uint64_t cap_records) {
  std::vector<uint32_t> rec;
  read_ovb_file(name, rec);
  const uint64_t n = rec.size() / 6;
  if (out && n <= cap_records) memcpy(out, rec.data(), rec.size() * 4);
  return n;
}

// sorted store records -> <path>/{info,index,0001..,evalueLen,overlapScores}; per-read filter inputs come back in the arrays
void sh_write_ovlstore(const char *path, const char *gkp_path, const uint32_t *rec, uint64_t n) {
  GkStore gkp(gkp_path);
  OvlStoreWriter w(path, gkp);
  w.write(rec, n);
  w.finish();
}
uint32_t sh_store_read_info(const char *gkp_path, uint32_t *length, uint8_t *skip_dup, uint8_t *library) {
  GkStore gkp(gkp_path);
  const StoreReadInfo ri = store_read_info(gkp, gkp_path);
  const uint32_t n = (uint32_t)ri.length.size();
  if (length) { memcpy(length, ri.length.data(), n * 4); memcpy(skip_dup, ri.skip_dup.data(), n); memcpy(library, ri.library.data(), n); }
  return n;
}
}
