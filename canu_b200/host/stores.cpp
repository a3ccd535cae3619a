// stores.cpp -- gkpStore reader and .ovb / .counts writer (see stores.h for the reference citations).
#include "stores.h"

#include "snappy_lite.h"

#include <errno.h>
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace oichost {

namespace {

[[noreturn]] void die(const char *fmt, const std::string &a, const char *b = "") {
  fprintf(stderr, fmt, a.c_str(), b);
  fputc('\n', stderr);
  exit(1);
}

const void *map_file(const std::string &name, size_t &bytes) {
  int fd = open(name.c_str(), O_RDONLY);
  if (fd < 0) die("gkStore: failed to open '%s': %s", name, strerror(errno));
  struct stat st;
  if (fstat(fd, &st) != 0) die("gkStore: failed to stat '%s': %s", name, strerror(errno));
  bytes = (size_t)st.st_size;
  const void *p = bytes ? mmap(nullptr, bytes, PROT_READ, MAP_SHARED, fd, 0) : nullptr;
  close(fd);
  if (bytes && p == MAP_FAILED) die("gkStore: failed to mmap '%s': %s", name, strerror(errno));
  return p;
}

struct GkInfo {   // gkStoreInfo, ST/gkStore.H:345-401 (56 bytes on disk)
  uint64_t magic, version;
  uint32_t library_size, read_size, max_libraries_bits, library_name_size, max_read_bits, max_readlen_bits, unused;
  uint32_t num_libraries, num_reads;
  uint32_t pad;
};

}  // namespace

GkStore::GkStore(const std::string &path) : path_(path) {
  FILE *f = fopen((path + "/info").c_str(), "r");
  if (!f) die("gkStore: failed to open '%s/info': %s", path, strerror(errno));
  GkInfo info;
  memset(&info, 0, sizeof(info));
  size_t got = fread(&info, 1, sizeof(info), f);
  fclose(f);
  if (got < 52) die("gkStore: '%s/info' is truncated%s", path);
  if (info.magic != 0x504b473a756e6163ull || info.version != 1) die("gkStore: '%s' is not a gkpStore (bad magic/version)%s", path);
  // the same sanity checks the reference makes before trusting the files (ST/gkStore.C:640-690)
  if (info.library_size != 168 || info.read_size != 16 || info.max_libraries_bits != 6 || info.library_name_size != 128 ||
      info.max_read_bits != 37 || info.max_readlen_bits != 21)
    die("gkStore: '%s' was built with different AS_MAX_* sizes than this overlapper supports (need 168/16/6/128/37/21)%s", path);
  num_libraries_ = info.num_libraries;
  num_reads_ = info.num_reads;
  reads_ = static_cast<const GkRead *>(map_file(path + "/reads", reads_bytes_));
  if (reads_bytes_ < (size_t)(num_reads_ + 1) * sizeof(GkRead)) die("gkStore: '%s/reads' is too short%s", path);
  blobs_ = static_cast<const uint8_t *>(map_file(path + "/blobs", blobs_bytes_));
}

GkStore::~GkStore() {
  if (reads_) munmap(const_cast<GkRead *>(reads_), reads_bytes_);
  if (blobs_) munmap(const_cast<uint8_t *>(blobs_), blobs_bytes_);
}

void GkStore::load_sequence(uint32_t id, char *out) const {
  const GkRead &r = reads_[id];
  const uint32_t len = r.seq_len();
  const uint64_t at = r.m_ptr();
  if (at + 8 > blobs_bytes_ || memcmp(blobs_ + at, "BLOB", 4) != 0) {
    fprintf(stderr, "gkStore: read %u: no BLOB at offset %llu of '%s/blobs'\n", id, (unsigned long long)at, path_.c_str());
    exit(1);
  }
  const uint8_t *b = blobs_ + at + 8;
  bool have = false;
  while (memcmp(b, "STOP", 4) != 0) {
    uint32_t clen;
    memcpy(&clen, b + 4, 4);
    const uint8_t *d = b + 8;
    if (memcmp(b, "2SEQ", 4) == 0) {          // 4 bases per byte, first base in bits 7-6 (ST/gkStoreEncode.C:81-114)
      static const char acgt[4] = {'A', 'C', 'G', 'T'};
      for (uint32_t i = 0; i < len; i++) out[i] = acgt[(d[i >> 2] >> (6 - 2 * (i & 3))) & 3];
      have = true;
    } else if (memcmp(b, "USEQ", 4) == 0) {
      memcpy(out, d, len);
      have = true;
    } else if (memcmp(b, "VERS", 4) && memcmp(b, "NAME", 4) && memcmp(b, "QSEQ", 4) && memcmp(b, "UQLT", 4) && memcmp(b, "3SEQ", 4) &&
               memcmp(b, "4QLT", 4) && memcmp(b, "5QLT", 4) && memcmp(b, "QVAL", 4)) {
      fprintf(stderr, "gkStore: read %u: unknown chunk type '%c%c%c%c'\n", id, b[0], b[1], b[2], b[3]);
      exit(1);
    }
    b += 8 + clen;
    if ((size_t)(b - blobs_) + 4 > blobs_bytes_) { fprintf(stderr, "gkStore: read %u: blob runs past the end of the file\n", id); exit(1); }
  }
  if (!have && len) { fprintf(stderr, "gkStore: read %u has no sequence chunk\n", id); exit(1); }
}

const uint8_t *GkStore::sequence_2bit(uint32_t id) const {
  const GkRead &r = reads_[id];
  const uint64_t at = r.m_ptr();
  if (at + 8 > blobs_bytes_ || memcmp(blobs_ + at, "BLOB", 4) != 0) {
    fprintf(stderr, "gkStore: read %u: no BLOB at offset %llu of '%s/blobs'\n", id, (unsigned long long)at, path_.c_str());
    exit(1);
  }
  const uint8_t *b = blobs_ + at + 8;
  while (memcmp(b, "STOP", 4) != 0) {
    uint32_t clen;
    memcpy(&clen, b + 4, 4);
    if (memcmp(b, "2SEQ", 4) == 0) return ((uint64_t)clen * 4 >= r.seq_len()) ? b + 8 : nullptr;
    b += 8 + clen;
    if ((size_t)(b - blobs_) + 4 > blobs_bytes_) { fprintf(stderr, "gkStore: read %u: blob runs past the end of the file\n", id); exit(1); }
  }
  return nullptr;
}

void GkStore::load_qualities(uint32_t id, uint8_t *out) const {
  const GkRead &r = reads_[id];
  const uint32_t len = r.seq_len();
  const uint64_t at = r.m_ptr();
  memset(out, 0, len);
  if (at + 8 > blobs_bytes_ || memcmp(blobs_ + at, "BLOB", 4) != 0) {
    fprintf(stderr, "gkStore: read %u: no BLOB at offset %llu of '%s/blobs'\n", id, (unsigned long long)at, path_.c_str());
    exit(1);
  }
  const uint8_t *b = blobs_ + at + 8;
  while (memcmp(b, "STOP", 4) != 0) {
    uint32_t clen;
    memcpy(&clen, b + 4, 4);
    if (memcmp(b, "QVAL", 4) == 0 && clen >= 4) {
      uint32_t qv;
      memcpy(&qv, b + 8, 4);
      memset(out, (int)(uint8_t)qv, len);
    } else if (memcmp(b, "UQLT", 4) == 0 && clen >= len) memcpy(out, b + 8, len);
    b += 8 + clen;
    if ((size_t)(b - blobs_) + 4 > blobs_bytes_) { fprintf(stderr, "gkStore: read %u: blob runs past the end of the file\n", id); exit(1); }
  }
}

// ---------------------------------------------------------------------------------------------------------
std::string counts_name_for(const std::string &name) {
  // AS_UTL_findBaseFileName: cut at the first '.' after the last '/', then add ".counts"
  struct stat st;
  if (stat(name.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) return name + "/counts";
  size_t slash = name.rfind('/');
  size_t dot = name.find('.', slash == std::string::npos ? 0 : slash);
  std::string base = dot == std::string::npos ? name : name.substr(0, dot);
  return base + ".counts";
}

OvbWriter::OvbWriter(const std::string &name) : name_(name) {
  f_ = fopen(name.c_str(), "w");
  if (!f_) die("ERROR: failed to open overlap output '%s': %s", name, strerror(errno));
  buf_.reserve(kBufferMax);
}

void OvbWriter::flush() {
  if (buf_.empty()) return;
  // one snappy raw block per buffer (ST/ovStoreFile.C:157-185): size_t compressed length, then the block
  const size_t n = buf_.size() * sizeof(uint32_t);
  std::vector<unsigned char> out(snappy_max_compressed(n));
  const size_t bl = snappy_compress(reinterpret_cast<const unsigned char *>(buf_.data()), n, out.data());
  if (fwrite(&bl, sizeof(size_t), 1, f_) != 1 || fwrite(out.data(), 1, bl, f_) != bl) die("ERROR: write to '%s' failed: %s", name_, strerror(errno));
  buf_.clear();
}

void OvbWriter::write(const uint32_t *rec, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) {
    const uint32_t *w = rec + 6 * i;
    const uint32_t a = w[0], b = w[1], mx = a > b ? a : b;
    if (opr_.size() < (size_t)mx + 1) opr_.resize((size_t)mx + 1 + mx / 2, 0);
    if (opr_len_ < mx + 1) opr_len_ = mx + 1;
    opr_[a]++;
    opr_[b]++;
    buf_.insert(buf_.end(), w, w + 6);
    if (buf_.size() >= kBufferMax) flush();
  }
}

void OvbWriter::close() {
  if (!f_) return;
  flush();
  fclose(f_);
  f_ = nullptr;
  const std::string cn = counts_name_for(name_);
  FILE *c = fopen(cn.c_str(), "w");
  if (!c) die("failed to open counts file '%s' for writing: %s", cn, strerror(errno));
  fwrite(&opr_len_, sizeof(uint32_t), 1, c);
  if (opr_len_) fwrite(opr_.data(), sizeof(uint32_t), opr_len_, c);
  fclose(c);
}

OvbWriter::~OvbWriter() { close(); }

}  // namespace oichost
