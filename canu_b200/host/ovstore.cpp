// ovstore.cpp -- .ovb reader and sequential overlap-store writer (see ovstore.h for the reference citations).
#include "ovstore.h"

#include <errno.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <algorithm>
#include <functional>

#include "snappy_lite.h"

namespace oichost {

namespace {
[[noreturn]] void die2(const char *fmt, const std::string &a, const char *b = "") {
  fprintf(stderr, fmt, a.c_str(), b);
  fputc('\n', stderr);
  exit(1);
}
const uint64_t kStoreMagic = 0x53564f3a756e6163ull, kStoreMagicIncomplete = 0x50564f3a756e6163ull;   // ST/ovStore.H:48-50
const uint32_t kEvalueMax = 4095, kBasesPerBucket = 250, kNScores = 16;
}  // namespace

void read_ovb_file(const std::string &name, std::vector<uint32_t> &rec) {
  FILE *f = fopen(name.c_str(), "r");
  if (!f) die2("ERROR: failed to open overlap file '%s': %s", name, strerror(errno));
  std::vector<uint8_t> comp, raw;
  for (;;) {
    size_t bl = 0;
    const size_t got = fread(&bl, 1, sizeof(size_t), f);
    if (got == 0) break;
    if (got != sizeof(size_t) || bl > (1ull << 30)) die2("ERROR: overlap file '%s' is truncated or not snappy-framed%s", name);
    comp.resize(bl);
    if (fread(comp.data(), 1, bl, f) != bl) die2("ERROR: overlap file '%s' ends inside a block%s", name);
    if (!snappy_uncompress(comp.data(), bl, raw) || raw.size() % 24 != 0) die2("ERROR: overlap file '%s' holds a malformed block%s", name);
    const size_t old = rec.size();
    rec.resize(old + raw.size() / 4);
    memcpy(rec.data() + old, raw.data(), raw.size());
  }
  fclose(f);
}

StoreReadInfo store_read_info(const GkStore &gkp, const std::string &gkp_path) {
  // gkLibrary on disk (ST/gkStore.H:167-196, 168 bytes): name[128], then uint32 libraryID, isNonRandom, readType, readCorrection,
  // finalTrim, removeDuplicateReads, removeSpurReads, removeChimericReads, checkForSubReads, defaultQV
  const uint32_t nl = gkp.num_libraries();
  std::vector<uint8_t> lib((size_t)(nl + 1) * 168, 0);
  FILE *f = fopen((gkp_path + "/libraries").c_str(), "r");
  if (!f) die2("gkStore: failed to open '%s/libraries': %s", gkp_path, strerror(errno));
  if (fread(lib.data(), 168, nl + 1, f) != nl + 1) die2("gkStore: '%s/libraries' is too short%s", gkp_path);
  fclose(f);
  StoreReadInfo r;
  const uint32_t max_id = gkp.num_reads() + 1;
  r.length.assign(max_id, 0);
  r.skip_dup.assign(max_id, 0);
  r.library.assign(max_id, 0);
  for (uint32_t id = 0; id < max_id; id++) {
    const GkRead &rd = gkp.read(id);
    const uint32_t l = rd.library_id();
    r.length[id] = rd.seq_len();
    r.library[id] = (uint8_t)l;
    uint32_t remove_dup = 0;
    if (l <= nl) memcpy(&remove_dup, lib.data() + (size_t)l * 168 + 128 + 5 * 4, 4);
    r.skip_dup[id] = remove_dup == 0;   // ST/ovStoreFilter.C:62-65
  }
  return r;
}

// ---------------------------------------------------------------------------------------------------------
OvlStoreWriter::OvlStoreWriter(const std::string &path, const GkStore &gkp) : path_(path), gkp_(gkp) {
  {   // fail if this is a valid (complete) store (ST/ovStoreWriter.C:95-98)
    FILE *t = fopen((path_ + "/info").c_str(), "r");
    if (t) {
      uint64_t magic = 0;
      const size_t got = fread(&magic, 8, 1, t);
      fclose(t);
      if (got == 1 && magic == kStoreMagic) die2("ERROR:  '%s' is a valid ovStore; cannot create a new one.%s", path_);
    }
  }
  mkdir(path_.c_str(), 0777);
  info_[0] = kStoreMagicIncomplete; info_[1] = 2; info_[2] = 0; info_[3] = UINT64_MAX; info_[4] = 0; info_[5] = 0; info_[6] = 0; info_[7] = 21;
  save_info(false);
  memset(&offt_, 0, sizeof(offt_));
  memset(&offm_, 0, sizeof(offm_));
  index_ = fopen((path_ + "/index").c_str(), "w");
  if (!index_) die2("AS_OVS_createOverlapStore()-- failed to open offset file '%s/index': %s", path_, strerror(errno));
  // ovStoreHistogram(gkp, ovFileNormalWrite): the longest read among IDs 1 .. numReads-1 sizes the length axis (ST/ovStoreHistogram.C:118-122)
  uint32_t longest = 0;
  for (uint32_t ii = 1; ii < gkp_.num_reads(); ii++) longest = std::max(longest, gkp_.read(ii).seq_len());
  opel_len_ = (uint32_t)(longest * 1.40 / kBasesPerBucket + 1);   // "the overlap could have 40% insertions"
  opel_.assign(kEvalueMax + 1, std::vector<uint32_t>());
  scores_.assign(65535, Sco());
  for (Sco &s : scores_) memset(&s, 0, sizeof(Sco));
}

void OvlStoreWriter::save_info(bool complete) {
  if (complete) { info_[0] = kStoreMagic; info_[1] = 2; info_[6] = file_index_; }
  FILE *f = fopen((path_ + "/info").c_str(), "w");
  if (!f) die2("ERROR: failed to save '%s/info': %s", path_, strerror(errno));
  fwrite(info_, 8, 8, f);
  fclose(f);
}

void OvlStoreWriter::process_scores(uint32_t aid) {   // ovStoreHistogram::processScores, ST/ovStoreHistogram.C:161-218
  if (scores_bgn_ == 0) { score_aid_ = aid; scores_bgn_ = aid; return; }
  if (score_list_.empty()) return;
  const uint32_t scoff = score_aid_ - scores_bgn_;
  if (scoff >= scores_.size()) {
    Sco z;
    memset(&z, 0, sizeof(z));
    scores_.resize((size_t)scoff + 65536, z);
  }
  std::sort(score_list_.begin(), score_list_.end(), std::greater<uint16_t>());
  const uint32_t n = (uint32_t)score_list_.size();
  double step = (double)n / kNScores, point = 0;
  if (step < 1) step = 1;
  if (step > 10) step = 10;
  Sco &s = scores_[scoff];
  for (uint32_t ii = 0; ii < kNScores; ii++) s.points[ii] = (uint16_t)(n - 1);
  for (uint32_t ii = 0; ii < kNScores - 1 && point < n; ii++, point += step) s.points[ii] = (uint16_t)floor(point + 0.5);
  for (uint32_t ii = 0; ii < kNScores; ii++) s.scores[ii] = score_list_[s.points[ii]];
  score_list_.clear();
  score_aid_ = aid;
  scores_len_ = scoff + 1;
}

void OvlStoreWriter::write(const uint32_t *rec, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) {
    const uint32_t *w = rec + 6 * i;
    const uint32_t a = w[0], b = w[1];
    if (offt_.a_iid > a) { fprintf(stderr, "LAST:  a:%u\nTHIS:  a:%u b:%u\novStoreWriter: overlaps are not sorted\n", offt_.a_iid, a, b); exit(1); }
    // a new data file when there is none yet or the current one is full (1 GiB of 20-byte records)
    if (data_ && this_file_ >= this_file_max_) { flush_data(); fclose(data_); data_ = nullptr; this_file_ = this_file_max_ = 0; }
    if (!data_) {
      char nm[32];
      snprintf(nm, sizeof(nm), "/%04d", (int)++file_index_);
      data_ = fopen((path_ + nm).c_str(), "w");
      if (!data_) die2("ovStoreWriter: failed to open '%s' for writing: %s", path_ + nm, strerror(errno));
      this_file_ = 0;
      this_file_max_ = 1024ull * 1024 * 1024 / 20;
    }
    // the index: one entry per read ID, gaps filled with empty entries that point at the next block (ST/ovStoreWriter.C:166-197)
    if (offt_.num_olaps != 0 && offt_.a_iid != a) {
      while (offm_.a_iid < offt_.a_iid) {
        offm_.fileno = offt_.fileno; offm_.offset = offt_.offset; offm_.overlap_id = offt_.overlap_id;
        fwrite(&offm_, sizeof(Offt), 1, index_);
        offm_.a_iid++;
      }
      offm_.a_iid++;
      fwrite(&offt_, sizeof(Offt), 1, index_);
      offt_.num_olaps = 0;
    }
    if (offt_.num_olaps == 0) { offt_.a_iid = a; offt_.fileno = file_index_; offt_.offset = (uint32_t)this_file_; offt_.overlap_id = info_[5]; }
    dbuf_.insert(dbuf_.end(), w + 1, w + 6);   // ovFileNormalWrite: b_iid + the four data words, uncompressed (ST/ovStoreFile.C:101-106,196-221)
    if (dbuf_.size() >= (1u << 20)) flush_data();
    offt_.num_olaps++;
    if (info_[3] > a) info_[3] = a;
    if (info_[4] < a) info_[4] = a;
    info_[5]++;
    this_file_++;

    // ---- ovStoreHistogram::addOverlap (ST/ovStoreHistogram.C:243-295)
    const uint64_t d0 = ((uint64_t)w[2] << 32) | w[3], d1 = ((uint64_t)w[4] << 32) | w[5];
    const uint64_t m21 = (1ull << 21) - 1;
    const uint64_t ahg5 = d0 & m21, ahg3 = (d0 >> 21) & m21, bhg5 = d1 & m21, bhg3 = (d1 >> 21) & m21;
    const uint32_t alen = gkp_.read(a).seq_len(), blen = gkp_.read(b).seq_len();
    uint32_t ev = (uint32_t)((d0 >> 42) & 4095);
    uint32_t len = (uint32_t)(((uint64_t)alen - ahg5 - ahg3 + blen - bhg5 - bhg3) / 2);
    if (max_evalue_ < ev) max_evalue_ = ev;
    if (max_olength_ < len) max_olength_ = len;
    len /= kBasesPerBucket;
    if (opel_[ev].empty()) opel_[ev].assign(opel_len_, 0);
    if (len < opel_len_) opel_[ev][len]++;
    else fprintf(stderr, "overlap %8u (len %6d) %8u (len %6d) -- BOGUS\n", a, (int)alen, b, (int)blen);
    if (score_aid_ != a) process_scores(a);
    // ovOverlap::overlapScore (ST/ovOverlap.H:306-310): floor(16384 * identity * a_len / read length) as uint16
    const double identity = 1 - ev / 10000.0;
    const uint32_t a_len = (uint32_t)((uint64_t)alen - ahg3 - ahg5);
    score_list_.push_back((uint16_t)floor(16384.0 * identity * a_len / alen));
  }
}

void OvlStoreWriter::finish() {
  if (finished_) return;
  finished_ = true;
  if (offt_.num_olaps > 0) {   // the last index element, filling in gaps (ST/ovStoreWriter.C:52-64)
    for (; offm_.a_iid < offt_.a_iid; offm_.a_iid++) {
      offm_.fileno = offt_.fileno; offm_.offset = offt_.offset; offm_.num_olaps = 0;
      fwrite(&offm_, sizeof(Offt), 1, index_);
    }
    fwrite(&offt_, sizeof(Offt), 1, index_);
  }
  fclose(index_);
  index_ = nullptr;
  save_info(true);
  if (data_) { flush_data(); fclose(data_); data_ = nullptr; }
  {   // evalueLen (ST/ovStoreHistogram.C:345-376)
    FILE *f = fopen((path_ + "/evalueLen").c_str(), "w");
    if (!f) die2("failed to open evalueLen file '%s/evalueLen' for writing: %s", path_, strerror(errno));
    uint32_t n_arr = 0;
    for (const auto &v : opel_) n_arr += !v.empty();
    const uint32_t hdr[6] = {opel_len_, max_olength_, max_evalue_, 1, kBasesPerBucket, n_arr};
    fwrite(hdr, 4, 6, f);
    for (uint32_t ii = 0; ii <= kEvalueMax; ii++) {
      if (opel_[ii].empty()) continue;
      fwrite(&ii, 4, 1, f);
      fwrite(opel_[ii].data(), 4, opel_len_, f);
    }
    fclose(f);
  }
  {   // overlapScores (ST/ovStoreHistogram.C:380-397)
    process_scores(UINT32_MAX);
    FILE *f = fopen((path_ + "/overlapScores").c_str(), "w");
    if (!f) die2("failed to open overlapScores file '%s/overlapScores' for writing: %s", path_, strerror(errno));
    fwrite(&scores_bgn_, 4, 1, f);
    fwrite(&scores_len_, 4, 1, f);
    if (scores_len_) fwrite(scores_.data(), sizeof(Sco), scores_len_, f);
    fclose(f);
  }
  fprintf(stderr, "Created ovStore '%s' with %llu overlaps for reads from %u to %u.\n", path_.c_str(), (unsigned long long)info_[5],
          (uint32_t)info_[3], (uint32_t)info_[4]);
}

OvlStoreWriter::~OvlStoreWriter() { finish(); }

}  // namespace oichost
