/*
 * oracle/oic_oracle.cpp -- CPU restatement of Canu's overlapInCore hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oic_oracle.h).  Single-threaded, plain C++.
 * Reference citations use OIC/ = src/overlapInCore/, LIBO/ = src/overlapInCore/liboverlap/,
 * ST/ = src/stores/, UTL/ = src/AS_UTL/ under /root/reference.
 *
 * Parity status: PINNED against the reference binary (oracle/_ref/overlapInCore, built by
 * oracle/build_ref.sh from the reference sources) -- tests/golden/make_golden.py ran it on the seeded cases and tests/test_oracle.py compares the
 * sorted overlap records and the .stats counters on seeded synthetic stores.  The reference
 * ships no golden vectors for this path.
 */
#include "oic_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace {

const int32_t MAX_READLEN = (1 << 21) - 1;   // ST/gkStore.H:61-62
const int32_t HOPELESS_MATCH = 90;           // OIC/overlapInCore.H:122
const int32_t MAX_DISTINCT_OLAPS = 3;        // OIC/overlapInCore.H:151
const int32_t MIN_INTERSECTION = 10;         // OIC/overlapInCore.H:171
const int32_t SHIFT_SLACK = 1;               // OIC/overlapInCore.H:197
const int32_t QUALITY_CUTOFF = 20;           // OIC/overlapInCore.H:190
const int32_t BAD_WINDOW_LEN = 50;           // OIC/overlapInCore.H:80
const int32_t BAD_WINDOW_VALUE = 8 * QUALITY_CUTOFF;   // OIC/overlapInCore.H:84 (the long window: 100 / 240, Process_String_Overlaps.C:616)
const int32_t MIN_BRANCH_END_DIST = 20;      // LIBO/prefixEditDistance.C:40

// ---------------------------------------------------------------------------------------
//  Host tables.  LIBO/Binomial_Bound.C:47-107 (bound search), :119-194 (limit table).
// ---------------------------------------------------------------------------------------

// Smallest n >= start with P[>= e errors in n trials at rate p] > 1e-4.
int binomial_bound(int e, double p, int start) {
  const double PROB_BOUND = 1e-4, Z_THOLD = 3.62;
  double q = 1.0 - p;
  if (start < e) start = e;
  for (int n = start; n < MAX_READLEN; n++) {
    if (n <= 35) {
      double sum = 0.0, ppow = 1.0, qpow = pow(q, n);
      int coeff = 1, ct = 0;
      for (int k = 0; k < e && 1.0 - sum > PROB_BOUND; k++) {
        sum += coeff * ppow * qpow;
        coeff *= n - ct;
        coeff /= ++ct;
        ppow *= p;
        qpow /= q;
      }
      if (1.0 - sum > PROB_BOUND) return n;
    } else {
      double z = (e - 0.5 - n * p) / sqrt(n * p * q);
      if (z <= Z_THOLD) return n;
      double sum = 0.0, mu = 1.0, fact = 1.0, pc = exp(-n * p);
      for (int k = 0; k < e; k++) {
        sum += mu * pc / fact;
        mu *= n * p;
        fact *= k + 1;
      }
      if (1.0 - sum > PROB_BOUND) return n;
    }
  }
  return MAX_READLEN;
}

int32_t max_errors_for(double erate) {  // LIBO/prefixEditDistance.C:39
  return 1 + (int32_t)ceil(erate * MAX_READLEN);
}

// ml[0..max_errors-1]; LIBO/Binomial_Bound.C:119-194 (AS_MAX_READLEN_BITS == 21 slope).
void match_limit(double erate, int32_t max_errors, int32_t *ml) {
  int32_t e = 0, s = 1;
  int32_t exact = std::min(max_errors, 2000);
  while (e <= 1) ml[e++] = 0;  // ERRORS_FOR_FREE = 1
  while (e < exact) {
    s = binomial_bound(e - 1, erate, s);
    ml[e] = s - 1;
    e++;
  }
  double sl = 0.982064188397525 / erate + 0.067835741959926;
  double vl = ml[e - 1] + sl;
  while (e < max_errors) {
    ml[e] = (int32_t)ceil(vl);
    vl += sl;
    e++;
  }
}

inline int32_t error_bound(double erate, int32_t len) {  // LIBO/prefixEditDistance.C:67-70
  return (int32_t)ceil(len * erate);
}

// ---------------------------------------------------------------------------------------
//  Prefix edit distance.  LIBO/prefixEditDistance-forward.C, -reverse.C, -extend.C.
//
//  Band rows are stored explicitly: row e keeps [lo, hi] and the values computed for it.
//  The reference writes -2 sentinels at the four cells just outside the trimmed band of
//  row e-1 before computing row e (forward.C:155-161) and never reads further out, so a
//  lookup outside the trimmed band returning -2 is equivalent.
// ---------------------------------------------------------------------------------------

struct Row {
  int32_t lo, hi;            // computed extent (values valid here)
  int32_t tlo, thi;          // trimmed extent (cells outside read as -2 from the next row on)
  std::vector<int32_t> v;    // v[d - lo]
};

struct EditDist {
  double erate;
  bool partial;
  double branch_match_value;
  double min_branch_tail_slope;
  std::vector<int32_t> limit;  // Edit_Match_Limit
  int32_t max_errors;
  std::vector<Row> rows;
  std::vector<int32_t> left_delta, right_delta;
  uint64_t cells = 0, words = 0;

  EditDist(double erate_, bool partial_) : erate(erate_), partial(partial_) {
    max_errors = max_errors_for(erate);
    limit.resize(max_errors + 1);
    match_limit(erate, max_errors, limit.data());
    branch_match_value = erate / (1 + erate);               // prefixEditDistance.C:95
    min_branch_tail_slope = (erate > 0.06) ? 1.0 : 0.20;    // prefixEditDistance.C:41
  }

  // value of row k at diagonal d as the reference would read it while computing row k+1 or tracing back
  inline int32_t get(int32_t k, int32_t d) const {
    const Row &r = rows[k];
    if (d < r.tlo || d > r.thi) return -2;
    return r.v[d - r.lo];
  }
  // Walk rows e..1 backwards choosing the predecessor by the same max rule that built the row
  // (ties: d, then d-1 if strictly greater, then d+1 if strictly greater).  Shared core of
  // Set_Right_Delta (forward.C:48-87) and Set_Left_Delta (reverse.C:52-109).  (e,d) itself is read raw:
  // it is either the cell where the DP stopped or a best cell inside its row's trimmed band.
  int32_t traceback(int32_t e, int32_t d, std::vector<int32_t> &stack) const {
    int32_t last = rows[e].v[d - rows[e].lo];
    for (int32_t k = e; k > 0; k--) {
      int32_t from = d;
      int32_t mx = 1 + get(k - 1, d);
      int32_t j;
      if ((j = get(k - 1, d - 1)) > mx) { from = d - 1; mx = j; }
      if ((j = 1 + get(k - 1, d + 1)) > mx) { from = d + 1; mx = j; }
      if (from == d - 1) {
        stack.push_back(mx - last - 1);
        d--;
        last = get(k - 1, from);
      } else if (from == d + 1) {
        stack.push_back(last - (mx - 1));
        d++;
        last = get(k - 1, from);
      }
    }
    return last;
  }

  void set_right_delta(int32_t e, int32_t d) {  // forward.C:48-87
    std::vector<int32_t> st;
    int32_t last = traceback(e, d, st);
    st.push_back(last + 1);
    right_delta.clear();
    for (int32_t i = (int32_t)st.size() - 1; i > 0; i--) {
      int32_t sgn = (st[i - 1] > 0) - (st[i - 1] < 0);
      right_delta.push_back(abs(st[i]) * sgn);
    }
  }

  void set_left_delta(int32_t e, int32_t d, int32_t &leftover, int32_t &t_end, int32_t t_len) {  // reverse.C:52-109
    left_delta.clear();
    leftover = traceback(e, d, left_delta);
    // a leading +1 indel becomes a substitution by pulling t_end in by one (reverse.C:92-108)
    if (left_delta.size() > 1 && left_delta[0] == 1 && t_end + t_len > 0) {
      if (left_delta[1] > 0) left_delta[0] = left_delta[1] + 1;
      else left_delta[0] = left_delta[1] - 1;
      left_delta.erase(left_delta.begin() + 1);
      t_end--;
    }
  }

  // One direction of the DP.  dir = +1: A[i], T[i] (forward.C:110-285); dir = -1: A[-i], T[-i]
  // (reverse.C:130-300).  Returns errors; sets a_end/t_end (already signed for reverse), leftover (reverse only).
  int32_t run(const char *A, int32_t m, const char *T, int32_t n, int32_t error_limit, int dir,
              int32_t &a_end, int32_t &t_end, int32_t &leftover, bool &match_to_end) {
    auto eq = [&](int32_t i, int32_t j) {
      char a = A[dir * i], t = T[dir * j];
      return a == t || a == 'n' || t == 'n';
    };
    rows.clear();
    if (dir > 0) right_delta.clear(); else left_delta.clear();

    int32_t row = 0;
    while (row < m && eq(row, row)) row++;
    words += (uint64_t)(row + 1 + 15) / 16;
    cells += 1;
    rows.push_back(Row{0, 0, 0, 0, {row}});

    if (row == m) {  // exact
      if (dir > 0) { a_end = t_end = m; }
      else { a_end = t_end = -m; leftover = m; }
      match_to_end = true;
      return 0;
    }

    int32_t left = 0, right = 0;
    double max_score = 0.0;
    int32_t max_score_len = 0, max_score_best_d = 0, max_score_best_e = 0;
    int32_t best_d = 0, best_e = 0, longest = 0;

    auto finish = [&](int32_t e, int32_t d) {
      if (dir > 0) set_right_delta(e, d);
      else set_left_delta(e, d, leftover, t_end, n);
    };

    for (int32_t e = 1; e <= error_limit; e++) {
      left = std::max(left - 1, -e);
      right = std::min(right + 1, e);
      rows.push_back(Row{left, right, left, right, std::vector<int32_t>(right - left + 1, 0)});
      Row &cur = rows[e];

      for (int32_t d = left; d <= right; d++) {
        int32_t r = 1 + get(e - 1, d);
        int32_t j;
        if ((j = get(e - 1, d - 1)) > r) r = j;
        if ((j = 1 + get(e - 1, d + 1)) > r) r = j;
        int32_t r0 = r;
        while (r < m && r + d < n && eq(r, r + d)) r++;
        cells += 1;
        words += (uint64_t)(r - r0 + 1 + 15) / 16;
        cur.v[d - cur.lo] = r;

        if (r == m || r + d == n) {
          double score = r * branch_match_value - e;
          int32_t tail_len = r - max_score_len;
          bool abort = false;
          double slope = (double)(max_score - score) / tail_len;
          if (partial && score < max_score) abort = true;
          if (e > MIN_BRANCH_END_DIST / 2 && tail_len >= MIN_BRANCH_END_DIST && slope >= min_branch_tail_slope)
            abort = true;
          if (abort) {
            a_end = dir * max_score_len;
            t_end = dir * (max_score_len + max_score_best_d);
            finish(max_score_best_e, max_score_best_d);
            match_to_end = false;
            return max_score_best_e;
          }
          if (dir > 0) {
            // force last error to be a mismatch rather than an insertion (forward.C:213-219)
            if (r == m && 1 + get(e - 1, d + 1) == cur.v[d - cur.lo] && d < right) {
              d++;
              cur.v[d - cur.lo] = cur.v[d - 1 - cur.lo];
            }
          }
          a_end = dir * r;
          t_end = dir * (r + d);
          finish(e, d);
          match_to_end = true;
          return e;
        }
      }

      // band trimming (forward.C:235-254)
      int32_t lim = limit[e];
      while (left <= right && left < 0 && cur.v[left - cur.lo] < lim) left++;
      if (left >= 0)
        while (left <= right && cur.v[left - cur.lo] + left < lim) left++;
      if (left > right) { cur.tlo = left; cur.thi = right; break; }
      while (right > 0 && cur.v[right - cur.lo] + right < lim) right--;
      if (right <= 0)
        while (cur.v[right - cur.lo] < lim) right--;
      cur.tlo = left;
      cur.thi = right;

      for (int32_t d = left; d <= right; d++)
        if (cur.v[d - cur.lo] > longest) { best_d = d; best_e = e; longest = cur.v[d - cur.lo]; }

      double score = longest * branch_match_value - e;
      if (score > max_score) {
        max_score = score;
        max_score_len = longest;
        max_score_best_d = best_d;
        max_score_best_e = best_e;
      }
    }

    a_end = dir * max_score_len;
    t_end = dir * (max_score_len + max_score_best_d);
    finish(max_score_best_e, max_score_best_d);
    match_to_end = false;
    return max_score_best_e;
  }

  // LIBO/prefixEditDistance-extend.C:87-229.  Result deltas end up in left_delta.
  int32_t extend(const char *S, int32_t s_len, const char *T, int32_t t_len, int32_t m_start, int32_t m_offset,
                 int32_t m_len, int32_t &s_lo, int32_t &s_hi, int32_t &t_lo, int32_t &t_hi, int32_t &errors) {
    int32_t right_errors = 0, left_errors = 0, leftover = 0;
    bool r_to_end = true, l_to_end = true;
    int32_t s_left_begin = m_start - 1, s_right_begin = m_start + m_len, s_right_len = s_len - s_right_begin;
    int32_t t_left_begin = m_offset - 1, t_right_begin = m_offset + m_len, t_right_len = t_len - t_right_begin;
    int32_t total_olap = std::min(m_start, m_offset) + m_len + std::min(s_right_len, t_right_len);
    int32_t err_limit = error_bound(erate, total_olap);
    int32_t dummy = 0;

    left_delta.clear();
    right_delta.clear();

    if (s_right_len == 0 || t_right_len == 0) {
      s_hi = t_hi = 0;
    } else if (s_right_len <= t_right_len) {
      right_errors = run(S + s_right_begin, s_right_len, T + t_right_begin, t_right_len, err_limit, +1, s_hi, t_hi, dummy, r_to_end);
      for (auto &x : right_delta) x = -x;
    } else {
      right_errors = run(T + t_right_begin, t_right_len, S + s_right_begin, s_right_len, err_limit, +1, t_hi, s_hi, dummy, r_to_end);
    }
    s_hi += s_right_begin - 1;
    t_hi += t_right_begin - 1;

    if (s_left_begin < 0 || t_left_begin < 0) {
      s_lo = t_lo = 0;
    } else if (s_right_begin <= t_right_begin) {
      left_errors = run(S + s_left_begin, s_left_begin + 1, T + t_left_begin, t_left_begin + 1, err_limit - right_errors, -1, s_lo, t_lo, leftover, l_to_end);
    } else {
      left_errors = run(T + t_left_begin, t_left_begin + 1, S + s_left_begin, s_left_begin + 1, err_limit - right_errors, -1, t_lo, s_lo, leftover, l_to_end);
      for (auto &x : left_delta) x = -x;
    }
    s_lo += s_left_begin + 1;
    t_lo += t_left_begin + 1;

    errors = left_errors + right_errors;
    int32_t kind = !r_to_end ? (!l_to_end ? 0 : 2) : (!l_to_end ? 1 : 3);

    if (!right_delta.empty()) {
      if (right_delta[0] > 0) left_delta.push_back(-(right_delta[0] + leftover + m_len));
      else left_delta.push_back(-(right_delta[0] - leftover - m_len));
    }
    for (size_t i = 1; i < right_delta.size(); i++) left_delta.push_back(-right_delta[i]);
    right_delta.clear();
    return kind;
  }
};

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------
//  The overlapper proper.
// ---------------------------------------------------------------------------------------

struct Occ { uint32_t str, off; };   // string index within the hash block, offset in that string

struct KmerEntry {
  std::vector<Occ> occ;   // DESCENDING (str, off): the reference pushes on a chain head (Build_Hash_Index.C:317-323)
  bool empty = false;     // Empty bit: screened k-mer (Build_Hash_Index.C:200-202)
};

struct HashString {
  uint32_t length = 0;
  bool lscreen = true, rscreen = true;   // skipped slots are "screened" (Build_Hash_Index.C:546-550)
  const char *seq = nullptr;
};

struct MatchNode { int32_t offset, len, start; };

struct PairAcc {   // String_Olap_t, OIC/overlapInCore.H:245-255
  uint32_t str;
  std::vector<MatchNode> list;  // front() is the list head
  double diag_sum = 0.0;
  int32_t diag_ct = 0, diag_bgn = MAX_READLEN, diag_end = 0;
  bool consistent = true;
  uint32_t slot;      // primary slot (0..255) or 256+k for the k-th overflow node
};

struct OlapInfo {  // Olap_Info_t, OIC/overlapInCore.H:258-267
  int32_t s_lo, s_hi, t_lo, t_hi;
  double quality;
  int32_t delta_ct;
  int32_t s_left_boundary, s_right_boundary, t_left_boundary, t_right_boundary;
  int32_t min_diag, max_diag;
};

inline int code_of(char c) {   // Bit_Equivalent, OIC/overlapInCore.C:517-520 (anything else codes as 0)
  switch (c) { case 'c': return 1; case 'g': return 2; case 't': return 3; default: return 0; }
}
inline bool is_bad(char c) { return !(c == 'a' || c == 'c' || c == 'g' || c == 't'); }

// UTL/AS_UTL_reverseComplement.C:39-72,114-143 : non-ACGT maps to NUL
inline char rc_char(char c) {
  switch (c) {
    case 'a': return 't'; case 'c': return 'g'; case 'g': return 'c'; case 't': return 'a';
    case 'A': return 'T'; case 'C': return 'G'; case 'G': return 'C'; case 'T': return 'A';
    default: return 0;
  }
}

}  // namespace

struct oracle_ctx {
  oracle_params_t P;
  uint32_t n_reads;
  std::vector<uint64_t> offsets;
  std::vector<uint32_t> lengths;
  std::vector<char> bases;          // lower-cased copy, reads concatenated with a NUL after each
  std::vector<uint8_t> hash_ok, ref_ok;
  std::vector<std::string> skip_kmers;
  uint32_t dump_mask = 0;

  // current hash block
  uint32_t hash_string_num_offset = 1;
  std::vector<HashString> strings;
  std::unordered_map<uint64_t, KmerEntry> index;
  std::vector<std::string> extra_strings;   // screened k-mers that were absent (Add_Extra_Hash_String)

  // per query strand state
  std::vector<PairAcc> pairs;
  std::unordered_map<uint32_t, uint32_t> pair_of;  // str -> index into pairs
  bool slot_used[256];
  uint32_t n_overflow;
  bool left_end_screened, right_end_screened;
  int64_t a_olaps_for_frag, b_olaps_for_frag;
  // -w
  std::vector<uint8_t> quals;          // per base, same offsets as the bases
  bool use_window_filter = false;
  const uint8_t *s_quality = nullptr;  // qualities of the current query strand (reversed on the reverse strand)

  EditDist *ed = nullptr;

  std::vector<oracle_overlap_t> overlaps;
  oracle_stats_t stats;
  std::vector<oracle_pair_t> d_pairs;
  std::vector<oracle_match_t> d_matches;
  std::vector<oracle_ext_t> d_exts;
  std::vector<uint64_t> d_keys, d_occ_start, d_occ;
  std::vector<uint8_t> d_flags;

  const char *read_seq(uint32_t id) const { return bases.data() + offsets[id - 1]; }

  uint32_t build_hash_index(uint32_t bgn, uint32_t end);
  void mark_skip_kmers();
  void hash_mark_empty(uint64_t key, const std::string &s);
  void find_overlaps(const char *frag, int32_t frag_len, uint32_t frag_num, int dir);
  void add_ref(uint32_t str, uint32_t hoff, int32_t offset);
  void process_string_olaps(const char *S, int32_t len, uint32_t id, int dir);
  void process_matches(PairAcc &pa, const char *S, int32_t s_len, uint32_t s_id, int dir);
  void output_overlap(uint32_t s_id, int32_t s_len, int dir, uint32_t t_id, int32_t t_len, const OlapInfo &o);
  void output_partial(uint32_t s_id, uint32_t t_id, int dir, const OlapInfo &o, int32_t s_len, int32_t t_len);
  void dump_index();
};

// OIC/overlapInCore-Build_Hash_Index.C:443-675 (+ Put_String_In_Hash :360-433, Hash_Insert :296-350).
// Loads reads bgn..end into string slots 0.. until a limit trips; returns the last ID loaded.
uint32_t oracle_ctx::build_hash_index(uint32_t bgn, uint32_t end) {
  const int K = P.kmer_len;
  hash_string_num_offset = bgn;
  strings.clear();
  index.clear();
  extra_strings.clear();
  uint64_t total_len = 0;
  uint64_t hash_entries = 0;
  // hash_entry_limit = Max_Hash_Load * HASH_TABLE_SIZE * ENTRIES_PER_BUCKET (double -> uint64), :474
  uint64_t hash_entry_limit = (uint64_t)(P.hash_load * (double)((uint64_t)1 << P.hash_bits) * 21);

  uint32_t cur;
  for (cur = bgn; strings.size() < P.max_hash_strings && total_len < P.max_hash_data_len &&
                  hash_entries < hash_entry_limit && cur <= end; cur++) {
    HashString hs;
    strings.push_back(hs);
    uint32_t si = (uint32_t)strings.size() - 1;
    if (!hash_ok[cur - 1]) continue;
    uint32_t len = lengths[cur - 1];
    if ((int64_t)len < (int64_t)P.min_olap_len) continue;
    const char *seq = read_seq(cur);
    strings[si].length = len;
    strings[si].lscreen = strings[si].rscreen = false;
    strings[si].seq = seq;
    total_len += len + 1;
    stats.n_hash_bases += len;

    // every window without a non-acgt character is inserted, first base in the low bits (:374-428)
    if ((int)len >= K) {
      uint64_t key = 0, bad = 0;
      const uint64_t kmask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
      for (uint32_t i = 0; i < len; i++) {
        key = (key >> 2) | ((uint64_t)code_of(seq[i]) << (2 * (K - 1)));
        bad = (bad >> 1) | ((uint64_t)is_bad(seq[i]) << (K - 1));
        if ((int)i + 1 < K) continue;
        if (bad) continue;
        KmerEntry &ke = index[key & kmask];
        if (ke.occ.empty()) hash_entries++;
        ke.occ.push_back(Occ{si, i + 1 - (uint32_t)K});
        stats.n_inserts++;
      }
    }
    // NOTE: a read shorter than K would walk off its end in the reference (Put_String_In_Hash reads K chars
    // unconditionally); --minlength >= K in every real invocation.
  }
  cur--;
  stats.n_distinct += hash_entries;

  for (auto &kv : index) std::reverse(kv.second.occ.begin(), kv.second.occ.end());   // newest first

  if (strings.empty()) return end;   // :627-630

  if (!skip_kmers.empty()) mark_skip_kmers();
  return cur;
}

// Hash_Mark_Empty + Mark_Screened_Ends_*: OIC/overlapInCore-Build_Hash_Index.C:147-226
void oracle_ctx::hash_mark_empty(uint64_t key, const std::string &s) {
  const int K = P.kmer_len;
  auto it = index.find(key);
  if (it != index.end()) {
    KmerEntry &ke = it->second;
    if (!ke.empty) {
      for (const Occ &o : ke.occ) {
        if (o.str >= strings.size()) continue;  // an extra string
        HashString &hs = strings[o.str];
        if ((int64_t)o.off < HOPELESS_MATCH) hs.lscreen = true;
        if ((int64_t)hs.length - (int64_t)o.off - K + 1 < HOPELESS_MATCH) hs.rscreen = true;
      }
    }
    ke.empty = true;
    return;
  }
  if (P.use_hopeless_check) {   // absent: add as an Empty entry so that queries can see it (:207-219)
    KmerEntry ke;
    ke.empty = true;
    ke.occ.push_back(Occ{(uint32_t)(strings.size() + extra_strings.size()), 0});
    extra_strings.push_back(s);
    index[key] = ke;
  }
}

// Mark_Skip_Kmers: OIC/overlapInCore-Build_Hash_Index.C:233-286 (k-mer, then its reverse complement)
void oracle_ctx::mark_skip_kmers() {
  const int K = P.kmer_len;
  for (const std::string &raw : skip_kmers) {
    std::string s(raw);
    uint64_t key = 0;
    for (int i = 0; i < K; i++) { s[i] = (char)tolower(s[i]); key |= (uint64_t)code_of(s[i]) << (2 * i); }
    hash_mark_empty(key, s);
    std::string r(K, 'a');
    for (int i = 0; i < K; i++) r[i] = rc_char(s[K - 1 - i]);
    key = 0;
    for (int i = 0; i < K; i++) key |= (uint64_t)code_of(r[i]) << (2 * i);
    hash_mark_empty(key, r);
  }
}

// Add_Ref + Add_Match: OIC/overlapInCore-Find_Overlaps.C:79-212
void oracle_ctx::add_ref(uint32_t str, uint32_t hoff, int32_t offset) {
  const int K = P.kmer_len;
  stats.n_add_ref++;
  uint32_t pi;
  auto it = pair_of.find(str);
  if (it == pair_of.end()) {
    PairAcc pa;
    pa.str = str;
    // 256-slot open hash on the string number, overflow nodes appended in creation order (:163-187);
    // Process_String_Olaps later walks the array in index order (Process_String_Overlaps.C:701-741)
    uint32_t sub = (str ^ (str >> 8)) & 255;
    if (!slot_used[sub]) { slot_used[sub] = true; pa.slot = sub; }
    else pa.slot = 256 + n_overflow++;
    pi = (uint32_t)pairs.size();
    pairs.push_back(pa);
    pair_of[str] = pi;
  } else pi = it->second;
  PairAcc &pa = pairs[pi];

  pa.diag_sum += (double)hoff - offset;
  pa.diag_ct++;
  if (pa.diag_bgn > offset) pa.diag_bgn = offset;
  if (pa.diag_end < offset) pa.diag_end = offset;

  // Add_Match (:79-149).  list.front() is the head.
  int32_t new_diag = (int32_t)hoff - offset;
  int32_t diag = 0, expected_start = 0, num_checked = 0;
  bool move_to_front = false;
  for (size_t i = 0; i < pa.list.size(); i++) {
    MatchNode &nd = pa.list[i];
    expected_start = nd.start + nd.len - K + 1;
    diag = nd.offset - nd.start;
    if (expected_start < offset) break;
    if (expected_start == offset) {
      if (new_diag == diag) {
        nd.len += 1;
        if (move_to_front) {
          MatchNode save = nd;
          pa.list.erase(pa.list.begin() + i);
          pa.list.insert(pa.list.begin(), save);
        }
        return;
      } else move_to_front = true;
    }
    num_checked++;
  }
  if (!pa.list.empty() && (num_checked > 0 || abs(diag - new_diag) > 3 || offset < expected_start + K - 2))
    pa.consistent = false;
  pa.list.insert(pa.list.begin(), MatchNode{(int32_t)hoff, K, offset});
}

// Find_Overlaps + Hash_Find: OIC/overlapInCore-Find_Overlaps.C:226-385
void oracle_ctx::find_overlaps(const char *frag, int32_t frag_len, uint32_t frag_num, int dir) {
  const int K = P.kmer_len;
  pairs.clear();
  pair_of.clear();
  memset(slot_used, 0, sizeof(slot_used));
  n_overflow = 0;
  left_end_screened = right_end_screened = false;
  a_olaps_for_frag = b_olaps_for_frag = 0;   // Find_Overlaps.C:306-307

  // The scan is NUL-driven: the first K characters are taken unconditionally (:309-311), then `while (*P != 0)` (:341) stops
  // at the first NUL at index >= K.  Only the reverse strand can hold NULs (RC turns non-ACGT into NUL).
  int32_t eff_len = frag_len;
  for (int32_t i = K; i < frag_len; i++)
    if (frag[i] == 0) { eff_len = i; break; }
  stats.n_query_bases += eff_len;
  if (eff_len >= K) {
    uint64_t key = 0, bad = 0;
    const uint64_t kmask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
    for (int32_t i = 0; i < eff_len; i++) {
      key = (key >> 2) | ((uint64_t)code_of(frag[i]) << (2 * (K - 1)));
      bad = (bad >> 1) | ((uint64_t)is_bad(frag[i]) << (K - 1));
      if (i + 1 < K) continue;
      int32_t offset = i + 1 - K;
      stats.n_windows++;
      if (bad) continue;   // the strncmp verify fails against any (all-acgt) table string
      auto it = index.find(key & kmask);
      if (it == index.end()) continue;
      stats.n_found++;
      const KmerEntry &ke = it->second;
      if (ke.empty) {   // hi_hits (:323-325 first window, :359-366 the rest)
        if (offset == 0) left_end_screened = true;
        else {
          if (offset < HOPELESS_MATCH) left_end_screened = true;
          if (frag_len - offset - K + 1 < HOPELESS_MATCH) right_end_screened = true;
        }
        continue;
      }
      for (const Occ &o : ke.occ) {
        stats.n_refs++;
        if (frag_num < o.str + hash_string_num_offset) add_ref(o.str, o.off, offset);
      }
    }
  }
  process_string_olaps(frag, frag_len, frag_num, dir);
}

// Process_String_Olaps: OIC/overlapInCore-Process_String_Overlaps.C:687-800, the -l path (:743-797) included
void oracle_ctx::process_string_olaps(const char *S, int32_t len, uint32_t id, int dir) {
  if (pairs.empty()) return;
  // compaction of String_Olap_Space (:701-718): array index order = the 256 primary slots, then the overflow nodes in creation order
  std::vector<uint32_t> order(pairs.size());
  for (uint32_t i = 0; i < pairs.size(); i++) order[i] = i;
  std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return pairs[a].slot < pairs[b].slot; });

  for (uint32_t oi : order) {
    PairAcc &pa = pairs[oi];
    pa.diag_sum /= pa.diag_ct;
    if (dump_mask & 1) {
      oracle_pair_t dp;
      dp.q_id = id; dp.h_id = pa.str + hash_string_num_offset; dp.strand = (uint32_t)dir;
      dp.consistent = pa.consistent; dp.diag_ct = pa.diag_ct; dp.diag_bgn = pa.diag_bgn; dp.diag_end = pa.diag_end;
      dp.n_matches = (uint32_t)pa.list.size(); dp.first_match = d_matches.size();
      for (const MatchNode &m : pa.list) d_matches.push_back(oracle_match_t{m.offset, m.len, m.start});
      d_pairs.push_back(dp);
    }
  }
  // one partner: the --minkmers filter (computeMinimumKmers :75-86,725), then Process_Matches
  auto one = [&](PairAcc &pa) {
    if (P.filter_by_kmer_count != 0) {
      double ovl = (double)(pa.diag_end - pa.diag_bgn);
      if (ovl < 0) ovl = -ovl;
      uint64_t expected = 0;
      if (!(ovl < (double)P.kmer_len))
        expected = (uint64_t)(int)floor(exp(-1.0 * (double)P.kmer_len * P.max_erate) * (ovl - P.kmer_len + 1));
      uint64_t need = std::max(P.filter_by_kmer_count, expected);
      if (need > (uint64_t)pa.diag_ct) { stats.kmer_hits_skipped++; return; }
    }
    process_matches(pa, S, len, id, dir);
  };
  const uint64_t ct = order.size();
  if (ct <= P.frag_olap_limit) {   // :722-741 (always, without -l)
    for (uint32_t oi : order) one(pairs[oi]);
    return;
  }
  // more partners than the limit (:743-797): by mean diagonal -- qsort(By_Diag_Sum), which on glibc is a merge sort, so partners on
  // the same mean diagonal keep their array order -- then outward from diagonal zero, each side until its end of the read is full
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return pairs[a].diag_sum < pairs[b].diag_sum; });
  uint64_t start = 0;
  while (start < ct && pairs[order[start]].diag_sum < 0) start++;
  for (uint64_t i = start; i < ct && (uint64_t)a_olaps_for_frag < P.frag_olap_limit; i++) one(pairs[order[i]]);
  for (int64_t i = (int64_t)start - 1; i >= 0 && (uint64_t)b_olaps_for_frag < P.frag_olap_limit; i--) one(pairs[order[i]]);
}

namespace {

// Add_Overlap: OIC/overlapInCore-Process_String_Overlaps.C:220-289
// returns the slot whose alignment -- and therefore delta list (:253-256,278-282) -- was set, or -1
int add_overlap(bool partial, int32_t s_lo, int32_t s_hi, int32_t t_lo, int32_t t_hi, double qual, int32_t delta_ct,
                OlapInfo *olap, int &ct) {
  if (!partial) {
    int32_t new_diag = t_lo - s_lo;
    for (int i = 0; i < ct; i++) {
      OlapInfo &o = olap[i];
      int32_t old_diag = o.t_lo - o.s_lo;
      if ((new_diag > 0 && old_diag > 0 && o.t_right_boundary - new_diag - o.s_left_boundary >= MIN_INTERSECTION) ||
          (new_diag <= 0 && old_diag <= 0 && o.s_right_boundary + new_diag - o.t_left_boundary >= MIN_INTERSECTION)) {
        if (new_diag < o.min_diag) o.min_diag = new_diag;
        if (new_diag > o.max_diag) o.max_diag = new_diag;
        if (s_lo < o.s_left_boundary) o.s_left_boundary = s_lo;
        if (s_hi > o.s_right_boundary) o.s_right_boundary = s_hi;
        if (t_lo < o.t_left_boundary) o.t_left_boundary = t_lo;
        if (t_hi > o.t_right_boundary) o.t_right_boundary = t_hi;
        if (qual < o.quality) {
          o.s_lo = s_lo; o.s_hi = s_hi; o.t_lo = t_lo; o.t_hi = t_hi;
          o.quality = qual;
          o.delta_ct = delta_ct;
          return i;
        }
        return -1;
      }
    }
  }
  if (ct >= MAX_DISTINCT_OLAPS) return -1;
  OlapInfo &o = olap[ct];
  o.s_lo = o.s_left_boundary = s_lo;
  o.s_hi = o.s_right_boundary = s_hi;
  o.t_lo = o.t_left_boundary = t_lo;
  o.t_hi = o.t_right_boundary = t_hi;
  o.quality = qual;
  o.delta_ct = delta_ct;
  o.min_diag = o.max_diag = t_lo - s_lo;
  ct++;
  return ct - 1;
}

// Has_Bad_Window (Process_String_Overlaps.C:365-388): does any window_len-long stretch of a[0 .. n) sum to threshold or more
bool has_bad_window(const std::vector<int8_t> &a, int window_len, int threshold) {
  const int n = (int)a.size();
  if (n < window_len) return false;
  int32_t sum = 0;
  int i = 0, j = 0;
  for (i = 0; i < window_len; i++) sum += a[i];
  if (sum >= threshold) return true;
  while (i < n) {
    sum -= a[j++];
    sum += a[i++];
    if (sum >= threshold) return true;
  }
  return false;
}

// Lies_On_Alignment: :305-326
bool lies_on_alignment(int32_t start, int32_t offset, int32_t s_lo, int32_t t_lo, const std::vector<int32_t> &delta) {
  int32_t diag = t_lo - s_lo, new_diag = offset - start;
  for (int32_t dl : delta) {
    s_lo += abs(dl);
    if (start < s_lo) return abs(new_diag - diag) <= SHIFT_SLACK;
    if (dl < 0) diag++;
    else { s_lo++; diag--; }
  }
  return abs(new_diag - diag) <= SHIFT_SLACK;
}

// Combine_Into_One_Olap :93-141
void combine_into_one(OlapInfo *olap, int ct, int *deleted) {
  int best = 0;
  OlapInfo acc = olap[0];
  for (int i = 1; i < ct; i++) {
    if (olap[i].quality < olap[best].quality) best = i;
    acc.min_diag = std::min(acc.min_diag, olap[i].min_diag);
    acc.max_diag = std::max(acc.max_diag, olap[i].max_diag);
    acc.s_left_boundary = std::min(acc.s_left_boundary, olap[i].s_left_boundary);
    acc.s_right_boundary = std::max(acc.s_right_boundary, olap[i].s_right_boundary);
    acc.t_left_boundary = std::min(acc.t_left_boundary, olap[i].t_left_boundary);
    acc.t_right_boundary = std::max(acc.t_right_boundary, olap[i].t_right_boundary);
  }
  olap[best].min_diag = acc.min_diag; olap[best].max_diag = acc.max_diag;
  olap[best].s_left_boundary = acc.s_left_boundary; olap[best].s_right_boundary = acc.s_right_boundary;
  olap[best].t_left_boundary = acc.t_left_boundary; olap[best].t_right_boundary = acc.t_right_boundary;
  for (int i = 0; i < ct; i++) deleted[i] = (i != best);
}

// Merge_Intersecting_Olaps :151-207
void merge_intersecting(OlapInfo *p, int ct, int *deleted) {
  for (int i = 0; i < ct - 1; i++)
    for (int j = i + 1; j < ct; j++) {
      if (deleted[i] || deleted[j]) continue;
      int lo = p[i].min_diag, hi = p[i].max_diag;
      if ((lo <= 0 && p[j].min_diag > 0) || (lo > 0 && p[j].min_diag <= 0)) continue;
      if ((lo >= 0 && p[j].t_right_boundary - lo - p[j].s_left_boundary >= MIN_INTERSECTION) ||
          (lo <= 0 && p[j].s_right_boundary + lo - p[j].t_left_boundary >= MIN_INTERSECTION) ||
          (hi >= 0 && p[j].t_right_boundary - hi - p[j].s_left_boundary >= MIN_INTERSECTION) ||
          (hi <= 0 && p[j].s_right_boundary + hi - p[j].t_left_boundary >= MIN_INTERSECTION)) {
        OlapInfo *keep, *discard;
        if (p[i].quality < p[j].quality) { keep = p + i; discard = p + j; deleted[j] = 1; }
        else { keep = p + j; discard = p + i; deleted[i] = 1; }
        keep->min_diag = std::min(keep->min_diag, discard->min_diag);
        keep->max_diag = std::max(keep->max_diag, discard->max_diag);
        keep->s_left_boundary = std::min(keep->s_left_boundary, discard->s_left_boundary);
        keep->s_right_boundary = std::max(keep->s_right_boundary, discard->s_right_boundary);
        keep->t_left_boundary = std::min(keep->t_left_boundary, discard->t_left_boundary);
        keep->t_right_boundary = std::max(keep->t_right_boundary, discard->t_right_boundary);
      }
    }
}

// Choose_Best_Partial :334-356
void choose_best_partial(OlapInfo *olap, int ct, int *deleted) {
  int best = 0;
  double matching = (1.0 - olap[0].quality) * (2 + olap[0].s_hi - olap[0].s_lo + olap[0].t_hi - olap[0].t_lo);
  for (int i = 1; i < ct; i++) {
    double mb = (1.0 - olap[i].quality) * (2 + olap[i].s_hi - olap[i].s_lo + olap[i].t_hi - olap[i].t_lo);
    if (matching < mb || (matching == mb && olap[i].quality < olap[best].quality)) best = i;
  }
  for (int i = 0; i < ct; i++) deleted[i] = (i != best);
}

inline uint32_t encode_evalue(double q) {   // ST/ovOverlap.H:41-47
  return (q < (4095 / 10000.0)) ? (uint32_t)(int)(10000.0 * q + 0.5) : 4095u;
}
inline void set_a_hang(oracle_overlap_t &o, int32_t a) { o.ahg5 = (a < 0) ? 0 : a; o.bhg5 = (a < 0) ? -a : 0; }  // ovOverlap.H:230
inline void set_b_hang(oracle_overlap_t &o, int32_t b) { o.bhg3 = (b < 0) ? 0 : b; o.ahg3 = (b < 0) ? -b : 0; }  // ovOverlap.H:231

}  // namespace

// Process_Matches: OIC/overlapInCore-Process_String_Overlaps.C:398-653 (the -w window filter is not restated)
void oracle_ctx::process_matches(PairAcc &pa, const char *S, int32_t s_len, uint32_t s_id, int dir) {
  const HashString &hs = strings[pa.str];
  const char *T = hs.seq;
  int32_t t_len = (int32_t)hs.length;
  uint32_t t_id = pa.str + hash_string_num_offset;
  std::vector<MatchNode> &list = pa.list;
  const bool partial = P.partial != 0;

  if (P.use_hopeless_check && list.size() == 1 && !partial) {   // :433-464
    const MatchNode &m = list[0];
    bool hopeless = false;
    int32_t s_head = m.start, t_head = m.offset;
    if (s_head <= t_head) { if (s_head > HOPELESS_MATCH && !left_end_screened) hopeless = true; }
    else { if (t_head > HOPELESS_MATCH && !hs.lscreen) hopeless = true; }
    int32_t s_tail = s_len - s_head - m.len + 1, t_tail = t_len - t_head - m.len + 1;
    if (s_tail <= t_tail) { if (s_tail > HOPELESS_MATCH && !right_end_screened) hopeless = true; }
    else { if (t_tail > HOPELESS_MATCH && !hs.rscreen) hopeless = true; }
    if (hopeless) { list.clear(); stats.kmer_hits_without_olap++; return; }
  }

  OlapInfo distinct[MAX_DISTINCT_OLAPS];
  std::vector<int32_t> olap_delta[MAX_DISTINCT_OLAPS];   // Olap_Info_t.delta: the delta list of the alignment each slot holds
  int distinct_ct = 0;
  int overlaps_output = 0;

  // Kind_Of_Olap starts as NONE and, like S_Lo .. T_Hi, keeps its value through an iteration that -l skips (:415,484-490): such a
  // seed only removes itself (re-applying the previous alignment's pruning removes nothing new).  In -G mode a skipped FIRST seed
  // makes the reference read uninitialised bounds; here they are zero (documented in DESIGN.md).
  int32_t kind = 0, s_lo = 0, s_hi = 0, t_lo = 0, t_hi = 0, errors = 0;
  while (!list.empty()) {
    size_t li = 0;
    for (size_t i = 1; i < list.size(); i++)
      if (list[i].len > list[li].len) li = i;     // first of the longest, in list order (:473-480)
    MatchNode longest = list[li];

    const int32_t a_hang = longest.start - longest.offset, b_hang = a_hang + s_len - t_len;
    const bool hit_limit = ((uint64_t)a_olaps_for_frag >= P.frag_olap_limit && a_hang <= 0) ||
                           ((uint64_t)b_olaps_for_frag >= P.frag_olap_limit && b_hang <= 0);   // :482-490
    if (hit_limit) {
      if (pa.consistent) { list.clear(); break; }
      std::vector<MatchNode> keep;
      for (size_t i = 0; i < list.size(); i++) {
        const MatchNode &nd = list[i];
        bool remove = (i == li) ||
                      ((kind == 3 || partial) && s_lo - SHIFT_SLACK <= nd.start &&
                       nd.start + nd.len <= (s_hi + 1) + SHIFT_SLACK - 1 &&
                       lies_on_alignment(nd.start, nd.offset, s_lo, t_lo, ed->left_delta));
        if (!remove) keep.push_back(nd);
      }
      list.swap(keep);
      continue;
    }
    stats.ext_calls++;
    kind = ed->extend(S, s_len, T, t_len, longest.start, longest.offset, longest.len, s_lo, s_hi, t_lo, t_hi, errors);

    if (dump_mask & 2) {
      oracle_ext_t x;
      x.q_id = s_id; x.h_id = t_id; x.strand = (uint32_t)dir;
      x.m_offset = longest.offset; x.m_len = longest.len; x.m_start = longest.start;
      x.s_lo = s_lo; x.s_hi = s_hi; x.t_lo = t_lo; x.t_hi = t_hi; x.errors = errors; x.kind = kind;
      x.delta_len = (int32_t)ed->left_delta.size();
      uint32_t h = 2166136261u;
      for (int32_t dl : ed->left_delta) { h ^= (uint32_t)dl; h *= 16777619u; }
      x.delta_hash = (int32_t)h;
      d_exts.push_back(x);
    }

    if (kind == 3 || partial) {
      if (1 + s_hi - s_lo >= P.min_olap_len && 1 + t_hi - t_lo >= P.min_olap_len) {
        int32_t olap_len = 1 + std::min(s_hi - s_lo, t_hi - t_lo);
        double quality = (double)errors / olap_len;
        if (errors <= error_bound(P.max_erate, olap_len)) {
          const int slot = add_overlap(partial, s_lo, s_hi, t_lo, t_hi, quality, (int32_t)ed->left_delta.size(), distinct, distinct_ct);
          if (slot >= 0) olap_delta[slot] = ed->left_delta;
        }
      }
    }

    if (pa.consistent) { list.clear(); break; }

    std::vector<MatchNode> keep;
    for (size_t i = 0; i < list.size(); i++) {
      const MatchNode &nd = list[i];
      bool remove = (i == li) ||
                    ((kind == 3 || partial) && s_lo - SHIFT_SLACK <= nd.start &&
                     nd.start + nd.len <= (s_hi + 1) + SHIFT_SLACK - 1 &&
                     lies_on_alignment(nd.start, nd.offset, s_lo, t_lo, ed->left_delta));
      if (!remove) keep.push_back(nd);
    }
    list.swap(keep);
  }

  if (distinct_ct > 0) {
    int deleted[MAX_DISTINCT_OLAPS] = {0, 0, 0};
    if (partial) { if (P.unique_olap) choose_best_partial(distinct, distinct_ct, deleted); }
    else { if (P.unique_olap) combine_into_one(distinct, distinct_ct, deleted); else merge_intersecting(distinct, distinct_ct, deleted); }
    for (int i = 0; i < distinct_ct; i++) {
      if (deleted[i]) continue;
      if (use_window_filter) {   // :562-620: quality-weighted differences along the alignment, then two sliding windows
        const OlapInfo &p = distinct[i];
        const uint8_t *tq = quals.data() + offsets[t_id - 1];
        std::vector<int8_t> q_diff;
        int32_t si = p.s_lo, tj = p.t_lo;
        auto same = [&](int32_t a_, int32_t b_) { return S[a_] == T[b_] || S[a_] == 'n' || T[b_] == 'n'; };
        for (int32_t dl : olap_delta[i]) {
          const int32_t len = dl < 0 ? -dl : dl;
          for (int32_t n = 1; n < len; n++) {
            int32_t d = same(si, tj) ? 0 : std::min((int32_t)std::min(s_quality[si], tq[tj]), (int32_t)QUALITY_CUTOFF);
            q_diff.push_back((int8_t)d);
            si++; tj++;
          }
          int32_t d;
          if (dl > 0) { d = s_quality[si]; si++; } else { d = tq[tj]; tj++; }
          q_diff.push_back((int8_t)std::min(d, (int32_t)QUALITY_CUTOFF));
        }
        while (si <= p.s_hi) {
          int32_t d = same(si, tj) ? 0 : std::min((int32_t)std::min(s_quality[si], tq[tj]), (int32_t)QUALITY_CUTOFF);
          q_diff.push_back((int8_t)d);
          si++; tj++;
        }
        if (has_bad_window(q_diff, BAD_WINDOW_LEN, BAD_WINDOW_VALUE)) { stats.bad_short_window++; continue; }
        if (has_bad_window(q_diff, 100, 240)) { stats.bad_long_window++; continue; }
      }
      if (partial) output_partial(s_id, t_id, dir, distinct[i], s_len, t_len);
      else output_overlap(s_id, s_len, dir, t_id, t_len, distinct[i]);
      overlaps_output++;
      if (distinct[i].s_lo == 0) a_olaps_for_frag++;               // :631-635
      if (distinct[i].s_hi >= s_len - 1) b_olaps_for_frag++;
    }
  }
  if (overlaps_output == 0) stats.kmer_hits_without_olap++;
  else { stats.kmer_hits_with_olap++; if (overlaps_output > 1) stats.multi_overlap++; }
}

// Output_Overlap: OIC/overlapInCore-Output.C:75-240
void oracle_ctx::output_overlap(uint32_t s_id, int32_t s_len, int dir, uint32_t t_id, int32_t t_len, const OlapInfo &o) {
  oracle_overlap_t r;
  memset(&r, 0, sizeof(r));
  r.for_utg = 1;
  r.span = (uint32_t)((o.s_hi - o.s_lo) + (o.t_hi - o.t_lo) + o.delta_ct) / 2;
  int32_t s_rh = s_len - o.s_hi - 1, t_rh = t_len - o.t_hi - 1;
  bool sleft = (o.s_lo > o.t_lo) || (o.s_lo == o.t_lo && s_rh > t_rh);
  if (sleft) { r.a_iid = s_id; r.b_iid = t_id; } else { r.a_iid = t_id; r.b_iid = s_id; }
  char orient; int32_t ahg, bhg;
  if (sleft) { orient = (dir == 0) ? 'N' : 'O'; ahg = o.s_lo; bhg = t_rh - s_rh; }
  else       { orient = (dir == 0) ? 'N' : 'I'; ahg = o.t_lo; bhg = s_rh - t_rh; }
  if (orient == 'O' && s_rh >= t_rh) { orient = 'I'; ahg = -(t_rh - s_rh); bhg = -(o.s_lo); }
  r.evalue = encode_evalue(o.quality);
  switch (orient) {
    case 'N': set_a_hang(r, ahg); set_b_hang(r, bhg); r.flipped = 0; break;
    case 'I': set_a_hang(r, ahg); set_b_hang(r, bhg); r.flipped = 1; break;
    default:  set_a_hang(r, -bhg); set_b_hang(r, -ahg); r.flipped = 1; break;   // 'O'
  }
  stats.total_overlaps++;
  if (bhg <= 0) stats.contained++; else stats.dovetail++;
  overlaps.push_back(r);
}

// Output_Partial_Overlap: OIC/overlapInCore-Output.C:244-313
void oracle_ctx::output_partial(uint32_t s_id, uint32_t t_id, int dir, const OlapInfo &o, int32_t s_len, int32_t t_len) {
  oracle_overlap_t r;
  memset(&r, 0, sizeof(r));
  r.a_iid = s_id; r.b_iid = t_id;
  r.for_obt = 1; r.for_dup = 1;
  r.span = (uint32_t)((o.s_hi - o.s_lo) + (o.t_hi - o.t_lo) + o.delta_ct) / 2;
  if (dir == 0) {
    r.ahg5 = o.s_lo; r.ahg3 = s_len - (o.s_hi + 1); r.bhg5 = o.t_lo; r.bhg3 = t_len - (o.t_hi + 1); r.flipped = 0;
  } else {
    r.ahg5 = s_len - (o.s_hi + 1); r.ahg3 = o.s_lo; r.bhg5 = t_len - (o.t_hi + 1); r.bhg3 = o.t_lo; r.flipped = 1;
  }
  r.evalue = encode_evalue(o.quality);
  stats.total_overlaps++;
  overlaps.push_back(r);
}

void oracle_ctx::dump_index() {
  d_keys.clear(); d_occ_start.clear(); d_occ.clear(); d_flags.clear();
  std::vector<uint64_t> keys;
  keys.reserve(index.size());
  for (auto &kv : index) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  for (uint64_t k : keys) {
    const KmerEntry &ke = index[k];
    d_keys.push_back(k);
    d_occ_start.push_back(d_occ.size());
    d_flags.push_back(ke.empty ? 1 : 0);
    for (const Occ &o : ke.occ) d_occ.push_back(((uint64_t)o.str << 32) | o.off);
  }
  d_occ_start.push_back(d_occ.size());
}

extern "C" {

oracle_ctx_t *oracle_create(const oracle_params_t *p, uint32_t n_reads, const uint64_t *offsets, const uint32_t *lengths,
                            const char *bases, const uint8_t *hash_ok, const uint8_t *ref_ok) {
  oracle_ctx *c = new oracle_ctx();
  c->P = *p;
  c->n_reads = n_reads;
  c->lengths.assign(lengths, lengths + n_reads);
  c->offsets.resize(n_reads);
  uint64_t total = 0;
  for (uint32_t i = 0; i < n_reads; i++) total += lengths[i] + 1;
  c->bases.resize(total + 1);
  uint64_t at = 0;
  for (uint32_t i = 0; i < n_reads; i++) {
    c->offsets[i] = at;
    for (uint32_t j = 0; j < lengths[i]; j++) c->bases[at + j] = (char)tolower(bases[offsets[i] + j]);
    c->bases[at + lengths[i]] = 0;
    at += lengths[i] + 1;
  }
  c->hash_ok.assign(n_reads, 1);
  c->ref_ok.assign(n_reads, 1);
  if (hash_ok) c->hash_ok.assign(hash_ok, hash_ok + n_reads);
  if (ref_ok) c->ref_ok.assign(ref_ok, ref_ok + n_reads);
  memset(&c->stats, 0, sizeof(c->stats));
  return c;
}

void oracle_set_qualities(oracle_ctx_t *c, const uint64_t *offsets, const uint8_t *quals) {
  // same layout as the bases (oracle_create): read i at c->offsets[i], one spare byte behind every read
  c->quals.assign(c->bases.size(), 0);
  for (uint32_t i = 0; i < c->n_reads; i++)
    for (uint32_t j = 0; j < c->lengths[i]; j++) c->quals[c->offsets[i] + j] = quals[offsets[i] + j];
  c->use_window_filter = !(c->P.max_erate > 0.06);   // OIC/overlapInCore.C:416-421
}

void oracle_set_skip_kmers(oracle_ctx_t *c, uint64_t n, const char *kmers) {
  c->skip_kmers.clear();
  for (uint64_t i = 0; i < n; i++) c->skip_kmers.push_back(std::string(kmers + i * c->P.kmer_len, c->P.kmer_len));
}

void oracle_set_dump(oracle_ctx_t *c, uint32_t mask) { c->dump_mask = mask; }

// OverlapDriver (OIC/overlapInCore.C:190-300) + Process_Overlaps (OIC/overlapInCore-Process_Overlaps.C:78-181)
int oracle_run(oracle_ctx_t *c) {
  oracle_params_t &P = c->P;
  if (c->ed) delete c->ed;
  c->ed = new EditDist(P.max_erate, P.partial != 0);

  uint32_t g_bgn_hash = P.bgn_hash_id, g_end_hash = P.end_hash_id;
  if (g_bgn_hash < 1) g_bgn_hash = 1;
  if (c->n_reads < g_end_hash) g_end_hash = c->n_reads;

  uint64_t bgn_hash = g_bgn_hash;
  uint64_t end_hash = (uint64_t)g_bgn_hash + P.max_hash_strings - 1;

  while (bgn_hash < g_end_hash) {    // strict '<' : a block of exactly the last read is never built (:222)
    if (end_hash > g_end_hash) end_hash = g_end_hash;
    end_hash = c->build_hash_index((uint32_t)bgn_hash, (uint32_t)end_hash);
    c->stats.n_hash_blocks++;
    c->stats.last_end_hash_id = (uint32_t)end_hash;
    if (c->dump_mask & 4) c->dump_index();

    uint32_t bgn_ref = P.bgn_ref_id, end_ref = P.end_ref_id;
    if (bgn_ref < 1) bgn_ref = 1;
    if (end_ref > c->n_reads) end_ref = c->n_reads;

    // Work is handed out in blocks of perThread reads starting at bgn_ref; a block that would START at end_ref
    // is dropped by the `bgnID < endRefID` guard (Process_Overlaps.C:86), so the last read is skipped exactly
    // when (end_ref - bgn_ref) is a multiple of perThread (OIC/overlapInCore.C:249,264-269).
    uint32_t per_thread = 1 + (end_ref - bgn_ref) / P.num_threads / 8;
    uint32_t last_processed = end_ref;
    if (end_ref >= bgn_ref && (end_ref - bgn_ref) % per_thread == 0) last_processed = end_ref - 1;
    if (end_ref < bgn_ref) last_processed = 0;

    std::string rc;
    std::vector<uint8_t> qrev;
    for (uint32_t fi = bgn_ref; fi <= last_processed && fi >= 1; fi++) {
      if (!c->ref_ok[fi - 1]) continue;
      uint32_t len = c->lengths[fi - 1];
      if ((int64_t)len < (int64_t)P.min_olap_len) continue;
      const char *seq = c->read_seq(fi);
      if (c->use_window_filter) c->s_quality = c->quals.data() + c->offsets[fi - 1];
      c->find_overlaps(seq, (int32_t)len, fi, 0);
      rc.assign(len, 0);
      for (uint32_t i = 0; i < len; i++) rc[i] = rc_char(seq[len - 1 - i]);
      if (c->use_window_filter) {   // reverseComplement(seq, qlt, len) reverses the qualities with the bases (AS_UTL_reverseComplement.C:114-143)
        qrev.assign(len, 0);
        for (uint32_t i = 0; i < len; i++) qrev[i] = c->s_quality[len - 1 - i];
        c->s_quality = qrev.data();
      }
      c->find_overlaps(rc.c_str(), (int32_t)len, fi, 1);
    }

    bgn_hash = end_hash + 1;
    end_hash = bgn_hash + P.max_hash_strings - 1;
  }
  c->stats.ext_cells = c->ed->cells;
  c->stats.ext_words = c->ed->words;
  return 0;
}

uint64_t oracle_num_overlaps(const oracle_ctx_t *c) { return c->overlaps.size(); }
const oracle_overlap_t *oracle_overlaps(const oracle_ctx_t *c) { return c->overlaps.data(); }
const oracle_stats_t *oracle_stats(const oracle_ctx_t *c) { return &c->stats; }
uint64_t oracle_num_pairs(const oracle_ctx_t *c) { return c->d_pairs.size(); }
const oracle_pair_t *oracle_pairs(const oracle_ctx_t *c) { return c->d_pairs.data(); }
uint64_t oracle_num_matches(const oracle_ctx_t *c) { return c->d_matches.size(); }
const oracle_match_t *oracle_matches(const oracle_ctx_t *c) { return c->d_matches.data(); }
uint64_t oracle_num_exts(const oracle_ctx_t *c) { return c->d_exts.size(); }
const oracle_ext_t *oracle_exts(const oracle_ctx_t *c) { return c->d_exts.data(); }
uint64_t oracle_index_num_keys(const oracle_ctx_t *c) { return c->d_keys.size(); }
const uint64_t *oracle_index_keys(const oracle_ctx_t *c) { return c->d_keys.data(); }
const uint64_t *oracle_index_occ_start(const oracle_ctx_t *c) { return c->d_occ_start.data(); }
const uint64_t *oracle_index_occ(const oracle_ctx_t *c) { return c->d_occ.data(); }
const uint8_t *oracle_index_flags(const oracle_ctx_t *c) { return c->d_flags.data(); }
void oracle_destroy(oracle_ctx_t *c) { if (c) { delete c->ed; delete c; } }

int32_t oracle_max_errors(double erate) { return max_errors_for(erate); }
void oracle_match_limit(double erate, int32_t max_errors, int32_t *out) { match_limit(erate, max_errors, out); }
int32_t oracle_error_bound(double erate, int32_t len) { return error_bound(erate, len); }

int32_t oracle_extend(double erate, int32_t partial, const char *S, int32_t s_len, const char *T, int32_t t_len,
                      int32_t m_start, int32_t m_offset, int32_t m_len, int32_t *out, int32_t *delta, int32_t delta_cap,
                      int32_t *delta_len, uint64_t *cells, uint64_t *words) {
  static EditDist *ed = nullptr;
  if (!ed || ed->erate != erate || ed->partial != (partial != 0)) { delete ed; ed = new EditDist(erate, partial != 0); }
  ed->cells = ed->words = 0;
  int32_t kind = ed->extend(S, s_len, T, t_len, m_start, m_offset, m_len, out[0], out[1], out[2], out[3], out[4]);
  *delta_len = (int32_t)ed->left_delta.size();
  for (int32_t i = 0; i < *delta_len && i < delta_cap; i++) delta[i] = ed->left_delta[i];
  if (cells) *cells = ed->cells;
  if (words) *words = ed->words;
  return kind;
}

}  // extern "C"
