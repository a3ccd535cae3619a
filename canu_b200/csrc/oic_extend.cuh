/*
 * oic_extend.cuh -- K3/K4: one warp per read pair: Process_Matches -> Extend_Alignment -> forward / reverse
 * prefix edit distance -> Add_Overlap -> Output_Overlap.
 *
 * Replaces (reference, OIC/ = src/overlapInCore/, LIBO/ = src/overlapInCore/liboverlap/):
 *   Process_String_Olaps / Process_Matches      OIC/overlapInCore-Process_String_Overlaps.C:400-800
 *   prefixEditDistance::Extend_Alignment        LIBO/prefixEditDistance-extend.C:87-229
 *   prefixEditDistance::forward / reverse       LIBO/prefixEditDistance-forward.C:110-285, -reverse.C:130-300
 *   Set_Right_Delta / Set_Left_Delta            LIBO/prefixEditDistance-forward.C:48-87, -reverse.C:52-109
 *   Output_Overlap / Output_Partial_Overlap     OIC/overlapInCore-Output.C:75-313
 *
 * The DP is the "furthest reaching row per diagonal" recurrence: row e holds, for every diagonal d of the
 * pruned band, the longest prefix of A matched with e errors.  The lanes of the warp own consecutive
 * diagonals of the current row; the slide along a diagonal compares 32 bases per step on the 2-bit packed
 * strings; end-of-string detection, band trimming and the best-cell search are warp ballots / reductions.
 * All rows are kept (band-compressed: only the computed extent of each row, plus 2 direction bits per cell) in the
 * warp's private slice of a global scratch arena because the traceback needs them for the delta list; the two
 * live rows are staged in shared memory (oic_extend_fast.cuh).  Integer only, except the
 * branch-point score, which is IEEE double without FMA contraction (OIC_DMUL/OIC_DSUB).
 */
#ifndef OIC_EXTEND_CUH
#define OIC_EXTEND_CUH
#include "oic_search.cuh"

namespace oic {

struct ExtArgs {
  const uint64_t *fwd, *rc;
  uint64_t n_off;             // word offset of the N-mask streams behind the base streams (0: no read holds an N)
  const int64_t *rd_pos;
  const uint32_t *rd_len;
  const uint32_t *qs_read;
  uint32_t first_id, hash_ri_bgn, hash_bgn_id;
  const uint8_t *str_flags;   // bit0 lscreen, bit1 rscreen per hash string
  const uint8_t *qs_screen;   // same per query strand
  Seed *seeds;
  const Pair *pairs;
  const PairInfo *info;
  uint32_t n_pairs;
  const uint32_t *task_list;  // optional indirection (retry pass)
  uint32_t n_tasks;
  // parameters
  int K, partial, unique_olap, use_hopeless, min_olap_len;
  unsigned long long filter_by_kmer_count;
  double erate, minkmer_factor, branch_match_value, min_branch_tail_slope;
  const int32_t *match_limit;   // Edit_Match_Limit[0 .. max_rows)
  int max_rows;
  // per-warp scratch
  int32_t *vals; unsigned long long rowcap;   // row values, rowcap ints per warp
  int4 *rowmeta;                              // max_rows per warp: {lo, tlo, thi, off}
  int32_t *deltas;                            // per warp: stack[max_rows+2] rdelta[max_rows+2] ldelta[2*max_rows+4]
  // output
  uint32_t *out; uint32_t out_cap;            // 6 words per record
  uint32_t *retry_list;
  // speculative round: pairs that still have seeds after their first extension (see k_extend_tasks / k_commit_pairs)
  int handoff;                                // k_process_pairs: 1 = stop after the first extension and hand the pair over
  int first_lean;   // k_process_pairs tries the lean row loop first (0: generic loop only)
  struct PairState *pstate;                   // per pair
  uint32_t *pend_list;                        // pairs handed over
  uint2 *tasks;                               // {pair, seed index}
  struct TaskResult *results;                 // per seed (index = pairs[pi].first + j)
  uint8_t *alive;                             // per seed, scratch of the commit pass
  uint32_t *wide_list;                        // tasks the lean kernel could not run (a row wider than SW)
  int32_t *dpool; unsigned long long dpool_cap;   // delta lists of the speculative results
  // -w (window filter): one quality value per base at the base's stream position (quals[rd_pos[read] + i]); nullptr = filter off
  const uint8_t *quals;
  DevCounters *ctr;
};

struct PairState {        // what k_process_pairs knows about a pair after its first extension
  OlapInfo distinct[MAX_DISTINCT_OLAPS];
  int32_t distinct_ct;
  uint32_t n_rem;         // seeds left, compacted in place at pairs[pi].first
  unsigned long long cells, words;   // work of the first extension (counted when the pair is finalised)
};
struct TaskResult {       // one speculative Extend_Alignment
  int32_t s_lo, s_hi, t_lo, t_hi, errors, kind, dlen;   // kind < 0: the extension could not be run (scratch / pool overflow)
  uint32_t pad;
  unsigned long long delta_off;      // first int of the delta list in dpool (only stored when kind allows pruning)
  unsigned long long cells, words;
};

struct WarpScratch {
  int32_t *vals;
  int4 *meta;
  int32_t *stack, *rdelta, *ldelta;
  unsigned long long rowcap;
  int max_rows;
  const int32_t *match_limit;
  unsigned long long cells, words;   // per-lane partial counters
  long long t_tb;
  // reads with N (generic row loop only): offset of the N-mask stream in 32-bit words, and which string of the current run
  // is the query's REVERSE strand, where the reference holds NUL instead of 'n' (0 none, 1 = A, 2 = T)
  long long n_off32;
  int n_mode;
};

}  // namespace oic
#include "oic_extend_fast.cuh"
#include "oic_extend_lean.cuh"
namespace oic {

// shared-memory path first; the generic global-memory path only for runs with a row wider than SW
/* Extend_Alignment.  kind: 0 NONE, 1 LEFT_BRANCH_PT, 2 RIGHT_BRANCH_PT, 3 DOVETAIL; -1 = scratch overflow.
 * The spliced delta list is left in ws.ldelta[0 .. delta_len). */
/* LSW = 0: generic row loop (dp_run_fast); LSW > 0: lean loop with rows of up to LSW diagonals staged in shared memory */
template <int LSW>
__device__ __forceinline__ int extend_alignment(const ExtArgs &a, WarpScratch &ws, const FastScratch &fs, const uint64_t *S, int64_t sp, int s_len,
                                const uint64_t *T, int64_t tp, int t_len, int m_start, int m_offset, int m_len,
                                int &s_lo, int &s_hi, int &t_lo, int &t_hi, int &errors, int &delta_len) {
  const int l = lane_id();
  int right_errors = 0, left_errors = 0, leftover = 0;
  int r_to_end = 1, l_to_end = 1;
  const int s_left_begin = m_start - 1, s_right_begin = m_start + m_len, s_right_len = s_len - s_right_begin;
  const int t_left_begin = m_offset - 1, t_right_begin = m_offset + m_len, t_right_len = t_len - t_right_begin;
  const int total_olap = imin(m_start, m_offset) + m_len + imin(s_right_len, t_right_len);
  const int err_limit = (int)ceil(OIC_DMUL((double)total_olap, a.erate));
  int rlen = 0, llen = 0;
  bool neg_right = false, neg_left = false;

  // Two runs, right of the seed (forward) then left of it (reverse); in each the string with the shorter tail plays
  // "A" (extend.C:136-187).  One call site (direction is a run-time argument of the shared-memory path) keeps a single
  // copy of the DP in the kernel.
  s_hi = t_hi = s_lo = t_lo = 0;
#pragma unroll 1
  for (int side = 0; side < 2; side++) {
    const bool skip = side == 0 ? (s_right_len == 0 || t_right_len == 0) : (s_left_begin < 0 || t_left_begin < 0);
    if (skip) continue;
    const bool s_is_a = side == 0 ? (s_right_len <= t_right_len) : (s_right_begin <= t_right_begin);
    const int s_from = side == 0 ? s_right_begin : s_left_begin, t_from = side == 0 ? t_right_begin : t_left_begin;
    const int s_n = side == 0 ? s_right_len : s_left_begin + 1, t_n = side == 0 ? t_right_len : t_left_begin + 1;
    const uint64_t *A = s_is_a ? S : T, *B = s_is_a ? T : S;
    const int64_t a0 = s_is_a ? sp + s_from : tp + t_from, b0 = s_is_a ? tp + t_from : sp + s_from;
    const int m = s_is_a ? s_n : t_n, n = s_is_a ? t_n : s_n;
    const int limit = side == 0 ? err_limit : err_limit - right_errors;
    int a_end = 0, b_end = 0, lo = 0, to_end = 1, dl = 0, errs;
    ws.n_mode = (S == a.rc) ? (s_is_a ? 1 : 2) : 0;
    if constexpr (LSW != 0) {
      // the lean loop only walks forward: the leftward run compares the same bases on the reverse-complement streams
      // (base i of a read is base len-1-i of its reverse complement, stored at the same stream offset)
      const uint64_t *Ax = A, *Bx = B;
      int64_t ax = a0, bx = b0;
      if (side == 1) {
        const uint64_t *So = (S == a.fwd) ? a.rc : a.fwd, *To = (T == a.fwd) ? a.rc : a.fwd;
        const int64_t s0 = sp + (s_len - 1 - s_from), t0 = tp + (t_len - 1 - t_from);
        Ax = s_is_a ? So : To; Bx = s_is_a ? To : So;
        ax = s_is_a ? s0 : t0; bx = s_is_a ? t0 : s0;
      }
      errs = dp_run_lean<LSW>(side == 0 ? 1 : -1, a, ws, fs.row0, fs.row1, Ax, ax, m, Bx, bx, n, limit, a_end, b_end, lo, to_end, dl);
    } else errs = dp_run_fast(side == 0 ? 1 : -1, a, ws, fs, A, a0, m, B, b0, n, limit, a_end, b_end, lo, to_end, dl);
    if (errs < 0) return errs;   // -1 history overflow, -2 (lean only) row wider than the shared-memory staging
    if (side == 0) {
      right_errors = errs; r_to_end = to_end; rlen = dl; neg_right = s_is_a;
      s_hi = s_is_a ? a_end : b_end; t_hi = s_is_a ? b_end : a_end;
    } else {
      left_errors = errs; l_to_end = to_end; llen = dl; neg_left = !s_is_a; leftover = lo;
      s_lo = s_is_a ? a_end : b_end; t_lo = s_is_a ? b_end : a_end;
    }
  }
  s_hi += s_right_begin - 1;
  t_hi += t_right_begin - 1;
  s_lo += s_left_begin + 1;
  t_lo += t_left_begin + 1;
  errors = left_errors + right_errors;
  const int kind = !r_to_end ? (!l_to_end ? 0 : 2) : (!l_to_end ? 1 : 3);

  {
    if (neg_left) for (int i = l; i < llen; i += 32) ws.ldelta[i] = -ws.ldelta[i];
    const int rs = neg_right ? -1 : 1;   // sign applied to right_delta before the splice
    for (int i = l; i < rlen; i += 32) {
      const int r = rs * ws.rdelta[i];
      ws.ldelta[llen + i] = (i == 0) ? ((r > 0) ? -(r + leftover + m_len) : -(r - leftover - m_len)) : -r;
    }
  }
  delta_len = llen + rlen;
  __syncwarp();
  return kind;
}

constexpr int EXT_WARPS = 4;
// Two instances of the lean task kernel: staging width x resident blocks per SM.  Measured on B200 (c3 job, round-2 row loop):
// 768 x 6 (80 registers) 64.6 ms; 640 x 7 (72 registers) 61.7 ms; 512 x 8 (64 registers) 65.0 ms plus 3.3 ms of generic reruns
// (105 tasks meet a row wider than 512); 768 x 5 (96 registers, hardly any spill) 74.8 ms; 1024 x 4 69.2 ms.  What is left of
// the 256 KB L1/shared memory for the sequence windows matters as much as the resident warps.  Low error rates keep the band
// narrow (no row beyond 640 diagonals at --maxerate 0.06 on 5-15 kb reads); at high rates the wide instance runs.
#ifndef OIC_SW_NARROW
#define OIC_SW_NARROW 640
#endif
#ifndef OIC_TASK_BLOCKS
#define OIC_TASK_BLOCKS 7
#endif
constexpr int LSW_NARROW = OIC_SW_NARROW, LSW_WIDE = SW;
constexpr int TASK_BLOCKS_NARROW = OIC_TASK_BLOCKS, TASK_BLOCKS_WIDE = 4;

struct PairCtx {   // the strings and IDs of one pair
  const uint64_t *S, *T;
  int64_t sp, tp;
  int s_len, t_len, dir;
  uint32_t s_id, t_id;
};
__device__ __forceinline__ PairCtx pair_ctx(const ExtArgs &a, const Pair &pr) {
  PairCtx x;
  const uint32_t ri = a.qs_read[pr.qs >> 1];
  x.dir = pr.qs & 1;
  x.S = x.dir ? a.rc : a.fwd;
  x.sp = a.rd_pos[ri];
  x.s_len = (int)a.rd_len[ri];
  x.s_id = a.first_id + ri;
  const uint32_t hri = a.hash_ri_bgn + pr.str;
  x.T = a.fwd;
  x.tp = a.rd_pos[hri];
  x.t_len = (int)a.rd_len[hri];
  x.t_id = a.hash_bgn_id + pr.str;
  return x;
}

__device__ __forceinline__ WarpScratch warp_scratch(const ExtArgs &a, uint32_t gw) {
  WarpScratch ws;
  ws.vals = a.vals + (unsigned long long)gw * a.rowcap;
  ws.meta = a.rowmeta + (unsigned long long)gw * a.max_rows;
  int32_t *dbase = a.deltas + (unsigned long long)gw * (4ull * a.max_rows + 8);
  ws.stack = dbase;
  ws.rdelta = dbase + (a.max_rows + 2);
  ws.ldelta = dbase + 2 * (a.max_rows + 2);
  ws.rowcap = a.rowcap;
  ws.max_rows = a.max_rows;
  ws.match_limit = a.match_limit;
  ws.cells = ws.words = 0;
  ws.t_tb = 0;
  ws.n_off32 = 2 * (long long)a.n_off;
  ws.n_mode = 0;
  return ws;
}

/* first of the longest seeds, in list order (Process_String_Overlaps.C:473-480); alive == nullptr: all of them */
__device__ __forceinline__ int longest_seed(const Seed *seeds, const uint8_t *alive, int n) {
  const int l = lane_id();
  long long bk = -1;
  for (int i = l; i < n; i += 32) {
    if (alive && !alive[i]) continue;
    const long long k = (long long)seeds[i].len * 4294967296LL + (long long)(0x7fffffff - i);
    if (k > bk) bk = k;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const long long ok = __shfl_xor_sync(FULL, bk, o);
    if (ok > bk) bk = ok;
  }
  return bk < 0 ? -1 : 0x7fffffff - (int)(bk & 0xffffffffLL);
}

/* the accept test of Process_Matches (Process_String_Overlaps.C:488-512) */
/* returns the slot of `distinct` whose alignment this extension became, or -1 */
__device__ __forceinline__ int accept_overlap(const ExtArgs &a, int kind, int s_lo, int s_hi, int t_lo, int t_hi, int errors, int dlen,
                                              OlapInfo *distinct, int &distinct_ct) {
  if (kind == 3 || a.partial) {
    if (1 + s_hi - s_lo >= a.min_olap_len && 1 + t_hi - t_lo >= a.min_olap_len) {
      const int olap_len = 1 + imin(s_hi - s_lo, t_hi - t_lo);
      const double quality = OIC_DDIV((double)errors, (double)olap_len);
      if (errors <= (int)ceil(OIC_DMUL((double)olap_len, a.erate)))
        return add_overlap(a.partial, s_lo, s_hi, t_lo, t_hi, quality, dlen, distinct, distinct_ct);
    }
  }
  return -1;
}

/* Combine_Into_One_Olap / Merge_Intersecting_Olaps / Choose_Best_Partial (Process_String_Overlaps.C:540-556): which of the
 * distinct overlaps survive.  Scalar, the same on every lane that calls it. */
__device__ __forceinline__ void reduce_olaps(const ExtArgs &a, OlapInfo *distinct, int distinct_ct, int *deleted) {
  if (a.partial) { if (a.unique_olap) choose_best_partial(distinct, distinct_ct, deleted); }
  else { if (a.unique_olap) combine_into_one(distinct, distinct_ct, deleted); else merge_intersecting(distinct, distinct_ct, deleted); }
}

/* end of Process_Matches for one pair: write the records of the overlaps that are not deleted, bump the -s counters
 * (Process_String_Overlaps.C:558-650, Output.C).  Lane 0 only.
 * a_inc / b_inc (-l only): overlaps output that reach the 5' / the 3' end of the query strand (:631-635). */
__device__ __forceinline__ void emit_olaps(const ExtArgs &a, const PairCtx &x, OlapInfo *distinct, int distinct_ct, const int *deleted, int *a_inc,
                                           int *b_inc) {
  int overlaps_output = 0;
  if (distinct_ct > 0) {
    int nout = 0;
    for (int i = 0; i < distinct_ct; i++) nout += !deleted[i];
    uint32_t ob = nout ? atomicAdd(&a.ctr->n_out, (unsigned)nout) : 0u;
    for (int i = 0; i < distinct_ct; i++) {
      if (deleted[i]) continue;
      if (ob >= a.out_cap) { atomicOr(&a.ctr->out_overflow, 1u); break; }
      uint32_t *w = a.out + 6ull * ob;
      if (a.partial) output_partial(w, x.s_id, x.t_id, x.dir, distinct[i], x.s_len, x.t_len);
      else {
        const int contained = output_overlap(w, x.s_id, x.s_len, x.dir, x.t_id, x.t_len, distinct[i]);
        if (contained) atomicAdd(&a.ctr->contained, 1ull); else atomicAdd(&a.ctr->dovetail, 1ull);
      }
      ob++;
      overlaps_output++;
      if (a_inc && distinct[i].s_lo == 0) (*a_inc)++;
      if (b_inc && distinct[i].s_hi >= x.s_len - 1) (*b_inc)++;
    }
    atomicAdd(&a.ctr->total_overlaps, (unsigned long long)overlaps_output);
  }
  if (overlaps_output == 0) atomicAdd(&a.ctr->kmer_hits_without_olap, 1ull);
  else {
    atomicAdd(&a.ctr->kmer_hits_with_olap, 1ull);
    if (overlaps_output > 1) atomicAdd(&a.ctr->multi_overlap, 1ull);
  }
}

__device__ __forceinline__ void finalize_pair(const ExtArgs &a, const PairCtx &x, OlapInfo *distinct, int distinct_ct, int *a_inc = nullptr,
                                              int *b_inc = nullptr) {
  int deleted[MAX_DISTINCT_OLAPS] = {0, 0, 0};
  if (distinct_ct > 0) reduce_olaps(a, distinct, distinct_ct, deleted);
  emit_olaps(a, x, distinct, distinct_ct, deleted, a_inc, b_inc);
}

/* ---- -w, the window filter (OIC/overlapInCore.C:398,417-419; Process_String_Overlaps.C:560-620, Has_Bad_Window :365-388)
 * An overlap about to be output is walked along its alignment: every mismatch contributes min(query quality, hash-read quality,
 * 20), every indel min(quality of the extra base, 20); the overlap is rejected when some 50 consecutive positions sum to 160 or
 * more ("short window"), else when some 100 sum to 240 or more ("long window").
 * The alignment is the delta list of the extension that set the overlap's slot.  The reference copies that list into the slot;
 * here the slot remembers the SEED of that extension and the extension is run again when the overlap is output (deterministic, one
 * more extension per output overlap, -w only) -- no delta lists are kept.  The walk itself is serial (lane 0). */
__device__ __forceinline__ bool base_same(const PairCtx &x, uint64_t n_off, int i, int j) {
  const int64_t ps = x.sp + i, pt = x.tp + j;
  if (n_off) {
    const bool sn = ((x.S[(ps >> 5) + n_off] >> ((ps & 31) * 2)) & 3u) != 0, tn = ((x.T[(pt >> 5) + n_off] >> ((pt & 31) * 2)) & 3u) != 0;
    if (tn) return true;                 // T[j] == 'n' matches anything (:578)
    if (sn) return x.dir == 0;           // S[i] == 'n' on the forward strand; on the reverse strand the reference holds NUL there, which equals nothing
  }
  return (((x.S[ps >> 5] >> ((ps & 31) * 2)) ^ (x.T[pt >> 5] >> ((pt & 31) * 2))) & 3u) == 0;
}
/* 0 = keep, 1 = rejected by the short window, 2 = by the long window.  delta[0 .. delta_ct) = the alignment's delta list. */
__device__ __forceinline__ int window_filter(const ExtArgs &a, const PairCtx &x, const OlapInfo &p, const int32_t *delta, int delta_ct, signed char *ring) {
  // query qualities are reversed with the bases on the reverse strand (reverseComplement(bases, quals, len), Process_Overlaps.C:139)
  const uint8_t *sq = a.quals + x.sp, *tq = a.quals + x.tp;
  auto s_q = [&](int i) { return (int)sq[x.dir ? x.s_len - 1 - i : i]; };
  WindowScan ws;
  ws.reset();
  int i = p.s_lo, j = p.t_lo;
  for (int k = 0; k < delta_ct; k++) {
    const int dl = delta[k], len = iabs(dl);
    for (int n = 1; n < len; n++) {
      ws.push(ring, base_same(x, a.n_off, i, j) ? 0 : imin(imin(s_q(i), (int)tq[j]), QUALITY_CUTOFF));
      i++; j++;
    }
    int d;
    if (dl > 0) { d = s_q(i); i++; } else { d = (int)tq[j]; j++; }
    ws.push(ring, imin(d, QUALITY_CUTOFF));
  }
  while (i <= p.s_hi) {
    ws.push(ring, base_same(x, a.n_off, i, j) ? 0 : imin(imin(s_q(i), (int)tq[j]), QUALITY_CUTOFF));
    i++; j++;
  }
  return (ws.bad & 1) ? 1 : ((ws.bad & 2) ? 2 : 0);
}

/*
 * Pass 1 of the pair stage: one warp per pair (atomic work queue, heavy pairs first).
 * a.handoff == 0: the whole Process_Matches loop (used for the retry list, with worst-case scratch).
 * a.handoff == 1: the filters and the FIRST extension only.  Extensions of one pair do not depend on each other --
 *   only whether a seed is still in the list depends on the alignments accepted before it -- so a pair that still
 *   has seeds after its first extension is handed over: its remaining seeds are all extended speculatively, one warp
 *   each, by k_extend_tasks, and k_commit_pairs then replays the loop over the stored results.  This removes the serial
 *   chain of a pair (up to ~100 extensions of the same 15 kb overlap) from the critical path of the stage.
 */
constexpr int PP_DONE = 0, PP_OVERFLOW = 1, PP_HANDED = 2;

/* Process_Matches for pair pi by one warp (filters, the extension loop, seed pruning, output).  LIMIT = false: the default path
 * (a.handoff decides whether the pair is parked after its first extension).  LIMIT = true: the -l path (Process_String_Overlaps.C:
 * 482-490): a_full / b_full say that the 5' / 3' end of the query strand already has its -l overlaps; a seed whose extension
 * would add to a full end is not extended -- it only removes itself, because Kind_Of_Olap and the alignment bounds keep the
 * values of the previous iteration (:415) and re-applying that alignment's pruning removes nothing new.  (With -G a skipped FIRST
 * seed makes the reference read uninitialised bounds; here, as in the oracle, they are zero and nothing else is pruned.) */
#ifndef OIC_FIRST_LEAN
#define OIC_FIRST_LEAN 896   /* staging width of the lean row loop inside k_process_pairs (0: first extensions run the generic loop) */
#endif
template <bool LIMIT, int LSWF = 0>
__device__ __forceinline__ int process_pair(const ExtArgs &a, WarpScratch &ws, const FastScratch &fs, uint32_t pi, unsigned long long &ext_calls,
                                            bool a_full, bool b_full, int &a_inc, int &b_inc) {
  const int l = lane_id();
  a_inc = b_inc = 0;
  const Pair pr = a.pairs[pi];
  const PairInfo inf = a.info[pi];
  const PairCtx x = pair_ctx(a, pr);
  Seed *seeds = a.seeds + pr.first;
  int n = (int)pr.n;

  // --minkmers filter: computeMinimumKmers (Process_String_Overlaps.C:75-86,725)
  if (a.filter_by_kmer_count != 0) {
    double ovl = (double)(inf.diag_end - inf.diag_bgn);
    if (ovl < 0) ovl = -ovl;
    unsigned long long expected = 0;
    if (!(ovl < (double)a.K)) expected = (unsigned long long)(int)floor(OIC_DMUL(a.minkmer_factor, ovl - a.K + 1));
    unsigned long long need = a.filter_by_kmer_count > expected ? a.filter_by_kmer_count : expected;
    if (need > (unsigned long long)inf.diag_ct) {
      if (l == 0) atomicAdd(&a.ctr->kmer_hits_skipped, 1ull);
      return PP_DONE;
    }
  }
  // hopeless check for a singleton seed far from both ends (Process_String_Overlaps.C:433-464)
  if (a.use_hopeless && n == 1 && !a.partial) {
    const Seed m = seeds[0];
    const unsigned qsc = a.qs_screen[pr.qs], hsc = a.str_flags[pr.str];
    bool hopeless = false;
    const int s_head = m.start, t_head = m.offset;
    if (s_head <= t_head) { if (s_head > HOPELESS_MATCH && !(qsc & 1)) hopeless = true; }
    else { if (t_head > HOPELESS_MATCH && !(hsc & 1)) hopeless = true; }
    const int s_tail = x.s_len - s_head - m.len + 1, t_tail = x.t_len - t_head - m.len + 1;
    if (s_tail <= t_tail) { if (s_tail > HOPELESS_MATCH && !(qsc & 2)) hopeless = true; }
    else { if (t_tail > HOPELESS_MATCH && !(hsc & 2)) hopeless = true; }
    if (hopeless) {
      if (l == 0) atomicAdd(&a.ctr->kmer_hits_without_olap, 1ull);
      return PP_DONE;
    }
  }

  // a pair with an N in either read stays here for its whole loop: only this kernel's generic row loop knows the N masks
  bool pair_n = false;
  if (a.n_off) {
    uint64_t acc = 0;
    const int64_t sw = (x.sp >> 5) + (int64_t)a.n_off, tw = (x.tp >> 5) + (int64_t)a.n_off;
    for (int w = l; w * 32 < x.s_len; w += 32) acc |= x.S[sw + w];
    for (int w = l; w * 32 < x.t_len; w += 32) acc |= x.T[tw + w];
    pair_n = __any_sync(FULL, acc != 0);
  }

  OlapInfo distinct[MAX_DISTINCT_OLAPS];
  Seed olap_seed[MAX_DISTINCT_OLAPS];   // -w only
  int distinct_ct = 0;
  bool overflow = false, handed = false;
  const unsigned long long snap_cells = ws.cells, snap_words = ws.words, snap_calls = ext_calls;
  int kind = 0, s_lo = 0, s_hi = 0, t_lo = 0, t_hi = 0, errors = 0, dlen = 0;   // (LIMIT: a skipped seed sees the previous iteration's)

  while (n > 0) {
    const int li = longest_seed(seeds, nullptr, n);
    const Seed lg = seeds[li];
    bool skip = false;
    if (LIMIT) {
      const int a_hang = lg.start - lg.offset, b_hang = a_hang + x.s_len - x.t_len;
      skip = (a_full && a_hang <= 0) || (b_full && b_hang <= 0);
    }
    if (!skip) {
      ext_calls++;
      bool done = false;
      if constexpr (LSWF != 0) {
        // the lean row loop first (same results, about two thirds of the instructions); it knows no N masks and no row wider
        // than its staging: those extensions run again in the generic loop, counters rewound
        if (!pair_n && a.first_lean) {
          const unsigned long long c0 = ws.cells, w0 = ws.words;
          kind = extend_alignment<LSWF>(a, ws, fs, x.S, x.sp, x.s_len, x.T, x.tp, x.t_len, lg.start, lg.offset, lg.len, s_lo, s_hi, t_lo, t_hi, errors, dlen);
          done = kind != LEAN_WIDE;
          if (!done) { ws.cells = c0; ws.words = w0; }
        }
      }
      if (!done) kind = extend_alignment<0>(a, ws, fs, x.S, x.sp, x.s_len, x.T, x.tp, x.t_len, lg.start, lg.offset, lg.len, s_lo, s_hi, t_lo, t_hi, errors, dlen);
      if (kind < 0) { overflow = true; break; }
      const int slot = accept_overlap(a, kind, s_lo, s_hi, t_lo, t_hi, errors, dlen, distinct, distinct_ct);
      if (a.quals && slot >= 0) olap_seed[slot] = lg;   // -w: the slot's alignment (and delta list) is this seed's extension
    }
    if (inf.consistent) break;

    // drop the extended seed and every seed that lies on this alignment (Process_String_Overlaps.C:514-531)
    int newn = 0;
    for (int i0 = 0; i0 < n; i0 += 32) {
      const int i = i0 + l;
      bool keep = false;
      Seed nd;
      if (i < n) {
        nd = seeds[i];
        bool remove = (i == li) ||
                      ((kind == 3 || a.partial) && s_lo - SHIFT_SLACK <= nd.start && nd.start + nd.len <= s_hi + SHIFT_SLACK &&
                       lies_on_alignment(nd.start, nd.offset, s_lo, t_lo, ws.ldelta, dlen));
        keep = !remove;
      }
      const unsigned km = __ballot_sync(FULL, keep);
      __syncwarp();
      if (keep) seeds[newn + __popc(km & ((1u << l) - 1))] = nd;
      newn += __popc(km);
      __syncwarp();
    }
    n = newn;
    if (!LIMIT && a.handoff && n > 0 && !pair_n) { handed = true; break; }
  }

  if (overflow) {   // the retry pass redoes this pair from scratch: forget its partial work counters
    ws.cells = snap_cells; ws.words = snap_words; ext_calls = snap_calls;
    return PP_OVERFLOW;
  }
  if (handed) {
    // park the pair: its first extension is counted when (and if) k_commit_pairs finalises it
    const unsigned long long cw = warp_sum64(ws.words - snap_words);
    PairState *ps = a.pstate + pi;
    if (l == 0) {
      for (int i = 0; i < distinct_ct; i++) ps->distinct[i] = distinct[i];
      ps->distinct_ct = distinct_ct;
      ps->n_rem = (uint32_t)n;
      ps->cells = ws.cells - snap_cells;
      ps->words = cw;
      a.pend_list[atomicAdd(&a.ctr->n_pend, 1u)] = pi;
    }
    uint32_t tb = 0;
    if (l == 0) tb = atomicAdd(&a.ctr->n_spec_tasks, (unsigned)n);
    tb = __shfl_sync(FULL, tb, 0);
    for (int i = l; i < n; i += 32) a.tasks[tb + i] = make_uint2(pi, (uint32_t)i);
    ws.cells = snap_cells; ws.words = snap_words; ext_calls = snap_calls;
    return PP_HANDED;
  }
  if (a.quals && distinct_ct > 0) {
    // -w: the survivors of the reduction go through the window filter before they are output (:558-620)
    int deleted[MAX_DISTINCT_OLAPS] = {0, 0, 0};
    reduce_olaps(a, distinct, distinct_ct, deleted);   // (every lane holds the same `distinct`)
    const unsigned long long keep_cells = ws.cells, keep_words = ws.words;   // the re-extensions below are not the reference's work
    for (int i = 0; i < distinct_ct; i++) {
      if (deleted[i]) continue;
      const Seed sd = olap_seed[i];
      int s_lo2, s_hi2, t_lo2, t_hi2, errors2, dlen2;
      if (extend_alignment<0>(a, ws, fs, x.S, x.sp, x.s_len, x.T, x.tp, x.t_len, sd.start, sd.offset, sd.len, s_lo2, s_hi2, t_lo2, t_hi2, errors2, dlen2) < 0) {
        ws.cells = snap_cells; ws.words = snap_words; ext_calls = snap_calls;
        return PP_OVERFLOW;   // (it fitted the first time: cannot happen)
      }
      int rej = 0;
      if (l == 0) {
        signed char ring[BAD_LONG_WINDOW_LEN];
        rej = window_filter(a, x, distinct[i], ws.ldelta, dlen2, ring);
        if (rej) atomicAdd(rej == 1 ? &a.ctr->bad_short_window : &a.ctr->bad_long_window, 1ull);
      }
      rej = __shfl_sync(FULL, rej, 0);
      if (rej) deleted[i] = 1;
    }
    ws.cells = keep_cells; ws.words = keep_words;
    if (l == 0) emit_olaps(a, x, distinct, distinct_ct, deleted, LIMIT ? &a_inc : nullptr, LIMIT ? &b_inc : nullptr);
  } else if (l == 0) finalize_pair(a, x, distinct, distinct_ct, LIMIT ? &a_inc : nullptr, LIMIT ? &b_inc : nullptr);
  if (LIMIT) { a_inc = __shfl_sync(FULL, a_inc, 0); b_inc = __shfl_sync(FULL, b_inc, 0); }
  return PP_DONE;
}

__global__ void __launch_bounds__(EXT_WARPS * 32, 4) k_process_pairs(ExtArgs a) {
  const int l = lane_id();
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  WarpScratch ws = warp_scratch(a, gw);
  unsigned long long ext_calls = 0;
  __shared__ int32_t s_rows[EXT_WARPS][2][SW + 4];
  FastScratch fs;
  fs.row0 = s_rows[threadIdx.x >> 5][0];
  fs.row1 = s_rows[threadIdx.x >> 5][1];

  while (true) {
    uint32_t ti = 0;
    if (l == 0) ti = atomicAdd(&a.ctr->pair_cursor, 1u);
    ti = __shfl_sync(FULL, ti, 0);
    if (ti >= a.n_tasks) break;
    const uint32_t pi = a.task_list ? a.task_list[ti] : ti;
    int a_inc, b_inc;
    if (process_pair<false, (OIC_FIRST_LEAN + 8 + 32 * LU <= SW + 4) ? OIC_FIRST_LEAN : 0>(a, ws, fs, pi, ext_calls, false, false, a_inc, b_inc) == PP_OVERFLOW && l == 0)
      a.retry_list[atomicAdd(&a.ctr->n_retry, 1u)] = pi;
  }

  unsigned long long c = warp_sum64(ws.cells), w = warp_sum64(ws.words);
  if (l == 0) {
    if (c) atomicAdd(&a.ctr->ext_cells, c);
    if (w) atomicAdd(&a.ctr->ext_words, w);
    if (ext_calls) atomicAdd(&a.ctr->ext_calls, ext_calls);
  }
}

/* ---- the -l path: a per-end cap on the overlaps of a query strand (OIC/overlapInCore.C:345-347, Process_String_Overlaps.C:743-797)
 * With -l the partners of a strand are no longer independent: whether a partner is processed at all, and which of its seeds
 * are extended, depends on how many overlaps the partners BEFORE it have put on each end of the read.  So this path gives a
 * strand to ONE warp, which walks the partners in the reference's order with the two counters in registers; the extensions of a
 * partner run serially (process_pair<true>).  Strands are still independent of each other, which is where the parallelism is.
 *
 * The order (Process_String_Olaps): the String_Olap array is compacted in index order -- the 256 primary slots
 * (StrNum ^ StrNum >> 8) & 255, first come first served, then the overflow nodes in creation order (Find_Overlaps.C:163-187) -- and,
 * when there are more partners than the limit, sorted by mean diagonal with qsort(By_Diag_Sum): glibc's qsort is a merge sort,
 * so partners on the same mean diagonal keep their array order (pinned against the reference binary on this image, glibc 2.39).
 * Then: from diagonal zero upward while the 5' end has room, then from just below zero downward while the 3' end has room.
 * Creation order of a strand's partners = arrival order of their first k-mer hit = (query offset ascending, hash string
 * descending: the chain order of one k-mer's occurrences), read off the first run of every pair.
 * Ranks are found by counting (O(partners^2 / 32) per strand): fine for the tens to hundreds of partners of a read, slow for a
 * read with many thousands -- this is the -l path only, which the pipeline never takes. */
struct StrandArgs {
  const Seed *src;                    // runs in arrival order, grouped by pair (output of k_group_runs)
  uint32_t *strand_first, *strand_n;  // per listed strand: its pair range (pairs of a strand are contiguous)
  unsigned long long *ckey;           // [n_pairs] creation key
  double *mean;                       // [n_pairs] mean diagonal
  uint32_t *slot, *rank, *by_creation, *order;   // [n_pairs]
  unsigned long long limit;           // G.Frag_Olap_Limit
};

__global__ void __launch_bounds__(256) k_strand_ranges(const Pair *__restrict__ pairs, uint32_t n_pairs, StrandArgs sa, DevCounters *ctr) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pairs) return;
  const uint32_t qs = pairs[i].qs;
  if (i > 0 && pairs[i - 1].qs == qs) return;
  uint32_t n = 1;
  while (i + n < n_pairs && pairs[i + n].qs == qs) n++;
  const uint32_t w = atomicAdd(&ctr->n_strands, 1u);
  sa.strand_first[w] = i;
  sa.strand_n[w] = n;
}

__global__ void __launch_bounds__(EXT_WARPS * 32, 4) k_process_strands(ExtArgs a, StrandArgs sa) {
  const int l = lane_id();
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  WarpScratch ws = warp_scratch(a, gw);
  unsigned long long ext_calls = 0;
  __shared__ int32_t s_rows[EXT_WARPS][2][SW + 4];
  FastScratch fs;
  fs.row0 = s_rows[threadIdx.x >> 5][0];
  fs.row1 = s_rows[threadIdx.x >> 5][1];
  const uint32_t n_strands = a.ctr->n_strands;

  while (true) {
    uint32_t si = 0;
    if (l == 0) si = atomicAdd(&a.ctr->pair_cursor, 1u);
    si = __shfl_sync(FULL, si, 0);
    if (si >= n_strands) break;
    const uint32_t p0 = sa.strand_first[si], np = sa.strand_n[si];
    // mean diagonal and creation key of every partner
    for (uint32_t i = l; i < np; i += 32) {
      const Pair pr = a.pairs[p0 + i];
      const Seed *runs = sa.src + pr.first;
      long long sum = 0;                                   // sum of (hash offset - query offset) over the k-mer hits: exact
      for (uint32_t j = 0; j < pr.n; j++) sum += (long long)(runs[j].offset - runs[j].start) * (long long)(runs[j].len - a.K + 1);
      sa.mean[p0 + i] = OIC_DDIV((double)sum, (double)a.info[p0 + i].diag_ct);
      sa.ckey[p0 + i] = ((unsigned long long)(uint32_t)runs[0].start << 32) | (0xffffffffu - pr.str);
    }
    __syncwarp();
    for (uint32_t i = l; i < np; i += 32) {
      const unsigned long long k = sa.ckey[p0 + i];
      uint32_t r = 0;
      for (uint32_t j = 0; j < np; j++) r += sa.ckey[p0 + j] < k;
      sa.by_creation[p0 + r] = i;
    }
    __syncwarp();
    if (l == 0) {   // String_Olap slots, in creation order
      uint32_t used[8] = {0, 0, 0, 0, 0, 0, 0, 0}, n_over = 0;
      for (uint32_t r = 0; r < np; r++) {
        const uint32_t i = sa.by_creation[p0 + r], str = a.pairs[p0 + i].str, sub = (str ^ (str >> 8)) & 255u;
        if (!((used[sub >> 5] >> (sub & 31)) & 1u)) { used[sub >> 5] |= 1u << (sub & 31); sa.slot[p0 + i] = sub; }
        else sa.slot[p0 + i] = 256u + n_over++;
      }
    }
    __syncwarp();
    for (uint32_t i = l; i < np; i += 32) {   // position in the compacted array
      const uint32_t k = sa.slot[p0 + i];
      uint32_t r = 0;
      for (uint32_t j = 0; j < np; j++) r += sa.slot[p0 + j] < k;
      sa.rank[p0 + i] = r;
    }
    __syncwarp();
    const bool all = (unsigned long long)np <= sa.limit;   // few partners: all of them, in array order (:722-741)
    uint32_t n_neg = 0;
    for (uint32_t i = l; i < np; i += 32) {
      uint32_t r = sa.rank[p0 + i];
      if (!all) {                                          // stable order by mean diagonal
        const double m = sa.mean[p0 + i];
        const uint32_t mine = r;
        r = 0;
        for (uint32_t j = 0; j < np; j++) {
          const double mj = sa.mean[p0 + j];
          r += (mj < m) || (mj == m && sa.rank[p0 + j] < mine);
        }
        n_neg += m < 0.0;
      }
      sa.order[p0 + r] = i;
    }
    n_neg = warp_sum(n_neg);
    __syncwarp();

    int A = 0, B = 0;
    bool failed = false;
    auto one = [&](uint32_t r) {
      int a_inc, b_inc;
      const uint32_t pi = p0 + sa.order[p0 + r];
      if (process_pair<true>(a, ws, fs, pi, ext_calls, (unsigned long long)A >= sa.limit, (unsigned long long)B >= sa.limit, a_inc, b_inc) != PP_DONE) failed = true;
      A += a_inc; B += b_inc;
    };
    if (all) for (uint32_t r = 0; r < np && !failed; r++) one(r);
    else {
      for (uint32_t r = n_neg; r < np && (unsigned long long)A < sa.limit && !failed; r++) one(r);
      for (uint32_t r = n_neg; r > 0 && (unsigned long long)B < sa.limit && !failed; r--) one(r - 1);
    }
    if (failed && l == 0) atomicAdd(&a.ctr->n_retry, 1u);   // cannot happen with worst-case scratch; the host reports it
  }

  unsigned long long c = warp_sum64(ws.cells), w = warp_sum64(ws.words);
  if (l == 0) {
    if (c) atomicAdd(&a.ctr->ext_cells, c);
    if (w) atomicAdd(&a.ctr->ext_words, w);
    if (ext_calls) atomicAdd(&a.ctr->ext_calls, ext_calls);
  }
}

/*
 * Pass 2: one warp per (pair, remaining seed): a single Extend_Alignment, result + delta list stored for the commit pass.
 * Tasks are taken from the END of the list: heavy pairs start first in pass 1 but finish (and emit their tasks) last.
 * LSW > 0: the shared-memory-only row loop; a task that meets a row wider than LSW is queued for the LSW = 0
 * instance of the kernel (generic row loop), which runs right after on just those tasks.
 */
template <int LSW, int BLOCKS>
__global__ void __launch_bounds__(EXT_WARPS * 32, BLOCKS) k_extend_tasks(ExtArgs a) {
  constexpr bool LEAN = LSW != 0;
  const int l = lane_id();
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  WarpScratch ws = warp_scratch(a, gw);
  __shared__ int32_t s_rows[EXT_WARPS][2][LEAN ? LSW + 8 + 32 * LU : SW + 4];
  FastScratch fs;
  fs.row0 = s_rows[threadIdx.x >> 5][0];
  fs.row1 = s_rows[threadIdx.x >> 5][1];
  const uint32_t n_all = a.ctr->n_spec_tasks;
  const uint32_t n_tasks = LEAN ? n_all : a.ctr->n_wide_tasks;

  while (true) {
    uint32_t ti = 0;
    if (l == 0) ti = atomicAdd(LEAN ? &a.ctr->task_cursor : &a.ctr->wide_cursor, 1u);
    ti = __shfl_sync(FULL, ti, 0);
    if (ti >= n_tasks) break;
    const uint32_t tidx = LEAN ? n_all - 1 - ti : a.wide_list[ti];
    const uint2 tk = a.tasks[tidx];
    const Pair pr = a.pairs[tk.x];
    const PairCtx x = pair_ctx(a, pr);
    const Seed sd = a.seeds[pr.first + tk.y];
    ws.cells = ws.words = 0;
    TaskResult r;
    r.pad = 0; r.delta_off = 0;
    r.kind = extend_alignment<LSW>(a, ws, fs, x.S, x.sp, x.s_len, x.T, x.tp, x.t_len, sd.start, sd.offset, sd.len, r.s_lo, r.s_hi, r.t_lo, r.t_hi,
                                    r.errors, r.dlen);
    if (LEAN && r.kind == LEAN_WIDE) {
      if (l == 0) a.wide_list[atomicAdd(&a.ctr->n_wide_tasks, 1u)] = tidx;
      continue;
    }
    if (r.kind == 3 || (a.partial && r.kind >= 0)) {   // the delta list is only read to prune seeds on an accepted-kind alignment
      unsigned long long off = 0;
      if (l == 0) off = atomicAdd(&a.ctr->dpool_used, (unsigned long long)r.dlen);
      off = __shfl_sync(FULL, off, 0);
      if (off + (unsigned long long)r.dlen > a.dpool_cap) r.kind = -1;
      else {
        r.delta_off = off;
        for (int i = l; i < r.dlen; i += 32) a.dpool[off + i] = ws.ldelta[i];
      }
    }
    r.words = warp_sum64(ws.words);
    r.cells = ws.cells;   // lane 0 holds the count
    if (l == 0) a.results[pr.first + tk.y] = r;
  }
}

/*
 * Pass 3: one warp per handed-over pair: the rest of the Process_Matches loop over the stored results.  A result is
 * consumed only if its seed is still in the list when its turn comes, exactly as the serial loop would have extended it;
 * results of seeds pruned before their turn are discarded (speculation waste).  Any unusable result sends the pair to
 * the retry list (full serial loop with worst-case scratch); nothing of the pair has been output or counted until here.
 */
constexpr int COMMIT_DCAP = 1024;   // longest delta list indexed in shared memory (longer ones are walked serially)

__global__ void __launch_bounds__(128) k_commit_pairs(ExtArgs a) {
  // Lies_On_Alignment (Process_String_Overlaps.C:307-326) walks the delta list until the alignment passes the seed's start
  // and compares diagonals there.  The walk is the same for every seed, so it is done once per alignment as a warp scan:
  // s_pos[i] = query position at which delta i is tested (strictly increasing), s_diag[i] = diagonal at that test;
  // each seed then finds its delta by binary search.
  __shared__ int32_t s_pos[4][COMMIT_DCAP], s_diag[4][COMMIT_DCAP + 1];
  const int l = lane_id(), wib = threadIdx.x >> 5;
  const uint32_t n_pend = a.ctr->n_pend;
  while (true) {
    uint32_t ti = 0;
    if (l == 0) ti = atomicAdd(&a.ctr->pend_cursor, 1u);
    ti = __shfl_sync(FULL, ti, 0);
    if (ti >= n_pend) break;
    const uint32_t pi = a.pend_list[ti];
    const Pair pr = a.pairs[pi];
    const PairCtx x = pair_ctx(a, pr);
    const PairState *ps = a.pstate + pi;
    const Seed *seeds = a.seeds + pr.first;
    const TaskResult *res = a.results + pr.first;
    uint8_t *alive = a.alive + pr.first;
    const int n = (int)ps->n_rem;
    OlapInfo distinct[MAX_DISTINCT_OLAPS];
    int distinct_ct = ps->distinct_ct;
    for (int i = 0; i < distinct_ct; i++) distinct[i] = ps->distinct[i];
    for (int i = l; i < n; i += 32) alive[i] = 1;
    __syncwarp();
    unsigned long long calls = 1, cells = ps->cells, words = ps->words;
    unsigned used = 0;
    bool failed = false;
    while (true) {
      const int li = longest_seed(seeds, alive, n);
      if (li < 0) break;
      const TaskResult r = res[li];
      if (r.kind < 0) { failed = true; break; }
      calls++; cells += r.cells; words += r.words; used++;
      accept_overlap(a, r.kind, r.s_lo, r.s_hi, r.t_lo, r.t_hi, r.errors, r.dlen, distinct, distinct_ct);
      const bool prune = (r.kind == 3 || a.partial);
      const int32_t *delta = a.dpool + r.delta_off;
      const bool indexed = prune && r.dlen <= COMMIT_DCAP;
      if (indexed) {
        int pos0 = r.s_lo, diag0 = r.t_lo - r.s_lo;   // running values in front of the current chunk of 32 deltas
        for (int i0 = 0; i0 < r.dlen; i0 += 32) {
          const int i = i0 + l;
          const int dl = (i < r.dlen) ? delta[i] : 0;
          const uint32_t ax = (uint32_t)iabs(dl) + (dl > 0 ? 1u : 0u);            // what delta i adds to the position in total
          const uint32_t pq = (dl > 0 ? 0x10000u : 0u) | (dl < 0 ? 1u : 0u);      // counts of positive / negative deltas
          const uint32_t ix = warp_incl_scan(ax), ipq = warp_incl_scan(pq);
          const uint32_t epq = ipq - pq;                                          // exclusive
          if (i < r.dlen) {
            s_pos[wib][i] = pos0 + (int)(ix - (dl > 0 ? 1u : 0u));                // tested before the +1 of a positive delta
            s_diag[wib][i] = diag0 + (int)(epq & 0xffffu) - (int)(epq >> 16);
          }
          const uint32_t tx = __shfl_sync(FULL, ix, 31), tpq = __shfl_sync(FULL, ipq, 31);
          pos0 += (int)tx;
          diag0 += (int)(tpq & 0xffffu) - (int)(tpq >> 16);
        }
        if (l == 0) s_diag[wib][r.dlen] = diag0;
        __syncwarp();
      }
      for (int i = l; i < n; i += 32) {
        if (!alive[i]) continue;
        const Seed nd = seeds[i];
        bool remove = (i == li);
        if (!remove && prune && r.s_lo - SHIFT_SLACK <= nd.start && nd.start + nd.len <= r.s_hi + SHIFT_SLACK) {
          if (indexed) {
            int lo = 0, hi = r.dlen;                   // first delta whose test position is beyond the seed's start
            while (lo < hi) {
              const int mid = (lo + hi) >> 1;
              if (nd.start < s_pos[wib][mid]) hi = mid; else lo = mid + 1;
            }
            remove = iabs((nd.offset - nd.start) - s_diag[wib][lo]) <= SHIFT_SLACK;
          } else remove = lies_on_alignment(nd.start, nd.offset, r.s_lo, r.t_lo, delta, r.dlen);
        }
        if (remove) alive[i] = 0;
      }
      __syncwarp();
    }
    if (l == 0) {
      if (failed) a.retry_list[atomicAdd(&a.ctr->n_retry, 1u)] = pi;
      else {
        finalize_pair(a, x, distinct, distinct_ct);
        atomicAdd(&a.ctr->ext_calls, calls);
        atomicAdd(&a.ctr->ext_cells, cells);
        atomicAdd(&a.ctr->ext_words, words);
        atomicAdd(&a.ctr->n_spec_used, used);
      }
    }
  }
}

}  // namespace oic
#endif
