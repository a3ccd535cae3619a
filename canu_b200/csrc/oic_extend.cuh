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
  uint32_t *dbg_pair;   // optional: per pair {kcycles, extension calls, lane-0 cells}
  DevCounters *ctr;
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
};

}  // namespace oic
#include "oic_extend_fast.cuh"
namespace oic {

// shared-memory path first; the generic global-memory path only for runs with a row wider than SW
/* Extend_Alignment.  kind: 0 NONE, 1 LEFT_BRANCH_PT, 2 RIGHT_BRANCH_PT, 3 DOVETAIL; -1 = scratch overflow.
 * The spliced delta list is left in ws.ldelta[0 .. delta_len). */
__device__ int extend_alignment(const ExtArgs &a, WarpScratch &ws, const FastScratch &fs, const uint64_t *S, int64_t sp, int s_len,
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
    errs = dp_run_fast(side == 0 ? 1 : -1, a, ws, fs, A, a0, m, B, b0, n, limit, a_end, b_end, lo, to_end, dl);
    if (errs < 0) return -1;
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

__global__ void __launch_bounds__(EXT_WARPS * 32, 4) k_process_pairs(ExtArgs a) {
  const int l = lane_id();
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
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
  long long t_ext = 0, t_prune = 0, t_pair_max = 0;
  const long long t_begin = OIC_CLOCK();
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
    const Pair pr = a.pairs[pi];
    const PairInfo inf = a.info[pi];
    const uint32_t ri = a.qs_read[pr.qs >> 1];
    const int dir = pr.qs & 1;
    const uint64_t *S = dir ? a.rc : a.fwd;
    const int64_t sp = a.rd_pos[ri];
    const int s_len = (int)a.rd_len[ri];
    const uint32_t s_id = a.first_id + ri;
    const uint32_t hri = a.hash_ri_bgn + pr.str;
    const uint64_t *T = a.fwd;
    const int64_t tp = a.rd_pos[hri];
    const int t_len = (int)a.rd_len[hri];
    const uint32_t t_id = a.hash_bgn_id + pr.str;
    Seed *seeds = a.seeds + pr.first;
    int n = (int)pr.n;
    const long long t_pair0 = OIC_CLOCK();

    // --minkmers filter: computeMinimumKmers (Process_String_Overlaps.C:75-86,725)
    if (a.filter_by_kmer_count != 0) {
      double ovl = (double)(inf.diag_end - inf.diag_bgn);
      if (ovl < 0) ovl = -ovl;
      unsigned long long expected = 0;
      if (!(ovl < (double)a.K)) expected = (unsigned long long)(int)floor(OIC_DMUL(a.minkmer_factor, ovl - a.K + 1));
      unsigned long long need = a.filter_by_kmer_count > expected ? a.filter_by_kmer_count : expected;
      if (need > (unsigned long long)inf.diag_ct) {
        if (l == 0) atomicAdd(&a.ctr->kmer_hits_skipped, 1ull);
        continue;
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
      const int s_tail = s_len - s_head - m.len + 1, t_tail = t_len - t_head - m.len + 1;
      if (s_tail <= t_tail) { if (s_tail > HOPELESS_MATCH && !(qsc & 2)) hopeless = true; }
      else { if (t_tail > HOPELESS_MATCH && !(hsc & 2)) hopeless = true; }
      if (hopeless) {
        if (l == 0) atomicAdd(&a.ctr->kmer_hits_without_olap, 1ull);
        continue;
      }
    }

    OlapInfo distinct[MAX_DISTINCT_OLAPS];
    int distinct_ct = 0;
    bool overflow = false;
    const unsigned long long snap_cells = ws.cells, snap_words = ws.words, snap_calls = ext_calls;

    while (n > 0) {
      // first of the longest seeds, in list order (Process_String_Overlaps.C:473-480)
      long long bk = -1;
      for (int i = l; i < n; i += 32) {
        long long k = (long long)seeds[i].len * 4294967296LL + (long long)(0x7fffffff - i);
        if (k > bk) bk = k;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        long long ok = __shfl_xor_sync(FULL, bk, o);
        if (ok > bk) bk = ok;
      }
      const int li = 0x7fffffff - (int)(bk & 0xffffffffLL);
      const Seed lg = seeds[li];
      int s_lo, s_hi, t_lo, t_hi, errors, dlen;
      ext_calls++;
      const long long te0 = OIC_CLOCK();
      const int kind = extend_alignment(a, ws, fs, S, sp, s_len, T, tp, t_len, lg.start, lg.offset, lg.len, s_lo, s_hi, t_lo, t_hi,
                                        errors, dlen);
      t_ext += OIC_CLOCK() - te0;
      if (kind < 0) { overflow = true; break; }

      if (kind == 3 || a.partial) {
        if (1 + s_hi - s_lo >= a.min_olap_len && 1 + t_hi - t_lo >= a.min_olap_len) {
          const int olap_len = 1 + imin(s_hi - s_lo, t_hi - t_lo);
          const double quality = OIC_DDIV((double)errors, (double)olap_len);
          if (errors <= (int)ceil(OIC_DMUL((double)olap_len, a.erate)))
            add_overlap(a.partial, s_lo, s_hi, t_lo, t_hi, quality, dlen, distinct, distinct_ct);
        }
      }
      if (inf.consistent) break;

      // drop the extended seed and every seed that lies on this alignment (Process_String_Overlaps.C:514-531)
      const long long tp0 = OIC_CLOCK();
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
      t_prune += OIC_CLOCK() - tp0;
    }
    { long long tp = OIC_CLOCK() - t_pair0; if (tp > t_pair_max) t_pair_max = tp;
      if (l == 0 && a.dbg_pair) { a.dbg_pair[3ull * pi] = (uint32_t)(tp >> 10); a.dbg_pair[3ull * pi + 1] = (uint32_t)(ext_calls - snap_calls); a.dbg_pair[3ull * pi + 2] = (uint32_t)((ws.cells - snap_cells) ); } }

    if (overflow) {   // the retry pass redoes this pair from scratch: forget its partial work counters
      ws.cells = snap_cells; ws.words = snap_words; ext_calls = snap_calls;
      if (l == 0) a.retry_list[atomicAdd(&a.ctr->n_retry, 1u)] = pi;
      continue;
    }

    if (l == 0) {
      int overlaps_output = 0;
      if (distinct_ct > 0) {
        int deleted[MAX_DISTINCT_OLAPS] = {0, 0, 0};
        if (a.partial) { if (a.unique_olap) choose_best_partial(distinct, distinct_ct, deleted); }
        else { if (a.unique_olap) combine_into_one(distinct, distinct_ct, deleted); else merge_intersecting(distinct, distinct_ct, deleted); }
        int nout = 0;
        for (int i = 0; i < distinct_ct; i++) nout += !deleted[i];
        uint32_t ob = atomicAdd(&a.ctr->n_out, (unsigned)nout);
        for (int i = 0; i < distinct_ct; i++) {
          if (deleted[i]) continue;
          if (ob >= a.out_cap) { atomicOr(&a.ctr->out_overflow, 1u); break; }
          uint32_t *w = a.out + 6ull * ob;
          if (a.partial) output_partial(w, s_id, t_id, dir, distinct[i], s_len, t_len);
          else {
            int contained = output_overlap(w, s_id, s_len, dir, t_id, t_len, distinct[i]);
            if (contained) atomicAdd(&a.ctr->contained, 1ull); else atomicAdd(&a.ctr->dovetail, 1ull);
          }
          ob++;
          overlaps_output++;
        }
        atomicAdd(&a.ctr->total_overlaps, (unsigned long long)overlaps_output);
      }
      if (overlaps_output == 0) atomicAdd(&a.ctr->kmer_hits_without_olap, 1ull);
      else {
        atomicAdd(&a.ctr->kmer_hits_with_olap, 1ull);
        if (overlaps_output > 1) atomicAdd(&a.ctr->multi_overlap, 1ull);
      }
    }
  }

  unsigned long long c = warp_sum64(ws.cells), w = warp_sum64(ws.words);
  if (l == 0) {
    const long long tw = OIC_CLOCK() - t_begin;
    atomicAdd(&a.ctr->t_dp, (unsigned long long)(t_ext - ws.t_tb));
    atomicAdd(&a.ctr->t_tb, (unsigned long long)ws.t_tb);
    atomicAdd(&a.ctr->t_prune, (unsigned long long)t_prune);
    atomicAdd(&a.ctr->t_warp_sum, (unsigned long long)tw);
    atomicMax(&a.ctr->t_warp_max, (unsigned long long)tw);
    atomicMax(&a.ctr->t_pair_max, (unsigned long long)t_pair_max);
    if (c) atomicAdd(&a.ctr->ext_cells, c);
    if (w) atomicAdd(&a.ctr->ext_words, w);
    if (ext_calls) atomicAdd(&a.ctr->ext_calls, ext_calls);
  }
}

}  // namespace oic
#endif
