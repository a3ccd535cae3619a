/*
 * oic_extend_fast.cuh -- one direction of the banded prefix edit distance, warp-synchronous.
 *
 * Recurrence, exits and tie-breaks of prefixEditDistance::forward / ::reverse
 * (LIBO/prefixEditDistance-forward.C:110-285, -reverse.C:130-300) and of Set_Right_Delta / Set_Left_Delta
 * (forward.C:48-87, reverse.C:52-109):
 *   - row e holds, per diagonal d of the pruned band, the longest prefix of A matched with e errors; the lanes own
 *     consecutive diagonals, UNR chunks of 32 diagonals are in flight together (predecessor rule + first 32-base
 *     compare of every chunk issued back to back), longer slides finish in a rare-path loop;
 *   - rows e-1 and e are staged in the warp's slice of shared memory while they are at most SW wide (the common
 *     case); wider rows (spurious seeds whose band is never trimmed) are read back from the global history;
 *   - every row is appended to the warp's global history (values, then 2 direction bits per cell as two ballot
 *     words per 32 diagonals): the traceback loads 32 rows at a time (one row per lane), walks the diagonal with
 *     shuffles, and fetches the few predecessor values the deltas need in one parallel round trip;
 *   - the band-trimming predicate (progress >= Edit_Match_Limit[e]) is a ballot taken while the cells are computed;
 *     the best cell of the trimmed band is a redux.max + ballot per chunk.
 * Integer only, except the branch-point score: IEEE double without FMA contraction.
 * Returns -1 when the rows do not fit the warp's history (the pair is rerun with worst-case scratch).
 */
#ifndef OIC_EXTEND_FAST_CUH
#define OIC_EXTEND_FAST_CUH

namespace oic {

#ifndef OIC_SW
#define OIC_SW 1024
#endif
constexpr int SW = OIC_SW;            // widest row staged in shared memory (lane c of the lean loop owns chunk c: at most 32 chunks)
constexpr int FROM_NONE = -9;
constexpr int UNR = 2;              // chunks of a row in flight together

struct FastScratch {
  int32_t *row0, *row1;   // shared memory, SW + 4 ints each (2-int margins for the -2 sentinels)
};

/* Packed strings as 32-bit words (16 bases each); p is a signed base offset from the word pointer. */
__device__ __forceinline__ uint32_t ext16(const uint32_t *__restrict__ Wp, int p) {
  const int w = p >> 4;
  return __funnelshift_r(Wp[w], Wp[w + 1], (p & 15) * 2);
}
/* 16-base compare step in either direction: number of equal bases starting at offsets pa / pt and walking by DIR
 * (16 if all equal) */
/* Reads with N (noff != 0: the N-mask stream lies noff words behind the bases): 'n' in either string matches anything
 * (prefixEditDistance-forward.C:172) -- except on the query's reverse strand, where the reference holds NUL instead of 'n'
 * (AS_UTL_reverseComplement.C:39-72), which equals nothing but is still matched by an 'n' of the other string.
 * nmode: 0 = both strings hold 'n', 1 = A holds NUL, 2 = T holds NUL. */
__device__ __forceinline__ int step16(const uint32_t *__restrict__ Ap, int pa, const uint32_t *__restrict__ Tp, int pt, int DIR,
                                      long long noff = 0, int nmode = 0) {
  const int back = DIR > 0 ? 0 : 15;
  uint32_t x = ext16(Ap, pa - back) ^ ext16(Tp, pt - back);
  if (noff) {
    const uint32_t ma = ext16(Ap + noff, pa - back), mt = ext16(Tp + noff, pt - back);
    const uint32_t wild = (nmode == 1 ? 0u : ma) | (nmode == 2 ? 0u : mt);
    const uint32_t forced = (nmode == 1 ? ma : 0u) | (nmode == 2 ? mt : 0u);
    x = (x | forced) & ~wild;
  }
  if (x == 0) return 16;
  return DIR > 0 ? ((__ffs((int)x) - 1) >> 1) : (__clz((int)x) >> 1);
}
__device__ __forceinline__ int match_dir(const uint32_t *__restrict__ Ap, int pa, const uint32_t *__restrict__ Tp, int pt, int limit, int DIR,
                                         long long noff = 0, int nmode = 0) {
  int r = 0;
  while (r < limit) {
    const int adv = step16(Ap, pa + DIR * r, Tp, pt + DIR * r, DIR, noff, nmode);
    r += adv;
    if (adv < 16) break;
  }
  return r < limit ? r : limit;
}

// event i of the traceback (an indel step): encoded (pred + 4) * 4 + (from + 1) in ws.stack[i]
__device__ __forceinline__ int tb_entry(const int32_t *ev, int i, int start_val) {
  const int e = ev[i];
  const int pred = (e >> 2) - 4, from = (e & 3) - 1;
  const int lastp = i ? ((ev[i - 1] >> 2) - 4) : start_val;
  const int mx = pred + (from == -1 ? 0 : 1);
  return (from == -1) ? mx - lastp - 1 : lastp - (mx - 1);
}

/* Cooperative traceback from cell (fin_e, fin_d).  from0 != FROM_NONE overrides the stored direction of the
 * first row (the exit row's cell may have been rewritten by the "last error is a mismatch" tweak).
 * Leaves the events in ws.stack[0 .. nev) and returns nev. */
__device__ __forceinline__ int traceback_fast(const WarpScratch &ws, int fin_e, int fin_d, int from0) {
  // Row k of the history: meta[k] = {lo, trimmed lo, trimmed hi, offset}; values at vals[offset ..], the direction
  // words right behind them (the row's computed width follows from row k-1: hi = min(trimmed hi of k-1 + 1, k)).
  const int l = lane_id();
  int nev = 0;
  int dcur = fin_d;
  for (int kb = fin_e; kb >= 1; kb -= 32) {
    const int k = kb - l;
    const bool valid = k >= 1;
    int lo = 0, cs = 0;
    uint2 w0 = make_uint2(0, 0), w1 = w0, w2 = w0;
    int4 mp = make_int4(0, 0, 0, 0);
    if (valid) {
      const int4 mk = ws.meta[k];
      mp = ws.meta[k - 1];
      lo = mk.x;
      const int width = imin(mp.z + 1, k) - lo + 1, nch = (width + 31) >> 5;
      const int imin_ = dcur - 31 - lo;
      cs = (imin_ > 0 ? imin_ : 0) >> 5;
      const uint32_t *dp = reinterpret_cast<const uint32_t *>(ws.vals + mk.w + width + 2);
      if (cs < nch) w0 = make_uint2(dp[2 * cs], dp[2 * cs + 1]);
      if (cs + 1 < nch) w1 = make_uint2(dp[2 * cs + 2], dp[2 * cs + 3]);
      if (cs + 2 < nch) w2 = make_uint2(dp[2 * cs + 4], dp[2 * cs + 5]);
    }
    int my_from = 0, my_d = 0;
    const int steps = kb < 32 ? kb : 32;
#pragma unroll 1
    for (int j = 0; j < steps; j++) {
      int f = 0;
      if (l == j) {
        const int i = dcur - lo;
        const int c = (i >> 5) - cs, bit = i & 31;
        const uint2 w = (c == 0) ? w0 : ((c == 1) ? w1 : w2);
        f = ((w.x >> bit) & 1) ? -1 : (((w.y >> bit) & 1) ? 1 : 0);
        if (from0 != FROM_NONE && k == fin_e) f = from0;
        my_from = f;
        my_d = dcur;
      }
      f = __shfl_sync(FULL, f, j);
      dcur += f;
    }
    const bool ev = valid && my_from != 0;
    const unsigned em = __ballot_sync(FULL, ev);
    if (ev) {
      const int pv = ws.vals[mp.w + (my_d + my_from) - mp.x];
      ws.stack[nev + __popc(em & ((1u << l) - 1))] = (pv + 4) * 4 + (my_from + 1);
    }
    nev += __popc(em);
  }
  __syncwarp();
  return nev;
}

__device__ __forceinline__ int traceback_vals(const WarpScratch &ws, int fin_e, int fin_d, int from0);   // oic_extend_lean.cuh

/* After the row loop of one direction: pick the exit cell (or fall back to the best cell seen), trace back, build the
 * delta list (Set_Right_Delta / Set_Left_Delta).  Shared by the generic and the lean row loops. */
__device__ __forceinline__ int dp_finish(const int DIR, WarpScratch &ws, int n, int row0, int fin_e, int fin_d, int fin_val, int from0, int errors,
                                         int max_score_len, int max_score_best_d, int max_score_best_e, int &a_end, int &t_end,
                                         int &leftover, int &to_end, int &delta_len, bool lean_history = false) {
  // lean_history: the history was written by the lean loop (values only, no direction words) -> traceback_vals
  const int l = lane_id();

  if (fin_e < 0) {   // ran out of errors / band, or branch-point abort: back to the best cell seen
    a_end = DIR * max_score_len;
    t_end = DIR * (max_score_len + max_score_best_d);
    fin_e = max_score_best_e; fin_d = max_score_best_d;
    fin_val = fin_e ? max_score_len : row0;
    from0 = FROM_NONE;
    errors = max_score_best_e;
    to_end = 0;
  }
  __syncwarp();
  const long long tb0 = OIC_CLOCK();
  const int ns = lean_history ? traceback_vals(ws, fin_e, fin_d, from0) : traceback_fast(ws, fin_e, fin_d, from0);
  const int last = ns ? ((ws.stack[ns - 1] >> 2) - 4) : fin_val;
  if (DIR > 0) {
    // Set_Right_Delta: st = entries, then last + 1; delta[x] = |st[ns - x]| * sign(st[ns - x - 1])
    for (int x = l; x < ns; x += 32) {
      const int i = ns - x;
      const int sti = (i == ns) ? last + 1 : tb_entry(ws.stack, i, fin_val);
      const int stm = tb_entry(ws.stack, i - 1, fin_val);
      ws.rdelta[x] = iabs(sti) * ((stm > 0) - (stm < 0));
    }
    delta_len = ns;
  } else {
    // Set_Left_Delta; a leading +1 indel becomes a substitution by pulling t_end in by one (reverse.C:92-108)
    leftover = last;
    bool fix = false;
    int e1 = 0;
    if (ns > 1 && tb_entry(ws.stack, 0, fin_val) == 1 && t_end + n > 0) { fix = true; e1 = tb_entry(ws.stack, 1, fin_val); }
    if (!fix) {
      for (int x = l; x < ns; x += 32) ws.ldelta[x] = tb_entry(ws.stack, x, fin_val);
      delta_len = ns;
    } else {
      for (int x = l; x < ns - 1; x += 32) ws.ldelta[x] = (x == 0) ? ((e1 > 0) ? e1 + 1 : e1 - 1) : tb_entry(ws.stack, x + 1, fin_val);
      delta_len = ns - 1;
      t_end--;
    }
  }
  __syncwarp();
  ws.t_tb += OIC_CLOCK() - tb0;
  return errors;
}

/* History layout of row e (ints, in the warp's slice `vals`): [2 margin][width values][2 margin][2 * nch direction words];
 * meta[e] = {lo, trimmed lo, trimmed hi, offset of the first value}.  The live rows have the same 2-int margins: after
 * trimming, the four cells just outside the band are overwritten with the reference's -2 sentinel (forward.C:155-161),
 * so the predecessor rule needs no bounds checks. */
__device__ __forceinline__ int dp_run_fast(const int DIR, const ExtArgs &a, WarpScratch &ws, const FastScratch &fs, const uint64_t *__restrict__ A, int64_t a0,
                           int m, const uint64_t *__restrict__ T, int64_t t0, int n, int error_limit, int &a_end, int &t_end,
                           int &leftover, int &to_end, int &delta_len) {
  const int l = lane_id();
  const uint32_t *Ap = reinterpret_cast<const uint32_t *>(A) + (a0 >> 4);
  const uint32_t *Tp = reinterpret_cast<const uint32_t *>(T) + (t0 >> 4);
  const int aph = (int)(a0 & 15), tph = (int)(t0 & 15);
  const int row0 = match_dir(Ap, aph, Tp, tph, m, DIR, ws.n_off32, ws.n_mode);
  int32_t *sm_prev = fs.row0 + 2, *sm_cur = fs.row1 + 2;   // shared-memory staging (rows up to SW wide)
  const int32_t *prev = sm_prev;                           // row e-1: shared memory, or the global history when wide
  if (l < 5) sm_prev[l - 2] = (l == 2) ? row0 : -2;
  if (l == 0) {
    ws.cells += 1;
    ws.words += (unsigned)(row0 + 1 + 15) / 16;
    ws.meta[0] = make_int4(0, 0, 0, 2);
    ws.vals[2] = row0;
  }
  delta_len = 0;
  leftover = 0;
  if (row0 == m) {
    a_end = t_end = DIR * m;
    leftover = m;
    to_end = 1;
    return 0;
  }
  int left = 0, right = 0;
  double max_score = 0.0;
  int max_score_len = 0, max_score_best_d = 0, max_score_best_e = 0;
  int best_d = 0, best_e = 0, longest = 0;
  int plo = 0;
  int next_off = 5;
  int fin_e = -1, fin_d = 0, fin_val = 0, from0 = FROM_NONE;
  int errors = 0;
  unsigned wsum = 0;    // per-lane count of 16-base compare words (W of the roofline formula)
  unsigned ncell = 0;   // warp-uniform count of cells (C)
  to_end = 0;
  __syncwarp();

  for (int e = 1; e <= error_limit; e++) {
    left = imax(left - 1, -e);
    right = imin(right + 1, e);
    const int width = right - left + 1;
    const int nch = (width + 31) >> 5;
    if (e >= ws.max_rows || (unsigned long long)next_off + width + 4 + 2 * nch > ws.rowcap) { errors = -1; break; }
    const int off = next_off + 2;
    next_off += width + 4 + 2 * nch;
    const bool wide = width > SW;                        // too wide for the staging buffer: use the history itself
    int32_t *cur = wide ? ws.vals + off : sm_cur;
    uint32_t *dirs = reinterpret_cast<uint32_t *>(ws.vals + off + width + 2);
    const int lim = ws.match_limit[e];
    const int pshift = left - plo;                       // index of diagonal `left` in the previous row

    bool ended = false;
    int d_end = 0, r_end = 0;
    int nl = right + 1, nr = left - 1;
    int gmax = -1, gpos = 0;                             // highest cell of the row, lowest diagonal on ties
    for (int c0 = 0; c0 < nch && !ended; c0 += UNR) {
      int r[UNR], r0[UNR], fr[UNR], cap[UNR];
#pragma unroll
      for (int u = 0; u < UNR; u++) {                    // predecessor rule (forward.C:163-170), out-of-band = -2 sentinel
        const int i = 32 * (c0 + u) + l, d = left + i;
        r0[u] = -1; fr[u] = 0; cap[u] = -1;
        if (i < width) {
          const int ip = i + pshift;
          const int g0 = prev[ip], gm = prev[ip - 1], gp = prev[ip + 1];
          int x = 1 + g0, f = 0;
          if (gm > x) { x = gm; f = -1; }
          if (1 + gp > x) { x = 1 + gp; f = 1; }
          r0[u] = x; fr[u] = f;
          cap[u] = imin(m, n - d);
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {                    // first 16 bases of every slide, loads back to back
        const int d = left + 32 * (c0 + u) + l;
        r[u] = r0[u];
        if (r0[u] < cap[u]) {                            // (inactive lanes have r0 = cap = -1)
          r[u] = r0[u] + step16(Ap, aph + DIR * r0[u], Tp, tph + DIR * (r0[u] + d), DIR, ws.n_off32, ws.n_mode);
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const int c = c0 + u;
        if (c < nch && !ended) {                         // warp-uniform
          const int i = 32 * c + l, d = left + i;
          const bool act = i < width;
          if (r[u] - r0[u] == 16 && r[u] < cap[u])       // rare: the slide goes on past 16 bases
            r[u] += match_dir(Ap, aph + DIR * r[u], Tp, tph + DIR * (r[u] + d), cap[u] - r[u], DIR, ws.n_off32, ws.n_mode);
          if (r[u] > cap[u]) r[u] = cap[u];
          if (act) {
            cur[i] = r[u];
            if (!wide) ws.vals[off + i] = r[u];
          }
          const unsigned bm = __ballot_sync(FULL, fr[u] == -1), bp = __ballot_sync(FULL, fr[u] == 1);
          if (l == 0) { dirs[2 * c] = bm; dirs[2 * c + 1] = bp; }
          const unsigned em = __ballot_sync(FULL, act && (r[u] == m || r[u] + d == n));
          const unsigned wc = act ? (unsigned)(r[u] - r0[u] + 16) >> 4 : 0u;
          if (em) {
            const int first = __ffs(em) - 1;
            if (l <= first) wsum += wc;
            ncell += 32 * c + first + 1;
            d_end = left + 32 * c + first;
            r_end = __shfl_sync(FULL, r[u], first);
            ended = true;
          } else {
            wsum += wc;
            const unsigned okm = __ballot_sync(FULL, act && (r[u] + imax(d, 0) >= lim));
            if (okm) {
              if (nl > right) nl = left + 32 * c + __ffs(okm) - 1;
              nr = left + 32 * c + (31 - __clz(okm));
            }
            const int mx = __reduce_max_sync(FULL, act ? r[u] : -1);
            if (mx > gmax) {
              gmax = mx;
              gpos = left + 32 * c + __ffs(__ballot_sync(FULL, act && r[u] == mx)) - 1;
            }
          }
        }
      }
    }
    __syncwarp();

    if (ended) {
      const double score = OIC_DSUB(OIC_DMUL((double)r_end, a.branch_match_value), (double)e);
      const int tail_len = r_end - max_score_len;
      bool abort = false;
      if (a.partial && score < max_score) abort = true;
      if (e > MIN_BRANCH_END_DIST / 2 && tail_len >= MIN_BRANCH_END_DIST) {
        const double slope = OIC_DDIV(OIC_DSUB(max_score, score), (double)tail_len);
        if (slope >= a.min_branch_tail_slope) abort = true;
      }
      if (l == 0) ws.meta[e] = make_int4(left, left, right, off);
      if (abort) break;
      if (DIR > 0) {   // force the last error to be a mismatch rather than an insertion (forward.C:213-219)
        const int gp = prev[d_end + 1 - plo];
        if (r_end == m && 1 + gp == r_end && d_end < right) {
          d_end++;
          if (l == 0) ws.vals[off + d_end - left] = r_end;
        }
      }
      a_end = DIR * r_end;
      t_end = DIR * (r_end + d_end);
      fin_e = e; fin_d = d_end; fin_val = r_end;
      {   // direction of the exit cell, by the rule (the cell may be one the row never computed)
        const int ip = d_end - plo;
        const int g0 = prev[ip], gm = prev[ip - 1], gp = prev[ip + 1];
        int x = 1 + g0;
        from0 = 0;
        if (gm > x) { x = gm; from0 = -1; }
        if (1 + gp > x) from0 = 1;
      }
      errors = e;
      to_end = 1;
      break;
    }
    ncell += width;

    // band trimming (forward.C:235-254): [nl, nr] = first / last cell whose progress reaches Edit_Match_Limit[e]
    if (nl > right) {
      if (l == 0) ws.meta[e] = make_int4(left, nl, right, off);
      break;
    }
    if (l == 0) ws.meta[e] = make_int4(left, nl, nr, off);

    // best cell of the trimmed band: highest row, lowest diagonal on ties; only if it beats `longest`
    if (gmax > longest) {
      if (gpos < nl || gpos > nr) {   // rare: the row's top cell was trimmed away -> exact search inside the band
        gmax = -1;
        for (int c = (nl - left) >> 5; c <= ((nr - left) >> 5); c++) {
          const int i = 32 * c + l, d = left + i;
          const int v = (d >= nl && d <= nr) ? cur[i] : -1;
          const int mx = __reduce_max_sync(FULL, v);
          if (mx > gmax) {
            gmax = mx;
            gpos = left + 32 * c + __ffs(__ballot_sync(FULL, v == mx)) - 1;
          }
        }
      }
      if (gmax > longest) { longest = gmax; best_d = gpos; best_e = e; }
    }
    const double score = OIC_DSUB(OIC_DMUL((double)longest, a.branch_match_value), (double)e);
    if (score > max_score) {
      max_score = score;
      max_score_len = longest;
      max_score_best_d = best_d;
      max_score_best_e = best_e;
    }
    // -2 sentinels on the four cells just outside the trimmed band
    __syncwarp();
    if (l < 4) cur[(l < 2 ? nl - 1 - l : nr - 1 + l) - left] = -2;
    plo = left;
    left = nl; right = nr;
    prev = cur;
    if (!wide) { int32_t *t = sm_prev; sm_prev = sm_cur; sm_cur = t; }
    __syncwarp();
  }
  ws.cells += (l == 0) ? ncell : 0;
  ws.words += wsum;
  if (errors < 0) return -1;
  return dp_finish(DIR, ws, n, row0, fin_e, fin_d, fin_val, from0, errors, max_score_len, max_score_best_d, max_score_best_e, a_end, t_end,
                   leftover, to_end, delta_len);
}

}  // namespace oic
#endif
