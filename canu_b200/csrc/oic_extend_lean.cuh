/*
 * oic_extend_lean.cuh -- the row loop of the banded prefix edit distance for the common case: every row at most SW
 * diagonals wide, both live rows in shared memory.  Same recurrence, exits, tie-breaks, counters and history layout as
 * dp_run_fast (oic_extend_fast.cuh; LIBO/prefixEditDistance-forward.C:110-285, -reverse.C:130-300) -- the results are
 * bit-identical -- but written for instruction count: no generic pointers, two chunks of 32 diagonals per iteration with
 * branch-free bookkeeping (selects on ballot words instead of warp-uniform branches), one uniform branch per chunk pair for
 * the rare "a cell reached the end of a string" exit, direction words gathered in registers and stored once per row.
 * Returns -1 when the rows do not fit the warp's history, -2 when a row is wider than SW (the caller reruns the
 * extension with dp_run_fast).
 */
#ifndef OIC_EXTEND_LEAN_CUH
#define OIC_EXTEND_LEAN_CUH

namespace oic {

constexpr int LSTRIDE = SW + 72;   // ints per live row: 2 margin + SW + idle-lane slack of the last chunk pair
constexpr int LEAN_WIDE = -2;

__device__ __forceinline__ int dp_run_lean(const int DIR, const ExtArgs &a, WarpScratch &ws, int32_t *srow0, int32_t *srow1,
                                           const uint64_t *__restrict__ A, int64_t a0, int m, const uint64_t *__restrict__ T, int64_t t0,
                                           int n, int error_limit, int &a_end, int &t_end, int &leftover, int &to_end, int &delta_len) {
  const int l = lane_id();
  const uint32_t *Ap = reinterpret_cast<const uint32_t *>(A) + (a0 >> 4);
  const uint32_t *Tp = reinterpret_cast<const uint32_t *>(T) + (t0 >> 4);
  const int aph = (int)(a0 & 15), tph = (int)(t0 & 15);
  const int row0 = match_dir(Ap, aph, Tp, tph, m, DIR);
  int32_t *prev = srow0 + 2, *cur = srow1 + 2;
  if (l < 5) prev[l - 2] = (l == 2) ? row0 : -2;
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
  unsigned wsum = 0;    // per-lane count of 16-base compare words (W)
  unsigned ncell = 0;   // warp-uniform count of cells (C)
  to_end = 0;
  int lim_next = ws.match_limit[1];
  __syncwarp();

  for (int e = 1; e <= error_limit; e++) {
    left = imax(left - 1, -e);
    right = imin(right + 1, e);
    const int width = right - left + 1;
    const int nch = (width + 31) >> 5;
    if (width > SW) { errors = LEAN_WIDE; break; }
    if (e >= ws.max_rows || (unsigned long long)next_off + width + 4 + 2 * nch > ws.rowcap) { errors = -1; break; }
    const int off = next_off + 2;
    next_off += width + 4 + 2 * nch;
    int32_t *hist = ws.vals + off;
    const int lim = lim_next;
    if (e + 1 < ws.max_rows) lim_next = ws.match_limit[e + 1];
    const int pshift = left - plo;                       // index of diagonal `left` in the previous row

    bool ended = false;
    int d_end = 0, r_end = 0;
    int nl = 0x7fffffff, nr = (int)0x80000000;
    int gmax = -1, gpos = 0;                             // highest cell of the row, lowest diagonal on ties
    unsigned my_bm = 0, my_bp = 0;                       // lane c keeps the direction words of chunk c

#pragma unroll 1
    for (int c0 = 0; c0 < nch; c0 += 2) {
      int r[2], r0[2], cap[2], dd[2];
      bool fm[2], fp[2], cont[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {                      // predecessor rule (forward.C:163-170), out-of-band = -2 sentinel
        const int i = 32 * (c0 + u) + l, d = left + i;
        const int ip = i + pshift;
        const int g0 = prev[ip] + 1, gm = prev[ip - 1], gp = prev[ip + 1] + 1;
        const int x = imax(g0, gm);
        fp[u] = gp > x;
        fm[u] = gm > g0;
        r0[u] = imax(x, gp);
        dd[u] = d;
        cap[u] = (i < width) ? imin(m, n - d) : (int)0x80000000;
      }
#pragma unroll
      for (int u = 0; u < 2; u++) {                      // first 16 bases of every slide, loads back to back
        r[u] = r0[u];
        cont[u] = false;
        if (r0[u] < cap[u]) {
          const int adv = step16(Ap, aph + DIR * r0[u], Tp, tph + DIR * (r0[u] + dd[u]), DIR);
          r[u] = r0[u] + adv;
          cont[u] = (adv == 16) && (r[u] < cap[u]);
        }
      }
      while (__any_sync(FULL, cont[0] || cont[1])) {     // slides longer than 16 bases
#pragma unroll
        for (int u = 0; u < 2; u++) {
          if (cont[u]) {
            const int adv = step16(Ap, aph + DIR * r[u], Tp, tph + DIR * (r[u] + dd[u]), DIR);
            r[u] += adv;
            cont[u] = (adv == 16) && (r[u] < cap[u]);
          }
        }
      }
      unsigned em[2], okm[2], pm[2], wc[2];
      int mx[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int i = 32 * (c0 + u) + l;
        const bool act = i < width;
        r[u] = imin(r[u], cap[u]);
        if (act) { cur[i] = r[u]; hist[i] = r[u]; }
        const unsigned bp = __ballot_sync(FULL, act && fp[u]);
        const unsigned bm = __ballot_sync(FULL, act && fm[u]) & ~bp;
        if (l == c0 + u) { my_bm = bm; my_bp = bp; }
        em[u] = __ballot_sync(FULL, act && (r[u] == m || r[u] + dd[u] == n));
        wc[u] = act ? (unsigned)(r[u] - r0[u] + 16) >> 4 : 0u;
        okm[u] = __ballot_sync(FULL, act && (r[u] + imax(dd[u], 0) >= lim));
        mx[u] = __reduce_max_sync(FULL, act ? r[u] : -1);
        pm[u] = __ballot_sync(FULL, act && r[u] == mx[u]);
      }
      if (em[0] | em[1]) {                               // a cell reached the end of A or T: the row stops at the first one
        const int u = em[0] ? 0 : 1;
        const int first = __ffs(em[u]) - 1;
        if (u == 1) wsum += wc[0];
        if (l <= first) wsum += wc[u];
        ncell += 32 * (c0 + u) + first + 1;
        d_end = left + 32 * (c0 + u) + first;
        r_end = __shfl_sync(FULL, u ? r[1] : r[0], first);
        ended = true;
        break;
      }
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int base = left + 32 * (c0 + u);
        wsum += wc[u];
        nl = imin(nl, okm[u] ? base + __ffs(okm[u]) - 1 : 0x7fffffff);
        nr = imax(nr, okm[u] ? base + 31 - __clz(okm[u]) : (int)0x80000000);
        const bool better = mx[u] > gmax;
        gpos = better ? base + __ffs(pm[u]) - 1 : gpos;
        gmax = better ? mx[u] : gmax;
      }
    }
    {   // direction words of the row: two per chunk, right behind the values (and their 2-int margin)
      uint32_t *dirs = reinterpret_cast<uint32_t *>(hist + width + 2);
      if (l < nch) { dirs[2 * l] = my_bm; dirs[2 * l + 1] = my_bp; }
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
          if (l == 0) hist[d_end - left] = r_end;
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
      if (l == 0) ws.meta[e] = make_int4(left, right + 1, right, off);
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
          const int mxv = __reduce_max_sync(FULL, v);
          if (mxv > gmax) {
            gmax = mxv;
            gpos = left + 32 * c + __ffs(__ballot_sync(FULL, v == mxv)) - 1;
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
    if (l < 4) cur[(l < 2 ? nl - 1 - l : nr - 1 + l) - left] = -2;
    plo = left;
    left = nl; right = nr;
    { int32_t *t = prev; prev = cur; cur = t; }
    __syncwarp();
  }
  ws.cells += (l == 0) ? ncell : 0;
  ws.words += wsum;
  if (errors < 0) return errors;
  return dp_finish(DIR, ws, n, row0, fin_e, fin_d, fin_val, from0, errors, max_score_len, max_score_best_d, max_score_best_e, a_end, t_end,
                   leftover, to_end, delta_len);
}

}  // namespace oic
#endif
