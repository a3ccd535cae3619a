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

constexpr int SW = 512;             // widest row staged in shared memory
constexpr int FROM_NONE = -9;
constexpr int UNR = 2;              // chunks of a row in flight together

struct FastScratch {
  int32_t *row0, *row1;   // shared memory, SW + 2 ints each
};

// predecessor rule of the recurrence at diagonal d of the row after `prev` (values in shared memory)
__device__ __forceinline__ int rule_from(const int32_t *prev, int plo, int ptlo, int pthi, int d, int &r0) {
  const int ip = d - plo;
  const int g0 = (d < ptlo || d > pthi) ? -2 : prev[ip];
  const int gm = (d - 1 < ptlo || d - 1 > pthi) ? -2 : prev[ip - 1];
  const int gp = (d + 1 < ptlo || d + 1 > pthi) ? -2 : prev[ip + 1];
  int r = 1 + g0, from = 0;
  if (gm > r) { r = gm; from = -1; }
  if (1 + gp > r) { r = 1 + gp; from = 1; }
  r0 = r;
  return from;
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
      const uint32_t *dp = reinterpret_cast<const uint32_t *>(ws.vals + mk.w + width);
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

/* 32-base compare step in either direction: number of equal bases starting at A[pa], T[pt] and walking by DIR
 * (32 if all equal).  The reverse direction bit-reverses the XOR so one find-first-set serves both: the two bits of
 * a base swap places, which does not matter for "first base that differs". */
__device__ __forceinline__ int step32(const uint64_t *__restrict__ A, int64_t pa, const uint64_t *__restrict__ T, int64_t pt, int DIR) {
  const int back = DIR > 0 ? 0 : 31;
  uint64_t x = ext32(A, pa - back) ^ ext32(T, pt - back);
  if (DIR < 0) x = __brevll(x);
  return x ? ((__ffsll((long long)x) - 1) >> 1) : 32;
}
__device__ __forceinline__ int match_dir(const uint64_t *__restrict__ A, int64_t pa, const uint64_t *__restrict__ T, int64_t pt, int limit, int DIR) {
  int r = 0;
  while (r < limit) {
    const int adv = step32(A, pa + DIR * r, T, pt + DIR * r, DIR);
    r += adv;
    if (adv < 32) break;
  }
  return r < limit ? r : limit;
}

__device__ __forceinline__ int dp_run_fast(const int DIR, const ExtArgs &a, WarpScratch &ws, const FastScratch &fs, const uint64_t *__restrict__ A, int64_t a0,
                           int m, const uint64_t *__restrict__ T, int64_t t0, int n, int error_limit, int &a_end, int &t_end,
                           int &leftover, int &to_end, int &delta_len) {
  const int l = lane_id();
  const int row0 = match_dir(A, a0, T, t0, m, DIR);
  int32_t *sm_prev = fs.row0, *sm_cur = fs.row1;   // shared-memory staging (rows up to SW wide)
  const int32_t *prev = sm_prev;                   // where row e-1 is read from: shared memory or the global history
  if (l == 0) {
    ws.cells += 1;
    ws.words += (unsigned)(row0 + 1 + 15) / 16;
    ws.meta[0] = make_int4(0, 0, 0, 0);
    ws.vals[0] = row0;
    sm_prev[0] = row0;
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
  int plo = 0, ptlo = 0, pthi = 0;
  int next_off = 1;
  int fin_e = -1, fin_d = 0, fin_val = 0, from0 = FROM_NONE;
  int errors = 0;
  to_end = 0;
  __syncwarp();

  for (int e = 1; e <= error_limit; e++) {
    left = imax(left - 1, -e);
    right = imin(right + 1, e);
    const int width = right - left + 1;
    const int nch = (width + 31) >> 5;
    if (e >= ws.max_rows || (unsigned long long)next_off + width + 2 * nch > ws.rowcap) return -1;
    const int off = next_off;
    next_off += width + 2 * nch;
    const bool wide = width > SW;                              // too wide for the staging buffer: use the history itself
    int32_t *cur = wide ? ws.vals + off : sm_cur;
    uint32_t *dirs = reinterpret_cast<uint32_t *>(ws.vals + off + width);

    // Cells of the row, UNR chunks (32 diagonals each) per pass: the predecessor rule and the first 32-base compare of
    // every chunk are issued back to back (independent loads in flight), longer slides finish in the rare-path loop.
    // The trimming predicate (progress >= Edit_Match_Limit[e]) is evaluated on the fly.
    const int lim = ws.match_limit[e];
    bool ended = false;
    int d_end = 0, r_end = 0;
    int nl = right + 1, nr = left - 1;
    for (int c0 = 0; c0 < nch && !ended; c0 += UNR) {
      int r[UNR], r0[UNR], fr[UNR], cap[UNR];
      bool more[UNR];
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const int i = 32 * (c0 + u) + l, d = left + i;
        r0[u] = -1; fr[u] = 0; cap[u] = -1;
        if (i < width) {
          fr[u] = rule_from(prev, plo, ptlo, pthi, d, r0[u]);
          cap[u] = imin(m, n - d);
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const int d = left + 32 * (c0 + u) + l;
        r[u] = r0[u];
        more[u] = false;
        if (r0[u] < cap[u]) {   // (inactive lanes have r0 = cap = -1)
          const int adv = step32(A, a0 + DIR * r0[u], T, t0 + DIR * (r0[u] + d), DIR);
          r[u] = r0[u] + adv;
          more[u] = (adv == 32) && r[u] < cap[u];
          if (r[u] > cap[u]) r[u] = cap[u];
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; u++) {
        const int c = c0 + u;
        if (c < nch && !ended) {   // warp-uniform
          const int i = 32 * c + l, d = left + i;
          const bool act = i < width;
          if (more[u])
            r[u] += match_dir(A, a0 + DIR * r[u], T, t0 + DIR * (r[u] + d), cap[u] - r[u], DIR);
          int wc = 0;
          if (act) {
            cur[i] = r[u];
            if (!wide) ws.vals[off + i] = r[u];
            wc = (r[u] - r0[u] + 1 + 15) / 16;
          }
          const unsigned bm = __ballot_sync(FULL, fr[u] == -1), bp = __ballot_sync(FULL, fr[u] == 1);
          if (l == 0) { dirs[2 * c] = bm; dirs[2 * c + 1] = bp; }
          const unsigned em = __ballot_sync(FULL, act && (r[u] == m || r[u] + d == n));
          if (em) {
            const int first = __ffs(em) - 1;
            if (l <= first) { ws.cells += 1; ws.words += wc; }
            d_end = left + 32 * c + first;
            r_end = __shfl_sync(FULL, r[u], first);
            ended = true;
          } else {
            if (act) { ws.cells += 1; ws.words += wc; }
            const unsigned okm = __ballot_sync(FULL, act && (r[u] + imax(d, 0) >= lim));
            if (okm) {
              if (nl > right) nl = left + 32 * c + __ffs(okm) - 1;
              nr = left + 32 * c + (31 - __clz(okm));
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
        const int gp = (d_end + 1 < ptlo || d_end + 1 > pthi) ? -2 : prev[d_end + 1 - plo];
        if (r_end == m && 1 + gp == r_end && d_end < right) {
          d_end++;
          if (l == 0) ws.vals[off + d_end - left] = r_end;
        }
      }
      a_end = DIR * r_end;
      t_end = DIR * (r_end + d_end);
      fin_e = e; fin_d = d_end; fin_val = r_end;
      int r0x;
      from0 = rule_from(prev, plo, ptlo, pthi, d_end, r0x);
      errors = e;
      to_end = 1;
      break;
    }

    // band trimming (forward.C:235-254): [nl, nr] = first / last cell whose progress reaches the limit
    if (nl > right) {
      if (l == 0) ws.meta[e] = make_int4(left, nl, right, off);
      break;
    }
    if (l == 0) ws.meta[e] = make_int4(left, nl, nr, off);

    // best cell of the trimmed band: highest row, lowest diagonal on ties; only if it beats `longest`
    int rowbest = -0x7fffffff, rowbest_d = 0;
    for (int c = (nl - left) >> 5; c <= ((nr - left) >> 5); c++) {
      const int i = 32 * c + l, d = left + i;
      const int v = (d >= nl && d <= nr) ? cur[i] : -0x7fffffff;
      const int mx = __reduce_max_sync(FULL, v);
      if (mx > rowbest) {
        rowbest = mx;
        rowbest_d = left + 32 * c + __ffs(__ballot_sync(FULL, v == mx)) - 1;
      }
    }
    if (rowbest > longest) { longest = rowbest; best_d = rowbest_d; best_e = e; }
    const double score = OIC_DSUB(OIC_DMUL((double)longest, a.branch_match_value), (double)e);
    if (score > max_score) {
      max_score = score;
      max_score_len = longest;
      max_score_best_d = best_d;
      max_score_best_e = best_e;
    }
    plo = left; ptlo = nl; pthi = nr;
    left = nl; right = nr;
    prev = cur;
    if (!wide) { int32_t *t = sm_prev; sm_prev = sm_cur; sm_cur = t; }
    __syncwarp();
  }

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
  const int ns = traceback_fast(ws, fin_e, fin_d, from0);
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

}  // namespace oic
#endif
