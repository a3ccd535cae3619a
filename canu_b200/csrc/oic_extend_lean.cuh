/*
 * oic_extend_lean.cuh -- the row loop of the banded prefix edit distance for the common case: every row at most LSW
 * diagonals wide (LSW = the staging width the kernel was instantiated with, <= SW), both live rows in shared memory.  Same recurrence, exits, tie-breaks, counters and history layout as
 * dp_run_fast (oic_extend_fast.cuh; LIBO/prefixEditDistance-forward.C:110-285, -reverse.C:130-300) -- the results are
 * bit-identical -- but written for instruction count: no generic pointers, two chunks of 32 diagonals per iteration with
 * branch-free bookkeeping (selects on ballot words instead of warp-uniform branches), one uniform branch per chunk pair for
 * the rare "a cell reached the end of a string" exit; no direction bits are stored (the traceback recomputes the
 * predecessors from the stored rows, traceback_vals), the band is trimmed by scanning the finished row from both ends.
 * Returns -1 when the rows do not fit the warp's history, -2 when a row is wider than LSW (the caller reruns the
 * extension with dp_run_fast).
 */
#ifndef OIC_EXTEND_LEAN_CUH
#define OIC_EXTEND_LEAN_CUH

#include <type_traits>

#ifndef OIC_TAIL_PASS
#define OIC_TAIL_PASS 1   /* an odd last chunk of a row runs a single-chunk copy of the pass (0: always pairs) */
#endif
#ifndef LEAN_U
#define LEAN_U 2   /* chunks of 32 diagonals in flight together */
#endif
namespace oic {

constexpr int LU = LEAN_U;
// ints per live row of the lean loop: 2 margin + staging width + idle-lane slack of the last chunk group = LSW + 8 + 32 * LU
constexpr int LEAN_WIDE = -2;
constexpr int LEAN_IDLE = (int)0x80000000;

/* Equal leading bases of the 16-base windows at A[pa ..] and T[pt ..], walking FORWARD (16 if all equal).  The lean loop
 * only walks forward: the reverse direction is run on the reverse-complement streams (same comparisons, mirrored). */
__device__ __forceinline__ uint32_t ldg_nc(const uint32_t *p) {   // the bases are "laundered" 64-bit registers: say they are global
  uint32_t v;
#if OIC_SEQ_LD == 1
  asm("ld.global.nc.L1::evict_last.u32 %0, [%1];" : "=r"(v) : "l"(p));
#else
  asm("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
#endif
  return v;
}
#ifndef OIC_SEQ_LD
#define OIC_SEQ_LD 0    /* 1: sequence window loads ask L1 to keep their lines (evict_last); measured: no change (88.38 vs 88.37 ms) */
#endif
#ifndef OIC_HIST_ST
#define OIC_HIST_ST 2   /* cache operator of the row-history stores: 0 default, 1 .cg (L2 only), 2 .cs (streaming: written once, read once by the traceback); measured 88.37 / 88.34 / 88.03 ms */
#endif
/* history store under a predicate (no branch around the store: an asm volatile inside an `if` keeps the compiler from predicating it) */
__device__ __forceinline__ void stg_s32_if(int32_t *p, int32_t v, bool on) {
#if OIC_HIST_ST == 2
  asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.cs.s32 [%0], %1; }" ::"l"(p), "r"(v), "r"((int)on) : "memory");
#elif OIC_HIST_ST == 1
  asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.cg.s32 [%0], %1; }" ::"l"(p), "r"(v), "r"((int)on) : "memory");
#else
  asm volatile("{ .reg .pred q; setp.ne.s32 q, %2, 0; @q st.global.s32 [%0], %1; }" ::"l"(p), "r"(v), "r"((int)on) : "memory");
#endif
}
/* both history stores of a chunk pair from ONE address operand (p and p + 32 ints) */
__device__ __forceinline__ void stg_s32_if2(int32_t *p, int32_t v0, bool on0, int32_t v1, bool on1) {
#if OIC_HIST_ST == 2
  asm volatile("{ .reg .pred q0, q1; setp.ne.s32 q0, %2, 0; setp.ne.s32 q1, %4, 0; @q0 st.global.cs.s32 [%0], %1; @q1 st.global.cs.s32 [%0+128], %3; }"
               ::"l"(p), "r"(v0), "r"((int)on0), "r"(v1), "r"((int)on1) : "memory");
#else
  stg_s32_if(p, v0, on0);
  stg_s32_if(p + 32, v1, on1);
#endif
}
__device__ __forceinline__ void stg_s32(int32_t *p, int32_t v) {
#if OIC_HIST_ST == 1
  asm volatile("st.global.cg.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#elif OIC_HIST_ST == 2
  asm volatile("st.global.cs.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#else
  asm volatile("st.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
#endif
}
__device__ __forceinline__ uint32_t lean_ext16(const uint32_t *Wp, int p) {
  const uint32_t *q = Wp + (p >> 4);
  uint32_t r;
  const uint32_t lo = ldg_nc(q), hi = ldg_nc(q + 1);
  asm("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(p + p));   // the shifter takes 2p mod 32 = 2 (p mod 16)
  return r;
}
__device__ __forceinline__ int lean_step(const uint32_t *Ap, int pa, const uint32_t *Tp, int pt) {
  const uint32_t x = lean_ext16(Ap, pa) ^ lean_ext16(Tp, pt);
  uint32_t y = x - 1u;
  asm("" : "+r"(y));                   // (hides the idiom: the compiler would turn it back into a count-trailing-zeros with a zero test)
  return __popc(~x & y) >> 1;          // trailing zero bits / 2; x == 0 gives 32 -> 16 without a special case
}

/* Traceback of the lean loop: the predecessor of every cell on the path is recomputed from the STORED rows with the rule of
 * the forward pass (forward.C:60-78 / 163-170: prefer d, then d-1 if strictly greater, then d+1 if strictly greater), reading
 * -2 for every cell outside the trimmed band of the row above, which is what the -2 sentinels of forward.C:155-161 make the
 * forward pass see.  The forward pass therefore stores no direction bits at all.  32 rows per batch, one row per lane: the
 * lane fetches a window of TB_WIN cells of the row above around the current diagonal (the path drifts by at most one diagonal
 * per row) and works out the direction of EVERY cell of the window (2 bits each, one word); the path is then a chain of 32
 * shuffles through those words, and the batch is re-centred early when the path leaves a window.  Leaves the events in
 * ws.stack[0 .. nev), same encoding as traceback_fast. */
constexpr int TB_HALF = 7, TB_WIN = 2 * TB_HALF + 1;   // 13 interior cells x 2 bits = 26 bits of a word
__device__ __forceinline__ int traceback_vals(const WarpScratch &ws, int fin_e, int fin_d, int from0) {
  const int l = lane_id();
  int nev = 0, dcur = fin_d, k = fin_e;
  while (k >= 1) {
    const int kj = k - l;                    // this lane's row of the batch
    const int centre = dcur;
    uint32_t fw = 0;                         // direction of the cell at window index rel: bits [2 rel, 2 rel + 2): 0 = d, 1 = d-1, 2 = d+1
    const int32_t *row = ws.vals;
    if (kj >= 1) {
      const int4 mp = ws.meta[kj - 1];       // row above: {lo, trimmed lo, trimmed hi, offset}
      row = ws.vals + mp.w - mp.x;
      int v[TB_WIN];
#pragma unroll
      for (int w = 0; w < TB_WIN; w++) {
        const int x = centre - TB_HALF + w;
        v[w] = (x >= mp.y && x <= mp.z) ? row[x] : -2;
      }
#pragma unroll
      for (int rel = 1; rel < TB_WIN - 1; rel++) {
        const int x = 1 + v[rel];
        const uint32_t c = (1 + v[rel + 1] > imax(x, v[rel - 1])) ? 2u : ((v[rel - 1] > x) ? 1u : 0u);
        fw |= c << (2 * rel);
      }
      if (from0 != FROM_NONE && kj == fin_e) fw = (from0 < 0 ? 0x55555555u : (from0 > 0 ? 0xaaaaaaaau : 0u));
    }
    const int steps = k < 32 ? k : 32;
    int rel = TB_HALF, myrel = TB_HALF, j = 0;
#pragma unroll 1
    for (; j < steps; j++) {
      if (rel < 1 || rel > TB_WIN - 2) break;            // the path left the windows: fetch again around it
      const uint32_t w = __shfl_sync(FULL, fw, j);
      myrel = (l == j) ? rel : myrel;
      const uint32_t c = (w >> (2 * rel)) & 3u;
      rel += (int)(c >> 1) - (int)(c & 1u);
    }
    // rows 0 .. j-1 of the batch are decided: an event for every row whose predecessor is on another diagonal
    const uint32_t c = (l < j) ? (fw >> (2 * myrel)) & 3u : 0u;
    const int f = (int)(c >> 1) - (int)(c & 1u);
    const unsigned em = __ballot_sync(FULL, c != 0);
    if (c != 0) {
      const int pv = row[centre - TB_HALF + myrel + f];  // the predecessor itself (never a sentinel: it won the comparison)
      ws.stack[nev + __popc(em & ((1u << l) - 1))] = (pv + 4) * 4 + (f + 1);
    }
    nev += __popc(em);
    dcur = centre + rel - TB_HALF;
    k -= j;
  }
  __syncwarp();
  return nev;
}

template <int LSW>
__device__ __forceinline__ int dp_run_lean(const int DIR, const ExtArgs &a, WarpScratch &ws, int32_t *srow0, int32_t *srow1,
                                           const uint64_t *__restrict__ A, int64_t a0, int m, const uint64_t *__restrict__ T, int64_t t0,
                                           int n, int error_limit, int &a_end, int &t_end, int &leftover, int &to_end, int &delta_len) {
  static_assert(LSW <= 1024, "the best-cell key packs the cell index into 10 bits");
  int l = lane_id();
  asm volatile("" : "+r"(l));                          // keep the lane in a register (S2R inside the row loop otherwise)
  const uint32_t *Ap = reinterpret_cast<const uint32_t *>(A) + (a0 >> 4);
  const uint32_t *Tp = reinterpret_cast<const uint32_t *>(T) + (t0 >> 4);
  asm volatile("" : "+l"(Ap));   // keep the two string bases as plain 64-bit registers: one IMAD.WIDE per window address
  asm volatile("" : "+l"(Tp));
  const int aph = (int)(a0 & 15), tph = (int)(t0 & 15);
  const int row0 = match_dir(Ap, aph, Tp, tph, m, 1);
  int32_t *prev = srow0 + 2, *cur = srow1 + 2;
  if (l < 5) prev[l - 2] = (l == 2) ? row0 : -2;
  if (l == 0) {
    ws.cells += 1;
    ws.words += (unsigned)(row0 + 1 + 15) / 16;
    ws.meta[0] = make_int4(0, 0, 0, 0);
    ws.vals[0] = row0;
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
  int next_off = 1;     // history of the warp: the rows back to back, values only (meta[e] = {lo, trimmed lo, trimmed hi, offset})
  int fin_e = -1, fin_d = 0, fin_val = 0, from0 = FROM_NONE;
  int errors = 0;
  unsigned wsum = 0;    // per-lane count of 16-base compare words (W), without the first word of the cells that ncell counts
  unsigned w_adj = 0;   // cells of the chunk group that ended the run (their words are counted in full by wsum)
  unsigned ncell = 0;   // warp-uniform count of cells (C)
  to_end = 0;
  int lim_next = ws.match_limit[1];
  const int rowcap = (int)(ws.rowcap > 0x7fffffffull ? 0x7fffffffull : ws.rowcap);
  const int max_e = imin(error_limit, ws.max_rows - 1);
  __syncwarp();

  for (int e = 1; e <= error_limit; e++) {
    left = imax(left - 1, -e);
    right = imin(right + 1, e);
    const int width = right - left + 1;
    const int nch = (width + 31) >> 5;
    if (width > LSW) { errors = LEAN_WIDE; break; }
    if (e > max_e || next_off + width > rowcap) { errors = -1; break; }
    const int off = next_off;
    next_off += width;
    int32_t *hp = ws.vals + (off + l);                   // this lane's cell of chunk 0 in the history
#ifndef OIC_HP_NOPIN
    asm volatile("" : "+l"(hp));                         // (one 64-bit register advanced per pass, not an index re-scaled per store)
#endif
    const int lim = lim_next;
    lim_next = ws.match_limit[imin(e + 1, max_e)];
    const int32_t *pp = prev + (left - plo) + l;         // this lane's cell of chunk 0 in the previous row
    int32_t *cp = cur + l;

    bool ended = false;
    int d_end = 0, r_end = 0;
    int key_mx = -1;                                     // this lane's best cell: (value << 10) | (1023 - index), see below
    int koff = 1023 - l;

    // One pass over NU chunks of 32 diagonals starting at chunk c0; returns true when the row ended in it.  Two chunks per
    // pass keep two independent dependency chains in flight; an odd last chunk takes the single-chunk copy of the pass instead
    // of dragging 32 idle lanes through a pair (rows average 3.8 chunks).
    int c0 = 0;
    auto pass = [&](auto nu_tag, auto full_tag) -> bool {
      constexpr int NU = decltype(nu_tag)::value;
      constexpr bool FULLP = decltype(full_tag)::value;   // every cell of the pass lies inside the row: no idle lanes, no predicates for them
      int r[NU], r0[NU], cap[NU], dd[NU];
#pragma unroll
      for (int u = 0; u < NU; u++) {                     // predecessor rule (forward.C:163-170), out-of-band = -2 sentinel
        const int i = 32 * (c0 + u) + l, d = left + i;
        const int g0 = pp[32 * u] + 1, gm = pp[32 * u - 1], gp = pp[32 * u + 1] + 1;
        r0[u] = imax(imax(g0, gm), gp);
        dd[u] = d;
        cap[u] = (FULLP || i < width) ? imin(m, n - d) : LEAN_IDLE;
      }
      int more = 0;                                      // bit u: the slide of chunk u goes on
      // Branch-free: a cell that does not slide (idle lane, or already at the end of a string) compares the first windows
      // of the two strings instead and discards the result, so the 4*NU loads of a pass issue back to back.
#pragma unroll
      for (int u = 0; u < NU; u++) {
        const bool go = r0[u] < cap[u];
        const int p = go ? r0[u] : 0, q = go ? r0[u] + dd[u] : 0;
        int adv = lean_step(Ap, aph + p, Tp, tph + q);
        adv = go ? adv : 0;
        r[u] = r0[u] + adv;
        more |= ((adv == 16) && (r[u] < cap[u])) ? (1 << u) : 0;
      }
      // Slides longer than 16 bases (the cells on the alignment path): few lanes, seldom both chunks of a lane, so every
      // step advances ONE chunk per lane (the lowest one still sliding) through a single copy of the compare.
      unsigned nst = 0;                                  // continuation steps of this lane in this pass
      while (__any_sync(FULL, more != 0)) {
        static_assert(NU <= 2, "the merged continuation loop is written for at most two chunks in flight");
        const bool has = more != 0, s1 = (NU == 2) && (more == 2);
        nst += has ? 1u : 0u;
        int rr = s1 ? r[NU - 1] : r[0];
        const int ds = s1 ? dd[NU - 1] : dd[0], cs = s1 ? cap[NU - 1] : cap[0];
        const int p = has ? rr : 0, q = has ? rr + ds : 0;
        int adv = lean_step(Ap, aph + p, Tp, tph + q);
        adv = has ? adv : 0;
        rr += adv;
        const bool stop = !((adv == 16) && (rr < cs));
        r[0] = (has && !s1) ? rr : r[0];
        if (NU == 2) r[NU - 1] = s1 ? rr : r[NU - 1];
        more &= (has && stop) ? (s1 ? ~2 : ~1) : ~0;
      }
      bool endp = false;
#pragma unroll
      for (int u = 0; u < NU; u++) {
        const bool act = FULLP || cap[u] != LEAN_IDLE;
        r[u] = imin(r[u], cap[u]);
        if (act) cp[32 * u] = r[u];
        if (NU == 1) stg_s32_if(hp, r[u], act);
        endp = endp || (act && r[u] == cap[u]);          // reached the end of A (r == m) or of T (r + d == n)
        // best cell of the lane: highest value, lowest index on ties, as ONE integer (idle lanes: INT_MIN << 10 = 0, never above a real cell)
        key_mx = imax(key_mx, r[u] * 1024 + (koff - 32 * u));
      }
      if (NU == 2) {
        if (FULLP) { stg_s32(hp, r[0]); stg_s32(hp + 32, r[NU - 1]); }
        else stg_s32_if2(hp, r[0], cap[0] != LEAN_IDLE, r[NU - 1], cap[NU - 1] != LEAN_IDLE);
      }
      if (__any_sync(FULL, endp)) {                      // a cell reached the end of A or T: the row stops at the first one
        unsigned em[NU];
#pragma unroll
        for (int u = 0; u < NU; u++) em[u] = __ballot_sync(FULL, (FULLP || cap[u] != LEAN_IDLE) && r[u] == cap[u]);
        // W of this pass from the slide lengths: ceil((run + 1) / 16) compare words per cell
        unsigned wc[NU];
#pragma unroll
        for (int u = 0; u < NU; u++) wc[u] = (FULLP || cap[u] != LEAN_IDLE) ? (unsigned)(r[u] - r0[u] + 16) >> 4 : 0u;
        int uu = 0;
#pragma unroll
        for (int u = NU - 1; u >= 0; u--) if (em[u]) uu = u;
        unsigned emu = 0, wcu = 0;
        int ru = 0;
#pragma unroll
        for (int u = 0; u < NU; u++) {
          if (u < uu) wsum += wc[u];
          if (u == uu) { emu = em[u]; wcu = wc[u]; ru = r[u]; }
        }
        const int first = __ffs(emu) - 1;
        if (l <= first) wsum += wcu;
        ncell += 32 * (c0 + uu) + first + 1;
        w_adj = 32 * uu + first + 1;                     // the cells of this pass are counted by wc above, not by ncell
        d_end = left + 32 * (c0 + uu) + first;
        r_end = __shfl_sync(FULL, ru, first);
        return true;
      }
      // no cell of the pass reached the end of a string: a cell slid 16 * (its continuation steps) + less than 16 bases, i.e.
      // it cost 1 + steps compare words; the "1" of every cell is added once per run from the cell count (below)
      wsum += nst;
      pp += 32 * NU; cp += 32 * NU; hp += 32 * NU; koff -= 32 * NU; c0 += NU;
      return false;
    };
#if OIC_TAIL_PASS
    int nfull = width >> 6;                              // chunk pairs that lie entirely inside the row
    asm volatile("" : "+r"(nfull));                      // (a countdown in a register: the compiler would re-derive it from the width every pass)
#pragma unroll 1
    for (; nfull > 0; nfull--) {
      if (pass(std::integral_constant<int, 2>{}, std::true_type{})) { ended = true; break; }
    }
    if (!ended && c0 + 1 < nch) ended = pass(std::integral_constant<int, 2>{}, std::false_type{});   // at most one partial pair
    if (!ended && c0 < nch) ended = pass(std::integral_constant<int, 1>{}, std::false_type{});
#else
#pragma unroll 1
    while (c0 < nch) {
      if (pass(std::integral_constant<int, 2>{}, std::false_type{})) { ended = true; break; }
    }
#endif
    int32_t *hist = ws.vals + off;
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

    // band trimming (forward.C:235-254): [nl, nr] = first / last cell whose progress reaches Edit_Match_Limit[e].  The row is
    // complete in shared memory: scan it from both ends, one chunk per step (the ends move by a few cells per row, so the
    // first step almost always decides) instead of a ballot per chunk in the loop above.
    int nl = right + 1, nr = left - 1;
    {
      unsigned b = __ballot_sync(FULL, l < width && cur[l] + imax(left + l, 0) >= lim);
      int c = 0;
      while (b == 0 && ++c < nch) {   // rare: the whole first chunk fell below the limit
        const int i = 32 * c + l;
        b = __ballot_sync(FULL, i < width && cur[i] + imax(left + i, 0) >= lim);
      }
      if (b) nl = left + 32 * c + __ffs(b) - 1;
    }
    if (nl > right) {
      if (l == 0) ws.meta[e] = make_int4(left, right + 1, right, off);
      break;
    }
    {
      int c = nch - 1;
      const int i0 = 32 * c + l;
      unsigned b = __ballot_sync(FULL, i0 < width && cur[i0] + imax(left + i0, 0) >= lim);
      while (b == 0) {                // rare; ends at the chunk that holds nl at the latest
        c--;
        const int i = 32 * c + l;
        b = __ballot_sync(FULL, i < width && cur[i] + imax(left + i, 0) >= lim);
      }
      nr = left + 32 * c + 31 - __clz(b);
    }
    if (l == 0) ws.meta[e] = make_int4(left, nl, nr, off);
    const int gkey = __reduce_max_sync(FULL, key_mx);    // highest cell of the row, lowest diagonal on ties
    int gmax = gkey >> 10;

    // best cell of the trimmed band: highest row, lowest diagonal on ties; only if it beats `longest`
    if (gmax > longest) {
      int gpos = left + 1023 - (gkey & 1023);
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
  ws.words += wsum + ((l == 0) ? ncell - w_adj : 0u);
  if (errors < 0) return errors;
  return dp_finish(DIR, ws, n, row0, fin_e, fin_d, fin_val, from0, errors, max_score_len, max_score_best_d, max_score_best_e, a_end, t_end,
                   leftover, to_end, delta_len, true);
}

}  // namespace oic
#endif
