/*
 * oic_core.cuh -- scalar (one-lane) pieces of the overlapInCore path, written once for device and host.
 *
 * On the GPU these run on one lane of the warp that owns a read pair (or one thread per pair);
 * compiled as plain C++ (g++ -DOIC_HOST_ONLY) they are unit-tested on the CPU against the stage dumps of
 * the oracle (tests/test_core_host.py).  Nothing here is a CPU fallback of the product: the host build is
 * a test harness for this header only.
 *
 * Reference semantics followed (OIC/ = src/overlapInCore/, ST/ = src/stores/ under the reference tree):
 *   seed-run bookkeeping  Add_Ref/Add_Match          OIC/overlapInCore-Find_Overlaps.C:79-212
 *   candidate merge       Add_Overlap                OIC/overlapInCore-Process_String_Overlaps.C:220-289
 *   reductions            Combine_Into_One_Olap :93-141, Merge_Intersecting_Olaps :151-207, Choose_Best_Partial :334-356
 *   seed pruning          Lies_On_Alignment          :305-326
 *   record encode         Output_Overlap / Output_Partial_Overlap  OIC/overlapInCore-Output.C:75-313, ST/ovOverlap.H:41-124,230-231
 */
#ifndef OIC_CORE_CUH
#define OIC_CORE_CUH
#include <stdint.h>

#if defined(__CUDACC__) && !defined(OIC_HOST_ONLY)
#define OIC_HD __host__ __device__ __forceinline__
#else
#define OIC_HD inline
#endif

/* IEEE double arithmetic without FMA contraction: the reference is built for baseline x86-64 (no FMA),
 * so a*b-c rounds twice (SURVEY.md 8.1 / row E-fp). */
#if defined(__CUDA_ARCH__)
#define OIC_DMUL(a, b) __dmul_rn((a), (b))
#define OIC_DSUB(a, b) __dsub_rn((a), (b))
#define OIC_DADD(a, b) __dadd_rn((a), (b))
#define OIC_DDIV(a, b) __ddiv_rn((a), (b))
#else
#define OIC_DMUL(a, b) ((double)(a) * (double)(b))
#define OIC_DSUB(a, b) ((double)(a) - (double)(b))
#define OIC_DADD(a, b) ((double)(a) + (double)(b))
#define OIC_DDIV(a, b) ((double)(a) / (double)(b))
#endif

namespace oic {

constexpr int HOPELESS_MATCH = 90;       // OIC/overlapInCore.H:122
constexpr int MAX_DISTINCT_OLAPS = 3;    // OIC/overlapInCore.H:151
constexpr int MIN_INTERSECTION = 10;     // OIC/overlapInCore.H:171
constexpr int SHIFT_SLACK = 1;           // OIC/overlapInCore.H:197
constexpr int MIN_BRANCH_END_DIST = 20;  // LIBO/prefixEditDistance.C:40
constexpr int MAX_READLEN = (1 << 21) - 1;

/* one exact-match run between a query strand and a hash read ("Match_Node_t", LIBO/prefixEditDistance.H:59-64) */
struct Seed {
  int32_t start;   // offset in the query strand
  int32_t offset;  // offset in the hash read
  int32_t len;     // bases
  int32_t str;     // hash-block string index (grouping key while sorting; unused afterwards)
};

OIC_HD int iabs(int x) { return x < 0 ? -x : x; }
OIC_HD int imin(int a, int b) { return a < b ? a : b; }
OIC_HD int imax(int a, int b) { return a > b ? a : b; }

/* ----------------------------------------------------------------------------------------------------
 * Add_Match replayed at run granularity.
 *
 * The reference feeds every k-mer hit of a pair, in (query offset ascending, hash offset descending) order,
 * into a linked list of exact-match nodes: a hit that continues a node on the same diagonal lengthens it,
 * anything else pushes a new node at the head, and a continued node is moved to the front when another
 * still-open node precedes it.  The final list order (it breaks ties when the longest seed is picked) and the
 * `consistent` flag depend on that interleaving.  Here the input is the set of maximal runs (one per node,
 * in arrival order of their first hit); the list is replayed only at the query offsets where something can
 * change: a run starts, a run ends, or the open nodes are not yet in their steady (diagonal-descending) order.
 *
 * runs[0..n)  : arrival order; node len = K + hits - 1
 * out[0..n)   : nodes in final list order, head first
 * scr[0..4n)  : scratch
 * ---------------------------------------------------------------------------------------------------- */
OIC_HD void replay_add_match(const Seed *runs, int n, int K, Seed *out, int32_t *scr, int &consistent,
                             int &diag_ct, int &diag_bgn, int &diag_end) {
  consistent = 1;
  diag_ct = 0;
  diag_bgn = MAX_READLEN;
  diag_end = 0;
  for (int i = 0; i < n; i++) {
    int hits = runs[i].len - K + 1;
    diag_ct += hits;
    diag_bgn = imin(diag_bgn, runs[i].start);
    diag_end = imax(diag_end, runs[i].start + hits - 1);
  }
  if (n == 1) { out[0] = runs[0]; return; }

  int32_t *act = scr;          // open nodes in list order                  [n]
  int32_t *dq = scr + n;       // next order being built (deque)            [2n]
  int32_t *R = scr + 3 * n;    // working copy of act with tombstones (-1)  [n]
  int nact = 0, ir = 0, out_tail = n, last_dead_head = -1;

#define OIC_ES(x) (runs[x].start + runs[x].len - K + 1)      /* expected_start once the run is complete */
#define OIC_DG(x) (runs[x].offset - runs[x].start)

  int o = runs[0].start;
  while (ir < n || nact > 0) {
    // ---- one step at query offset o
    int nR = nact, firstR = 0, liveR = nact;
    for (int i = 0; i < nact; i++) R[i] = act[i];
    int head = n, tail = n;  // deque occupies dq[head..tail)
    int had_new = 0;

    // continuing nodes in diagonal-descending order: selection by repeated max (nact is tiny in practice)
    int last_diag = 0x7fffffff;
    int jn = ir;  // new runs at this offset: ir..; already diagonal-descending
    while (true) {
      int bx = -1, bd = -0x7fffffff;
      for (int i = 0; i < nR; i++) {
        int x = R[i];
        if (x < 0 || OIC_ES(x) <= o) continue;  // tombstone or dying
        int d = OIC_DG(x);
        if (d < last_diag && d > bd) { bd = d; bx = i; }
      }
      int have_new = (jn < n && runs[jn].start == o);
      if (bx < 0 && !have_new) break;
      int take_new = have_new && (bx < 0 || OIC_DG(jn) > bd);
      if (take_new) {
        int nd = OIC_DG(jn);
        if (tail - head + liveR > 0) consistent = 0;  // num_checked > 0
        else if (last_dead_head >= 0) {
          int dl = OIC_DG(last_dead_head), esl = OIC_ES(last_dead_head);
          if (iabs(dl - nd) > 3 || o < esl + K - 2) consistent = 0;
        }
        dq[--head] = jn;
        jn++;
        had_new = 1;
        last_diag = nd;
      } else {
        int x = R[bx];
        while (firstR < nR && R[firstR] < 0) firstR++;
        if (bx == firstR) dq[tail++] = x;  // stays where it is: joins the processed prefix
        else dq[--head] = x;               // moved to the front
        R[bx] = -1;
        liveR--;
        last_diag = bd;
      }
    }
    ir = jn;

    // whatever is left in R ended before this offset: it sinks behind the open nodes, order kept
    int ndead = liveR;
    if (ndead > 0) {
      int w = out_tail - ndead, first = 1;
      for (int i = 0; i < nR; i++)
        if (R[i] >= 0) {
          if (first) { last_dead_head = R[i]; first = 0; }
          out[w++] = runs[R[i]];
        }
      out_tail -= ndead;
    }

    int changed = had_new || ndead > 0;
    int nn = tail - head;
    if (!changed)
      for (int i = 0; i < nn; i++)
        if (act[i] != dq[head + i]) { changed = 1; break; }
    for (int i = 0; i < nn; i++) act[i] = dq[head + i];
    nact = nn;

    // ---- next offset that needs an explicit step
    if (nact > 0 && changed) { o = o + 1; continue; }
    int nxt = 0x7fffffff;
    if (ir < n) nxt = runs[ir].start;
    for (int i = 0; i < nact; i++) nxt = imin(nxt, OIC_ES(act[i]));
    o = nxt;
  }
#undef OIC_ES
#undef OIC_DG
  // (the loop only ends once every node has sunk into out[])
}

/* ---------------------------------------------------------------------------------------------------- */
struct OlapInfo {   // Olap_Info_t, OIC/overlapInCore.H:258-267 (without the 8 MB delta array)
  int32_t s_lo, s_hi, t_lo, t_hi;
  double quality;
  int32_t delta_ct;
  int32_t s_left_boundary, s_right_boundary, t_left_boundary, t_right_boundary;
  int32_t min_diag, max_diag;
};

/* returns the slot whose alignment -- and with it the delta list the reference copies into Olap_Info_t.delta
 * (Process_String_Overlaps.C:253-256,278-282; read only by the -w window filter) -- was set, or -1 */
OIC_HD int add_overlap(int partial, int s_lo, int s_hi, int t_lo, int t_hi, double qual, int delta_ct,
                       OlapInfo *olap, int &ct) {
  if (!partial) {
    int new_diag = t_lo - s_lo;
    for (int i = 0; i < ct; i++) {
      OlapInfo &o = olap[i];
      int old_diag = o.t_lo - o.s_lo;
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

constexpr int QUALITY_CUTOFF = 20;                          // OIC/overlapInCore.H:190
constexpr int BAD_WINDOW_LEN = 50, BAD_WINDOW_VALUE = 8 * QUALITY_CUTOFF;     // OIC/overlapInCore.H:80-84
constexpr int BAD_LONG_WINDOW_LEN = 100, BAD_LONG_WINDOW_VALUE = 240;         // Process_String_Overlaps.C:616

/* Has_Bad_Window (Process_String_Overlaps.C:365-388) for BOTH windows in one streaming pass: feed the q_diff values in order;
 * `bad` gets bit 0 when some 50-long stretch sums to >= 160 and bit 1 when some 100-long stretch sums to >= 240.  ring[] holds
 * the last 100 values. */
struct WindowScan {
  int n, sum_s, sum_l, bad;
  OIC_HD void reset() { n = sum_s = sum_l = bad = 0; }
  OIC_HD void push(signed char *ring, int d) {
    const int at = n % BAD_LONG_WINDOW_LEN;
    if (n >= BAD_LONG_WINDOW_LEN) sum_l -= ring[at];
    if (n >= BAD_WINDOW_LEN) sum_s -= ring[(n - BAD_WINDOW_LEN) % BAD_LONG_WINDOW_LEN];
    ring[at] = (signed char)d;
    sum_s += d; sum_l += d;
    n++;
    if (n >= BAD_WINDOW_LEN && sum_s >= BAD_WINDOW_VALUE) bad |= 1;
    if (n >= BAD_LONG_WINDOW_LEN && sum_l >= BAD_LONG_WINDOW_VALUE) bad |= 2;
  }
};

OIC_HD void combine_into_one(OlapInfo *olap, int ct, int *deleted) {
  int best = 0;
  int mn = olap[0].min_diag, mx = olap[0].max_diag;
  int sl = olap[0].s_left_boundary, sr = olap[0].s_right_boundary;
  int tl = olap[0].t_left_boundary, tr = olap[0].t_right_boundary;
  for (int i = 1; i < ct; i++) {
    if (olap[i].quality < olap[best].quality) best = i;
    mn = imin(mn, olap[i].min_diag); mx = imax(mx, olap[i].max_diag);
    sl = imin(sl, olap[i].s_left_boundary); sr = imax(sr, olap[i].s_right_boundary);
    tl = imin(tl, olap[i].t_left_boundary); tr = imax(tr, olap[i].t_right_boundary);
  }
  olap[best].min_diag = mn; olap[best].max_diag = mx;
  olap[best].s_left_boundary = sl; olap[best].s_right_boundary = sr;
  olap[best].t_left_boundary = tl; olap[best].t_right_boundary = tr;
  for (int i = 0; i < ct; i++) deleted[i] = (i != best);
}

OIC_HD void merge_intersecting(OlapInfo *p, int ct, int *deleted) {
  for (int i = 0; i < ct - 1; i++)
    for (int j = i + 1; j < ct; j++) {
      if (deleted[i] || deleted[j]) continue;
      int lo = p[i].min_diag, hi = p[i].max_diag;
      if ((lo <= 0 && p[j].min_diag > 0) || (lo > 0 && p[j].min_diag <= 0)) continue;
      if ((lo >= 0 && p[j].t_right_boundary - lo - p[j].s_left_boundary >= MIN_INTERSECTION) ||
          (lo <= 0 && p[j].s_right_boundary + lo - p[j].t_left_boundary >= MIN_INTERSECTION) ||
          (hi >= 0 && p[j].t_right_boundary - hi - p[j].s_left_boundary >= MIN_INTERSECTION) ||
          (hi <= 0 && p[j].s_right_boundary + hi - p[j].t_left_boundary >= MIN_INTERSECTION)) {
        int k, x;
        if (p[i].quality < p[j].quality) { k = i; x = j; } else { k = j; x = i; }
        deleted[x] = 1;
        p[k].min_diag = imin(p[k].min_diag, p[x].min_diag);
        p[k].max_diag = imax(p[k].max_diag, p[x].max_diag);
        p[k].s_left_boundary = imin(p[k].s_left_boundary, p[x].s_left_boundary);
        p[k].s_right_boundary = imax(p[k].s_right_boundary, p[x].s_right_boundary);
        p[k].t_left_boundary = imin(p[k].t_left_boundary, p[x].t_left_boundary);
        p[k].t_right_boundary = imax(p[k].t_right_boundary, p[x].t_right_boundary);
      }
    }
}

OIC_HD void choose_best_partial(OlapInfo *olap, int ct, int *deleted) {
  int best = 0;
  double matching = OIC_DMUL(OIC_DSUB(1.0, olap[0].quality),
                             (double)(2 + olap[0].s_hi - olap[0].s_lo + olap[0].t_hi - olap[0].t_lo));
  for (int i = 1; i < ct; i++) {
    double mb = OIC_DMUL(OIC_DSUB(1.0, olap[i].quality),
                         (double)(2 + olap[i].s_hi - olap[i].s_lo + olap[i].t_hi - olap[i].t_lo));
    if (matching < mb || (matching == mb && olap[i].quality < olap[best].quality)) best = i;
  }
  for (int i = 0; i < ct; i++) deleted[i] = (i != best);
}

OIC_HD bool lies_on_alignment(int start, int offset, int s_lo, int t_lo, const int32_t *delta, int delta_ct) {
  int diag = t_lo - s_lo, new_diag = offset - start;
  for (int i = 0; i < delta_ct; i++) {
    int dl = delta[i];
    s_lo += iabs(dl);
    if (start < s_lo) return iabs(new_diag - diag) <= SHIFT_SLACK;
    if (dl < 0) diag++;
    else { s_lo++; diag--; }
  }
  return iabs(new_diag - diag) <= SHIFT_SLACK;
}

/* ---------------------------------------------------------------------------------------------------- */
OIC_HD uint32_t encode_evalue(double q) {   // ST/ovOverlap.H:41-47
  return (q < (4095 / 10000.0)) ? (uint32_t)(int)OIC_DADD(OIC_DMUL(10000.0, q), 0.5) : 4095u;
}

/* file words: a_iid, b_iid, dat0.hi, dat0.lo, dat1.hi, dat1.lo  (ST/ovStoreFile.C:196-221) */
OIC_HD void pack_record(uint32_t *w, uint32_t a, uint32_t b, uint32_t ahg5, uint32_t ahg3, uint32_t bhg5, uint32_t bhg3,
                        uint32_t span, uint32_t evalue, uint32_t flipped, uint32_t obt, uint32_t dup, uint32_t utg) {
  const uint64_t M21 = (1ull << 21) - 1;
  uint64_t d0 = ((uint64_t)ahg5 & M21) | (((uint64_t)ahg3 & M21) << 21) | (((uint64_t)evalue & 4095) << 42) |
                ((uint64_t)(flipped & 1) << 54) | ((uint64_t)(obt & 1) << 55) | ((uint64_t)(dup & 1) << 56) |
                ((uint64_t)(utg & 1) << 57);
  uint64_t d1 = ((uint64_t)bhg5 & M21) | (((uint64_t)bhg3 & M21) << 21) | (((uint64_t)span & M21) << 42);
  w[0] = a; w[1] = b;
  w[2] = (uint32_t)(d0 >> 32); w[3] = (uint32_t)d0;
  w[4] = (uint32_t)(d1 >> 32); w[5] = (uint32_t)d1;
}

/* Output_Overlap.  Returns 1 if the overlap is a containment (bhg <= 0), 0 for a dovetail. */
OIC_HD int output_overlap(uint32_t *w, uint32_t s_id, int s_len, int dir, uint32_t t_id, int t_len, const OlapInfo &o) {
  uint32_t span = (uint32_t)((o.s_hi - o.s_lo) + (o.t_hi - o.t_lo) + o.delta_ct) / 2;
  int s_rh = s_len - o.s_hi - 1, t_rh = t_len - o.t_hi - 1;
  bool sleft = (o.s_lo > o.t_lo) || (o.s_lo == o.t_lo && s_rh > t_rh);
  uint32_t a = sleft ? s_id : t_id, b = sleft ? t_id : s_id;
  int orient, ahg, bhg;   // 0 N, 1 I, 2 O
  if (sleft) { orient = (dir == 0) ? 0 : 2; ahg = o.s_lo; bhg = t_rh - s_rh; }
  else       { orient = (dir == 0) ? 0 : 1; ahg = o.t_lo; bhg = s_rh - t_rh; }
  if (orient == 2 && s_rh >= t_rh) { orient = 1; ahg = -(t_rh - s_rh); bhg = -(o.s_lo); }
  int A, B, flipped;
  if (orient == 2) { A = -bhg; B = -ahg; flipped = 1; }
  else { A = ahg; B = bhg; flipped = (orient == 1); }
  uint32_t ahg5 = (A < 0) ? 0 : A, bhg5 = (A < 0) ? -A : 0;   // a_hang(), ST/ovOverlap.H:230
  uint32_t bhg3 = (B < 0) ? 0 : B, ahg3 = (B < 0) ? -B : 0;   // b_hang(), ST/ovOverlap.H:231
  pack_record(w, a, b, ahg5, ahg3, bhg5, bhg3, span, encode_evalue(o.quality), flipped, 0, 0, 1);
  return bhg <= 0;
}

OIC_HD void output_partial(uint32_t *w, uint32_t s_id, uint32_t t_id, int dir, const OlapInfo &o, int s_len, int t_len) {
  uint32_t span = (uint32_t)((o.s_hi - o.s_lo) + (o.t_hi - o.t_lo) + o.delta_ct) / 2;
  uint32_t ahg5, ahg3, bhg5, bhg3;
  if (dir == 0) { ahg5 = o.s_lo; ahg3 = s_len - (o.s_hi + 1); bhg5 = o.t_lo; bhg3 = t_len - (o.t_hi + 1); }
  else { ahg5 = s_len - (o.s_hi + 1); ahg3 = o.s_lo; bhg5 = t_len - (o.t_hi + 1); bhg3 = o.t_lo; }
  pack_record(w, s_id, t_id, ahg5, ahg3, bhg5, bhg3, span, encode_evalue(o.quality), dir != 0, 1, 1, 0);
}

/* occurrence entry of the k-mer index: [63:33] string, [32:12] offset, [2] previous base valid, [1:0] previous base */
OIC_HD uint64_t occ_pack(uint32_t str, uint32_t off, uint32_t prev_valid, uint32_t prev_base) {
  return ((uint64_t)str << 33) | ((uint64_t)off << 12) | ((uint64_t)(prev_valid & 1) << 2) | (prev_base & 3);
}
OIC_HD uint32_t occ_str(uint64_t o) { return (uint32_t)(o >> 33); }
OIC_HD uint32_t occ_off(uint64_t o) { return (uint32_t)(o >> 12) & 0x1fffff; }

OIC_HD uint64_t hash_key(uint64_t k) {   // murmur3 finaliser
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
  return k;
}

}  // namespace oic
#endif
