/* oracle/wfa_oracle.c — TEST INFRASTRUCTURE ONLY.  See wfa_oracle.h for scope and parity pin.
 *
 * A from-scratch scalar restatement of what the reference computes; it is written for clarity
 * (one dense int32 array indexed by diagonal, no slabs, no modular buffers) and shares no code
 * with WFA2-lib.  Each block cites the reference lines whose behaviour it reproduces
 * (paths relative to /root/reference/include/WFA2-lib/ unless they start with src/).
 */
#include "wfa_oracle.h"

#include <stdlib.h>
#include <string.h>

#define NUL WFO_OFFSET_NULL
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }

/* Heuristic parameters: wavefront/wavefront_attributes.c:72-77 (wf_heuristic_wfadaptive). */
enum { WFO_MIN_WF_LEN = 10, WFO_MAX_DIST = 50, WFO_PRUNE_LEN = 1000 };

/* In-range test used by compute and trim: wavefront_compute_edit.c:88-92,
 * wavefront_compute.c:386-391 (unsigned compare so negative h or v fail too). */
static inline int in_bounds(int off, int k, int plen, int tlen) {
  return (uint32_t)off <= (uint32_t)tlen && (uint32_t)(off - k) <= (uint32_t)plen;
}

int wfo_edit_score(const char* pattern, int plen, const char* text, int tlen, int endsfree,
                   int pb, int pe, int tb, int te, wfo_counters* ctr) {
  /* Diagonals k = h - v live in [-plen, tlen]; two guard cells each side. */
  const int kmin = -plen - 3, kmax = tlen + 3, span = kmax - kmin + 1;
  int32_t* bufA = (int32_t*)malloc(sizeof(int32_t) * (size_t)span);
  int32_t* bufB = (int32_t*)malloc(sizeof(int32_t) * (size_t)span);
  int32_t* M = bufA - kmin;  /* current wavefront, M[k]  */
  int32_t* Mn = bufB - kmin; /* next wavefront           */
  const int ak = tlen - plen; /* target diagonal, wavefront_extend.c:57 */
  int lo, hi, s = 0, k, result = -1;
  int dm_ext = 0; /* minimum cut-off distance at the last extend (0 = unknown: the gate stays open) */
  if (ctr) memset(ctr, 0, sizeof(*ctr));
  for (k = kmin; k <= kmax; ++k) M[k] = Mn[k] = NUL;
  if (!endsfree) { /* wavefront_align.c:113-135 */
    lo = hi = 0;
    M[0] = 0;
  } else { /* wavefront_align.c:190-237 */
    lo = -pb;
    hi = tb;
    for (k = 1; k <= tb; ++k) M[k] = k;
    for (k = -pb; k <= 0; ++k) M[k] = 0;
  }
  for (;;) {
    /* ---- extend (wavefront_extend.c:172-260): sentinels == clamp at either sequence end ---- */
    if (ctr) {
      ctr->width_sum += hi - lo + 1;
      if (hi - lo + 1 > ctr->max_width) ctr->max_width = hi - lo + 1;
      {
        static const int edge[7] = {30, 61, 125, 189, 253, 509, 1021};
        int b = 0;
        while (b < 7 && hi - lo + 1 > edge[b]) ++b;
        ctr->width_hist[b]++;
      }
    }
    for (k = lo; k <= hi; ++k) {
      int h = M[k], v, n = 0;
      if (h == NUL) continue;
      v = h - k;
      while (v + n < plen && h + n < tlen && pattern[v + n] == text[h + n]) ++n;
      M[k] = h + n;
      if (ctr) {
        ctr->extend_bases += n;
        ctr->extend_words += n / 32 + 1;
        {
          static const int edge[7] = {31, 62, 93, 124, 248, 1116, 4000};
          int b = 0;
          while (b < 7 && n >= edge[b]) ++b;
          ctr->run_hist[b]++;
        }
      }
      if (endsfree) { /* wavefront_extend.c:122-168, tested per k in ascending order */
        const int hh = h + n, vv = hh - k;
        if ((hh >= tlen && plen - vv <= pe) || (vv >= plen && tlen - hh <= te)) {
          result = s;
          goto done;
        }
      }
    }
    if (!endsfree) { /* wavefront_extend.c:45-68 */
      if (lo <= ak && ak <= hi && M[ak] >= tlen) {
        result = s;
        goto done;
      }
    }
    /* ---- WF-adaptive cut-off, every step (wavefront_heuristic.c:470-502 -> 285-328) ---- */
    if (lo <= hi && hi - lo + 1 >= WFO_MIN_WF_LEN) {
      int dmin = imax(plen, tlen), nlo, nhi, top, bot; /* :230 */
#define DIST(kk) (M[kk] >= 0 ? imax(plen - (M[kk] - (kk)), tlen - M[kk]) : -NUL) /* :111-119 */
      for (k = lo; k <= hi; ++k) dmin = imin(dmin, DIST(k));
      dm_ext = dmin;
      top = imin(ak, hi); /* :269 */
      nlo = lo;
      for (k = lo; k < top; ++k) {
        if (DIST(k) - dmin <= WFO_MAX_DIST) break;
        ++nlo;
      }
      bot = imax(ak, nlo); /* :277, uses the already-reduced lo */
      nhi = hi;
      for (k = hi; k > bot; --k) {
        if (DIST(k) - dmin <= WFO_MAX_DIST) break;
        --nhi;
      }
#undef DIST
      if (ctr && (nlo != lo || nhi != hi)) ctr->cutoffs++;
      lo = nlo;
      hi = nhi;
    }
    /* ---- compute s+1 (wavefront_compute_edit.c:294-356, kernel :70-97) ---- */
    ++s;
    /* The kernel's gate for splitting a step around the prune (wfa_edit.cuh, wfa_edit_run): a
     * NECESSARY condition for the prune below to change anything, from the wavefront before this
     * compute.  Counted so that tests can check "prune fired => gate was open". */
    const int gate = !endsfree && hi - lo + 3 >= WFO_PRUNE_LEN &&
                     imax(abs(lo - 1 - ak), abs(hi + 1 - ak)) >= dm_ext - 1;
    if (ctr && gate) ctr->prune_gate++;
    {
      const int nlo = lo - 1, nhi = hi + 1;
      for (k = nlo; k <= nhi; ++k) {
        /* everything outside [lo,hi] reads as NULL (:311-318) */
        const int ins = (k - 1 >= lo && k - 1 <= hi) ? M[k - 1] : NUL;
        const int mis = (k >= lo && k <= hi) ? M[k] : NUL;
        const int del = (k + 1 >= lo && k + 1 <= hi) ? M[k + 1] : NUL;
        int m = imax(del, imax(ins, mis) + 1);
        if (!in_bounds(m, k, plen, tlen)) m = NUL;
        Mn[k] = m;
      }
      lo = nlo;
      hi = nhi;
      { int32_t* t = M; M = Mn; Mn = t; }
    }
    /* ---- trim ends (wavefront_compute.c:376-409) ---- */
    while (hi >= lo && !in_bounds(M[hi], hi, plen, tlen)) --hi;
    while (lo <= hi && !in_bounds(M[lo], lo, plen, tlen)) ++lo;
    if (lo > hi) { /* num_null_steps = INT_MAX -> WF_STATUS_UNFEASIBLE (edit.c:349) */
      result = -1;
      goto done;
    }
    /* ---- exact prune, end-to-end only (wavefront_compute_edit.c:214-269) ---- */
    if (!endsfree && hi - lo + 1 >= WFO_PRUNE_LEN) {
      const int ks = lo + (hi - lo) / 2;
      if (M[ks] >= 0) {
#define WORST(kk) imax(plen - (M[kk] - (kk)), tlen - M[kk])
#define BEST(kk) abs((plen - (M[kk] - (kk))) - (tlen - M[kk]))
        const int smax = WORST(ks);
        if (!(BEST(lo) <= smax && BEST(hi) <= smax)) {
          int w = 0x7fffffff, nlo = lo, nhi = hi;
          if (ctr) ctr->prunes++;
          if (ctr && !gate) ctr->prune_gate = -(1 << 30); /* would disprove the gate */
          for (k = lo; k <= hi; ++k)
            if (M[k] >= 0 && WORST(k) < w) w = WORST(k);
          for (k = lo; k <= hi; ++k) {
            if (BEST(k) <= w) break;
            ++nlo;
          }
          for (k = hi; k > nlo; --k) {
            if (BEST(k) <= w) break;
            --nhi;
          }
          lo = nlo;
          hi = nhi;
        }
#undef WORST
#undef BEST
      }
    }
  }
done:
  if (ctr) ctr->steps = s;
  free(bufA);
  free(bufB);
  return result;
}

/* src/analignments.cpp:62-101 */
double wfo_align_anreads(const char* x, uint32_t xlen, int xspan, const char* y, uint32_t ylen,
                         int yspan, int* score_out) {
  const int x_sp = (xspan & 3) == 3, y_sp = (yspan & 3) == 3;
  int score = -1;
  double d;
  if (xlen == ylen && memcmp(x, y, xlen) == 0) {
    d = 0.0; /* :64 */
  } else if ((x_sp && y_sp) || (y_sp && xlen >= ylen)) { /* :65-72 */
    const int x_is_smallest = xlen < ylen;
    const double largest = x_is_smallest ? (double)ylen : (double)xlen;
    if (x_is_smallest)
      score = wfo_edit_score(y, (int)ylen, x, (int)xlen, 0, 0, 0, 0, 0, NULL);
    else
      score = wfo_edit_score(x, (int)xlen, y, (int)ylen, 0, 0, 0, 0, 0, NULL);
    d = score / largest;
  } else if (y_sp) { /* :82-98; here xlen < ylen so length_diff > 0 */
    const int diff = (int)ylen - (int)xlen;
    int pb, pe;
    if (xspan & 1) { pb = 0; pe = diff; }
    else if (xspan & 2) { pb = diff; pe = 0; }
    else { pb = diff / 2; pe = diff / 2; }
    score = wfo_edit_score(y, (int)ylen, x, (int)xlen, 1, pb, pe, 0, 0, NULL);
    d = score / (double)xlen;
  } else {
    d = -1.0; /* :100 */
  }
  if (score_out) *score_out = score;
  return d;
}

/* src/analignments.cpp:117-124 + src/andistmat.cpp:15-21 (row-major condensed upper triangle) */
void wfo_fill_dist_matrix(const char* bases, const uint64_t* off, const uint32_t* len,
                          const uint8_t* span, const int* indeces, uint32_t n_idx, double* out) {
  uint32_t i, j;
  size_t p = 0;
  for (i = 0; i < n_idx; ++i)
    for (j = i + 1; j < n_idx; ++j) {
      const int a = indeces[i], b = indeces[j];
      out[p++] = wfo_align_anreads(bases + off[a], len[a], span ? span[a] : 3, bases + off[b],
                                   len[b], span ? span[b] : 3, NULL);
    }
}
