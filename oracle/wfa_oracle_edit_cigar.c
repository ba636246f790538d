/* oracle/wfa_oracle_edit_cigar.c — TEST INFRASTRUCTURE ONLY.  See wfa_oracle.h.
 *
 * Plain-C restatement of edit-distance WFA WITH CIGAR, as otter's `compare` runs it
 * (wfa::WFAlignerEdit(Alignment, MemoryMed), reference src/compare.cpp:95, used by get_distances
 * src/compare.cpp:50-66 for the score and the CIGAR length):
 *   compute      wavefront/wavefront_compute_edit.c:152-195 (piggyback kernel: the offsets of the
 *                score-only kernel plus the predecessor, mismatch > deletion > insertion on ties)
 *   extend/term, cut-off, trim, exact prune: identical to the score-only path (wfa_oracle.c)
 *   CIGAR        wavefront/wavefront_backtrace_buffer.c:266-312 + wavefront_pcigar.c:151-199
 *                (linear unpack: before every operation the matches are re-extended)
 * The reference carries a packed CIGAR with every cell; the CIGAR it emits depends only on each
 * cell's predecessor choice, kept here as one byte per (score, diagonal).
 */
#include "wfa_oracle.h"

#include <stdlib.h>
#include <string.h>

#define NUL WFO_OFFSET_NULL
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int in_bounds(int off, int k, int plen, int tlen) {
  return (uint32_t)off <= (uint32_t)tlen && (uint32_t)(off - k) <= (uint32_t)plen;
}

typedef struct {
  int lo, hi;     /* computed range of this score step (before trim) */
  uint8_t* cells; /* per diagonal: 0 mismatch, 1 insertion, 2 deletion, 3 start */
} estep_t;

int wfo_edit_cigar(const char* pattern, int plen, const char* text, int tlen, int endsfree, int pb,
                   int pe, int tb, int te, int* score_out, char* cigar, int cap) {
  const int kmin = -plen - 3, kmax = tlen + 3, span = kmax - kmin + 1;
  int32_t* bufA = (int32_t*)malloc(sizeof(int32_t) * (size_t)span);
  int32_t* bufB = (int32_t*)malloc(sizeof(int32_t) * (size_t)span);
  int32_t* M = bufA - kmin;
  int32_t* Mn = bufB - kmin;
  const int ak = tlen - plen;
  estep_t* steps = NULL;
  int n_steps = 0, cap_steps = 0;
  int lo, hi, s = 0, k, i, result = -1, end_k = 0, end_off = 0;
#define PUSH_STEP(LO, HI)                                                    \
  do {                                                                       \
    if (n_steps == cap_steps) {                                              \
      cap_steps = cap_steps ? cap_steps * 2 : 1024;                          \
      steps = (estep_t*)realloc(steps, sizeof(estep_t) * (size_t)cap_steps); \
    }                                                                        \
    steps[n_steps].lo = (LO);                                                \
    steps[n_steps].hi = (HI);                                                \
    steps[n_steps].cells = (uint8_t*)calloc((size_t)((HI) - (LO) + 1), 1);   \
    ++n_steps;                                                               \
  } while (0)

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
  PUSH_STEP(lo, hi);
  for (k = lo; k <= hi; ++k) steps[0].cells[k - lo] = 3;

  for (;;) {
    /* ---- extend + termination (wavefront_extend.c:172-260, 45-68, 122-168) ---- */
    for (k = lo; k <= hi; ++k) {
      int h = M[k], v, n = 0;
      if (h == NUL) continue;
      v = h - k;
      while (v + n < plen && h + n < tlen && pattern[v + n] == text[h + n]) ++n;
      h += n;
      M[k] = h;
      if (endsfree) {
        const int vv = h - k;
        if ((h >= tlen && plen - vv <= pe) || (vv >= plen && tlen - h <= te)) {
          end_k = k; end_off = h; result = s;
          goto backtrace;
        }
      }
    }
    if (!endsfree && lo <= ak && ak <= hi && M[ak] >= tlen) {
      end_k = ak; end_off = tlen; result = s;
      goto backtrace;
    }
    /* ---- WF-adaptive cut-off (wavefront_heuristic.c:285-328) ---- */
    if (lo <= hi && hi - lo + 1 >= 10) {
      int dmin = imax(plen, tlen), nlo, nhi, top, bot;
#define DIST(kk) (M[kk] >= 0 ? imax(plen - (M[kk] - (kk)), tlen - M[kk]) : -NUL)
      for (k = lo; k <= hi; ++k) dmin = imin(dmin, DIST(k));
      top = imin(ak, hi);
      nlo = lo;
      for (k = lo; k < top; ++k) { if (DIST(k) - dmin <= 50) break; ++nlo; }
      bot = imax(ak, nlo);
      nhi = hi;
      for (k = hi; k > bot; --k) { if (DIST(k) - dmin <= 50) break; --nhi; }
#undef DIST
      lo = nlo;
      hi = nhi;
    }
    /* ---- compute s+1 with predecessor (wavefront_compute_edit.c:152-195) ---- */
    ++s;
    {
      const int nlo = lo - 1, nhi = hi + 1;
      PUSH_STEP(nlo, nhi);
      for (k = nlo; k <= nhi; ++k) {
        const int ins = ((k - 1 >= lo && k - 1 <= hi) ? M[k - 1] : NUL) + 1;
        const int mis = ((k >= lo && k <= hi) ? M[k] : NUL) + 1;
        const int del = (k + 1 >= lo && k + 1 <= hi) ? M[k + 1] : NUL;
        int m = imax(del, imax(ins, mis));
        int src = 0;
        if (m == ins) src = 1;
        if (m == del) src = 2;
        if (m == mis) src = 0; /* last assignment wins (:176-187) */
        steps[s].cells[k - nlo] = (uint8_t)src;
        if (!in_bounds(m, k, plen, tlen)) m = NUL;
        Mn[k] = m;
      }
      lo = nlo;
      hi = nhi;
      { int32_t* t = M; M = Mn; Mn = t; }
    }
    /* ---- trim (wavefront_compute.c:376-409) ---- */
    while (hi >= lo && !in_bounds(M[hi], hi, plen, tlen)) --hi;
    while (lo <= hi && !in_bounds(M[lo], lo, plen, tlen)) ++lo;
    if (lo > hi) { result = -1; goto finish; }
    /* ---- exact prune, end-to-end only (wavefront_compute_edit.c:214-269) ---- */
    if (!endsfree && hi - lo + 1 >= 1000) {
      const int ks = lo + (hi - lo) / 2;
      if (M[ks] >= 0) {
#define WORST(kk) imax(plen - (M[kk] - (kk)), tlen - M[kk])
#define BEST(kk) abs((plen - (M[kk] - (kk))) - (tlen - M[kk]))
        const int smax = WORST(ks);
        if (!(BEST(lo) <= smax && BEST(hi) <= smax)) {
          int w = 0x7fffffff, nlo = lo, nhi = hi;
          for (k = lo; k <= hi; ++k)
            if (M[k] >= 0 && WORST(k) < w) w = WORST(k);
          for (k = lo; k <= hi; ++k) { if (BEST(k) <= w) break; ++nlo; }
          for (k = hi; k > nlo; --k) { if (BEST(k) <= w) break; --nhi; }
          lo = nlo;
          hi = nhi;
        }
#undef WORST
#undef BEST
      }
    }
  }

backtrace : {
  int ss = result, kk = end_k, n_ops = 0, cap_ops = 1024, v, h, len = 0;
  const int ev = end_off - end_k, eh = end_off;
  char* ops = (char*)malloc((size_t)cap_ops);
#define EMIT(c) do { if (len < cap) cigar[len] = (c); ++len; } while (0)
  while (ss > 0) {
    const uint8_t c = steps[ss].cells[kk - steps[ss].lo];
    if (n_ops + 1 >= cap_ops) { cap_ops *= 2; ops = (char*)realloc(ops, (size_t)cap_ops); }
    if (c == 0) ops[n_ops++] = 'X';
    else if (c == 1) { ops[n_ops++] = 'I'; kk -= 1; }
    else { ops[n_ops++] = 'D'; kk += 1; }
    --ss;
  }
  h = kk > 0 ? kk : 0; /* start cell of wavefront 0 (wavefront_align.c:216-237) */
  v = kk > 0 ? 0 : -kk;
  for (i = 0; i < h; ++i) EMIT('I'); /* wavefront_backtrace_buffer.c:283-284 */
  for (i = 0; i < v; ++i) EMIT('D');
  for (i = n_ops - 1; i >= 0; --i) { /* wavefront_pcigar.c:172-193 */
    while (v < plen && h < tlen && pattern[v] == text[h]) { EMIT('M'); ++v; ++h; }
    EMIT(ops[i]);
    if (ops[i] == 'X') { ++v; ++h; }
    else if (ops[i] == 'I') ++h;
    else ++v;
  }
  {
    const int nm = imin(ev - v, eh - h); /* last stroke (:300-303) */
    for (i = 0; i < nm; ++i) EMIT('M');
    if (nm > 0) { v += nm; h += nm; }
  }
  while (h < tlen) { EMIT('I'); ++h; }
  while (v < plen) { EMIT('D'); ++v; }
#undef EMIT
  free(ops);
  if (score_out) *score_out = result; /* edit: the distance itself (wavefront_align.c:289) */
  result = len;
  goto cleanup;
}
finish:
  if (score_out) *score_out = 0;
cleanup:
  for (i = 0; i < n_steps; ++i) free(steps[i].cells);
  free(steps);
  free(bufA);
  free(bufB);
  return result;
}
