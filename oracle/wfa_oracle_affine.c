/* oracle/wfa_oracle_affine.c — TEST INFRASTRUCTURE ONLY.  See wfa_oracle.h.
 *
 * Plain-C restatement of gap-affine WFA with full CIGAR as otter's `aligner2` runs it
 * (WFAlignerGapAffine(4,6,2, Alignment, MemoryMed), reference src/assemble.cpp:50):
 *   compute      wavefront/wavefront_compute_affine.c:88-248 (piggyback kernel + driver)
 *   input fetch  wavefront/wavefront_compute.c:39-84 (limits), :88-174 (null substitution)
 *   init ends    wavefront/wavefront_compute.c:325-372  (outside a wavefront's [lo,hi] == NULL)
 *   trim         wavefront/wavefront_compute.c:376-423  (M, I1, D1 separately)
 *   extend/term  wavefront/wavefront_extend.c:45-68,122-168,203-260,355-452
 *   cut-off      wavefront/wavefront_heuristic.c:285-328,470-509 (on M; I1/D1 clamped to M)
 *   CIGAR        wavefront/wavefront_backtrace_buffer.c:314-362 + wavefront_pcigar.c:203-266
 * The reference carries a packed CIGAR with every cell and spills it to a block list; the CIGAR it
 * finally emits depends only on each cell's choice of predecessor.  Here those choices are kept
 * as one byte per (score, diagonal) and walked back at the end — a different store, the same
 * choices, tie-breaks and unpack order.
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
  int32_t* buf; /* offsets, indexed buf[k - kmin] */
  int lo, hi;
  int exists; /* wavefront pointer != NULL  */
  int null;   /* ->null flag                */
} wf_t;

typedef struct {
  int lo, hi;       /* computed range of this score step (before trim) */
  uint8_t* cells;   /* per diagonal: bits0-1 M source (0 misms,1 ins,2 del,3 start); bit2 I ext; bit3 D ext */
} step_t;

/* value of a wavefront at k with the init-ends rule: outside [lo,hi] (or null) reads NULL */
static inline int rd(const wf_t* w, int k, int kmin) {
  if (!w || k < w->lo || k > w->hi) return NUL;
  return w->buf[k - kmin];
}

static void trim(wf_t* w, int kmin, int plen, int tlen) { /* wavefront_compute.c:376-409 */
  int k;
  for (k = w->hi; k >= w->lo; --k)
    if (in_bounds(w->buf[k - kmin], k, plen, tlen)) break;
  w->hi = k;
  for (k = w->lo; k <= w->hi; ++k)
    if (in_bounds(w->buf[k - kmin], k, plen, tlen)) break;
  w->lo = k;
  w->null = w->lo > w->hi;
}

int wfo_affine_cigar(const char* pattern, int plen, const char* text, int tlen, int endsfree,
                     int pb, int pe, int tb, int te, int x, int o, int e, int* score_out,
                     char* cigar, int cap) {
  const int scope = imax(o + e, x) + 1; /* wavefront_components.c:80-99 */
  const int kmin = -plen - 8, kmax = tlen + 8, span = kmax - kmin + 1;
  const int ak = tlen - plen;
  wf_t *M = calloc(scope, sizeof(wf_t)), *I = calloc(scope, sizeof(wf_t)), *D = calloc(scope, sizeof(wf_t));
  int32_t* tmpI = malloc(sizeof(int32_t) * span);
  int32_t* tmpD = malloc(sizeof(int32_t) * span);
  step_t* steps = NULL;
  int n_steps = 0, cap_steps = 0;
  int score = 0, null_steps = 0, result = -1, end_k = 0, end_off = 0, k, i;
  for (i = 0; i < scope; ++i) {
    M[i].buf = malloc(sizeof(int32_t) * span);
    I[i].buf = malloc(sizeof(int32_t) * span);
    D[i].buf = malloc(sizeof(int32_t) * span);
  }
#define PUSH_STEP(LO, HI)                                                         \
  do {                                                                            \
    if (n_steps == cap_steps) {                                                   \
      cap_steps = cap_steps ? cap_steps * 2 : 1024;                               \
      steps = realloc(steps, sizeof(step_t) * cap_steps);                         \
    }                                                                             \
    steps[n_steps].lo = (LO);                                                     \
    steps[n_steps].hi = (HI);                                                     \
    steps[n_steps].cells = (HI) >= (LO) ? calloc((HI) - (LO) + 1, 1) : NULL;      \
    ++n_steps;                                                                    \
  } while (0)

  /* ---- wavefront 0 (wavefront_align.c:113-135 / :190-245) ---- */
  M[0].exists = 1;
  if (!endsfree) {
    M[0].lo = M[0].hi = 0;
    M[0].buf[0 - kmin] = 0;
  } else {
    M[0].lo = -pb;
    M[0].hi = tb;
    for (k = -pb; k <= tb; ++k) M[0].buf[k - kmin] = k > 0 ? k : 0;
  }
  PUSH_STEP(M[0].lo, M[0].hi);
  for (k = M[0].lo; k <= M[0].hi; ++k) steps[0].cells[k - M[0].lo] = 3;

  for (;;) {
    const int sm = score % scope;
    /* ---- extend + termination + cut-off (wavefront_extend.c:355-452) ---- */
    if (!M[sm].exists) {
      if (null_steps > scope) { result = -1; goto finish; } /* WF_STATUS_UNFEASIBLE */
    } else {
      wf_t* w = &M[sm];
      for (k = w->lo; k <= w->hi; ++k) {
        int h = w->buf[k - kmin], v, n = 0;
        if (h == NUL) continue;
        v = h - k;
        while (v + n < plen && h + n < tlen && pattern[v + n] == text[h + n]) ++n;
        h += n;
        w->buf[k - kmin] = h;
        if (endsfree) {
          const int vv = h - k;
          if ((h >= tlen && plen - vv <= pe) || (vv >= plen && tlen - h <= te)) {
            end_k = k; end_off = h; result = score;
            goto backtrace;
          }
        }
      }
      if (!endsfree && w->lo <= ak && ak <= w->hi && w->buf[ak - kmin] >= tlen) {
        end_k = ak; end_off = tlen; result = score;
        goto backtrace;
      }
      if (w->lo <= w->hi) { /* wavefront_heuristic.c:470-509 */
        if (w->hi - w->lo + 1 >= 10) {
          int dmin = imax(plen, tlen), nlo, nhi, top, bot;
#define DIST(kk) (w->buf[(kk)-kmin] >= 0 ? imax(plen - (w->buf[(kk)-kmin] - (kk)), tlen - w->buf[(kk)-kmin]) : -NUL)
          for (k = w->lo; k <= w->hi; ++k) dmin = imin(dmin, DIST(k));
          top = imin(ak, w->hi);
          nlo = w->lo;
          for (k = w->lo; k < top; ++k) { if (DIST(k) - dmin <= 50) break; ++nlo; }
          bot = imax(ak, nlo);
          nhi = w->hi;
          for (k = w->hi; k > bot; --k) { if (DIST(k) - dmin <= 50) break; --nhi; }
#undef DIST
          w->lo = nlo;
          w->hi = nhi;
        }
        if (w->lo > w->hi) w->null = 1;
        /* equate I1/D1 to M (wavefront_heuristic.c:154-165) */
        if (I[sm].exists) {
          if (w->lo > I[sm].lo) I[sm].lo = w->lo;
          if (w->hi < I[sm].hi) I[sm].hi = w->hi;
          if (I[sm].lo > I[sm].hi) I[sm].null = 1;
        }
        if (D[sm].exists) {
          if (w->lo > D[sm].lo) D[sm].lo = w->lo;
          if (w->hi < D[sm].hi) D[sm].hi = w->hi;
          if (D[sm].lo > D[sm].hi) D[sm].null = 1;
        }
      }
    }
    /* ---- compute score+1 (wavefront_compute_affine.c:192-248) ---- */
    ++score;
    {
      const int so = score % scope;
      const int s_mis = score - x, s_open = score - o - e, s_ext = score - e;
      const wf_t* mm = (s_mis >= 0 && M[s_mis % scope].exists && !M[s_mis % scope].null) ? &M[s_mis % scope] : NULL;
      const wf_t* mo = (s_open >= 0 && M[s_open % scope].exists && !M[s_open % scope].null) ? &M[s_open % scope] : NULL;
      const wf_t* ie = (s_ext >= 0 && I[s_ext % scope].exists && !I[s_ext % scope].null) ? &I[s_ext % scope] : NULL;
      const wf_t* de = (s_ext >= 0 && D[s_ext % scope].exists && !D[s_ext % scope].null) ? &D[s_ext % scope] : NULL;
      int lo, hi, has_i, has_d;
      int32_t *oi, *od;
      if (!mm && !mo && !ie && !de) { /* null step (:199-206) */
        ++null_steps;
        M[so].exists = I[so].exists = D[so].exists = 0;
        PUSH_STEP(1, -1);
        continue;
      }
      null_steps = 0;
      /* limits (wavefront_compute.c:39-69); a null input counts as lo=1, hi=-1 */
      lo = mm ? mm->lo : 1;
      hi = mm ? mm->hi : -1;
      lo = imin(lo, (mo ? mo->lo : 1) - 1);
      hi = imax(hi, (mo ? mo->hi : -1) + 1);
      lo = imin(lo, (ie ? ie->lo : 1) + 1);
      hi = imax(hi, (ie ? ie->hi : -1) + 1);
      lo = imin(lo, (de ? de->lo : 1) - 1);
      hi = imax(hi, (de ? de->hi : -1) - 1);
      has_i = mo || ie; /* allocate_output, wavefront_compute.c:262-283 */
      has_d = mo || de;
      M[so].exists = 1; M[so].null = 0; M[so].lo = lo; M[so].hi = hi;
      I[so].exists = has_i; I[so].null = 0; I[so].lo = lo; I[so].hi = hi;
      D[so].exists = has_d; D[so].null = 0; D[so].lo = lo; D[so].hi = hi;
      oi = has_i ? I[so].buf : tmpI; /* else: victim wavefront */
      od = has_d ? D[so].buf : tmpD;
      PUSH_STEP(lo, hi);
      for (k = lo; k <= hi; ++k) { /* kernel :130-187 */
        const int ins_o = rd(mo, k - 1, kmin), ins_e = rd(ie, k - 1, kmin);
        const int del_o = rd(mo, k + 1, kmin), del_e = rd(de, k + 1, kmin);
        const int i_ext = ins_e >= ins_o, d_ext = del_e >= del_o; /* ties prefer extend */
        const int ins = (i_ext ? ins_e : ins_o) + 1;
        const int del = d_ext ? del_e : del_o;
        const int mis = rd(mm, k, kmin) + 1;
        int mx = imax(del, imax(mis, ins));
        int src = 0;
        if (mx == ins) src = 1;
        if (mx == del) src = 2;
        if (mx == mis) src = 0; /* last assignment wins: misms > del > ins (:166-177) */
        oi[k - kmin] = ins;
        od[k - kmin] = del;
        if (!in_bounds(mx, k, plen, tlen)) mx = NUL;
        M[so].buf[k - kmin] = mx;
        steps[score].cells[k - lo] = (uint8_t)(src | (i_ext << 2) | (d_ext << 3));
      }
      trim(&M[so], kmin, plen, tlen); /* wavefront_compute.c:410-423 */
      if (has_i) trim(&I[so], kmin, plen, tlen);
      if (has_d) trim(&D[so], kmin, plen, tlen);
    }
  }

backtrace : {
  /* walk the choices back from the end cell (equivalent of following bt_prev + pcigars) */
  int s = result, kk = end_k, st = 0 /* 0 M, 1 I, 2 D */, n_ops = 0, cap_ops = 1024;
  char* ops = malloc(cap_ops);
  int v, h, len = 0, ev = end_off - end_k, eh = end_off, mstate = 1;
#define EMIT(c) do { if (len < cap) cigar[len] = (c); ++len; } while (0)
  for (;;) {
    uint8_t c;
    if (st == 0 && s == 0) break;
    c = steps[s].cells[kk - steps[s].lo];
    if (n_ops + 1 >= cap_ops) { cap_ops *= 2; ops = realloc(ops, cap_ops); }
    if (st == 0) {
      if ((c & 3) == 0) { ops[n_ops++] = 'X'; s -= x; }
      else { /* the X pushed on top of an I/D is a fake gap-close: keep it as a marker */
        ops[n_ops++] = 'c';
        st = (c & 3) == 1 ? 1 : 2;
      }
    } else if (st == 1) {
      ops[n_ops++] = 'I';
      if (c & 4) { s -= e; kk -= 1; } else { s -= o + e; kk -= 1; st = 0; }
    } else {
      ops[n_ops++] = 'D';
      if (c & 8) { s -= e; kk += 1; } else { s -= o + e; kk += 1; st = 0; }
    }
  }
  /* start cell of wavefront 0 (wavefront_align.c:216-237) */
  h = kk > 0 ? kk : 0;
  v = kk > 0 ? 0 : -kk;
  for (i = 0; i < h; ++i) EMIT('I'); /* wavefront_backtrace_buffer.c:334-335 */
  for (i = 0; i < v; ++i) EMIT('D');
  for (i = n_ops - 1; i >= 0; --i) { /* forward replay, wavefront_pcigar.c:203-266 */
    if (mstate) { /* matches are re-extended only in the M state (:225-238) */
      while (v < plen && h < tlen && pattern[v] == text[h]) { EMIT('M'); ++v; ++h; }
    }
    if (ops[i] == 'c') { mstate = 1; continue; } /* gap-close: back to M, nothing emitted (:250-254) */
    EMIT(ops[i]);
    if (ops[i] == 'X') { ++v; ++h; mstate = 1; }
    else if (ops[i] == 'I') { ++h; mstate = 0; }
    else { ++v; mstate = 0; }
  }
  {
    const int nm = imin(ev - v, eh - h); /* last stroke, wavefront_backtrace_buffer.c:352-358 */
    for (i = 0; i < nm; ++i) EMIT('M');
    v += nm > 0 ? nm : 0;
    h += nm > 0 ? nm : 0;
  }
  while (h < tlen) { EMIT('I'); ++h; }
  while (v < plen) { EMIT('D'); ++v; }
#undef EMIT
  free(ops);
  if (score_out) *score_out = -result; /* WF_PENALTIES_GET_SW_SCORE with match 0 */
  result = len;
  goto cleanup;
}
finish:
  if (score_out) *score_out = 0;
cleanup:
  for (i = 0; i < n_steps; ++i) free(steps[i].cells);
  free(steps);
  for (i = 0; i < scope; ++i) { free(M[i].buf); free(I[i].buf); free(D[i].buf); }
  free(M); free(I); free(D); free(tmpI); free(tmpD);
  return result;
}
