/* oracle/tools/prune_search.c — TEST INFRASTRUCTURE ONLY.
 *
 * Adversarial search for inputs on which the reference's exact prune
 * (WFA2-lib wavefront/wavefront_compute_edit.c:214-269) changes the wavefront under otter's
 * WF-adaptive(10,50,1) edit aligner.  Runs the instrumented C restatement (wfa_oracle.c, pinned to
 * the unmodified reference by tests/test_oracle.py) over several families of sequence pairs and
 * reports, per family: pairs, pairs whose wavefront reached 1000 diagonals, steps on which the
 * kernel's necessary condition ("gate") held, and actual prune firings.  Firing pairs are written
 * as FASTA-like records so that they can be replayed through the unmodified reference and
 * committed to tests/golden/.
 *
 *   gcc -O2 -o prune_search prune_search.c ../wfa_oracle.c && ./prune_search <seed> <pairs> <out> [swap 0|1] [lmax]
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../wfa_oracle.h"

static uint64_t rs;
static inline uint64_t rnd(void) { /* xorshift64* */
  rs ^= rs >> 12; rs ^= rs << 25; rs ^= rs >> 27;
  return rs * 2685821657736338717ull;
}
static inline int rint_(int lo, int hi) { return lo + (int)(rnd() % (uint64_t)(hi - lo + 1)); }
static inline double runif(void) { return (rnd() >> 11) * (1.0 / 9007199254740992.0); }
static int g_lmax = 7000; /* longest base sequence drawn */
static const char ACGT[4] = {'A', 'C', 'G', 'T'};

typedef struct { char* p; int n, cap; } Seq;
static void push(Seq* s, char c) {
  if (s->n == s->cap) { s->cap = s->cap ? 2 * s->cap : 4096; s->p = (char*)realloc(s->p, s->cap); }
  s->p[s->n++] = c;
}
static void rand_seq(Seq* s, int n, int alphabet) {
  for (int i = 0; i < n; ++i) push(s, ACGT[rnd() % alphabet]);
}
static void repeat_seq(Seq* s, const char* motif, int mlen, int n) {
  for (int i = 0; i < n; ++i) push(s, motif[i % mlen]);
}
/* copy src[a..b) into dst with substitution/insertion/deletion noise */
static void mutate(Seq* dst, const char* src, int a, int b, double sub, double ins, double del) {
  for (int i = a; i < b; ++i) {
    double u = runif();
    if (u < del) continue;
    if (u < del + ins) { push(dst, ACGT[rnd() & 3]); --i; continue; }
    if (u < del + ins + sub) { push(dst, ACGT[rnd() & 3]); continue; }
    push(dst, src[i]);
  }
}

enum { N_FAM = 10 };
static const char* fam_name[N_FAM] = {
    "random unrelated", "noisy copy 10-30%", "copy + length difference (TR expansion)",
    "shift A+B vs B+C", "low complexity / homopolymer blocks", "mosaic of blocks",
    "tail exhausted: copy + unrelated long tail", "two-letter alphabet", "periodic with phase slips",
    "prefix copy then divergent"};

static void gen(int fam, Seq* P, Seq* T) {
  P->n = T->n = 0;
  Seq base = {0, 0, 0};
  const int L = rint_(1500, g_lmax);
  switch (fam) {
    case 0:
      rand_seq(P, L, 4);
      rand_seq(T, rint_(L / 4, L), 4);
      break;
    case 1: {
      rand_seq(&base, L, 4);
      const double e = 0.10 + 0.20 * runif();
      mutate(P, base.p, 0, base.n, e / 3, e / 3, e / 3);
      mutate(T, base.p, 0, base.n, e / 3, e / 3, e / 3);
      break;
    }
    case 2: { /* flank + motif^a + flank vs flank + motif^b + flank */
      char motif[64];
      const int ml = rint_(1, 60);
      for (int i = 0; i < ml; ++i) motif[i] = ACGT[rnd() & 3];
      Seq f1 = {0, 0, 0}, f2 = {0, 0, 0};
      rand_seq(&f1, rint_(50, 2000), 4);
      rand_seq(&f2, rint_(50, 2000), 4);
      const int ra = rint_(500, 6000), rb = rint_(100, ra);
      const double e = 0.15 * runif();
      Seq a = {0, 0, 0}, b = {0, 0, 0};
      for (int i = 0; i < f1.n; ++i) { push(&a, f1.p[i]); push(&b, f1.p[i]); }
      repeat_seq(&a, motif, ml, ra);
      repeat_seq(&b, motif, ml, rb);
      for (int i = 0; i < f2.n; ++i) { push(&a, f2.p[i]); push(&b, f2.p[i]); }
      mutate(P, a.p, 0, a.n, e / 3, e / 3, e / 3);
      mutate(T, b.p, 0, b.n, e / 3, e / 3, e / 3);
      free(a.p); free(b.p); free(f1.p); free(f2.p);
      break;
    }
    case 3: { /* P = A+B, T = B+C */
      const int la = rint_(400, 2500), lb = rint_(500, 5000), lc = rint_(0, 2500);
      rand_seq(&base, la + lb + lc, 4);
      const double e = 0.08 * runif();
      mutate(P, base.p, 0, la + lb, e / 3, e / 3, e / 3);
      mutate(T, base.p, la, la + lb + lc, e / 3, e / 3, e / 3);
      break;
    }
    case 4: { /* blocks of homopolymers / short motifs of random lengths, independently drawn */
      for (int side = 0; side < 2; ++side) {
        Seq* S = side ? T : P;
        const int want = side ? rint_(L / 3, L) : L;
        while (S->n < want) {
          char motif[4];
          const int ml = rint_(1, 3), bl = rint_(20, 1500);
          for (int i = 0; i < ml; ++i) motif[i] = ACGT[rnd() % 2 + (rnd() & 1) * 2 % 4];
          repeat_seq(S, motif, ml, bl);
        }
      }
      break;
    }
    case 5: { /* mosaic: both reads assembled from a shared pool of blocks in different order */
      Seq pool[6];
      for (int i = 0; i < 6; ++i) { pool[i].p = 0; pool[i].n = pool[i].cap = 0; rand_seq(&pool[i], rint_(200, 1500), rint_(1, 4)); }
      for (int side = 0; side < 2; ++side) {
        Seq* S = side ? T : P;
        const int nb = rint_(2, 7);
        for (int j = 0; j < nb; ++j) {
          Seq* b = &pool[rnd() % 6];
          mutate(S, b->p, 0, b->n, 0.01, 0.01, 0.01);
        }
      }
      for (int i = 0; i < 6; ++i) free(pool[i].p);
      break;
    }
    case 6: { /* shared part then an unrelated tail on the pattern: text runs out early */
      const int ls = rint_(200, 3000);
      rand_seq(&base, ls, 4);
      const double e = 0.1 * runif();
      mutate(P, base.p, 0, ls, e / 3, e / 3, e / 3);
      mutate(T, base.p, 0, ls, e / 3, e / 3, e / 3);
      rand_seq(P, rint_(1000, 5000), rint_(1, 4));
      if (rnd() & 1) rand_seq(T, rint_(0, 900), rint_(1, 4));
      break;
    }
    case 7:
      rand_seq(P, L, 2);
      rand_seq(T, rint_(L / 4, L), 2);
      break;
    case 8: { /* same motif, phase slips every few hundred bases */
      char motif[16];
      const int ml = rint_(2, 12);
      for (int i = 0; i < ml; ++i) motif[i] = ACGT[rnd() & 3];
      for (int side = 0; side < 2; ++side) {
        Seq* S = side ? T : P;
        const int want = side ? rint_(L / 4, L) : L;
        while (S->n < want) {
          repeat_seq(S, motif + 0, ml, rint_(50, 800));
          if (rnd() & 1) rand_seq(S, rint_(1, 40), 4);
        }
      }
      break;
    }
    default: { /* prefix copy, then unrelated */
      const int lp = rint_(100, 3000);
      rand_seq(&base, lp, 4);
      mutate(P, base.p, 0, lp, 0.01, 0.01, 0.01);
      mutate(T, base.p, 0, lp, 0.01, 0.01, 0.01);
      rand_seq(P, rint_(500, 4000), 4);
      rand_seq(T, rint_(500, 4000), 4);
      break;
    }
  }
  free(base.p);
  if (P->n < T->n) { Seq t = *P; *P = *T; *T = t; } /* otter: pattern = the longer read */
  if (T->n == 0) push(T, 'A');
}

int main(int argc, char** argv) {
  const uint64_t seed = argc > 1 ? strtoull(argv[1], 0, 10) : 1;
  const long n_pairs = argc > 2 ? atol(argv[2]) : 1000;
  FILE* out = argc > 3 ? fopen(argv[3], "w") : NULL;
  const int swap_ok = argc > 4 ? atoi(argv[4]) : 1; /* also try text longer than pattern (facade use) */
  if (argc > 5) g_lmax = atoi(argv[5]);
  rs = seed * 0x9E3779B97F4A7C15ull + 1;
  long pairs[N_FAM] = {0}, wide[N_FAM] = {0}, gates[N_FAM] = {0}, fires[N_FAM] = {0}, bad = 0;
  Seq P = {0, 0, 0}, T = {0, 0, 0};
  for (long i = 0; i < n_pairs; ++i) {
    const int fam = (int)(i % N_FAM);
    gen(fam, &P, &T);
    const int swap = swap_ok && (rnd() % 4 == 0);
    wfo_counters c;
    const int sc = swap ? wfo_edit_score(T.p, T.n, P.p, P.n, 0, 0, 0, 0, 0, &c)
                        : wfo_edit_score(P.p, P.n, T.p, T.n, 0, 0, 0, 0, 0, &c);
    pairs[fam]++;
    if (c.max_width >= 1000) wide[fam]++;
    if (c.prune_gate < 0) { ++bad; fprintf(stderr, "GATE VIOLATION fam %d pair %ld\n", fam, i); }
    else gates[fam] += c.prune_gate;
    if (c.prunes > 0) {
      fires[fam]++;
      if (out) {
        fprintf(out, ">fam%d pair%ld swap%d score %d prunes %d\n", fam, i, swap, sc, c.prunes);
        fwrite(swap ? T.p : P.p, 1, swap ? T.n : P.n, out); fputc('\n', out);
        fwrite(swap ? P.p : T.p, 1, swap ? P.n : T.n, out); fputc('\n', out);
        fflush(out);
      }
    }
  }
  long tp = 0, tw = 0, tg = 0, tf = 0;
  for (int f = 0; f < N_FAM; ++f) {
    printf("%-45s pairs %8ld  width>=1000 %8ld  gate-open steps %8ld  pairs with prune %6ld\n", fam_name[f],
           pairs[f], wide[f], gates[f], fires[f]);
    tp += pairs[f]; tw += wide[f]; tg += gates[f]; tf += fires[f];
  }
  printf("TOTAL seed %llu: pairs %ld  width>=1000 %ld  gate-open steps %ld  pairs with prune %ld  gate violations %ld\n",
         (unsigned long long)seed, tp, tw, tg, tf, bad);
  return bad ? 2 : 0;
}
