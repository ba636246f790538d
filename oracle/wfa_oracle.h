/* oracle/wfa_oracle.h — TEST INFRASTRUCTURE ONLY (CPU restatement of the reference algorithm).
 *
 * Plain-C restatement of the part of WFA2-lib v2.2 (vendored in the reference at
 * include/WFA2-lib, pinned 52e8ac1d) that otter's hot path executes:
 *   - edit distance, score only, end-to-end and ends-free, under the default WF-adaptive(10,50,1)
 *     heuristic                           (wfo_edit_score)
 *   - gap-affine (x,o,e) with CIGAR       (wfo_affine_cigar)   [see wfa_oracle.c]
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it; the product
 * under otter_b200/ never links or loads it.
 *
 * Parity pin: asserted equal to the compiled reference (oracle/_ref/libotter_ref.so) by
 * tests/test_oracle.py in the build container and to the committed fixtures under
 * tests/golden/ (generated from the reference by tests/golden/make_golden.py) everywhere.
 */
#ifndef WFA_ORACLE_H
#define WFA_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define WFO_OFFSET_NULL (INT32_MIN / 2) /* wavefront_offset.h:46 */

typedef struct {
  int64_t steps;        /* score steps executed (= final score)                        */
  int64_t width_sum;    /* W: sum over steps of live width at extend time              */
  int64_t extend_bases; /* total matched bases found by extension                      */
  int64_t extend_words; /* E: 32-base word comparisons an LCP over 32-base words costs */
  int32_t max_width;    /* widest wavefront seen at extend time                        */
  int32_t cutoffs;      /* steps on which the adaptive cut-off shrank the wavefront    */
  int32_t prunes;       /* exact-prune triggers (width >= 1000)                        */
  int32_t prune_gate;   /* steps on which the kernel's necessary condition for a prune held
                           (negative: a prune fired with the gate closed - never)           */
  int64_t width_hist[8]; /* score steps by live width: <=30, <=61, <=125, <=189, <=253, <=509,
                            <=1021, wider (the kernel's narrow / mid / 2, 3, 4, 8, 16 rounds)   */
  int64_t run_hist[8];   /* extensions by matched bases: <31, <62, <93, <124, <248, <1116, <4000, more */
} wfo_counters;

/* Edit-distance score under WF-adaptive(10,50,1).  endsfree=0: alignEnd2End(pattern,text);
 * endsfree=1: alignEndsFree(pattern,pb,pe,text,tb,te).  Returns the score (>=0) or -1 if the
 * wavefront became empty (unreachable for edit distance).  counters may be NULL. */
int wfo_edit_score(const char* pattern, int plen, const char* text, int tlen, int endsfree,
                   int pb, int pe, int tb, int te, wfo_counters* counters);

/* align_anreads (reference src/analignments.cpp:62-101) on top of wfo_edit_score.
 * span bits: 1 = is_spanning_l, 2 = is_spanning_r.  *score_out (may be NULL) receives the raw
 * integer score, or -1 when no alignment is run (equal strings / not comparable). */
double wfo_align_anreads(const char* x, uint32_t xlen, int xspan, const char* y, uint32_t ylen,
                         int yspan, int* score_out);

/* fill_dist_matrix (src/analignments.cpp:117-124) for one region, ignore_haps = true. */
void wfo_fill_dist_matrix(const char* bases, const uint64_t* off, const uint32_t* len,
                          const uint8_t* span, const int* indeces, uint32_t n_idx, double* out);

/* Gap-affine (mismatch x, open o, extend e; match 0) with full CIGAR, MemoryMed piggyback
 * semantics, WF-adaptive(10,50,1).  Returns CIGAR length (may exceed cap; only cap chars are
 * written) or -1 if unfeasible.  *score_out = -penalty (as getAlignmentScore() reports). */
int wfo_affine_cigar(const char* pattern, int plen, const char* text, int tlen, int endsfree,
                     int pb, int pe, int tb, int te, int x, int o, int e, int* score_out,
                     char* cigar, int cap);

/* Edit distance with full CIGAR (WFAlignerEdit(Alignment, MemoryMed), src/compare.cpp:95):
 * same offsets, heuristic and prune as wfo_edit_score; predecessor ties mismatch > deletion >
 * insertion; linear CIGAR unpack.  Returns the CIGAR length (may exceed cap) or -1;
 * *score_out = the edit score getAlignmentScore() reports.  [wfa_oracle_edit_cigar.c] */
int wfo_edit_cigar(const char* pattern, int plen, const char* text, int tlen, int endsfree, int pb,
                   int pe, int tb, int te, int* score_out, char* cigar, int cap);

#ifdef __cplusplus
}
#endif
#endif
