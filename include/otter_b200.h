/* include/otter_b200.h — the drop-in boundary: a plain C ABI over the sm_100a CUDA hot path.
 *
 * What it replaces in the reference (paths relative to holstegelab/otter @ 2573c48e):
 *   - the per-pair call chain  align_anreads (src/analignments.cpp:62-101)
 *       -> wfa::WFAligner::alignEnd2End / alignEndsFree / getAlignmentScore
 *          (include/WFA2-lib/bindings/cpp/WFAligner.cpp:75-137,208-210)
 *       -> wavefront_align (include/WFA2-lib/wavefront/wavefront_align.c:465-479)
 *     with the aligner otter constructs at src/assemble.cpp:49
 *     (WFAlignerEdit(Score, MemoryMed), default WF-adaptive(10,50,1) heuristic);
 *   - the pair loop of fill_dist_matrix (src/analignments.cpp:117-124) over a region's
 *     valid reads, batched over thousands of regions per launch.
 * Scores are the reference's integer scores (heuristic included), bit-exact.  The FP64
 * dissimilarity is formed on the host with the reference's own expression
 * (otter_b200/host/analignments_b200.cpp), never on the device.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every call returns an int status
 * (OTGPU_OK == 0); nothing throws or exits; inputs are borrowed for the duration of the call and
 * outputs are caller-allocated.  There is NO CPU fallback: without a CUDA device every compute
 * entry point fails with OTGPU_ERR_NO_DEVICE.
 *
 * A context owns one CUDA stream, grow-only device buffers and pinned staging; it is not
 * thread-safe (use one context per host thread / per device, as otter uses one aligner per
 * thread, src/assemble.cpp:45-50).
 */
#ifndef OTTER_B200_H
#define OTTER_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OTGPU_ABI_VERSION 1

enum {
  OTGPU_OK = 0,
  OTGPU_ERR_ARGS = 1,      /* NULL / inconsistent arguments                                   */
  OTGPU_ERR_NO_DEVICE = 2, /* no CUDA device / driver (there is no CPU fallback)              */
  OTGPU_ERR_CUDA = 3,      /* a CUDA call or kernel failed; see otgpu_last_cuda_error         */
  OTGPU_ERR_OOM = 4,       /* device or pinned-host allocation failed                         */
  OTGPU_ERR_STATE = 5,     /* call sequence error (e.g. align before upload)                  */
  OTGPU_ERR_LIMIT = 6      /* batch exceeds the fixed limits of otgpu_edit_scores_static      */
};

/* Score written for a pair whose wavefront became empty (WF_STATUS_UNFEASIBLE in
 * include/WFA2-lib/wavefront/wavefront_aligner.h:53-57; unreachable for edit distance). */
#define OTGPU_SCORE_UNFEASIBLE (-1)

typedef struct otgpu_ctx otgpu_ctx;

/* One alignment request.  pattern/text are read indices into the batch.
 * pattern_begin_free < 0  => alignEnd2End(pattern, text)            (WFAligner.cpp:75-90)
 * otherwise               => alignEndsFree(pattern, pbf, pef, text, tbf, tef) (:103-137). */
typedef struct {
  uint32_t pattern;
  uint32_t text;
  int32_t pattern_begin_free;
  int32_t pattern_end_free;
  int32_t text_begin_free;
  int32_t text_end_free;
} otgpu_pair;

/* One BAM record against one region, as parse_anreads hands it to parse_alignment
 * (reference src/anseqs.cpp:440-451): core.pos, core.l_qseq, core.n_cigar, the record's CIGAR
 * (cigar[cigar_off .. +n_cigar), BAM encoding len << 4 | op) and its 4-bit sequence
 * (seq4[seq_off .. +(l_qseq+1)/2), two bases per byte, high nibble first), and the region
 * [start, end] (bed.start, bed.end). */
typedef struct {
  int32_t pos;
  int32_t l_qseq;
  uint32_t n_cigar;
  uint32_t reserved;
  uint64_t cigar_off;
  uint64_t seq_off;
  int32_t start;
  int32_t end;
} otgpu_bam_read;

/* status bits of otgpu_extract_reads */
#define OTGPU_READ_OK 1u         /* PARSEMSG.successful: the record reaches the region        */
#define OTGPU_READ_SPAN_L 2u     /* PARSEMSG.spanning_l                                       */
#define OTGPU_READ_SPAN_R 4u     /* PARSEMSG.spanning_r                                       */
#define OTGPU_READ_BAD_COORDS 8u /* the reference would exit(1) or read outside the record    */

/* Per-launch work counters (diagnostics for the roofline / issue-bound argument). */
typedef struct {
  uint64_t n_pairs;
  uint64_t n_pairs_8bit;   /* pairs routed to the byte-compare variant (non-ACGT present)  */
  uint64_t n_pairs_wide;   /* pairs re-run with the wavefront in HBM (width > smem window) */
  uint64_t kernel_launches;/* kernels launched by this context since creation              */
} otgpu_stats;

/* ---- lifecycle ------------------------------------------------------------------------- */
int otgpu_abi_version(void);
int otgpu_device_count(void); /* 0 when no device/driver; never fails */
int otgpu_create(int device, otgpu_ctx** out);
void otgpu_destroy(otgpu_ctx* ctx);
const char* otgpu_strerror(int status);
const char* otgpu_last_cuda_error(const otgpu_ctx* ctx);
void* otgpu_stream(otgpu_ctx* ctx); /* the cudaStream_t every kernel of ctx is launched on */
int otgpu_get_stats(const otgpu_ctx* ctx, otgpu_stats* out);

/* Pinned host memory so callers can stage reads where H2D runs at full PCIe rate. */
int otgpu_host_alloc(size_t bytes, void** out);
void otgpu_host_free(void* p);

/* ---- one-shot calls on HOST buffers (H2D + kernels + D2H inside the call) --------------- */

/* Edit-distance scores for an explicit pair list.
 * bases: concatenated read bytes (as std::string holds them; any byte values), read r occupies
 * bases[read_off[r] .. read_off[r]+read_len[r]).  scores_out[n_pairs]. */
int otgpu_edit_scores(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                      const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                      const otgpu_pair* pairs, uint64_t n_pairs, int32_t* scores_out);

/* fill_dist_matrix for many regions at once (all listed reads valid and spanning).
 * Region g owns reads [region_start[g], region_start[g+1]); its n(n-1)/2 scores are written in
 * DistMatrix order (src/andistmat.cpp:15-21: (0,1),(0,2)..(n-2,n-1)) at
 * scores_out + score_off[g].  For pair (i,j), i<j: pattern = the longer read, ties -> read i
 * (src/analignments.cpp:67-70).  Equal strings score 0 like the reference's shortcut (:64). */
int otgpu_region_edit_scores(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len,
                             uint32_t n_reads, const uint32_t* region_start, uint32_t n_regions,
                             const uint64_t* score_off, uint64_t n_scores, int32_t* scores_out);
/* The same call for a caller that is still filling `bases` (and wants the scores as they come):
 * the batch is cut into chunks of consecutive regions (about 64 MB of reads each; layout as above,
 * region after region) and the chunks alternate between two sets of device buffers, so the
 * caller's staging copy of chunk i+1 and its use of chunk i-1's scores overlap the kernels of
 * chunk i.  Replaces the per-region loop body's call into fill_dist_matrix
 * (src/assemble.cpp:103-110 -> src/analignments.cpp:117-124) when one host thread feeds a device.
 *   wait_bases(user, upto)        called before the library first reads bases[0..upto): return
 *                                  once those bytes are in place (the read, region and score
 *                                  tables must be complete when the call is made);
 *   scores_ready(user, first, n)   called once scores_out[first..first+n) are final.
 * Either hook may be NULL.  Hooks run on the calling thread.  Scores are identical to
 * otgpu_region_edit_scores. */
typedef struct otgpu_stream_hooks {
  void (*wait_bases)(void* user, uint64_t upto_byte);
  void (*scores_ready)(void* user, uint64_t first_score, uint64_t n_scores);
  void* user;
} otgpu_stream_hooks;
int otgpu_region_edit_scores_hooked(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                    const uint64_t* read_off, const uint32_t* read_len,
                                    uint32_t n_reads, const uint32_t* region_start,
                                    uint32_t n_regions, const uint64_t* score_off,
                                    uint64_t n_scores, int32_t* scores_out,
                                    const otgpu_stream_hooks* hooks);
/* Batches larger than ~1.5 x 256 MB of reads (laid out region after region, as above) are streamed
 * through two sets of device buffers in chunks of consecutive regions, so that the upload of one
 * chunk overlaps the alignment of the previous one and device memory stays bounded; scores are
 * identical to a single batch.  Number of chunks the last otgpu_region_edit_scores call used: */
uint64_t otgpu_last_chunk_count(const otgpu_ctx* ctx);
/* The score kernel for reads up to 8 kb comes in two flavours that differ only in speed: 64-bit
 * first comparison windows (noisy reads), or 32-bit first windows with 128-bit continuations
 * (low-error reads, about 5 % faster on HiFi-shaped batches, slower on noisy ones).  Each call
 * counts its low-error pairs (>= 32 bases per score step) and the NEXT call on the context takes
 * the second flavour when they were at least 30 %; a fresh context starts with the first.
 * Returns the flavour the next call will use (0 or 1).  No counterpart in the reference: both
 * flavours compute the reference's scores. */
int otgpu_first_window(const otgpu_ctx* ctx);


/* ---- split-phase calls (inputs stay resident in HBM between phases) ---------------------- */
/* upload: async H2D of the batch on ctx's stream.  align: pack + pair expansion + alignment
 * kernels, async on ctx's stream (may be called repeatedly on the resident batch).
 * download: D2H of the scores + stream synchronize. */
int otgpu_batch_upload_regions(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                               const uint64_t* read_off, const uint32_t* read_len,
                               uint32_t n_reads, const uint32_t* region_start,
                               uint32_t n_regions, const uint64_t* score_off, uint64_t n_scores);
int otgpu_batch_upload_pairs(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                             const otgpu_pair* pairs, uint64_t n_pairs);
int otgpu_batch_align(otgpu_ctx* ctx);
int otgpu_batch_download(otgpu_ctx* ctx, int32_t* scores_out, uint64_t n_scores);
int otgpu_synchronize(otgpu_ctx* ctx);
/* Device time (CUDA events on ctx's stream) of the last otgpu_batch_align, per stage:
 * ms[0] pack, ms[1] pair expansion + work ordering, ms[2] k_align (shared-memory wavefronts, the dominant
 * kernel), ms[3] k_align wide fallback.  Waits for that align to finish. */
int otgpu_last_align_ms(otgpu_ctx* ctx, float ms[4]);

/* ---- gap-affine alignment with full CIGAR ------------------------------------------------- */
/* What otter's `aligner2` does (src/assemble.cpp:50: WFAlignerGapAffine(4,6,2, Alignment,
 * MemoryMed)) for local_realignment (src/analignments.cpp:24,30) and rapid_consensus (:262-281):
 * alignEnd2End / alignEndsFree + getAlignmentCigar, for a list of pairs in one launch.
 * Pair i's CIGAR (chars M X I D, WFA2 conventions: I consumes text, D consumes pattern) is
 * written at cigars_out + cigar_off[i], at most cigar_off[i+1]-cigar_off[i] chars (give each pair
 * pattern_len + text_len); cigar_len_out[i] = its length, or OTGPU_SCORE_UNFEASIBLE (-1), or
 * OTGPU_CIGAR_OOM (-3, WF_STATUS_OOM).  scores_out[i] = getAlignmentScore() (= -penalty).
 * Penalties: match 0, mismatch > 0, gap_open >= 0, gap_ext > 0, max(open+ext, mismatch) <= 15. */
#define OTGPU_CIGAR_OOM (-3)
int otgpu_affine_cigars(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                        const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                        const otgpu_pair* pairs, uint64_t n_pairs, int mismatch, int gap_open,
                        int gap_ext, int32_t* scores_out, char* cigars_out,
                        const uint64_t* cigar_off, int32_t* cigar_len_out);

/* ---- edit-distance alignment with full CIGAR ----------------------------------------------- */
/* What `otter compare` does per allele pair (src/compare.cpp:50-66,95: WFAlignerEdit(Alignment,
 * MemoryMed), alignEnd2End + getAlignmentScore + getAlignmentCigar): same wavefronts, heuristic
 * and exact prune as otgpu_edit_scores, plus the CIGAR (ties mismatch > deletion > insertion,
 * wavefront_compute_edit.c:152-195; linear unpack, wavefront_backtrace_buffer.c:266-312).
 * Arguments and outputs as otgpu_affine_cigars; scores_out[i] = the edit score.
 * For both functions cigars_out = cigar_off = NULL asks for scores and CIGAR lengths only
 * (what get_distances uses, src/compare.cpp:60-61): no CIGAR bytes are written or copied. */
int otgpu_edit_cigars(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                      const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                      const otgpu_pair* pairs, uint64_t n_pairs, int32_t* scores_out,
                      char* cigars_out, const uint64_t* cigar_off, int32_t* cigar_len_out);

/* ---- packed CIGAR output --------------------------------------------------------------------- */
/* The same two alignments for a caller that wants the strings themselves in bulk (rapid_consensus
 * feeding every CIGAR to the partial-order graph, src/analignments.cpp:280-282): the CIGARs are
 * packed end to end on the device and only the bytes they occupy are copied, into pinned memory
 * the context owns.  CIGAR i is (*cigars)[(*cigar_begin)[i] .. (*cigar_begin)[i] + cigar_len_out[i])
 * (no terminator; failed alignments occupy nothing); (*cigar_begin)[n_pairs] is the total.  Both
 * arrays stay valid until the next call on ctx.  scores_out / cigar_len_out as above. */
int otgpu_affine_cigars_packed(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                               const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                               const otgpu_pair* pairs, uint64_t n_pairs, int mismatch, int gap_open,
                               int gap_ext, int32_t* scores_out, int32_t* cigar_len_out,
                               const char** cigars, const uint64_t** cigar_begin);
int otgpu_edit_cigars_packed(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                             const otgpu_pair* pairs, uint64_t n_pairs, int32_t* scores_out,
                             int32_t* cigar_len_out, const char** cigars,
                             const uint64_t** cigar_begin);

/* ---- batched read extraction ------------------------------------------------------------------ */
/* get_breakpoints + parse_alignment (reference src/anseqs.cpp:286-436) for a batch of BAM records:
 * per record the CIGAR is walked on the device, the query is cut to the region and its 4-bit
 * bases are decoded to characters (seq_nt16_str).  Outputs, all indexed like `reads`:
 *   status[i]      OTGPU_READ_* bits (records without OTGPU_READ_OK yield nothing);
 *   coords[2i..]   PARSEMSG.alignment_coords after parse_alignment (what becomes ANREAD.ccoords);
 *   out_len[i]     length of the extracted sequence ("N" = 1 for a deleted region / empty cut);
 *   *bases, *begin the sequences packed end to end in pinned memory the context owns (valid
 *                  until the next call on ctx): record i is (*bases)[(*begin)[i] .. +out_len[i]),
 *                  (*begin)[n_reads] the total — directly usable as the bases / read_off /
 *                  read_len arguments of the alignment entry points above.
 * Record-level filters (mapq, flags, rq; :444,454) stay with the caller. */
int otgpu_extract_reads(otgpu_ctx* ctx, const otgpu_bam_read* reads, uint32_t n_reads,
                        const uint32_t* cigar, uint64_t n_cigar_ops, const uint8_t* seq4,
                        uint64_t n_seq_bytes, uint8_t* status, int32_t* coords, uint32_t* out_len,
                        const char** bases, const uint64_t** begin);

/* fill_dist_matrix over reads that the LAST otgpu_extract_reads call on ctx produced, without the
 * round trip through the host: the extracted sequences stay on the device, and this call builds
 * the alignment batch from them device to device (parse_alignment -> fill_dist_matrix,
 * src/anseqs.cpp:431 -> src/analignments.cpp:117-124, with nothing but the tables crossing PCIe).
 * read_index[r] = the extracted record that is read r of the batch (the caller keeps the records
 * it wants: status, spanning flags, rq filter, coverage cap); region g owns reads
 * [region_start[g], region_start[g+1]); scores as otgpu_region_edit_scores.
 * OTGPU_ERR_STATE when ctx holds no extraction result. */
int otgpu_region_edit_scores_extracted(otgpu_ctx* ctx, const uint32_t* read_index, uint32_t n_reads,
                                       const uint32_t* region_start, uint32_t n_regions,
                                       const uint64_t* score_off, uint64_t n_scores,
                                       int32_t* scores_out);

/* ---- partial-order consensus ------------------------------------------------------------------ */
/* otter's PPOA (reference src/anppoa.hpp:63-369, as rapid_consensus drives it,
 * src/analignments.cpp:258-290) for many alleles at once, one warp per graph: the backbone read
 * seeds the graph (init :63-83), every member is threaded through it along its CIGAR over
 * (backbone = pattern, member = text) with its spanning flags (insert_alignment :112-233), edge
 * weights are damped with (c, t) (adjust_weights :235-244) and the consensus is the heaviest path
 * into an ending node (consensus :305-369).  Byte-identical to the reference's string.
 *   graphs[g]   backbone read, its members [first_member, first_member + n_members), c and t;
 *   members[m]  read index, flags (bit0 is_spanning_l, bit1 is_spanning_r), CIGAR = cigars[cigar_off
 *               .. cigar_off + cigar_len) over the characters M X I D.
 * consensus_len[g] = length of graph g's consensus (may be 0: the caller substitutes "N" as the
 * reference does, src/analignments.cpp:290); the strings themselves are packed end to end in
 * memory the context owns (valid until the next call on ctx): graph g at
 * (*consensus)[(*begin)[g] .. (*begin)[g] + consensus_len[g]). */
typedef struct {
  uint32_t backbone;
  uint32_t first_member;
  uint32_t n_members;
  float c, t;
} otgpu_ppoa_graph;
typedef struct {
  uint32_t read;
  uint32_t flags;
  uint64_t cigar_off;
  uint32_t cigar_len;
  uint32_t reserved;
} otgpu_ppoa_member;
int otgpu_ppoa_consensus(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                         const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                         const otgpu_ppoa_graph* graphs, uint32_t n_graphs,
                         const otgpu_ppoa_member* members, uint32_t n_members, const char* cigars,
                         uint64_t n_cigar_bytes, int32_t* consensus_len, const char** consensus,
                         const uint64_t** begin);

/* ---- edit-distance scores without any allocation --------------------------------------------- */
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and scores as otgpu_edit_scores, but no context and NO
 * allocation of any kind (no cudaMalloc / cudaFree / cudaMallocHost, no stream or event creation):
 * the staged batch (padded bases, 2-bit copy, tables, pair list, scores) and the alignment kernels'
 * scratch (work-queue counters, heaviest-first order, overflow list, per-warp wavefront rings in
 * HBM for the pairs that outgrow the shared-memory ring) live in ONE file-scope
 * `static __device__ __align__(256)` array of OTGPU_STATIC_SCRATCH_BYTES (40 MiB), which every GPU
 * that loads the library holds until the process exits.  Hence the fixed limits below; a larger
 * batch is rejected with OTGPU_ERR_LIMIT before anything is launched.
 *   device   CUDA device ordinal; stream   a cudaStream_t of that device, or NULL for the default
 *   stream.  bases / read_off / read_len / pairs / scores_out: HOST buffers (pinned or pageable),
 *   as for otgpu_edit_scores.  The call returns after the scores have arrived (it synchronises
 *   `stream`).
 * Every call on a device shares the one array, which keeps the last call's contents; the call
 * clears what must start at zero itself.  Calls on the same device are serialised inside the
 * library per process, but work that overlaps on the GPU through any other route (another process
 * has its own copy; the same process cannot) would race: do not rely on the array between calls. */
#define OTGPU_STATIC_MAX_BASES (4u << 20)   /* bytes of reads per call                          */
#define OTGPU_STATIC_MAX_READS 4096u
#define OTGPU_STATIC_MAX_PAIRS 131072u
#define OTGPU_STATIC_MAX_READ_LEN 8192u     /* longest single read                              */
#define OTGPU_STATIC_SCRATCH_BYTES (40u << 20)
int otgpu_edit_scores_static(int device, void* stream, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                             const otgpu_pair* pairs, uint64_t n_pairs, int32_t* scores_out);

/* ---- kernel density estimates for the clustering cut ------------------------------------------ */
/* KDE::f (reference src/ankde.cpp:9-23) as otter_find_clustering_dist uses it
 * (src/otterclust.cpp:20-29): for every region g the density of its pairwise distances
 * values[value_off[g] .. value_off[g+1]) with bandwidth[g], at the n_grid abscissae of `grid`
 * (the caller builds them the way the reference does: x = 0, x += 0.0025 while x <= 1).
 * densities_out[g * n_grid + i] = f(grid[i]), bit-identical to the host computation (same
 * summation order, IEEE operations, the host libm's exp: exp_host_libm.cuh).  n_regions <= 65535
 * per call. */
int otgpu_kde_densities(otgpu_ctx* ctx, const double* values, const uint64_t* value_off,
                        uint32_t n_regions, const double* bandwidth, const double* grid,
                        uint32_t n_grid, double* densities_out);

#ifdef __cplusplus
}
#endif
#endif /* OTTER_B200_H */
