// integration/otter_shim.cpp — HARNESS, not product: one set of extern "C" entry points over
// otter's own call surface, so that tests and bench.py can drive the same calls in two builds.
//
// Nothing here re-implements an algorithm: every entry point constructs otter's own objects
// (wfa::WFAlignerEdit / wfa::WFAlignerGapAffine, ANREAD, DistMatrix) exactly the way
// src/assemble.cpp:49-50 does and calls the function of that name.  Built twice:
//   oracle/_ref/libotter_ref.so             (oracle/Makefile) against the UNMODIFIED reference:
//       WFA2-lib's own WFAligner.hpp + src/analignments.cpp - the oracle and the CPU baseline.
//       Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
//       load it; the product (otter_b200/) never does.
//   integration/_build/libotter_intree.so   (integration/Makefile) -DOTTER_B200_IN_TREE: the same
//       reference pipeline objects (otterclust, ankde, anppoa, anseqs, hclust-cpp) with
//       otter_b200/host/{analignments,realign,wfaligner,otterclust}_b200.cpp in place of
//       analignments.cpp + WFA2, i.e. otter running on the GPU hot path.  The GPU tests compare
//       whole-region results (labels, alleles) between the two libraries.
#include "bindings/cpp/WFAligner.hpp"
#ifdef OTTER_B200_IN_TREE
#include "analignments_b200.hpp"
#include "otterclust_b200.hpp"
#include "wave_b200.hpp"
#else
#include "analignments.hpp"
#endif
#include "andistmat.hpp"
#include "anfahelper.hpp"
#include "ankde.hpp"
#include "anppoa.hpp"
#include "anseqs.hpp"
#include "fastcluster.h"
#ifndef OTTER_B200_IN_TREE  // (no include guard in that header; otterclust_b200.hpp brings it in-tree)
#include "otterclust.hpp"
#endif

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

// Defined (external linkage, not in the header) at src/analignments.cpp:62.
double align_anreads(wfa::WFAligner& aligner, ANREAD& read_x, ANREAD& read_y);
// src/anseqs.cpp:218-227,412: the class is local to that source file, so its declaration is
// repeated here to reach parse_alignment (members are defined in anseqs.o)
class PARSEMSG {
 public:
  bool successful;
  bool spanning_l;
  bool spanning_r;
  std::pair<int, int> alignment_coords;
  PARSEMSG();
  bool is_spanning() const;
  void transfer_status(ANREAD&) const;
};
void parse_alignment(const int& rstart, const int& rend, bam1_t* read, PARSEMSG& msg, std::string& seq);

namespace {
struct Aligners {
  // src/assemble.cpp:49-50
  wfa::WFAlignerEdit edit{wfa::WFAligner::Score, wfa::WFAligner::MemoryMed};
  wfa::WFAlignerGapAffine affine{4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed};
  // src/compare.cpp:95
  wfa::WFAlignerEdit edit_cigar{wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed};
};
Aligners& tls() {
  thread_local Aligners a;
  return a;
}

ANREAD make_read(const char* s, uint32_t len, int span_l, int span_r) {
  ANREAD r;
  r.seq.assign(s, len);
  r.is_spanning_l = span_l != 0;
  r.is_spanning_r = span_r != 0;
  r.rq = 1.0;
  r.ccoords = {0, (int)len};
  return r;
}

void build_reads(const char* bases, const uint64_t* off, const uint32_t* len, const uint8_t* span,
                 uint32_t first, uint32_t last, std::vector<ANREAD>& reads) {
  reads.clear();
  reads.reserve(last - first);
  for (uint32_t i = first; i < last; ++i) {
    const int sl = span ? (span[i] & 1) : 1, sr = span ? ((span[i] >> 1) & 1) : 1;
    reads.emplace_back(make_read(bases + off[i], len[i], sl, sr));
  }
}
}  // namespace

extern "C" {

int ref_edit_e2e(const char* p, int plen, const char* t, int tlen) {
  auto& a = tls().edit;
  a.alignEnd2End(p, plen, t, tlen);
  return a.getAlignmentScore();
}

int ref_edit_endsfree(const char* p, int plen, int pb, int pe, const char* t, int tlen, int tb,
                      int te) {
  auto& a = tls().edit;
  a.alignEndsFree(p, plen, pb, pe, t, tlen, tb, te);
  return a.getAlignmentScore();
}

// which: 0 = gap-affine (4,6,2) aligner2 of assemble.cpp:50 ; 1 = edit+CIGAR of compare.cpp:95.
// endsfree: 0 = alignEnd2End, 1 = alignEndsFree.  Returns the CIGAR length (chars written
// without terminator, truncated to cap); *score = getAlignmentScore(); *status = return value.
int ref_cigar(int which, int endsfree, const char* p, int plen, int pb, int pe, const char* t,
              int tlen, int tb, int te, int* score, int* status, char* cigar, int cap) {
  wfa::WFAligner* a = which == 0 ? static_cast<wfa::WFAligner*>(&tls().affine)
                                 : static_cast<wfa::WFAligner*>(&tls().edit_cigar);
  int st = endsfree ? a->alignEndsFree(p, plen, pb, pe, t, tlen, tb, te)
                    : a->alignEnd2End(p, plen, t, tlen);
  if (status) *status = st;
  if (score) *score = a->getAlignmentScore();
  std::string c = a->getAlignmentCigar();
  int n = (int)c.size();
  if (cigar && cap > 0) std::memcpy(cigar, c.data(), n < cap ? n : cap);
  return n;
}

#ifndef OTTER_B200_IN_TREE
}  // extern "C"
// src/compare.cpp:17-24 and :50 (defined there, not declared in a header)
class DistCompare {
 public:
  int i;
  int j;
  double edit;
  double ops;
  DistCompare(int _i, int _j, double _e, double _o) : i(_i), j(_j), edit(_e), ops(_o) {}
};
void get_distances(wfa::WFAligner& aligner, std::vector<ANALLELE>& subjs,
                   std::vector<ANALLELE>& querys, std::vector<DistCompare>& distances);
extern "C" {
// The reference's own get_distances (src/compare.cpp:50-66) with the aligner compare builds
// (:95).  Reads [0,n_subj) are the subjects, the rest the queries; out = rows {i, j, edit, ops}.
void ref_get_distances(const char* bases, const uint64_t* off, const uint32_t* len, uint32_t n_subj,
                       uint32_t n_query, double* out) {
  std::vector<ANALLELE> subjs, querys;
  for (uint32_t i = 0; i < n_subj; ++i) subjs.emplace_back(std::string(bases + off[i], len[i]));
  for (uint32_t i = n_subj; i < n_subj + n_query; ++i)
    querys.emplace_back(std::string(bases + off[i], len[i]));
  std::vector<DistCompare> d;
  get_distances(tls().edit_cigar, subjs, querys, d);
  for (size_t r = 0; r < d.size(); ++r) {
    out[4 * r] = d[r].i; out[4 * r + 1] = d[r].j; out[4 * r + 2] = d[r].edit; out[4 * r + 3] = d[r].ops;
  }
}
#endif

// span bits: 1 = is_spanning_l, 2 = is_spanning_r.
double ref_align_anreads(const char* x, uint32_t xlen, int xspan, const char* y, uint32_t ylen,
                         int yspan) {
  ANREAD rx = make_read(x, xlen, xspan & 1, xspan & 2);
  ANREAD ry = make_read(y, ylen, yspan & 1, yspan & 2);
  return align_anreads(tls().edit, rx, ry);
}

// One region: reads [0,n_reads) given as concatenated bases + offsets + lengths (+ span bits,
// NULL = all spanning); indeces = positions of the valid reads; out = n_idx*(n_idx-1)/2 doubles.
void ref_fill_dist_matrix(const char* bases, const uint64_t* off, const uint32_t* len,
                          const uint8_t* span, uint32_t n_reads, const int* indeces,
                          uint32_t n_idx, double* out) {
  std::vector<ANREAD> reads;
  build_reads(bases, off, len, span, 0, n_reads, reads);
  std::vector<int> idx(indeces, indeces + n_idx);
  DistMatrix dm(n_idx);
  fill_dist_matrix(true, tls().edit, reads, idx, dm);
  if (!dm.values.empty()) std::memcpy(out, dm.values.data(), dm.values.size() * sizeof(double));
}

// Many regions, all reads valid+spanning: region r owns reads [region_start[r], region_start[r+1])
// and writes its condensed matrix at out + out_off[r].  `threads` workers take static contiguous
// blocks of regions, one WFAlignerEdit each — the scheme of BS::thread_pool::parallelize_loop as
// used by src/assemble.cpp:43-50.  Returns wall seconds spent inside fill_dist_matrix calls
// (read construction excluded: otter builds ANREADs in parse_anreads, outside the hot path).
double ref_fill_dist_matrix_regions(const char* bases, const uint64_t* off, const uint32_t* len,
                                    const uint32_t* region_start, uint32_t n_regions, double* out,
                                    const uint64_t* out_off, int threads) {
  if (threads < 1) threads = 1;
  if ((uint32_t)threads > n_regions) threads = n_regions ? (int)n_regions : 1;
  std::vector<std::vector<std::vector<ANREAD>>> per_thread(threads);
  std::vector<std::pair<uint32_t, uint32_t>> blocks(threads);
  for (int t = 0; t < threads; ++t) {
    uint32_t a = (uint64_t)n_regions * t / threads, b = (uint64_t)n_regions * (t + 1) / threads;
    blocks[t] = {a, b};
    per_thread[t].resize(b - a);
    for (uint32_t r = a; r < b; ++r)
      build_reads(bases, off, len, nullptr, region_start[r], region_start[r + 1],
                  per_thread[t][r - a]);
  }
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    pool.emplace_back([&, t]() {
      wfa::WFAlignerEdit aligner(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed);
      for (uint32_t r = blocks[t].first; r < blocks[t].second; ++r) {
        auto& reads = per_thread[t][r - blocks[t].first];
        std::vector<int> idx(reads.size());
        for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
        DistMatrix dm((uint32_t)reads.size());
        fill_dist_matrix(true, aligner, reads, idx, dm);
        if (!dm.values.empty())
          std::memcpy(out + out_off[r], dm.values.data(), dm.values.size() * sizeof(double));
      }
    });
  }
  for (auto& th : pool) th.join();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// invalid_reassignment (src/analignments.cpp:126-177) on one region; labels updated in place.
void ref_invalid_reassignment(const char* bases, const uint64_t* off, const uint32_t* len,
                              const uint8_t* span, uint32_t n_reads, double min_sim,
                              double max_error, int total_alleles, int* labels) {
  std::vector<ANREAD> reads;
  build_reads(bases, off, len, span, 0, n_reads, reads);
  std::vector<int> lab(labels, labels + n_reads);
  invalid_reassignment(true, min_sim, max_error, total_alleles, reads, tls().edit, lab);
  std::memcpy(labels, lab.data(), n_reads * sizeof(int));
}

int ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

// test/ppoa_test.cpp:10-37 (runtest) with the 4-argument insert_alignment the current header has:
// every sequence aligned end-to-end to the first with aligner2, inserted into the partial-order
// graph, consensus with weights (n*0.4, 0.3).  Returns the consensus length.
int ref_ppoa_consensus(const char* bases, const uint64_t* off, const uint32_t* len, uint32_t n,
                       char* out, int cap) {
  std::vector<std::string> seqs;
  for (uint32_t i = 0; i < n; ++i) seqs.emplace_back(bases + off[i], len[i]);
  PPOA poa;
  poa.init(seqs[0]);
  auto& aligner = tls().affine;
  bool yes = true;
  for (auto& seq : seqs) {
    aligner.alignEnd2End(seqs[0], seq);
    std::string cigar = aligner.getAlignmentCigar();
    poa.insert_alignment(seq, cigar, yes, yes);
  }
  poa.adjust_weights((float)seqs.size() * 0.4, 0.3);
  std::string consensus;
  poa.consensus(consensus);
  std::memcpy(out, consensus.data(), (int)consensus.size() < cap ? consensus.size() : cap);
  return (int)consensus.size();
}

// The reference PPOA (src/anppoa.hpp) from given CIGARs: reads 1..n-1 inserted into the graph of
// read 0 along cigars[coff[i], coff[i+1]), then adjust_weights(c, t) and consensus.
int ref_ppoa_cigars(const char* bases, const uint64_t* off, const uint32_t* len, const uint8_t* span,
                    uint32_t n, const char* cigars, const uint64_t* coff, float c, float t, char* out,
                    int cap) {
  PPOA poa;
  std::string backbone(bases + off[0], len[0]);
  poa.init(backbone);
  for (uint32_t i = 1; i < n; ++i) {
    std::string seq(bases + off[i], len[i]), cigar(cigars + coff[i], coff[i + 1] - coff[i]);
    bool sl = span ? (span[i] & 1) != 0 : true, sr = span ? (span[i] & 2) != 0 : true;
    poa.insert_alignment(seq, cigar, sl, sr);
  }
  poa.adjust_weights(c, t);
  std::string consensus;
  poa.consensus(consensus);
  if (out && cap > 0) std::memcpy(out, consensus.data(), (int)consensus.size() < cap ? consensus.size() : cap);
  return (int)consensus.size();
}

// local_realignment (src/analignments.cpp:11-60) against an indexed FASTA on disk.
// ccoords = pairs (first, second) per read.  Outputs: new sequences concatenated in out_bases
// (lengths in out_len) and new span bits.  Returns total bytes needed.
uint64_t ref_local_realignment(const char* fasta, const char* chr, int start, int end, int flank,
                               double min_sim, const char* bases, const uint64_t* off,
                               const uint32_t* len, const uint8_t* span, const int* ccoords,
                               uint32_t n_reads, char* out_bases, uint64_t cap, uint32_t* out_len,
                               uint8_t* out_span) {
  std::vector<ANREAD> reads;
  build_reads(bases, off, len, span, 0, n_reads, reads);
  for (uint32_t i = 0; i < n_reads; ++i) reads[i].ccoords = {ccoords[2 * i], ccoords[2 * i + 1]};
  FaidxInstance fa;
  fa.init(fasta);
  local_realignment(chr, start, end, flank, min_sim, fa, reads, tls().affine);
  fa.destroy();
  uint64_t used = 0;
  for (uint32_t i = 0; i < n_reads; ++i) {
    out_len[i] = (uint32_t)reads[i].seq.size();
    out_span[i] = (reads[i].is_spanning_l ? 1 : 0) | (reads[i].is_spanning_r ? 2 : 0);
    if (used + reads[i].seq.size() <= cap) std::memcpy(out_bases + used, reads[i].seq.data(), reads[i].seq.size());
    used += reads[i].seq.size();
  }
  return used;
}

// One region through the body of assemble_process (src/assemble.cpp:103-142) with the default
// options of command_assemble.cpp:34-45 (no --haps): partition_valid_reads -> fill_dist_matrix ->
// otter_hclust -> invalid_reassignment -> rapid_consensus.  Outputs the per-read labels and the
// alleles (sequence, acov, scov, tcov, se, ic).  Returns the number of alleles (0 if no spanning
// read), allele sequences concatenated in out_seq (lengths out_seq_len).
int ref_assemble_region(const char* bases, const uint64_t* off, const uint32_t* len,
                        const uint8_t* span, uint32_t n_reads, int max_alleles, int* out_labels,
                        char* out_seq, uint64_t cap, uint32_t* out_seq_len, int* out_cov /*3 per*/,
                        float* out_se, int* out_ic) {
  std::vector<ANREAD> reads;
  build_reads(bases, off, len, span, 0, n_reads, reads);
  for (uint32_t i = 0; i < n_reads; ++i) out_labels[i] = -1;
  bool ignore_haps = true;
  std::vector<int> valid, invalid;
  for (int i = 0; i < (int)reads.size(); ++i)  // partition_valid_reads, src/assemble.cpp:27-37
    (reads[i].is_spanning() ? valid : invalid).emplace_back(i);
  if (valid.empty()) return 0;
  auto& A = tls();
  DistMatrix distmatrix(valid.size());
  if (max_alleles != 1) fill_dist_matrix(ignore_haps, A.edit, reads, valid, distmatrix);
  ClusteringStatus clust;
  otter_hclust(ignore_haps, max_alleles, 0.01, 500, 0.015, 0.01, 0.2, 500, 0.1, valid, distmatrix,
               A.edit, reads, clust);
  std::vector<int> labels(reads.size(), -1);
  for (uint32_t i = 0; i < clust.labels.size(); ++i) labels[valid[i]] = clust.labels[i];
  if (!invalid.empty()) invalid_reassignment(ignore_haps, 0.9, 0.01, clust.fc, reads, A.edit, labels);
  std::vector<ANALLELE> alleles(clust.fc);
  rapid_consensus(ignore_haps, reads, labels, valid, clust.fc, distmatrix, A.affine, alleles);
  uint64_t used = 0;
  for (int l = 0; l < clust.fc; ++l) {
    out_seq_len[l] = (uint32_t)alleles[l].seq.size();
    if (used + alleles[l].seq.size() <= cap) std::memcpy(out_seq + used, alleles[l].seq.data(), alleles[l].seq.size());
    used += alleles[l].seq.size();
    out_cov[3 * l] = alleles[l].acov;
    out_cov[3 * l + 1] = alleles[l].scov;
    out_cov[3 * l + 2] = alleles[l].tcov;
    out_se[l] = alleles[l].se;
  }
  *out_ic = clust.ic;
  for (uint32_t i = 0; i < n_reads; ++i) out_labels[i] = labels[i];
  return clust.fc;
}

// parse_alignment (src/anseqs.cpp:393-418: get_breakpoints :286-391 + the 4-bit decode) on one
// synthetic BAM record: pos, BAM-encoded CIGAR (len << 4 | op), query sequence.  Outputs what
// parse_anreads keeps: msg.successful, the spanning flags, msg.alignment_coords and the extracted
// sequence.  Returns its length (0 when not successful).
int ref_parse_alignment(int pos, const uint32_t* cigar, uint32_t n_cigar, const char* seq, int l_qseq,
                        int start, int end, int* successful, int* span, int* coords, char* out,
                        int cap) {
  bam1_t* b = bam_init1();
  const char qname[4] = "r\0\0";
  b->core.pos = pos;
  b->core.l_qname = 4;
  b->core.n_cigar = n_cigar;
  b->core.l_qseq = l_qseq;
  b->l_data = b->m_data = 4 + 4 * (int)n_cigar + (l_qseq + 1) / 2 + l_qseq;
  b->data = (uint8_t*)calloc(b->m_data + 8, 1);
  std::memcpy(b->data, qname, 4);
  std::memcpy(b->data + 4, cigar, 4 * (size_t)n_cigar);
  uint8_t* t = bam_get_seq(b);
  for (int i = 0; i < l_qseq; ++i)  // sam.c:598
    t[i >> 1] |= seq_nt16_table[(unsigned char)seq[i]] << ((~i & 1) << 2);
  PARSEMSG msg;
  std::string got;
  parse_alignment(start, end, b, msg, got);
  *successful = msg.successful;
  *span = (msg.spanning_l ? 1 : 0) | (msg.spanning_r ? 2 : 0);
  coords[0] = msg.alignment_coords.first;
  coords[1] = msg.alignment_coords.second;
  bam_destroy1(b);
  if (!msg.successful) return 0;
  std::memcpy(out, got.data(), (int)got.size() < cap ? got.size() : cap);
  return (int)got.size();
}

// parse_alignment over a batch of records (8 int64 each: pos, l_qseq, n_cigar, cigar_off, seq_off,
// region, start, end; seq4 already in BAM's 4-bit form) on `threads` host threads, static
// contiguous blocks.  The bam1_t records are built before the clock starts (htslib hands them to
// parse_anreads ready-made).  out_len[i] = length of the extracted sequence, -1 if not successful.
// Returns wall seconds of the parse_alignment calls.
double ref_parse_alignments_batch(const int64_t* recs, uint32_t n, const uint32_t* cigar,
                                  const uint8_t* seq4, int threads, int32_t* out_len) {
  std::vector<bam1_t*> bs(n);
  for (uint32_t i = 0; i < n; ++i) {
    const int64_t* r = recs + 8 * (size_t)i;
    bam1_t* b = bam_init1();
    b->core.pos = (int32_t)r[0];
    b->core.l_qseq = (int32_t)r[1];
    b->core.n_cigar = (uint32_t)r[2];
    b->core.l_qname = 4;
    const size_t nb = ((size_t)r[1] + 1) / 2;
    b->l_data = b->m_data = (int)(4 + 4 * (size_t)r[2] + nb + (size_t)r[1]);
    b->data = (uint8_t*)calloc((size_t)b->m_data + 8, 1);
    b->data[0] = 'r';
    std::memcpy(b->data + 4, cigar + r[3], 4 * (size_t)r[2]);
    std::memcpy(bam_get_seq(b), seq4 + r[4], nb);
    bs[i] = b;
  }
  if (threads < 1) threads = 1;
  const auto t0 = std::chrono::steady_clock::now();
  {
    std::vector<std::thread> pool;
    const uint32_t per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
      const uint32_t a = (uint32_t)t * per, e = a + per < n ? a + per : n;
      if (a >= e) break;
      pool.emplace_back([&, a, e] {
        for (uint32_t i = a; i < e; ++i) {
          const int64_t* r = recs + 8 * (size_t)i;
          PARSEMSG msg;
          std::string got;
          parse_alignment((int)r[6], (int)r[7], bs[i], msg, got);
          out_len[i] = msg.successful ? (int32_t)got.size() : -1;
        }
      });
    }
    for (auto& th : pool) th.join();
  }
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (bam1_t* b : bs) bam_destroy1(b);
  return sec;
}

// KDE::f (src/ankde.cpp:19-23) over a grid built as otter_find_clustering_dist builds it
// (src/otterclust.cpp:25-28): out[i] = kde.f(x_i), x_0 = 0, x_{i+1} = x_i + dinterval while <= 1.
// Returns the number of grid points.
int ref_kde_densities(const double* values, uint64_t n, double bandwidth, double dinterval, double* out,
                      int cap) {
  KDE kde(bandwidth);
  for (uint64_t i = 0; i < n; ++i) kde.values.emplace_back(values[i]);
  int k = 0;
  for (double x = 0.0; x <= 1.0; x += dinterval) {
    const double f = kde.f(x);
    if (k < cap) out[k] = f;
    ++k;
  }
  return k;
}

// otter_hclust (src/otterclust.cpp:118-320) on one distance matrix; read lengths stand in for the
// reads (only seq.size() is used).  Outputs labels (per valid read), ic, fc.
void ref_otter_hclust(const double* dist, uint32_t n, const uint32_t* read_len, int max_alleles,
                      int* labels, int* ic, int* fc) {
  std::vector<ANREAD> reads(n);
  std::vector<int> indeces(n);
  for (uint32_t i = 0; i < n; ++i) {
    reads[i].seq.assign(read_len[i], 'A');
    indeces[i] = (int)i;
  }
  DistMatrix dm(n);
  for (size_t i = 0; i < dm.values.size(); ++i) dm.values[i] = dist[i];
  ClusteringStatus clust;
#ifdef OTTER_B200_IN_TREE  // the same call through the batched, device-assisted restatement
  std::vector<HclustJob> jobs{{&indeces, &dm, &reads, &clust}};
  otter_hclust_batch(true, max_alleles, 0.01, 500, 0.015, 0.01, 0.2, 500, 0.1, tls().edit, jobs, 1);
#else
  otter_hclust(true, max_alleles, 0.01, 500, 0.015, 0.01, 0.2, 500, 0.1, indeces, dm, tls().edit, reads, clust);
#endif
  for (uint32_t i = 0; i < n; ++i) labels[i] = clust.labels[i];
  *ic = clust.ic;
  *fc = clust.fc;
}

// hclust-cpp as otter calls it (src/otterclust.cpp:183-187,226,241): average linkage on a copy of
// the condensed matrix, then both cuts.  merge: 2(n-1) ints, height: n-1, labels_k / labels_h: n.
void ref_hclust_average(const double* dist, int n, int k, double cut, int* merge, double* height,
                        int* labels_k, int* labels_h) {
  std::vector<double> copy(dist, dist + (size_t)n * (n - 1) / 2);
  hclust_fast(n, copy.data(), HCLUST_METHOD_AVERAGE, merge, height);
  cutree_k(n, merge, k, labels_k);
  cutree_cdist(n, merge, height, cut, labels_h);
}

// ---- many regions through the body of assemble_process, timed --------------------------------
extern "C++" {
namespace {
template <class F>
void parallel_blocks(uint32_t n, int threads, F&& f) {  // static contiguous blocks, as parallelize_loop
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  const uint32_t per = (n + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    const uint32_t a = (uint32_t)t * per, b = a + per < n ? a + per : n;
    if (a >= b) break;
    pool.emplace_back([a, b, &f] { for (uint32_t g = a; g < b; ++g) f(g); });
  }
  for (auto& th : pool) th.join();
}

uint64_t fnv(uint64_t h, const void* p, size_t n) {
  const unsigned char* c = static_cast<const unsigned char*>(p);
  for (size_t i = 0; i < n; ++i) { h ^= c[i]; h *= 1099511628211ull; }
  return h;
}
uint64_t allele_digest(const std::vector<ANALLELE>& alleles, int ic) {
  uint64_t h = 1469598103934665603ull;
  h = fnv(h, &ic, sizeof ic);
  for (const auto& a : alleles) {
    h = fnv(h, a.seq.data(), a.seq.size());
    const int cov[3] = {a.acov, a.scov, a.tcov};
    h = fnv(h, cov, sizeof cov);
    const float se = a.se;
    h = fnv(h, &se, sizeof se);
  }
  return h;
}

struct RegionWork {
  std::vector<ANREAD> reads;
  std::vector<int> valid, invalid, labels;
  std::unique_ptr<DistMatrix> dm;
  ClusteringStatus clust;
  std::vector<ANALLELE> alleles;
};
}  // namespace
}  // extern "C++"

// assemble_process's loop body (src/assemble.cpp:103-142, defaults of command_assemble.cpp:34-45,
// no --haps) over n_regions regions on `threads` host threads.  Per region: the labels of its
// reads (out_labels, indexed like the reads), the number of alleles and a digest of the alleles
// (sequence, coverages, se) + initial cluster count.  Returns wall seconds.
//   reference build: one region at a time per thread, static contiguous blocks (parallelize_loop);
//   in-tree build  : the same stages in one wave — every region's distance matrix in ONE batch on
//                    the GPU, clustering per region on the host threads, re-assignment and the
//                    consensus alignments batched on the GPU, partial-order consensus per region
//                    on the host threads (the seam INTEGRATION.md describes).
double ref_assemble_regions(const char* bases, const uint64_t* off, const uint32_t* len,
                            const uint8_t* span, const uint32_t* region_start, uint32_t n_regions,
                            int max_alleles, int threads, int32_t* out_labels,
                            int32_t* out_n_alleles, uint64_t* out_digest) {
  bool ignore_haps = true;
  std::vector<RegionWork> W(n_regions);
  const auto t0 = std::chrono::steady_clock::now();
  auto prepare = [&](uint32_t g) {
    RegionWork& w = W[g];
    build_reads(bases, off, len, span, region_start[g], region_start[g + 1], w.reads);
    for (int i = 0; i < (int)w.reads.size(); ++i)  // partition_valid_reads, src/assemble.cpp:27-37
      (w.reads[i].is_spanning() ? w.valid : w.invalid).emplace_back(i);
    w.labels.assign(w.reads.size(), -1);
    if (!w.valid.empty()) w.dm.reset(new DistMatrix(w.valid.size()));
  };
  auto cluster = [&](uint32_t g, wfa::WFAligner& unused) {
    RegionWork& w = W[g];
    if (w.valid.empty()) return;
    otter_hclust(ignore_haps, max_alleles, 0.01, 500, 0.015, 0.01, 0.2, 500, 0.1, w.valid, *w.dm,
                 unused, w.reads, w.clust);
    for (uint32_t i = 0; i < w.clust.labels.size(); ++i) w.labels[w.valid[i]] = w.clust.labels[i];
    w.alleles.resize(w.clust.fc);
  };
  auto publish = [&](uint32_t g) {
    RegionWork& w = W[g];
    for (uint32_t i = 0; i < w.labels.size(); ++i) out_labels[region_start[g] + i] = w.labels[i];
    out_n_alleles[g] = w.valid.empty() ? 0 : w.clust.fc;
    out_digest[g] = w.valid.empty() ? 0 : allele_digest(w.alleles, w.clust.ic);
  };
#ifndef OTTER_B200_IN_TREE
  parallel_blocks(n_regions, threads, [&](uint32_t g) {
    prepare(g);
    RegionWork& w = W[g];
    if (!w.valid.empty()) {
      auto& A = tls();
      if (max_alleles != 1) fill_dist_matrix(ignore_haps, A.edit, w.reads, w.valid, *w.dm);
      cluster(g, A.edit);
      if (!w.invalid.empty())
        invalid_reassignment(ignore_haps, 0.9, 0.01, w.clust.fc, w.reads, A.edit, w.labels);
      rapid_consensus(ignore_haps, w.reads, w.labels, w.valid, w.clust.fc, *w.dm, A.affine, w.alleles);
    }
    publish(g);
  });
#else
  // one wave at a time through device / host stages, overlapped (otter_b200/host/wave_b200.hpp)
  auto& A = tls();
  const bool stages = getenv("OTREF_STAGES") != nullptr;
  std::mutex log_mu;
  auto timed = [&](const char* what, auto&& body) {
    const auto a0 = std::chrono::steady_clock::now();
    body();
    if (stages) {
      std::lock_guard<std::mutex> lk(log_mu);
      fprintf(stderr, "  [wave] %-28s %8.2f ms\n", what,
              std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a0).count());
    }
  };
  // host-stage pool size: with two issuers the two host stages run at once next to two issuing
  // threads, so each gets a share of the cores instead of all of them (OTREF_HOST_THREADS overrides)
  int issuers = 1;
  if (const char* e = getenv("OTREF_ISSUERS")) issuers = atoi(e);
  int pool_threads = issuers >= 2 ? std::max(1, (threads - 2) / 2) : threads;
  if (const char* e = getenv("OTREF_HOST_THREADS")) pool_threads = std::max(1, atoi(e));
  auto blocks = [&](uint32_t a, uint32_t b, auto&& f) {
    parallel_blocks(b - a, pool_threads, [&](uint32_t i) { f(a + i); });
  };
  std::vector<ConsensusJob> cjobs(n_regions);  // one slot per region (unused when no valid read)
  std::vector<std::vector<double>> dens(n_regions);
  std::vector<double> bandw(n_regions, 0.01);
  const bool host_kde = getenv("OTREF_HOST_KDE") != nullptr;  // A/B: the reference's own otter_hclust
  const bool device_ppoa = getenv("OTREF_DEVICE_PPOA") != nullptr;  // A/B: the graphs on the GPU too
  // the second device stage is issued from its own thread (wave_b200.hpp): its own edit aligner,
  // kept for the life of the process like the thread's other aligners (a context is not free)
  static wfa::WFAlignerEdit edit2(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed);
  otb200::WaveStages st;
  st.device1 = [&](uint32_t a, uint32_t b) {
    timed("build reads", [&] { blocks(a, b, prepare); });
    timed("distance matrices (GPU)", [&] {
      if (max_alleles == 1) return;
      std::vector<RegionDistJob> jobs;
      for (uint32_t g = a; g < b; ++g)
        if (!W[g].valid.empty()) jobs.push_back({&W[g].reads, &W[g].valid, W[g].dm.get()});
      fill_dist_matrices(ignore_haps, A.edit, jobs);
    });
    if (!host_kde) timed("density estimates (GPU)", [&] {
      std::vector<HclustJob> jobs;
      std::vector<uint32_t> owner;
      for (uint32_t g = a; g < b; ++g)
        if (!W[g].valid.empty()) {
          jobs.push_back({&W[g].valid, W[g].dm.get(), &W[g].reads, &W[g].clust});
          owner.push_back(g);
        }
      std::vector<std::vector<double>> d;
      std::vector<double> h;
      hclust_densities(max_alleles, 0.01, 500, 0.015, A.edit, jobs, d, h);
      for (size_t i = 0; i < owner.size(); ++i) {
        dens[owner[i]] = std::move(d[i]);
        bandw[owner[i]] = h[i];
      }
    });
  };
  st.host1 = [&](uint32_t a, uint32_t b) {
    timed("otter_hclust (host)", [&] {
      blocks(a, b, [&](uint32_t g) {
        if (host_kde) return cluster(g, A.edit);
        RegionWork& w = W[g];
        if (w.valid.empty()) return;
        HclustJob job{&w.valid, w.dm.get(), &w.reads, &w.clust};
        hclust_finish(max_alleles, 0.01, 0.2, 500, 0.1, job, dens[g], bandw[g]);
        for (uint32_t i = 0; i < w.clust.labels.size(); ++i) w.labels[w.valid[i]] = w.clust.labels[i];
        w.alleles.resize(w.clust.fc);
      });
    });
  };
  st.device2 = [&](uint32_t a, uint32_t b) {
    timed("re-assignment + consensus alignments (GPU)", [&] {
      std::vector<RegionReassignJob> rjobs;
      for (uint32_t g = a; g < b; ++g)
        if (!W[g].valid.empty() && !W[g].invalid.empty())
          rjobs.push_back({&W[g].reads, &W[g].labels, W[g].clust.fc});
      if (!rjobs.empty()) invalid_reassignments(ignore_haps, 0.9, 0.01, edit2, rjobs);
      std::vector<ConsensusJob> batch;
      std::vector<uint32_t> owner;
      for (uint32_t g = a; g < b; ++g) {
        RegionWork& w = W[g];
        if (w.valid.empty()) continue;
        ConsensusJob j;
        j.reads = &w.reads;
        j.labels = &w.labels;
        j.valid_indeces = &w.valid;
        j.total_alleles = w.clust.fc;
        j.valid_distmatrix = w.dm.get();
        batch.push_back(std::move(j));
        owner.push_back(g);
      }
      rapid_consensus_alignments(A.affine, batch);
      for (size_t i = 0; i < batch.size(); ++i) cjobs[owner[i]] = std::move(batch[i]);
    });
    if (device_ppoa) timed("partial-order consensus (GPU)", [&] {
      std::vector<ConsensusJob*> jp;
      std::vector<std::vector<ANALLELE>*> ap;
      for (uint32_t g = a; g < b; ++g)
        if (!W[g].valid.empty()) {
          jp.push_back(&cjobs[g]);
          ap.push_back(&W[g].alleles);
        }
      rapid_consensus_finish_device(ignore_haps, A.affine, jp, ap);
    });
  };
  st.host2 = [&](uint32_t a, uint32_t b) {
    timed("partial-order consensus (host)", [&] {
      blocks(a, b, [&](uint32_t g) {
        if (!W[g].valid.empty() && !device_ppoa) rapid_consensus_finish(ignore_haps, cjobs[g], W[g].alleles);
        publish(g);
      });
    });
  };
  // wave size: the consensus alignments need a few hundred regions per launch before their
  // persistent warps stop waiting for the longest alignment (profiles/README.md r1n: 34.5 ms per
  // 100 regions in waves of 100, 18 ms per 100 in waves of 400), while four waves are enough to
  // overlap the host stages
  uint32_t wave = std::min(400u, std::max(100u, n_regions / 4));
  if (const char* e = getenv("OTREF_WAVE")) wave = (uint32_t)atoi(e);
  otb200::run_waves(n_regions, wave, st, issuers);
#endif
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// aligner2 over a list of (pattern, text, pb, pe, tb, te) requests (pb < 0: end-to-end), static
// contiguous blocks over `threads` workers, one WFAlignerGapAffine(4,6,2,Alignment,MemoryMed)
// each.  pairs = 6 int32 per request (pattern read, text read, pb, pe, tb, te).  Returns wall
// seconds; cigar_len[i] = length of request i's CIGAR (the strings themselves are dropped).
double ref_cigars_batch_impl(bool edit, const char* bases, const uint64_t* off, const uint32_t* len,
                             const int32_t* pairs, uint64_t n_pairs, int threads, int32_t* cigar_len);
double ref_affine_cigars_batch(const char* bases, const uint64_t* off, const uint32_t* len,
                               const int32_t* pairs, uint64_t n_pairs, int threads,
                               int32_t* cigar_len) {
  return ref_cigars_batch_impl(false, bases, off, len, pairs, n_pairs, threads, cigar_len);
}
// The same with compare's aligner (WFAlignerEdit(Alignment, MemoryMed), src/compare.cpp:95).
double ref_edit_cigars_batch(const char* bases, const uint64_t* off, const uint32_t* len,
                             const int32_t* pairs, uint64_t n_pairs, int threads,
                             int32_t* cigar_len) {
  return ref_cigars_batch_impl(true, bases, off, len, pairs, n_pairs, threads, cigar_len);
}
double ref_cigars_batch_impl(bool edit, const char* bases, const uint64_t* off, const uint32_t* len,
                             const int32_t* pairs, uint64_t n_pairs, int threads, int32_t* cigar_len) {
  if (threads < 1) threads = 1;
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    pool.emplace_back([&, t]() {
      wfa::WFAlignerGapAffine aff(4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed);
      wfa::WFAlignerEdit ed(wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed);
      wfa::WFAligner& a = edit ? static_cast<wfa::WFAligner&>(ed) : static_cast<wfa::WFAligner&>(aff);
      const uint64_t lo = n_pairs * t / threads, hi = n_pairs * (t + 1) / threads;
      for (uint64_t i = lo; i < hi; ++i) {
        const int32_t* q = pairs + 6 * i;
        const char* p = bases + off[q[0]];
        const char* x = bases + off[q[1]];
        if (q[2] < 0) a.alignEnd2End(p, (int)len[q[0]], x, (int)len[q[1]]);
        else a.alignEndsFree(p, (int)len[q[0]], q[2], q[3], x, (int)len[q[1]], q[4], q[5]);
        cigar_len[i] = (int32_t)a.getAlignmentCigar().size();
      }
    });
  }
  for (auto& th : pool) th.join();
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
}
