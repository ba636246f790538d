// otter_b200/host/host_capi.cpp — flat extern "C" entry points over the C++ host layer so the
// Python parity tests (ctypes) can drive the reference-shaped C++ functions
// (fill_dist_matrix, invalid_reassignment, align_anreads) exactly as otter would.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>

#include "analignments_b200.hpp"
#include "anseqs_b200.hpp"
#include "compare_b200.hpp"
#include "linkage_b200.hpp"
#include "ppoa_b200.hpp"

namespace {
std::vector<ANREAD> build_reads(const char* bases, const uint64_t* off, const uint32_t* len,
                                const uint8_t* span, uint32_t first, uint32_t last) {
  std::vector<ANREAD> reads;
  reads.reserve(last - first);
  for (uint32_t i = first; i < last; ++i) {
    ANREAD r("r" + std::to_string(i), std::string(bases + off[i], len[i]));
    r.is_spanning_l = span ? (span[i] & 1) != 0 : true;
    r.is_spanning_r = span ? (span[i] & 2) != 0 : true;
    reads.emplace_back(std::move(r));
  }
  return reads;
}
wfa::WFAlignerEdit& aligner() {
  static thread_local std::unique_ptr<wfa::WFAlignerEdit> a;
  if (!a) a.reset(new wfa::WFAlignerEdit(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed));
  return *a;
}
}  // namespace

extern "C" {

int othost_have_device() { return aligner().context() != nullptr; }

double othost_align_anreads(const char* x, uint32_t xlen, int xspan, const char* y, uint32_t ylen,
                            int yspan) {
  ANREAD rx("x", std::string(x, xlen)), ry("y", std::string(y, ylen));
  rx.is_spanning_l = xspan & 1; rx.is_spanning_r = xspan & 2;
  ry.is_spanning_l = yspan & 1; ry.is_spanning_r = yspan & 2;
  return align_anreads(aligner(), rx, ry);
}

// One region through the reference-shaped fill_dist_matrix.
void othost_fill_dist_matrix(const char* bases, const uint64_t* off, const uint32_t* len,
                             const uint8_t* span, uint32_t n_reads, const int* indeces,
                             uint32_t n_idx, double* out) {
  auto reads = build_reads(bases, off, len, span, 0, n_reads);
  std::vector<int> idx(indeces, indeces + n_idx);
  DistMatrix dm(n_idx);
  fill_dist_matrix(true, aligner(), reads, idx, dm);
  if (!dm.values.empty()) std::memcpy(out, dm.values.data(), dm.values.size() * sizeof(double));
}

// Many regions (all reads valid) through fill_dist_matrices: one launch.
void othost_fill_dist_matrices(const char* bases, const uint64_t* off, const uint32_t* len,
                               const uint8_t* span, const uint32_t* region_start,
                               uint32_t n_regions, double* out, const uint64_t* out_off) {
  std::vector<std::vector<ANREAD>> reads(n_regions);
  std::vector<std::vector<int>> idx(n_regions);
  std::vector<DistMatrix> dms;
  dms.reserve(n_regions);
  std::vector<RegionDistJob> jobs;
  for (uint32_t g = 0; g < n_regions; ++g) {
    reads[g] = build_reads(bases, off, len, span, region_start[g], region_start[g + 1]);
    idx[g].resize(reads[g].size());
    for (size_t i = 0; i < idx[g].size(); ++i) idx[g][i] = (int)i;
    dms.emplace_back((uint32_t)reads[g].size());
  }
  for (uint32_t g = 0; g < n_regions; ++g) jobs.push_back({&reads[g], &idx[g], &dms[g]});
  fill_dist_matrices(true, aligner(), jobs);
  for (uint32_t g = 0; g < n_regions; ++g)
    if (!dms[g].values.empty())
      std::memcpy(out + out_off[g], dms[g].values.data(), dms[g].values.size() * sizeof(double));
}

// The same over n_aligners aligners from this one process: aligner a is bound to device
// devices[a] (several may name the same device).  part_out (optional) receives the aligner of
// each region.
void othost_fill_dist_matrices_multi(const char* bases, const uint64_t* off, const uint32_t* len,
                                     const uint32_t* region_start, uint32_t n_regions,
                                     const int* devices, int n_aligners, double* out,
                                     const uint64_t* out_off, int* part_out) {
  std::vector<std::vector<ANREAD>> reads(n_regions);
  std::vector<std::vector<int>> idx(n_regions);
  std::vector<DistMatrix> dms;
  dms.reserve(n_regions);
  std::vector<RegionDistJob> jobs;
  for (uint32_t g = 0; g < n_regions; ++g) {
    reads[g] = build_reads(bases, off, len, nullptr, region_start[g], region_start[g + 1]);
    idx[g].resize(reads[g].size());
    for (size_t i = 0; i < idx[g].size(); ++i) idx[g][i] = (int)i;
    dms.emplace_back((uint32_t)reads[g].size());
  }
  for (uint32_t g = 0; g < n_regions; ++g) jobs.push_back({&reads[g], &idx[g], &dms[g]});
  std::vector<std::unique_ptr<wfa::WFAlignerEdit>> own;
  std::vector<wfa::WFAligner*> aligners;
  for (int a = 0; a < n_aligners; ++a) {
    wfa::WFAligner::setDefaultDevice(devices[a]);
    own.emplace_back(new wfa::WFAlignerEdit(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed));
    aligners.push_back(own.back().get());
  }
  wfa::WFAligner::setDefaultDevice(0);
  if (part_out) {
    const std::vector<int> part = partition_regions_by_work(jobs, n_aligners);
    for (uint32_t g = 0; g < n_regions; ++g) part_out[g] = part[g];
  }
  fill_dist_matrices_multi(true, aligners, jobs);
  for (uint32_t g = 0; g < n_regions; ++g)
    if (!dms[g].values.empty())
      std::memcpy(out + out_off[g], dms[g].values.data(), dms[g].values.size() * sizeof(double));
}

// Timed runs of the seam otter calls: fill_dist_matrices (n_devices == 1) or
// fill_dist_matrices_multi over `devices`, from std::string-shaped reads that exist before the
// clock starts (as otter's ANREADs do), `warmup` untimed calls first.  times_out holds 6 doubles
// per timed call: wall seconds of the call, then max over the devices of {whole share, device
// call, wait for the staging copy inside it, score -> dissimilarity conversion}, then the number
// of chunks of device 0.  `out` receives the last call's matrices.
void othost_bench_fill(const char* bases, const uint64_t* off, const uint32_t* len,
                       const uint32_t* region_start, uint32_t n_regions, const int* devices,
                       int n_devices, int warmup, int reps, double* out, const uint64_t* out_off,
                       double* times_out) {
  std::vector<std::vector<ANREAD>> reads(n_regions);
  std::vector<std::vector<int>> idx(n_regions);
  std::vector<DistMatrix> dms;
  dms.reserve(n_regions);
  std::vector<RegionDistJob> jobs;
  for (uint32_t g = 0; g < n_regions; ++g) {
    reads[g] = build_reads(bases, off, len, nullptr, region_start[g], region_start[g + 1]);
    idx[g].resize(reads[g].size());
    for (size_t i = 0; i < idx[g].size(); ++i) idx[g][i] = (int)i;
    dms.emplace_back((uint32_t)reads[g].size());
  }
  for (uint32_t g = 0; g < n_regions; ++g) jobs.push_back({&reads[g], &idx[g], &dms[g]});
  std::vector<std::unique_ptr<wfa::WFAlignerEdit>> own;
  std::vector<wfa::WFAligner*> aligners;
  for (int a = 0; a < n_devices; ++a) {
    wfa::WFAligner::setDefaultDevice(devices[a]);
    own.emplace_back(new wfa::WFAlignerEdit(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed));
    aligners.push_back(own.back().get());
  }
  wfa::WFAligner::setDefaultDevice(0);
  std::vector<FillTimes> times;
  for (int r = 0; r < warmup + reps; ++r) {
    const auto t0 = std::chrono::steady_clock::now();
    fill_dist_matrices_multi(true, aligners, jobs, &times);
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (r < warmup) continue;
    double* t = times_out + 6 * (size_t)(r - warmup);
    t[0] = wall; t[1] = t[2] = t[3] = t[4] = 0.0; t[5] = times[0].chunks;
    for (const FillTimes& f : times) {
      t[1] = std::max(t[1], f.total_s); t[2] = std::max(t[2], f.device_call_s);
      t[3] = std::max(t[3], f.stage_wait_s); t[4] = std::max(t[4], f.unpack_s);
    }
  }
  for (uint32_t g = 0; g < n_regions; ++g)
    if (!dms[g].values.empty())
      std::memcpy(out + out_off[g], dms[g].values.data(), dms[g].values.size() * sizeof(double));
}

// rapid_consensus_alignments for one region (reads with labels[i] >= 0 are clustered; valid =
// spanning reads, their distance matrix filled first): the member CIGARs of every label, in plan
// order, each terminated by '\n', written to `out` (truncated to cap).  Returns the bytes needed.
uint64_t othost_consensus_cigars(const char* bases, const uint64_t* off, const uint32_t* len,
                                 const uint8_t* span, uint32_t n_reads, const int* labels_in,
                                 int total_alleles, char* out, uint64_t cap) {
  auto reads = build_reads(bases, off, len, span, 0, n_reads);
  std::vector<int> labels(labels_in, labels_in + n_reads), valid;
  for (int i = 0; i < (int)n_reads; ++i)
    if (reads[i].is_spanning()) valid.push_back(i);
  DistMatrix dm((uint32_t)valid.size());
  fill_dist_matrix(true, aligner(), reads, valid, dm);
  static thread_local std::unique_ptr<wfa::WFAlignerGapAffine> a2;
  if (!a2) a2.reset(new wfa::WFAlignerGapAffine(4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed));
  std::vector<ConsensusJob> jobs(1);
  jobs[0].reads = &reads;
  jobs[0].labels = &labels;
  jobs[0].valid_indeces = &valid;
  jobs[0].total_alleles = total_alleles;
  jobs[0].valid_distmatrix = &dm;
  rapid_consensus_alignments(*a2, jobs);
  std::string all;
  for (const ConsensusPlan& P : jobs[0].plans)
    for (const std::string& c : P.cigars) { all += c; all += '\n'; }
  if (out && cap) std::memcpy(out, all.data(), std::min<uint64_t>(cap, all.size()));
  return all.size();
}

// The partition alone (no device needed): part_out[g] = share of region g, loads_out[p] = the
// estimated work of share p.
void othost_partition_regions(const uint32_t* len, const uint32_t* region_start, uint32_t n_regions,
                              int n_parts, int* part_out, double* loads_out) {
  std::vector<std::vector<ANREAD>> reads(n_regions);
  std::vector<std::vector<int>> idx(n_regions);
  std::vector<RegionDistJob> jobs;
  for (uint32_t g = 0; g < n_regions; ++g) {
    for (uint32_t r = region_start[g]; r < region_start[g + 1]; ++r) {
      reads[g].emplace_back("r", std::string(len[r], 'A'));
      idx[g].push_back((int)(r - region_start[g]));
    }
  }
  for (uint32_t g = 0; g < n_regions; ++g) jobs.push_back({&reads[g], &idx[g], nullptr});
  std::vector<double> loads;
  const std::vector<int> part = partition_regions_by_work(jobs, n_parts, &loads);
  for (uint32_t g = 0; g < n_regions; ++g) part_out[g] = part[g];
  for (int p = 0; p < n_parts && loads_out; ++p) loads_out[p] = loads[p];
}

void othost_invalid_reassignment(const char* bases, const uint64_t* off, const uint32_t* len,
                                 const uint8_t* span, uint32_t n_reads, double min_sim,
                                 double max_error, int total_alleles, int* labels) {
  auto reads = build_reads(bases, off, len, span, 0, n_reads);
  std::vector<int> lab(labels, labels + n_reads);
  invalid_reassignment(true, min_sim, max_error, total_alleles, reads, aligner(), lab);
  std::memcpy(labels, lab.data(), n_reads * sizeof(int));
}

// aligner2 facade (WFAlignerGapAffine(4,6,2, Alignment, MemoryMed)), one pair.
int othost_affine_cigar(const char* p, int plen, const char* t, int tlen, int endsfree, int pb,
                        int pe, int tb, int te, int* score, int* status, char* cigar, int cap) {
  static thread_local std::unique_ptr<wfa::WFAlignerGapAffine> a;
  if (!a) a.reset(new wfa::WFAlignerGapAffine(4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed));
  const int st = endsfree ? a->alignEndsFree(p, plen, pb, pe, t, tlen, tb, te)
                          : a->alignEnd2End(p, plen, t, tlen);
  if (status) *status = st;
  if (score) *score = a->getAlignmentScore();
  const std::string c = a->getAlignmentCigar();
  if (cigar && cap > 0) std::memcpy(cigar, c.data(), (int)c.size() < cap ? c.size() : cap);
  return (int)c.size();
}

// local_realignment with pre-fetched flanks (local_realignments, one job).  ccoords = (first,
// second) per read.  New sequences are concatenated into out_bases; returns total bytes needed.
uint64_t othost_local_realignment(int flank, double min_sim, const char* ref_left, uint32_t ll,
                                  const char* ref_right, uint32_t rl, const char* bases,
                                  const uint64_t* off, const uint32_t* len, const uint8_t* span,
                                  const int* ccoords, uint32_t n_reads, char* out_bases,
                                  uint64_t cap, uint32_t* out_len, uint8_t* out_span) {
  static thread_local std::unique_ptr<wfa::WFAlignerGapAffine> a;
  if (!a) a.reset(new wfa::WFAlignerGapAffine(4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed));
  auto reads = build_reads(bases, off, len, span, 0, n_reads);
  for (uint32_t i = 0; i < n_reads; ++i) reads[i].ccoords = {ccoords[2 * i], ccoords[2 * i + 1]};
  const std::string L(ref_left, ll), R(ref_right, rl);
  std::vector<RealignJob> one{{flank, min_sim, &L, &R, &reads}};
  local_realignments(*a, one);
  uint64_t used = 0;
  for (uint32_t i = 0; i < n_reads; ++i) {
    out_len[i] = (uint32_t)reads[i].seq.size();
    out_span[i] = (reads[i].is_spanning_l ? 1 : 0) | (reads[i].is_spanning_r ? 2 : 0);
    if (used + reads[i].seq.size() <= cap) std::memcpy(out_bases + used, reads[i].seq.data(), reads[i].seq.size());
    used += reads[i].seq.size();
  }
  return used;
}

// `compare`'s aligner facade (WFAlignerEdit(Alignment, MemoryMed), src/compare.cpp:95), one pair.
int othost_edit_cigar(const char* p, int plen, const char* t, int tlen, int endsfree, int pb, int pe,
                      int tb, int te, int* score, int* status, char* cigar, int cap) {
  static thread_local std::unique_ptr<wfa::WFAlignerEdit> a;
  if (!a) a.reset(new wfa::WFAlignerEdit(wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed));
  const int st = endsfree ? a->alignEndsFree(p, plen, pb, pe, t, tlen, tb, te)
                          : a->alignEnd2End(p, plen, t, tlen);
  if (status) *status = st;
  if (score) *score = a->getAlignmentScore();
  const std::string c = a->getAlignmentCigar();
  if (cigar && cap > 0) std::memcpy(cigar, c.data(), (int)c.size() < cap ? c.size() : cap);
  return (int)c.size();
}

// get_distances (src/compare.cpp:50-66): reads [0,n_subj) are the subjects, the rest the queries.
// out = n_subj * n_query rows of {i, j, edit, ops} as doubles, subject-major.
void othost_get_distances(const char* bases, const uint64_t* off, const uint32_t* len,
                          uint32_t n_subj, uint32_t n_query, double* out) {
  static thread_local std::unique_ptr<wfa::WFAlignerEdit> a;
  if (!a) a.reset(new wfa::WFAlignerEdit(wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed));
  std::vector<ANALLELE> subjs, querys;
  for (uint32_t i = 0; i < n_subj; ++i) subjs.emplace_back(std::string(bases + off[i], len[i]));
  for (uint32_t i = n_subj; i < n_subj + n_query; ++i) querys.emplace_back(std::string(bases + off[i], len[i]));
  std::vector<DistCompare> d;
  get_distances(*a, subjs, querys, d);
  for (size_t r = 0; r < d.size(); ++r) {
    out[4 * r] = d[r].i; out[4 * r + 1] = d[r].j; out[4 * r + 2] = d[r].edit; out[4 * r + 3] = d[r].ops;
  }
}

// Partial-order consensus (ppoa_b200.hpp; reference src/anppoa.hpp) from given CIGARs: reads
// 1..n-1 threaded through the graph of read 0 (the backbone) along cigars[i] (CIGAR i occupies
// cigars[coff[i], coff[i+1]); entry 0 is unused), span bits as elsewhere, then
// adjust_weights(c, t) and consensus.  No device involved.  Returns the consensus length.
int othost_ppoa_cigars(const char* bases, const uint64_t* off, const uint32_t* len,
                       const uint8_t* span, uint32_t n, const char* cigars, const uint64_t* coff,
                       float c, float t, char* out, int cap) {
  otb200::Ppoa poa;
  poa.init(std::string(bases + off[0], len[0]));
  for (uint32_t i = 1; i < n; ++i)
    poa.insert_alignment(std::string(bases + off[i], len[i]),
                         std::string(cigars + coff[i], coff[i + 1] - coff[i]),
                         span ? (span[i] & 1) != 0 : true, span ? (span[i] & 2) != 0 : true);
  poa.adjust_weights(c, t);
  std::string consensus;
  poa.consensus(consensus);
  if (out && cap > 0) std::memcpy(out, consensus.data(), (int)consensus.size() < cap ? consensus.size() : cap);
  return (int)consensus.size();
}

// ReadExtractor (anseqs_b200.hpp) over flat arrays: record i = (pos, l_qseq, n_cigar, cigar words
// at cigar_off, 4-bit bases at seq_off, region id, start, end), 8 int64 per record.  Output per
// kept read (in region order, then input order): region, span bits, ccoords, length; sequences
// concatenated in out_bases.  Returns the number of reads kept.
uint32_t othost_extract_reads(const int64_t* recs, uint32_t n, const uint32_t* cigar, const uint8_t* seq4,
                              int omitnonspanning, uint32_t n_regions, int32_t* out_meta /*5 per read*/,
                              char* out_bases, uint64_t cap) {
  ReadExtractor ex;
  for (uint32_t i = 0; i < n; ++i) {
    const int64_t* r = recs + 8 * (size_t)i;
    BamRecordView v{(int32_t)r[0], (int32_t)r[1], (uint32_t)r[2], cigar + r[3], seq4 + r[4], "r"};
    ex.add((uint32_t)r[5], (int)r[6], (int)r[7], v);
  }
  std::vector<std::vector<ANREAD>> blocks(n_regions);
  ex.extract(aligner(), omitnonspanning != 0, blocks);
  uint32_t kept = 0;
  uint64_t used = 0;
  for (uint32_t g = 0; g < blocks.size(); ++g)
    for (const ANREAD& a : blocks[g]) {
      int32_t* m = out_meta + 5 * (size_t)kept;
      m[0] = (int32_t)g;
      m[1] = (a.is_spanning_l ? 1 : 0) | (a.is_spanning_r ? 2 : 0);
      m[2] = a.ccoords.first;
      m[3] = a.ccoords.second;
      m[4] = (int32_t)a.seq.size();
      if (used + a.seq.size() <= cap) std::memcpy(out_bases + used, a.seq.data(), a.seq.size());
      used += a.seq.size();
      ++kept;
    }
  return kept;
}

// linkage_b200.hpp: average linkage on a copy of the condensed matrix + both cuts (host only).
void othost_hclust_average(const double* dist, int n, int k, double cut, int* merge, double* height,
                           int* labels_k, int* labels_h) {
  std::vector<double> copy(dist, dist + (size_t)n * (n - 1) / 2);
  otb200::average_linkage(n, copy.data(), merge, height);
  otb200::cut_tree_k(n, merge, k, labels_k);
  otb200::cut_tree_height(n, merge, height, cut, labels_h);
}
}
