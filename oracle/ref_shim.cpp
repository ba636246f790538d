// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY.
//
// Thin extern "C" shim over the UNMODIFIED reference (otter + vendored WFA2-lib v2.2), compiled
// against the reference headers where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libotter_ref.so.  Nothing here re-implements an algorithm: every entry point
// constructs the reference's own objects (wfa::WFAlignerEdit / wfa::WFAlignerGapAffine, ANREAD,
// DistMatrix) exactly the way src/assemble.cpp:49-50 does and calls the reference function.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (otter_b200/) never does.
#include "bindings/cpp/WFAligner.hpp"
#include "analignments.hpp"
#include "andistmat.hpp"
#include "anseqs.hpp"
#include "otterclust.hpp"

#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

// Defined (external linkage, not in the header) at src/analignments.cpp:62.
double align_anreads(wfa::WFAligner& aligner, ANREAD& read_x, ANREAD& read_y);

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
}
