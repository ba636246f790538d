// otter_b200/host/analignments_b200.hpp — C++17 host side of the hot path: the reference's own
// call surface (src/analignments.hpp:37-60), same names / argument meaning / error behaviour,
// routed to the GPU through the C ABI, plus batched variants that put many regions in one launch.
#pragma once
#include <string>
#include <vector>

#include "bindings/cpp/WFAligner.hpp"  // the shim in otter_b200/host/shim (or WFA2's, in-tree)
#include "otter_types.hpp"

// ---- reference signatures (drop-in for src/analignments.cpp:62-177) -----------------------------
double align_anreads(wfa::WFAligner& aligner, ANREAD& read_x, ANREAD& read_y);

double get_dist_anreads(const bool& ignore_haps, wfa::WFAligner& aligner, ANREAD& read_x,
                        ANREAD& read_y);

void fill_dist_matrix(const bool& ignore_haps, wfa::WFAligner& aligner, std::vector<ANREAD>& reads,
                      std::vector<int>& indeces, DistMatrix& distmatrix);

void invalid_reassignment(const bool& ignore_haps, const double& min_sim, const double& max_error,
                          const int& total_alleles, std::vector<ANREAD>& reads,
                          wfa::WFAligner& aligner, std::vector<int>& labels);

// ---- batched entry points (the seam INTEGRATION.md describes for assemble_process) --------------
struct RegionDistJob {
  std::vector<ANREAD>* reads;
  std::vector<int>* indeces;
  DistMatrix* distmatrix;
};
// fill_dist_matrix for every job with ONE upload + launch + download.  Results per region are
// identical to calling fill_dist_matrix region by region.
void fill_dist_matrices(const bool& ignore_haps, wfa::WFAligner& aligner,
                        std::vector<RegionDistJob>& jobs);

struct RegionReassignJob {
  std::vector<ANREAD>* reads;
  std::vector<int>* labels;
  int total_alleles;
};
void invalid_reassignments(const bool& ignore_haps, const double& min_sim,
                           const double& max_error, wfa::WFAligner& aligner,
                           std::vector<RegionReassignJob>& jobs);
