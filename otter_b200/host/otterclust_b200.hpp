// otter_b200/host/otterclust_b200.hpp — otter_hclust (reference src/otterclust.cpp:118-320) for a
// wave of regions with the kernel density estimates computed on the device.
//
// otter_hclust spends its time in KDE::f: 401 grid points x n(n-1)/2 distances x one exp per
// region.  Here the densities of every region of the wave come from ONE launch
// (otgpu_kde_densities, bit-identical to the host computation: csrc/kde.cuh), and the rest —
// normalisation, maxima / minima, the decision bound, average-linkage clustering and the tree
// cuts (linkage_b200.hpp, bit-identical to hclust-cpp's hclust_fast / cutree_cdist / cutree_k),
// the coverage-based relabelling — runs per region on the host in the reference's order of
// operations.  ClusteringStatus is otter's own (src/otterclust.hpp:7-14) in the in-tree build and
// its mirror in otter_types.hpp stand-alone.
#pragma once
#include <vector>

#include "bindings/cpp/WFAligner.hpp"
#include "otter_types.hpp"
#ifdef OTTER_B200_IN_TREE
#include "otterclust.hpp"
#endif

struct HclustJob {
  const std::vector<int>* indeces;   // valid (spanning) reads
  DistMatrix* distmatrix;
  std::vector<ANREAD>* reads;
  ClusteringStatus* clustering;      // out
};

// otter_hclust for every job, same parameters and results as the reference function called job
// by job.  Phase 1 (calling thread): one device launch for all density estimates; phase 2: the
// jobs are finished on `threads` host threads.
void otter_hclust_batch(const bool& ignore_haps, const int& max_alleles, const double& bandwidth_short,
                        const int& bandwidth_length, const double& bandwidth_long,
                        const double& max_tolerable_diff, const double& min_cov_fraction,
                        const int& min_cov_fraction2_l, const double& min_cov_fraction2_f,
                        wfa::WFAligner& aligner, std::vector<HclustJob>& jobs, int threads);

// The two halves, for callers that overlap them with other work (wave_b200.hpp):
// densities[j] = KDE::f on the reference's grid for job j (empty when the job needs none).
void hclust_densities(const int& max_alleles, const double& bandwidth_short, const int& bandwidth_length,
                      const double& bandwidth_long, wfa::WFAligner& aligner,
                      const std::vector<HclustJob>& jobs, std::vector<std::vector<double>>& densities,
                      std::vector<double>& bandwidths);
void hclust_finish(const int& max_alleles, const double& max_tolerable_diff, const double& min_cov_fraction,
                   const int& min_cov_fraction2_l, const double& min_cov_fraction2_f, HclustJob& job,
                   std::vector<double>& densities, double bandwidth);
