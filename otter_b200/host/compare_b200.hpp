// otter_b200/host/compare_b200.hpp — the alignment step of `otter compare`
// (reference src/compare.cpp:17-24 DistCompare, :50-66 get_distances) on the GPU.
// allele x allele edit distance with CIGAR length.
#pragma once
#include <string>
#include <vector>

#include "bindings/cpp/WFAligner.hpp"
#include "otter_types.hpp"

class DistCompare {  // reference src/compare.cpp:17-24
 public:
  int i;
  int j;
  double edit;
  double ops;
  DistCompare(int _i, int _j, double _e, double _o) : i(_i), j(_j), edit(_e), ops(_o) {}
};

// Reference signature (src/compare.cpp:50): every (subject, query) pair, subject-major, appended
// to `distances`.  The special cases ("N", "NDNNN", equal strings) are decided on the host as the
// reference does; all remaining pairs of the call go to the device in ONE launch.
// `aligner` must be a WFAlignerEdit(Alignment, ...) as compare builds it (src/compare.cpp:95).
void get_distances(wfa::WFAligner& aligner, std::vector<ANALLELE>& subjs,
                   std::vector<ANALLELE>& querys, std::vector<DistCompare>& distances);

// Many regions in one launch (the seam for compare's parallelize_loop, src/compare.cpp:93-114).
struct CompareJob {
  std::vector<ANALLELE>* subjs;
  std::vector<ANALLELE>* querys;
  std::vector<DistCompare>* distances;  // out, appended
};
void get_distances_batch(wfa::WFAligner& aligner, std::vector<CompareJob>& jobs);
