// otter_b200/host/compare_b200.cpp — get_distances of `otter compare` (reference
// src/compare.cpp:50-66): plan on the host, ONE batched otgpu_edit_cigars launch, fill in.
#include "compare_b200.hpp"

#include "analignments_b200.hpp"

void get_distances_batch(wfa::WFAligner& aligner, std::vector<CompareJob>& jobs) {
  struct Slot { size_t job, at; };
  std::vector<CigarJob> work;
  std::vector<Slot> slots;
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& subjs = *jobs[g].subjs;
    auto& querys = *jobs[g].querys;
    auto& out = *jobs[g].distances;
    for (uint32_t i = 0; i < subjs.size(); ++i) {
      const std::string& subj = subjs[i].seq;
      for (uint32_t j = 0; j < querys.size(); ++j) {
        const std::string& query = querys[j].seq;
        if (subj == query || (subj == "N" && query == "NDNNN") || (query == "N" && subj == "NDNNN")) {
          out.emplace_back(i, j, 0, query.size());  // src/compare.cpp:56
        } else if (subj == "N" || query == "N" || subj == "NDNNN" || query == "NDNNN") {
          out.emplace_back(i, j, query.size() - 1, query.size());  // :57
        } else {
          // alignEnd2End(longer, shorter); equal lengths -> (query, subj)  (:59)
          const bool subj_longer = subj.size() > query.size();
          work.push_back({subj_longer ? &subj : &query, subj_longer ? &query : &subj, -1, 0, 0, 0});
          slots.push_back({g, out.size()});
          out.emplace_back(i, j, 0, 0);
        }
      }
    }
  }
  std::vector<int> scores, lengths;  // the CIGAR itself is not used, only its length
  batched_cigars(aligner, work, nullptr, &scores, &lengths);
  for (size_t w = 0; w < work.size(); ++w) {
    DistCompare& d = (*jobs[slots[w].job].distances)[slots[w].at];
    d.edit = scores[w];   // getAlignmentScore()          (:60)
    d.ops = lengths[w];   // getAlignmentCigar().size()   (:61)
  }
}

void get_distances(wfa::WFAligner& aligner, std::vector<ANALLELE>& subjs,
                   std::vector<ANALLELE>& querys, std::vector<DistCompare>& distances) {
  std::vector<CompareJob> one{{&subjs, &querys, &distances}};
  get_distances_batch(aligner, one);
}
