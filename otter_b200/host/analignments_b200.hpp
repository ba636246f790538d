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

// Where the time of this thread's last all-spanning fill_dist_matrices call went (seconds): the
// whole call, the device call inside it (uploads, kernels, downloads, and the waits below), the
// part of the device call spent waiting for the staging copy, the score -> dissimilarity
// conversion (runs under the kernels of later chunks), and the number of chunks.
struct FillTimes { double total_s, device_call_s, stage_wait_s, unpack_s, chunks; };
FillTimes last_fill_times();

// The same over several devices of one box from ONE host process: `aligners`
// holds one WFAlignerEdit per device (WFAligner::setDefaultDevice(d) before constructing it).
// Regions are independent, so the jobs are partitioned across the aligners by estimated work
// (greedy longest-processing-time over sum of |x|*|y| per region), each share runs
// fill_dist_matrices on its own host thread / device, and every job's DistMatrix is filled in
// place: the gather is the host memory the caller already owns, there is no collective.  Results
// do not depend on the partition.
void fill_dist_matrices_multi(const bool& ignore_haps, std::vector<wfa::WFAligner*>& aligners,
                              std::vector<RegionDistJob>& jobs,
                              std::vector<FillTimes>* times = nullptr);  // times: one per aligner
// Which aligner each job goes to (exposed for tests): loads[] receives the per-aligner estimate.
std::vector<int> partition_regions_by_work(const std::vector<RegionDistJob>& jobs, int n_parts,
                                           std::vector<double>* loads = nullptr);

struct RegionReassignJob {
  std::vector<ANREAD>* reads;
  std::vector<int>* labels;
  int total_alleles;
};
void invalid_reassignments(const bool& ignore_haps, const double& min_sim,
                           const double& max_error, wfa::WFAligner& aligner,
                           std::vector<RegionReassignJob>& jobs);

// ---- gap-affine + CIGAR callers (reference src/analignments.cpp:11-60, 192-298) -----------------
// One alignment request of `aligner2` (src/assemble.cpp:50).  pb < 0: alignEnd2End(pattern,text);
// else alignEndsFree(pattern, pb, pe, text, tb, te).
struct CigarJob {
  const std::string* pattern;
  const std::string* text;
  int pb, pe, tb, te;
};
// All jobs in ONE launch; cigars[i] is what aligner2.getAlignmentCigar() returns after job i.
void affine_cigars(wfa::WFAligner& aligner2, const std::vector<CigarJob>& jobs,
                   std::vector<std::string>& cigars);

// The same for either aligner flavour that produces CIGARs: gap-affine (aligner2) or edit
// distance with CIGAR (`compare`'s WFAlignerEdit(Alignment), src/compare.cpp:95).
// Each output is optional: cigars (the strings), scores_out (getAlignmentScore()), lengths_out
// (getAlignmentCigar().size()).  With cigars == nullptr no CIGAR bytes leave the device.
void batched_cigars(wfa::WFAligner& aligner, const std::vector<CigarJob>& jobs,
                    std::vector<std::string>* cigars, std::vector<int>* scores_out,
                    std::vector<int>* lengths_out);

// local_realignment (src/analignments.cpp:11-60) with the two reference flanks already fetched
// (ref_left = ref[start-flank,start), ref_right = ref[end,end+flank), upper-cased as
// FaidxInstance::fetch does).  Mutates read.seq / spanning flags exactly like the reference.
struct RealignJob {
  int flank;
  double min_sim;
  const std::string* ref_left;
  const std::string* ref_right;
  std::vector<ANREAD>* reads;
};
void local_realignments(wfa::WFAligner& aligner2, std::vector<RealignJob>& jobs);

// The alignment half of rapid_consensus (src/analignments.cpp:192-281): per allele label the
// medoid representative and, for alleles with >= 3 reads, each member's CIGAR against it, in the
// order the reference inserts them into the partial-order graph.
struct ConsensusPlan {
  int rep = -1;                      // index into reads of the representative (medoid)
  int rep_pos = -1;                  // the same read as a position in valid_indeces
  std::vector<int> members;          // reads with this label, excluding rep, ascending (:228)
  std::vector<uint32_t> valid_reads; // valid (spanning) reads with this label, indexed to reads
  std::vector<uint32_t> valid_pos;   // the same, indexed to valid_indeces
  std::vector<std::string> cigars;   // one per member when members.size() + 1 > 2, else empty
};
struct ConsensusJob {
  std::vector<ANREAD>* reads;
  std::vector<int>* labels;
  std::vector<int>* valid_indeces;
  int total_alleles;
  DistMatrix* valid_distmatrix;
  std::vector<ConsensusPlan> plans;  // out: one per label
};
void rapid_consensus_alignments(wfa::WFAligner& aligner2, std::vector<ConsensusJob>& jobs);

// rapid_consensus with the reference signature (src/analignments.hpp:62-71): alignments batched
// on the GPU, allele bookkeeping as the reference, partial-order consensus by ppoa_b200.hpp.
void rapid_consensus(const bool& ignore_haps, std::vector<ANREAD>& reads, std::vector<int>& labels,
                     std::vector<int>& valid_indeces, int& total_alleles,
                     DistMatrix& valid_distmatrix, wfa::WFAligner& aligner,
                     std::vector<ANALLELE>& alleles);
// Its host half (src/analignments.cpp:227-298: coverage, standard error, haplotag check,
// partial-order consensus) for one job whose plans rapid_consensus_alignments has filled — lets a
// caller batch the alignments of many regions and finish each region on its own host thread.
void rapid_consensus_finish(const bool& ignore_haps, ConsensusJob& job,
                            std::vector<ANALLELE>& alleles);

// The same host half for many jobs at once, with the partial-order consensus of every allele of
// the batch in ONE device call (otgpu_ppoa_consensus): alleles[j] receives job j's alleles.
void rapid_consensus_finish_device(const bool& ignore_haps, wfa::WFAligner& aligner2,
                                   std::vector<ConsensusJob*>& jobs,
                                   std::vector<std::vector<ANALLELE>*>& alleles);

#ifdef OTTER_B200_IN_TREE
#include "anfahelper.hpp"
// Reference signature; the alignment calls inside run batched on the GPU.
void local_realignment(const std::string& chr, const int& start, const int& end, const int& flank,
                       const double& min_sim, const FaidxInstance& faidx_inst,
                       std::vector<ANREAD>& reads, wfa::WFAligner& aligner);
#endif
