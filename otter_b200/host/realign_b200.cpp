// otter_b200/host/realign_b200.cpp — the gap-affine + CIGAR callers of the hot path
// (reference src/analignments.cpp:11-60 local_realignment, :192-298 rapid_consensus), organised
// as plan -> ONE batched GPU launch (otgpu_affine_cigars) -> apply.  The host decides which
// alignment the reference would run and consumes the CIGAR the way the reference does; every
// alignment itself runs on the device.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>

#include <cmath>

#include "analignments_b200.hpp"
#include "device_turn.hpp"
#include "otter_b200.h"
#include "ppoa_b200.hpp"

namespace {

[[noreturn]] void die(const char* what, int st, otgpu_ctx* ctx) {
  std::cerr << "ERROR: otter_b200 " << what << ": " << otgpu_strerror(st) << " ("
            << (ctx ? otgpu_last_cuda_error(ctx) : "") << ")" << std::endl;
  exit(1);
}

}  // namespace

// -------------------------------------------------------------------------------------------------
void affine_cigars(wfa::WFAligner& aligner2, const std::vector<CigarJob>& jobs,
                   std::vector<std::string>& cigars) {
  batched_cigars(aligner2, jobs, &cigars, nullptr, nullptr);
}

// Gap-affine aligner -> otgpu_affine_cigars; edit aligner (compare's) -> otgpu_edit_cigars.
void batched_cigars(wfa::WFAligner& aligner2, const std::vector<CigarJob>& jobs,
                    std::vector<std::string>* cigars, std::vector<int>* scores_out,
                    std::vector<int>* lengths_out) {
  if (cigars) cigars->assign(jobs.size(), std::string());
  if (scores_out) scores_out->assign(jobs.size(), 0);
  if (lengths_out) lengths_out->assign(jobs.size(), 0);
  if (jobs.empty()) return;
  otgpu_ctx* ctx = aligner2.context();
  if (!ctx) die("affine_cigars", OTGPU_ERR_NO_DEVICE, nullptr);
  // each distinct string is staged once (a representative is the pattern of many jobs)
  std::map<const std::string*, uint32_t> slot;
  std::vector<const std::string*> order;
  size_t total = 0;
  for (const auto& j : jobs)
    for (const std::string* s : {j.pattern, j.text})
      if (slot.emplace(s, (uint32_t)order.size()).second) {
        order.push_back(s);
        total += s->size();
      }
  // pinned staging that lives as long as the calling thread (a pinned allocation per call costs
  // more than the alignments of a small batch)
  struct Pinned {
    void* p = nullptr;
    size_t cap = 0;
    ~Pinned() { if (p) otgpu_host_free(p); }
  };
  thread_local Pinned pin;
  int st = OTGPU_OK;
  if (total + 64 > pin.cap) {
    if (pin.p) otgpu_host_free(pin.p);
    pin.p = nullptr;
    pin.cap = 0;
    const size_t want = total + total / 4 + 4096;
    st = otgpu_host_alloc(want, &pin.p);
    if (st != OTGPU_OK) die("pinned allocation", st, ctx);
    pin.cap = want;
  }
  uint8_t* bases = static_cast<uint8_t*>(pin.p);
  std::vector<uint64_t> off(order.size());
  std::vector<uint32_t> len(order.size());
  uint64_t used = 0;
  for (size_t i = 0; i < order.size(); ++i) {
    off[i] = used;
    len[i] = (uint32_t)order[i]->size();
    std::memcpy(bases + used, order[i]->data(), order[i]->size());
    used += order[i]->size();
  }
  std::vector<otgpu_pair> pairs(jobs.size());
  for (size_t i = 0; i < jobs.size(); ++i) {
    const CigarJob& j = jobs[i];
    pairs[i].pattern = slot[j.pattern];
    pairs[i].text = slot[j.text];
    pairs[i].pattern_begin_free = j.pb < 0 ? -1 : j.pb;
    pairs[i].pattern_end_free = j.pb < 0 ? 0 : j.pe;
    pairs[i].text_begin_free = j.pb < 0 ? 0 : j.tb;
    pairs[i].text_end_free = j.pb < 0 ? 0 : j.te;
  }
  std::vector<int32_t> scores(jobs.size()), clen(jobs.size());
  const char* packed = nullptr;        // context-owned pinned memory, valid until the next call
  const uint64_t* begin = nullptr;
  std::unique_lock<otb200::DeviceTurn> turn(otb200::device_turn(aligner2.device()));
  if (cigars) {  // the strings: packed on the device, only their bytes are copied back
    st = aligner2.isEdit()
        ? otgpu_edit_cigars_packed(ctx, bases, used, off.data(), len.data(), (uint32_t)order.size(),
                                   pairs.data(), pairs.size(), scores.data(), clen.data(), &packed, &begin)
        : otgpu_affine_cigars_packed(ctx, bases, used, off.data(), len.data(), (uint32_t)order.size(),
                                     pairs.data(), pairs.size(), aligner2.mismatch(),
                                     aligner2.gapOpening(), aligner2.gapExtension(), scores.data(),
                                     clen.data(), &packed, &begin);
  } else {  // scores and lengths only: no CIGAR bytes leave the device
    st = aligner2.isEdit()
        ? otgpu_edit_cigars(ctx, bases, used, off.data(), len.data(), (uint32_t)order.size(),
                            pairs.data(), pairs.size(), scores.data(), nullptr, nullptr, clen.data())
        : otgpu_affine_cigars(ctx, bases, used, off.data(), len.data(), (uint32_t)order.size(),
                              pairs.data(), pairs.size(), aligner2.mismatch(), aligner2.gapOpening(),
                              aligner2.gapExtension(), scores.data(), nullptr, nullptr, clen.data());
  }
  turn.unlock();  // (the packed CIGARs sit in pinned memory this context owns until its next call)
  if (st != OTGPU_OK) die("batched_cigars", st, ctx);
  for (size_t i = 0; i < jobs.size(); ++i) {
    if (cigars && clen[i] >= 0) (*cigars)[i].assign(packed + begin[i], (size_t)clen[i]);
    if (scores_out) (*scores_out)[i] = scores[i];
    if (lengths_out) (*lengths_out)[i] = clen[i] >= 0 ? clen[i] : 0;
  }
  // clen < 0 (unfeasible / OOM): the reference ignores the status and reads whatever CIGAR the
  // aligner holds; an empty string is the closest defined behaviour and is not reachable for the
  // end-to-end / ends-free forms otter issues.
}

// -------------------------------------------------------------------------------------------------
// local_realignment, src/analignments.cpp:11-60.  Per half-spanning read with >= flank clipped
// bases on its open side: align the clipped part (pattern) end-to-end to the reference flank
// (text), walk the CIGAR over pattern positions (M = +1, X/D = -1, I ignored) keeping a running
// sum that restarts at 0, and accept the best-scoring prefix if score/flank >= min_sim.
namespace {
struct RealignPlan {
  size_t job;       // index into the RealignJob vector
  int read;         // index into that job's reads
  bool left;        // left_realignment (else right_realignment)
  std::string subseq;
};
}  // namespace

void local_realignments(wfa::WFAligner& aligner2, std::vector<RealignJob>& jobs) {
  std::vector<RealignPlan> plans;
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& reads = *jobs[g].reads;
    const int flank = jobs[g].flank;
    for (int i = 0; i < (int)reads.size(); ++i) {
      const ANREAD& r = reads[i];
      if (r.is_spanning() || !(r.is_spanning_l || r.is_spanning_r)) continue;  // :17
      const bool left = r.is_spanning_r && r.ccoords.first >= flank;           // :18
      const bool right = r.is_spanning_l && (int)r.seq.size() - r.ccoords.second >= flank;  // :19
      if (left) plans.push_back({g, i, true, r.seq.substr(0, r.ccoords.first)});   // :23
      else if (right) plans.push_back({g, i, false, r.seq.substr(r.ccoords.second)});  // :29
    }
  }
  std::vector<CigarJob> cj;
  std::vector<size_t> which;  // plan index of each submitted job
  for (size_t p = 0; p < plans.size(); ++p) {
    if (plans[p].subseq.empty()) continue;  // :34 (nothing to score; the reference still aligns)
    const RealignJob& J = jobs[plans[p].job];
    cj.push_back({&plans[p].subseq, plans[p].left ? J.ref_left : J.ref_right, -1, 0, 0, 0});
    which.push_back(p);
  }
  std::vector<std::string> cigars;
  affine_cigars(aligner2, cj, cigars);
  for (size_t q = 0; q < which.size(); ++q) {
    const RealignPlan& P = plans[which[q]];
    const RealignJob& J = jobs[P.job];
    ANREAD& read = (*J.reads)[P.read];
    std::vector<int> scores(P.subseq.size(), 0);
    int j = 0;
    for (const char op : cigars[q]) {  // :37-47
      if (op == 'I') continue;
      if (j >= (int)scores.size()) break;  // cannot happen: non-I ops == pattern length
      if (op == 'M') scores[j] = (j == 0) ? 1 : scores[j - 1] + 1;
      else if (j > 0 && scores[j - 1] > 0) scores[j] = scores[j - 1] - 1;
      ++j;
    }
    int best = 0;
    for (j = 0; j < (int)scores.size(); ++j)
      if (scores[j] > scores[best]) best = j;  // :49-50, first maximum
    int start_i = best;
    while (start_i > 0 && scores[start_i] > 0) --start_i;  // :51-52
    if ((scores[best] / (double)J.flank) >= J.min_sim) {    // :53
      if (P.left) read.seq = read.seq.substr(best);
      else read.seq = read.seq.substr(0, read.ccoords.second + start_i);
      read.set_is_spanning();
    }
  }
}

// -------------------------------------------------------------------------------------------------
// The alignment half of rapid_consensus, src/analignments.cpp:192-281.
namespace {
// Which aligner2 call the reference issues for (rep, read) at :262-279.  Returns false for the
// one combination where it issues none (read longer than rep and spanning on neither side).
bool consensus_form(const ANREAD& rep, const ANREAD& read, CigarJob& j) {
  const int length_diff = (int)rep.seq.size() - (int)read.seq.size();  // :261
  j.pattern = &rep.seq;
  j.text = &read.seq;
  j.pb = j.pe = j.tb = j.te = 0;
  if (read.is_spanning() || length_diff < 0) {
    if (length_diff >= 0) { j.pb = -1; return true; }                     // alignEnd2End (:263)
    if (read.is_spanning_l) { j.te = -length_diff; return true; }          // (:265)
    if (read.is_spanning_r) { j.tb = -length_diff; return true; }          // (:266)
    return false;
  }
  if (read.is_spanning_l) j.pe = length_diff;                              // (:271)
  else if (read.is_spanning_r) j.pb = length_diff;                         // (:272)
  else j.pb = j.pe = length_diff / 2;                                      // (:273)
  return true;
}
}  // namespace

void rapid_consensus_alignments(wfa::WFAligner& aligner2, std::vector<ConsensusJob>& jobs) {
  std::vector<CigarJob> cj;
  struct Ref { size_t job; int label; size_t member; bool aligned; };
  std::vector<Ref> refs;  // every (job,label,member) that feeds the partial-order graph, in order
  for (size_t g = 0; g < jobs.size(); ++g) {
    ConsensusJob& J = jobs[g];
    auto& reads = *J.reads;
    auto& labels = *J.labels;
    auto& valid = *J.valid_indeces;
    if (valid.empty()) {  // :194-197
      std::cerr << "ERROR: empty vector of valid read-indeces" << std::endl;
      exit(1);
    }
    J.plans.assign(J.total_alleles, ConsensusPlan());
    for (int label = 0; label < J.total_alleles; ++label) {
      ConsensusPlan& P = J.plans[label];
      for (uint32_t i = 0; i < valid.size(); ++i)
        if (label == labels[valid[i]]) {  // :205-210
          P.valid_reads.emplace_back(valid[i]);
          P.valid_pos.emplace_back(i);
        }
      if (P.valid_reads.empty()) {  // :211-214
        std::cerr << "ERROR: empty vector of valid read-indeces for allele cluster " << label << std::endl;
        exit(1);
      }
      P.rep_pos = (int)J.valid_distmatrix->get_medoid(P.valid_pos);  // :216
      P.rep = valid[P.rep_pos];                                       // :218
      for (int i = 0; i < (int)reads.size(); ++i)
        if (i != P.rep && labels[i] == label) P.members.emplace_back(i);  // :222-223
      if (P.members.size() + 1 <= 2) continue;  // :258: no graph, allele = first valid read
      P.cigars.assign(P.members.size(), std::string());
      for (size_t m = 0; m < P.members.size(); ++m) {
        CigarJob j;
        const bool aligned = consensus_form(reads[P.rep], reads[P.members[m]], j);
        if (aligned) cj.push_back(j);
        refs.push_back({g, label, m, aligned});
      }
    }
  }
  std::vector<std::string> cigars;
  affine_cigars(aligner2, cj, cigars);
  // getAlignmentCigar() after a skipped call returns the previous alignment's CIGAR (:280): within a
  // batch that is replayed with a running "last CIGAR", in the order a single reference thread
  // would have run (each CIGAR is moved to its plan, not copied: they are the bulk of the
  // host-side bytes).  DELIBERATE DIVERGENCE for a skipped call that comes FIRST in a batch: the
  // reference then inserts whatever its thread-local aligner produced last - in an earlier region
  // of that thread's block, or in local_realignment - i.e. a value that depends on the thread
  // schedule; here it is the facade's own last single-call CIGAR (empty unless the caller made
  // one), which is deterministic.  tests/test_gpu_affine.py pins this.
  const std::string initial = aligner2.getAlignmentCigar();
  const std::string* last = &initial;
  size_t q = 0;
  for (const Ref& r : refs) {
    std::string& slot = jobs[r.job].plans[r.label].cigars[r.member];
    if (r.aligned) {
      slot = std::move(cigars[q++]);
      last = &slot;
    } else {
      slot = *last;
    }
  }
}

#ifdef OTTER_B200_IN_TREE
// -------------------------------------------------------------------------------------------------
// Reference signature (src/analignments.hpp:26-35) for an in-tree build: the flank fetch is
// otter's own code (FaidxInstance), the alignments run batched on the GPU.
void local_realignment(const std::string& chr, const int& start, const int& end, const int& flank,
                       const double& min_sim, const FaidxInstance& faidx_inst,
                       std::vector<ANREAD>& reads, wfa::WFAligner& aligner) {
  std::string ref_left, ref_right;
  bool need_left = false, need_right = false;
  for (const auto& r : reads) {
    if (r.is_spanning() || !(r.is_spanning_l || r.is_spanning_r)) continue;
    const bool left = r.is_spanning_r && r.ccoords.first >= flank;
    const bool right = r.is_spanning_l && (int)r.seq.size() - r.ccoords.second >= flank;
    if (left) need_left = true;
    else if (right) need_right = true;
  }
  if (need_left) faidx_inst.fetch(chr, start - flank, start, ref_left);   // :22
  if (need_right) faidx_inst.fetch(chr, end, end + flank, ref_right);     // :28
  std::vector<RealignJob> one{{flank, min_sim, &ref_left, &ref_right, &reads}};
  local_realignments(aligner, one);
}
#endif  // OTTER_B200_IN_TREE

// -------------------------------------------------------------------------------------------------
// rapid_consensus (reference src/analignments.cpp:192-298, signature src/analignments.hpp:62-71):
// the alignments run batched on the GPU, the allele bookkeeping follows the reference line by
// line, the partial-order consensus is ppoa_b200.hpp (same result as otter's PPOA, linear time).
namespace {
double standard_error(const std::vector<double>& values) {  // compute_se, :179-190
  if (values.empty()) return -1.0;
  double u = 0.0, n = 0.0;
  for (const auto& v : values) u += v;
  u /= values.size();
  for (const auto& v : values) n += std::pow(v - u, 2.0);
  return std::sqrt(n / (values.size() - 1)) / std::sqrt(values.size());
}
}  // namespace

void rapid_consensus(const bool& ignore_haps, std::vector<ANREAD>& reads, std::vector<int>& labels,
                     std::vector<int>& valid_indeces, int& total_alleles,
                     DistMatrix& valid_distmatrix, wfa::WFAligner& aligner,
                     std::vector<ANALLELE>& alleles) {
  std::vector<ConsensusJob> one(1);
  one[0].reads = &reads;
  one[0].labels = &labels;
  one[0].valid_indeces = &valid_indeces;
  one[0].total_alleles = total_alleles;
  one[0].valid_distmatrix = &valid_distmatrix;
  rapid_consensus_alignments(aligner, one);
  rapid_consensus_finish(ignore_haps, one[0], alleles);
}

namespace {
// Everything rapid_consensus does for one allele except the graph (src/analignments.cpp:227-258):
// coverages, standard error, haplotag check.  Returns true when the allele needs the partial-order
// consensus (three reads or more); otherwise allele.seq is already final.
bool allele_bookkeeping(const bool& ignore_haps, ConsensusJob& job, int label, ANALLELE& allele) {
  std::vector<ANREAD>& reads = *job.reads;
  DistMatrix& valid_distmatrix = *job.valid_distmatrix;
  const ConsensusPlan& P = job.plans[label];
  allele.tcov = reads.size();               // :227-229
  allele.acov = P.members.size() + 1;
  allele.scov = P.valid_reads.size();
  if (P.valid_pos.size() == 1) allele.se = 0;  // :231-237
  else if (P.valid_pos.size() == 2) allele.se = valid_distmatrix.get_dist(P.valid_pos[0], P.valid_pos[1]);
  else {
    std::vector<double> d;
    for (const auto& i : P.valid_pos)
      if ((int)i != P.rep_pos) d.emplace_back(valid_distmatrix.get_dist(i, P.rep_pos));
    allele.se = standard_error(d);
  }
  if (!ignore_haps) {  // :239-256
    int ps = -1, hp = -1;
    bool conflicting = false;
    for (const auto& i : P.valid_reads) {
      if (ps < 0) ps = reads[i].hpt.ps;
      else if (ps != reads[i].hpt.ps) conflicting = true;
      if (hp < 0) hp = reads[i].hpt.hp;
      else if (hp != reads[i].hpt.hp) conflicting = true;
    }
    if (conflicting) {
      std::cerr << "ERROR: conflicting haplotag information:\n";
      for (const auto& i : P.valid_reads)
        std::cerr << reads[i].name << '\t' << reads[i].hpt.ps << '\t' << reads[i].hpt.hp << '\n';
      std::cerr << std::endl;
      exit(1);
    }
    allele.hpt = reads[P.rep].hpt;
  }
  if (P.members.size() + 1 <= 2) {  // :258
    allele.seq = reads[P.valid_reads.front()].seq;
    return false;
  }
  return true;
}
// adjust_weights parameters of an allele with n reads (:285-288)
void damping(size_t n_reads_of_allele, float& c, float& t) {
  c = n_reads_of_allele * 0.4;
  t = 0.3;
  if (n_reads_of_allele < 4) c = 1.0;
}
}  // namespace

void rapid_consensus_finish(const bool& ignore_haps, ConsensusJob& job,
                            std::vector<ANALLELE>& alleles) {
  std::vector<ANREAD>& reads = *job.reads;
  const int total_alleles = job.total_alleles;
  for (int label = 0; label < total_alleles; ++label) {
    const ConsensusPlan& P = job.plans[label];
    auto& allele = alleles[label];
    if (!allele_bookkeeping(ignore_haps, job, label, allele)) continue;
    otb200::Ppoa poa;  // src/anppoa.hpp restated (ppoa_b200.hpp)
    poa.init(reads[P.rep].seq);
    for (size_t m = 0; m < P.members.size(); ++m) {
      ANREAD& read = reads[P.members[m]];
      std::string cigar = P.cigars[m];
      poa.insert_alignment(read.seq, cigar, read.is_spanning_l, read.is_spanning_r);  // :282
    }
    float c, t;
    damping(P.members.size() + 1, c, t);
    poa.adjust_weights(c, t);
    poa.consensus(allele.seq);
    if (allele.seq.empty()) allele.seq = "N";
  }
}

// The same for many jobs, with the partial-order consensus of every allele in ONE device call
// (otgpu_ppoa_consensus, csrc/ppoa.cuh): the host keeps the bookkeeping, the graphs are built and
// walked on the GPU.  Same alleles as rapid_consensus_finish job by job.
void rapid_consensus_finish_device(const bool& ignore_haps, wfa::WFAligner& aligner2,
                                   std::vector<ConsensusJob*>& jobs,
                                   std::vector<std::vector<ANALLELE>*>& alleles) {
  otgpu_ctx* ctx = aligner2.context();
  if (!ctx) die("rapid_consensus", OTGPU_ERR_NO_DEVICE, nullptr);
  struct Slot { size_t job; int label; };
  std::vector<Slot> slots;
  std::vector<otgpu_ppoa_graph> graphs;
  std::vector<otgpu_ppoa_member> members;
  std::vector<uint64_t> off;
  std::vector<uint32_t> len;
  std::string bases, cigars;
  for (size_t j = 0; j < jobs.size(); ++j) {
    ConsensusJob& job = *jobs[j];
    std::vector<ANREAD>& reads = *job.reads;
    size_t first_read = (size_t)-1;  // this job's reads enter the batch once, on first use
    for (int label = 0; label < job.total_alleles; ++label) {
      if (!allele_bookkeeping(ignore_haps, job, label, (*alleles[j])[label])) continue;
      if (first_read == (size_t)-1) {
        first_read = off.size();
        for (const ANREAD& r : reads) {
          off.push_back(bases.size());
          len.push_back((uint32_t)r.seq.size());
          bases += r.seq;
        }
      }
      const ConsensusPlan& P = job.plans[label];
      otgpu_ppoa_graph g;
      g.backbone = (uint32_t)(first_read + P.rep);
      g.first_member = (uint32_t)members.size();
      g.n_members = (uint32_t)P.members.size();
      damping(P.members.size() + 1, g.c, g.t);
      for (size_t m = 0; m < P.members.size(); ++m) {
        const ANREAD& read = reads[P.members[m]];
        otgpu_ppoa_member M;
        M.read = (uint32_t)(first_read + P.members[m]);
        M.flags = (read.is_spanning_l ? 1u : 0u) | (read.is_spanning_r ? 2u : 0u);
        M.cigar_off = cigars.size();
        M.cigar_len = (uint32_t)P.cigars[m].size();
        M.reserved = 0;
        cigars += P.cigars[m];
        members.push_back(M);
      }
      graphs.push_back(g);
      slots.push_back({j, label});
    }
  }
  if (graphs.empty()) return;
  std::vector<int32_t> clen(graphs.size());
  const char* cons = nullptr;
  const uint64_t* begin = nullptr;
  int st;
  {
    std::lock_guard<otb200::DeviceTurn> turn(otb200::device_turn(aligner2.device()));
    st = otgpu_ppoa_consensus(ctx, reinterpret_cast<const uint8_t*>(bases.data()), bases.size(), off.data(),
                              len.data(), (uint32_t)off.size(), graphs.data(), (uint32_t)graphs.size(),
                              members.data(), (uint32_t)members.size(), cigars.data(), cigars.size(),
                              clen.data(), &cons, &begin);
  }
  if (st != OTGPU_OK) die("partial-order consensus", st, ctx);
  for (size_t i = 0; i < slots.size(); ++i) {
    ANALLELE& allele = (*alleles[slots[i].job])[slots[i].label];
    allele.seq.assign(cons + begin[i], (size_t)clen[i]);
    if (allele.seq.empty()) allele.seq = "N";
  }
}
