// otter_b200/host/analignments_b200.cpp — see analignments_b200.hpp.
//
// Every alignment below runs on the GPU (include/otter_b200.h); the host only decides WHICH
// alignment the reference would run for a pair (src/analignments.cpp:62-101) and forms the FP64
// result with the reference's own expressions, so dissimilarities are bit-identical.
#include "analignments_b200.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "device_turn.hpp"
#include "otter_b200.h"

namespace {

// What align_anreads (src/analignments.cpp:62-101) does for (x, y), minus the alignment itself.
struct PairPlan {
  enum Kind { kNone, kE2E, kEndsFree } kind;
  bool pattern_is_y;  // which read is the pattern
  int pb, pe;         // ends-free: pattern begin/end free
  double denom;       // score / denom is the returned distance
};

PairPlan plan_pair(const ANREAD& x, const ANREAD& y) {
  PairPlan p{PairPlan::kNone, false, 0, 0, 1.0};
  if ((x.is_spanning() && y.is_spanning()) || (y.is_spanning() && x.seq.size() >= y.seq.size())) {
    const bool x_is_smallest = x.seq.size() < y.seq.size();  // :67
    p.kind = PairPlan::kE2E;
    p.pattern_is_y = x_is_smallest;  // alignEnd2End(longer, shorter), tie -> x is the pattern (:70)
    p.denom = x_is_smallest ? (double)y.seq.size() : (double)x.seq.size();
  } else if (y.is_spanning()) {  // x shorter and not spanning (:82-98; length_diff > 0 here)
    const int length_diff = (int)y.seq.size() - (int)x.seq.size();
    p.kind = PairPlan::kEndsFree;
    p.pattern_is_y = true;
    if (x.is_spanning_l) { p.pb = 0; p.pe = length_diff; }
    else if (x.is_spanning_r) { p.pb = length_diff; p.pe = 0; }
    else { p.pb = length_diff / 2; p.pe = length_diff / 2; }
    p.denom = (double)x.seq.size();
  }
  return p;
}

[[noreturn]] void die(const char* what, int st, otgpu_ctx* ctx) {
  std::cerr << "ERROR: otter_b200 " << what << ": " << otgpu_strerror(st) << " ("
            << (ctx ? otgpu_last_cuda_error(ctx) : "") << ")" << std::endl;
  exit(1);  // the reference's convention for fatal conditions (src/analignments.cpp:146-153)
}

// Reads of a batch laid out for the C ABI (pinned when the device is there).
struct Staging {
  uint8_t* bases = nullptr;
  size_t cap = 0;
  std::vector<uint64_t> off;
  std::vector<uint32_t> len;
  uint64_t used = 0;
  ~Staging() { if (bases) otgpu_host_free(bases); }
  void reserve(size_t n) {
    if (n <= cap) return;
    if (bases) otgpu_host_free(bases);
    void* p = nullptr;
    size_t want = n + n / 4 + 4096;
    int st = otgpu_host_alloc(want, &p);
    if (st != OTGPU_OK) die("pinned allocation", st, nullptr);
    bases = static_cast<uint8_t*>(p);
    cap = want;
  }
  void clear() { off.clear(); len.clear(); used = 0; }
  uint32_t add(const std::string& s) {
    off.push_back(used);
    len.push_back((uint32_t)s.size());
    std::memcpy(bases + used, s.data(), s.size());
    used += s.size();
    return (uint32_t)(off.size() - 1);
  }
};

// One staging area per device context (an aligner owns its context and is used by one thread at
// a time, like the reference's per-thread aligners): the pinned buffer is allocated once and
// re-used by every later call on that aligner, whichever host thread makes it.
Staging& staging(otgpu_ctx* ctx) {
  static std::mutex m;
  static std::map<otgpu_ctx*, std::unique_ptr<Staging>> by_ctx;
  std::lock_guard<std::mutex> lk(m);
  auto& p = by_ctx[ctx];
  if (!p) p.reset(new Staging());
  return *p;
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
thread_local FillTimes g_last_fill = {0, 0, 0, 0, 0};
// how many fill_dist_matrices calls fill_dist_matrices_multi is about to run side by side
std::atomic<int> g_fill_shares{1};

// Producer side of the hooked C-ABI call: worker threads copy the reads into the pinned staging
// buffer front to back and publish how far they got; the library waits for the bytes of a chunk
// right before it uploads them, and hands back each chunk's scores as soon as they are final, so
// the copy of chunk i+1 and the conversion of chunk i-1's scores run under the kernels of chunk i.

}  // namespace

// ---------------------------------------------------------------------------------------------
double align_anreads(wfa::WFAligner& aligner, ANREAD& read_x, ANREAD& read_y) {
  if (read_x.seq == read_y.seq) return 0.0;  // :64
  const PairPlan p = plan_pair(read_x, read_y);
  if (p.kind == PairPlan::kNone) return -1.0;  // :100
  std::string& pat = p.pattern_is_y ? read_y.seq : read_x.seq;
  std::string& txt = p.pattern_is_y ? read_x.seq : read_y.seq;
  if (p.kind == PairPlan::kE2E) aligner.alignEnd2End(pat, txt);
  else aligner.alignEndsFree(pat, p.pb, p.pe, txt, 0, 0);
  return aligner.getAlignmentScore() / p.denom;
}

double get_dist_anreads(const bool& ignore_haps, wfa::WFAligner& aligner, ANREAD& read_x,
                        ANREAD& read_y) {
  if (ignore_haps) return align_anreads(aligner, read_x, read_y);
  if (read_x.hpt.is_defined() && read_y.hpt.is_defined()) {  // :109-111
    if (read_x.hpt == read_y.hpt) return 0;
    return 1.0;
  }
  return 1.0;  // ambiguous, set them apart (:113)
}

void fill_dist_matrix(const bool& ignore_haps, wfa::WFAligner& aligner, std::vector<ANREAD>& reads,
                      std::vector<int>& indeces, DistMatrix& distmatrix) {
  std::vector<RegionDistJob> one{{&reads, &indeces, &distmatrix}};
  fill_dist_matrices(ignore_haps, aligner, one);
}

// Estimated alignment work of one region: the wavefront work of a pair grows with the product of
// the two lengths at a given divergence, so sum |x|*|y| over the pairs = ((sum l)^2 - sum l^2) / 2.
std::vector<int> partition_regions_by_work(const std::vector<RegionDistJob>& jobs, int n_parts,
                                           std::vector<double>* loads) {
  std::vector<int> part(jobs.size(), 0);
  std::vector<double> load(n_parts > 0 ? n_parts : 1, 0.0);
  if (n_parts > 1) {
    std::vector<std::pair<double, size_t>> work(jobs.size());
    for (size_t g = 0; g < jobs.size(); ++g) {
      double sum = 0.0, sq = 0.0;
      for (int i : *jobs[g].indeces) {
        const double l = (double)(*jobs[g].reads)[i].seq.size();
        sum += l;
        sq += l * l;
      }
      work[g] = {(sum * sum - sq) / 2, g};
    }
    // heaviest region first, each to the least loaded part (ties: lower index, so it is stable)
    std::stable_sort(work.begin(), work.end(),
                     [](const auto& a, const auto& b) { return a.first > b.first; });
    for (const auto& w : work) {
      int best = 0;
      for (int p = 1; p < n_parts; ++p)
        if (load[p] < load[best]) best = p;
      part[w.second] = best;
      load[best] += w.first;
    }
  }
  if (loads) *loads = load;
  return part;
}

void fill_dist_matrices_multi(const bool& ignore_haps, std::vector<wfa::WFAligner*>& aligners,
                              std::vector<RegionDistJob>& jobs, std::vector<FillTimes>* times) {
  if (aligners.empty()) die("fill_dist_matrices_multi", OTGPU_ERR_ARGS, nullptr);
  if (times) times->assign(aligners.size(), FillTimes{0, 0, 0, 0, 0});
  if (aligners.size() == 1 || !ignore_haps) {
    fill_dist_matrices(ignore_haps, *aligners[0], jobs);
    if (times) (*times)[0] = last_fill_times();
    return;
  }
  for (auto* a : aligners)
    if (!a || !a->context()) die("fill_dist_matrices_multi", OTGPU_ERR_NO_DEVICE, nullptr);
  const std::vector<int> part = partition_regions_by_work(jobs, (int)aligners.size());
  std::vector<std::vector<RegionDistJob>> share(aligners.size());
  for (size_t g = 0; g < jobs.size(); ++g) share[part[g]].push_back(jobs[g]);
  // one host thread per device; staging buffers and contexts are per aligner
  int n_shares = 0;
  for (const auto& sh : share) n_shares += !sh.empty();
  g_fill_shares.store(std::max(1, n_shares), std::memory_order_relaxed);
  std::vector<std::thread> pool;
  for (size_t d = 1; d < aligners.size(); ++d)
    if (!share[d].empty())
      pool.emplace_back([&, d] {
        fill_dist_matrices(ignore_haps, *aligners[d], share[d]);
        if (times) (*times)[d] = last_fill_times();
      });
  if (!share[0].empty()) {
    fill_dist_matrices(ignore_haps, *aligners[0], share[0]);
    if (times) (*times)[0] = last_fill_times();
  }
  for (auto& th : pool) th.join();
  g_fill_shares.store(1, std::memory_order_relaxed);
}

void fill_dist_matrices(const bool& ignore_haps, wfa::WFAligner& aligner,
                        std::vector<RegionDistJob>& jobs) {
  if (!ignore_haps) {  // haplotag rule only, no alignment (:108-114)
    for (auto& J : jobs) {
      auto& idx = *J.indeces;
      for (uint32_t i = 0; i < idx.size(); ++i)
        for (uint32_t j = i + 1; j < idx.size(); ++j)
          J.distmatrix->set_dist(i, j, get_dist_anreads(false, aligner, (*J.reads)[idx[i]],
                                                         (*J.reads)[idx[j]]));
    }
    return;
  }
  otgpu_ctx* ctx = aligner.context();
  if (!ctx) die("fill_dist_matrix", OTGPU_ERR_NO_DEVICE, nullptr);
  Staging& S = staging(ctx);
  size_t total = 0;
  bool all_spanning = true;
  for (auto& J : jobs)
    for (int i : *J.indeces) {
      total += (*J.reads)[i].seq.size();
      all_spanning &= (*J.reads)[i].is_spanning();
    }
  S.reserve(total + 64);
  S.clear();
  std::vector<uint32_t> region_start(jobs.size() + 1, 0);
  std::vector<uint64_t> score_off(jobs.size(), 0);
  uint64_t n_scores = 0;
  std::vector<const std::string*> seqs;
  for (size_t g = 0; g < jobs.size(); ++g) {
    for (int i : *jobs[g].indeces) {  // lay the reads out; the bytes are copied below
      const std::string& q = (*jobs[g].reads)[i].seq;
      S.off.push_back(S.used);
      S.len.push_back((uint32_t)q.size());
      S.used += q.size();
      seqs.push_back(&q);
    }
    region_start[g + 1] = (uint32_t)S.off.size();
    score_off[g] = n_scores;
    const uint64_t n = jobs[g].indeces->size();
    n_scores += n * (n ? n - 1 : 0) / 2;
  }
  const double t_begin = now_s();
  // The copy into pinned staging is the longest host step of a wave.  Worker threads run it front
  // to back and publish a watermark; in the all-spanning case (below) the device call starts at
  // once and waits, chunk by chunk, only for the bytes it is about to upload.
  // Up to 6 copy threads per call, fewer when several calls run at once (fill_dist_matrices_multi:
  // one call per device): with 8 devices on 32 cores, 8 x (6 + 1) threads oversubscribe the host
  // and the slowest share waits for its copy (profiles/README.md r2v: 15 of 52 ms).
  struct InFlight {
    std::atomic<int>& n;
    explicit InFlight(std::atomic<int>& c) : n(c) { n.fetch_add(1, std::memory_order_relaxed); }
    ~InFlight() { n.fetch_sub(1, std::memory_order_relaxed); }
  };
  static std::atomic<int> calls_in_flight{0};
  const InFlight in_flight(calls_in_flight);
  int workers = 1;
  if (total > ((size_t)8 << 20)) {
    const int hw = (int)std::thread::hardware_concurrency();
    const int sharing = std::max(g_fill_shares.load(std::memory_order_relaxed),
                                 calls_in_flight.load(std::memory_order_relaxed));
    workers = std::min(6, std::max(2, (hw > 0 ? hw : 8) / sharing - 1));
  }
  const size_t n_seqs = seqs.size();
  constexpr size_t kBlock = 64;  // reads per unit of work
  const size_t n_blocks = (n_seqs + kBlock - 1) / kBlock;
  std::vector<std::atomic<uint8_t>> block_done(n_blocks);
  for (auto& d : block_done) d.store(0, std::memory_order_relaxed);
  std::atomic<size_t> next_block{0};
  auto copy = [&](int) {  // blocks are taken in order, so the buffer fills front to back
    for (;;) {
      const size_t b = next_block.fetch_add(1, std::memory_order_relaxed);
      if (b >= n_blocks) return;
      const size_t r1 = std::min(n_seqs, (b + 1) * kBlock);
      for (size_t r = b * kBlock; r < r1; ++r) std::memcpy(S.bases + S.off[r], seqs[r]->data(), seqs[r]->size());
      block_done[b].store(1, std::memory_order_release);
    }
  };
  double stage_wait_s = 0.0;
  size_t blocks_seen = 0;  // every block below this one is known to be complete
  auto wait_reads = [&](size_t upto_read) {  // returns once reads [0, upto_read) are in place
    const size_t need = std::min(n_blocks, (upto_read + kBlock - 1) / kBlock);
    if (blocks_seen >= need) return;
    const double t0 = now_s();
    for (; blocks_seen < need; ++blocks_seen)
      while (!block_done[blocks_seen].load(std::memory_order_acquire)) std::this_thread::yield();
    stage_wait_s += now_s() - t0;
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < workers; ++t) pool.emplace_back(copy, t);
  struct Joiner {
    std::vector<std::thread>& p;
    ~Joiner() { for (auto& th : p) if (th.joinable()) th.join(); }
  } joiner{pool};
  std::vector<int32_t> scores(n_scores);
  if (all_spanning) {
    // every pair is alignEnd2End(longer, shorter): pairs are expanded on the device
    double unpack_s = 0.0;
    // scores [first, first+n) are final: form the dissimilarities of the regions they belong to
    size_t g_next = 0;
    auto unpack = [&](uint64_t first, uint64_t n) {
      const double t0 = now_s();
      const uint64_t end = first + n;
      for (; g_next < jobs.size(); ++g_next) {
        const size_t g = g_next;
        const uint64_t cnt = jobs[g].indeces->size();
        if (score_off[g] + cnt * (cnt ? cnt - 1 : 0) / 2 > end) break;
        auto& idx = *jobs[g].indeces;
        auto& reads = *jobs[g].reads;
        const int32_t* sc = scores.data() + score_off[g];
        // (i,j) in this order is the order of the condensed store itself (src/andistmat.cpp:15-21),
        // so element q is written directly instead of through set_dist's index arithmetic
        double* out = jobs[g].distmatrix->values.data();
        size_t q = 0;
        for (uint32_t i = 0; i < idx.size(); ++i)
          for (uint32_t j = i + 1; j < idx.size(); ++j, ++q) {
            const size_t lx = reads[idx[i]].seq.size(), ly = reads[idx[j]].seq.size();
            const double largest = lx < ly ? (double)ly : (double)lx;  // :68
            // equal strings score 0, and 0/largest == 0.0 == the :64 shortcut
            out[q] = sc[q] / largest;                                  // :71-72
          }
      }
      unpack_s += now_s() - t0;
    };
    struct Hooks {
      decltype(wait_reads)* wait;
      decltype(unpack)* done;
      const Staging* S;
      size_t cursor;  // first read whose end lies beyond the last byte asked for
    } H{&wait_reads, &unpack, &S, 0};
    otgpu_stream_hooks hooks;
    hooks.user = &H;
    hooks.wait_bases = [](void* u, uint64_t upto) {
      Hooks* h = static_cast<Hooks*>(u);
      while (h->cursor < h->S->off.size() && h->S->off[h->cursor] < upto) ++h->cursor;
      (*h->wait)(h->cursor);
    };
    hooks.scores_ready = [](void* u, uint64_t first, uint64_t n) {
      Hooks* h = static_cast<Hooks*>(u);
      (*h->done)(first, n);
    };
    int st;
    double t_dev0, t_dev1;
    // a small batch is copied inline; a large one by the helpers while this thread drives the device
    if (workers == 1) copy(0);
    else pool.emplace_back(copy, 0);
    {
      std::lock_guard<otb200::DeviceTurn> turn(otb200::device_turn(aligner.device()));
      t_dev0 = now_s();
      st = otgpu_region_edit_scores_hooked(ctx, S.bases, S.used, S.off.data(), S.len.data(),
                                           (uint32_t)S.off.size(), region_start.data(),
                                           (uint32_t)jobs.size(), score_off.data(), n_scores,
                                           scores.data(), &hooks);
      t_dev1 = now_s();
    }
    if (st != OTGPU_OK) die("fill_dist_matrix", st, ctx);
    unpack(n_scores, 0);  // regions without pairs at the tail
    const double t_end = now_s();
    g_last_fill = FillTimes{t_end - t_begin, t_dev1 - t_dev0, stage_wait_s, unpack_s,
                            (double)otgpu_last_chunk_count(ctx)};
    return;
  }
  copy(0);
  wait_reads(n_seqs);
  // general case: some valid reads are not spanning -> per-pair plan, explicit pair list
  std::vector<otgpu_pair> pairs;
  std::vector<PairPlan> plans;
  pairs.reserve(n_scores);
  plans.reserve(n_scores);
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& idx = *jobs[g].indeces;
    auto& reads = *jobs[g].reads;
    for (uint32_t i = 0; i < idx.size(); ++i)
      for (uint32_t j = i + 1; j < idx.size(); ++j) {
        const PairPlan p = plan_pair(reads[idx[i]], reads[idx[j]]);
        const uint32_t rx = region_start[g] + i, ry = region_start[g] + j;
        plans.push_back(p);
        if (p.kind == PairPlan::kNone) continue;
        otgpu_pair pr;
        pr.pattern = p.pattern_is_y ? ry : rx;
        pr.text = p.pattern_is_y ? rx : ry;
        pr.pattern_begin_free = p.kind == PairPlan::kEndsFree ? p.pb : -1;
        pr.pattern_end_free = p.pe;
        pr.text_begin_free = pr.text_end_free = 0;
        pairs.push_back(pr);
      }
  }
  std::vector<int32_t> sc(pairs.size());
  int st;
  {
    std::lock_guard<otb200::DeviceTurn> turn(otb200::device_turn(aligner.device()));
    st = otgpu_edit_scores(ctx, S.bases, S.used, S.off.data(), S.len.data(),
                           (uint32_t)S.off.size(), pairs.data(), pairs.size(), sc.data());
  }
  if (st != OTGPU_OK) die("fill_dist_matrix", st, ctx);
  size_t q = 0, a = 0;
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& idx = *jobs[g].indeces;
    auto& reads = *jobs[g].reads;
    for (uint32_t i = 0; i < idx.size(); ++i)
      for (uint32_t j = i + 1; j < idx.size(); ++j, ++q) {
        double d;
        if (reads[idx[i]].seq == reads[idx[j]].seq) {
          d = 0.0;  // :64 comes first, whatever the spanning flags say
          if (plans[q].kind != PairPlan::kNone) ++a;
        } else if (plans[q].kind == PairPlan::kNone) {
          d = -1.0;
        } else {
          d = sc[a++] / plans[q].denom;
        }
        jobs[g].distmatrix->set_dist(i, j, d);
      }
  }
}

// ---------------------------------------------------------------------------------------------
void invalid_reassignment(const bool& ignore_haps, const double& min_sim, const double& max_error,
                          const int& total_alleles, std::vector<ANREAD>& reads,
                          wfa::WFAligner& aligner, std::vector<int>& labels) {
  std::vector<RegionReassignJob> one{{&reads, &labels, total_alleles}};
  invalid_reassignments(ignore_haps, min_sim, max_error, aligner, one);
}

void invalid_reassignments(const bool& ignore_haps, const double& min_sim,
                           const double& max_error, wfa::WFAligner& aligner,
                           std::vector<RegionReassignJob>& jobs) {
  (void)ignore_haps;  // the reference always aligns here: get_dist_anreads(true, ...) (:144)
  otgpu_ctx* ctx = aligner.context();
  if (!ctx) die("invalid_reassignment", OTGPU_ERR_NO_DEVICE, nullptr);
  Staging& S = staging(ctx);
  size_t total = 0;
  for (auto& J : jobs)
    for (auto& r : *J.reads) total += r.seq.size();
  S.reserve(total + 64);
  S.clear();
  std::vector<uint32_t> base(jobs.size());
  for (size_t g = 0; g < jobs.size(); ++g) {
    base[g] = (uint32_t)S.off.size();
    for (auto& r : *jobs[g].reads) S.add(r.seq);
  }
  // pass 1: every alignment the reference could run: unlabeled read i against every spanning
  // read j.  The reference walks reads in order and updates labels in place (:131-176), so a
  // spanning read labelled earlier in the loop becomes a target for later reads (possible with
  // --haps, where spanning reads without haplotag start unlabeled); aligning against ALL spanning
  // reads up front covers that, and pass 2 replays the reference's loop on the live labels.
  std::vector<otgpu_pair> pairs;
  std::vector<PairPlan> plans;             // one per candidate (i,j)
  std::vector<int64_t> plan_score;         // index into sc[] or -1
  std::vector<std::vector<int64_t>> table(jobs.size());  // per region: [u * n + j] -> plan index
  std::vector<std::vector<int>> unl(jobs.size());
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& reads = *jobs[g].reads;
    auto& labels = *jobs[g].labels;
    const int n = (int)labels.size();
    for (int i = 0; i < n; ++i)
      if (labels[i] < 0) unl[g].push_back(i);
    table[g].assign(unl[g].size() * (size_t)n, -1);
    for (size_t u = 0; u < unl[g].size(); ++u) {
      const int i = unl[g][u];
      for (int j = 0; j < n; ++j) {
        if (i == j || !reads[j].is_spanning()) continue;
        PairPlan p = plan_pair(reads[i], reads[j]);
        if (reads[i].seq == reads[j].seq) p.kind = PairPlan::kNone, p.denom = 0.0;  // :64 -> 0.0
        table[g][u * n + j] = (int64_t)plans.size();
        plans.push_back(p);
        if (p.kind == PairPlan::kNone) {
          plan_score.push_back(-1);
          continue;
        }
        plan_score.push_back((int64_t)pairs.size());
        otgpu_pair pr;
        pr.pattern = base[g] + (p.pattern_is_y ? j : i);
        pr.text = base[g] + (p.pattern_is_y ? i : j);
        pr.pattern_begin_free = p.kind == PairPlan::kEndsFree ? p.pb : -1;
        pr.pattern_end_free = p.pe;
        pr.text_begin_free = pr.text_end_free = 0;
        pairs.push_back(pr);
      }
    }
  }
  std::vector<int32_t> sc(pairs.size());
  if (!pairs.empty()) {
    std::unique_lock<otb200::DeviceTurn> turn(otb200::device_turn(aligner.device()));
    int st = otgpu_edit_scores(ctx, S.bases, S.used, S.off.data(), S.len.data(),
                               (uint32_t)S.off.size(), pairs.data(), pairs.size(), sc.data());
    turn.unlock();
    if (st != OTGPU_OK) die("invalid_reassignment", st, ctx);
  }
  // pass 2: the reference's loop (:131-176) on the live labels, scores looked up from pass 1
  for (size_t g = 0; g < jobs.size(); ++g) {
    auto& reads = *jobs[g].reads;
    auto& labels = *jobs[g].labels;
    const int total_alleles = jobs[g].total_alleles;
    const int n = (int)labels.size();
    for (size_t u = 0; u < unl[g].size(); ++u) {
      const int i = unl[g][u];
      std::vector<double> max_sim(total_alleles, 0.0);
      for (int j = 0; j < n; ++j) {
        if (i == j || labels[j] < 0 || !reads[j].is_spanning()) continue;
        const int64_t pi = table[g][u * n + j];
        const PairPlan& p = plans[pi];
        double dist;
        if (p.kind == PairPlan::kNone) dist = (p.denom == 0.0) ? 0.0 : -1.0;
        else dist = sc[plan_score[pi]] / p.denom;
        if (dist < 0) {  // :145-153
          std::cerr << "ERROR: unexpected distance for the following alignment:\n";
          std::cerr << reads[i].name << '\t' << reads[i].is_spanning() << '\n';
          std::cerr << reads[i].seq << '\n';
          std::cerr << reads[j].name << '\t' << reads[j].is_spanning() << '\n';
          std::cerr << reads[j].seq << std::endl;
          exit(1);
        }
        const double sim = 1 - dist;
        if (sim > max_sim[labels[j]]) max_sim[labels[j]] = sim;
      }
      int max_sim_label = 0;
      for (int j = 1; j < total_alleles; ++j)
        if (max_sim[j] > max_sim[max_sim_label]) max_sim_label = j;
      int same_max_sim = 0;
      for (const auto& s : max_sim)
        if (s == max_sim[max_sim_label]) ++same_max_sim;
      if (same_max_sim == 1 && max_sim[max_sim_label] >= min_sim) {
        double min_diff = 1.0;
        for (int j = 0; j < total_alleles; ++j)
          if (max_sim_label != j) {
            const double diff = max_sim[max_sim_label] - max_sim[j];
            if (diff < min_diff) min_diff = diff;
          }
        if (min_diff >= max_error) labels[i] = max_sim_label;
      }
    }
  }
}

FillTimes last_fill_times() { return g_last_fill; }
