// integration/assemble_b200.cpp — `otter assemble` on the B200 hot path: assemble_process as waves.
//
// A maintainer adds this file to otter's src/ and applies integration/assemble_waves.patch (six
// lines in src/assemble.cpp that route assemble() here when built with -DOTTER_B200); everything
// else in otter is compiled as it is, against otter_b200/host/shim/bindings/cpp/WFAligner.hpp
// instead of WFA2-lib's header, and linked with otter_b200/host/*_b200.cpp + libotter_b200.so
// in place of src/analignments.cpp + libwfacpp (integration/Makefile, target `binaries`).
//
// The reference (src/assemble.cpp:39-158) takes one region at a time per host thread:
//   parse_anreads -> local_realignment -> fill_dist_matrix -> otter_hclust ->
//   invalid_reassignment -> rapid_consensus -> print.
// A region's alignments only pay off on the GPU in batches, so the same per-region steps, in the
// same order, with the same parameters, messages and output, are run here over WAVES of regions:
//   device 1   [host threads] parse_anreads, coverage check, reference flanks
//              [GPU] local_realignments of the wave; every region's distance matrix in one batch
//                    (fill_dist_matrices); every region's density estimate (hclust_densities)
//   host 1     [host threads] otter_hclust's decision per region (hclust_finish), labels
//   device 2   [GPU] invalid_reassignments + rapid_consensus_alignments of the wave
//   host 2     [host threads] rapid_consensus_finish (coverage, se, partial-order consensus);
//              then the wave's alleles are printed in BED order
// with consecutive waves overlapped by otb200::run_waves (wave_b200.hpp).  Host stages use otter's
// own BS::thread_pool, one BAM / FASTA handle per pool thread like the reference.
#include "bindings/cpp/WFAligner.hpp"
#include "assemble.hpp"
#include "antimestamp.hpp"
#include "anbed.hpp"
#include "anbamfilehelper.hpp"
#include "otter_opts.hpp"
#include "BS_thread_pool.hpp"
#include "anseqs.hpp"
#include "anfahelper.hpp"
#include "analignments_b200.hpp"
#include "andistmat.hpp"
#include "otterclust_b200.hpp"
#include "wave_b200.hpp"

#include <cstdlib>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

// defined in src/assemble.cpp (:20-37)
uint32_t count_spanning_reads(const std::vector<ANREAD>& anread_block);
void partition_valid_reads(bool& ignore_haps, std::vector<ANREAD>& anread_block,
                           std::vector<int>& valid_indeces, std::vector<int>& invalid_indeces);

namespace {

struct RegionWork {
  enum State { kSkip, kReadsOnly, kCluster } state = kSkip;
  BED mod_bed;
  std::vector<ANREAD> reads;
  std::string ref_left, ref_right;  // reference flanks for local_realignment (fetched when needed)
  bool local_ignore_haps = true;
  std::vector<int> valid, invalid, labels;
  std::unique_ptr<DistMatrix> dm;
  ClusteringStatus clust;
  std::vector<double> density;
  double bandwidth = 0.0;
  ConsensusJob cjob;
  std::vector<ANALLELE> alleles;
};

// one BAM handle and one FASTA handle per host thread, opened on first use (src/assemble.cpp:45-48)
struct ThreadFiles {
  BamInstance bam;
  FaidxInstance fa;
  bool bam_open = false, fa_open = false;
  ~ThreadFiles() {
    if (bam_open) bam.destroy();
    if (fa_open) fa.destroy();
  }
};
ThreadFiles& thread_files(const std::string& bam, const std::string& reference) {
  thread_local ThreadFiles f;
  if (!f.bam_open) { f.bam.init(bam, true); f.bam_open = true; }
  if (!reference.empty() && !f.fa_open) { f.fa.init(reference); f.fa_open = true; }
  return f;
}

}  // namespace

void assemble_process_b200(const OtterOpts& params, const std::string& bam,
                           const std::vector<BED>& bed_regions, const std::string& reference,
                           bool& reads_only, BS::thread_pool& pool) {
  std::cerr << '(' << antimestamp() << "): Processing " << bam << " (" << params.read_group << ')' << std::endl;
  std::mutex std_out_mtx;
  const uint32_t n_regions = (uint32_t)bed_regions.size();
  std::vector<RegionWork> W(n_regions);
  // the issuing thread's aligners (src/assemble.cpp:49-50): all batched device work goes through them
  wfa::WFAlignerEdit aligner(wfa::WFAligner::Score, wfa::WFAligner::MemoryMed);
  wfa::WFAlignerGapAffine aligner2(4, 6, 2, wfa::WFAligner::Alignment, wfa::WFAligner::MemoryMed);
  auto blocks = [&](uint32_t a, uint32_t b, auto&& f) {
    pool.parallelize_loop(a, b, [&](const int x, const int y) {
      for (int g = x; g < y; ++g) f((uint32_t)g);
    }).wait();
  };

  // ---- per-region host work before the first device stage (src/assemble.cpp:52-124) ----
  auto prepare = [&](uint32_t g) {
    RegionWork& w = W[g];
    const BED& local_bed = bed_regions[g];
    w.mod_bed = local_bed;
    w.mod_bed.start -= (int)params.offset_l;
    w.mod_bed.end += (int)params.offset_r;
    ThreadFiles& files = thread_files(bam, reference);
    if (params.is_debug) {
      std::lock_guard<std::mutex> lk(std_out_mtx);
      std::cerr << '(' << antimestamp() << "): [DEBUG] Processing " << local_bed.toScString() << std::endl;
    }
    parse_anreads(params, w.mod_bed, files.bam, w.reads);
    if (params.is_debug) {
      std::lock_guard<std::mutex> lk(std_out_mtx);
      std::cerr << '(' << antimestamp() << "): [DEBUG] Loaded " << w.reads.size() << " reads" << std::endl;
    }
    if ((int)w.reads.size() > params.max_cov) {
      std::cerr << '(' << antimestamp() << "): [WARNING] Skipping region with abnormal coverage: "
                << local_bed.toScString() << " (" << w.reads.size() << ")" << std::endl;
      w.reads.clear();
      return;
    }
    w.state = RegionWork::kReadsOnly;  // (refined after the realignment)
    if (!reference.empty()) {  // the flanks local_realignment would fetch (src/analignments.cpp:22,28)
      bool need_l = false, need_r = false;
      for (const ANREAD& r : w.reads) {
        if (r.is_spanning() || !(r.is_spanning_l || r.is_spanning_r)) continue;
        if (r.is_spanning_r && r.ccoords.first >= params.flank) need_l = true;
        else if (r.is_spanning_l && (int)r.seq.size() - r.ccoords.second >= params.flank) need_r = true;
      }
      if (need_l) files.fa.fetch(w.mod_bed.chr, w.mod_bed.start - params.flank, w.mod_bed.start, w.ref_left);
      if (need_r) files.fa.fetch(w.mod_bed.chr, w.mod_bed.end, w.mod_bed.end + params.flank, w.ref_right);
    }
  };
  // ---- after the realignment: which reads are valid (src/assemble.cpp:88-124) ----
  auto partition = [&](uint32_t g) {
    RegionWork& w = W[g];
    if (w.state == RegionWork::kSkip || reads_only) return;
    const BED& local_bed = bed_regions[g];
    w.state = RegionWork::kSkip;
    const uint32_t spanning_reads = count_spanning_reads(w.reads);
    if (spanning_reads == 0) {
      std::lock_guard<std::mutex> lk(std_out_mtx);
      std::cerr << '(' << antimestamp() << "): [WARNING] No spanning reads for " << local_bed.toScString() << std::endl;
      return;
    }
    w.local_ignore_haps = params.ignore_haps;
    partition_valid_reads(w.local_ignore_haps, w.reads, w.valid, w.invalid);
    if (w.valid.size() < 2) {  // not enough haplotagged reads: 'ignore-haps' mode for this region
      w.local_ignore_haps = true;
      w.valid.clear();
      w.invalid.clear();
      partition_valid_reads(w.local_ignore_haps, w.reads, w.valid, w.invalid);
      if (spanning_reads != w.valid.size()) {
        std::lock_guard<std::mutex> lk(std_out_mtx);
        std::cerr << '(' << antimestamp() << "): [ERROR] Unexpected number of valid reads after switching to 'ignore-haps' mode: "
                  << spanning_reads << " vs " << w.valid.size() << std::endl;
        exit(1);
      }
    }
    if (w.valid.empty()) {
      std::lock_guard<std::mutex> lk(std_out_mtx);
      std::cerr << '(' << antimestamp() << "): [WARNING] No spanning reads for " << local_bed.toScString() << std::endl;
      return;
    }
    w.dm.reset(new DistMatrix(w.valid.size()));
    w.labels.assign(w.reads.size(), -1);
    w.state = RegionWork::kCluster;
  };

  otb200::WaveStages st;
  st.device1 = [&](uint32_t a, uint32_t b) {
    blocks(a, b, prepare);
    if (!reference.empty()) {  // local_realignment of the whole wave in one launch
      std::vector<RealignJob> rj;
      for (uint32_t g = a; g < b; ++g)
        if (W[g].state != RegionWork::kSkip && (!W[g].ref_left.empty() || !W[g].ref_right.empty()))
          rj.push_back({params.flank, params.min_sim, &W[g].ref_left, &W[g].ref_right, &W[g].reads});
      if (!rj.empty()) local_realignments(aligner2, rj);
      if (params.is_debug)
        for (uint32_t g = a; g < b; ++g)
          if (W[g].state != RegionWork::kSkip)
            std::cerr << '(' << antimestamp() << "): [DEBUG] Locally realigned " << W[g].reads.size() << " reads" << std::endl;
    }
    if (reads_only) return;
    blocks(a, b, partition);
    if (params.max_alleles != 1) {  // every distance matrix of the wave (src/assemble.cpp:127)
      std::vector<RegionDistJob> on, off;  // by the region's ignore-haps mode
      for (uint32_t g = a; g < b; ++g)
        if (W[g].state == RegionWork::kCluster)
          (W[g].local_ignore_haps ? on : off).push_back({&W[g].reads, &W[g].valid, W[g].dm.get()});
      if (!on.empty()) fill_dist_matrices(true, aligner, on);
      if (!off.empty()) fill_dist_matrices(false, aligner, off);
    }
    // otter_hclust's density estimates for the wave (src/otterclust.cpp:20-57 via KDE::f)
    std::vector<HclustJob> jobs;
    std::vector<uint32_t> owner;
    for (uint32_t g = a; g < b; ++g)
      if (W[g].state == RegionWork::kCluster) {
        jobs.push_back({&W[g].valid, W[g].dm.get(), &W[g].reads, &W[g].clust});
        owner.push_back(g);
      }
    std::vector<std::vector<double>> d;
    std::vector<double> h;
    hclust_densities(params.max_alleles, params.bandwidth_short, params.bandwidth_length,
                     params.bandwidth_long, aligner, jobs, d, h);
    for (size_t i = 0; i < owner.size(); ++i) {
      W[owner[i]].density = std::move(d[i]);
      W[owner[i]].bandwidth = h[i];
    }
  };
  st.host1 = [&](uint32_t a, uint32_t b) {  // src/assemble.cpp:129-134
    if (reads_only) return;
    blocks(a, b, [&](uint32_t g) {
      RegionWork& w = W[g];
      if (w.state != RegionWork::kCluster) return;
      HclustJob job{&w.valid, w.dm.get(), &w.reads, &w.clust};
      hclust_finish(params.max_alleles, params.max_error, params.min_cov_fraction,
                    params.min_cov_fraction2_l, params.min_cov_fraction2_f, job, w.density, w.bandwidth);
      for (uint32_t i = 0; i < w.clust.labels.size(); ++i) w.labels[w.valid[i]] = w.clust.labels[i];
      w.alleles.resize(w.clust.fc);
    });
  };
  st.device2 = [&](uint32_t a, uint32_t b) {  // src/assemble.cpp:136-142
    if (reads_only) return;
    for (int mode = 0; mode < 2; ++mode) {  // regions in 'ignore-haps' mode, then the others
      const bool ig = mode == 0;
      std::vector<RegionReassignJob> rjobs;
      for (uint32_t g = a; g < b; ++g)
        if (W[g].state == RegionWork::kCluster && W[g].local_ignore_haps == ig && !W[g].invalid.empty())
          rjobs.push_back({&W[g].reads, &W[g].labels, W[g].clust.fc});
      if (!rjobs.empty()) invalid_reassignments(ig, params.min_sim, params.max_error, aligner, rjobs);
    }
    std::vector<ConsensusJob> batch;
    std::vector<uint32_t> owner;
    for (uint32_t g = a; g < b; ++g) {
      RegionWork& w = W[g];
      if (w.state != RegionWork::kCluster) continue;
      ConsensusJob j;
      j.reads = &w.reads;
      j.labels = &w.labels;
      j.valid_indeces = &w.valid;
      j.total_alleles = w.clust.fc;
      j.valid_distmatrix = w.dm.get();
      batch.push_back(std::move(j));
      owner.push_back(g);
    }
    rapid_consensus_alignments(aligner2, batch);
    for (size_t i = 0; i < batch.size(); ++i) W[owner[i]].cjob = std::move(batch[i]);
  };
  st.host2 = [&](uint32_t a, uint32_t b) {
    if (!reads_only)
      blocks(a, b, [&](uint32_t g) {
        RegionWork& w = W[g];
        if (w.state == RegionWork::kCluster) rapid_consensus_finish(w.local_ignore_haps, w.cjob, w.alleles);
      });
    std::lock_guard<std::mutex> lk(std_out_mtx);  // the wave's output, in BED order
    for (uint32_t g = a; g < b; ++g) {
      RegionWork& w = W[g];
      const BED& local_bed = bed_regions[g];
      if (reads_only) {
        if (w.state != RegionWork::kSkip)
          for (const auto& read : w.reads) {
            if (params.is_fa) read.stdout_fa(local_bed.toScString());
            else read.stdout_sam(local_bed.chr, local_bed.start, local_bed.end, params.read_group);
          }
      } else if (w.state == RegionWork::kCluster) {
        for (int l = 0; l < w.clust.fc; ++l) {
          w.alleles[l].ic = w.clust.ic;
          if (params.is_fa) w.alleles[l].stdout_fa(params.read_group, local_bed.toScString() + '#' + std::to_string(l));
          else w.alleles[l].stdout_sam(local_bed.toScString() + "_" + std::to_string(l), local_bed.chr,
                                       local_bed.start, local_bed.end, params.read_group);
        }
      }
      RegionWork().reads.swap(w.reads);  // the region is done: give its memory back
      w.alleles.clear();
      w.dm.reset();
      w.cjob = ConsensusJob();
    }
  };
  uint32_t wave = 256;
  if (const char* e = getenv("OTTER_B200_WAVE")) wave = (uint32_t)std::max(1, atoi(e));
  otb200::run_waves(n_regions, wave, st, 1);
}
