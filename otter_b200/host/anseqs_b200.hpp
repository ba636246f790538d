// otter_b200/host/anseqs_b200.hpp — batched read extraction on the device:
// the per-record half of parse_anreads (reference src/anseqs.cpp:438-460), i.e. parse_alignment
// (:412-436) + PARSEMSG::transfer_status (:233-239), for every record of many regions in one
// launch.  BAM iteration, the record-level filters (mapq, secondary/supplementary, :444) and the
// aux tags (parse_standard_auxs, rq filter :453-454) stay with the caller, which owns htslib.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "bindings/cpp/WFAligner.hpp"
#include "otter_types.hpp"

// One BAM record as the bam1_t accessors expose it: core.pos, core.l_qseq, core.n_cigar,
// bam_get_cigar(b), bam_get_seq(b), bam_get_qname(b).  The arrays are copied during the call
// (htslib reuses the record buffer on the next bam_itr_next).
struct BamRecordView {
  int32_t pos;
  int32_t l_qseq;
  uint32_t n_cigar;
  const uint32_t* cigar;
  const uint8_t* seq4;
  const char* qname;
};

// Grow-only array in pinned host memory (otgpu_host_alloc): the record data goes to the device
// at PCIe speed instead of through the driver's pageable staging.
template <class T>
class PinnedBuf {
 public:
  PinnedBuf() {}
  PinnedBuf(const PinnedBuf&) = delete;
  PinnedBuf& operator=(const PinnedBuf&) = delete;
  ~PinnedBuf();
  void append(const T* src, size_t n);
  void clear() { size_ = 0; }
  size_t size() const { return size_; }
  const T* data() const { return p_; }

 private:
  T* p_ = nullptr;
  size_t size_ = 0, cap_ = 0;
};

// Accumulates the records of a wave of regions, then extracts them all at once.
class ReadExtractor {
 public:
  // parse_anreads's loop body up to parse_alignment: remember one record of region `region`
  // (an index the caller chooses) with bounds [start, end] = (bed.start, bed.end).
  void add(uint32_t region, int start, int end, const BamRecordView& rec);
  size_t size() const { return names_.size(); }
  // Run get_breakpoints + parse_alignment for every remembered record on the device behind
  // `aligner` and append, per region and in the order added, the reads the reference would have
  // kept (msg.successful && (!omitnonspanning || spanning), :451) to blocks[region].
  // Clears the accumulator.
  void extract(wfa::WFAligner& aligner, bool omitnonspanning, std::vector<std::vector<ANREAD>>& blocks);

 private:
  struct Rec { int32_t pos, l_qseq; uint32_t n_cigar, region; uint64_t cigar_off, seq_off; int32_t start, end; };
  std::vector<Rec> recs_;
  PinnedBuf<uint32_t> cigar_;
  PinnedBuf<uint8_t> seq4_;
  std::vector<std::string> names_;
};
