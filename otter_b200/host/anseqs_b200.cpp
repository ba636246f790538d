// otter_b200/host/anseqs_b200.cpp — see anseqs_b200.hpp.
#include "anseqs_b200.hpp"

#include <cstdlib>
#include <cstring>
#include <iostream>

#include <mutex>

#include "device_turn.hpp"
#include "otter_b200.h"

template <class T>
PinnedBuf<T>::~PinnedBuf() {
  if (p_) otgpu_host_free(p_);
}

template <class T>
void PinnedBuf<T>::append(const T* src, size_t n) {
  if (size_ + n > cap_) {
    const size_t want = (size_ + n) * 2 + 4096;
    void* q = nullptr;
    const int st = otgpu_host_alloc(want * sizeof(T), &q);
    if (st != OTGPU_OK) {  // no device or no memory: there is no CPU path to fall back to
      std::cerr << "ERROR: otter_b200 read extraction: pinned allocation: " << otgpu_strerror(st) << std::endl;
      exit(1);
    }
    if (size_) std::memcpy(q, p_, size_ * sizeof(T));
    if (p_) otgpu_host_free(p_);
    p_ = static_cast<T*>(q);
    cap_ = want;
  }
  if (n) std::memcpy(p_ + size_, src, n * sizeof(T));
  size_ += n;
}
template class PinnedBuf<uint32_t>;
template class PinnedBuf<uint8_t>;

void ReadExtractor::add(uint32_t region, int start, int end, const BamRecordView& rec) {
  Rec r;
  r.pos = rec.pos;
  r.l_qseq = rec.l_qseq;
  r.n_cigar = rec.n_cigar;
  r.region = region;
  r.cigar_off = cigar_.size();
  r.seq_off = seq4_.size();
  r.start = start;
  r.end = end;
  cigar_.append(rec.cigar, rec.n_cigar);
  const size_t nb = ((size_t)(rec.l_qseq > 0 ? rec.l_qseq : 0) + 1) / 2;
  seq4_.append(rec.seq4, nb);
  recs_.push_back(r);
  names_.emplace_back(rec.qname ? rec.qname : "");
}

void ReadExtractor::extract(wfa::WFAligner& aligner, bool omitnonspanning,
                            std::vector<std::vector<ANREAD>>& blocks) {
  const uint32_t n = (uint32_t)recs_.size();
  if (n == 0) return;
  otgpu_ctx* ctx = aligner.context();
  std::vector<otgpu_bam_read> in(n);
  for (uint32_t i = 0; i < n; ++i) {
    const Rec& r = recs_[i];
    in[i] = otgpu_bam_read{r.pos, r.l_qseq, r.n_cigar, 0, r.cigar_off, r.seq_off, r.start, r.end};
  }
  std::vector<uint8_t> status(n);
  std::vector<int32_t> coords(2 * (size_t)n);
  std::vector<uint32_t> len(n);
  const char* bases = nullptr;
  const uint64_t* begin = nullptr;
  int st = OTGPU_ERR_NO_DEVICE;
  if (ctx) {
    std::lock_guard<otb200::DeviceTurn> turn(otb200::device_turn(aligner.device()));
    st = otgpu_extract_reads(ctx, in.data(), n, cigar_.data(), cigar_.size(), seq4_.data(), seq4_.size(),
                             status.data(), coords.data(), len.data(), &bases, &begin);
  }
  if (st != OTGPU_OK) {
    std::cerr << "ERROR: otter_b200 read extraction: " << otgpu_strerror(st) << " ("
              << (ctx ? otgpu_last_cuda_error(ctx) : "") << ")" << std::endl;
    exit(1);  // the reference's convention for fatal conditions
  }
  for (uint32_t i = 0; i < n; ++i) {
    const uint8_t s = status[i];
    if (s & OTGPU_READ_BAD_COORDS) {  // src/anseqs.cpp:418-421
      std::cerr << "ERROR: unexpected querty start/end coords found for read " << names_[i] << std::endl;
      exit(1);
    }
    if (!(s & OTGPU_READ_OK)) continue;
    const bool sl = (s & OTGPU_READ_SPAN_L) != 0, sr = (s & OTGPU_READ_SPAN_R) != 0;
    if (omitnonspanning && !(sl && sr)) continue;  // :451
    const uint32_t g = recs_[i].region;
    if (g >= blocks.size()) blocks.resize(g + 1);
    ANREAD read;
    read.name = names_[i];
    read.seq.assign(bases + begin[i], len[i]);
    if (sl && sr) read.set_is_spanning();  // transfer_status :233-239
    else if (sl) read.is_spanning_l = true;
    else if (sr) read.is_spanning_r = true;
    read.ccoords = {coords[2 * (size_t)i], coords[2 * (size_t)i + 1]};
    blocks[g].emplace_back(std::move(read));
  }
  recs_.clear();
  cigar_.clear();
  seq4_.clear();
  names_.clear();
}
