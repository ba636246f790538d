// otter_b200/host/wfaligner_b200.cpp — implementation of the wfa::WFAligner facade
// (shim/bindings/cpp/WFAligner.hpp) on top of the C ABI.  One pair per call.
#include <cstdint>
#include <cstring>
#include <mutex>
#include <vector>

#include "bindings/cpp/WFAligner.hpp"
#include "device_turn.hpp"
#include "otter_b200.h"

namespace wfa {

static int g_default_device = 0;
void WFAligner::setDefaultDevice(int device) { g_default_device = device; }

WFAligner::WFAligner(Kind kind, AlignmentScope scope, int mismatch, int gapOpening,
                     int gapExtension)
    : kind_(kind), scope_(scope), x_(mismatch), o_(gapOpening), e_(gapExtension) {
  // No device -> ctx_ stays null and every align call reports StatusDeviceError: there is no CPU
  // implementation to fall back to.
  device_ = g_default_device;
  if (otgpu_create(device_, &ctx_) != OTGPU_OK) ctx_ = nullptr;
}

WFAligner::~WFAligner() { otgpu_destroy(ctx_); }

WFAligner::AlignmentStatus WFAligner::run(const char* p, int plen, const char* t, int tlen,
                                          bool endsfree, int pb, int pe, int tb, int te) {
  cigar_.clear();
  if (!ctx_) { status_ = StatusDeviceError; return StatusDeviceError; }
  std::vector<uint8_t> bases((size_t)plen + (size_t)tlen + 1);
  std::memcpy(bases.data(), p, plen);
  std::memcpy(bases.data() + plen, t, tlen);
  const uint64_t off[2] = {0, (uint64_t)plen};
  const uint32_t len[2] = {(uint32_t)plen, (uint32_t)tlen};
  otgpu_pair pr{0, 1, endsfree ? pb : -1, pe, tb, te};
  int32_t sc = 0;
  // one whole-GPU call at a time per device, also for single alignments (device_turn.hpp)
  std::lock_guard<otb200::DeviceTurn> turn(otb200::device_turn(device_));
  if (kind_ == kGapAffine || scope_ == Alignment) {
    // score + CIGAR in one launch: gap-affine (wfa_affine.cuh) or, for the aligner `compare`
    // builds (reference src/compare.cpp:95), edit distance with CIGAR (wfa_edit_cigar.cuh)
    std::vector<char> cig((size_t)plen + tlen + 1);
    const uint64_t coff[2] = {0, (uint64_t)plen + tlen};
    int32_t clen = 0;
    const int st = kind_ == kGapAffine
        ? otgpu_affine_cigars(ctx_, bases.data(), (uint64_t)plen + tlen, off, len, 2, &pr, 1, x_, o_,
                              e_, &sc, cig.data(), coff, &clen)
        : otgpu_edit_cigars(ctx_, bases.data(), (uint64_t)plen + tlen, off, len, 2, &pr, 1, &sc,
                            cig.data(), coff, &clen);
    if (st != OTGPU_OK) { status_ = StatusDeviceError; return StatusDeviceError; }
    if (clen == OTGPU_SCORE_UNFEASIBLE) { status_ = StatusUnfeasible; return StatusUnfeasible; }
    if (clen < 0) { status_ = StatusOOM; return StatusOOM; }
    score_ = sc;
    if (scope_ == Alignment) cigar_.assign(cig.data(), (size_t)clen);
    status_ = StatusSuccessful;
    return StatusSuccessful;
  }
  const int st = otgpu_edit_scores(ctx_, bases.data(), (uint64_t)plen + tlen, off, len, 2, &pr, 1, &sc);
  if (st != OTGPU_OK) { status_ = StatusDeviceError; return StatusDeviceError; }
  if (sc < 0) { status_ = StatusUnfeasible; return StatusUnfeasible; }  // score_ keeps its old value
  score_ = sc;
  status_ = StatusSuccessful;
  return StatusSuccessful;
}

WFAligner::AlignmentStatus WFAligner::alignEnd2End(const char* pattern, int patternLength,
                                                   const char* text, int textLength) {
  return run(pattern, patternLength, text, textLength, false, 0, 0, 0, 0);
}
WFAligner::AlignmentStatus WFAligner::alignEnd2End(std::string& pattern, std::string& text) {
  return run(pattern.data(), (int)pattern.size(), text.data(), (int)text.size(), false, 0, 0, 0, 0);
}
WFAligner::AlignmentStatus WFAligner::alignEndsFree(const char* pattern, int patternLength,
                                                    int patternBeginFree, int patternEndFree,
                                                    const char* text, int textLength,
                                                    int textBeginFree, int textEndFree) {
  return run(pattern, patternLength, text, textLength, true, patternBeginFree, patternEndFree,
             textBeginFree, textEndFree);
}
WFAligner::AlignmentStatus WFAligner::alignEndsFree(std::string& pattern, int patternBeginFree,
                                                    int patternEndFree, std::string& text,
                                                    int textBeginFree, int textEndFree) {
  return run(pattern.data(), (int)pattern.size(), text.data(), (int)text.size(), true,
             patternBeginFree, patternEndFree, textBeginFree, textEndFree);
}

}  // namespace wfa
