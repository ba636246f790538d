// otter_b200/host/shim/bindings/cpp/WFAligner.hpp — drop-in for the header otter includes as
// "bindings/cpp/WFAligner.hpp" (reference include/WFA2-lib/bindings/cpp/WFAligner.hpp:43-235).
//
// Same namespace, class names, enums and the member functions otter calls
// (alignEnd2End, alignEndsFree, getAlignmentScore, getAlignmentCigar, getAlignmentStatus —
// call sites: reference src/analignments.cpp:24,30,35,70,87-96,264-276; src/assemble.cpp:49-50;
// src/compare.cpp:55-60,95), implemented on the GPU through include/otter_b200.h.  Put this
// directory on the include path instead of include/WFA2-lib and otter's sources compile
// unchanged.  The facade aligns ONE pair per call (one kernel launch): correct, but the speed
// comes from the batched entry points in analignments_b200.hpp that share one launch across
// thousands of regions.  There is no CPU implementation behind this class.
#pragma once
#include <string>

struct otgpu_ctx;

namespace wfa {

class WFAligner {
 public:
  enum MemoryModel { MemoryHigh, MemoryMed, MemoryLow, MemoryUltralow };
  enum AlignmentScope { Score, Alignment };
  enum AlignmentStatus {
    StatusSuccessful = 0,         // WF_STATUS_SUCCESSFUL
    StatusUnfeasible = -1,        // WF_STATUS_UNFEASIBLE
    StatusMaxScoreReached = -2,   // WF_STATUS_MAX_SCORE_REACHED
    StatusOOM = -3,               // WF_STATUS_OOM
    StatusDeviceError = -100,     // CUDA failure (no reference equivalent; there is no fallback)
    StatusNotOnDevice = -101      // reserved: an aligner flavour without a device kernel (none now)
  };
  AlignmentStatus alignEnd2End(const char* pattern, int patternLength, const char* text,
                               int textLength);
  AlignmentStatus alignEnd2End(std::string& pattern, std::string& text);
  AlignmentStatus alignEndsFree(const char* pattern, int patternLength, int patternBeginFree,
                                int patternEndFree, const char* text, int textLength,
                                int textBeginFree, int textEndFree);
  AlignmentStatus alignEndsFree(std::string& pattern, int patternBeginFree, int patternEndFree,
                                std::string& text, int textBeginFree, int textEndFree);
  int getAlignmentScore() { return score_; }
  int getAlignmentStatus() { return status_; }
  std::string getAlignmentCigar() { return cigar_; }
  // B200 additions (not in the reference): the device context behind this aligner, shared by
  // the batched entry points; which CUDA device new aligners bind to (default 0).
  otgpu_ctx* context() { return ctx_; }
  int device() const { return device_; }
  bool isEdit() const { return kind_ == kEdit; }
  int mismatch() const { return x_; }
  int gapOpening() const { return o_; }
  int gapExtension() const { return e_; }
  static void setDefaultDevice(int device);

 protected:
  enum Kind { kEdit, kGapAffine };
  WFAligner(Kind kind, AlignmentScope scope, int mismatch, int gapOpening, int gapExtension);
  ~WFAligner();
  AlignmentStatus run(const char* p, int plen, const char* t, int tlen, bool endsfree, int pb,
                      int pe, int tb, int te);

 private:
  WFAligner(const WFAligner&);
  Kind kind_;
  AlignmentScope scope_;
  int x_, o_, e_;
  otgpu_ctx* ctx_ = nullptr;
  int device_ = 0;
  int score_ = 0;
  int status_ = 0;
  std::string cigar_;
};

class WFAlignerEdit : public WFAligner {
 public:
  WFAlignerEdit(const AlignmentScope alignmentScope, const MemoryModel memoryModel = MemoryHigh)
      : WFAligner(kEdit, alignmentScope, 0, 0, 0) { (void)memoryModel; }
};

class WFAlignerGapAffine : public WFAligner {
 public:
  WFAlignerGapAffine(const int mismatch, const int gapOpening, const int gapExtension,
                     const AlignmentScope alignmentScope,
                     const MemoryModel memoryModel = MemoryHigh)
      : WFAligner(kGapAffine, alignmentScope, mismatch, gapOpening, gapExtension) {
    (void)memoryModel;
  }
};

}  // namespace wfa
