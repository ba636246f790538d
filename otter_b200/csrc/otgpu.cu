// otter_b200/csrc/otgpu.cu — C ABI of include/otter_b200.h + batch kernels (sm_100a).
//
// Pipeline of one batch on the context's stream (no host sync until download):
//   H2D(raw bases, read table, region table | pair list)
//   -> k_pack_reads      raw bytes -> 2-bit words, flags reads holding a non-ACGT byte
//   -> k_expand_regions  region table -> explicit (pattern,text) list in DistMatrix order,
//                        pattern = longer read (reference src/analignments.cpp:67-70)
//   -> k_align<1024|2048> persistent warps, one pair per warp, wavefront ring in shared memory
//   -> k_align<0>         the few pairs whose live width outgrew the ring, wavefront in HBM
//   D2H(int32 scores)
// There is no CPU path: every entry point needs a CUDA device.
#include "../../include/otter_b200.h"
#include "wfa_edit.cuh"
#include "wfa_affine.cuh"
#include "wfa_edit_cigar.cuh"
#include "bam_extract.cuh"
#include "kde.cuh"
#include "ppoa.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

using namespace otgpu;

namespace {

constexpr int kWarpsPerCta = 8;
constexpr int kCtaThreads = kWarpsPerCta * 32;
constexpr uint32_t kEmptySlot = 0xffffffffu;
constexpr uint32_t kLongReadLen = 8000;  // batches with longer reads use the 2048-slot ring

struct DevBatch {
  const uint8_t* bases;      // raw bytes, preceded and followed by kPadBases guard bytes
  const uint64_t* read_off;  // byte offset of each read in bases
  const uint32_t* read_len;
  const uint64_t* pk_off;    // word offset of each read in packed
  const uint32_t* packed;    // 2-bit words, 16 bases per word
  uint32_t* read_flag;       // bit0: read holds a byte outside {A,C,G,T}
  const otgpu_pair* pairs;
  int32_t* scores;
  uint64_t n_pairs;
  uint32_t n_reads;
};

// counters[0] main queue, [1] wide (overflow) count, [2] wide queue, [3] 8-bit pairs,
// [4] pairs re-run on an HBM ring inside the main kernel
struct Counters { unsigned long long v[8]; };

__global__ void __launch_bounds__(128) k_pack_reads(DevBatch b, uint32_t* __restrict__ packed) {
  const uint32_t r = blockIdx.x;
  const uint32_t len = b.read_len[r];
  const uint8_t* __restrict__ src = b.bases + kPadBases + b.read_off[r];
  uint32_t* __restrict__ dst = packed + b.pk_off[r];
  const uint32_t nwords = (len + 15) / 16 + 2;  // + guard words the window loads may touch
  bool bad = false;
  for (uint32_t w = threadIdx.x; w < nwords; w += blockDim.x) {
    uint32_t word = 0;
    const uint32_t p0 = w * 16;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const uint32_t p = p0 + i;
      if (p < len) {
        const uint32_t c = src[p];
        bad |= !(c == 'A' || c == 'C' || c == 'G' || c == 'T');
        word |= ((c >> 1) & 3u) << (2 * i);
      }
    }
    dst[w] = word;
  }
  if (bad) atomicOr(&b.read_flag[r], 1u);
}

__global__ void __launch_bounds__(128) k_expand_regions(const uint32_t* __restrict__ region_start,
                                                        const uint64_t* __restrict__ score_off,
                                                        const uint32_t* __restrict__ read_len,
                                                        otgpu_pair* __restrict__ pairs) {
  const uint32_t g = blockIdx.x;
  const uint32_t r0 = region_start[g];
  const uint32_t n = region_start[g + 1] - r0;
  otgpu_pair* out = pairs + score_off[g];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (uint32_t i = warp; i + 1 < n; i += nw) {
    const uint64_t row = (uint64_t)i * (2ull * n - i - 1) / 2;  // src/andistmat.cpp:20
    const uint32_t li = read_len[r0 + i];
    for (uint32_t j = i + 1 + lane; j < n; j += 32) {
      const uint32_t lj = read_len[r0 + j];
      otgpu_pair p;
      const bool x_is_smallest = li < lj;  // src/analignments.cpp:67
      p.pattern = x_is_smallest ? r0 + j : r0 + i;
      p.text = x_is_smallest ? r0 + i : r0 + j;
      p.pattern_begin_free = -1;
      p.pattern_end_free = p.text_begin_free = p.text_end_free = 0;
      out[row + (j - i - 1)] = p;
    }
  }
}

// ---- heaviest-first work order -------------------------------------------------------------------
// Pairs differ by 3-4 orders of magnitude in work.  The persistent warps
// pull pairs from one queue, so the pairs are queued by descending estimated work (64 log-scale
// buckets of |len difference| + len/16) to keep the tail of the launch short.
constexpr int kBuckets = 64;
struct OrderCtl { unsigned int hist[kBuckets]; unsigned int cursor[kBuckets]; };

__device__ __forceinline__ int work_bucket(const otgpu_pair& p, const uint32_t* __restrict__ read_len) {
  if (p.pattern == kEmptySlot) return 0;
  const uint32_t a = read_len[p.pattern], b = read_len[p.text];
  const uint32_t est = (a > b ? a - b : b - a) + (max(a, b) >> 4) + 1;
  const int lg = 31 - __clz(est);                       // 0..31
  const int half = lg > 0 ? (est >> (lg - 1)) & 1 : 0;  // one more bit of resolution
  return min(kBuckets - 1, 2 * lg + half);
}

__global__ void __launch_bounds__(256) k_order_count(const otgpu_pair* __restrict__ pairs, uint64_t n,
                                                     const uint32_t* __restrict__ read_len,
                                                     OrderCtl* __restrict__ ctl) {
  __shared__ unsigned int h[kBuckets];
  if (threadIdx.x < kBuckets) h[threadIdx.x] = 0;
  __syncthreads();
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    atomicAdd(&h[work_bucket(pairs[i], read_len)], 1u);
  __syncthreads();
  if (threadIdx.x < kBuckets && h[threadIdx.x]) atomicAdd(&ctl->hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void k_order_scan(OrderCtl* __restrict__ ctl) {  // heaviest bucket first
  unsigned int run = 0;
  for (int b = kBuckets - 1; b >= 0; --b) {
    ctl->cursor[b] = run;
    run += ctl->hist[b];
  }
}

__global__ void __launch_bounds__(256) k_order_scatter(const otgpu_pair* __restrict__ pairs, uint64_t n,
                                                       const uint32_t* __restrict__ read_len,
                                                       OrderCtl* __restrict__ ctl,
                                                       uint32_t* __restrict__ order) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    order[atomicAdd(&ctl->cursor[work_bucket(pairs[i], read_len)], 1u)] = (uint32_t)i;
}

__device__ __forceinline__ SeqView view2(const DevBatch& b, uint32_t r) {
  return SeqView{(uint32_t)b.pk_off[r], 0};
}
__device__ __forceinline__ SeqView view8(const DevBatch& b, uint32_t r) {
  const uint64_t o = b.read_off[r] + kPadBases;
  return SeqView{(uint32_t)(o >> 2), (int)(o & 3ull)};
}

// A pair whose wavefront outgrew the shared ring, started again on this warp's HBM ring without
// leaving the kernel (out of line, so that the hot kernel's register allocation is its own).
template <int BITS>
__device__ __noinline__ int wfa_edit_pair_hbm(int32_t* ws, int mask, const EditJob J, const int lane) {
  Ring<0> ring;
  ring.p = ws;
  ring.mask = mask;
  return wfa_edit_pair<BITS, 0>(ring, J, lane);
}

#ifndef OTGPU_ALIGN_MIN_CTAS
#define OTGPU_ALIGN_MIN_CTAS 4
#endif
// One pair per warp.  CAP > 0: pairs named by `list` (count in ctr->v[count_slot], or all
// b.n_pairs of them when count_slot < 0), wavefront ring of CAP slots in shared memory; a pair that
// outgrows it goes to ovf_list — except on the 2048-slot tier (batches of long reads, where such
// pairs actually occur), which starts the pair again at once on the warp's own HBM ring when the
// launch has one per warp (wide_ws != nullptr): the heaviest pairs come first in the queue, so
// their long single-warp runs overlap everything else instead of forming a tail.  (The call costs
// the 64-register 1024-slot kernel spills and 10 % of its speed, so that tier keeps two passes.)
// CAP == 0 ("wide"): the pairs of ovf_list, ring in HBM.
// CAP = 1024 keeps 4 CTAs (32 warps) per SM; 2048 (3 CTAs) is for batches of long reads, whose
// wavefronts pass a thousand diagonals.
constexpr size_t ring_bytes(int cap) { return (size_t)(kWarpsPerCta + 1) * cap * 4; }
// FW: first window of an extension (wfa_edit.cuh, first_window): 0 = 64 bits, 1 = 32 bits and
// 128-bit continuations, for batches of low-error reads.  Scores are the same; the host picks per
// launch from what the previous batch on the context looked like (ctr->v[5], below).
template <int CAP, int FW = 0>
__global__ void __launch_bounds__(kCtaThreads, CAP == 2048 ? 3 : OTGPU_ALIGN_MIN_CTAS)
k_align(DevBatch b, Counters* __restrict__ ctr, const uint32_t* __restrict__ list, int count_slot,
        int cursor_slot, uint32_t* __restrict__ ovf_list, int32_t* __restrict__ wide_ws,
        int wide_cap) {
  // one extra ring of slack: the rings are aligned to their size at run time (Ring<false>::addr
  // ORs the slot offset into the base), whatever offset the static allocation starts at
  constexpr bool WIDE = CAP == 0;
  extern __shared__ int32_t ring_smem[];  // ring_bytes(CAP) of dynamic shared memory
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Ring<CAP> ring;
  if constexpr (WIDE) {
    ring.p = wide_ws + (size_t)(blockIdx.x * kWarpsPerCta + warp) * wide_cap;
    ring.mask = wide_cap - 1;
  } else {
    const unsigned a0 = ((unsigned)__cvta_generic_to_shared(ring_smem) + CAP * 4 - 1) & ~(CAP * 4 - 1u);
    ring.base = a0 + warp * (CAP * 4);
  }
  const unsigned long long limit = WIDE ? ctr->v[1] : (count_slot < 0 ? b.n_pairs : ctr->v[count_slot]);
  unsigned fast = 0;  // pairs of this warp that advanced >= kFwRun bases per score step
  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(&ctr->v[WIDE ? 2 : cursor_slot], 1ull);
    q = __shfl_sync(kFull, q, 0);
    if (q >= limit) break;
    const uint64_t idx = WIDE ? ovf_list[q] : list[q];
    const otgpu_pair pr = b.pairs[idx];
    if (pr.pattern == kEmptySlot) {
      if (lane == 0) b.scores[idx] = 0;
      continue;
    }
    EditJob J;
    J.plen = (int)b.read_len[pr.pattern];
    J.tlen = (int)b.read_len[pr.text];
    J.pb = pr.pattern_begin_free; J.pe = pr.pattern_end_free;
    J.tb = pr.text_begin_free;    J.te = pr.text_end_free;
    const bool bytes = ((b.read_flag[pr.pattern] | b.read_flag[pr.text]) & 1u) != 0;
    int score;
    if (bytes) {
      J.words = reinterpret_cast<const uint32_t*>(b.bases);
      J.P = view8(b, pr.pattern); J.T = view8(b, pr.text);
      score = wfa_edit_pair<8, CAP, FW>(ring, J, lane);
      if (!WIDE && lane == 0) atomicAdd(&ctr->v[3], 1ull);
    } else {
      J.words = b.packed;
      J.P = view2(b, pr.pattern); J.T = view2(b, pr.text);
      score = wfa_edit_pair<2, CAP, FW>(ring, J, lane);
    }
    __syncwarp();
    if (CAP == 1024 && score >= 0 && max(J.plen, J.tlen) >= kFwRun * score) ++fast;
    if (CAP == 2048 && score == kScoreOverflow && wide_ws != nullptr) {
      int32_t* ws = wide_ws + (size_t)(blockIdx.x * kWarpsPerCta + warp) * wide_cap;
      score = bytes ? wfa_edit_pair_hbm<8>(ws, wide_cap - 1, J, lane)
                    : wfa_edit_pair_hbm<2>(ws, wide_cap - 1, J, lane);
      if (lane == 0) atomicAdd(&ctr->v[4], 1ull);
      __syncwarp();
    }
    if (lane == 0) {
      if (!WIDE && score == kScoreOverflow) {
        const unsigned long long pos = atomicAdd(&ctr->v[1], 1ull);
        ovf_list[pos] = (uint32_t)idx;
      } else {
        b.scores[idx] = score;
      }
    }
  }
  if (CAP == 1024 && lane == 0 && fast) atomicAdd(&ctr->v[5], (unsigned long long)fast);
}

// ------------------------------------------------------------------------------------------
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

}  // namespace

struct otgpu_ctx {
  int device = 0;
  int n_sm = 0;
  cudaStream_t stream = nullptr;
  char err[256] = {0};
  DevBuf bases, read_off, read_len, pk_off, read_flag, packed, region_start, score_off, pairs,
      scores, ovf_list, counters, wide_ws, order, order_ctl;
  DevBuf aff_rings, aff_slab, aff_slab2, aff_ops, cigar_off, cigars, cigar_len;  // gap-affine path
  DevBuf cigar_begin, cigars_packed;            // packed CIGAR output (device side)
  DevBuf bam_reads, bam_cigar, bam_seq4, bam_status, bam_coords, bam_plan;  // read extraction
  DevBuf ppoa_in, ppoa_ws, ppoa_out;              // partial-order consensus: inputs, graph pools, strings
  std::vector<char> h_ppoa;                       // packed consensus strings of the last call
  std::vector<uint64_t> h_ppoa_begin;
  DevBuf bam_out, bam_begin, bam_len, gather_index;  // extracted reads, kept on the device
  uint32_t extracted_reads = 0;              // records of the last otgpu_extract_reads call (0: none)
  std::vector<uint32_t> extracted_len;       // their lengths (host copy)
  DevBuf kde_values, kde_off, kde_grid, kde_h, kde_inv_h, kde_out;          // density estimates
  void* h_cigars = nullptr;                     // pinned: packed CIGARs of the last call
  size_t h_cigars_cap = 0;
  void* h_cigar_begin = nullptr;                // pinned: n_pairs + 1 start offsets
  size_t h_cigar_begin_cap = 0;
  void* h_reads = nullptr;                      // pinned: extracted reads of the last call
  size_t h_reads_cap = 0;
  void* h_read_begin = nullptr;                 // pinned: n_reads + 1 start offsets
  size_t h_read_begin_cap = 0;
  void* h_pk_off = nullptr;  // pinned
  size_t h_pk_off_cap = 0;
  void* h_counters = nullptr;  // pinned, 4 x u64
  // resident batch
  bool resident = false, by_regions = false;
  uint64_t n_bases = 0, n_pairs = 0, n_packed_words = 0;
  uint32_t n_reads = 0, n_regions = 0, max_len = 0;
  int grid_main = 0, grid_main_long = 0, grid_wide = 0, wide_cap = 0, wide_per_sm = 2;
  int force_ring = 0;  // OTGPU_RING_SLOTS=1024|2048 at otgpu_create: pin the ring tier (tests)
  // First window of the 1,024-slot kernel (k_align<1024, FW>): the short one when at least 30 % of
  // the previous batch on this context were low-error pairs (>= kFwRun bases per score step: reads
  // of the same allele at <= 1.5 % error each; noisy reads reach 10-16).  Only speed depends on it.
  // OTGPU_FIRST_WINDOW=0|1 at otgpu_create pins it (tests, A/B).
  int fw_force = -1;
  bool fw_next = false, fw_counted = false;
  bool wide_two_pass = false;  // OTGPU_WIDE_TWO_PASS=1 at otgpu_create: always use k_align<0> (tests)
  // streaming of large region batches (otgpu_region_edit_scores): a second context on the same
  // device, made on first use, so that chunk i+1 uploads while chunk i aligns
  otgpu_ctx* peer = nullptr;
  uint64_t chunk_bases = (uint64_t)256 << 20;  // OTGPU_CHUNK_MB at otgpu_create (tests use 1)
  uint64_t hook_chunk_bases = (uint64_t)64 << 20;  // chunk size of the hooked call (OTGPU_CHUNK_MB too)
  uint64_t n_chunks_last = 0;
  otgpu_stats stats = {0, 0, 0, 0};
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // pack|expand|align|wide|end
  bool timed = false;
  bool gather = false;  // upload_reads: the bases come from the device (k_gather_reads), not the host
};

namespace {

int fail_cuda(otgpu_ctx* c, cudaError_t e, const char* what) {
  snprintf(c->err, sizeof(c->err), "%s: %s", what, cudaGetErrorString(e));
  return e == cudaErrorMemoryAllocation ? OTGPU_ERR_OOM : OTGPU_ERR_CUDA;
}
#define CK(call)                                              \
  do {                                                        \
    cudaError_t e_ = (call);                                  \
    if (e_ != cudaSuccess) return fail_cuda(ctx, e_, #call);  \
  } while (0)

int reserve(otgpu_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return OTGPU_OK;
  if (b.p) CK(cudaFree(b.p));
  b.p = nullptr;
  b.cap = 0;
  size_t want = bytes + bytes / 8 + 256;
  CK(cudaMalloc(&b.p, want));
  b.cap = want;
  return OTGPU_OK;
}
#define RS(buf, bytes)                              \
  do {                                              \
    int s_ = reserve(ctx, ctx->buf, (bytes));       \
    if (s_ != OTGPU_OK) return s_;                  \
  } while (0)

int next_pow2(int x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}

// Shared front half of both upload flavours: reads + packed layout.
int upload_reads(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases, const uint64_t* read_off,
                 const uint32_t* read_len, uint32_t n_reads) {
  if (!ctx || (!bases && n_bases && !ctx->gather) || ((!read_off || !read_len) && n_reads)) return OTGPU_ERR_ARGS;
  CK(cudaSetDevice(ctx->device));
  // the previous batch (device buffers, pinned staging) must be finished before it is replaced
  CK(cudaStreamSynchronize(ctx->stream));
  if ((size_t)n_reads * 8 > ctx->h_pk_off_cap) {
    if (ctx->h_pk_off) cudaFreeHost(ctx->h_pk_off);
    ctx->h_pk_off = nullptr;
    ctx->h_pk_off_cap = 0;
    size_t want = (size_t)n_reads * 8 + (size_t)n_reads + 4096;
    CK(cudaMallocHost(&ctx->h_pk_off, want));
    ctx->h_pk_off_cap = want;
  }
  if (n_bases + 2ull * kPadBases >= ((uint64_t)1 << 34)) {  // kernels address reads by 32-bit word index
    snprintf(ctx->err, sizeof(ctx->err), "batch exceeds 16 GiB of bases; split it");
    return OTGPU_ERR_ARGS;
  }
  uint64_t* pk = static_cast<uint64_t*>(ctx->h_pk_off);
  // guard words either side: the extension's window loads are unguarded and a null-derived
  // offset addresses up to kPadBases positions before a read (wfa_edit.cuh)
  uint64_t words = kPadBases / 16;
  uint32_t max_len = 0;
  for (uint32_t r = 0; r < n_reads; ++r) {
    if (read_off[r] + read_len[r] > n_bases) {
      snprintf(ctx->err, sizeof(ctx->err), "read %u exceeds the bases buffer", r);
      return OTGPU_ERR_ARGS;
    }
    if (read_len[r] > kMaxReadLen) {
      snprintf(ctx->err, sizeof(ctx->err), "read %u is longer than %u bases", r, kMaxReadLen);
      return OTGPU_ERR_ARGS;
    }
    pk[r] = words;
    words += (((uint64_t)read_len[r] + 15) / 16 + 2 + 3) & ~3ull;  // 16 B aligned, >= 2 guards
    if (read_len[r] > max_len) max_len = read_len[r];
  }
  RS(bases, n_bases + 2ull * kPadBases + 16);
  RS(read_off, (size_t)n_reads * 8 + 8);
  RS(read_len, (size_t)n_reads * 4 + 4);
  RS(pk_off, (size_t)n_reads * 8 + 8);
  RS(read_flag, (size_t)n_reads * 4 + 4);
  RS(packed, (words + kPadBases / 16 + 4) * 4);
  RS(counters, sizeof(Counters));
  if (n_bases && !ctx->gather)
    CK(cudaMemcpyAsync(static_cast<uint8_t*>(ctx->bases.p) + kPadBases, bases, n_bases,
                       cudaMemcpyHostToDevice, ctx->stream));
  if (n_reads) {
    CK(cudaMemcpyAsync(ctx->read_off.p, read_off, (size_t)n_reads * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->read_len.p, read_len, (size_t)n_reads * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->pk_off.p, pk, (size_t)n_reads * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  ctx->n_bases = n_bases;
  ctx->n_reads = n_reads;
  ctx->n_packed_words = words;
  ctx->max_len = max_len;
  return OTGPU_OK;
}

int size_launch(otgpu_ctx* ctx) {
  if (ctx->grid_main == 0) {
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_align<1024>, kCtaThreads, ring_bytes(1024)));
    if (per_sm < 1) per_sm = 1;
    ctx->grid_main = ctx->n_sm * per_sm;
    CK(cudaFuncSetAttribute(k_align<2048>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)ring_bytes(2048)));
    cudaFuncSetAttribute(k_align<2048>, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_align<2048>, kCtaThreads, ring_bytes(2048)));
    if (per_sm < 1) per_sm = 1;
    ctx->grid_main_long = ctx->n_sm * per_sm;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_align<0>, kCtaThreads, 0));
    if (per_sm < 1) per_sm = 1;
    ctx->wide_per_sm = per_sm;
  }
  // wide fallback: ring must hold every diagonal of the longest possible pair
  const int need = next_pow2(2 * (int)ctx->max_len + 256);
  int grid_wide = ctx->n_sm * ctx->wide_per_sm;  // full occupancy: the wide pairs are the heaviest
  const size_t budget = (size_t)4 << 30;
  while (grid_wide > 1 && (size_t)grid_wide * kWarpsPerCta * need * 4 > budget) grid_wide /= 2;
  ctx->wide_cap = need;
  ctx->grid_wide = grid_wide;
  RS(wide_ws, (size_t)grid_wide * kWarpsPerCta * need * 4);
  return OTGPU_OK;
}

}  // namespace

extern "C" {

int otgpu_abi_version(void) { return OTGPU_ABI_VERSION; }

int otgpu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char* otgpu_strerror(int status) {
  switch (status) {
    case OTGPU_OK: return "ok";
    case OTGPU_ERR_ARGS: return "invalid arguments";
    case OTGPU_ERR_NO_DEVICE: return "no CUDA device (otter_b200 has no CPU fallback)";
    case OTGPU_ERR_CUDA: return "CUDA error";
    case OTGPU_ERR_OOM: return "out of memory";
    case OTGPU_ERR_STATE: return "call sequence error";
    case OTGPU_ERR_LIMIT:
      return "batch exceeds the fixed limits of otgpu_edit_scores_static (OTGPU_STATIC_MAX_* in "
             "otter_b200.h: 4 MiB of bases, 4,096 reads, 131,072 pairs, reads of at most 8,192 bases)";
    default: return "unknown status";
  }
}

int otgpu_create(int device, otgpu_ctx** out) {
  if (!out) return OTGPU_ERR_ARGS;
  *out = nullptr;
  if (device < 0 || device >= otgpu_device_count()) return OTGPU_ERR_NO_DEVICE;
  otgpu_ctx* ctx = new (std::nothrow) otgpu_ctx();
  if (!ctx) return OTGPU_ERR_OOM;
  ctx->device = device;
  if (const char* rs = getenv("OTGPU_RING_SLOTS")) {
    const int v = atoi(rs);
    if (v == 1024 || v == 2048) ctx->force_ring = v;
  }
  if (const char* fw = getenv("OTGPU_FIRST_WINDOW")) {
    if (fw[0] == '0' || fw[0] == '1') ctx->fw_force = fw[0] - '0';
  }
  if (const char* tp = getenv("OTGPU_WIDE_TWO_PASS")) ctx->wide_two_pass = atoi(tp) != 0;
  if (const char* cm = getenv("OTGPU_CHUNK_MB")) {
    const long v = atol(cm);
    if (v > 0) ctx->chunk_bases = ctx->hook_chunk_bases = (uint64_t)v << 20;
  }
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&ctx->n_sm, cudaDevAttrMultiProcessorCount, device);
  if (e == cudaSuccess) e = cudaMallocHost(&ctx->h_counters, sizeof(Counters));
  for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&ctx->ev[i]);
  if (e != cudaSuccess) {
    delete ctx;
    return OTGPU_ERR_CUDA;
  }
  *out = ctx;
  return OTGPU_OK;
}

void otgpu_destroy(otgpu_ctx* ctx) {
  if (!ctx) return;
  if (ctx->peer) otgpu_destroy(ctx->peer);
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->bases, &ctx->read_off, &ctx->read_len, &ctx->pk_off, &ctx->read_flag,
                    &ctx->packed, &ctx->region_start, &ctx->score_off, &ctx->pairs, &ctx->scores,
                    &ctx->ovf_list, &ctx->counters, &ctx->wide_ws, &ctx->order, &ctx->order_ctl, &ctx->aff_rings,
                    &ctx->aff_slab, &ctx->aff_slab2, &ctx->aff_ops, &ctx->cigar_off, &ctx->cigars,
                    &ctx->cigar_len, &ctx->cigar_begin, &ctx->cigars_packed, &ctx->bam_reads,
                    &ctx->bam_cigar, &ctx->bam_seq4, &ctx->bam_status, &ctx->bam_coords, &ctx->bam_plan,
                    &ctx->bam_out, &ctx->bam_begin, &ctx->bam_len, &ctx->gather_index,
                    &ctx->ppoa_in, &ctx->ppoa_ws, &ctx->ppoa_out,
                    &ctx->kde_values, &ctx->kde_off, &ctx->kde_grid, &ctx->kde_h, &ctx->kde_inv_h,
                    &ctx->kde_out};
  for (DevBuf* b : bufs)
    if (b->p) cudaFree(b->p);
  if (ctx->h_pk_off) cudaFreeHost(ctx->h_pk_off);
  if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
  if (ctx->h_cigars) cudaFreeHost(ctx->h_cigars);
  if (ctx->h_cigar_begin) cudaFreeHost(ctx->h_cigar_begin);
  if (ctx->h_reads) cudaFreeHost(ctx->h_reads);
  if (ctx->h_read_begin) cudaFreeHost(ctx->h_read_begin);
  for (cudaEvent_t e : ctx->ev)
    if (e) cudaEventDestroy(e);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* otgpu_last_cuda_error(const otgpu_ctx* ctx) { return ctx ? ctx->err : ""; }
void* otgpu_stream(otgpu_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

uint64_t otgpu_last_chunk_count(const otgpu_ctx* ctx) { return ctx ? ctx->n_chunks_last : 0; }
int otgpu_first_window(const otgpu_ctx* ctx) {
  return ctx ? (ctx->fw_force >= 0 ? ctx->fw_force : (ctx->fw_next ? 1 : 0)) : 0;
}

int otgpu_get_stats(const otgpu_ctx* ctx, otgpu_stats* out) {
  if (!ctx || !out) return OTGPU_ERR_ARGS;
  *out = ctx->stats;
  return OTGPU_OK;
}

int otgpu_host_alloc(size_t bytes, void** out) {
  if (!out) return OTGPU_ERR_ARGS;
  *out = nullptr;
  if (otgpu_device_count() == 0) return OTGPU_ERR_NO_DEVICE;
  cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return OTGPU_ERR_OOM;
  }
  return OTGPU_OK;
}
void otgpu_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

int otgpu_batch_upload_regions(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                               const uint64_t* read_off, const uint32_t* read_len,
                               uint32_t n_reads, const uint32_t* region_start,
                               uint32_t n_regions, const uint64_t* score_off, uint64_t n_scores) {
  if (!ctx || !region_start || (!score_off && n_regions)) return OTGPU_ERR_ARGS;
  ctx->resident = false;
  for (uint32_t g = 0; g < n_regions; ++g) {
    if (region_start[g + 1] < region_start[g] || region_start[g + 1] > n_reads) return OTGPU_ERR_ARGS;
    const uint64_t n = region_start[g + 1] - region_start[g];
    if (score_off[g] + n * (n - (n ? 1 : 0)) / 2 > n_scores) return OTGPU_ERR_ARGS;
  }
  if (n_scores >= 0xffffffffull) return OTGPU_ERR_ARGS;  // overflow list holds 32-bit slots
  int s = upload_reads(ctx, bases, n_bases, read_off, read_len, n_reads);
  if (s != OTGPU_OK) return s;
  RS(region_start, (size_t)(n_regions + 1) * 4);
  RS(score_off, (size_t)n_regions * 8 + 8);
  RS(pairs, (size_t)n_scores * sizeof(otgpu_pair) + 32);
  RS(scores, (size_t)n_scores * 4 + 4);
  RS(ovf_list, (size_t)n_scores * 4 + 4);
  RS(order, (size_t)n_scores * 4 + 4);
  RS(order_ctl, sizeof(OrderCtl));
  CK(cudaMemcpyAsync(ctx->region_start.p, region_start, (size_t)(n_regions + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (n_regions)
    CK(cudaMemcpyAsync(ctx->score_off.p, score_off, (size_t)n_regions * 8, cudaMemcpyHostToDevice, ctx->stream));
  ctx->n_regions = n_regions;
  ctx->n_pairs = n_scores;
  ctx->by_regions = true;
  s = size_launch(ctx);
  if (s != OTGPU_OK) return s;
  ctx->resident = true;
  return OTGPU_OK;
}

int otgpu_batch_upload_pairs(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                             const otgpu_pair* pairs, uint64_t n_pairs) {
  if (!ctx || (!pairs && n_pairs)) return OTGPU_ERR_ARGS;
  ctx->resident = false;
  if (n_pairs >= 0xffffffffull) return OTGPU_ERR_ARGS;
  for (uint64_t i = 0; i < n_pairs; ++i) {
    const otgpu_pair& p = pairs[i];
    if (p.pattern >= n_reads || p.text >= n_reads) return OTGPU_ERR_ARGS;
    if (p.pattern_begin_free >= 0) {
      // wavefront_check_endsfree_form (wavefront_align.c:40-61) would exit(1) on these
      if (p.pattern_end_free < 0 || p.text_begin_free < 0 || p.text_end_free < 0 ||
          (uint32_t)p.pattern_begin_free > read_len[p.pattern] ||
          (uint32_t)p.pattern_end_free > read_len[p.pattern] ||
          (uint32_t)p.text_begin_free > read_len[p.text] ||
          (uint32_t)p.text_end_free > read_len[p.text])
        return OTGPU_ERR_ARGS;
    }
  }
  int s = upload_reads(ctx, bases, n_bases, read_off, read_len, n_reads);
  if (s != OTGPU_OK) return s;
  RS(pairs, (size_t)n_pairs * sizeof(otgpu_pair) + 32);
  RS(scores, (size_t)n_pairs * 4 + 4);
  RS(ovf_list, (size_t)n_pairs * 4 + 4);
  RS(order, (size_t)n_pairs * 4 + 4);
  RS(order_ctl, sizeof(OrderCtl));
  if (n_pairs)
    CK(cudaMemcpyAsync(ctx->pairs.p, pairs, (size_t)n_pairs * sizeof(otgpu_pair), cudaMemcpyHostToDevice, ctx->stream));
  ctx->n_pairs = n_pairs;
  ctx->n_regions = 0;
  ctx->by_regions = false;
  s = size_launch(ctx);
  if (s != OTGPU_OK) return s;
  ctx->resident = true;
  return OTGPU_OK;
}

#ifdef OTGPU_CHECK_BOUNDS
// Checking build only (tests): tell the kernels which two buffers a window load may touch, and let
// the test read back how many loads were checked and how many would have left them.
static int check_set_ranges(otgpu_ctx* ctx) {
  CheckRange r[2];
  r[0].lo = static_cast<const uint32_t*>(ctx->packed.p);
  r[0].hi = r[0].lo + ctx->packed.cap / 4;
  r[1].lo = static_cast<const uint32_t*>(ctx->bases.p);
  r[1].hi = r[1].lo + ctx->bases.cap / 4;
  CK(cudaMemcpyToSymbolAsync(otgpu_check_range, r, sizeof r, 0, cudaMemcpyHostToDevice, ctx->stream));
  return OTGPU_OK;
}
int otgpu_check_counts(otgpu_ctx* ctx, unsigned long long* loads, unsigned long long* oob) {
  if (!ctx || !loads || !oob) return OTGPU_ERR_ARGS;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaMemcpyFromSymbol(loads, otgpu_check_loads, sizeof *loads));
  CK(cudaMemcpyFromSymbol(oob, otgpu_check_oob, sizeof *oob));
  return OTGPU_OK;
}
#endif

int otgpu_batch_align(otgpu_ctx* ctx) {
  if (!ctx) return OTGPU_ERR_ARGS;
  if (!ctx->resident) return OTGPU_ERR_STATE;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
#ifdef OTGPU_CHECK_BOUNDS
  if (int cs = check_set_ranges(ctx)) return cs;
#endif
  DevBatch b;
  b.bases = static_cast<const uint8_t*>(ctx->bases.p);
  b.read_off = static_cast<const uint64_t*>(ctx->read_off.p);
  b.read_len = static_cast<const uint32_t*>(ctx->read_len.p);
  b.pk_off = static_cast<const uint64_t*>(ctx->pk_off.p);
  b.packed = static_cast<const uint32_t*>(ctx->packed.p);
  b.read_flag = static_cast<uint32_t*>(ctx->read_flag.p);
  b.pairs = static_cast<const otgpu_pair*>(ctx->pairs.p);
  b.scores = static_cast<int32_t*>(ctx->scores.p);
  b.n_pairs = ctx->n_pairs;
  b.n_reads = ctx->n_reads;
  CK(cudaMemsetAsync(ctx->counters.p, 0, sizeof(Counters), st));
  CK(cudaEventRecord(ctx->ev[0], st));
  if (ctx->n_reads) {
    CK(cudaMemsetAsync(ctx->read_flag.p, 0, (size_t)ctx->n_reads * 4, st));
    k_pack_reads<<<ctx->n_reads, 128, 0, st>>>(b, static_cast<uint32_t*>(ctx->packed.p));
    ctx->stats.kernel_launches++;
  }
  CK(cudaEventRecord(ctx->ev[1], st));
  if (ctx->by_regions && ctx->n_pairs) {
    CK(cudaMemsetAsync(ctx->pairs.p, 0xff, (size_t)ctx->n_pairs * sizeof(otgpu_pair), st));
    if (ctx->n_regions) {
      k_expand_regions<<<ctx->n_regions, 128, 0, st>>>(
          static_cast<const uint32_t*>(ctx->region_start.p),
          static_cast<const uint64_t*>(ctx->score_off.p), b.read_len,
          static_cast<otgpu_pair*>(ctx->pairs.p));
      ctx->stats.kernel_launches++;
    }
  }
  if (ctx->n_pairs) {  // heaviest-first queue order
    OrderCtl* ctl = static_cast<OrderCtl*>(ctx->order_ctl.p);
    CK(cudaMemsetAsync(ctl, 0, sizeof(OrderCtl), st));
    const int g = (int)((ctx->n_pairs + 255) / 256 < (uint64_t)ctx->n_sm * 8 ? (ctx->n_pairs + 255) / 256
                                                                             : (uint64_t)ctx->n_sm * 8);
    k_order_count<<<g, 256, 0, st>>>(b.pairs, ctx->n_pairs, b.read_len, ctl);
    k_order_scan<<<1, 1, 0, st>>>(ctl);
    k_order_scatter<<<g, 256, 0, st>>>(b.pairs, ctx->n_pairs, b.read_len, ctl,
                                       static_cast<uint32_t*>(ctx->order.p));
    ctx->stats.kernel_launches += 3;
  }
  CK(cudaEventRecord(ctx->ev[2], st));
  if (ctx->n_pairs) {
    Counters* ctr = static_cast<Counters*>(ctx->counters.p);
    uint32_t* ovf = static_cast<uint32_t*>(ctx->ovf_list.p);
    uint32_t* order = static_cast<uint32_t*>(ctx->order.p);
    // wavefront width grows with read length: batches of long reads get the 2048-slot ring
    const bool long_ring = ctx->force_ring ? ctx->force_ring == 2048 : ctx->max_len > kLongReadLen;
    const int grid = long_ring ? ctx->grid_main_long : ctx->grid_main;
    // one HBM ring per warp of this launch (the wide workspace holds grid_wide CTAs' worth): pairs
    // that outgrow the shared ring continue inside the kernel; else they wait for k_align<0>
    int32_t* ws = grid <= ctx->grid_wide && !ctx->wide_two_pass ? static_cast<int32_t*>(ctx->wide_ws.p) : nullptr;
    ctx->fw_counted = !long_ring;
    if (long_ring)
      k_align<2048><<<grid, kCtaThreads, ring_bytes(2048), st>>>(b, ctr, order, -1, 0, ovf, ws, ctx->wide_cap);
    else if (ctx->fw_force >= 0 ? ctx->fw_force == 1 : ctx->fw_next)
      k_align<1024, 1><<<grid, kCtaThreads, ring_bytes(1024), st>>>(b, ctr, order, -1, 0, ovf, ws, ctx->wide_cap);
    else
      k_align<1024><<<grid, kCtaThreads, ring_bytes(1024), st>>>(b, ctr, order, -1, 0, ovf, ws, ctx->wide_cap);
    CK(cudaEventRecord(ctx->ev[3], st));
    k_align<0><<<ctx->grid_wide, kCtaThreads, 0, st>>>(
        b, ctr, ovf, 1, 2, ovf, static_cast<int32_t*>(ctx->wide_ws.p), ctx->wide_cap);
    ctx->stats.kernel_launches += 2;
  } else {
    CK(cudaEventRecord(ctx->ev[3], st));
  }
  CK(cudaEventRecord(ctx->ev[4], st));
  ctx->timed = true;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(ctx->h_counters, ctx->counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, st));
  return OTGPU_OK;
}

int otgpu_last_align_ms(otgpu_ctx* ctx, float ms[4]) {
  if (!ctx || !ms) return OTGPU_ERR_ARGS;
  if (!ctx->timed) return OTGPU_ERR_STATE;
  CK(cudaSetDevice(ctx->device));
  CK(cudaEventSynchronize(ctx->ev[4]));
  for (int i = 0; i < 4; ++i) CK(cudaEventElapsedTime(&ms[i], ctx->ev[i], ctx->ev[i + 1]));
  return OTGPU_OK;
}

int otgpu_synchronize(otgpu_ctx* ctx) {
  if (!ctx) return OTGPU_ERR_ARGS;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  return OTGPU_OK;
}

int otgpu_batch_download(otgpu_ctx* ctx, int32_t* scores_out, uint64_t n_scores) {
  if (!ctx || (!scores_out && n_scores)) return OTGPU_ERR_ARGS;
  if (!ctx->resident || n_scores != ctx->n_pairs) return OTGPU_ERR_STATE;
  CK(cudaSetDevice(ctx->device));
  if (n_scores)
    CK(cudaMemcpyAsync(scores_out, ctx->scores.p, (size_t)n_scores * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  const Counters* hc = static_cast<const Counters*>(ctx->h_counters);
  ctx->stats.n_pairs = ctx->n_pairs;
  ctx->stats.n_pairs_wide = hc->v[1] + hc->v[4];
  ctx->stats.n_pairs_8bit = hc->v[3];
  if (ctx->fw_counted && ctx->n_pairs) ctx->fw_next = hc->v[5] * 10 >= (unsigned long long)ctx->n_pairs * 3;
  return OTGPU_OK;
}

int otgpu_edit_scores(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                      const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                      const otgpu_pair* pairs, uint64_t n_pairs, int32_t* scores_out) {
  int s = otgpu_batch_upload_pairs(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs);
  if (s == OTGPU_OK) s = otgpu_batch_align(ctx);
  if (s == OTGPU_OK) s = otgpu_batch_download(ctx, scores_out, n_pairs);
  return s;
}

// Large batches are streamed: consecutive regions are cut into chunks of about chunk_bases bytes
// of reads and the chunks alternate between two contexts (two streams, two sets of buffers) on the
// same device, so the H2D copy and the host-side table building of chunk i+1 overlap the kernels
// of chunk i, and device memory stays bounded by two chunks.  Requires the usual layout (reads and
// scores of consecutive regions are consecutive); anything else takes the single-batch path.
// Results do not depend on the chunking (regions are independent).
static int region_scores_streamed(otgpu_ctx* ctx, const uint8_t* bases, const uint64_t* read_off,
                                  const uint32_t* read_len, const uint32_t* region_start,
                                  uint32_t n_regions, const uint64_t* score_off,
                                  int32_t* scores_out, uint64_t chunk_bases,
                                  const otgpu_stream_hooks* hooks) {
  if (!ctx->peer) {
    const int s = otgpu_create(ctx->device, &ctx->peer);
    if (s != OTGPU_OK) return s;
    ctx->peer->force_ring = ctx->force_ring;
  }
  otgpu_ctx* slot[2] = {ctx, ctx->peer};
  struct Pending { bool on = false; uint64_t s0 = 0, ns = 0; } pend[2];
  std::vector<uint64_t> off_c, so_c;
  std::vector<uint32_t> rs_c;
  uint64_t chunks = 0, pairs_total = 0, wide_total = 0, b8_total = 0;
  auto finish = [&](int k) -> int {
    if (!pend[k].on) return OTGPU_OK;
    pend[k].on = false;
    const int s = otgpu_batch_download(slot[k], scores_out + pend[k].s0, pend[k].ns);
    if (s == OTGPU_OK && hooks && hooks->scores_ready) hooks->scores_ready(hooks->user, pend[k].s0, pend[k].ns);
    pairs_total += slot[k]->stats.n_pairs;
    wide_total += slot[k]->stats.n_pairs_wide;
    b8_total += slot[k]->stats.n_pairs_8bit;
    if (k == 1) ctx->stats.kernel_launches += slot[1]->stats.kernel_launches, slot[1]->stats.kernel_launches = 0;
    return s;
  };
  uint32_t g0 = 0;
  while (g0 < n_regions) {
    uint32_t g1 = g0;
    // (trailing regions without reads: region_start[g0] == n_reads, nothing left to address)
    const uint64_t base0 = region_start[g0] < region_start[n_regions] ? read_off[region_start[g0]] : 0;
    uint64_t bytes = 0;
    // with hooks the first chunks are small (1/8, 1/4, 1/2 of the chunk size): the caller's staging
    // copy of the first chunk is the one wait nothing can hide
    const uint64_t want = hooks ? chunk_bases >> (chunks < 3 ? 3 - chunks : 0) : chunk_bases;
    while (g1 < n_regions && (g1 == g0 || bytes < want || region_start[g1] == region_start[n_regions])) {
      const uint32_t ra = region_start[g1], rb = region_start[g1 + 1];
      if (rb > ra) bytes = read_off[rb - 1] + read_len[rb - 1] - base0;
      ++g1;
    }
    const uint32_t r0 = region_start[g0], r1 = region_start[g1];
    const uint64_t s0 = score_off[g0];
    off_c.resize(r1 - r0);
    for (uint32_t r = r0; r < r1; ++r) off_c[r - r0] = read_off[r] - base0;
    rs_c.resize(g1 - g0 + 1);
    so_c.resize(g1 - g0);
    uint64_t ns = 0;
    for (uint32_t g = g0; g < g1; ++g) {
      rs_c[g - g0] = region_start[g] - r0;
      so_c[g - g0] = score_off[g] - s0;
      const uint64_t n = region_start[g + 1] - region_start[g];
      ns = so_c[g - g0] + n * (n ? n - 1 : 0) / 2;
    }
    rs_c[g1 - g0] = r1 - r0;
    const int k = (int)(chunks & 1);
    int s = finish(k);  // the chunk before last on this slot: its buffers are about to be reused
    if (hooks && hooks->wait_bases) hooks->wait_bases(hooks->user, base0 + bytes);
    if (s == OTGPU_OK)
      s = otgpu_batch_upload_regions(slot[k], bases + base0, bytes, off_c.data(), read_len + r0, r1 - r0,
                                     rs_c.data(), g1 - g0, so_c.data(), ns);
    if (s == OTGPU_OK) s = otgpu_batch_align(slot[k]);
    if (s != OTGPU_OK) {
      if (k == 1) snprintf(ctx->err, sizeof(ctx->err), "%s", slot[1]->err);
      return s;
    }
    pend[k].on = true;
    pend[k].s0 = s0;
    pend[k].ns = ns;
    ++chunks;
    g0 = g1;
  }
  int s = finish((int)(chunks & 1));
  if (s == OTGPU_OK) s = finish((int)((chunks + 1) & 1));
  ctx->stats.n_pairs = pairs_total;
  ctx->stats.n_pairs_wide = wide_total;
  ctx->stats.n_pairs_8bit = b8_total;
  ctx->n_chunks_last = chunks;
  return s;
}

// Layout check of the streamed path: reads and scores of consecutive regions are consecutive.
static bool regions_are_consecutive(const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                                    uint64_t n_bases, const uint32_t* region_start, uint32_t n_regions,
                                    const uint64_t* score_off, uint64_t n_scores) {
  bool ok = region_start[0] == 0 && region_start[n_regions] == n_reads && score_off[0] == 0;
  uint64_t at = 0, sc = 0;
  for (uint32_t g = 0; ok && g < n_regions; ++g) {
    if (region_start[g + 1] < region_start[g] || region_start[g + 1] > n_reads) return false;
    const uint64_t n = region_start[g + 1] - region_start[g];
    ok = score_off[g] == sc;
    sc += n * (n ? n - 1 : 0) / 2;
    for (uint32_t r = region_start[g]; ok && r < region_start[g + 1]; ++r) {
      ok = read_off[r] >= at && read_off[r] + read_len[r] <= n_bases;
      at = read_off[r] + read_len[r];
    }
  }
  return ok && sc == n_scores;
}

int otgpu_region_edit_scores_hooked(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                    const uint64_t* read_off, const uint32_t* read_len,
                                    uint32_t n_reads, const uint32_t* region_start,
                                    uint32_t n_regions, const uint64_t* score_off,
                                    uint64_t n_scores, int32_t* scores_out,
                                    const otgpu_stream_hooks* hooks) {
  if (ctx) ctx->n_chunks_last = 1;
  const bool hooked = hooks && (hooks->wait_bases || hooks->scores_ready);
  // with hooks the chunks are smaller: their purpose is overlap with the caller's own work
  const uint64_t chunk = hooked ? (ctx ? ctx->hook_chunk_bases : 0) : (ctx ? ctx->chunk_bases : 0);
  if (ctx && bases && read_off && read_len && region_start && score_off && scores_out &&
      n_regions > 1 && n_bases > (hooked ? chunk / 4 : chunk + chunk / 2)) {
    // streaming needs reads and scores laid out region after region, which is how every caller
    // here builds them; verify (the tables only - the bases may still be on their way), else fall
    // through to the single batch
    if (regions_are_consecutive(read_off, read_len, n_reads, n_bases, region_start, n_regions, score_off,
                                n_scores))
      return region_scores_streamed(ctx, bases, read_off, read_len, region_start, n_regions, score_off,
                                    scores_out, chunk, hooked ? hooks : nullptr);
  }
  if (hooked && hooks->wait_bases) hooks->wait_bases(hooks->user, n_bases);
  int s = otgpu_batch_upload_regions(ctx, bases, n_bases, read_off, read_len, n_reads,
                                     region_start, n_regions, score_off, n_scores);
  if (s == OTGPU_OK) s = otgpu_batch_align(ctx);
  if (s == OTGPU_OK) s = otgpu_batch_download(ctx, scores_out, n_scores);
  if (s == OTGPU_OK && hooked && hooks->scores_ready) hooks->scores_ready(hooks->user, 0, n_scores);
  return s;
}

int otgpu_region_edit_scores(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                             const uint64_t* read_off, const uint32_t* read_len,
                             uint32_t n_reads, const uint32_t* region_start, uint32_t n_regions,
                             const uint64_t* score_off, uint64_t n_scores, int32_t* scores_out) {
  return otgpu_region_edit_scores_hooked(ctx, bases, n_bases, read_off, read_len, n_reads, region_start,
                                         n_regions, score_off, n_scores, scores_out, nullptr);
}

}  // extern "C"

namespace {

// Gap-affine + CIGAR: persistent warps, one pair per warp, per-warp HBM workspace.
// RETRY = second pass over the pairs whose trace outgrew the first-tier slab.
constexpr int kRingPad = 32;  // int32: 128 B between the power-of-two rings of a warp
#ifndef OTGPU_AFF_MIN_CTAS
#define OTGPU_AFF_MIN_CTAS 1
#endif
template <bool RETRY, bool EDIT>
__global__ void __launch_bounds__(kCtaThreads, OTGPU_AFF_MIN_CTAS)
k_affine(DevBatch b, Counters* __restrict__ ctr, const uint32_t* __restrict__ order,
         uint32_t* __restrict__ ovf_list, AffPen pen,
         int32_t* __restrict__ rings, int wcap, int n_rings, uint8_t* __restrict__ slab, uint32_t slab_bytes,
         uint8_t* __restrict__ ops, uint32_t ops_cap, const uint64_t* __restrict__ cigar_off,
         char* __restrict__ cigars, int32_t* __restrict__ cigar_len, int max_warps) {
  __shared__ AffSlot meta_smem[kWarpsPerCta][EDIT ? 1 : kAffMaxScope];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t gw = (size_t)blockIdx.x * kWarpsPerCta + warp;
  if (max_warps > 0 && gw >= (size_t)max_warps) return;  // retry tier: only the warps that own a slab
  AffWork W;
  W.rstride = wcap + kRingPad;
  W.rings = rings + gw * (size_t)n_rings * W.rstride;
  W.wmask = wcap - 1;
  W.slab = slab + gw * (size_t)slab_bytes;
  W.slab_bytes = slab_bytes;
  W.ops = ops + gw * (size_t)ops_cap;
  W.ops_cap = ops_cap;
  const unsigned long long limit = RETRY ? ctr->v[1] : b.n_pairs;
  for (;;) {
    unsigned long long q = 0;
    if (lane == 0) q = atomicAdd(&ctr->v[RETRY ? 2 : 0], 1ull);
    q = __shfl_sync(kFull, q, 0);
    if (q >= limit) break;
    const uint64_t idx = RETRY ? ovf_list[q] : order[q];  // first pass: heaviest pairs first
    const otgpu_pair pr = b.pairs[idx];
    EditJob J;
    J.plen = (int)b.read_len[pr.pattern];
    J.tlen = (int)b.read_len[pr.text];
    J.pb = pr.pattern_begin_free; J.pe = pr.pattern_end_free;
    J.tb = pr.text_begin_free;    J.te = pr.text_end_free;
    const bool bytes = ((b.read_flag[pr.pattern] | b.read_flag[pr.text]) & 1u) != 0;
    // cigar_off == nullptr: the caller wants scores and CIGAR lengths only, nothing is written
    const uint64_t c0 = cigar_off ? cigar_off[idx] : 0;
    const int cap = cigar_off ? (int)(cigar_off[idx + 1] - c0) : 0;
    int score = 0, len;
    if (bytes) {
      J.words = reinterpret_cast<const uint32_t*>(b.bases);
      J.P = view8(b, pr.pattern); J.T = view8(b, pr.text);
      len = EDIT ? wfa_editcigar_pair<8>(W, J, lane, cigars + c0, cap, &score)
                 : wfa_affine_pair<8>(W, meta_smem[warp], J, pen, lane, cigars + c0, cap, &score);
    } else {
      J.words = b.packed;
      J.P = view2(b, pr.pattern); J.T = view2(b, pr.text);
      len = EDIT ? wfa_editcigar_pair<2>(W, J, lane, cigars + c0, cap, &score)
                 : wfa_affine_pair<2>(W, meta_smem[warp], J, pen, lane, cigars + c0, cap, &score);
    }
    __syncwarp();
    if (lane == 0) {
      if (len == kAffOverflow && !RETRY) {
        const unsigned long long pos = atomicAdd(&ctr->v[1], 1ull);
        ovf_list[pos] = (uint32_t)idx;
      } else {
        if (len == kAffOverflow) len = kAffOom;
        cigar_len[idx] = len;
        b.scores[idx] = len >= 0 ? score : 0;
      }
    }
  }
}

}  // namespace

namespace {
// Packed CIGAR output: start offsets = exclusive scan of the CIGAR lengths (failed alignments
// count 0), one CTA; then one warp per CIGAR moves it from its capacity slot to its packed place.
__global__ void __launch_bounds__(1024)
k_cigar_scan(const int32_t* __restrict__ len, const uint64_t n, uint64_t* __restrict__ begin) {
  __shared__ uint64_t part[1024];
  const uint64_t per = (n + 1023) / 1024, a = threadIdx.x * per, b = a + per < n ? a + per : n;
  uint64_t sum = 0;
  for (uint64_t i = a; i < b; ++i) sum += len[i] > 0 ? (uint64_t)len[i] : 0;
  part[threadIdx.x] = sum;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const uint64_t v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  uint64_t at = part[threadIdx.x] - sum;
  for (uint64_t i = a; i < b; ++i) {
    begin[i] = at;
    at += len[i] > 0 ? (uint64_t)len[i] : 0;
  }
  if (threadIdx.x == 1023) begin[n] = part[1023];
}

__global__ void __launch_bounds__(kCtaThreads)
k_cigar_pack(const char* __restrict__ cigars, const uint64_t* __restrict__ cap_off,
             const uint64_t* __restrict__ begin, const int32_t* __restrict__ len, const uint64_t n,
             char* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const uint64_t warps = (uint64_t)gridDim.x * kWarpsPerCta;
  for (uint64_t i = (uint64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5); i < n; i += warps) {
    const int l = len[i];
    const char* src = cigars + cap_off[i];
    char* dst = out + begin[i];
    for (int k = lane; k < l; k += 32) dst[k] = src[k];
  }
}
}  // namespace

namespace {
// Alignments with CIGAR, gap-affine (edit = false) or edit distance (edit = true): same batch
// upload, per-warp HBM workspace, two-tier trace slab and result download.
int cigars_impl(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases, const uint64_t* read_off,
                const uint32_t* read_len, uint32_t n_reads, const otgpu_pair* pairs,
                uint64_t n_pairs, const bool edit, AffPen pen, int32_t* scores_out, char* cigars_out,
                const uint64_t* cigar_off, int32_t* cigar_len_out,
                const char** packed_out = nullptr, const uint64_t** packed_begin = nullptr) {
  if (!ctx || ((!scores_out || !cigar_len_out) && n_pairs)) return OTGPU_ERR_ARGS;
  if (!cigar_off && cigars_out) return OTGPU_ERR_ARGS;  // both null = lengths only
  const bool packed = packed_out != nullptr;
  if (packed && (!packed_begin || cigars_out || cigar_off)) return OTGPU_ERR_ARGS;
  int s = otgpu_batch_upload_pairs(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs);
  if (s != OTGPU_OK) return s;
  if (packed) {
    *packed_out = nullptr;
    *packed_begin = nullptr;
  }
  if (n_pairs == 0) return OTGPU_OK;
  std::vector<uint64_t> own_off;
  if (packed) {  // capacity slots on the device only: a CIGAR is never longer than |pattern|+|text|
    try {
      own_off.resize(n_pairs + 1);
    } catch (...) {  // nothing may be thrown across the C ABI
      return OTGPU_ERR_OOM;
    }
    own_off[0] = 0;
    for (uint64_t i = 0; i < n_pairs; ++i)
      own_off[i + 1] = own_off[i] + read_len[pairs[i].pattern] + read_len[pairs[i].text];
    cigar_off = own_off.data();
  }
  const uint64_t cig_bytes = cigar_off ? cigar_off[n_pairs] : 0;
  for (uint64_t i = 0; cigar_off && i < n_pairs; ++i)
    if (cigar_off[i + 1] < cigar_off[i]) return OTGPU_ERR_ARGS;
  if (cig_bytes && !cigars_out && !packed) return OTGPU_ERR_ARGS;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
#ifdef OTGPU_CHECK_BOUNDS
  if (int cs = check_set_ranges(ctx)) return cs;
#endif
  // per-warp workspace: rings cover every diagonal of the longest pair, so they never alias
  const int wcap = next_pow2(2 * (int)ctx->max_len + 64);
  const uint32_t ops_cap = 4u * ctx->max_len + 64u;
  // rings per warp: edit + CIGAR ping-pongs two; gap-affine keeps max(x, o+e)/g + 1 wavefronts of
  // M and e/g + 1 of I1 and of D1 (wfa_affine.cuh)
  const int n_rings = edit ? 2
                           : (pen.x > pen.o + pen.e ? pen.x : pen.o + pen.e) / pen.step + 1 +
                                 2 * (pen.e / pen.step + 1);
  const size_t ring_bytes = (size_t)n_rings * (wcap + kRingPad) * 4;
  // first-tier trace slab per warp: edit distance stores one byte per (score, diagonal) of ONE
  // component and its wavefronts are narrower, so half the slab and four times the warps
  const uint32_t slab1 = edit ? (1u << 20) : (2u << 20);
  int per_sm = 1;  // latency-bound on its per-warp HBM rings: as many resident warps as fit
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
      &per_sm, edit ? k_affine<false, true> : k_affine<false, false>, kCtaThreads, 0));
  if (const char* e = getenv("OTGPU_AFF_CTAS")) per_sm = atoi(e) > 0 ? atoi(e) : per_sm;
  int grid = ctx->n_sm * (per_sm > 0 ? per_sm : 1);
  while (grid > 8 && (size_t)grid * kWarpsPerCta * (ring_bytes + slab1 + ops_cap) > ((size_t)24 << 30)) grid /= 2;
  const size_t nw = (size_t)grid * kWarpsPerCta;
  RS(aff_rings, nw * ring_bytes);
  RS(aff_slab, nw * slab1);
  RS(aff_ops, nw * ops_cap);
  RS(cigar_off, (n_pairs + 1) * 8);
  RS(cigars, cig_bytes + 16);
  RS(cigar_len, n_pairs * 4);
  if (cigar_off)
    CK(cudaMemcpyAsync(ctx->cigar_off.p, cigar_off, (n_pairs + 1) * 8, cudaMemcpyHostToDevice, st));
  const uint64_t* d_cigar_off = cigar_off ? static_cast<const uint64_t*>(ctx->cigar_off.p) : nullptr;
  DevBatch b;
  b.bases = static_cast<const uint8_t*>(ctx->bases.p);
  b.read_off = static_cast<const uint64_t*>(ctx->read_off.p);
  b.read_len = static_cast<const uint32_t*>(ctx->read_len.p);
  b.pk_off = static_cast<const uint64_t*>(ctx->pk_off.p);
  b.packed = static_cast<const uint32_t*>(ctx->packed.p);
  b.read_flag = static_cast<uint32_t*>(ctx->read_flag.p);
  b.pairs = static_cast<const otgpu_pair*>(ctx->pairs.p);
  b.scores = static_cast<int32_t*>(ctx->scores.p);
  b.n_pairs = n_pairs;
  b.n_reads = n_reads;
  CK(cudaMemsetAsync(ctx->counters.p, 0, sizeof(Counters), st));
  CK(cudaMemsetAsync(ctx->read_flag.p, 0, (size_t)n_reads * 4, st));
  k_pack_reads<<<n_reads, 128, 0, st>>>(b, static_cast<uint32_t*>(ctx->packed.p));
  Counters* ctr = static_cast<Counters*>(ctx->counters.p);
  uint32_t* ovf = static_cast<uint32_t*>(ctx->ovf_list.p);
  uint32_t* order = static_cast<uint32_t*>(ctx->order.p);
  {  // heaviest-first queue order, as for the score-only kernel: the last wave of warps is short
    OrderCtl* ctl = static_cast<OrderCtl*>(ctx->order_ctl.p);
    CK(cudaMemsetAsync(ctl, 0, sizeof(OrderCtl), st));
    const int g = (int)((n_pairs + 255) / 256 < (uint64_t)ctx->n_sm * 8 ? (n_pairs + 255) / 256
                                                                         : (uint64_t)ctx->n_sm * 8);
    k_order_count<<<g, 256, 0, st>>>(b.pairs, n_pairs, b.read_len, ctl);
    k_order_scan<<<1, 1, 0, st>>>(ctl);
    k_order_scatter<<<g, 256, 0, st>>>(b.pairs, n_pairs, b.read_len, ctl, order);
    ctx->stats.kernel_launches += 3;
  }
  (edit ? k_affine<false, true> : k_affine<false, false>)<<<grid, kCtaThreads, 0, st>>>(
      b, ctr, order, ovf, pen, static_cast<int32_t*>(ctx->aff_rings.p), wcap, n_rings,
      static_cast<uint8_t*>(ctx->aff_slab.p), slab1, static_cast<uint8_t*>(ctx->aff_ops.p), ops_cap,
      d_cigar_off, static_cast<char*>(ctx->cigars.p), static_cast<int32_t*>(ctx->cigar_len.p), 0);
  ctx->stats.kernel_launches += 2;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(ctx->h_counters, ctx->counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const unsigned long long n_retry = static_cast<const Counters*>(ctx->h_counters)->v[1];
  bool free_slab2 = false;
  if (n_retry) {  // traces that outgrew 2 MiB: few warps, big slabs - one slab per retried pair, at
    // most 32 of them (8 GiB), never more warps than the rings and op buffers were sized for
    const uint32_t slab2 = 256u << 20;
    int warps2 = (int)(n_retry < 32 ? n_retry : 32);
    if (warps2 > grid * kWarpsPerCta) warps2 = grid * kWarpsPerCta;
    const int grid2 = (warps2 + kWarpsPerCta - 1) / kWarpsPerCta;
    RS(aff_slab2, (size_t)warps2 * slab2);
    free_slab2 = true;
    (edit ? k_affine<true, true> : k_affine<true, false>)<<<grid2, kCtaThreads, 0, st>>>(
        b, ctr, order, ovf, pen, static_cast<int32_t*>(ctx->aff_rings.p), wcap, n_rings,
        static_cast<uint8_t*>(ctx->aff_slab2.p), slab2, static_cast<uint8_t*>(ctx->aff_ops.p),
        ops_cap, d_cigar_off, static_cast<char*>(ctx->cigars.p),
        static_cast<int32_t*>(ctx->cigar_len.p), warps2);
    ctx->stats.kernel_launches++;
    CK(cudaGetLastError());
  }
  CK(cudaMemcpyAsync(scores_out, ctx->scores.p, n_pairs * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(cigar_len_out, ctx->cigar_len.p, n_pairs * 4, cudaMemcpyDeviceToHost, st));
  if (packed) {
    // only the bytes the CIGARs occupy leave the device, into pinned memory the context keeps
    RS(cigar_begin, (n_pairs + 1) * 8);
    RS(cigars_packed, cig_bytes + 16);
    if ((n_pairs + 1) * 8 > ctx->h_cigar_begin_cap) {
      if (ctx->h_cigar_begin) cudaFreeHost(ctx->h_cigar_begin);
      ctx->h_cigar_begin = nullptr;
      ctx->h_cigar_begin_cap = 0;
      const size_t want = (n_pairs + 1) * 8 + n_pairs + 4096;
      CK(cudaMallocHost(&ctx->h_cigar_begin, want));
      ctx->h_cigar_begin_cap = want;
    }
    uint64_t* d_begin = static_cast<uint64_t*>(ctx->cigar_begin.p);
    k_cigar_scan<<<1, 1024, 0, st>>>(static_cast<const int32_t*>(ctx->cigar_len.p), n_pairs, d_begin);
    k_cigar_pack<<<ctx->n_sm * 8, kCtaThreads, 0, st>>>(
        static_cast<const char*>(ctx->cigars.p), d_cigar_off, d_begin,
        static_cast<const int32_t*>(ctx->cigar_len.p), n_pairs, static_cast<char*>(ctx->cigars_packed.p));
    ctx->stats.kernel_launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(ctx->h_cigar_begin, d_begin, (n_pairs + 1) * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    const uint64_t total = static_cast<const uint64_t*>(ctx->h_cigar_begin)[n_pairs];
    if (total + 1 > ctx->h_cigars_cap) {
      if (ctx->h_cigars) cudaFreeHost(ctx->h_cigars);
      ctx->h_cigars = nullptr;
      ctx->h_cigars_cap = 0;
      const size_t want = total + total / 4 + 4096;
      CK(cudaMallocHost(&ctx->h_cigars, want));
      ctx->h_cigars_cap = want;
    }
    if (total) CK(cudaMemcpyAsync(ctx->h_cigars, ctx->cigars_packed.p, total, cudaMemcpyDeviceToHost, st));
    *packed_out = static_cast<const char*>(ctx->h_cigars);
    *packed_begin = static_cast<const uint64_t*>(ctx->h_cigar_begin);
  } else if (cig_bytes) {
    CK(cudaMemcpyAsync(cigars_out, ctx->cigars.p, cig_bytes, cudaMemcpyDeviceToHost, st));
  }
  CK(cudaStreamSynchronize(st));
  if (free_slab2 && ctx->aff_slab2.p) {  // gigabytes, needed by a handful of alignments: give them back
    cudaFree(ctx->aff_slab2.p);
    ctx->aff_slab2.p = nullptr;
    ctx->aff_slab2.cap = 0;
  }
  ctx->stats.n_pairs = n_pairs;
  ctx->stats.n_pairs_wide = n_retry;
  return OTGPU_OK;
}
}  // namespace

extern "C" int otgpu_affine_cigars(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                   const uint64_t* read_off, const uint32_t* read_len,
                                   uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
                                   int mismatch, int gap_open, int gap_ext, int32_t* scores_out,
                                   char* cigars_out, const uint64_t* cigar_off,
                                   int32_t* cigar_len_out) {
  // wavefronts_penalties_set_affine (wavefront_penalties.c:86-119): X>0, O>=0, E>0
  if (mismatch <= 0 || gap_open < 0 || gap_ext <= 0) return OTGPU_ERR_ARGS;
  AffPen pen{mismatch, gap_open, gap_ext, 0, 1};
  pen.scope = (gap_open + gap_ext > mismatch ? gap_open + gap_ext : mismatch) + 1;
  {
    auto gcd = [](int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; };
    pen.step = gcd(gcd(mismatch, gap_open + gap_ext), gap_ext);
  }
  if (pen.scope > kAffMaxScope) return OTGPU_ERR_ARGS;
  return cigars_impl(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs, false, pen,
                     scores_out, cigars_out, cigar_off, cigar_len_out);
}

extern "C" int otgpu_edit_cigars(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                 const uint64_t* read_off, const uint32_t* read_len,
                                 uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
                                 int32_t* scores_out, char* cigars_out, const uint64_t* cigar_off,
                                 int32_t* cigar_len_out) {
  return cigars_impl(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs, true,
                     AffPen{1, 0, 1, 2, 1}, scores_out, cigars_out, cigar_off, cigar_len_out);
}

extern "C" int otgpu_affine_cigars_packed(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                          const uint64_t* read_off, const uint32_t* read_len,
                                          uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
                                          int mismatch, int gap_open, int gap_ext, int32_t* scores_out,
                                          int32_t* cigar_len_out, const char** cigars,
                                          const uint64_t** cigar_begin) {
  if (mismatch <= 0 || gap_open < 0 || gap_ext <= 0 || !cigars || !cigar_begin) return OTGPU_ERR_ARGS;
  AffPen pen{mismatch, gap_open, gap_ext, 0, 1};
  pen.scope = (gap_open + gap_ext > mismatch ? gap_open + gap_ext : mismatch) + 1;
  {
    auto gcd = [](int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; };
    pen.step = gcd(gcd(mismatch, gap_open + gap_ext), gap_ext);
  }
  if (pen.scope > kAffMaxScope) return OTGPU_ERR_ARGS;
  return cigars_impl(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs, false, pen,
                     scores_out, nullptr, nullptr, cigar_len_out, cigars, cigar_begin);
}

extern "C" int otgpu_edit_cigars_packed(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                        const uint64_t* read_off, const uint32_t* read_len,
                                        uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
                                        int32_t* scores_out, int32_t* cigar_len_out,
                                        const char** cigars, const uint64_t** cigar_begin) {
  if (!cigars || !cigar_begin) return OTGPU_ERR_ARGS;
  return cigars_impl(ctx, bases, n_bases, read_off, read_len, n_reads, pairs, n_pairs, true,
                     AffPen{1, 0, 1, 2, 1}, scores_out, nullptr, nullptr, cigar_len_out, cigars,
                     cigar_begin);
}

// ---- batched read extraction (bam_extract.cuh) -------------------------------------------------
extern "C" int otgpu_extract_reads(otgpu_ctx* ctx, const otgpu_bam_read* reads, uint32_t n_reads,
                                   const uint32_t* cigar, uint64_t n_cigar_ops, const uint8_t* seq4,
                                   uint64_t n_seq_bytes, uint8_t* status, int32_t* coords,
                                   uint32_t* out_len, const char** bases, const uint64_t** begin) {
  if (!ctx || !bases || !begin) return OTGPU_ERR_ARGS;
  *bases = nullptr;
  *begin = nullptr;
  if (n_reads == 0) return OTGPU_OK;
  if (!reads || !status || !coords || !out_len || (!cigar && n_cigar_ops) || (!seq4 && n_seq_bytes))
    return OTGPU_ERR_ARGS;
  for (uint32_t i = 0; i < n_reads; ++i) {  // every record must lie inside the two arrays
    const otgpu_bam_read& r = reads[i];
    if (r.l_qseq < 0 || r.cigar_off > n_cigar_ops || r.n_cigar > n_cigar_ops - r.cigar_off ||
        r.seq_off > n_seq_bytes || ((uint64_t)r.l_qseq + 1) / 2 > n_seq_bytes - r.seq_off)
      return OTGPU_ERR_ARGS;
  }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  CK(cudaStreamSynchronize(st));
  RS(bam_reads, (size_t)n_reads * sizeof(otgpu_bam_read));
  RS(bam_cigar, n_cigar_ops * 4 + 4);
  RS(bam_seq4, n_seq_bytes + 4);
  RS(bam_status, n_reads);
  RS(bam_coords, (size_t)n_reads * 8);
  RS(bam_plan, (size_t)n_reads * sizeof(ExtractPlan));
  RS(bam_len, (size_t)n_reads * 4);            // per-record output lengths (scan input)
  RS(bam_begin, ((size_t)n_reads + 1) * 8);    // start offsets
  ctx->extracted_reads = 0;
  if (((size_t)n_reads + 1) * 8 > ctx->h_read_begin_cap) {
    if (ctx->h_read_begin) cudaFreeHost(ctx->h_read_begin);
    ctx->h_read_begin = nullptr;
    ctx->h_read_begin_cap = 0;
    const size_t want = ((size_t)n_reads + 1) * 8 + n_reads + 4096;
    CK(cudaMallocHost(&ctx->h_read_begin, want));
    ctx->h_read_begin_cap = want;
  }
  CK(cudaMemcpyAsync(ctx->bam_reads.p, reads, (size_t)n_reads * sizeof(otgpu_bam_read), cudaMemcpyHostToDevice, st));
  if (n_cigar_ops) CK(cudaMemcpyAsync(ctx->bam_cigar.p, cigar, n_cigar_ops * 4, cudaMemcpyHostToDevice, st));
  if (n_seq_bytes) CK(cudaMemcpyAsync(ctx->bam_seq4.p, seq4, n_seq_bytes, cudaMemcpyHostToDevice, st));
  const otgpu_bam_read* d_reads = static_cast<const otgpu_bam_read*>(ctx->bam_reads.p);
  int32_t* d_len = static_cast<int32_t*>(ctx->bam_len.p);
  uint64_t* d_begin = static_cast<uint64_t*>(ctx->bam_begin.p);
  ExtractPlan* d_plan = static_cast<ExtractPlan*>(ctx->bam_plan.p);
  k_bam_breakpoints<<<(n_reads + 255) / 256, 256, 0, st>>>(
      d_reads, n_reads, static_cast<const uint32_t*>(ctx->bam_cigar.p),
      static_cast<uint8_t*>(ctx->bam_status.p), static_cast<int32_t*>(ctx->bam_coords.p), d_plan, d_len);
  k_cigar_scan<<<1, 1024, 0, st>>>(d_len, n_reads, d_begin);
  ctx->stats.kernel_launches += 2;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(ctx->h_read_begin, d_begin, ((size_t)n_reads + 1) * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(status, ctx->bam_status.p, n_reads, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(coords, ctx->bam_coords.p, (size_t)n_reads * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out_len, d_len, (size_t)n_reads * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const uint64_t total = static_cast<const uint64_t*>(ctx->h_read_begin)[n_reads];
  RS(bam_out, total + 16);
  if (total + 1 > ctx->h_reads_cap) {
    if (ctx->h_reads) cudaFreeHost(ctx->h_reads);
    ctx->h_reads = nullptr;
    ctx->h_reads_cap = 0;
    const size_t want = total + total / 4 + 4096;
    CK(cudaMallocHost(&ctx->h_reads, want));
    ctx->h_reads_cap = want;
  }
  if (total) {
    k_bam_decode<<<ctx->n_sm * 8, 256, 0, st>>>(d_reads, n_reads, static_cast<const uint8_t*>(ctx->bam_seq4.p),
                                                d_plan, d_len, d_begin, static_cast<char*>(ctx->bam_out.p));
    ctx->stats.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(ctx->h_reads, ctx->bam_out.p, total, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  *bases = static_cast<const char*>(ctx->h_reads);
  *begin = static_cast<const uint64_t*>(ctx->h_read_begin);
  // the extracted reads stay on the device for otgpu_region_edit_scores_extracted
  try {
    ctx->extracted_len.assign(out_len, out_len + n_reads);
  } catch (...) {
    return OTGPU_ERR_OOM;
  }
  ctx->extracted_reads = n_reads;
  return OTGPU_OK;
}

// Extracted reads -> the resident batch, device to device: read r of the batch is extracted record
// index[r] (one CTA per read; HBM byte copy).
__global__ void __launch_bounds__(128)
k_gather_reads(const char* __restrict__ src, const uint64_t* __restrict__ begin,
               const uint32_t* __restrict__ index, const uint64_t* __restrict__ read_off,
               const uint32_t* __restrict__ read_len, uint8_t* __restrict__ bases) {
  const uint32_t r = blockIdx.x;
  const char* s = src + begin[index[r]];
  uint8_t* d = bases + kPadBases + read_off[r];
  for (uint32_t i = threadIdx.x; i < read_len[r]; i += blockDim.x) d[i] = (uint8_t)s[i];
}

extern "C" int otgpu_region_edit_scores_extracted(otgpu_ctx* ctx, const uint32_t* read_index,
                                                  uint32_t n_reads, const uint32_t* region_start,
                                                  uint32_t n_regions, const uint64_t* score_off,
                                                  uint64_t n_scores, int32_t* scores_out) {
  if (!ctx || (!read_index && n_reads) || !region_start || (!score_off && n_regions) ||
      (!scores_out && n_scores))
    return OTGPU_ERR_ARGS;
  if (ctx->extracted_reads == 0) return OTGPU_ERR_STATE;  // no extraction result on the device
  std::vector<uint64_t> off;
  std::vector<uint32_t> len;
  try {
    off.resize(n_reads);
    len.resize(n_reads);
  } catch (...) {
    return OTGPU_ERR_OOM;
  }
  uint64_t n_bases = 0;
  for (uint32_t r = 0; r < n_reads; ++r) {
    if (read_index[r] >= ctx->extracted_reads) return OTGPU_ERR_ARGS;
    off[r] = n_bases;
    len[r] = ctx->extracted_len[read_index[r]];
    n_bases += len[r];
  }
  ctx->n_chunks_last = 1;
  CK(cudaSetDevice(ctx->device));
  // tables as for any region batch, but the bases never leave the device
  ctx->gather = true;
  int s = otgpu_batch_upload_regions(ctx, nullptr, n_bases, off.data(), len.data(), n_reads, region_start,
                                     n_regions, score_off, n_scores);
  ctx->gather = false;
  if (s != OTGPU_OK) return s;
  if (n_reads) {
    RS(gather_index, (size_t)n_reads * 4);
    CK(cudaMemcpyAsync(ctx->gather_index.p, read_index, (size_t)n_reads * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_gather_reads<<<n_reads, 128, 0, ctx->stream>>>(
        static_cast<const char*>(ctx->bam_out.p), static_cast<const uint64_t*>(ctx->bam_begin.p),
        static_cast<const uint32_t*>(ctx->gather_index.p), static_cast<const uint64_t*>(ctx->read_off.p),
        static_cast<const uint32_t*>(ctx->read_len.p), static_cast<uint8_t*>(ctx->bases.p));
    ctx->stats.kernel_launches++;
    CK(cudaGetLastError());
  }
  s = otgpu_batch_align(ctx);
  if (s == OTGPU_OK) s = otgpu_batch_download(ctx, scores_out, n_scores);
  return s;
}

// ---- kernel density estimates (kde.cuh) ---------------------------------------------------------
extern "C" int otgpu_kde_densities(otgpu_ctx* ctx, const double* values, const uint64_t* value_off,
                                   uint32_t n_regions, const double* bandwidth, const double* grid,
                                   uint32_t n_grid, double* densities_out) {
  if (!ctx) return OTGPU_ERR_ARGS;
  if (n_regions == 0 || n_grid == 0) return OTGPU_OK;
  if (!value_off || !bandwidth || !grid || !densities_out || n_regions > 65535u) return OTGPU_ERR_ARGS;
  const uint64_t n_values = value_off[n_regions];
  if (value_off[0] != 0 || (!values && n_values)) return OTGPU_ERR_ARGS;
  for (uint32_t g = 0; g < n_regions; ++g)
    if (value_off[g + 1] < value_off[g] || !(bandwidth[g] > 0.0)) return OTGPU_ERR_ARGS;
  std::vector<double> inv_h;
  try {
    inv_h.resize(n_regions);
  } catch (...) {
    return OTGPU_ERR_OOM;
  }
  for (uint32_t g = 0; g < n_regions; ++g) inv_h[g] = 1 / bandwidth[g];  // KDE::k_h, ankde.cpp:14-17
  const double pi = 3.14159265358979323846;                             // ankde.hpp:17
  const double c1 = 1 / std::sqrt(2 * pi);                               // KDE::k, ankde.cpp:9-12
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  CK(cudaStreamSynchronize(st));
  RS(kde_values, n_values * 8 + 8);
  RS(kde_off, ((size_t)n_regions + 1) * 8);
  RS(kde_grid, (size_t)n_grid * 8);
  RS(kde_h, (size_t)n_regions * 8);
  RS(kde_inv_h, (size_t)n_regions * 8);
  RS(kde_out, (size_t)n_regions * n_grid * 8);
  if (n_values) CK(cudaMemcpyAsync(ctx->kde_values.p, values, n_values * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->kde_off.p, value_off, ((size_t)n_regions + 1) * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->kde_grid.p, grid, (size_t)n_grid * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->kde_h.p, bandwidth, (size_t)n_regions * 8, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(ctx->kde_inv_h.p, inv_h.data(), (size_t)n_regions * 8, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));  // inv_h leaves scope
  const dim3 grid_dim((n_grid + 127) / 128, n_regions);
  k_kde<<<grid_dim, 128, 0, st>>>(static_cast<const double*>(ctx->kde_values.p),
                                  static_cast<const uint64_t*>(ctx->kde_off.p),
                                  static_cast<const double*>(ctx->kde_grid.p), (int)n_grid,
                                  static_cast<const double*>(ctx->kde_h.p),
                                  static_cast<const double*>(ctx->kde_inv_h.p), c1,
                                  static_cast<double*>(ctx->kde_out.p));
  ctx->stats.kernel_launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(densities_out, ctx->kde_out.p, (size_t)n_regions * n_grid * 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return OTGPU_OK;
}

// ---- partial-order consensus (ppoa.cuh) ----------------------------------------------------------
extern "C" int otgpu_ppoa_consensus(otgpu_ctx* ctx, const uint8_t* bases, uint64_t n_bases,
                                    const uint64_t* read_off, const uint32_t* read_len, uint32_t n_reads,
                                    const otgpu_ppoa_graph* graphs, uint32_t n_graphs,
                                    const otgpu_ppoa_member* members, uint32_t n_members,
                                    const char* cigars, uint64_t n_cigar_bytes, int32_t* consensus_len,
                                    const char** consensus, const uint64_t** begin) {
  if (!ctx || !consensus || !begin) return OTGPU_ERR_ARGS;
  *consensus = nullptr;
  *begin = nullptr;
  if ((!bases && n_bases) || ((!read_off || !read_len) && n_reads) || (!graphs && n_graphs) ||
      (!members && n_members) || (!cigars && n_cigar_bytes) || (!consensus_len && n_graphs))
    return OTGPU_ERR_ARGS;
  for (uint32_t r = 0; r < n_reads; ++r)
    if (read_off[r] + read_len[r] > n_bases) return OTGPU_ERR_ARGS;
  std::vector<PpoaGraphDev> gd;
  std::vector<uint32_t> nonm;
  try {
    gd.resize(n_graphs);
    nonm.resize(n_members);
    ctx->h_ppoa_begin.assign((size_t)n_graphs + 1, 0);
  } catch (...) {
    return OTGPU_ERR_OOM;
  }
  for (uint32_t m = 0; m < n_members; ++m) {
    const otgpu_ppoa_member& M = members[m];
    if (M.read >= n_reads || M.cigar_off > n_cigar_bytes || M.cigar_len > n_cigar_bytes - M.cigar_off)
      return OTGPU_ERR_ARGS;
    uint32_t k = 0;
    const char* c = cigars + M.cigar_off;
    for (uint32_t i = 0; i < M.cigar_len; ++i) k += c[i] != 'M';
    nonm[m] = k;
  }
  // pool sizes: a graph gains at most one node per X / I and at most two edges per non-M operation
  uint64_t nodes = 0, edges = 0, starts = 0;
  for (uint32_t g = 0; g < n_graphs; ++g) {
    const otgpu_ppoa_graph& G = graphs[g];
    if (G.backbone >= n_reads || G.first_member > n_members || G.n_members > n_members - G.first_member)
      return OTGPU_ERR_ARGS;
    uint64_t k = 0;
    for (uint32_t m = 0; m < G.n_members; ++m) k += nonm[G.first_member + m];
    const uint64_t nb = read_len[G.backbone];
    const uint64_t ncap = nb + k + 1, ecap = nb + 2 * k + 2;
    if (ncap >= 0x7fffffffull || ecap >= 0x7fffffffull) return OTGPU_ERR_ARGS;
    PpoaGraphDev& D = gd[g];
    D.backbone = G.backbone; D.first_member = G.first_member; D.n_members = G.n_members;
    D.c = G.c; D.t = G.t;
    D.node_off = nodes; D.edge_off = edges; D.start_off = starts;
    D.node_cap = (uint32_t)ncap; D.edge_cap = (uint32_t)ecap;
    nodes += ncap; edges += ecap; starts += (uint64_t)G.n_members + 2;
  }
  if (n_graphs == 0) {
    *begin = ctx->h_ppoa_begin.data();
    return OTGPU_OK;
  }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  CK(cudaStreamSynchronize(st));
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  // inputs in one buffer
  const size_t i_bases = 0, i_off = i_bases + al(n_bases + 16), i_len = i_off + al((size_t)n_reads * 8 + 8),
               i_graphs = i_len + al((size_t)n_reads * 4 + 4), i_members = i_graphs + al(gd.size() * sizeof(PpoaGraphDev)),
               i_cigars = i_members + al((size_t)n_members * sizeof(PpoaMemberDev) + 16),
               i_cursor = i_cigars + al(n_cigar_bytes + 16), i_lens = i_cursor + 256,
               i_end = i_lens + al((size_t)n_graphs * 4);
  static_assert(sizeof(PpoaMemberDev) == sizeof(otgpu_ppoa_member), "member layout");
  RS(ppoa_in, i_end);
  uint8_t* I = static_cast<uint8_t*>(ctx->ppoa_in.p);
  // pools in one buffer
  PpoaPools P;
  size_t w = 0;
  uint8_t* W = nullptr;
  auto carve = [&](size_t bytes) { const size_t at = w; w += al(bytes); return at; };
  const size_t o_base = carve(nodes), o_ending = carve(nodes), o_done = carve(nodes), o_head = carve(nodes * 4),
               o_tail = carve(nodes * 4), o_inoff = carve((nodes + n_graphs + 1) * 4), o_fill = carve(nodes * 4),
               o_pending = carve(nodes * 4), o_weight = carve(nodes * 4), o_pred = carve(nodes * 4),
               o_queue = carve(nodes * 4), o_esink = carve(edges * 4), o_ew = carve(edges * 4),
               o_enext = carve(edges * 4), o_insrc = carve(edges * 4), o_inpos = carve(edges * 4),
               o_inw = carve(edges * 4), o_starts = carve(starts * 4);
  RS(ppoa_ws, w + 256);
  RS(ppoa_out, nodes + 16);
  W = static_cast<uint8_t*>(ctx->ppoa_ws.p);
  P.base = W + o_base; P.ending = W + o_ending; P.done = W + o_done;
  P.head = reinterpret_cast<int32_t*>(W + o_head); P.tail = reinterpret_cast<int32_t*>(W + o_tail);
  P.in_off = reinterpret_cast<uint32_t*>(W + o_inoff); P.fill = reinterpret_cast<uint32_t*>(W + o_fill);
  P.pending = reinterpret_cast<uint32_t*>(W + o_pending); P.weight = reinterpret_cast<float*>(W + o_weight);
  P.pred = reinterpret_cast<int32_t*>(W + o_pred); P.queue = reinterpret_cast<uint32_t*>(W + o_queue);
  P.e_sink = reinterpret_cast<uint32_t*>(W + o_esink); P.e_w = reinterpret_cast<float*>(W + o_ew);
  P.e_next = reinterpret_cast<int32_t*>(W + o_enext); P.in_src = reinterpret_cast<uint32_t*>(W + o_insrc);
  P.in_pos = reinterpret_cast<uint32_t*>(W + o_inpos); P.in_w = reinterpret_cast<float*>(W + o_inw);
  P.starts = reinterpret_cast<uint32_t*>(W + o_starts);
  if (n_bases) CK(cudaMemcpyAsync(I + i_bases, bases, n_bases, cudaMemcpyHostToDevice, st));
  if (n_reads) {
    CK(cudaMemcpyAsync(I + i_off, read_off, (size_t)n_reads * 8, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(I + i_len, read_len, (size_t)n_reads * 4, cudaMemcpyHostToDevice, st));
  }
  CK(cudaMemcpyAsync(I + i_graphs, gd.data(), gd.size() * sizeof(PpoaGraphDev), cudaMemcpyHostToDevice, st));
  if (n_members)
    CK(cudaMemcpyAsync(I + i_members, members, (size_t)n_members * sizeof(PpoaMemberDev), cudaMemcpyHostToDevice, st));
  if (n_cigar_bytes) CK(cudaMemcpyAsync(I + i_cigars, cigars, n_cigar_bytes, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(I + i_cursor, 0, 256, st));
  const int ctas = (int)std::min<uint64_t>(((uint64_t)n_graphs + 3) / 4, (uint64_t)ctx->n_sm * 8);
  k_ppoa<<<ctas, 128, 0, st>>>(reinterpret_cast<const PpoaGraphDev*>(I + i_graphs), n_graphs,
                               reinterpret_cast<const PpoaMemberDev*>(I + i_members), I + i_bases,
                               reinterpret_cast<const uint64_t*>(I + i_off),
                               reinterpret_cast<const uint32_t*>(I + i_len),
                               reinterpret_cast<const char*>(I + i_cigars), P,
                               reinterpret_cast<unsigned int*>(I + i_cursor),
                               static_cast<char*>(ctx->ppoa_out.p), reinterpret_cast<int32_t*>(I + i_lens));
  ctx->stats.kernel_launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(consensus_len, I + i_lens, (size_t)n_graphs * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  uint64_t total = 0;
  for (uint32_t g = 0; g < n_graphs; ++g) {
    if (consensus_len[g] < 0) {
      snprintf(ctx->err, sizeof(ctx->err), "partial-order graph %u outgrew its pools", g);
      return OTGPU_ERR_STATE;
    }
    ctx->h_ppoa_begin[g] = total;
    total += (uint64_t)consensus_len[g];
  }
  ctx->h_ppoa_begin[n_graphs] = total;
  try {
    ctx->h_ppoa.resize(total + 1);
  } catch (...) {
    return OTGPU_ERR_OOM;
  }
  for (uint32_t g = 0; g < n_graphs; ++g)  // only the strings leave the device
    if (consensus_len[g] > 0)
      CK(cudaMemcpyAsync(ctx->h_ppoa.data() + ctx->h_ppoa_begin[g],
                         static_cast<const char*>(ctx->ppoa_out.p) + gd[g].node_off, (size_t)consensus_len[g],
                         cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  *consensus = ctx->h_ppoa.data();
  *begin = ctx->h_ppoa_begin.data();
  return OTGPU_OK;
}

// ---- edit-distance scores without any allocation (otgpu_edit_scores_static) ----------------------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc.  The same kernels as otgpu_edit_scores (k_pack_reads, k_order_*,
// k_align<1024>, k_align<0>), but the staged batch and all kernel scratch live in the file-scope
// array below instead of the context's cudaMalloc'ed buffers.  In HOST code the array's name is
// only a stand-in for the symbol: its device address is taken once per device with
// cudaGetSymbolAddress (static_setup) and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_otgpu_static_scratch[OTGPU_STATIC_SCRATCH_BYTES];

namespace {

constexpr size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
// layout of the array (byte offsets, each 256-byte aligned)
constexpr size_t kStPackedWords = 2ull * (kPadBases / 16) + OTGPU_STATIC_MAX_BASES / 16 +
                                  8ull * OTGPU_STATIC_MAX_READS + 8;  // per read <= len/16 + 8 words
constexpr int kStWideCtas = 8;                                         // k_align<0> grid
constexpr int kStWideCap = 32768;  // int32 per warp: next_pow2(2 * OTGPU_STATIC_MAX_READ_LEN + 256)
constexpr size_t kStBases = 0;
constexpr size_t kStPacked = kStBases + al256(2ull * kPadBases + OTGPU_STATIC_MAX_BASES + 16);
constexpr size_t kStReadOff = kStPacked + al256(kStPackedWords * 4);
constexpr size_t kStReadLen = kStReadOff + al256(8ull * OTGPU_STATIC_MAX_READS + 8);
constexpr size_t kStPkOff = kStReadLen + al256(4ull * OTGPU_STATIC_MAX_READS + 4);
constexpr size_t kStFlag = kStPkOff + al256(8ull * OTGPU_STATIC_MAX_READS + 8);
constexpr size_t kStPairs = kStFlag + al256(4ull * OTGPU_STATIC_MAX_READS + 4);
constexpr size_t kStScores = kStPairs + al256(sizeof(otgpu_pair) * (size_t)OTGPU_STATIC_MAX_PAIRS + 32);
constexpr size_t kStOrder = kStScores + al256(4ull * OTGPU_STATIC_MAX_PAIRS + 4);
constexpr size_t kStOvf = kStOrder + al256(4ull * OTGPU_STATIC_MAX_PAIRS + 4);
constexpr size_t kStCounters = kStOvf + al256(4ull * OTGPU_STATIC_MAX_PAIRS + 4);
constexpr size_t kStOrderCtl = kStCounters + al256(sizeof(Counters));
constexpr size_t kStWide = kStOrderCtl + al256(sizeof(OrderCtl));
constexpr size_t kStEnd = kStWide + al256((size_t)kStWideCtas * kWarpsPerCta * kStWideCap * 4);
static_assert(kStEnd <= OTGPU_STATIC_SCRATCH_BYTES, "static scratch layout exceeds the array");
static_assert(2 * OTGPU_STATIC_MAX_READ_LEN + 256 <= kStWideCap, "wide ring too small for the longest read");
static_assert(OTGPU_STATIC_MAX_READ_LEN <= kMaxReadLen, "read length limit");

struct StaticDev {
  bool ready = false;
  unsigned char* base = nullptr;  // device address of g_otgpu_static_scratch on this device
  int n_sm = 0, grid_main = 0;
};
StaticDev g_static_dev[64];
std::mutex g_static_mu[64];                              // one call at a time per device
uint64_t g_static_pk[64][OTGPU_STATIC_MAX_READS + 1];    // host: packed-word offsets of the call

#define CKS(call)                                         \
  do {                                                    \
    if ((call) != cudaSuccess) {                          \
      cudaGetLastError();                                 \
      return OTGPU_ERR_CUDA;                              \
    }                                                     \
  } while (0)

// Once per device, with that device current.
int static_setup(int device, StaticDev& d) {
  if (d.ready) return OTGPU_OK;
  void* p = nullptr;
  CKS(cudaGetSymbolAddress(&p, g_otgpu_static_scratch));
  if (!p) return OTGPU_ERR_CUDA;
  CKS(cudaDeviceGetAttribute(&d.n_sm, cudaDevAttrMultiProcessorCount, device));
  int per_sm = 0;
  CKS(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_align<1024>, kCtaThreads, ring_bytes(1024)));
  if (per_sm < 1) per_sm = 1;
  d.grid_main = d.n_sm * per_sm;
  d.base = static_cast<unsigned char*>(p);
  d.ready = true;
  return OTGPU_OK;
}

int static_run(int device, cudaStream_t st, const uint8_t* bases, uint64_t n_bases, const uint64_t* read_off,
               const uint32_t* read_len, uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
               int32_t* scores_out) {
  StaticDev& d = g_static_dev[device];
  if (int s = static_setup(device, d)) return s;
  unsigned char* S = d.base;
  // packed layout, as upload_reads does it: guard words either side, reads 16-byte aligned
  uint64_t* pk = g_static_pk[device];
  uint64_t words = kPadBases / 16;
  uint32_t max_len = 0;
  for (uint32_t r = 0; r < n_reads; ++r) {
    pk[r] = words;
    words += (((uint64_t)read_len[r] + 15) / 16 + 2 + 3) & ~3ull;
    if (read_len[r] > max_len) max_len = read_len[r];
  }
  if (words + kPadBases / 16 + 4 > kStPackedWords) return OTGPU_ERR_LIMIT;  // (cannot happen within the limits)
  DevBatch b;
  b.bases = S + kStBases;
  b.read_off = reinterpret_cast<const uint64_t*>(S + kStReadOff);
  b.read_len = reinterpret_cast<const uint32_t*>(S + kStReadLen);
  b.pk_off = reinterpret_cast<const uint64_t*>(S + kStPkOff);
  b.packed = reinterpret_cast<const uint32_t*>(S + kStPacked);
  b.read_flag = reinterpret_cast<uint32_t*>(S + kStFlag);
  b.pairs = reinterpret_cast<const otgpu_pair*>(S + kStPairs);
  b.scores = reinterpret_cast<int32_t*>(S + kStScores);
  b.n_pairs = n_pairs;
  b.n_reads = n_reads;
  Counters* ctr = reinterpret_cast<Counters*>(S + kStCounters);
  OrderCtl* ctl = reinterpret_cast<OrderCtl*>(S + kStOrderCtl);
  uint32_t* order = reinterpret_cast<uint32_t*>(S + kStOrder);
  uint32_t* ovf = reinterpret_cast<uint32_t*>(S + kStOvf);
  int32_t* wide = reinterpret_cast<int32_t*>(S + kStWide);
  // inputs into the array (its guard regions keep whatever they held: their content is never used)
  if (n_bases) CKS(cudaMemcpyAsync(S + kStBases + kPadBases, bases, n_bases, cudaMemcpyHostToDevice, st));
  CKS(cudaMemcpyAsync(S + kStReadOff, read_off, (size_t)n_reads * 8, cudaMemcpyHostToDevice, st));
  CKS(cudaMemcpyAsync(S + kStReadLen, read_len, (size_t)n_reads * 4, cudaMemcpyHostToDevice, st));
  CKS(cudaMemcpyAsync(S + kStPkOff, pk, (size_t)n_reads * 8, cudaMemcpyHostToDevice, st));
  CKS(cudaMemcpyAsync(S + kStPairs, pairs, (size_t)n_pairs * sizeof(otgpu_pair), cudaMemcpyHostToDevice, st));
  // what must start at zero: the array keeps the last call's contents
  CKS(cudaMemsetAsync(ctr, 0, sizeof(Counters), st));
  CKS(cudaMemsetAsync(ctl, 0, sizeof(OrderCtl), st));
  CKS(cudaMemsetAsync(b.read_flag, 0, (size_t)n_reads * 4, st));
#ifdef OTGPU_CHECK_BOUNDS
  {
    CheckRange r[2];
    r[0].lo = b.packed;
    r[0].hi = b.packed + kStPackedWords;
    r[1].lo = reinterpret_cast<const uint32_t*>(S + kStBases);
    r[1].hi = r[1].lo + (kStPacked - kStBases) / 4;
    CKS(cudaMemcpyToSymbolAsync(otgpu_check_range, r, sizeof r, 0, cudaMemcpyHostToDevice, st));
  }
#endif
  k_pack_reads<<<n_reads, 128, 0, st>>>(b, reinterpret_cast<uint32_t*>(S + kStPacked));
  const int g = (int)((n_pairs + 255) / 256 < (uint64_t)d.n_sm * 8 ? (n_pairs + 255) / 256 : (uint64_t)d.n_sm * 8);
  k_order_count<<<g, 256, 0, st>>>(b.pairs, n_pairs, b.read_len, ctl);
  k_order_scan<<<1, 1, 0, st>>>(ctl);
  k_order_scatter<<<g, 256, 0, st>>>(b.pairs, n_pairs, b.read_len, ctl, order);
  // 1024-slot shared ring for every pair; the ones that outgrow it finish on the HBM rings
  int wide_cap = next_pow2(2 * (int)max_len + 256);
  if (wide_cap > kStWideCap) return OTGPU_ERR_LIMIT;
  k_align<1024><<<d.grid_main, kCtaThreads, ring_bytes(1024), st>>>(b, ctr, order, -1, 0, ovf, nullptr, wide_cap);
  k_align<0><<<kStWideCtas, kCtaThreads, 0, st>>>(b, ctr, ovf, 1, 2, ovf, wide, wide_cap);
  CKS(cudaGetLastError());
  CKS(cudaMemcpyAsync(scores_out, b.scores, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, st));
  CKS(cudaStreamSynchronize(st));
  return OTGPU_OK;
}

}  // namespace

extern "C" int otgpu_edit_scores_static(int device, void* stream, const uint8_t* bases, uint64_t n_bases,
                                        const uint64_t* read_off, const uint32_t* read_len,
                                        uint32_t n_reads, const otgpu_pair* pairs, uint64_t n_pairs,
                                        int32_t* scores_out) {
  if ((!bases && n_bases) || ((!read_off || !read_len) && n_reads) || (!pairs && n_pairs) ||
      (!scores_out && n_pairs))
    return OTGPU_ERR_ARGS;
  // the fixed limits of the static array come first: nothing is launched for a batch beyond them
  if (n_bases > OTGPU_STATIC_MAX_BASES || n_reads > OTGPU_STATIC_MAX_READS || n_pairs > OTGPU_STATIC_MAX_PAIRS)
    return OTGPU_ERR_LIMIT;
  for (uint32_t r = 0; r < n_reads; ++r) {
    if (read_len[r] > OTGPU_STATIC_MAX_READ_LEN) return OTGPU_ERR_LIMIT;
    if (read_off[r] + read_len[r] > n_bases) return OTGPU_ERR_ARGS;
  }
  for (uint64_t i = 0; i < n_pairs; ++i) {  // as otgpu_batch_upload_pairs
    const otgpu_pair& p = pairs[i];
    if (p.pattern >= n_reads || p.text >= n_reads) return OTGPU_ERR_ARGS;
    if (p.pattern_begin_free >= 0) {
      if (p.pattern_end_free < 0 || p.text_begin_free < 0 || p.text_end_free < 0 ||
          (uint32_t)p.pattern_begin_free > read_len[p.pattern] ||
          (uint32_t)p.pattern_end_free > read_len[p.pattern] ||
          (uint32_t)p.text_begin_free > read_len[p.text] || (uint32_t)p.text_end_free > read_len[p.text])
        return OTGPU_ERR_ARGS;
    }
  }
  if (device < 0 || device >= 64 || device >= otgpu_device_count()) return OTGPU_ERR_NO_DEVICE;
  if (n_pairs == 0) return OTGPU_OK;
  std::lock_guard<std::mutex> lk(g_static_mu[device]);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); return OTGPU_ERR_CUDA; }
  if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return OTGPU_ERR_CUDA; }
  const int s = static_run(device, static_cast<cudaStream_t>(stream), bases, n_bases, read_off, read_len,
                           n_reads, pairs, n_pairs, scores_out);
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return s;
}
