// otter_b200/csrc/wfa_edit.cuh — warp-per-pair wavefront edit distance for sm_100a.
//
// Computes exactly what the reference computes for one pair with
//   wfa::WFAlignerEdit(Score, MemoryMed)  (WFA2-lib v2.2 bindings/cpp/WFAligner.cpp:252-258)
// i.e. edit-distance WFA under the default WF-adaptive(10,50,1) cut-off:
//   extend        wavefront/wavefront_extend.c:172-260,355-452
//   termination   wavefront/wavefront_extend.c:45-68 (end2end), :122-168 (ends-free)
//   cut-off       wavefront/wavefront_heuristic.c:111-119,223-328,470-502
//   compute       wavefront/wavefront_compute_edit.c:70-97,294-356
//   trim          wavefront/wavefront_compute.c:376-409
//   exact prune   wavefront/wavefront_compute_edit.c:214-269
// but organised for a 32-lane warp instead of a scalar loop.  ncu shows the kernel is
// instruction-issue bound (profiles/), so the design goal is fewest warp instructions per live
// diagonal per score step:
//   * diagonal k lives at ring slot (k & mask) of a per-warp int32 ring (shared memory, or HBM
//     for the rare pair whose live width exceeds the ring) — no re-basing when lo/hi move and
//     bank = k mod 32, so every pass is conflict-free;
//   * ONE fused pass per score step: each lane computes its next offset from the k-1/k/k+1
//     neighbours (warp shuffles, written in place), extends it at once, and accumulates the
//     cut-off distance minimum and the trim limits in registers (REDUX once per step);
//   * the cut-off distance of diagonal k is plen + max(k, tlen-plen) - offset, and the end-to-end
//     alignment is finished exactly when that distance reaches 0;
//   * extension compares 64-bit windows of packed bases (32 bases at 2 bit, 8 at 8 bit) fetched
//     with three aligned loads + two funnel shifts; a run that outlives its private windows is
//     handed to the whole warp (32 windows per round, __ballot_sync + __ffs for the LCP).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace otgpu {

constexpr int kNull = -(1 << 30);  // WAVEFRONT_OFFSET_NULL, wavefront_offset.h:46
constexpr unsigned kFull = 0xffffffffu;
constexpr int kMinWfLen = 10;   // wavefront_attributes.c:74
constexpr int kMaxDist = 50;    // wavefront_attributes.c:75
constexpr int kPruneLen = 1000; // wavefront_compute_edit.c:224
constexpr int kScoreOverflow = -2;  // internal: live width exceeded the shared-memory ring
constexpr int kBigDist = 1 << 30;   // distance of a null offset, wavefront_heuristic.c:118
constexpr int kRingCap = 1024;      // int32 slots per warp in shared memory (4 KiB, 4 KiB aligned)

// ---- wavefront ring ---------------------------------------------------------------------------
__device__ __forceinline__ int lds32(unsigned a) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(unsigned a, int v) {
  asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

template <bool WIDE>
struct Ring;
template <>
struct Ring<false> {  // shared memory, fixed capacity, 32-bit shared-window addressing
  unsigned base;      // aligned to the ring size, so slot address = base | ((k*4) & mask): one LOP3
  static constexpr unsigned kMaskB = (kRingCap - 1) * 4;
  __device__ __forceinline__ int cap() const { return kRingCap; }
  __device__ __forceinline__ unsigned addr(int k) const { return base | (((unsigned)k << 2) & kMaskB); }
  __device__ __forceinline__ int ld(int k) const { return lds32(addr(k)); }
  __device__ __forceinline__ void st(int k, int v) const { sts32(addr(k), v); }
};
template <>
struct Ring<true> {  // HBM workspace (L1/L2 cached), capacity covers every diagonal of the pair
  int32_t* p;
  int mask;
  __device__ __forceinline__ int cap() const { return mask + 1; }
  __device__ __forceinline__ int ld(int k) const { return p[k & mask]; }
  __device__ __forceinline__ void st(int k, int v) const { p[k & mask] = v; }
};

// ---- packed reads -----------------------------------------------------------------------------
// A read as the kernel sees it: 32-bit word index of its first word inside the batch buffer
// (2-bit words, or the raw byte buffer viewed as words) + position of base 0 inside that word.
// (Index + one warp-uniform base pointer costs two instructions per window address, LEA.HI +
// IMAD.WIDE; a per-read 64-bit pointer costs five.)
struct SeqView {
  uint32_t word0;
  int sh0;  // always 0 for 2-bit reads (16-byte aligned); 0..3 for raw bytes
};

template <int BITS>
struct Pack {
  static constexpr int kBpw = 32 / BITS;  // bases per 32-bit word
  static constexpr int kLogBpw = (BITS == 2) ? 4 : (BITS == 4 ? 3 : 2);
  static constexpr int kLogBits = (BITS == 2) ? 1 : (BITS == 4 ? 2 : 3);
  static constexpr int kWin = 2 * kBpw;  // bases per 64-bit window
};

// 64-bit window of packed bases starting at base position p: three aligned 32-bit loads + two
// funnel shifts (shf.wrap takes the shift modulo 32, so no masking of the bit offset).
// Touches up to two words past the word holding p; buffers carry guards for that.
template <int BITS>
__device__ __forceinline__ void window64(const uint32_t* __restrict__ words, const SeqView& s,
                                         int p, bool on, uint32_t& lo, uint32_t& hi) {
  const int a = (BITS == 2) ? p : s.sh0 + p;
  const uint32_t* q = words + (s.word0 + (uint32_t)(a >> Pack<BITS>::kLogBpw));
  uint32_t w0 = 0, w1 = 0, w2 = 0;
  if (on) { w0 = __ldg(q); w1 = __ldg(q + 1); w2 = __ldg(q + 2); }  // predicated loads
  const int sh = a << Pack<BITS>::kLogBits;
  lo = __funnelshift_r(w0, w1, sh);
  hi = __funnelshift_r(w1, w2, sh);
}

// Matching bases at the front of P[v..] and T[h..] within one 64-bit window (kWin if identical).
template <int BITS>
__device__ __forceinline__ int match_window(const uint32_t* __restrict__ words, const SeqView& P,
                                            const SeqView& T, int v, int h, bool on = true) {
  uint32_t pl, ph, tl, th;
  window64<BITS>(words, P, v, on, pl, ph);
  window64<BITS>(words, T, h, on, tl, th);
  const uint32_t xl = pl ^ tl, xh = ph ^ th;
  const uint32_t sel = xl ? xl : xh;
  const int bit = (__ffs(sel) - 1) + (xl ? 0 : 32);
  return sel ? (bit >> Pack<BITS>::kLogBits) : Pack<BITS>::kWin;
}

struct EditJob {
  const uint32_t* __restrict__ words;  // packed 2-bit words, or the raw bytes viewed as words
  SeqView P, T;
  int plen, tlen;
  int pb, pe, tb, te;  // pb < 0 => end-to-end
};

// Longest common prefix of P[v..v+rem) and T[h..h+rem) for every lane at once (rem <= 0 for lanes
// with nothing to extend: they issue no loads).  All 32 lanes must call.
// First window is unconditional; after that, private windows while several lanes are still
// running, and once at most two are left each remaining run is finished by the whole warp.
template <int BITS>
__device__ __forceinline__ int lcp_warp(const EditJob& J, int v, int h, int rem, const int lane) {
  constexpr int WIN = Pack<BITS>::kWin;
  int m = match_window<BITS>(J.words, J.P, J.T, v, h, rem > 0);
  int n = max(min(m, rem), 0);  // clamp at a sequence end (the reference stops on its sentinels)
  bool cont = (m >= WIN) && (rem > WIN);
  unsigned pend = __ballot_sync(kFull, cont);
  if (pend == 0) return n;
  for (int it = 0;; ++it) {
    if (cont) {
      v += WIN; h += WIN; rem -= WIN;
      m = match_window<BITS>(J.words, J.P, J.T, v, h);
      n += min(m, rem);
      cont = (m >= WIN) && (rem > WIN);
    }
    pend = __ballot_sync(kFull, cont);
    if (pend == 0) return n;
    if (__popc(pend) <= 2 || it >= 3) break;
  }
  while (pend) {  // warp-uniform: one pending run at a time, 32 windows per round
    const int src = __ffs(pend) - 1;
    pend &= pend - 1;
    int bv = __shfl_sync(kFull, v, src) + WIN;
    int bh = __shfl_sync(kFull, h, src) + WIN;
    int brem = __shfl_sync(kFull, rem, src) - WIN;  // > 0
    int got = 0;
    for (;;) {
      const int o = lane * WIN;
      int mm = 0;
      bool stop = true;
      if (o < brem) {
        mm = match_window<BITS>(J.words, J.P, J.T, bv + o, bh + o);
        const int lim = brem - o;
        if (mm >= lim) mm = lim;
        else stop = mm < WIN;
      }
      const unsigned b = __ballot_sync(kFull, stop);
      if (b) {
        const int f = __ffs(b) - 1;
        got += f * WIN + __shfl_sync(kFull, mm, f);
        break;
      }
      got += 32 * WIN; bv += 32 * WIN; bh += 32 * WIN; brem -= 32 * WIN;
    }
    if (lane == src) n += got;
  }
  return n;
}

// Warp-uniform state of one alignment (lo/hi) + per-lane partials of the last pass.
struct WfState {
  int lo, hi;  // live diagonal range
  int dmin;    // per-lane partial minimum of the cut-off distance
  bool fin;    // per-lane: ends-free termination met on a diagonal this lane owns
};

// One pass over the wavefront.
//   COMPUTE: build wavefront s+1 over [lo-1,hi+1] from wavefront s in place and trim it
//            (wavefront_compute_edit.c:70-97 + wavefront_compute.c:376-409);
//   EXTEND : extend every live offset, gather the cut-off distance minimum and (ends-free) the
//            termination flag (wavefront_extend.c:203-260,122-168; wavefront_heuristic.c:223-239).
// Within a pass a ring slot is only ever touched by the lane that owns it, so no intra-pass
// synchronisation is needed; passes are separated by __syncwarp().
template <int BITS, bool WIDE, bool EF, bool COMPUTE, bool EXTEND>
__device__ __forceinline__ void wf_pass(const Ring<WIDE>& R, const EditJob& J, const int lane,
                                        WfState& S) {
  const int plen = J.plen, tlen = J.tlen;
  const int ak = tlen - plen;
  const int lo = S.lo, hi = S.hi;
  const unsigned span = (unsigned)(hi - lo);
  const int clo = COMPUTE ? lo - 1 : lo, chi = COMPUTE ? hi + 1 : hi;
  int kmin = 0x7fffffff, kmax = (int)0x80000000;  // per-lane trimmed limits
  int carry = kNull;                              // old M[kc-1]
  int dmin = kBigDist;
  bool fin = false;
  int k = clo + lane;
  int mid = R.ld(k);
  if ((unsigned)(k - lo) > span) mid = kNull;
#ifdef OTGPU_UNROLL2
#pragma unroll 2
#endif
  for (int kc = clo; kc <= chi; kc += 32, k += 32) {
    int m, nxt = R.ld(k + 32);  // this lane's slot in the next chunk (still old)
    if ((unsigned)(k + 32 - lo) > span) nxt = kNull;
    if (COMPUTE) {
      int left = __shfl_up_sync(kFull, mid, 1);
      int right = __shfl_down_sync(kFull, mid, 1);
      const int edge = __shfl_sync(kFull, nxt, 0);
      const int last = __shfl_sync(kFull, mid, 31);
      if (lane == 0) left = carry;
      if (lane == 31) right = edge;
      carry = last;
      m = max(right, max(left, mid) + 1);
    } else {
      m = mid;
    }
    mid = nxt;
    // in-bounds <=> m >= 0, v >= 0, and neither sequence overrun (lanes beyond chi see only
    // nulls and fail the first test)
    const int v = m - k;
    const int rem = min(plen - v, tlen - m);
    const bool ok = (m | v | rem) >= 0;
    if (EXTEND) {
      m += lcp_warp<BITS>(J, v, m, ok ? rem : 0, lane);
      if (ok) {
        dmin = min(dmin, plen + max(k, ak) - m);  // = max(plen - v', tlen - m')
        if (EF) {
          const int left_v = plen - (m - k), left_h = tlen - m;
          fin |= (left_h <= 0 && left_v <= J.pe) || (left_v <= 0 && left_h <= J.te);
        }
      }
    }
    R.st(k, ok ? m : kNull);  // slots just beyond chi get nulls: never read (range-guarded)
    if (COMPUTE && ok) {
      kmin = min(kmin, k);
      kmax = max(kmax, k);
    }
  }
  __syncwarp();  // the next pass maps diagonals to lanes differently
  if (COMPUTE) {
    S.lo = __reduce_min_sync(kFull, kmin);
    S.hi = __reduce_max_sync(kFull, kmax);
  }
  if (EXTEND) { S.dmin = dmin; S.fin = fin; }
}

// wf_compute_distance_end2end, wavefront_heuristic.c:111-119
__device__ __forceinline__ int wf_dist(int off, int k, int plen, int ak) {
  return off >= 0 ? plen + max(k, ak) - off : kBigDist;
}

// WF-adaptive cut-off (wavefront_heuristic.c:285-328), run after every extend.
// dmin = warp-wide minimum distance of the live wavefront.
template <bool WIDE>
__device__ __forceinline__ void wf_cutoff(const Ring<WIDE>& R, const int plen, const int tlen,
                                          const int lane, int dmin, WfState& S) {
  const int lo = S.lo, hi = S.hi, ak = tlen - plen;
  dmin = min(max(plen, tlen), dmin);  // :230-237
  const int top = min(ak, hi);        // :269
  int nlo = lo;
  if (lo < top) {
    nlo = top;
    for (int kc = lo; kc < top; kc += 32) {
      const int k = kc + lane;
      const int d = wf_dist(R.ld(k), k, plen, ak);
      const unsigned b = __ballot_sync(kFull, k < top && (d - dmin <= kMaxDist));
      if (b) { nlo = kc + __ffs(b) - 1; break; }
    }
  }
  const int bot = max(ak, nlo);  // :277 (uses the already reduced lo)
  int nhi = hi;
  if (hi > bot) {
    nhi = bot;
    for (int kc = hi; kc > bot; kc -= 32) {
      const int k = kc - lane;
      const int d = wf_dist(R.ld(k), k, plen, ak);
      const unsigned b = __ballot_sync(kFull, k > bot && (d - dmin <= kMaxDist));
      if (b) { nhi = kc - (__ffs(b) - 1); break; }
    }
  }
  S.lo = nlo;
  S.hi = nhi;
}

// Exact prune, end-to-end only, width >= 1000 (wavefront_compute_edit.c:214-269).
// Returns false if nothing survives (cannot happen: the minimum is attained inside [lo,hi]).
template <bool WIDE>
__device__ __forceinline__ bool wf_prune(const Ring<WIDE>& R, const int plen, const int tlen,
                                         const int lane, WfState& S) {
  const int lo = S.lo, hi = S.hi, ak = tlen - plen;
  if (hi - lo + 1 < kPruneLen) return true;
  const int ks = lo + (hi - lo) / 2;
  const int so = R.ld(ks);
  if (so < 0) return true;
  const int smax = max(plen - (so - ks), tlen - so);
  // best-case score of diagonal k is |k - ak| whatever its offset (:193-201)
  if (abs(lo - ak) <= smax && abs(hi - ak) <= smax) return true;
  int w = 0x7fffffff;
  for (int k = lo + lane; k <= hi; k += 32) {
    const int off = R.ld(k);
    if (off >= 0) w = min(w, max(plen - (off - k), tlen - off));
  }
  w = __reduce_min_sync(kFull, w);
  const int nlo = max(lo, ak - w);  // first k in [lo,hi] with |k-ak| <= w
  if (nlo > min(hi, ak + w)) return false;
  int nhi = min(hi, ak + w);  // last k in (nlo,hi] with |k-ak| <= w, else nlo
  if (nhi <= nlo) nhi = nlo;
  S.lo = nlo;
  S.hi = nhi;
  return true;
}

// One pair on one warp.  Returns score >= 0, OTGPU_SCORE_UNFEASIBLE (-1), or kScoreOverflow
// (-2, never when WIDE).
template <int BITS, bool WIDE, bool EF>
__device__ int wfa_edit_run(const Ring<WIDE>& R, const EditJob& J, const int lane) {
  const int plen = J.plen, tlen = J.tlen;
  const int cap = R.cap();
  WfState S;
  int s = 0;

  // ---- wavefront 0 (wavefront_align.c:113-135 / :190-237) ----
  if (!EF) {
    S.lo = S.hi = 0;
    if (lane == 0) R.st(0, 0);
  } else {
    S.lo = -J.pb;
    S.hi = J.tb;
    if (S.hi - S.lo + 40 > cap) return kScoreOverflow;
    for (int k = S.lo + lane; k <= S.hi; k += 32) R.st(k, k > 0 ? k : 0);
  }
  __syncwarp();
  wf_pass<BITS, WIDE, EF, false, true>(R, J, lane, S);

  for (;;) {
    // termination: ends-free per-diagonal flags (wavefront_extend.c:122-168); end-to-end is done
    // iff the cut-off distance hit 0, i.e. offset tlen on diagonal tlen-plen (:45-68)
    if (__any_sync(kFull, EF ? S.fin : (S.dmin == 0))) return s;
    if (S.hi - S.lo + 1 >= kMinWfLen)
      wf_cutoff<WIDE>(R, plen, tlen, lane, __reduce_min_sync(kFull, S.dmin), S);
    ++s;
    // the pass touches slots [lo-1, hi+1+31]; they must not alias inside the ring
    if (!WIDE && S.hi - S.lo + 40 > cap) return kScoreOverflow;
    if (WIDE && !EF && S.hi - S.lo + 3 >= kPruneLen) {
      // the exact prune sits between compute and extend: run them as separate passes
      wf_pass<BITS, WIDE, EF, true, false>(R, J, lane, S);
      if (S.lo > S.hi) return -1;
      if (!wf_prune<WIDE>(R, plen, tlen, lane, S)) return -1;
      wf_pass<BITS, WIDE, EF, false, true>(R, J, lane, S);
    } else {
      wf_pass<BITS, WIDE, EF, true, true>(R, J, lane, S);
      if (S.lo > S.hi) return -1;  // empty wavefront -> WF_STATUS_UNFEASIBLE
    }
  }
}

template <int BITS, bool WIDE>
__device__ __forceinline__ int wfa_edit_pair(const Ring<WIDE>& R, const EditJob& J, const int lane) {
  return J.pb >= 0 ? wfa_edit_run<BITS, WIDE, true>(R, J, lane)
                   : wfa_edit_run<BITS, WIDE, false>(R, J, lane);
}

}  // namespace otgpu
