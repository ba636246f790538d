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
// but organised for a 32-lane warp instead of a scalar loop:
//   * diagonal k lives at ring slot (k & mask) of a per-warp int32 ring (shared memory, or HBM
//     for the rare pair whose live width exceeds the ring) — no re-basing when lo/hi move,
//     bank = k mod 32 so every pass is conflict-free;
//   * lanes own 32 consecutive diagonals per chunk; k±1 neighbours come from warp shuffles and
//     the next wavefront is written in place (one carried value per chunk);
//   * extension compares 32-bit windows of packed bases (16 bases at 2 bit, 4 at 8 bit) fetched
//     with two aligned loads + funnel shift; a lane whose run outlives its private windows hands
//     the run to the whole warp (32 windows per round, __ballot_sync + __ffs for the LCP);
//   * cut-off = REDUX min + two ballot scans; trim falls out of the compute pass's ballots.
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

// A read as the kernel sees it: word pointer + index (in bases) of position 0 inside word 0.
struct SeqView {
  const uint32_t* __restrict__ w;
  int sh0;
};

template <int BITS>
struct Pack {
  static constexpr int kBpw = 32 / BITS;                   // bases per 32-bit window
  static constexpr int kLogBpw = (BITS == 2) ? 4 : (BITS == 4 ? 3 : 2);
  static constexpr int kLogBits = (BITS == 2) ? 1 : (BITS == 4 ? 2 : 3);
};

// 32-bit window of packed bases starting at base position p (two aligned loads + funnel shift).
// Reads at most one word past the word holding p; buffers carry guard words for that.
template <int BITS>
__device__ __forceinline__ uint32_t window32(const SeqView& s, int p) {
  const int a = s.sh0 + p;
  const int i = a >> Pack<BITS>::kLogBpw;
  const int sh = (a & (Pack<BITS>::kBpw - 1)) << Pack<BITS>::kLogBits;
  const uint32_t w0 = __ldg(s.w + i);
  const uint32_t w1 = __ldg(s.w + i + 1);
  return __funnelshift_r(w0, w1, sh);
}

__device__ __forceinline__ bool in_bounds(int off, int k, int plen, int tlen) {
  return (unsigned)off <= (unsigned)tlen && (unsigned)(off - k) <= (unsigned)plen;
}

// wf_compute_distance_end2end, wavefront_heuristic.c:111-119
__device__ __forceinline__ int wf_dist(int off, int k, int plen, int tlen) {
  return off >= 0 ? max(plen - (off - k), tlen - off) : (1 << 30);
}

// Longest common prefix of P[v..plen) and T[h..tlen), warp-cooperative for long runs.
// Every lane of the warp must call this (inactive lanes pass rem = 0).
// kPrivate = windows a lane checks on its own before asking the warp for help.
template <int BITS, int kPrivate>
__device__ __forceinline__ int lcp_warp(const SeqView& P, const SeqView& T, int v, int h, int rem,
                                        int lane) {
  constexpr int BPW = Pack<BITS>::kBpw;
  constexpr int LB = Pack<BITS>::kLogBits;
  int n = 0;
  bool cont = rem > 0;
#pragma unroll
  for (int it = 0; it < kPrivate; ++it) {
    if (cont) {
      const uint32_t x = window32<BITS>(P, v) ^ window32<BITS>(T, h);
      const int m = x ? ((__ffs(x) - 1) >> LB) : BPW;
      if (m >= rem) {  // ran into a sequence end (the reference stops on its sentinels)
        n += rem;
        cont = false;
      } else {
        n += m;
        if (m < BPW) cont = false;
        else { rem -= BPW; v += BPW; h += BPW; }
      }
    }
  }
  unsigned pend = __ballot_sync(kFull, cont);
  while (pend) {  // warp-uniform
    const int src = __ffs(pend) - 1;
    pend &= pend - 1;
    int bv = __shfl_sync(kFull, v, src);
    int bh = __shfl_sync(kFull, h, src);
    int brem = __shfl_sync(kFull, rem, src);
    int got = 0;
    for (;;) {
      const int o = lane * BPW;
      int m = 0;
      bool stop = true;
      if (o < brem) {
        const uint32_t x = window32<BITS>(P, bv + o) ^ window32<BITS>(T, bh + o);
        m = x ? ((__ffs(x) - 1) >> LB) : BPW;
        const int lim = brem - o;
        if (m >= lim) m = lim;
        else stop = m < BPW;
      }
      const unsigned b = __ballot_sync(kFull, stop);
      if (b) {
        const int f = __ffs(b) - 1;
        got += f * BPW + __shfl_sync(kFull, m, f);
        break;
      }
      got += 32 * BPW; bv += 32 * BPW; bh += 32 * BPW; brem -= 32 * BPW;
    }
    if (lane == src) n += got;
  }
  return n;
}

struct EditJob {
  SeqView P, T;
  int plen, tlen;
  int pb, pe, tb, te;  // pb < 0 => end-to-end
};

// One pair on one warp.  wf = ring of (mask+1) int32 slots private to the warp.
// Returns score >= 0, OTGPU_SCORE_UNFEASIBLE (-1), or kScoreOverflow (-2, never when WIDE).
template <int BITS, bool WIDE>
__device__ int wfa_edit_pair(int32_t* __restrict__ wf, const int mask, const EditJob& J,
                             const int lane) {
  const int plen = J.plen, tlen = J.tlen;
  const bool ef = J.pb >= 0;
  const int pe = J.pe, te = J.te;
  const int ak = tlen - plen;  // target diagonal
  const int cap = mask + 1;
  int lo, hi, s = 0;

  // ---- wavefront 0 (wavefront_align.c:113-135 / :190-237) ----
  if (!ef) {
    lo = hi = 0;
    if (lane == 0) wf[0] = 0;
  } else {
    lo = -J.pb;
    hi = J.tb;
    if (hi - lo + 3 > cap) return kScoreOverflow;
    for (int k = lo + lane; k <= hi; k += 32) wf[k & mask] = k > 0 ? k : 0;
  }
  __syncwarp();

  for (;;) {
    // ---- extend every live diagonal; collect the cut-off distance minimum on the way ----
    int dmin = max(plen, tlen);
    bool fin = false;
    for (int kc = lo; kc <= hi; kc += 32) {
      const int k = kc + lane;
      int off = (k <= hi) ? wf[k & mask] : kNull;
      const bool live = off >= 0;  // stored offsets are either in-bounds (>= 0) or kNull
      const int v = off - k;
      const int rem = live ? min(plen - v, tlen - off) : 0;
      const int n = lcp_warp<BITS, 2>(J.P, J.T, v, off, rem, lane);
      if (live) {
        off += n;
        wf[k & mask] = off;
        dmin = min(dmin, max(plen - (off - k), tlen - off));
        if (ef) {  // wavefront_extend.c:122-168
          const int vv = off - k;
          fin |= (off >= tlen && plen - vv <= pe) || (vv >= plen && tlen - off <= te);
        }
      }
    }
    __syncwarp();
    if (ef) {
      if (__any_sync(kFull, fin)) return s;
    } else if (lo <= ak && ak <= hi) {  // wavefront_extend.c:45-68
      if (wf[ak & mask] >= tlen) return s;
    }

    // ---- WF-adaptive cut-off (wavefront_heuristic.c:285-328), every step ----
    if (hi - lo + 1 >= kMinWfLen) {
      dmin = __reduce_min_sync(kFull, dmin);
      const int top = min(ak, hi);
      int nlo = lo;
      if (lo < top) {
        nlo = top;
        for (int kc = lo; kc < top; kc += 32) {
          const int k = kc + lane;
          const bool in = k < top;
          const int d = in ? wf_dist(wf[k & mask], k, plen, tlen) : 0;
          const unsigned b = __ballot_sync(kFull, in && (d - dmin <= kMaxDist));
          if (b) { nlo = kc + __ffs(b) - 1; break; }
        }
      }
      const int bot = max(ak, nlo);
      int nhi = hi;
      if (hi > bot) {
        nhi = bot;
        for (int kc = hi; kc > bot; kc -= 32) {
          const int k = kc - lane;
          const bool in = k > bot;
          const int d = in ? wf_dist(wf[k & mask], k, plen, tlen) : 0;
          const unsigned b = __ballot_sync(kFull, in && (d - dmin <= kMaxDist));
          if (b) { nhi = kc - (__ffs(b) - 1); break; }
        }
      }
      lo = nlo;
      hi = nhi;
    }

    // ---- compute wavefront s+1 in place; trim = first/last in-bounds diagonal ----
    ++s;
    const int clo = lo - 1, chi = hi + 1;
    if (!WIDE && chi - clo + 3 > cap) return kScoreOverflow;
    int tlo = chi + 1, thi = clo - 1;  // trimmed limits (empty until a valid cell is seen)
    int carry = kNull;                 // old M[kc-1]
    for (int kc = clo; kc <= chi; kc += 32) {
      const int k = kc + lane;
      const int mid = (k >= lo && k <= hi) ? wf[k & mask] : kNull;
      const int ke = kc + 32;
      const int edge = (ke >= lo && ke <= hi) ? wf[ke & mask] : kNull;  // broadcast read
      int left = __shfl_up_sync(kFull, mid, 1);
      int right = __shfl_down_sync(kFull, mid, 1);
      if (lane == 0) left = carry;
      if (lane == 31) right = edge;
      carry = __shfl_sync(kFull, mid, 31);
      int m = max(right, max(left, mid) + 1);
      const bool ok = (k <= chi) && in_bounds(m, k, plen, tlen);
      if (!ok) m = kNull;
      __syncwarp();  // all reads of this chunk (incl. the edge slot) done before it is rewritten
      if (k <= chi) wf[k & mask] = m;
      const unsigned b = __ballot_sync(kFull, ok);
      if (b) {
        if (tlo > chi) tlo = kc + __ffs(b) - 1;
        thi = kc + 31 - __clz(b);
      }
    }
    __syncwarp();
    lo = tlo;
    hi = thi;
    if (lo > hi) return -1;  // empty wavefront -> WF_STATUS_UNFEASIBLE

    // ---- exact prune, end-to-end only, width >= 1000 (wavefront_compute_edit.c:214-269) ----
    if (WIDE && !ef && hi - lo + 1 >= kPruneLen) {
      const int ks = lo + (hi - lo) / 2;
      const int so = wf[ks & mask];
      if (so >= 0) {
        const int smax = max(plen - (so - ks), tlen - so);
        // best-case score of diagonal k is |k - ak| whatever its offset
        if (!(abs(lo - ak) <= smax && abs(hi - ak) <= smax)) {
          int w = 0x7fffffff;
          for (int k = lo + lane; k <= hi; k += 32) {
            const int off = wf[k & mask];
            if (off >= 0) w = min(w, max(plen - (off - k), tlen - off));
          }
          w = __reduce_min_sync(kFull, w);
          // first k in [lo,hi] with |k-ak| <= w, else hi+1
          const int nlo = max(lo, ak - w);
          if (nlo > min(hi, ak + w)) return -1;  // nothing survives (cannot happen: w is attained)
          // last k in (nlo,hi] with |k-ak| <= w, else nlo
          int nhi = min(hi, ak + w);
          if (nhi <= nlo) nhi = nlo;
          lo = nlo;
          hi = nhi;
        }
      }
    }
  }
}

}  // namespace otgpu
