// otter_b200/csrc/wfa_edit_cigar.cuh — warp-per-pair edit-distance WFA WITH CIGAR for sm_100a.
//
// What `otter compare` computes per allele pair (reference src/compare.cpp:50-66,95:
// wfa::WFAlignerEdit(Alignment, MemoryMed), alignEnd2End + getAlignmentScore +
// getAlignmentCigar), default WF-adaptive(10,50,1) cut-off:
//   compute      wavefront/wavefront_compute_edit.c:152-195 (piggyback kernel: the score-only
//                offsets plus the predecessor; ties mismatch > deletion > insertion)
//   extend/term  wavefront/wavefront_extend.c:45-68,122-168,203-260
//   cut-off      wavefront/wavefront_heuristic.c:285-328; trim wavefront_compute.c:376-409;
//   exact prune  wavefront/wavefront_compute_edit.c:214-269 (end-to-end, width >= 1000)
//   CIGAR        wavefront/wavefront_backtrace_buffer.c:266-312 + wavefront_pcigar.c:151-199
// Same storage idea as the gap-affine path (wfa_affine.cuh): one byte of predecessor choice per
// (score, diagonal) in the warp's HBM slab, walked back from the end cell and replayed forward,
// re-extending matches before every operation.  Two ping-pong wavefront rings in HBM; this path
// is O(alleles^2) per region on short allele lists, written for clarity over issue count.
#pragma once
#include "wfa_affine.cuh"

namespace otgpu {

__device__ __forceinline__ int ec_rd(const int32_t* ring, int mask, int lo, int hi, int k) {
  // everything outside the previous wavefront's [lo,hi] reads as null (edit.c:311-318)
  return (k >= lo && k <= hi) ? ring[k & mask] : kNull;
}

// One pair on one warp.  Returns the CIGAR length (>= 0), OTGPU_SCORE_UNFEASIBLE (-1) or
// kAffOverflow (-2: trace slab too small).  *score_out = the edit score (wavefront_align.c:289).
template <int BITS>
__device__ int wfa_editcigar_pair(const AffWork& W, const EditJob& J, const int lane,
                                  char* __restrict__ cig, const int cig_cap, int* score_out) {
  const int plen = J.plen, tlen = J.tlen, ak = tlen - plen;
  const bool ef = J.pb >= 0;
  const int mask = W.wmask;
  int32_t* cur = W.rings;
  int32_t* nxt = W.rings + (size_t)W.rstride;
  int2* const steps = reinterpret_cast<int2*>(W.slab);  // {lo, cell base} per score
  uint32_t cells_top = W.slab_bytes;                    // cells grow downward from here
  int score = 0, end_k = 0, end_off = 0;
  WfState S;

  // ---- wavefront 0 (wavefront_align.c:113-135 / :190-237) ----
  {
    S.lo = ef ? -J.pb : 0;
    S.hi = ef ? J.tb : 0;
    const int w0 = S.hi - S.lo + 1;
    if ((uint32_t)w0 + 64 > W.slab_bytes || w0 + 8 > mask) return kAffOverflow;
    cells_top -= (uint32_t)w0;
    for (int k = S.lo + lane; k <= S.hi; k += 32) {
      cur[k & mask] = k > 0 ? k : 0;
      W.slab[cells_top + (k - S.lo)] = 3;  // start cell
    }
    if (lane == 0) steps[0] = make_int2(S.lo, (int)cells_top);
    __syncwarp();
  }

  for (;;) {
    // ---- extend + termination ----
    int dmin = kBigDist;
    for (int kc = S.lo; kc <= S.hi; kc += 32) {
      const int k = kc + lane;
      int off = (k <= S.hi) ? cur[k & mask] : kNull;
      const bool live = off >= 0;  // offsets are in bounds or exactly null
      const int v = off - k;
      const int rem = live ? min(plen - v, tlen - off) : 0;
      off += lcp_warp<BITS>(J, v, off, rem, lane);
      bool term = false;
      if (live) {
        cur[k & mask] = off;
        dmin = min(dmin, plen + max(k, ak) - off);
        if (ef) {
          const int left_v = plen - (off - k), left_h = tlen - off;
          term = (left_h <= 0 && left_v <= J.pe) || (left_v <= 0 && left_h <= J.te);
        }
      }
      if (ef) {  // first k in ascending order wins (wavefront_extend.c:238-260)
        const unsigned b = __ballot_sync(kFull, term);
        if (b) {
          const int f = __ffs(b) - 1;
          end_k = kc + f;
          end_off = __shfl_sync(kFull, off, f);
          goto backtrace;
        }
      }
    }
    __syncwarp();
    dmin = __reduce_min_sync(kFull, dmin);
    if (!ef && dmin == 0) {  // offset tlen on diagonal tlen-plen (wavefront_extend.c:45-68)
      end_k = ak;
      end_off = tlen;
      goto backtrace;
    }
    Ring<0> R;
    if (S.hi - S.lo + 1 >= kMinWfLen) {
      R.p = cur;
      R.mask = mask;
      wf_cutoff<0>(R, plen, tlen, lane, dmin, S);
    }

    // ---- compute score+1 with predecessor choice (wavefront_compute_edit.c:152-195) ----
    ++score;
    const int lo = S.lo - 1, hi = S.hi + 1, width = hi - lo + 1;
    if (width + 8 > mask) return kAffOverflow;
    if ((uint32_t)(score + 1) * 8u + 64u + (uint32_t)width > cells_top) return kAffOverflow;
    cells_top -= (uint32_t)width;
    int mn = 0x7fffffff, mx = (int)0x80000000;
    for (int kc = lo; kc <= hi; kc += 32) {
      const int k = kc + lane;
      if (k <= hi) {
        const int ins = ec_rd(cur, mask, S.lo, S.hi, k - 1) + 1;
        const int mis = ec_rd(cur, mask, S.lo, S.hi, k) + 1;
        const int del = ec_rd(cur, mask, S.lo, S.hi, k + 1);
        int best = max(del, max(ins, mis));
        int src = 0;
        if (best == ins) src = 1;
        if (best == del) src = 2;
        if (best == mis) src = 0;  // last assignment wins (:176-187)
        W.slab[cells_top + (k - lo)] = (uint8_t)src;
        if (!in_bounds(best, k, plen, tlen)) best = kNull;
        nxt[k & mask] = best;
        if (best >= 0) { mn = min(mn, k); mx = max(mx, k); }
      }
    }
    mn = __reduce_min_sync(kFull, mn);  // trim (wavefront_compute.c:376-409)
    mx = __reduce_max_sync(kFull, mx);
    if (lane == 0) steps[score] = make_int2(lo, (int)cells_top);
    __syncwarp();
    if (mn > mx) return -1;  // WF_STATUS_UNFEASIBLE (edit.c:349)
    S.lo = mn;
    S.hi = mx;
    { int32_t* t = cur; cur = nxt; nxt = t; }
    if (!ef && S.hi - S.lo + 1 >= kPruneLen) {  // exact prune (edit.c:214-269)
      R.p = cur;
      R.mask = mask;
      if (!wf_prune<0>(R, plen, tlen, lane, S)) return -1;
    }
  }

backtrace:
  __syncwarp();
  {
    // ---- walk the choices back from the end cell (all lanes in lock-step, lane 0 stores) ----
    int kk = end_k;
    if ((uint32_t)score + 2 > W.ops_cap) return kAffOverflow;
    for (int s = score; s > 0; --s) {
      const int2 se = steps[s];
      const int c = W.slab[(uint32_t)se.y + (uint32_t)(kk - se.x)];
      const char op = c == 0 ? 'X' : (c == 1 ? 'I' : 'D');
      kk += c == 1 ? -1 : (c == 2 ? 1 : 0);
      if (lane == 0) W.ops[score - s] = (uint8_t)op;
    }
    __syncwarp();
    // ---- forward replay (wavefront_backtrace_buffer.c:266-312, wavefront_pcigar.c:151-199) ----
    int h = kk > 0 ? kk : 0, v = kk > 0 ? 0 : -kk;  // start cell of wavefront 0
    int len = 0;
    auto fill = [&](char ch, int n) {
      for (int i = lane; i < n; i += 32)
        if (len + i < cig_cap) cig[len + i] = ch;
      len += n;
    };
    fill('I', h);  // :283-284
    fill('D', v);
    for (int i = score - 1; i >= 0; --i) {
      const int n = lcp_uniform<BITS>(J, v, h, min(plen - v, tlen - h), lane);  // :176-185
      fill('M', n);
      v += n;
      h += n;
      const char op = (char)W.ops[i];
      fill(op, 1);
      if (op == 'X') { ++v; ++h; }
      else if (op == 'I') ++h;
      else ++v;
    }
    {
      const int nm = min((end_off - end_k) - v, end_off - h);  // last stroke (:300-303)
      if (nm > 0) { fill('M', nm); v += nm; h += nm; }
    }
    if (h < tlen) fill('I', tlen - h);
    if (v < plen) fill('D', plen - v);
    if (score_out) *score_out = score;
    return len;
  }
}

}  // namespace otgpu
