// otter_b200/csrc/wfa_affine.cuh — warp-per-pair gap-affine WFA with full CIGAR for sm_100a.
//
// What otter's `aligner2` computes (reference src/assemble.cpp:50:
// WFAlignerGapAffine(4,6,2, Alignment, MemoryMed), default WF-adaptive(10,50,1) cut-off):
//   compute      wavefront/wavefront_compute_affine.c:88-248 (+ limits/null inputs/init-ends of
//                wavefront/wavefront_compute.c:39-174,325-372; per-component trim :376-423)
//   extend/term  wavefront/wavefront_extend.c:45-68,122-168,203-260,355-452
//   cut-off      wavefront/wavefront_heuristic.c:285-328,470-509 (on M; I1/D1 clamped to M)
//   CIGAR        wavefront/wavefront_backtrace_buffer.c:314-362 + wavefront_pcigar.c:203-266
// The reference drags a packed CIGAR along with every cell and spills it into a linked block
// list; what it finally emits depends only on each cell's choice of predecessor.  Here each
// (score, diagonal) keeps one byte of choices in an HBM slab (M source, I ext/open, D ext/open),
// the warp walks them back from the end cell and replays the ops forward with the same
// "re-extend matches only in M state / drop the gap-closing X" rules.  Tie-breaks are the
// reference's: gaps prefer extension on ties (:133,151), M prefers mismatch > deletion > insertion
// (:166-177).  This path is O(reads) per region, so it is written for clarity over issue count:
// wavefronts live in per-warp HBM rings (L1/L2 resident), lanes own consecutive diagonals.
#pragma once
#include "wfa_edit.cuh"

namespace otgpu {

constexpr int kAffMaxScope = 16;     // max(o+e, x) + 1 must fit
constexpr int kAffOverflow = -2;     // trace slab too small (retry with a bigger one)
constexpr int kAffOom = -3;          // even the big slab was too small (WF_STATUS_OOM)

struct AffSlot {  // one score slot (score % scope): ranges + flags of M, I1, D1
  int lo[3], hi[3];
  int flags;  // bit c: component c exists (pointer != NULL); bit 3+c: component c is ->null
  int pad;
};

struct AffWork {        // per-warp workspace in HBM
  int32_t* rings;       // gap-affine: nM rings of M, then nID of I1, then nID of D1 (see below);
                        // edit + CIGAR: two ping-pong rings
  int wmask;            // ring index mask (a ring covers every diagonal of the longest pair)
  int rstride;          // int32 between consecutive rings: wmask + 1 + padding, so that the rings
                        // of one warp do not fall into the same L1 sets
  uint8_t* slab;        // step table grows from the front (8 B / score), cells from the back
  uint32_t slab_bytes;
  uint8_t* ops;         // reversed op list for the CIGAR replay
  uint32_t ops_cap;
};

struct AffPen { int x, o, e, scope, step; };  // step = gcd(x, o+e, e)

__device__ __forceinline__ bool in_bounds(int off, int k, int plen, int tlen) {
  return (unsigned)off <= (unsigned)tlen && (unsigned)(off - k) <= (unsigned)plen;
}

__device__ __forceinline__ int aff_rd(const int32_t* ring, int mask, int lo, int hi, int k) {
  // init-ends rule (wavefront_compute.c:325-372): outside a wavefront's own [lo,hi] reads NULL;
  // a null input has lo=1, hi=-1 (wavefront.c:108-138) and so always reads NULL
  return (k >= lo && k <= hi) ? ring[k & mask] : kNull;
}

// One pair on one warp.  Returns the CIGAR length (>= 0; chars written to cig[0..min(len,cap)))
// or OTGPU_SCORE_UNFEASIBLE (-1) / kAffOverflow (-2).  *score_out = -(penalty), as
// getAlignmentScore() reports for match = 0 (wavefront_penalties.h:66-67).
template <int BITS>
__device__ int wfa_affine_pair(const AffWork& W, AffSlot* __restrict__ meta, const EditJob& J,
                               const AffPen pen, const int lane, char* __restrict__ cig,
                               const int cig_cap, int* score_out) {
  const int plen = J.plen, tlen = J.tlen, ak = tlen - plen;
  const bool ef = J.pb >= 0;
  const int scope = pen.scope, mask = W.wmask;
  const size_t rstride = (size_t)W.rstride;
  // Everything below counts score STEPS of g = gcd(x, o+e, e): wavefront t holds score t*g.  M is
  // read xs = x/g and oes = (o+e)/g steps back, I1/D1 es = e/g steps back, so nM = max(xs,oes)+1
  // M rings and nID = es+1 rings each of I1 and D1 are live at any time (5 + 2 + 2 for 4/6/2
  // instead of 3 x 9 score slots: a third of the cache footprint).
  const int xs = pen.x / pen.step, oes = (pen.o + pen.e) / pen.step, es = pen.e / pen.step;
  const int nM = (xs > oes ? xs : oes) + 1, nID = es + 1;
  int2* const steps = reinterpret_cast<int2*>(W.slab);  // {lo, cell base} per score
  uint32_t cells_top = W.slab_bytes;                    // cells grow downward from here
  int score = 0, sm = 0, si = 0, null_steps = 0, end_k = 0, end_off = 0;  // sm = t % nM, si = t % nID
  // every reachable score is a multiple of g = gcd(x, o+e, e): the scores in between are null
  // steps (wavefront_compute_affine.c:199-206) and are only counted, g at a time
  const int g = pen.step;

#define RING_M(slot) (W.rings + (size_t)(slot) * rstride)
#define RING_I(slot) (W.rings + (size_t)(nM + (slot)) * rstride)
#define RING_D(slot) (W.rings + (size_t)(nM + nID + (slot)) * rstride)
  // ---- wavefront 0 (wavefront_align.c:113-135 / :190-245) ----
  {
    const int lo0 = ef ? -J.pb : 0, hi0 = ef ? J.tb : 0;
    const int w0 = hi0 - lo0 + 1;
    if ((uint32_t)w0 + 64 > W.slab_bytes || w0 + 8 > mask) return kAffOverflow;
    cells_top -= (uint32_t)w0;
    int32_t* m0 = RING_M(0);
    for (int k = lo0 + lane; k <= hi0; k += 32) {
      m0[k & mask] = k > 0 ? k : 0;
      W.slab[cells_top + (k - lo0)] = 3;  // start cell
    }
    if (lane == 0) {
      for (int s = 0; s < nM; ++s) meta[s].flags = 0;
      meta[0].lo[0] = lo0; meta[0].hi[0] = hi0; meta[0].flags = 1;
      steps[0] = make_int2(lo0, (int)cells_top);
    }
    __syncwarp();
  }

  for (;;) {
    const int fl = meta[sm].flags;
    // ---- extend + termination + cut-off (wavefront_extend.c:355-452) ----
    if (!(fl & 1)) {
      if (null_steps > scope) return -1;  // WF_STATUS_UNFEASIBLE (:369-372)
    } else {
      int32_t* m = RING_M(sm);
      WfState S;
      S.lo = meta[sm].lo[0];
      S.hi = meta[sm].hi[0];
      int dmin = kBigDist;
      for (int kc = S.lo; kc <= S.hi; kc += 32) {
        const int k = kc + lane;
        int off = (k <= S.hi) ? m[k & mask] : kNull;
        const bool live = off >= 0;  // M offsets are in-bounds or exactly NULL
        const int v = off - k;
        const int rem = live ? min(plen - v, tlen - off) : 0;
        off += lcp_warp<BITS>(J, v, off, rem, lane);
        bool term = false;
        if (live) {
          m[k & mask] = off;
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
      if (!ef && __any_sync(kFull, dmin == 0)) {  // offset tlen on diagonal tlen-plen (:45-68)
        end_k = ak;
        end_off = tlen;
        goto backtrace;
      }
      if (S.lo <= S.hi) {  // wavefront_heuristic.c:470-509
        if (S.hi - S.lo + 1 >= kMinWfLen) {
          Ring<0> R;
          R.p = m;
          R.mask = mask;
          wf_cutoff<0>(R, plen, tlen, lane, __reduce_min_sync(kFull, dmin), S);
        }
        if (lane == 0) {
          AffSlot& a = meta[sm];
          a.lo[0] = S.lo;
          a.hi[0] = S.hi;
          int f2 = a.flags;
          for (int c = 1; c <= 2; ++c)  // equate I1/D1 to M (:154-165)
            if (f2 & (1 << c)) {
              if (S.lo > a.lo[c]) a.lo[c] = S.lo;
              if (S.hi < a.hi[c]) a.hi[c] = S.hi;
              if (a.lo[c] > a.hi[c]) f2 |= 8 << c;
            }
          a.flags = f2;
        }
        __syncwarp();
      }
    }

    // ---- compute score+1 (wavefront_compute_affine.c:192-248) ----
    if (g > 1) {  // the g-1 null scores in between (checked at their extend, :369-372)
      null_steps += g - 1;
      if (null_steps > scope) return -1;
    }
    score += g;
    if (++sm == nM) sm = 0;
    if (++si == nID) si = 0;
    {
      const int so = sm;
      if ((uint32_t)(score + 1) * 8u + 64u > cells_top) return kAffOverflow;
      const int s_mis = score - pen.x, s_open = score - pen.o - pen.e, s_ext = score - pen.e;
      // ring slots of the three source wavefronts
      const int q_mis = so - xs + (so < xs ? nM : 0);
      const int q_open = so - oes + (so < oes ? nM : 0);
      const int q_ext = so - es + (so < es ? nM : 0);  // M slot (and meta) of score - e
      const int qi_ext = si - es + (si < es ? nID : 0);  // I1/D1 ring of score - e
      // usable input = exists and not ->null (wavefront_compute.c:88-95); else lo=1, hi=-1
      int mm_lo = 1, mm_hi = -1, mo_lo = 1, mo_hi = -1, ie_lo = 1, ie_hi = -1, de_lo = 1, de_hi = -1;
      bool has_mm = false, has_mo = false, has_ie = false, has_de = false;
      if (s_mis >= 0) {
        const AffSlot& a = meta[q_mis];
        if ((a.flags & 1) && !(a.flags & 8)) { has_mm = true; mm_lo = a.lo[0]; mm_hi = a.hi[0]; }
      }
      if (s_open >= 0) {
        const AffSlot& a = meta[q_open];
        if ((a.flags & 1) && !(a.flags & 8)) { has_mo = true; mo_lo = a.lo[0]; mo_hi = a.hi[0]; }
      }
      if (s_ext >= 0) {
        const AffSlot& a = meta[q_ext];
        if ((a.flags & 2) && !(a.flags & 16)) { has_ie = true; ie_lo = a.lo[1]; ie_hi = a.hi[1]; }
        if ((a.flags & 4) && !(a.flags & 32)) { has_de = true; de_lo = a.lo[2]; de_hi = a.hi[2]; }
      }
      __syncwarp();
      if (!has_mm && !has_mo && !has_ie && !has_de) {  // null step (:199-206)
        ++null_steps;
        if (lane == 0) {
          meta[so].flags = 0;
          steps[score] = make_int2(1, 0);
        }
        __syncwarp();
        continue;
      }
      null_steps = 0;
      // limits (wavefront_compute.c:39-69)
      const int lo = min(min(mm_lo, mo_lo - 1), min(ie_lo + 1, de_lo - 1));
      const int hi = max(max(mm_hi, mo_hi + 1), max(ie_hi + 1, de_hi - 1));
      const int width = hi - lo + 1;
      if (width + 8 > mask) return kAffOverflow;
      if ((uint32_t)(score + 1) * 8u + 64u + (uint32_t)width > cells_top) return kAffOverflow;
      cells_top -= (uint32_t)width;
      const bool has_i = has_mo || has_ie, has_d = has_mo || has_de;  // wavefront_compute.c:262-283
      const int32_t* r_mm = RING_M(q_mis);
      const int32_t* r_mo = RING_M(q_open);
      const int32_t* r_ie = RING_I(qi_ext);
      const int32_t* r_de = RING_D(qi_ext);
      int32_t* o_m = RING_M(so);
      int32_t* o_i = RING_I(si);
      int32_t* o_d = RING_D(si);
      int mn[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
      int mx[3] = {(int)0x80000000, (int)0x80000000, (int)0x80000000};
      // Two adjacent diagonals per lane, 64 per round (k0 even: one 64-bit access per ring and
      // pair; twice the independent loads in flight per warp - this kernel waits on its ring
      // round trips, not on issue slots: profiles/README.md r2m; 41.6 -> 38.1 ms per 11,800 C2
      // alignments.  The same for the extend loop measured slower, 39.6 ms, and is not used).
      for (int kc = lo & ~1; kc <= hi; kc += 64) {
        const int k0 = kc + 2 * lane, k1 = k0 + 1;
        if (k0 <= hi) {
          const bool on0 = k0 >= lo, on1 = k1 <= hi;  // (kc may start one below lo; hi may be k0)
          const int i0 = k0 & mask;                    // even: slots i0, i0+1 share an 8-byte word
          // M[s-o-e] at k0-1 .. k0+2, I1[s-e] at k0-1, k0, D1[s-e] at k0+1, k0+2, M[s-x] at k0, k1
          const int2 mo_p = *reinterpret_cast<const int2*>(r_mo + i0);
          const int2 mm_p = *reinterpret_cast<const int2*>(r_mm + i0);
          const int mo_m1 = aff_rd(r_mo, mask, mo_lo, mo_hi, k0 - 1);
          const int mo_p2 = aff_rd(r_mo, mask, mo_lo, mo_hi, k0 + 2);
          const int ie_m1 = aff_rd(r_ie, mask, ie_lo, ie_hi, k0 - 1);
          const int ie_0 = aff_rd(r_ie, mask, ie_lo, ie_hi, k0);
          const int de_p1 = aff_rd(r_de, mask, de_lo, de_hi, k0 + 1);
          const int de_p2 = aff_rd(r_de, mask, de_lo, de_hi, k0 + 2);
          const int mo_0 = (k0 >= mo_lo && k0 <= mo_hi) ? mo_p.x : kNull;
          const int mo_1 = (k1 >= mo_lo && k1 <= mo_hi) ? mo_p.y : kNull;
          const int mm_0 = (k0 >= mm_lo && k0 <= mm_hi) ? mm_p.x : kNull;
          const int mm_1 = (k1 >= mm_lo && k1 <= mm_hi) ? mm_p.y : kNull;
          // diagonal k0: ins from k0-1, del from k0+1; diagonal k1: ins from k0, del from k0+2
          const bool ie0 = ie_m1 >= mo_m1, de0 = de_p1 >= mo_1;  // ties prefer extend
          const bool ie1 = ie_0 >= mo_0, de1 = de_p2 >= mo_p2;
          const int ins0 = (ie0 ? ie_m1 : mo_m1) + 1, del0 = de0 ? de_p1 : mo_1, mis0 = mm_0 + 1;
          const int ins1 = (ie1 ? ie_0 : mo_0) + 1, del1 = de1 ? de_p2 : mo_p2, mis1 = mm_1 + 1;
          int best0 = max(del0, max(mis0, ins0)), best1 = max(del1, max(mis1, ins1));
          // mismatch > deletion > insertion on ties (:166-177)
          const int src0 = best0 == mis0 ? 0 : (best0 == del0 ? 2 : 1);
          const int src1 = best1 == mis1 ? 0 : (best1 == del1 ? 2 : 1);
          if (!in_bounds(best0, k0, plen, tlen)) best0 = kNull;
          if (!in_bounds(best1, k1, plen, tlen)) best1 = kNull;
          const uint8_t c0 = (uint8_t)(src0 | (ie0 ? 4 : 0) | (de0 ? 8 : 0));
          const uint8_t c1 = (uint8_t)(src1 | (ie1 ? 4 : 0) | (de1 ? 8 : 0));
          if (on0 && on1) {
            if (has_i) *reinterpret_cast<int2*>(o_i + i0) = make_int2(ins0, ins1);
            if (has_d) *reinterpret_cast<int2*>(o_d + i0) = make_int2(del0, del1);
            *reinterpret_cast<int2*>(o_m + i0) = make_int2(best0, best1);
            W.slab[cells_top + (k0 - lo)] = c0;
            W.slab[cells_top + (k1 - lo)] = c1;
          } else if (on0) {
            if (has_i) o_i[i0] = ins0;
            if (has_d) o_d[i0] = del0;
            o_m[i0] = best0;
            W.slab[cells_top + (k0 - lo)] = c0;
          } else if (on1) {
            if (has_i) o_i[i0 + 1] = ins1;
            if (has_d) o_d[i0 + 1] = del1;
            o_m[i0 + 1] = best1;
            W.slab[cells_top + (k1 - lo)] = c1;
          }
          if (on0) {
            if (best0 >= 0) { mn[0] = min(mn[0], k0); mx[0] = max(mx[0], k0); }
            if (in_bounds(ins0, k0, plen, tlen)) { mn[1] = min(mn[1], k0); mx[1] = max(mx[1], k0); }
            if (in_bounds(del0, k0, plen, tlen)) { mn[2] = min(mn[2], k0); mx[2] = max(mx[2], k0); }
          }
          if (on1) {
            if (best1 >= 0) { mn[0] = min(mn[0], k1); mx[0] = max(mx[0], k1); }
            if (in_bounds(ins1, k1, plen, tlen)) { mn[1] = min(mn[1], k1); mx[1] = max(mx[1], k1); }
            if (in_bounds(del1, k1, plen, tlen)) { mn[2] = min(mn[2], k1); mx[2] = max(mx[2], k1); }
          }
        }
      }
      // trim each component to its first/last in-bounds diagonal (wavefront_compute.c:376-423)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        mn[c] = __reduce_min_sync(kFull, mn[c]);
        mx[c] = __reduce_max_sync(kFull, mx[c]);
      }
      if (lane == 0) {
        AffSlot& a = meta[so];
        int f2 = 1 | (has_i ? 2 : 0) | (has_d ? 4 : 0);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          a.lo[c] = mn[c];
          a.hi[c] = mx[c];
          if (mn[c] > mx[c]) f2 |= 8 << c;
        }
        a.flags = f2;
        steps[score] = make_int2(lo, (int)cells_top);
      }
      __syncwarp();
    }
  }

backtrace:
  __syncwarp();
  {
    // ---- walk the choices back from the end cell (all lanes in lock-step, lane 0 stores) ----
    const int result = score;
    int s = score, kk = end_k, st = 0, n_ops = 0;  // st: 0 M, 1 I1, 2 D1
    for (;;) {
      if (st == 0 && s == 0) break;
      const int2 se = steps[s];
      const int c = W.slab[(uint32_t)se.y + (uint32_t)(kk - se.x)];
      if ((uint32_t)n_ops + 2 > W.ops_cap) return kAffOverflow;
      char op;
      if (st == 0) {
        if ((c & 3) == 0) { op = 'X'; s -= pen.x; }
        else { op = 'c'; st = (c & 3) == 1 ? 1 : 2; }  // gap-closing X: marker only
      } else if (st == 1) {
        op = 'I';
        if (c & 4) { s -= pen.e; kk -= 1; } else { s -= pen.o + pen.e; kk -= 1; st = 0; }
      } else {
        op = 'D';
        if (c & 8) { s -= pen.e; kk += 1; } else { s -= pen.o + pen.e; kk += 1; st = 0; }
      }
      if (lane == 0) W.ops[n_ops] = (uint8_t)op;
      ++n_ops;
    }
    __syncwarp();
    // ---- forward replay (wavefront_backtrace_buffer.c:314-362, wavefront_pcigar.c:203-266) ----
    int h = kk > 0 ? kk : 0, v = kk > 0 ? 0 : -kk;  // start cell of wavefront 0
    int len = 0;
    bool mstate = true;
    auto fill = [&](char ch, int n) {
      for (int i = lane; i < n; i += 32)
        if (len + i < cig_cap) cig[len + i] = ch;
      len += n;
    };
    fill('I', h);  // :334-335
    fill('D', v);
    for (int i = n_ops - 1; i >= 0; --i) {
      if (mstate) {  // matches are re-extended only in the M state (:225-238)
        const int n = lcp_uniform<BITS>(J, v, h, min(plen - v, tlen - h), lane);
        fill('M', n);
        v += n;
        h += n;
      }
      const char op = (char)W.ops[i];
      if (op == 'c') { mstate = true; continue; }  // back to M, nothing emitted (:250-254)
      fill(op, 1);
      if (op == 'X') { ++v; ++h; mstate = true; }
      else if (op == 'I') { ++h; mstate = false; }
      else { ++v; mstate = false; }
    }
    {
      const int nm = min((end_off - end_k) - v, end_off - h);  // last stroke (:352-358)
      if (nm > 0) { fill('M', nm); v += nm; h += nm; }
    }
    if (h < tlen) fill('I', tlen - h);
    if (v < plen) fill('D', plen - v);
    if (score_out) *score_out = -result;
    return len;
  }
#undef RING_M
#undef RING_I
#undef RING_D
}

}  // namespace otgpu
