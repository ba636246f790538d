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

constexpr int kNull = -(1 << 30);  // WAVEFRONT_OFFSET_NULL, wavefront_offset.h:46 (gap-affine path)
// Null offset of the EDIT path.  Only its sign and being out of bounds matter there (a stored
// offset is either in bounds or exactly null; null + 1 is re-nulled by the same pass), so a
// smaller magnitude is equivalent — and keeps the positions that null-derived offsets address
// within a few MiB of the read, where the batch buffers are padded, so the extension's window
// loads need no guards.
constexpr int kENull = -(1 << 22);
constexpr uint32_t kMaxReadLen = 1u << 20;                       // enforced at upload
constexpr uint32_t kPadBases = (1u << 22) + 2 * kMaxReadLen + 4096;  // guard positions either side
constexpr unsigned kFull = 0xffffffffu;
constexpr int kMinWfLen = 10;   // wavefront_attributes.c:74
constexpr int kMaxDist = 50;    // wavefront_attributes.c:75
constexpr int kPruneLen = 1000; // wavefront_compute_edit.c:224
constexpr int kScoreOverflow = -2;  // internal: live width exceeded the shared-memory ring
constexpr int kFwRun = 32;  // a pair advancing >= this many bases per score step is a low-error pair (k_align's count)
constexpr int kBigDist = 1 << 30;   // distance of a null offset, wavefront_heuristic.c:118

// ---- wavefront ring ---------------------------------------------------------------------------
__device__ __forceinline__ int lds32(unsigned a) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(unsigned a, int v) {
  asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

// Ring<CAP>: CAP int32 slots in shared memory (CAP a power of two), or, CAP == 0, an HBM
// workspace whose capacity covers every diagonal of the pair ("wide").
template <int CAP>
struct Ring {         // shared memory, fixed capacity, 32-bit shared-window addressing
  unsigned base;      // aligned to the ring size, so slot address = base | ((k*4) & mask): one LOP3
  static constexpr unsigned kMaskB = (CAP - 1) * 4;
  __device__ __forceinline__ int cap() const { return CAP; }
  __device__ __forceinline__ unsigned addr(int k) const { return base | (((unsigned)k << 2) & kMaskB); }
  __device__ __forceinline__ int ld(int k) const { return lds32(addr(k)); }
  __device__ __forceinline__ void st(int k, int v) const { sts32(addr(k), v); }
  // pair of adjacent slots (k even), addressed through a token so that a pass computes each
  // address once (loaded as "next", stored one round later as "current")
  typedef unsigned Tok;
  __device__ __forceinline__ Tok tok(int k) const { return addr(k); }
  __device__ __forceinline__ int2 ld2(Tok t) const {
    int2 v;
    asm volatile("ld.shared.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(t) : "memory");
    return v;
  }
  __device__ __forceinline__ void st2(Tok t, int a, int b) const {
    asm volatile("st.shared.v2.s32 [%0], {%1, %2};" ::"r"(t), "r"(a), "r"(b) : "memory");
  }
  __device__ __forceinline__ void fill_null(int lane) const {
#pragma unroll
    for (int i = 0; i < CAP / 128; ++i)
      asm volatile("st.shared.v4.s32 [%0], {%1, %1, %1, %1};" ::"r"(base + (i * 32 + lane) * 16), "r"(kENull)
                   : "memory");
  }
};
template <>
struct Ring<0> {  // HBM workspace (L1/L2 cached), capacity covers every diagonal of the pair
  int32_t* p;
  int mask;
  __device__ __forceinline__ int cap() const { return mask + 1; }
  __device__ __forceinline__ int ld(int k) const { return p[k & mask]; }
  __device__ __forceinline__ void st(int k, int v) const { p[k & mask] = v; }
  typedef int2* Tok;
  __device__ __forceinline__ Tok tok(int k) const { return reinterpret_cast<int2*>(p + (k & mask)); }
  __device__ __forceinline__ int2 ld2(Tok t) const { return *t; }
  __device__ __forceinline__ void st2(Tok t, int a, int b) const { *t = make_int2(a, b); }
  __device__ __forceinline__ void fill_null(int lane) const {
    for (int i = lane; i <= mask; i += 32) p[i] = kENull;
  }
};

// ---- checking build (-DOTGPU_CHECK_BOUNDS; tests only) ------------------------------------------
// The window loads below are issued without guards (buffers carry guard regions instead, see
// DESIGN.md §4).  The checking build verifies that claim from the inside: every window address is
// compared with the two buffers it may fall into (the 2-bit copy, the raw bytes viewed as words),
// a load that would leave them is counted and redirected to the start of the buffer.
#ifdef OTGPU_CHECK_BOUNDS
struct CheckRange { const uint32_t* lo; const uint32_t* hi; };
__device__ CheckRange otgpu_check_range[2];
__device__ unsigned long long otgpu_check_oob;
__device__ unsigned long long otgpu_check_loads;
#endif

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
                                         int p, uint32_t& lo, uint32_t& hi) {
  const int a = (BITS == 2) ? p : s.sh0 + p;
#ifdef OTGPU_MADHI
  // word index = word0 + (a >> kLogBpw) as one multiply-high-add: FMA pipe instead of the ALU pipe
  uint32_t wi;
  asm("mad.hi.s32 %0, %1, %2, %3;" : "=r"(wi) : "r"(a), "r"(1 << (32 - Pack<BITS>::kLogBpw)), "r"(s.word0));
  const uint32_t* q = words + wi;
#else
  const uint32_t* q = words + (s.word0 + (uint32_t)(a >> Pack<BITS>::kLogBpw));
#endif
#ifdef OTGPU_CHECK_BOUNDS
  {
    const CheckRange r0 = otgpu_check_range[0], r1 = otgpu_check_range[1];
    const bool ok = (q >= r0.lo && q + 3 <= r0.hi) || (q >= r1.lo && q + 3 <= r1.hi);
    if (!ok) {
      atomicAdd(&otgpu_check_oob, 1ull);
      q = r0.lo;
    }
    if ((threadIdx.x & 31) == 0) atomicAdd(&otgpu_check_loads, 1ull);
  }
#endif
  const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2);
  const int sh = a << Pack<BITS>::kLogBits;
  lo = __funnelshift_r(w0, w1, sh);
  hi = __funnelshift_r(w1, w2, sh);
}

// Matching bases at the front of P[v..] and T[h..] within one 64-bit window (kWin if identical).
// v and h must be positions inside the reads (the loads are unconditional).
template <int BITS>
__device__ __forceinline__ int match_window(const uint32_t* __restrict__ words, const SeqView& P,
                                            const SeqView& T, int v, int h) {
  uint32_t pl, ph, tl, th;
  window64<BITS>(words, P, v, pl, ph);
  window64<BITS>(words, T, h, tl, th);
  const uint32_t xl = pl ^ tl, xh = ph ^ th;
  const uint32_t sel = xl ? xl : xh;
  // first differing bit; identical windows give ffs = 0 -> 0xffffffff, clamped to kWin
  const uint32_t bit = (uint32_t)(__ffs(sel) - 1) | (xl ? 0u : 32u);
  return (int)min(bit >> Pack<BITS>::kLogBits, (uint32_t)Pack<BITS>::kWin);
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
// In two halves, so that a caller with two diagonals per lane can issue both first windows before
// the first vote: lcp_first = the unconditional first window (idle lanes, rem <= 0, probe position
// 0 of both reads and discard the result), lcp_rest = private windows while several lanes are
// still running, then each remaining run finished by the whole warp.  cont = this lane's run
// matched its whole window and has more to compare.
template <int BITS>
__device__ __forceinline__ int lcp_first(const EditJob& J, int v, int h, int rem, bool& cont) {
  constexpr int WIN = Pack<BITS>::kWin;
  const bool on = rem > 0;
  const int m = match_window<BITS>(J.words, J.P, J.T, on ? v : 0, on ? h : 0);
  cont = (m >= WIN) && (rem > WIN);
  return max(min(m, rem), 0);  // clamp at a sequence end (the reference stops on its sentinels)
}
template <int BITS>
__device__ __forceinline__ int lcp_rest(const EditJob& J, int v, int h, int rem, int n, bool cont,
                                        const int lane) {
  constexpr int WIN = Pack<BITS>::kWin;
  unsigned pend = __ballot_sync(kFull, cont);
  if (pend == 0) return n;
  for (int it = 0;; ++it) {
    if (cont) {
      v += WIN; h += WIN; rem -= WIN;
      const int m = match_window<BITS>(J.words, J.P, J.T, v, h);
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
template <int BITS>
__device__ __forceinline__ int lcp_warp(const EditJob& J, int v, int h, int rem, const int lane) {
  bool cont;
  const int n = lcp_first<BITS>(J, v, h, rem, cont);
  return lcp_rest<BITS>(J, v, h, rem, n, cont, lane);
}

// A group of G lanes (the whole warp, or one of its halves) that works on one pair.  Every
// warp-level primitive below takes the group: sync masks name only its lanes, lane indices are
// relative to its first lane, so two half-warps can run two pairs through the same instruction
// stream (k_align_dual) while the full-warp kernels compile to what they were.
template <int G>
struct Grp {
  static constexpr int kLanes = G;
  unsigned mask;  // lanes of the group
  int gl;         // this lane's index inside the group
  int first;      // warp lane of the group's lane 0
  __device__ __forceinline__ explicit Grp(int lane)
      : mask(G == 32 ? kFull : (0xffffu << (lane & 16))), gl(lane & (G - 1)), first(lane & ~(G - 1)) {}
  // group-relative index of the first set bit of a ballot taken with `mask`
  __device__ __forceinline__ int pos(unsigned ballot) const { return __ffs(ballot) - 1 - first; }
};

// LCP of P[v..) and T[h..) for ONE group-uniform position, all lanes of the group cooperating
// (G windows per round).  rem = min(plen - v, tlen - h).
template <int BITS, int G>
__device__ __forceinline__ int lcp_uniform(const EditJob& J, int v, int h, int rem, const Grp<G>& g) {
  constexpr int WIN = Pack<BITS>::kWin;
  int got = 0;
  while (rem > 0) {
    const int o = g.gl * WIN;
    int mm = 0;
    bool stop = true;
    if (o < rem) {
      mm = match_window<BITS>(J.words, J.P, J.T, v + o, h + o);
      const int lim = rem - o;
      if (mm >= lim) mm = lim;
      else stop = mm < WIN;
    }
    const unsigned b = __ballot_sync(g.mask, stop);
    if (b) {
      const int f = __ffs(b) - 1;
      return got + (f - g.first) * WIN + __shfl_sync(g.mask, mm, f);
    }
    got += G * WIN; v += G * WIN; h += G * WIN; rem -= G * WIN;
  }
  return got;
}
template <int BITS>
__device__ __forceinline__ int lcp_uniform(const EditJob& J, int v, int h, int rem, const int lane) {
  return lcp_uniform<BITS, 32>(J, v, h, rem, Grp<32>(lane));
}

// Bases matched at the front of one 64-bit window whose LAST base is replaced by a sentinel
// difference: the result is 0..kWin-1, and kWin-1 means "all kWin-1 compared bases match" (the
// caller continues from there).  That makes the clamp a single min with the remaining length.
// Loads are unconditional: v and h may be any position a wavefront offset can take, null-derived
// garbage included (the batch buffers are padded for that).
template <int BITS>
__device__ __forceinline__ uint32_t first_diff(const EditJob& J, int v, int h) {
  uint32_t pl, ph, tl, th;
  window64<BITS>(J.words, J.P, v, pl, ph);
  window64<BITS>(J.words, J.T, h, tl, th);
  const uint32_t xl = pl ^ tl, xh = (ph ^ th) | 0x80000000u;
  const uint32_t sel = xl ? xl : xh;
  uint32_t bit = (uint32_t)(__ffs(sel) - 1);
  if (!xl) bit |= 32u;
  return bit >> Pack<BITS>::kLogBits;
}

// The same over one 32-bit window (two words per read: kWin/2 - 1 compared bases).
template <int BITS>
__device__ __forceinline__ uint32_t first_diff_half(const EditJob& J, int v, int h) {
  const int a = (BITS == 2) ? v : J.P.sh0 + v, b = (BITS == 2) ? h : J.T.sh0 + h;
  const uint32_t* p = J.words + (J.P.word0 + (uint32_t)(a >> Pack<BITS>::kLogBpw));
  const uint32_t* t = J.words + (J.T.word0 + (uint32_t)(b >> Pack<BITS>::kLogBpw));
#ifdef OTGPU_CHECK_BOUNDS
  {
    const CheckRange r0 = otgpu_check_range[0], r1 = otgpu_check_range[1];
    const bool okp = (p >= r0.lo && p + 2 <= r0.hi) || (p >= r1.lo && p + 2 <= r1.hi);
    const bool okt = (t >= r0.lo && t + 2 <= r0.hi) || (t >= r1.lo && t + 2 <= r1.hi);
    if (!okp || !okt) {
      atomicAdd(&otgpu_check_oob, 1ull);
      p = t = r0.lo;
    }
    if ((threadIdx.x & 31) == 0) atomicAdd(&otgpu_check_loads, 2ull);
  }
#endif
  const uint32_t pw = __funnelshift_r(__ldg(p), __ldg(p + 1), a << Pack<BITS>::kLogBits);
  const uint32_t tw = __funnelshift_r(__ldg(t), __ldg(t + 1), b << Pack<BITS>::kLogBits);
  return (uint32_t)(__ffs((pw ^ tw) | 0x80000000u) - 1) >> Pack<BITS>::kLogBits;
}
// First window of an extension.  FW = 0: 64 bits (31 compared bases at 2 bit), what noisy reads
// want - most of their runs end inside it.  FW = 1: 32 bits (15 bases: two loads and one funnel
// shift less per read) followed by 128-bit continuation windows, for pairs whose runs are either
// a few bases (off the alignment path) or hundreds (on it): low-error reads.  Same counts either
// way.  The choice is made per launch (k_align<1024, FW>): a pair picking its own window inside
// one kernel doubles the hot loop and ran 60 % slower (instruction cache), see profiles/README r3f.
template <int BITS, int FW>
__device__ __forceinline__ uint32_t first_window(const EditJob& J, int v, int h) {
  return FW ? first_diff_half<BITS>(J, v, h) : first_diff<BITS>(J, v, h);
}
template <int BITS, int FW>
struct FirstWin { static constexpr int kWE = FW ? Pack<BITS>::kWin / 2 - 1 : Pack<BITS>::kWin - 1; };

// The same over a 128-bit window (four words per read: 63 compared bases at 2 bit, 15 at 8 bit, the
// last base again a sentinel difference): continuation of a run that outlived its first window.
// Touches up to four words past the word holding the position (buffers carry guards for that).
template <int BITS>
__device__ __forceinline__ uint32_t first_diff_wide(const EditJob& J, int v, int h) {
  uint32_t x[4];
  {
    const int a = (BITS == 2) ? v : J.P.sh0 + v, b = (BITS == 2) ? h : J.T.sh0 + h;
    const uint32_t* p = J.words + (J.P.word0 + (uint32_t)(a >> Pack<BITS>::kLogBpw));
    const uint32_t* t = J.words + (J.T.word0 + (uint32_t)(b >> Pack<BITS>::kLogBpw));
#ifdef OTGPU_CHECK_BOUNDS
    {
      const CheckRange r0 = otgpu_check_range[0], r1 = otgpu_check_range[1];
      const bool okp = (p >= r0.lo && p + 5 <= r0.hi) || (p >= r1.lo && p + 5 <= r1.hi);
      const bool okt = (t >= r0.lo && t + 5 <= r0.hi) || (t >= r1.lo && t + 5 <= r1.hi);
      if (!okp || !okt) {
        atomicAdd(&otgpu_check_oob, 1ull);
        p = t = r0.lo;
      }
      if ((threadIdx.x & 31) == 0) atomicAdd(&otgpu_check_loads, 2ull);
    }
#endif
    const uint32_t p0 = __ldg(p), p1 = __ldg(p + 1), p2 = __ldg(p + 2), p3 = __ldg(p + 3), p4 = __ldg(p + 4);
    const uint32_t t0 = __ldg(t), t1 = __ldg(t + 1), t2 = __ldg(t + 2), t3 = __ldg(t + 3), t4 = __ldg(t + 4);
    const int sp = a << Pack<BITS>::kLogBits, st = b << Pack<BITS>::kLogBits;
    x[0] = __funnelshift_r(p0, p1, sp) ^ __funnelshift_r(t0, t1, st);
    x[1] = __funnelshift_r(p1, p2, sp) ^ __funnelshift_r(t1, t2, st);
    x[2] = __funnelshift_r(p2, p3, sp) ^ __funnelshift_r(t2, t3, st);
    x[3] = (__funnelshift_r(p3, p4, sp) ^ __funnelshift_r(t3, t4, st)) | 0x80000000u;
  }
  const uint32_t lo2 = x[0] | x[1];
  const uint32_t sel = lo2 ? (x[0] ? x[0] : x[1]) : (x[2] ? x[2] : x[3]);
  const uint32_t word = lo2 ? (x[0] ? 0u : 32u) : (x[2] ? 64u : 96u);
  return ((uint32_t)(__ffs(sel) - 1) + word) >> Pack<BITS>::kLogBits;
}

// Slow path of a pass: some lane's first window matched completely.  n = bases matched so far
// from (v,h), rem = the limit, c = this lane's run is still pending.  A few rounds of private
// windows, then every run still pending is finished by the whole group.  Returns the new n.
template <int BITS, int G, bool WIDE_ALL = false>
__device__ __forceinline__ int lcp_more(const EditJob& J, const Grp<G>& g, int v, int h, int rem,
                                        bool c, int n) {
  // the first continuation window is 64 bits (most runs of noisy reads end inside it), the later
  // ones 128 bits (63 bases at 2 bit: a long run of low-error reads needs half the iterations).
  // WIDE_ALL (every ring but the 1024-slot one, i.e. long reads and overflow pairs): 128 bits from the start - measured on C3, and the
  // two-flavour loop costs that kernel spills.  C1 25.3 -> 24.8 ms, C2/300 87.5 -> 86.9, C3/60 ~94 -> 91.
  constexpr int WE = Pack<BITS>::kWin - 1, WE2 = 2 * Pack<BITS>::kWin - 1;
  unsigned pend;
  for (int it = 0;; ++it) {
    if (c) {
      const int left = rem - n;
      if (!WIDE_ALL && it == 0) {
        const uint32_t fd = first_diff<BITS>(J, v + n, h + n);
        n += (int)min(fd, (uint32_t)left);
        c = fd == (uint32_t)WE && left > WE;
      } else {
        const uint32_t fd = first_diff_wide<BITS>(J, v + n, h + n);
        n += (int)min(fd, (uint32_t)left);
        c = fd == (uint32_t)WE2 && left > WE2;
      }
    }
    pend = __ballot_sync(g.mask, c);
    if (pend == 0) return n;
    if (__popc(pend) <= 2 || it >= 2) break;
  }
  while (pend) {
    const int src = __ffs(pend) - 1;
    pend &= pend - 1;
    const int nn = __shfl_sync(g.mask, n, src);
    const int got = lcp_uniform<BITS, G>(J, __shfl_sync(g.mask, v, src) + nn, __shfl_sync(g.mask, h, src) + nn,
                                         __shfl_sync(g.mask, rem, src) - nn, g);
    if (g.gl + g.first == src) n += got;
  }
  return n;
}

// (m & a) | (~m & b) in one LOP3 (the compiler otherwise re-derives a predicate + SEL per round)
__device__ __forceinline__ int bitsel(int m, int a, int b) {
  int d;
  asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(d) : "r"(m), "r"(a), "r"(b));
  return d;
}

// Warp-uniform state of one alignment (lo/hi) + per-lane partials of the last pass.
struct WfState {
  int lo, hi;  // live diagonal range
  int dmin;    // per-lane partial minimum of the cut-off distance
  int dm;      // warp-uniform minimum of the cut-off distance after the last extend
  bool fin;    // per-lane: ends-free termination met on a diagonal this lane owns
};

// A value the compiler cannot re-derive, so it stays in its register instead of being recomputed
// (ISETP + SEL) in every round of the loop.
__device__ __forceinline__ int opaque(int x) {
  asm volatile("" : "+r"(x));
  return x;
}

// One pass over the wavefront, 64 diagonals per round: lane L owns the adjacent diagonals
// kb + 2L and kb + 2L + 1 (one 64-bit ring access, two independent extensions in flight).
//   COMPUTE: build wavefront s+1 over [lo-1,hi+1] from wavefront s in place
//            (wavefront_compute_edit.c:70-97); the caller trims afterwards (wf_trim);
//   EXTEND : extend every live offset and gather the per-lane minimum of the cut-off distance
//            (wavefront_extend.c:203-260; wavefront_heuristic.c:223-239) and, ends-free, the
//            termination flag (wavefront_extend.c:122-168).
// Ring invariant (kept by every writer): slots outside [lo,hi] hold nulls, so the pass needs no
// range guards.  Within a pass a slot is only touched by its owning lane; passes are separated by
// __syncwarp().  The ALU pipe limits this loop (profiles/), hence: one rotating shuffle per
// neighbour with the round edges blended in by a LOP3, the sentinel window (first_diff), the
// cut-off distance taken from the stored value (a null gives a huge distance by itself), the
// slow-path test on the clamped counts alone, and no guards on the window loads.
struct LaneConst {  // per-lane constants of the pass, made once per pair
  int mk0, mk31;         // all-ones on the group's first / last lane
  int from_lo, from_hi;  // source lanes (warp lane ids) of the two rotating shuffles
  template <int G>
  __device__ __forceinline__ explicit LaneConst(const Grp<G>& g)
      : mk0(opaque(g.gl == 0 ? -1 : 0)), mk31(opaque(g.gl == G - 1 ? -1 : 0)),
        from_lo(opaque(g.first | ((g.gl + G - 1) & (G - 1)))),
        from_hi(opaque(g.first | ((g.gl + 1) & (G - 1)))) {}
};

template <int BITS, int CAP, bool EF, bool COMPUTE, bool EXTEND, int G, int FW = 0>
__device__ __forceinline__ void wf_pass(const Ring<CAP>& R, const EditJob& J, const Grp<G>& g,
                                        const LaneConst& C, WfState& S) {
  constexpr int WE = FirstWin<BITS, FW>::kWE;
  constexpr bool WIDE_ALL = CAP != 1024 || FW != 0;
  const int plen = J.plen, tlen = J.tlen, ak = tlen - plen;
  const int clo = COMPUTE ? S.lo - 1 : S.lo, chi = COMPUTE ? S.hi + 1 : S.hi;
  const int mk0 = C.mk0, mk31 = C.mk31, from_lo = C.from_lo, from_hi = C.from_hi;
  int kb = clo & ~1;  // even: the lane's pair of slots is 8-byte aligned
  int k0 = kb + 2 * g.gl;
  int dmin = kBigDist;
  bool fin = false;
  typename Ring<CAP>::Tok tc = R.tok(k0);
  int2 cur = R.ld2(tc);
  int prevy = kENull;  // lane 31: M[kb-1] of the next round
#ifdef OTGPU_UNROLL2
#pragma unroll 2
#endif
  for (; chi - kb >= G; kb += 2 * G, k0 += 2 * G) {  // full rounds while more than G diagonals remain
    const typename Ring<CAP>::Tok tn = R.tok(k0 + 2 * G);
    const int2 nxt = R.ld2(tn);  // this lane's slots in the next round (still old)
    int m0, m1;
    if (COMPUTE) {
      // left neighbour of slot 0: lane-1's slot 1; lane 0 takes lane 31's slot 1 of the last round
      const int left0 = __shfl_sync(g.mask, bitsel(mk31, prevy, cur.y), from_lo);
      // right neighbour of slot 1: lane+1's slot 0; lane 31 takes lane 0's slot 0 of the next round
      const int right1 = __shfl_sync(g.mask, bitsel(mk0, nxt.x, cur.x), from_hi);
      prevy = cur.y;
      m0 = max(cur.y, max(left0, cur.x) + 1);
      m1 = max(right1, max(cur.x, cur.y) + 1);
    } else {
      m0 = cur.x;
      m1 = cur.y;
    }
    cur = nxt;
    // in-bounds <=> offset >= 0, v >= 0 and neither sequence overrun
    const int v0 = m0 - k0, v1 = m1 - k0 - 1;
    const int pr0 = plen - v0, tr0 = tlen - m0, pr1 = plen - v1, tr1 = tlen - m1;
    const int rem0 = min(pr0, tr0), rem1 = min(pr1, tr1);
    const bool ok0 = (m0 | v0 | rem0) >= 0, ok1 = (m1 | v1 | rem1) >= 0;
    if (EXTEND) {
      int n0 = (int)min(first_window<BITS, FW>(J, v0, m0), (uint32_t)rem0);
      int n1 = (int)min(first_window<BITS, FW>(J, v1, m1), (uint32_t)rem1);
      // a count of WE = every compared base matched and more remain: rare, and tested on the
      // counts alone (lanes that are out of bounds may raise it spuriously; lcp_more re-checks)
      if (__any_sync(g.mask, (n0 == WE) | (n1 == WE))) {
        const bool c0 = ok0 && n0 == WE && rem0 > WE, c1 = ok1 && n1 == WE && rem1 > WE;
        if (__any_sync(g.mask, c0 | c1)) {  // each lane finishes one pending run per call
          const int n = lcp_more<BITS, G, WIDE_ALL>(J, g, c0 ? v0 : v1, c0 ? m0 : m1, c0 ? rem0 : rem1, c0 | c1,
                                                    c0 ? n0 : n1);
          if (c0) n0 = n; else n1 = n;
          if (__any_sync(g.mask, c0 & c1)) n1 = lcp_more<BITS, G, WIDE_ALL>(J, g, v1, m1, rem1, c0 & c1, n1);
        }
      }
      m0 += n0;
      m1 += n1;
      if (EF) {
        fin |= ok0 && ((tr0 - n0 <= 0 && pr0 - n0 <= J.pe) || (pr0 - n0 <= 0 && tr0 - n0 <= J.te));
        fin |= ok1 && ((tr1 - n1 <= 0 && pr1 - n1 <= J.pe) || (pr1 - n1 <= 0 && tr1 - n1 <= J.te));
      }
    }
    const int s0 = ok0 ? m0 : kENull, s1 = ok1 ? m1 : kENull;  // out of bounds / beyond range: nulls
    if (EXTEND) {
      // cut-off distance max(plen - v', tlen - h') = plen + max(k, tlen - plen) - offset
      dmin = min(dmin, plen + max(k0, ak) - s0);
      dmin = min(dmin, plen + max(k0 + 1, ak) - s1);
    }
    R.st2(tc, s0, s1);
    tc = tn;
  }
  if (kb <= chi) {
    // The last G diagonals or fewer: one diagonal per lane (kb + lane) instead of a 64-wide round
    // that would be at least half nulls (C1 26.4 -> 25.2 ms, C2/300 91.2 -> 87.6 ms).  Old values outside [clo,chi] are nulls (ring invariant):
    // lane G-1's right neighbour kb+G lies beyond chi; lane 0's left neighbour is the last
    // diagonal of the previous round, kept in lane G-1's prevy.
    const int k = kb + g.gl;
    const int o = R.ld(k);
    int m;
    if (COMPUTE) {
      const int left = __shfl_sync(g.mask, bitsel(mk31, prevy, o), from_lo);
      const int right = __shfl_sync(g.mask, bitsel(mk0, kENull, o), from_hi);
      m = max(right, max(left, o) + 1);
    } else {
      m = o;
    }
    const int v = m - k, pr = plen - v, tr = tlen - m, rem = min(pr, tr);
    const bool ok = (m | v | rem) >= 0;
    if (EXTEND) {
      int n = (int)min(first_window<BITS, FW>(J, v, m), (uint32_t)rem);
      if (__any_sync(g.mask, n == WE)) {
        const bool c = ok && n == WE && rem > WE;
        if (__any_sync(g.mask, c)) n = lcp_more<BITS, G, WIDE_ALL>(J, g, v, m, rem, c, n);
      }
      m += n;
      if (EF) fin |= ok && ((tr - n <= 0 && pr - n <= J.pe) || (pr - n <= 0 && tr - n <= J.te));
    }
    const int sv = ok ? m : kENull;
    if (EXTEND) dmin = min(dmin, plen + max(k, ak) - sv);
    R.st(k, sv);
  }
  __syncwarp(g.mask);  // the next pass (and the trim/cut-off probes) map diagonals to lanes differently
  if (EXTEND) { S.dmin = dmin; S.fin = fin; }
}

// Trim after a compute pass over [S.lo-1, S.hi+1] (wavefront_compute.c:376-409): the new limits
// are the first and last diagonal holding an in-bounds offset, i.e. a non-null slot.  Normally
// both are found by one 32-wide probe from each end.  Returns false if the wavefront is empty.
template <int CAP, int G>
__device__ __forceinline__ bool wf_trim(const Ring<CAP>& R, const Grp<G>& g, WfState& S) {
  const int clo = S.lo - 1, chi = S.hi + 1;
  // (slots outside [clo,chi] hold nulls, so the first probes need no range test)
  unsigned bl = __ballot_sync(g.mask, R.ld(clo + g.gl) >= 0);
  unsigned bh = __ballot_sync(g.mask, R.ld(chi - g.gl) >= 0);
  int lo = clo, hi = chi;
  while (bl == 0) {
    lo += G;
    if (lo > chi) return false;
    bl = __ballot_sync(g.mask, R.ld(lo + g.gl) >= 0 && lo + g.gl <= chi);
  }
  while (bh == 0) {  // a non-null slot exists, so this ends
    hi -= G;
    bh = __ballot_sync(g.mask, R.ld(hi - g.gl) >= 0 && hi - g.gl >= clo);
  }
  S.lo = lo + g.pos(bl);
  S.hi = hi - g.pos(bh);
  return true;
}

// wf_compute_distance_end2end, wavefront_heuristic.c:111-119
__device__ __forceinline__ int wf_dist(int off, int k, int plen, int ak) {
  return off >= 0 ? plen + max(k, ak) - off : kBigDist;
}

// WF-adaptive cut-off (wavefront_heuristic.c:285-328), run after every extend.
// dmin = warp-wide minimum distance of the live wavefront.
// FILL: write nulls over the diagonals that were cut off, so that every ring slot outside
// [lo,hi] is null (the invariant wf_pass relies on instead of range guards).
template <int CAP, bool FILL = false>
__device__ __forceinline__ void wf_cutoff(const Ring<CAP>& R, const int plen, const int tlen,
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
  if (FILL) {
    for (int k = lo + lane; k < nlo; k += 32) R.st(k, kENull);
    for (int k = hi - lane; k > nhi; k -= 32) R.st(k, kENull);
    __syncwarp();
  }
  S.lo = nlo;
  S.hi = nhi;
}

// Generic halves of the cut-off scan, for the rare step on which the 32-wide edge probes of
// wf_edges do not settle it (wavefront_heuristic.c:269-283).
template <int CAP, int G>
__device__ __noinline__ int wf_scan_lo(const Ring<CAP>& R, int from, int top, int plen, int ak, int dmin,
                                       const Grp<G>& g) {
  for (int kc = from; kc < top; kc += G) {
    const int k = kc + g.gl;
    const int d = wf_dist(R.ld(k), k, plen, ak);
    const unsigned b = __ballot_sync(g.mask, k < top && (d - dmin <= kMaxDist));
    if (b) return kc + g.pos(b);
  }
  return top;
}
template <int CAP, int G>
__device__ __noinline__ int wf_scan_hi(const Ring<CAP>& R, int from, int bot, int plen, int ak, int dmin,
                                       const Grp<G>& g) {
  for (int kc = from; kc > bot; kc -= G) {
    const int k = kc - g.gl;
    const int d = wf_dist(R.ld(k), k, plen, ak);
    const unsigned b = __ballot_sync(g.mask, k > bot && (d - dmin <= kMaxDist));
    if (b) return kc - g.pos(b);
  }
  return bot;
}

// Everything between two passes, on one pair of 32-wide probes of the ring (the first and the
// last 32 slots of the candidate range [clo,chi]):
//   trim         first / last non-null slot                      (wavefront_compute.c:376-409)
//   termination  end-to-end: cut-off distance 0; ends-free: flags (wavefront_extend.c:45-68,122-168)
//   cut-off      WF-adaptive(10,50,1) on the trimmed wavefront    (wavefront_heuristic.c:285-328)
// and nulls written over what the cut-off removed (ring invariant).  Returns 1 = alignment done,
// -1 = wavefront empty, 0 = go on with [S.lo,S.hi].
template <int CAP, bool EF, int G>
__device__ __forceinline__ int wf_edges(const Ring<CAP>& R, const int plen, const int tlen,
                                        const Grp<G>& g, const int clo, const int chi, WfState& S) {
  const int ak = tlen - plen;
  const int kl = clo + g.gl, kh = chi - g.gl;
  const int pl = R.ld(kl), ph = R.ld(kh);  // slots outside [clo,chi] hold nulls
  const unsigned bl = __ballot_sync(g.mask, pl >= 0), bh = __ballot_sync(g.mask, ph >= 0);
  int lo, hi;
  if (bl != 0 && bh != 0) {
    lo = clo + g.pos(bl);
    hi = chi - g.pos(bh);
  } else {  // more than G nulls at an end (or nothing left)
    S.lo = clo + 1;
    S.hi = chi - 1;
    if (!wf_trim<CAP, G>(R, g, S)) return -1;
    lo = S.lo;
    hi = S.hi;
  }
  const int dm = __reduce_min_sync(g.mask, S.dmin);
  S.dm = dm;
  if (EF ? __any_sync(g.mask, S.fin) : (dm == 0)) return 1;
  if (hi - lo + 1 >= kMinWfLen) {
    const int dmin = min(max(plen, tlen), dm);  // :230-237
    const int top = min(ak, hi);                // :269
    int nlo = lo;
    if (lo < top) {
      const int d = plen + max(kl, ak) - pl;  // a null slot gives a distance in the millions
      const unsigned b = __ballot_sync(g.mask, kl >= lo && kl < top && d - dmin <= kMaxDist);
      if (b) nlo = clo + g.pos(b);
      else if (clo + G >= top) nlo = top;
      else nlo = wf_scan_lo<CAP, G>(R, clo + G, top, plen, ak, dmin, g);
    }
    const int bot = max(ak, nlo);  // :277 (uses the already reduced lo)
    int nhi = hi;
    if (hi > bot) {
      const int d = plen + max(kh, ak) - ph;
      const unsigned b = __ballot_sync(g.mask, kh <= hi && kh > bot && d - dmin <= kMaxDist);
      if (b) nhi = chi - g.pos(b);
      else if (chi - G <= bot) nhi = bot;
      else nhi = wf_scan_hi<CAP, G>(R, chi - G, bot, plen, ak, dmin, g);
    }
    if (kl >= lo && kl < nlo) R.st(kl, kENull);
    if (kh <= hi && kh > nhi) R.st(kh, kENull);
    for (int k = kl + G; k < nlo; k += G) R.st(k, kENull);
    for (int k = kh - G; k > nhi; k -= G) R.st(k, kENull);
    __syncwarp(g.mask);
    lo = nlo;
    hi = nhi;
  }
  S.lo = lo;
  S.hi = hi;
  return 0;
}

// Consecutive score steps of a NARROW wavefront, held in registers: lane L owns diagonal kb + L
// for as long as the computed range [lo-1,hi+1] stays inside lanes 1..30 (lanes 0 and 31 are
// always null, so the neighbour shuffles need no edge masks); the window is re-centred when the
// range drifts to an edge.  No ring traffic, no __syncwarp: compute, extend, trim, termination and
// cut-off of a step are ~80 instructions.  Low-error reads (HiFi) spend most of their steps here.
// The ring is all nulls while the wavefront lives in registers; it is written back on the way out.
// Same order of operations as wf_pass + wf_edges.  Returns 1 = alignment done, -1 = wavefront
// empty, 0 = the next step needs more than 30 lanes (wavefront [S.lo,S.hi] is back in the ring).
template <int BITS, int CAP, bool EF, int FW = 0>
__device__ __forceinline__ int wf_narrow_run(const Ring<CAP>& R, const EditJob& J, const int lane,
                                             WfState& S, int& s) {
  constexpr int WE = FirstWin<BITS, FW>::kWE;
  constexpr bool WIDE_ALL = CAP != 1024 || FW != 0;
  const int plen = J.plen, tlen = J.tlen, ak = tlen - plen;
  const int dcap = max(plen, tlen);
  int lo = S.lo, hi = S.hi;
  // centre the computed range [lo-1,hi+1] (<= 30 wide) in lanes 1..30
  int kb = lo - 2 - ((28 - (hi - lo)) >> 1);
  int k = kb + lane;
  int m = R.ld(k);        // nulls outside [lo,hi] (ring invariant)
  R.st(k, kENull);        // the ring holds nothing while the wavefront is in registers
  int kdist = plen + max(k, ak);  // cut-off distance of an offset on diagonal k = kdist - offset
  int dm = S.dm;
  int st;
  for (;;) {
    ++s;
    // ---- compute (wavefront_compute_edit.c:70-97) ----
    const int left = __shfl_up_sync(kFull, m, 1), right = __shfl_down_sync(kFull, m, 1);
    m = max(right, max(left, m) + 1);
    const int v = m - k, pr = plen - v, tr = tlen - m, rem = min(pr, tr);
    const bool ok = (m | v | rem) >= 0;
    // ---- extend (wavefront_extend.c:203-260) ----
    int n = (int)min(first_window<BITS, FW>(J, v, m), (uint32_t)rem);
    if (__any_sync(kFull, n == WE)) {
      const bool c = ok && n == WE && rem > WE;
      if (__any_sync(kFull, c)) n = lcp_more<BITS, 32, WIDE_ALL>(J, Grp<32>(lane), v, m, rem, c, n);
    }
    bool fin = false;
    if (EF) fin = ok && ((tr - n <= 0 && pr - n <= J.pe) || (pr - n <= 0 && tr - n <= J.te));
    m = ok ? m + n : kENull;
    const int dist = kdist - m;  // in the millions for a null
    // ---- trim, termination, cut-off: three votes back to back, then straight-line lane-mask
    // arithmetic (wavefront_compute.c:376-409; wavefront_extend.c:45-68,122-168;
    // wavefront_heuristic.c:285-328) ----
    const unsigned okb = __ballot_sync(kFull, ok);
    dm = __reduce_min_sync(kFull, dist);
    const bool done = EF ? __any_sync(kFull, fin) : (dm == 0);
    const unsigned keepb = __ballot_sync(kFull, dist <= min(dcap, dm) + kMaxDist);  // :230-237
    if (okb == 0) { st = -1; break; }
    if (done) { st = 1; break; }
    int llo = __clz(__brev(okb)), lhi = 31 - __clz(okb);  // trim: first / last in-bounds diagonal
    // cut-off only from kMinWfLen diagonals on: below that every live diagonal is kept
    const unsigned kp = (lhi - llo + 1 >= kMinWfLen) ? keepb : okb;
    const int first = __clz(__brev(kp)), last = 31 - __clz(kp);  // kp != 0: the minimum is kept
    const int lak = ak - kb;  // lane of the target diagonal (may lie outside 0..31)
    // lo: first kept diagonal below top = min(ak, hi), else top; untouched when lo >= top (:269-275)
    const int nlo = max(llo, min(first, min(lak, lhi)));
    // hi: last kept diagonal above bot = max(ak, lo'), else bot; untouched when hi <= bot (:277-283)
    const int nhi = min(lhi, max(last, max(lak, nlo)));
    if ((unsigned)(lane - nlo) > (unsigned)(nhi - nlo)) m = kENull;
    llo = nlo;
    lhi = nhi;
    lo = kb + llo;
    hi = kb + lhi;
    // ---- room for the next step's range [lo-1,hi+1] in lanes 1..30? ----
    if (llo < 2 || lhi > 29) {
      if (lhi - llo > 27) { st = 0; break; }
      const int nkb = lo - 2 - ((28 - (hi - lo)) >> 1);
      const int src = lane + (nkb - kb);
      const int mv = __shfl_sync(kFull, m, src & 31);
      m = (unsigned)src < 32u ? mv : kENull;
      kb = nkb;
      k = kb + lane;
      kdist = plen + max(k, ak);
    }
  }
  S.lo = lo;
  S.hi = hi;
  S.dm = dm;
  if (st == 0) {
    R.st(k, m);
    __syncwarp();
  }
  return st;
}

// Exact prune, end-to-end only, width >= 1000 (wavefront_compute_edit.c:214-269).
// Returns false if nothing survives (cannot happen: the minimum is attained inside [lo,hi]).
template <int CAP>
__device__ __forceinline__ bool wf_prune(const Ring<CAP>& R, const int plen, const int tlen,
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
  for (int k = lo + lane; k < nlo; k += 32) R.st(k, kENull);  // keep the ring invariant
  for (int k = hi - lane; k > nhi; k -= 32) R.st(k, kENull);
  __syncwarp();
  S.lo = nlo;
  S.hi = nhi;
  return true;
}

// One pair on one warp.  Returns score >= 0, OTGPU_SCORE_UNFEASIBLE (-1), or kScoreOverflow
// (-2, never when WIDE).
template <int BITS, int CAP, bool EF, int FW = 0>
__device__ int wfa_edit_run(const Ring<CAP>& R, const EditJob& J, const int lane) {
  const int plen = J.plen, tlen = J.tlen, ak = tlen - plen;
  const int cap = R.cap();
  WfState S;
  S.dm = 0;
  int s = 0;

  const Grp<32> g(lane);
  // ring invariant: every slot outside the live range [lo,hi] holds a null
  R.fill_null(lane);
  __syncwarp();
  // ---- wavefront 0 (wavefront_align.c:113-135 / :190-237) ----
  if (!EF) {
    S.lo = S.hi = 0;
    if (lane == 0) R.st(0, 0);
  } else {
    S.lo = -J.pb;
    S.hi = J.tb;
    if (S.hi - S.lo + 136 > cap) return kScoreOverflow;
    for (int k = S.lo + lane; k <= S.hi; k += 32) R.st(k, k > 0 ? k : 0);
  }
  __syncwarp();
  const LaneConst C(g);
  wf_pass<BITS, CAP, EF, false, true>(R, J, g, C, S);
  int st = wf_edges<CAP, EF>(R, plen, tlen, g, S.lo, S.hi, S);

  while (st == 0) {
    if (S.hi - S.lo + 3 <= 30) {  // the computed range fits lanes 1..30: register-resident steps
      st = wf_narrow_run<BITS, CAP, EF, FW>(R, J, lane, S, s);
      continue;
    }
    ++s;
    // the pass touches slots [lo-2, hi+1+127] and relies on nulls outside [lo,hi]: no aliasing
    if (CAP != 0 && S.hi - S.lo + 136 > cap) return kScoreOverflow;
    // Exact prune (wavefront_compute_edit.c:214-269, call site :341-347), on every ring that can
    // hold 1000 diagonals.  It sits between compute and extend, which the fused pass cannot
    // separate, so the step is split only when the prune could do anything.  It acts iff
    // |lo'-ak| or |hi'-ak| (best case of an edge diagonal of the trimmed wavefront [lo',hi'],
    // inside [lo-1,hi+1]) exceeds the worst case max(plen-v, tlen-h) of the middle diagonal.
    // Every offset of wavefront s+1 is at most one past a neighbouring offset of wavefront s, so
    // that worst case is >= (minimum cut-off distance of wavefront s) - 1 = S.dm - 1: when both
    // edge bounds stay below it the reference's prune returns at :234 and the fused step is exact.
    if ((CAP == 0 || CAP >= kPruneLen + 136) && !EF && S.hi - S.lo + 3 >= kPruneLen &&
        max(abs(S.lo - 1 - ak), abs(S.hi + 1 - ak)) >= S.dm - 1) {
      wf_pass<BITS, CAP, EF, true, false>(R, J, g, C, S);
      if (!wf_trim<CAP>(R, g, S)) return -1;
      if (!wf_prune<CAP>(R, plen, tlen, lane, S)) return -1;
      wf_pass<BITS, CAP, EF, false, true>(R, J, g, C, S);
      st = wf_edges<CAP, EF>(R, plen, tlen, g, S.lo, S.hi, S);
    } else {
      wf_pass<BITS, CAP, EF, true, true, 32, FW>(R, J, g, C, S);
      st = wf_edges<CAP, EF>(R, plen, tlen, g, S.lo - 1, S.hi + 1, S);
    }
  }
  return st > 0 ? s : -1;  // done | empty wavefront -> WF_STATUS_UNFEASIBLE
}

template <int BITS, int CAP, int FW = 0>
__device__ __forceinline__ int wfa_edit_pair(const Ring<CAP>& R, const EditJob& J, const int lane) {
  return J.pb >= 0 ? wfa_edit_run<BITS, CAP, true, FW>(R, J, lane)
                   : wfa_edit_run<BITS, CAP, false, FW>(R, J, lane);
}

}  // namespace otgpu
