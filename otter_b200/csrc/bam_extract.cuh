// otter_b200/csrc/bam_extract.cuh — batched read extraction for sm_100a.
//
// What otter does per BAM record that overlaps a region (reference src/anseqs.cpp):
//   get_breakpoints  :286-391  walk the CIGAR, find the query positions aligned closest to the
//                              region start (from the right) and end (from the left), the leftmost
//                              and rightmost aligned reference positions, the spanning flags and
//                              the clipped-end re-adjustment;
//   parse_alignment  :412-436  cut the query to [first, second), decode BAM's 4-bit bases
//                              (bam_seqi + seq_nt16_str), "N" for a deleted region / empty cut.
// The reference walks every aligned base of every match operation; the quantities it keeps are
// monotone in the reference position, so one step per CIGAR operation gives the same values:
//   * the first aligned base with rpos >= start is the closest one to start (later ones are
//     farther and the update is a strict '<');
//   * every aligned base with rpos <= end is closer to end than the one before it, so the last
//     such base of the last such operation wins;
//   * leftmost = first aligned base, rightmost = last aligned base.
// One thread per record (CIGARs are short compared to reads), then one warp per record decodes
// the cut straight from the 4-bit sequence.  Byte/integer work, HBM-bound: 0.5 byte read and
// 1 byte written per extracted base.
#pragma once
#include <cstdint>

#include "../../include/otter_b200.h"

namespace otgpu {

struct ExtractPlan {  // per record, device side
  int32_t first;      // query offset of the cut, -1 = no bases ("N")
  int32_t length;     // bases to decode (0 with first = -1: the record yields "N")
};

__device__ __forceinline__ void bam_breakpoints(const otgpu_bam_read& r, const uint32_t* __restrict__ cigar,
                                                uint8_t& status, int32_t& c0, int32_t& c1,
                                                ExtractPlan& plan, uint32_t& out_len) {
  const int start = r.start, end = r.end, n = (int)r.n_cigar, l_qseq = r.l_qseq;
  bool clipped_l = false, clipped_r = false, have_start = false;
  int leftmost_q = -1, rightmost_q = -1, leftmost_r = -1, rightmost_r = -1;
  int qstart_q = -1, qend_q = -1, qstart_ci = 0, qend_ci = 0;
  int rpos = r.pos, qpos = 0;
  for (int i = 0; i < n; ++i) {  // :299-343
    const uint32_t c = cigar[i];
    const int op = (int)(c & 0xf), ol = (int)(c >> 4);
    if (op == 5 || op == 4) {  // hard / soft clip
      if (i == 0) clipped_l = true;
      if (i == n - 1) clipped_r = true;
      if (op == 4) qpos += ol;
    } else if (op == 0 || op == 7 || op == 8) {  // M = X: aligned bases rpos .. rpos+ol-1
      if (ol > 0) {
        if (leftmost_q == -1) { leftmost_q = qpos; leftmost_r = rpos; }
        if (rightmost_q == -1 || rpos + ol - 1 > rightmost_r) {
          rightmost_q = qpos + ol - 1;
          rightmost_r = rpos + ol - 1;
        }
        if (!have_start && rpos + ol - 1 >= start) {
          const int at = rpos > start ? rpos : start;
          have_start = true;
          qstart_q = qpos + (at - rpos);
          qstart_ci = i;
        }
        if (rpos <= end) {
          const int at = rpos + ol - 1 < end ? rpos + ol - 1 : end;
          qend_q = qpos + (at - rpos);
          qend_ci = i;
        }
        rpos += ol;
        qpos += ol;
      }
    } else if (op == 1) {
      qpos += ol;
    } else if (op == 2) {
      rpos += ol;
    }
  }
  bool ok, span_l, span_r;
  c0 = -1;
  c1 = -1;
  if (rightmost_r < start || leftmost_r > end) {  // :346-352 does not reach the region
    qstart_q = qend_q = -1;
    ok = span_l = span_r = false;
  } else if (qstart_q > -1 && qend_q > -1 && qstart_q > qend_q) {  // :354-360 region deleted
    qstart_q = qend_q = -1;
    ok = span_l = span_r = true;
  } else {
    c0 = qstart_q;
    c1 = qend_q;
    if (leftmost_r > start && clipped_l && qstart_ci == 1) {  // :364-376
      while (qstart_q > 0 && qstart_ci > 0) {
        const int op = (int)(cigar[qstart_ci - 1] & 0xf), ol = (int)(cigar[qstart_ci - 1] >> 4);
        if (op == 2) --qstart_ci;
        else if (op == 5 || op == 4 || op == 1) { qstart_q -= ol; --qstart_ci; }
        else break;
      }
    }
    if (rightmost_r < end && clipped_r && qend_ci == n - 1) {  // :378-390
      while (qend_q < l_qseq - 1 && qend_ci < n) {
        const int op = (int)(cigar[qend_ci - 1] & 0xf), ol = (int)(cigar[qend_ci - 1] >> 4);
        if (op == 2) ++qend_ci;
        else if (op == 5 || op == 4 || op == 1) { qend_q += ol; ++qend_ci; }
        else break;
      }
    }
    span_l = leftmost_q >= 0 && leftmost_r <= start;  // :392-393
    span_r = rightmost_q >= 0 && rightmost_r >= end;
    ok = true;
  }
  status = (uint8_t)((ok ? OTGPU_READ_OK : 0) | (span_l ? OTGPU_READ_SPAN_L : 0) | (span_r ? OTGPU_READ_SPAN_R : 0));
  plan.first = -1;
  plan.length = 0;
  out_len = 0;
  if (!ok) return;
  // :397-405 the cut handed to parse_alignment
  int first, second;
  if (span_l && span_r) { first = qstart_q; second = qend_q; }
  else if (span_l) { first = qstart_q; second = l_qseq; }
  else if (span_r) { first = 0; second = qend_q; }
  else { first = 0; second = l_qseq; }
  // parse_alignment :417-434
  if ((first == -1) != (second == -1)) {  // the reference stops with an error here
    status |= OTGPU_READ_BAD_COORDS;
    out_len = 1;
    return;
  }
  if (first == -1 || l_qseq < second - first) {
    out_len = 1;  // "N"
    return;
  }
  const int l_sub = second - first, l_og = c1 - c0;
  c0 = c0 - first;
  c1 = c0 + l_og;
  if (l_sub <= 0) {
    out_len = 1;  // empty cut -> "N"
    return;
  }
  if (first < 0 || first + l_sub > l_qseq) {  // would read outside the record in the reference
    status |= OTGPU_READ_BAD_COORDS;
    out_len = 1;
    return;
  }
  plan.first = first;
  plan.length = l_sub;
  out_len = (uint32_t)l_sub;
}

__global__ void __launch_bounds__(256)
k_bam_breakpoints(const otgpu_bam_read* __restrict__ reads, const uint32_t n_reads,
                  const uint32_t* __restrict__ cigar, uint8_t* __restrict__ status,
                  int32_t* __restrict__ coords, ExtractPlan* __restrict__ plan,
                  int32_t* __restrict__ out_len) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_reads) return;
  const otgpu_bam_read r = reads[i];
  uint8_t st;
  int32_t c0, c1;
  ExtractPlan p;
  uint32_t len;
  bam_breakpoints(r, cigar + r.cigar_off, st, c0, c1, p, len);
  status[i] = st;
  coords[2 * i] = c0;
  coords[2 * i + 1] = c1;
  plan[i] = p;
  out_len[i] = (int32_t)len;
}

// One warp per record: decode plan.length 4-bit bases starting at query offset plan.first
// (bam_seqi: high nibble first) into characters (seq_nt16_str), or write "N".
__global__ void __launch_bounds__(256)
k_bam_decode(const otgpu_bam_read* __restrict__ reads, const uint32_t n_reads,
             const uint8_t* __restrict__ seq4, const ExtractPlan* __restrict__ plan,
             const int32_t* __restrict__ out_len, const uint64_t* __restrict__ begin,
             char* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const uint32_t warps = gridDim.x * (blockDim.x >> 5);
  for (uint32_t i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n_reads; i += warps) {
    const int len = out_len[i];
    if (len <= 0) continue;
    char* dst = out + begin[i];
    const ExtractPlan p = plan[i];
    if (p.first < 0) {
      if (lane == 0) dst[0] = 'N';
      continue;
    }
    const uint8_t* src = seq4 + reads[i].seq_off;
    for (int k = lane; k < p.length; k += 32) {
      const int q = p.first + k;
      const int code = (src[q >> 1] >> ((~q & 1) << 2)) & 0xf;
      dst[k] = "=ACMGRSVTWYHKDBN"[code];
    }
  }
}

}  // namespace otgpu
