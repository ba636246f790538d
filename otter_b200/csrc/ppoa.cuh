// otter_b200/csrc/ppoa.cuh — otter's partial-order consensus on the device, one warp per allele.
//
// Same graph and same string as otter_b200/host/ppoa_b200.hpp (reference src/anppoa.hpp:63-369):
// the medoid read is the backbone, every member read is threaded through the graph along its
// CIGAR, edge weights are damped, the consensus is the heaviest path into an ending node.  The
// graph is the host restatement's flat layout (per-node insertion-ordered edge lists in one edge
// pool); what changes is who does the work:
//   * a member's CIGAR is read 32 operations at a time; a run of 'M' is ONE parallel step (the
//     backbone edge i-1 -> i is edge number i-1 by construction, so the run adds 1 to a range of
//     edge weights), every other operation is executed by the whole warp in lock-step on uniform
//     state, lane 0 storing;
//   * incoming edges are gathered with their (source, position in the source's list) key, so that
//     "the first maximum in the reference's scan order" is a deterministic (weight, key) maximum
//     whatever order the lanes add them in;
//   * nodes are settled frontier by frontier (Kahn), 32 nodes per round when the frontier is wide.
// Float arithmetic is the host's: one add per candidate, strict '>' on the weight, and the damping
// without fused multiply-add (__fmul_rn / __fsub_rn).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace otgpu {

// Per-graph description (host -> device).  Offsets index the pooled arrays of PpoaPools.
struct PpoaGraphDev {
  uint32_t backbone;      // read index of the representative
  uint32_t first_member;  // first entry of this graph in the member table
  uint32_t n_members;
  float c, t;             // adjust_weights parameters (src/anppoa.hpp:235-244)
  uint64_t node_off;      // first node slot; the graph may use node_cap of them
  uint64_t edge_off;
  uint32_t node_cap, edge_cap;
  uint64_t start_off;     // starting-node list (n_members + 2 slots)
};
struct PpoaMemberDev {
  uint32_t read;
  uint32_t flags;      // bit0 is_spanning_l, bit1 is_spanning_r
  uint64_t cigar_off;
  uint32_t cigar_len;
  uint32_t pad;
};
struct PpoaPools {
  // nodes
  uint8_t* base;
  uint8_t* ending;
  uint8_t* done;       // settled (consensus)
  int32_t* head;
  int32_t* tail;
  uint32_t* in_off;    // node_cap + 1 per graph: indexed node_off + graph + i
  uint32_t* fill;      // also the path buffer at the end
  uint32_t* pending;
  float* weight;
  int32_t* pred;
  uint32_t* queue;
  // edges
  uint32_t* e_sink;
  float* e_w;
  int32_t* e_next;
  uint32_t* in_src;
  uint32_t* in_pos;
  float* in_w;
  // starts
  uint32_t* starts;
};

// The graph under construction: counters are warp-uniform registers.
struct PpoaBuild {
  uint8_t* base;
  uint8_t* ending;
  int32_t* head;
  int32_t* tail;
  uint32_t* e_sink;
  float* e_w;
  int32_t* e_next;
  uint32_t* starts;
  uint32_t n, ne, ns, nb;       // nodes, edges, starts, backbone length
  uint32_t node_cap, edge_cap;
  bool overflow;
};

__device__ __forceinline__ uint32_t ppoa_new_node(PpoaBuild& G, const int lane, char b) {
  const uint32_t id = G.n;
  if (id >= G.node_cap) { G.overflow = true; return id; }
  if (lane == 0) {
    G.base[id] = (uint8_t)b;
    G.ending[id] = 0;
    G.head[id] = -1;
    G.tail[id] = -1;
  }
  ++G.n;
  return id;
}

// src/anppoa.hpp:95-110: one more read over an existing edge, or a new edge of weight 1.
// Uniform: every lane walks the list (same addresses), lane 0 stores; the caller syncs the warp.
__device__ __forceinline__ void ppoa_add_edge(PpoaBuild& G, const int lane, uint32_t source, uint32_t sink) {
  if (source >= G.n || sink >= G.n) return;  // malformed CIGAR
  for (int32_t e = G.head[source]; e >= 0; e = G.e_next[e])
    if (G.e_sink[e] == sink) {
      if (lane == 0) G.e_w[e] += 1.0f;
      return;
    }
  const uint32_t id = G.ne;
  if (id >= G.edge_cap) { G.overflow = true; return; }
  if (lane == 0) {
    G.e_sink[id] = sink;
    G.e_w[id] = 1.0f;
    G.e_next[id] = -1;
    const int32_t tl = G.tail[source];
    if (tl >= 0) G.e_next[tl] = (int32_t)id;
    else G.head[source] = (int32_t)id;
    G.tail[source] = (int32_t)id;
  }
  ++G.ne;
}

// :160-185 / :205-228: follow the first outgoing edge into an alternate (non-backbone) node that
// carries base b, else branch into a new node.
__device__ __forceinline__ uint32_t ppoa_follow_or_branch(PpoaBuild& G, const int lane, uint32_t from, char b) {
  if (from >= G.n) return from;  // malformed CIGAR
  for (int32_t e = G.head[from]; e >= 0; e = G.e_next[e]) {
    const uint32_t s = G.e_sink[e];
    if (s >= G.nb && G.base[s] == (uint8_t)b) {
      if (lane == 0) G.e_w[e] += 1.0f;
      return s;
    }
  }
  const uint32_t node = ppoa_new_node(G, lane, b);
  __syncwarp();
  if (!G.overflow) ppoa_add_edge(G, lane, from, node);
  return node;
}

// One operation of a member's CIGAR, executed by the whole warp on uniform state
// (src/anppoa.hpp:136-231).
struct PpoaCursor {
  uint32_t prev;
  int64_t ref_i;
  uint32_t tgt_i;
  bool first;
};
__device__ __forceinline__ void ppoa_op(PpoaBuild& G, const int lane, PpoaCursor& C, const char op, const char b,
                                        const bool span_r) {
  if (op == 'M') {  // :138-143
    if (C.first || (int64_t)C.prev == C.ref_i) C.first = false;
    else ppoa_add_edge(G, lane, C.prev, (uint32_t)C.ref_i);
    C.prev = (uint32_t)C.ref_i;
    ++C.ref_i;
    ++C.tgt_i;
  } else if (op == 'X') {  // :145-186
    if (C.first) {  // a new way into the graph unless a start with this base exists
      bool known = false;
      for (uint32_t i = 0; i < G.ns; ++i)
        if (G.base[G.starts[i]] == (uint8_t)b) { known = true; break; }
      if (!known) {
        C.prev = ppoa_new_node(G, lane, b);
        if (lane == 0 && !G.overflow) G.starts[G.ns] = C.prev;
        ++G.ns;
      }
      C.first = false;
    } else {
      C.prev = ppoa_follow_or_branch(G, lane, C.prev, b);
    }
    ++C.ref_i;
    ++C.tgt_i;
  } else if (op == 'D') {  // :188-195 (still "first" afterwards)
    ++C.ref_i;
    if (C.first) C.prev = (uint32_t)C.ref_i;
  } else if (op == 'I') {  // :196-229
    if (C.first) {
      C.prev = ppoa_new_node(G, lane, b);
      if (lane == 0 && !G.overflow) G.starts[G.ns] = C.prev;
      ++G.ns;
      C.first = false;
    } else {
      C.prev = ppoa_follow_or_branch(G, lane, C.prev, b);
    }
    ++C.tgt_i;
  }
  if ((int64_t)G.nb - C.ref_i <= 10 && span_r && C.prev < G.n && lane == 0) G.ending[C.prev] = 1;  // :231
  __syncwarp();
}

// insert_alignment (src/anppoa.hpp:112-233) for one member.
__device__ __forceinline__ void ppoa_insert(PpoaBuild& G, const int lane, const uint8_t* __restrict__ seq,
                                            const uint32_t seq_len, const char* __restrict__ cigar,
                                            const uint32_t n_ops, const bool span_l, const bool span_r) {
  PpoaCursor C{0u, 0, 0u, true};
  uint32_t ci = 0;
  if (!span_l) {  // the read starts inside the backbone: drop its leading gap (:116-130)
    C.first = false;
    for (; ci < n_ops; ++ci) {
      const char op = cigar[ci];
      if (op == 'D') C.prev = (uint32_t)++C.ref_i;
      else if (op == 'I') ++C.tgt_i;
      else break;
    }
  }
  while (ci < n_ops && !G.overflow) {
    // 32 operations at a time; bit j of nonm = operation ci + j is not an 'M' (or past the end)
    const uint32_t at = ci + (uint32_t)lane;
    const char mine = at < n_ops ? cigar[at] : '\0';
    const unsigned nonm = __ballot_sync(kFull, mine != 'M');
    const uint32_t valid = min(32u, n_ops - ci);
    uint32_t pos = 0;
    while (pos < valid && !G.overflow) {
      const unsigned rest = nonm >> pos;
      uint32_t run = rest ? (uint32_t)(__ffs(rest) - 1) : 32u - pos;
      run = min(run, valid - pos);
      if (run > 0) {
        // the first 'M' of a run takes the general path; the rest walk the backbone: element j adds
        // one read to edge (ref_i + j - 1 -> ref_i + j), which is edge number ref_i + j - 1
        ppoa_op(G, lane, C, 'M', '\0', span_r);
        const uint32_t more = run - 1;
        if (more > 0) {
          if (C.ref_i >= 1 && (uint64_t)C.ref_i + more <= (uint64_t)G.nb) {
            if ((uint32_t)lane < more) {
              const uint32_t sink = (uint32_t)C.ref_i + (uint32_t)lane;
              G.e_w[sink - 1] += 1.0f;
              if ((int64_t)G.nb - ((int64_t)sink + 1) <= 10 && span_r) G.ending[sink] = 1;  // :231, prev = sink
            }
            C.prev = (uint32_t)C.ref_i + more - 1;
            C.ref_i += more;
            C.tgt_i += more;
            __syncwarp();
          } else {
            for (uint32_t j = 0; j < more; ++j) ppoa_op(G, lane, C, 'M', '\0', span_r);
          }
        }
        pos += run;
      }
      if (pos < valid) {  // the operation that ended the run
        const char op = (char)__shfl_sync(kFull, (int)mine, (int)pos);
        const char b = C.tgt_i < seq_len ? (char)seq[C.tgt_i] : '\0';
        ppoa_op(G, lane, C, op, b, span_r);
        ++pos;
      }
    }
    ci += valid;
  }
}


// Packed tie-break key of an incoming edge: the reference scans edges by source id, then by the
// position in the source's list, and keeps the FIRST maximum (set_heaviest_path :250-275).
__device__ __forceinline__ unsigned long long ppoa_key(uint32_t src, uint32_t pos) {
  return ((unsigned long long)src << 32) | pos;
}

// One warp per graph.  out: consensus characters of graph g at out + node_off[g] (at most node_cap),
// out_len[g] = their number, or -1 if the graph outgrew its pools (cannot happen with the caps the
// host computes; reported, not hidden).
__global__ void __launch_bounds__(128)
k_ppoa(const PpoaGraphDev* __restrict__ graphs, const uint32_t n_graphs,
       const PpoaMemberDev* __restrict__ members, const uint8_t* __restrict__ bases,
       const uint64_t* __restrict__ read_off, const uint32_t* __restrict__ read_len,
       const char* __restrict__ cigars, PpoaPools P, unsigned int* __restrict__ cursor,
       char* __restrict__ out, int32_t* __restrict__ out_len) {
  __shared__ unsigned int s_qtail[4];  // queue tail of each warp's graph
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (;;) {
    unsigned int g = 0;
    if (lane == 0) g = atomicAdd(cursor, 1u);
    g = __shfl_sync(kFull, g, 0);
    if (g >= n_graphs) break;
    const PpoaGraphDev D = graphs[g];
    PpoaBuild G;
    G.base = P.base + D.node_off;     G.ending = P.ending + D.node_off;
    G.head = P.head + D.node_off;     G.tail = P.tail + D.node_off;
    G.e_sink = P.e_sink + D.edge_off; G.e_w = P.e_w + D.edge_off; G.e_next = P.e_next + D.edge_off;
    G.starts = P.starts + D.start_off;
    G.n = G.ne = G.ns = 0;
    G.node_cap = D.node_cap; G.edge_cap = D.edge_cap;
    G.overflow = false;
    // ---- init (src/anppoa.hpp:63-83): one node per backbone base, edge i-1 -> i is edge i-1 ----
    const uint8_t* bb = bases + read_off[D.backbone];
    const uint32_t nb = read_len[D.backbone];
    G.nb = nb;
    if (nb > D.node_cap || (nb > 0 && nb - 1 > D.edge_cap)) G.overflow = true;
    if (!G.overflow) {
      for (uint32_t i = lane; i < nb; i += 32) {
        G.base[i] = nb > 1 ? bb[i] : (uint8_t)0;  // a backbone of one base gets a node without a base
        G.ending[i] = (i >= 1 && nb - i <= 10) ? 1 : 0;
        G.head[i] = i + 1 < nb ? (int32_t)i : -1;
        G.tail[i] = i + 1 < nb ? (int32_t)i : -1;
        if (i + 1 < nb) {
          G.e_sink[i] = i + 1;
          G.e_w[i] = 1.0f;
          G.e_next[i] = -1;
        }
      }
      G.n = nb;
      G.ne = nb > 0 ? nb - 1 : 0;
      if (nb > 1) {
        if (lane == 0) G.starts[0] = 0;
        G.ns = 1;
      }
      __syncwarp();
      // ---- thread the members through the graph, in order (:112-233) ----
      for (uint32_t m = 0; m < D.n_members && !G.overflow; ++m) {
        const PpoaMemberDev M = members[D.first_member + m];
        ppoa_insert(G, lane, bases + read_off[M.read], read_len[M.read], cigars + M.cigar_off, M.cigar_len,
                    (M.flags & 1u) != 0, (M.flags & 2u) != 0);
      }
    }
    if (G.overflow) {
      if (lane == 0) out_len[g] = -1;
      continue;
    }
    const uint32_t n = G.n, ne = G.ne;
    // ---- adjust_weights (:235-244), no fused multiply-add ----
    for (uint32_t e = lane; e < ne; e += 32) {
      const float w = G.e_w[e];
      const float t_applied = __fmul_rn(D.t, w);
      const float cut = D.c > t_applied ? D.c : t_applied;
      G.e_w[e] = __fsub_rn(w, cut);
    }
    __syncwarp();
    // ---- consensus (:305-369) ----
    uint32_t* in_off = P.in_off + D.node_off + g;  // n + 1 entries
    uint32_t* fill = P.fill + D.node_off;
    uint32_t* pending = P.pending + D.node_off;
    float* weight = P.weight + D.node_off;
    int32_t* pred = P.pred + D.node_off;
    uint8_t* done = P.done + D.node_off;
    uint32_t* queue = P.queue + D.node_off;
    uint32_t* in_src = P.in_src + D.edge_off;
    uint32_t* in_pos = P.in_pos + D.edge_off;
    float* in_w = P.in_w + D.edge_off;
    if (n == 0) {
      if (lane == 0) out_len[g] = 0;
      continue;
    }
    for (uint32_t i = lane; i <= n; i += 32) in_off[i] = 0;
    __syncwarp();
    for (uint32_t e = lane; e < ne; e += 32) atomicAdd(&in_off[G.e_sink[e] + 1], 1u);
    __syncwarp();
    {  // exclusive prefix over in_off[1..n] (in_off[0] = 0), 32 entries per round
      uint32_t carry = 0;
      for (uint32_t i0 = 1; i0 <= n; i0 += 32) {
        const uint32_t i = i0 + lane;
        uint32_t v = i <= n ? in_off[i] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const uint32_t o = __shfl_up_sync(kFull, v, d);
          if (lane >= d) v += o;
        }
        if (i <= n) in_off[i] = v + carry;
        carry += __shfl_sync(kFull, v, 31);
      }
    }
    __syncwarp();
    for (uint32_t u = lane; u < n; u += 32) {
      fill[u] = in_off[u];
      pending[u] = in_off[u + 1] - in_off[u];
      done[u] = 0;
      weight[u] = 0.0f;
      pred[u] = -1;
    }
    __syncwarp();
    // incoming edges with their scan-order key; the order inside a node's list is arbitrary
    for (uint32_t u = lane; u < n; u += 32) {
      uint32_t pos = 0;
      for (int32_t e = G.head[u]; e >= 0; e = G.e_next[e], ++pos) {
        const uint32_t at = atomicAdd(&fill[G.e_sink[e]], 1u);
        in_src[at] = u;
        in_pos[at] = pos;
        in_w[at] = G.e_w[e];
      }
    }
    __syncwarp();
    // settle nodes frontier by frontier (:317-333): queue[qs..qe) is the current frontier.  A node
    // is released by the lane whose atomicSub takes its pending count to zero, i.e. after every
    // one of its sources has been settled; it is processed in the next round, after a __syncwarp.
    uint32_t qe = 0;
    for (uint32_t u0 = 0; u0 < n; u0 += 32) {  // the graph's sources, in id order
      const uint32_t u = u0 + lane;
      const bool src = u < n && pending[u] == 0;
      const unsigned b = __ballot_sync(kFull, src);
      if (src) queue[qe + __popc(b & ((1u << lane) - 1u))] = u;
      qe += __popc(b);
    }
    if (lane == 0) s_qtail[warp] = qe;
    __syncwarp();
    uint32_t qs = 0;
    while (qs < qe) {
      for (uint32_t q = qs + lane; q < qe; q += 32) {
        const uint32_t u = queue[q];
        bool have = false;
        float best = 0.0f;
        unsigned long long bkey = 0;
        int32_t bsrc = -1;
        for (uint32_t i = in_off[u]; i < in_off[u + 1]; ++i) {  // :262-275: first maximum in key order
          const float cand = weight[in_src[i]] + in_w[i];
          const unsigned long long key = ppoa_key(in_src[i], in_pos[i]);
          if (!have || cand > best || (cand == best && key < bkey)) {
            have = true;
            best = cand;
            bkey = key;
            bsrc = (int32_t)in_src[i];
          }
        }
        weight[u] = best;
        pred[u] = bsrc;
        done[u] = 1;
        for (int32_t e = G.head[u]; e >= 0; e = G.e_next[e]) {
          const uint32_t sink = G.e_sink[e];
          if (atomicSub(&pending[sink], 1u) == 1u) queue[atomicAdd(&s_qtail[warp], 1u)] = sink;
        }
      }
      __syncwarp();
      qs = qe;
      qe = s_qtail[warp];
      __syncwarp();
    }
    // heaviest settled ending node, lowest id first (:339-349)
    bool have_end = false;
    float bw = 0.0f;
    uint32_t bu = 0;
    for (uint32_t u = lane; u < n; u += 32)
      if (G.ending[u] && done[u]) {
        const float w = weight[u];
        if (!have_end || w > bw) {  // ascending u per lane: the first maximum is kept
          have_end = true;
          bw = w;
          bu = u;
        }
      }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
      const bool oh = __shfl_xor_sync(kFull, (int)have_end, d) != 0;
      const float ow = __shfl_xor_sync(kFull, bw, d);
      const uint32_t ou = __shfl_xor_sync(kFull, bu, d);
      if (oh && (!have_end || ow > bw || (ow == bw && ou < bu))) {
        have_end = true;
        bw = ow;
        bu = ou;
      }
    }
    // the path, end to start, into `fill` (free again), then written start to end without the
    // nameless node of a one-base backbone
    uint32_t plen = 0;
    if (lane == 0) {
      if (have_end) {
        for (int64_t u = bu; u >= 0 && plen < n; u = pred[u]) fill[plen++] = (uint32_t)u;
      } else {
        fill[plen++] = 0;  // the reference's default: node 0 alone (:335,357)
      }
    }
    plen = __shfl_sync(kFull, plen, 0);
    __syncwarp();
    char* dst = out + D.node_off;
    uint32_t written = 0;
    for (uint32_t j0 = 0; j0 < plen; j0 += 32) {
      const uint32_t j = j0 + lane;
      const uint8_t ch = j < plen ? G.base[fill[plen - 1 - j]] : (uint8_t)0;
      const unsigned keep = __ballot_sync(kFull, ch != 0);
      if (ch != 0) dst[written + __popc(keep & ((1u << lane) - 1u))] = (char)ch;
      written += __popc(keep);
    }
    if (lane == 0) out_len[g] = (int32_t)written;
    __syncwarp();
  }
}

}  // namespace otgpu
