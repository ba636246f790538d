// otter_b200/host/ppoa_b200.hpp — otter's partial-order consensus restated so
// that it runs in time linear in the graph, with bit-identical output.
//
// What the reference does (src/anppoa.hpp): the medoid read is the backbone, one node per base
// (init :63-83); every other read of the allele is threaded through the graph along its CIGAR
// against the backbone (insert_alignment :112-233: 'M' follows / adds a backbone edge, 'X' and 'I'
// follow or create an alternate node, 'D' skips a backbone node; reads that do not span the left
// end skip their leading gap, reads that span the right end mark ending nodes); edge weights are
// damped (adjust_weights :235-244) and the consensus is the heaviest path into an ending node
// (consensus :305-369, set_heaviest_path :246-283).
//
// The reference finds a node's incoming edges by scanning every edge of the graph and carries a
// full copy of the path with every node: O(nodes x edges) time and O(nodes^2) copies per allele —
// seconds for 3-5 kb reads, more than the alignments it consumes once those run on the GPU.  Here
// the graph is a flat edge pool with per-node insertion-ordered lists, incoming edges are gathered
// once in the order the reference's scan produces (source id, then position in the source's
// list), nodes are settled in topological order and each keeps one predecessor.  Same float
// arithmetic (one float add per candidate, strict '>' so the first maximum wins), same final
// choice (lowest-numbered ending node of maximal weight), hence the same string.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace otb200 {

class Ppoa {
 public:
  Ppoa() {}
  explicit Ppoa(const std::string& backbone) { init(backbone); }

  // src/anppoa.hpp:63-83
  void init(const std::string& backbone) {
    base_.clear(); head_.clear(); tail_.clear(); edges_.clear(); starts_.clear(); ending_.clear();
    n_backbone_ = (uint32_t)backbone.size();
    // a backbone of one base gets a node without a base (the reference never names it)
    for (uint32_t i = 0; i < n_backbone_; ++i) new_node(n_backbone_ > 1 ? backbone[i] : '\0');
    if (n_backbone_ > 1) starts_.push_back(0);
    for (uint32_t i = 1; i < n_backbone_; ++i) {
      add_edge(i - 1, i);
      if (n_backbone_ - i <= 10) ending_[i] = 1;
    }
  }

  // src/anppoa.hpp:112-233.  cigar: M X I D over (backbone = pattern, sequence = text).
  void insert_alignment(const std::string& sequence, const std::string& cigar, bool is_spanning_l,
                        bool is_spanning_r) {
    uint32_t prev = 0;
    int64_t ref_i = 0;
    size_t tgt_i = 0, ci = 0;
    bool first = true;  // no node of this read placed yet
    if (!is_spanning_l) {  // the read starts inside the backbone: drop its leading gap (:116-130)
      first = false;
      for (; ci < cigar.size() && (cigar[ci] == 'D' || cigar[ci] == 'I'); ++ci) {
        if (cigar[ci] == 'D') prev = (uint32_t)++ref_i;
        else ++tgt_i;
      }
    }
    for (; ci < cigar.size(); ++ci) {
      const char op = cigar[ci];
      const char b = tgt_i < sequence.size() ? sequence[tgt_i] : '\0';
      if (op == 'M') {  // :138-143
        if (first || (int64_t)prev == ref_i) first = false;
        else add_edge(prev, (uint32_t)ref_i);
        prev = (uint32_t)ref_i;
        ++ref_i;
        ++tgt_i;
      } else if (op == 'X') {  // :145-186
        if (first) {  // a new way into the graph unless a start with this base exists
          bool known = false;
          for (const uint32_t s : starts_)
            if (base_[s] == b) { known = true; break; }
          if (!known) {
            prev = new_node(b);
            starts_.push_back(prev);
          }
          first = false;
        } else {
          prev = follow_or_branch(prev, b);
        }
        ++ref_i;
        ++tgt_i;
      } else if (op == 'D') {  // :188-195 (still "first" afterwards)
        ++ref_i;
        if (first) prev = (uint32_t)ref_i;
      } else if (op == 'I') {  // :196-229
        if (first) {
          prev = new_node(b);
          starts_.push_back(prev);
          first = false;
        } else {
          prev = follow_or_branch(prev, b);
        }
        ++tgt_i;
      }
      if ((int64_t)n_backbone_ - ref_i <= 10 && is_spanning_r && prev < ending_.size())  // :231
        ending_[prev] = 1;
    }
  }

  // src/anppoa.hpp:235-244
  void adjust_weights(float c, float t) {
    for (Edge& e : edges_) {
      const float t_applied = t * e.weight;
      const float cut = c > t_applied ? c : t_applied;
      e.weight = e.weight - cut;
    }
  }

  // src/anppoa.hpp:305-369 (appends to consensus_seq, like the reference)
  void consensus(std::string& consensus_seq) const {
    const uint32_t n = (uint32_t)base_.size();
    if (n == 0) return;
    // incoming edges per node, in the order of the reference's scan: by source id, then by the
    // position in the source's list (set_heaviest_path :250-255)
    std::vector<uint32_t> in_off(n + 1, 0);
    for (const Edge& e : edges_) ++in_off[e.sink + 1];
    for (uint32_t i = 0; i < n; ++i) in_off[i + 1] += in_off[i];
    std::vector<uint32_t> in_src(edges_.size()), fill(in_off.begin(), in_off.end() - 1);
    std::vector<float> in_w(edges_.size());
    for (uint32_t u = 0; u < n; ++u)
      for (int32_t e = head_[u]; e >= 0; e = edges_[e].next) {
        const uint32_t at = fill[edges_[e].sink]++;
        in_src[at] = u;
        in_w[at] = edges_[e].weight;
      }
    // settle nodes once all their sources are settled (:317-333); the order does not matter
    std::vector<uint32_t> pending(n), order;
    order.reserve(n);
    for (uint32_t u = 0; u < n; ++u) {
      pending[u] = in_off[u + 1] - in_off[u];
      if (pending[u] == 0) order.push_back(u);
    }
    std::vector<float> weight(n, 0.0f);
    std::vector<int64_t> pred(n, -1);
    std::vector<uint8_t> settled(n, 0);
    for (size_t at = 0; at < order.size(); ++at) {
      const uint32_t u = order[at];
      bool have = false;
      float best = 0.0f;
      for (uint32_t i = in_off[u]; i < in_off[u + 1]; ++i) {  // :262-275
        const float cand = weight[in_src[i]] + in_w[i];
        if (!have || cand > best) {
          have = true;
          best = cand;
          pred[u] = in_src[i];
        }
      }
      weight[u] = best;
      settled[u] = 1;
      for (int32_t e = head_[u]; e >= 0; e = edges_[e].next)
        if (--pending[edges_[e].sink] == 0) order.push_back(edges_[e].sink);
    }
    // heaviest settled ending node, lowest id first (:339-349)
    uint32_t end = 0;
    bool have_end = false;
    for (uint32_t u = 0; u < n; ++u)
      if (ending_[u] && settled[u] && (!have_end || weight[u] > weight[end])) {
        have_end = true;
        end = u;
      }
    std::vector<uint32_t> path;
    if (have_end)
      for (int64_t u = end; u >= 0; u = pred[u]) path.push_back((uint32_t)u);
    else
      path.push_back(0);  // the reference's default: node 0 alone (:335,357)
    for (size_t i = path.size(); i-- > 0;)
      if (base_[path[i]]) consensus_seq.push_back(base_[path[i]]);
  }

  uint32_t n_nodes() const { return (uint32_t)base_.size(); }
  uint32_t n_edges() const { return (uint32_t)edges_.size(); }

 private:
  struct Edge {
    uint32_t sink;
    float weight;
    int32_t next;  // next edge of the same source, in insertion order
  };
  std::vector<char> base_;            // node -> base ('\0': the nameless node of a 1-base backbone)
  std::vector<int32_t> head_, tail_;  // node -> first / last outgoing edge
  std::vector<Edge> edges_;
  std::vector<uint32_t> starts_;      // starting nodes, in creation order
  std::vector<uint8_t> ending_;       // node -> may end the consensus
  uint32_t n_backbone_ = 0;

  uint32_t new_node(char b) {
    base_.push_back(b);
    head_.push_back(-1);
    tail_.push_back(-1);
    ending_.push_back(0);
    return (uint32_t)base_.size() - 1;
  }

  // src/anppoa.hpp:95-110: one more read over an existing edge, or a new edge of weight 1
  void add_edge(uint32_t source, uint32_t sink) {
    if (source >= base_.size() || sink >= base_.size()) return;  // malformed CIGAR
    for (int32_t e = head_[source]; e >= 0; e = edges_[e].next)
      if (edges_[e].sink == sink) {
        edges_[e].weight += 1.0f;
        return;
      }
    edges_.push_back(Edge{sink, 1.0f, -1});
    const int32_t id = (int32_t)edges_.size() - 1;
    if (tail_[source] >= 0) edges_[tail_[source]].next = id;
    else head_[source] = id;
    tail_[source] = id;
  }

  // :160-185 / :205-228: follow the first outgoing edge into an alternate (non-backbone) node
  // that carries base b, else branch into a new node
  uint32_t follow_or_branch(uint32_t from, char b) {
    if (from >= base_.size()) return from;  // malformed CIGAR
    for (int32_t e = head_[from]; e >= 0; e = edges_[e].next)
      if (edges_[e].sink >= n_backbone_ && base_[edges_[e].sink] == b) {
        edges_[e].weight += 1.0f;
        return edges_[e].sink;
      }
    const uint32_t node = new_node(b);
    add_edge(from, node);
    return node;
  }
};

}  // namespace otb200
