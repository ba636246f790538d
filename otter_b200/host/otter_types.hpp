// otter_b200/host/otter_types.hpp — the reference types that cross the hot-path seam.
//
// In-tree build (-DOTTER_B200_IN_TREE, compiled inside holstegelab/otter): otter's own headers
// are used, nothing is redefined.  Stand-alone build (tests, bench, the GPU box): minimal mirrors
// with the same names, public fields and semantics:
//   ANSEQ / HAPLOTAG / ANREAD   reference src/anseqs.hpp:22-76, src/anseqs.cpp (ctors, is_spanning)
//   DistMatrix                  reference src/andistmat.hpp:8-19, src/andistmat.cpp:8-52
#pragma once
#ifdef OTTER_B200_IN_TREE
#include "andistmat.hpp"
#include "anseqs.hpp"
#else
#include <cstddef>
#include <cstdint>
#include <exception>
#include <string>
#include <utility>
#include <vector>

class ANSEQ {
 public:
  std::string seq;
  ANSEQ() {}
  explicit ANSEQ(const std::string& s) : seq(s) {}
};

class HAPLOTAG {
 public:
  int ps = -1;
  int hp = -1;
  HAPLOTAG() {}
  HAPLOTAG(int _ps, int _hp) : ps(_ps), hp(_hp) {}
  bool operator==(const HAPLOTAG& x) const { return ps == x.ps && hp == x.hp; }
  bool operator!=(const HAPLOTAG& x) const { return !(*this == x); }
  bool is_defined() const { return ps >= 0 && hp >= 0; }
};

class ANREAD : public ANSEQ {
 public:
  std::string name;
  double rq = -1.0;
  bool is_spanning_l = false;
  bool is_spanning_r = false;
  HAPLOTAG hpt;
  std::pair<int, int> ccoords{0, 0};
  ANREAD() {}
  ANREAD(const std::string& _name, const std::string& _seq) : ANSEQ(_seq), name(_name) {}
  bool is_spanning() const { return is_spanning_l && is_spanning_r; }
  void set_is_spanning() { is_spanning_l = is_spanning_r = true; }
};

// An assembled allele (reference src/anseqs.hpp:40-54); only the fields the hot path reads.
class ANALLELE : public ANSEQ {
 public:
  int scov = 0, acov = 0, tcov = 0;
  float se = 0.f;
  int ic = 0;
  HAPLOTAG hpt;
  ANALLELE() {}
  explicit ANALLELE(const std::string& s) : ANSEQ(s) {}
};

// Condensed upper-triangular distance store, row-major (0,1),(0,2)..(n-2,n-1): the layout
// hclust-cpp consumes (reference include/hclust-cpp/fastcluster.h:45-51).
class DistMatrix {
 public:
  uint32_t n;
  std::vector<double> values;
  explicit DistMatrix(uint32_t _n) : n(_n) { values.resize((size_t)n * (n ? n - 1 : 0) / 2, 1.0); }
  static size_t index(uint32_t n, uint32_t i, uint32_t j) {
    const int a = i < j ? i : j, b = i > j ? i : j;
    return (size_t)((static_cast<std::ptrdiff_t>(2 * n - 3 - a) * a >> 1) + b - 1);
  }
  void set_dist(uint32_t i, uint32_t j, double d) {
    if (i == j) throw std::exception();
    values[index(n, i, j)] = d;
  }
  double get_dist(uint32_t i, uint32_t j) const {
    if (i == j) throw std::exception();
    return values[index(n, i, j)];
  }
  uint32_t get_medoid(std::vector<uint32_t>& indeces) const {
    uint32_t min_i = indeces.front();
    double min_dist_sum = 100000000.0;
    for (const auto& i : indeces) {
      double dist_sum = 0.0;
      for (const auto& j : indeces)
        if (i != j) dist_sum += get_dist(i, j);
      if (dist_sum < min_dist_sum) {
        min_i = i;
        min_dist_sum = dist_sum;
      }
    }
    return min_i;
  }
  uint32_t size() const { return (uint32_t)values.size(); }
};
// src/otterclust.hpp:7-14 (what otter_hclust fills)
class ClusteringStatus {
 public:
  int ic;
  int fc;
  std::vector<int> labels;
  ClusteringStatus() : ic(0), fc(0) {}
  void set_global_label(int l) {
    for (uint32_t i = 0; i < labels.size(); ++i) labels[i] = l;
  }
};
#endif
