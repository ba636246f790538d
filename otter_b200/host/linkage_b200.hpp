// otter_b200/host/linkage_b200.hpp — average-linkage clustering and tree cuts as otter uses them
// (reference src/otterclust.cpp:180-188,224-243: hclust_fast(…, HCLUST_METHOD_AVERAGE, …),
// cutree_cdist, cutree_k of the vendored hclust-cpp / fastcluster, include/hclust-cpp/).
//
// Same algorithm and, more to the point, the same results bit for bit: the nearest-neighbour
// chain (Murtagh) over the condensed matrix with strict '<' searches in index order, the
// Lance-Williams update for average linkage d(k, i∪j) = s·d(k,i) + t·d(k,j) with s, t the size
// fractions (two products and one sum, in that order), merges stably sorted by height and
// renumbered to R's convention (-(i+1) for a leaf, step number for a cluster, smaller id first).
// Restated so the clustering step needs nothing from the reference tree and can sit next to the
// batched density estimates; n <= 200 reads per region, so this is microseconds of host time.
#pragma once
#include <algorithm>
#include <cstddef>
#include <vector>

namespace otb200 {

// D: condensed upper-triangular distances (row-major pairs, n(n-1)/2), overwritten.
// merge: 2*(n-1) ints, column-major (merge[i], merge[i+n-1]); height: n-1 doubles.
inline void average_linkage(const int n, double* D, int* merge, double* height) {
  if (n < 2) return;
  auto at = [&](int r, int c) -> double& {  // r < c
    return D[(static_cast<std::ptrdiff_t>(2 * n - 3 - r) * r >> 1) + c - 1];
  };
  // active nodes as a linked list over indices (succ[i] == 0 marks a removed node; index n ends it)
  std::vector<int> succ(n + 1), pred(n + 1);
  for (int i = 0; i < n; ++i) { pred[i + 1] = i; succ[i] = i + 1; }
  int start = 0;
  auto drop = [&](int i) {
    if (i == start) start = succ[i];
    else { succ[pred[i]] = succ[i]; pred[succ[i]] = pred[i]; }
    succ[i] = 0;
  };
  struct Step { int a, b; double d; };
  std::vector<Step> steps;
  steps.reserve(n - 1);
  std::vector<double> size(n, 1.0);
  std::vector<int> chain(n);
  int tip = 0, a = 0, b = 0;
  double best = 0.0;
  for (int j = 0; j < n - 1; ++j) {
    if (tip <= 3) {  // restart the chain at the first active node and its nearest neighbour
      chain[0] = a = start;
      tip = 1;
      b = succ[a];
      best = at(a, b);
      for (int i = succ[b]; i < n; i = succ[i])
        if (at(a, i) < best) { best = at(a, i); b = i; }
    } else {         // continue from three links back
      tip -= 3;
      a = chain[tip - 1];
      b = chain[tip];
      best = a < b ? at(a, b) : at(b, a);
    }
    do {  // follow nearest neighbours until two nodes point at each other
      chain[tip] = b;
      for (int i = start; i < b; i = succ[i])
        if (at(i, b) < best) { best = at(i, b); a = i; }
      for (int i = succ[b]; i < n; i = succ[i])
        if (at(b, i) < best) { best = at(b, i); a = i; }
      b = a;
      a = chain[tip++];
    } while (b != chain[tip - 2]);
    steps.push_back({a, b, best});
    if (a > b) std::swap(a, b);
    const double sa = size[a], sb = size[b];
    size[b] += size[a];
    drop(a);
    // distances of every other active node to the merged node, kept under the larger index
    const double s = sa / (sa + sb), t = sb / (sa + sb);
    int i = start;
    for (; i < a; i = succ[i]) at(i, b) = s * at(i, a) + t * at(i, b);
    for (; i < b; i = succ[i]) at(i, b) = s * at(a, i) + t * at(i, b);
    for (i = succ[b]; i < n; i = succ[i]) at(b, i) = s * at(a, i) + t * at(b, i);
  }
  // merges in order of height (stable), leaves and clusters renumbered to R's convention
  std::stable_sort(steps.begin(), steps.end(), [](const Step& x, const Step& y) { return x.d < y.d; });
  std::vector<int> parent(2 * n - 1, 0);
  int next = n;
  auto root = [&](int i) {
    if (parent[i] != 0) {
      int p = i;
      i = parent[i];
      if (parent[i] != 0) {
        do i = parent[i]; while (parent[i] != 0);
        do { const int up = parent[p]; parent[p] = i; p = up; } while (parent[p] != i);
      }
    }
    return i;
  };
  for (int k = 0; k < n - 1; ++k) {
    int x = root(steps[k].a), y = root(steps[k].b);
    parent[x] = parent[y] = next++;
    if (x > y) std::swap(x, y);
    merge[k] = x < n ? -x - 1 : x - n + 1;
    merge[k + n - 1] = y < n ? -y - 1 : y - n + 1;
    height[k] = steps[k].d;
  }
}

// Labels 0..k-1 in order of first appearance after undoing the last k-1 merges.
inline void cut_tree_k(const int n, const int* merge, const int k, int* labels) {
  if (k > n || k < 2) {
    for (int j = 0; j < n; ++j) labels[j] = 0;
    return;
  }
  std::vector<int> last(n, 0);  // per leaf: the last merge step (1-based) it took part in
  for (int step = 1; step <= n - k; ++step) {
    int m1 = merge[step - 1], m2 = merge[n - 1 + step - 1];
    if (m1 < 0 && m2 < 0) {
      last[-m1 - 1] = last[-m2 - 1] = step;
    } else if (m1 < 0 || m2 < 0) {
      int leaf;
      if (m1 < 0) { leaf = -m1; m1 = m2; } else leaf = -m2;
      for (int l = 0; l < n; ++l)
        if (last[l] == m1) last[l] = step;
      last[leaf - 1] = step;
    } else {
      for (int l = 0; l < n; ++l)
        if (last[l] == m1 || last[l] == m2) last[l] = step;
    }
  }
  int label = 0;
  std::vector<int> of_step(n, -1);
  for (int j = 0; j < n; ++j) {
    if (last[j] == 0) labels[j] = label++;
    else {
      if (of_step[last[j]] < 0) of_step[last[j]] = label++;
      labels[j] = of_step[last[j]];
    }
  }
}

// Stop merging at the first step whose height is >= cut.
inline void cut_tree_height(const int n, const int* merge, const double* height, const double cut,
                            int* labels) {
  int k = 0;
  while (k < n - 1 && !(height[k] >= cut)) ++k;
  cut_tree_k(n, merge, n - k, labels);
}

}  // namespace otb200
