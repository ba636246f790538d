// otter_b200/host/otterclust_b200.cpp — see otterclust_b200.hpp.
#include "otterclust_b200.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <thread>
#include <utility>

#include "../csrc/exp_host_libm.cuh"
#include "linkage_b200.hpp"
#include "device_turn.hpp"
#include "otter_b200.h"

namespace {

constexpr double kInterval = 0.0025;  // error_intervals, src/otterclust.cpp:158

// the abscissae of otter_find_clustering_dist (:25): repeated addition, not i * step
const std::vector<double>& kde_grid() {
  static const std::vector<double> grid = [] {
    std::vector<double> g;
    for (double x = 0.0; x <= 1.0; x += kInterval) g.push_back(x);
    return g;
  }();
  return grid;
}

typedef std::vector<std::pair<int, double>> Extrema;

// KDE::maximas, src/ankde.cpp:25-64: windowed sums (centre, then left, then right neighbours),
// alternating search for the next fall (a maximum) and the next rise (a minimum).
void window_extrema(const int radius, const std::vector<double>& dens, Extrema& maxs, Extrema& mins) {
  const int n = (int)dens.size();
  bool rising = true;
  double prev = 0.0;
  int prev_i = 1;
  for (int i = 1; i < n - 1; ++i) {
    double sum = 0.0;
    sum += dens[i];
    for (int j = 1; j < radius && i - j >= 0; ++j) sum += dens[i - j];
    for (int j = 1; j < radius && i + j < n; ++j) sum += dens[i + j];
    if (rising) {
      if (sum < prev) { rising = false; maxs.emplace_back(prev_i, prev); }
    } else if (sum > prev) {
      rising = true;
      mins.emplace_back(prev_i, prev);
    }
    prev = sum;
    prev_i = i;
  }
  if (rising) maxs.emplace_back(prev_i, prev);
}

struct Bound { double dist0, dist1, cut0; };

[[noreturn]] void fatal(const char* what) {
  std::cerr << "ERROR: " << what << std::endl;
  exit(1);
}

// otter_find_clustering_dist, src/otterclust.cpp:20-116, from the raw densities f(x_i)
Bound decision_bound(const int radius, std::vector<double>& dens) {
  double total = 0.0;
  for (const double d : dens) total += d;
  for (double& d : dens) d = d / total;
  Extrema maxs, mins;
  window_extrema(radius, dens, maxs, mins);
  if (maxs.empty()) fatal("failed to obtain maximas");
  auto at = [](const std::pair<int, double>& e) { return e.first * kInterval; };
  if (maxs.size() == 1) return {at(maxs[0]), at(maxs[0]), -1.0};
  if (mins.empty()) fatal("failed to obtain minimas");
  if (maxs.size() == 2) return {at(maxs[0]), at(maxs[1]), at(mins[0])};
  // three or more peaks: rank them by height, near-equal heights (<= 0.01) by position (:57-63)
  std::vector<int> rank(maxs.size());
  for (int i = 0; i < (int)rank.size(); ++i) rank[i] = i;
  std::sort(rank.begin(), rank.end(), [&maxs](const auto& a, const auto& b) {
    double diff = maxs[a].second - maxs[b].second;
    diff = diff > 0 ? diff : -diff;
    if (diff <= 0.01) return maxs[a].first < maxs[b].first;
    return maxs[a].second > maxs[b].second;
  });
  // drop a peak that follows one of the same height (:70-83)
  int last = 0, cur = 1;
  while (cur < (int)rank.size()) {
    const int gap = cur > last ? cur - last : last - cur;
    double fd = maxs[rank[cur]].second - maxs[rank[last]].second;
    fd = fd < 0 ? -fd : fd;
    if (gap == 1 && fd <= 0.01) {
      rank.erase(rank.begin() + cur);
      last = cur;
    }
    ++cur;
  }
  if (rank.size() < 2) return {at(maxs[0]), at(maxs[1]), at(mins[0])};
  int first = rank[0], second = rank[1];
  if (first > second) std::swap(first, second);
  int boundary = second - 1;
  if (boundary < 0 || boundary >= (int)mins.size()) {
    std::cerr << "ERROR: unexpected index for minimas: " << boundary << std::endl;
    exit(1);
  }
  if (second - first > 1 && second - 2 >= 0 &&
      (maxs[second].first * kInterval - mins[boundary].first * kInterval <= 0.01)) {
    boundary = second - 2;
    if (boundary < 0 || boundary >= (int)mins.size()) {
      std::cerr << "ERROR: unexpected index for minimas after correction: " << boundary << std::endl;
      exit(1);
    }
  }
  const int middle = first + (second - first) / 2;
  if (middle >= (int)mins.size()) fatal("unexpected index for minimas");  // (unchecked in the reference)
  return {at(maxs[first]), at(maxs[second]), at(mins[middle])};
}

// The device evaluates exp with the roundings of glibc's FMA variant (csrc/exp_host_libm.cuh).
// If this host's libm rounds differently (no FMA, another libm), device and host densities could
// differ in the last bit and move a cut: checked once, on the KDE's own range of arguments, and
// refused loudly rather than answered differently.
void check_host_exp_once() {
  static const bool ok = [] {
    static const uint64_t tab[256] = {
#include "../csrc/exp_table.inc"
    };
    uint64_t seed = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 20000; ++i) {
      seed = seed * 6364136223846793005ull + 1442695040888963407ull;
      const double t = (double)(seed >> 11) / 9007199254740992.0 * 42.0;  // |x - v| / h up to 42
      volatile double x = -(t * t / 2);
      const double a = std::exp(x), b = otb200::exp_host_libm(x, tab);
      if (std::memcmp(&a, &b, sizeof a) != 0) return false;
    }
    return true;
  }();
  if (!ok) fatal("otter_b200: this host's exp() is not the one the device KDE reproduces "
                 "(glibc >= 2.28 with FMA); use the reference otter_hclust on this machine");
}

bool needs_kde(const HclustJob& job, int max_alleles) {
  return job.indeces->size() >= 3 && max_alleles != 1;
}

}  // namespace

void hclust_densities(const int& max_alleles, const double& bandwidth_short, const int& bandwidth_length,
                      const double& bandwidth_long, wfa::WFAligner& aligner,
                      const std::vector<HclustJob>& jobs, std::vector<std::vector<double>>& densities,
                      std::vector<double>& bandwidths) {
  check_host_exp_once();
  const std::vector<double>& grid = kde_grid();
  densities.assign(jobs.size(), std::vector<double>());
  bandwidths.assign(jobs.size(), bandwidth_short);
  std::vector<double> values, h;
  std::vector<uint64_t> off{0};
  std::vector<size_t> owner;
  for (size_t j = 0; j < jobs.size(); ++j) {
    if (!needs_kde(jobs[j], max_alleles)) continue;
    for (const auto& i : *jobs[j].indeces)  // :164-170
      if ((int)(*jobs[j].reads)[i].seq.size() >= bandwidth_length) {
        bandwidths[j] = bandwidth_long;
        break;
      }
    const std::vector<double>& v = jobs[j].distmatrix->values;
    values.insert(values.end(), v.begin(), v.end());
    off.push_back(values.size());
    h.push_back(bandwidths[j]);
    owner.push_back(j);
  }
  // at most 65,535 regions per launch (grid y)
  for (size_t a = 0; a < owner.size(); a += 60000) {
    const size_t b = std::min(owner.size(), a + 60000);
    std::vector<uint64_t> rel(b - a + 1);
    for (size_t k = a; k <= b; ++k) rel[k - a] = off[k] - off[a];
    std::vector<double> out((b - a) * grid.size());
    otgpu_ctx* ctx = aligner.context();
    std::unique_lock<otb200::DeviceTurn> turn(otb200::device_turn(aligner.device()));
    const int st = ctx ? otgpu_kde_densities(ctx, values.data() + off[a], rel.data(), (uint32_t)(b - a),
                                             h.data() + a, grid.data(), (uint32_t)grid.size(), out.data())
                       : OTGPU_ERR_NO_DEVICE;
    turn.unlock();
    if (st != OTGPU_OK) {
      std::cerr << "ERROR: otter_b200 density estimates: " << otgpu_strerror(st) << " ("
                << (ctx ? otgpu_last_cuda_error(ctx) : "") << ")" << std::endl;
      exit(1);
    }
    for (size_t k = a; k < b; ++k)
      densities[owner[k]].assign(out.begin() + (k - a) * grid.size(), out.begin() + (k - a + 1) * grid.size());
  }
}

void hclust_finish(const int& max_alleles, const double& max_tolerable_diff, const double& min_cov_fraction,
                   const int& min_cov_fraction2_l, const double& min_cov_fraction2_f, HclustJob& job,
                   std::vector<double>& densities, double bandwidth) {
  const std::vector<int>& idx = *job.indeces;
  DistMatrix& dm = *job.distmatrix;
  std::vector<ANREAD>& reads = *job.reads;
  ClusteringStatus& out = *job.clustering;
  const int n = (int)idx.size();
  out.labels.resize(n, -1);
  auto one_cluster = [&] {
    out.set_global_label(0);
    out.ic = out.fc = 1;
  };
  if (n == 1) {  // :123-127
    out.labels[0] = 0;
    out.ic = out.fc = 1;
    return;
  }
  if (n == 2) {  // :129-148
    out.labels[0] = out.labels[1] = 0;
    if (max_alleles == 1 || dm.get_dist(0, 1) <= max_tolerable_diff) out.ic = out.fc = 1;
    else {
      out.labels[1] = 1;
      out.ic = out.fc = 2;
    }
    return;
  }
  if (max_alleles == 1) return one_cluster();  // :152-156
  int radius = int(max_tolerable_diff / kInterval);  // :160-161
  radius = radius < 1 ? 1 : radius;
  const Bound bound = decision_bound(radius, densities);
  if (bound.dist1 - bound.dist0 <= max_tolerable_diff) return one_cluster();  // :174-178
  // average linkage on a copy of the matrix, cut at the decision bound (:180-188)
  std::vector<int> labels(n), merge(2 * (n - 1));
  std::vector<double> height(n - 1), copy = dm.values;
  otb200::average_linkage(n, copy.data(), merge.data(), height.data());
  const double cut = bound.dist1 == bandwidth ? bound.dist1 : bound.cut0 + 0.0025;
  otb200::cut_tree_height(n, merge.data(), height.data(), cut, labels.data());
  int n_clusters = 0;
  for (int i = 0; i < n; ++i) n_clusters = std::max(n_clusters, labels[i]);
  ++n_clusters;
  out.ic = n_clusters;
  const int min_cov1 = int(n * min_cov_fraction + 0.5), min_cov2 = int(n * min_cov_fraction2_f + 0.5);
  if (max_alleles != 0) {
    // per cluster: size, longest read, the coverage it needs (:194-213)
    std::vector<int> count(n_clusters), longest(n_clusters), need(n_clusters);
    for (int i = 0; i < n; ++i) {
      ++count[labels[i]];
      const int len = (int)reads[idx[i]].seq.size();
      if (len > longest[labels[i]]) longest[labels[i]] = len;
    }
    for (int l = 0; l < n_clusters; ++l) need[l] = longest[l] < min_cov_fraction2_l ? min_cov1 : min_cov2;
    std::vector<int> seeds, outliers;
    for (int l = 0; l < n_clusters; ++l) (count[l] < need[l] ? outliers : seeds).push_back(l);
    if (seeds.empty()) {  // nothing is covered well enough: cut into max_alleles as is (:224-230)
      std::vector<int> again(n);
      otb200::cut_tree_k(n, merge.data(), max_alleles, again.data());
      labels = again;
      out.fc = max_alleles;
    } else if ((int)seeds.size() > max_alleles) {  // :240-243
      otb200::cut_tree_k(n, merge.data(), max_alleles, labels.data());
      out.fc = max_alleles;
    } else {
      // seeds are renumbered 0.., reads of outlier clusters join the cluster of their nearest
      // already labelled read, in read order (:245-300)
      std::vector<int> renumber(n_clusters, -1);
      for (int s = 0; s < (int)seeds.size(); ++s) renumber[seeds[s]] = s;
      for (int i = 0; i < n; ++i) labels[i] = renumber[labels[i]];
      for (int i = 0; i < n; ++i) {
        if (labels[i] != -1) continue;
        int closest = -1;
        double best = 100000.0;
        for (int j = 0; j < n; ++j)
          if (i != j && labels[j] != -1) {
            const double d = dm.get_dist(i, j);
            if (d < best) { closest = j; best = d; }
          }
        if (closest >= 0) labels[i] = labels[closest];
      }
      out.fc = (int)seeds.size();
    }
  }
  for (int i = 0; i < n; ++i) out.labels[i] = labels[i];
}

void otter_hclust_batch(const bool& ignore_haps, const int& max_alleles, const double& bandwidth_short,
                        const int& bandwidth_length, const double& bandwidth_long,
                        const double& max_tolerable_diff, const double& min_cov_fraction,
                        const int& min_cov_fraction2_l, const double& min_cov_fraction2_f,
                        wfa::WFAligner& aligner, std::vector<HclustJob>& jobs, int threads) {
  (void)ignore_haps;  // unused by the reference function as well
  std::vector<std::vector<double>> densities;
  std::vector<double> bandwidths;
  hclust_densities(max_alleles, bandwidth_short, bandwidth_length, bandwidth_long, aligner, jobs, densities,
                   bandwidths);
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  const size_t per = (jobs.size() + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    const size_t a = (size_t)t * per, b = std::min(jobs.size(), a + per);
    if (a >= b) break;
    pool.emplace_back([&, a, b] {
      for (size_t j = a; j < b; ++j)
        hclust_finish(max_alleles, max_tolerable_diff, min_cov_fraction, min_cov_fraction2_l,
                      min_cov_fraction2_f, jobs[j], densities[j], bandwidths[j]);
    });
  }
  for (auto& th : pool) th.join();
}
