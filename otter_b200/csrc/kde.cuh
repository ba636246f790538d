// otter_b200/csrc/kde.cuh — the kernel density estimate behind otter's clustering cut, batched
// over regions (reference src/ankde.cpp:9-23, src/otterclust.cpp:20-29).
//
// Per region the reference evaluates f(x) = (1/n) * sum_v (1/h) * (1/sqrt(2 pi)) * exp(-((x-v)/h)^2 / 2)
// over the n(n-1)/2 pairwise distances v, in their stored order, at every grid point x (401 of
// them): 0.7 M exp per 60-read region, the whole cost of otter_hclust.  Here one thread owns one
// (region, grid point) and runs the same sum in the same order with the same roundings: IEEE
// double add / mul / div without contraction and exp_host_libm() (exp_host_libm.cuh) for the
// exponential, so the densities are the host's bits and every comparison downstream
// (KDE::maximas, the cut) falls the same way.  All threads of a block read the same distance at
// the same time (one broadcast load); FP64 pipe bound.
#pragma once
#include "exp_host_libm.cuh"

namespace otgpu {

__device__ const uint64_t kExpTab[256] = {
#include "exp_table.inc"
};

// values: all regions' distances back to back, region g = values[off[g] .. off[g+1]).
// grid  : the n_grid abscissae (the reference builds them by repeated addition; passed in).
// inv_h, h: per region 1/h and h; c1 = 1/sqrt(2*pi) as the host computes it.
// out   : n_regions x n_grid densities f(x) (before the reference's normalisation by their sum).
__global__ void __launch_bounds__(128)
k_kde(const double* __restrict__ values, const uint64_t* __restrict__ off,
      const double* __restrict__ grid, const int n_grid, const double* __restrict__ h,
      const double* __restrict__ inv_h, const double c1, double* __restrict__ out) {
  __shared__ uint64_t tab[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) tab[i] = kExpTab[i];
  __syncthreads();
  const int g = blockIdx.y, p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_grid) return;
  const uint64_t a = off[g], b = off[g + 1];
  const double x = grid[p], hh = h[g], c2 = inv_h[g];
  double total = 0.0;
  for (uint64_t i = a; i < b; ++i) {
    const double t = __ddiv_rn(__dadd_rn(x, -values[i]), hh);         // (x - v) / h
    const double e = otb200::exp_host_libm(-__dmul_rn(__dmul_rn(t, t), 0.5), tab);  // exp(-(t*t/2))
    total = __dadd_rn(total, __dmul_rn(c2, __dmul_rn(c1, e)));        // += (1/h) * ((1/sqrt(2pi)) * e)
  }
  out[(size_t)g * n_grid + p] = __ddiv_rn(total, (double)(b - a));     // / values.size()
}

}  // namespace otgpu
