// otter_b200/csrc/exp_host_libm.cuh — exp(double) with the results of the HOST's libm.
//
// otter's clustering cut depends on a kernel density estimate built from std::exp (reference
// src/ankde.cpp:9-23) and compared value against value (KDE::maximas :25-64), so a device
// version is only a drop-in if every exp() returns the host's bits.  glibc >= 2.28 computes
// double exp with the Arm Optimized Routines algorithm: k = round(x * 128/ln2), r = x - k ln2/128
// (two-part constant), 2^(k/128) from a 128-entry table, a degree-5 polynomial in r, and a
// subnormal-safe final scaling.  On x86-64 CPUs with FMA the dynamic loader picks the variant
// compiled with -mfma, whose roundings are reproduced here operation by operation (each fma
// below is one fused instruction there, each mul/add a separate one; __dmul_rn / __dadd_rn /
// __fma_rn keep nvcc from re-associating or fusing on its own).  The constants are the published
// ones of that algorithm; the table is regenerated from first principles (gen_exp_table.py).
// tests/test_kde.py checks this function against the host's exp() bit for bit, so a host whose
// libm differs is caught there rather than in a clustering result.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define OTB_HD __host__ __device__ __forceinline__
#else
#define OTB_HD inline
#endif

namespace otb200 {

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ double x_fma(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ double x_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double x_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ uint64_t x_bits(double a) { return (uint64_t)__double_as_longlong(a); }
__device__ __forceinline__ double x_dbl(uint64_t a) { return __longlong_as_double((long long)a); }
#else
inline double x_fma(double a, double b, double c) { return __builtin_fma(a, b, c); }
inline double x_mul(double a, double b) { volatile double r = a * b; return r; }
inline double x_add(double a, double b) { volatile double r = a + b; return r; }
inline uint64_t x_bits(double a) { uint64_t u; __builtin_memcpy(&u, &a, 8); return u; }
inline double x_dbl(uint64_t a) { double d; __builtin_memcpy(&d, &a, 8); return d; }
#endif

// tab points to the 256-entry table of exp_table.inc (constant or global memory on the device).
OTB_HD double exp_host_libm(double x, const uint64_t* __restrict__ tab) {
  const double InvLn2N = 0x1.71547652b82fep+7, Shift = 0x1.8p52;
  const double NegLn2hiN = -0x1.62e42fefa0000p-8, NegLn2loN = -0x1.cf79abc9e3b3ap-47;
  const double C2 = 0x1.ffffffffffdbdp-2, C3 = 0x1.555555555543cp-3, C4 = 0x1.55555cf172b91p-5,
               C5 = 0x1.1111167a4d017p-7;
  unsigned abstop = (unsigned)(x_bits(x) >> 52) & 0x7ff;
  if (abstop - 0x3c9u > 0x3eu) {          // |x| < 2^-54 or |x| >= 512
    if ((int)(abstop - 0x3c9u) < 0) return x_add(x, 1.0);
    if (abstop > 0x408u) {                // |x| >= 1024, inf, nan
      if (x_bits(x) == 0xfff0000000000000ull) return 0.0;
      if (abstop == 0x7ffu) return x_add(x, 1.0);
      return (x_bits(x) >> 63) ? 0.0 : x_dbl(0x7ff0000000000000ull);  // underflow / overflow
    }
    abstop = 0;                           // 512 <= |x| < 1024: scaled below
  }
  double kd = x_fma(x, InvLn2N, Shift);
  const uint64_t ki = x_bits(kd);
  kd = x_add(kd, -Shift);
  double r = x_fma(kd, NegLn2hiN, x);
  r = x_fma(kd, NegLn2loN, r);
  const unsigned idx = 2u * (unsigned)(ki & 127u);
  const double p23 = x_fma(C3, r, C2);
  const double tr = x_add(r, x_dbl(tab[idx]));            // tail + r
  uint64_t sbits = tab[idx + 1] + (ki << 45);
  const double r2 = x_mul(r, r);
  const double p45 = x_fma(r, C5, C4);
  const double t1 = x_fma(p23, r2, tr);
  const double r4 = x_mul(r2, r2);
  const double tmp = x_fma(r4, p45, t1);
  if (abstop != 0) {
    const double scale = x_dbl(sbits);
    return x_fma(scale, tmp, scale);
  }
  if ((ki & 0x80000000ull) == 0) {        // k > 0: result may overflow
    sbits -= 1009ull << 52;
    const double scale = x_dbl(sbits);
    return x_mul(0x1p1009, x_fma(scale, tmp, scale));
  }
  sbits += 1022ull << 52;                 // k < 0: result may be subnormal
  const double scale = x_dbl(sbits);
  const double prod = x_mul(scale, tmp);
  double y = x_add(scale, prod);
  if (y < 1.0) {                          // round once, at the subnormal precision
    double lo = x_add(x_add(scale, -y), prod);
    const double hi = x_add(1.0, y);
    lo = x_add(x_add(x_add(1.0, -hi), y), lo);
    y = x_add(x_add(hi, lo), -1.0);
    if (y == 0.0) y = 0.0;
  }
  return x_mul(0x1p-1022, y);
}

}  // namespace otb200
