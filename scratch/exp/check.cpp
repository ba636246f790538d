#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include "../../otter_b200/csrc/exp_host_libm.cuh"
static const uint64_t TAB[256] = {
#include "../../otter_b200/csrc/exp_table.inc"
};
int main() {
  std::mt19937_64 rng(12345);
  long bad = 0, n = 0;
  auto test = [&](double x) {
    volatile double xv = x;
    double a = std::exp(xv), b = otb200::exp_host_libm(x, TAB);
    ++n;
    if (memcmp(&a, &b, 8) != 0) { if (bad < 10) printf("x=%a libm=%a mine=%a\n", x, a, b); ++bad; }
  };
  std::uniform_real_distribution<double> u1(-760.0, 0.0), u2(-1.0, 1.0), u3(-50.0, 50.0), u4(-746.0, -700.0), u5(-1100, 720);
  for (long i = 0; i < 20000000; ++i) { test(u1(rng)); test(u2(rng)); test(u3(rng)); test(u4(rng)); test(u5(rng)); }
  for (double x : {0.0, -0.0, 1e-300, -1e-300, 1e-17, -1e-17, -745.13321910194122, -745.2, -744.0, -708.3, -708.4, 709.7, 709.8, 710.0, -1024.0, -1e300, 512.0, -512.0, -1023.9999}) test(x);
  // the KDE's own arguments: -(t*t/2) with t = (x - v) / h on the 0.0025 grid
  for (double h : {0.01, 0.015}) for (double x = 0.0; x <= 1.0; x += 0.0025) for (int j = 0; j < 2000; ++j) {
    double v = (double)(rng() % 100000) / 300000.0; double t = (x - v) / h; test(-(t * t / 2));
  }
  printf("tested %ld, mismatches %ld\n", n, bad);
  return bad != 0;
}
