"""exp_host_libm (otter_b200/csrc/exp_host_libm.cuh) against the host libm's exp, bit for bit: the
device KDE is only a drop-in for KDE::f (reference src/ankde.cpp) if every
exponential has the host's bits.  Host code only."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHECK = r'''
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include "exp_host_libm.cuh"
static const uint64_t TAB[256] = {
#include "exp_table.inc"
};
int main() {
  std::mt19937_64 rng(2024);
  long bad = 0, n = 0;
  auto test = [&](double x) {
    volatile double xv = x;
    const double a = std::exp(xv), b = otb200::exp_host_libm(x, TAB);
    ++n;
    if (std::memcmp(&a, &b, 8) != 0) { if (bad < 5) std::printf("x=%a libm=%a port=%a\n", x, a, b); ++bad; }
  };
  std::uniform_real_distribution<double> u1(-760.0, 0.0), u2(-1.0, 1.0), u3(-746.0, -700.0), u4(-1100.0, 720.0);
  for (long i = 0; i < 1000000; ++i) { test(u1(rng)); test(u2(rng)); test(u3(rng)); test(u4(rng)); }
  for (double x : {0.0, -0.0, 1e-300, -1e-17, -745.13321910194122, -745.2, -744.0, -708.3, -708.4, 709.7,
                   709.8, 710.0, -1024.0, -1e300, 512.0, -512.0, -1023.9999})
    test(x);
  for (double h : {0.01, 0.015})  // the KDE's own arguments
    for (double x = 0.0; x <= 1.0; x += 0.0025)
      for (int j = 0; j < 500; ++j) {
        const double v = (double)(rng() % 100000) / 300000.0, t = (x - v) / h;
        test(-(t * t / 2));
      }
  std::printf("tested %ld mismatches %ld\n", n, bad);
  return bad != 0;
}
'''


def test_exp_table_is_reproducible(tmp_path):
    """exp_table.inc is what gen_exp_table.py derives from first principles."""
    src = os.path.join(ROOT, "otter_b200", "csrc")
    have = open(os.path.join(src, "exp_table.inc")).read()
    gen = tmp_path / "gen_exp_table.py"
    gen.write_text(open(os.path.join(src, "gen_exp_table.py")).read())
    subprocess.check_call([sys.executable, str(gen)])
    assert (tmp_path / "exp_table.inc").read_text() == have


def test_exp_port_equals_host_libm(tmp_path):
    (tmp_path / "check.cpp").write_text(CHECK)
    exe = str(tmp_path / "check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "otter_b200", "csrc"),
                           "-o", exe, str(tmp_path / "check.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-600:]
    assert "mismatches 0" in out.stdout
