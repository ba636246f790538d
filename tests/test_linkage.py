"""Average linkage and tree cuts (otter_b200/host/linkage_b200.hpp) against hclust-cpp as the
reference calls it (src/otterclust.cpp:183-187,226,241): merge matrix, heights and labels bit for
bit, on golden vectors everywhere and live where oracle/_ref is built.  Host code only."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARGS = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]


def run(fn, dist, n, k, cut):
    merge = np.zeros(2 * (n - 1), dtype=np.int32)
    height = np.zeros(n - 1, dtype=np.float64)
    lk, lh = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
    fn(S.ptr(dist), n, k, cut, S.ptr(merge), S.ptr(height), S.ptr(lk), S.ptr(lh))
    return merge.tolist(), [float(x).hex() for x in height], lk.tolist(), lh.tolist()


def matrix(rng, n, kind):
    m = n * (n - 1) // 2
    if kind == 0:
        d = rng.random(m)
    elif kind == 1:      # many exact ties
        d = rng.integers(0, 6, m) / 8.0
    elif kind == 2:      # two tight groups
        lab = rng.integers(0, 2, n)
        d = np.array([abs(rng.normal(0.03 if lab[i] == lab[j] else 0.3, 0.01))
                      for i in range(n) for j in range(i + 1, n)])
    else:                # quotients score / length
        d = rng.integers(0, 400, m) / rng.integers(400, 4000, m)
    return np.ascontiguousarray(d, dtype=np.float64)


@pytest.fixture(scope="module")
def host(built):
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    l.othost_hclust_average.restype = None
    l.othost_hclust_average.argtypes = ARGS
    return l


def test_linkage_golden(host):
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "linkage_golden.json")))["cases"]
    assert len(gold) >= 40
    for c in gold:
        dist = np.array([float.fromhex(x) for x in c["dist"]], dtype=np.float64)
        got = run(host.othost_hclust_average, dist, c["n"], c["k"], float.fromhex(c["cut"]))
        assert got == (c["merge"], c["height"], c["labels_k"], c["labels_h"]), (c["n"], c["k"])


@pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref is built in the build container only")
def test_linkage_live_vs_hclust_cpp(host):
    R = S.ref()
    R.ref_hclust_average.restype = None
    R.ref_hclust_average.argtypes = ARGS
    rng = np.random.default_rng(2718)
    for t in range(400):
        n = int(rng.integers(2, 120))
        dist = matrix(rng, n, t % 4)
        k = int(rng.integers(0, n + 2))
        cut = float(rng.choice(dist)) if t % 2 else float(rng.random())
        assert run(host.othost_hclust_average, dist, n, k, cut) == run(R.ref_hclust_average, dist, n, k, cut), (t, n, k)
