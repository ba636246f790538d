"""Full-shape soak on the GPU box (-m gpu, about a minute): regions of the three BASELINE.json read
shapes (C1 HiFi, C2 ONT-like, C3 long mixed reads) and a block of adversarial shapes, distance
matrices from the product's host layer (fill_dist_matrices -> CUDA) against the UNMODIFIED
reference's fill_dist_matrix (oracle/_ref/libotter_ref.so, all host threads), compared as doubles
bit for bit; plus gap-affine and edit-distance CIGAR lengths against the reference aligners."""
import ctypes as C
import os

import numpy as np
import pytest

import support as S
from harness import synth

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref/libotter_ref.so is prebuilt in the build container")]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def host_lib():
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    l.othost_fill_dist_matrices.argtypes = [vp, vp, vp, vp, vp, C.c_uint32, vp, vp]
    return l


def reference_matrices(B):
    R = S.ref()
    want = np.empty(B["n_scores"], dtype=np.float64)
    R.ref_fill_dist_matrix_regions(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]),
                                   S.ptr(B["region_start"]), B["n_regions"], S.ptr(want),
                                   S.ptr(B["score_off"]), os.cpu_count() or 1)
    return want


def product_matrices(B):
    got = np.full(B["n_scores"], -7.0)
    host_lib().othost_fill_dist_matrices(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]), None,
                                         S.ptr(B["region_start"]), B["n_regions"], S.ptr(got),
                                         S.ptr(B["score_off"]))
    return got


@pytest.mark.parametrize("name,n_regions", [("C1", 200), ("C2", 50), ("C3", 10)])
def test_baseline_shapes_equal_the_reference(built, name, n_regions):
    B = synth.region_batch(synth.make_regions(name, 77, n_regions))
    got, want = product_matrices(B), reference_matrices(B)
    bad = np.nonzero(got.view(np.uint64) != want.view(np.uint64))[0]
    assert len(bad) == 0, (name, len(bad), bad[:5], got[bad[:5]], want[bad[:5]])
    assert B["n_scores"] > 5000


def test_adversarial_shapes_equal_the_reference(built):
    rng = np.random.default_rng(2718)
    regions = []
    for _ in range(150):
        tl = int(rng.choice([20, 80, 300, 1000, 2500, 5000, 9000]))
        e = float(rng.choice([0.0, 0.002, 0.01, 0.05, 0.12, 0.3]))
        n = int(rng.integers(2, 22))
        reads = S.tr_region(rng, n, tl, e, extra_frac=float(rng.choice([0.0, 0.2, 0.6])),
                            n_rate=float(rng.choice([0, 0, 0.1, 0.5])))
        if rng.random() < 0.1:
            reads[int(rng.integers(0, n))] = reads[0].copy()       # identical pair
        if rng.random() < 0.05:
            reads[int(rng.integers(0, n))] = np.frombuffer(b"N", dtype=np.uint8)
        regions.append(reads)
    # unrelated reads of very different length: large scores, wavefronts past 1000 diagonals,
    # where the exact prune acts (tests/test_prune.py)
    for _ in range(6):
        regions.append([S.ALPHA[rng.integers(0, 4, int(rng.integers(800, 4500)))] for _ in range(5)])
    B = synth.region_batch(regions)
    got, want = product_matrices(B), reference_matrices(B)
    bad = np.nonzero(got.view(np.uint64) != want.view(np.uint64))[0]
    assert len(bad) == 0, (len(bad), bad[:5], got[bad[:5]], want[bad[:5]])


def test_cigar_lengths_equal_the_reference(ctx):
    from otter_b200.capi import PAIR_DTYPE
    R = S.ref()
    rng = np.random.default_rng(31)
    B = synth.region_batch(synth.make_regions("C2", 78, 12) + synth.make_regions("C1", 78, 30))
    rs, lens = B["region_start"], B["lens"]
    rows = []
    for g in range(B["n_regions"]):
        a = int(rs[g])
        for r in range(a + 1, int(rs[g + 1])):
            p, t = (a, r) if lens[a] >= lens[r] else (r, a)
            d = int(lens[p]) - int(lens[t])
            rows.append((p, t, -1, 0, 0, 0) if rng.random() < 0.6 else (p, t, d // 2, d - d // 2, 0, 0))
    pairs = np.array(rows, dtype=PAIR_DTYPE)
    req = np.array(rows, dtype=np.int32)
    for edit, fn in ((False, R.ref_affine_cigars_batch), (True, R.ref_edit_cigars_batch)):
        sc, ln, begin, packed = ctx.cigars_packed(B["bases"], B["off"], lens, pairs, edit=edit)
        fn.restype = C.c_double
        clen = np.zeros(len(rows), dtype=np.int32)
        fn(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(lens), S.ptr(req), C.c_uint64(len(rows)),
           os.cpu_count() or 1, S.ptr(clen))
        assert np.array_equal(clen, ln), (edit, int((clen != ln).sum()))
