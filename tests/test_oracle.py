"""CPU tests of the oracle itself: the C restatement (oracle/wfa_oracle.c) against the committed
golden vectors (generated from the unmodified reference) and, where the reference library was
built (this container), against the reference directly on fresh random inputs."""
import itertools

import numpy as np
import pytest

import support as S


def test_port_matches_golden_pairs(golden):
    P = S.port()
    for g in golden["pairs"]:
        p, t = g["p"].encode(), g["t"].encode()
        got = P.wfo_edit_score(p, len(p), t, len(t), g["ef"], g["pb"], g["pe"], g["tb"], g["te"], None)
        assert got == g["score"], (len(p), len(t), g["ef"], g["pb"], g["pe"], g["tb"], g["te"])


def test_port_matches_golden_align_anreads(golden):
    P = S.port()
    for g in golden["anreads"]:
        x, y = g["x"].encode(), g["y"].encode()
        d = P.wfo_align_anreads(x, len(x), g["xspan"], y, len(y), g["yspan"], None)
        assert d == float.fromhex(g["dist"])  # bit-exact doubles


def test_port_matches_golden_regions(golden):
    P = S.port()
    for g in golden["regions"]:
        reads = [np.frombuffer(r.encode(), dtype=np.uint8) for r in g["reads"]]
        bases, off, lens = S.pack_reads(reads)
        span = np.array(g["span"], dtype=np.uint8)
        idx = np.array(g["indeces"], dtype=np.int32)
        out = np.zeros(len(g["dist"]), dtype=np.float64)
        P.wfo_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(span), S.ptr(idx), len(idx),
                               S.ptr(out))
        exp = np.array([float.fromhex(h) for h in g["dist"]])
        assert np.array_equal(out, exp)


def test_golden_covers_the_hazards(golden):
    """The fixture must keep the parity hazards of SURVEY.md §8c in it."""
    pairs = golden["pairs"]
    assert any(g["ef"] == 1 and g["pb"] > 0 and g["pe"] > 0 for g in pairs)  # delta/2 form
    assert any(g["ef"] == 1 and g["tb"] + g["te"] > 0 for g in pairs)        # text free ends
    assert any("N" in g["p"] for g in pairs) and any(g["p"] == "N" for g in pairs)
    assert any(g["t"] == "" for g in pairs)
    assert any("max_width" in g.get("note", "") and int(g["note"].split("=")[1].split()[0]) >= 1000
               for g in pairs)  # wide wavefront, past the shared-memory ring and the prune length
    assert any(float.fromhex(g["dist"]) == -1.0 for g in golden["anreads"])
    assert any(float.fromhex(g["dist"]) == 0.0 for g in golden["anreads"])
    # heuristic: at least one score differs from the exact edit distance lower bound check below
    assert len(golden["regions"]) >= 5 and any(len(g["reads"]) == 1 for g in golden["regions"])


def test_heuristic_score_is_not_always_exact():
    """Finding 1 of the survey: scores are WF-adaptive heuristic scores (>= exact)."""
    P = S.port()
    rng = np.random.default_rng(3)

    def exact(a, b):
        prev = np.arange(len(b) + 1)
        for i in range(1, len(a) + 1):
            cur = np.empty_like(prev)
            cur[0] = i
            sub = prev[:-1] + (b != a[i - 1])
            cur[1:] = np.minimum(sub, prev[1:] + 1)
            for j in range(1, len(b) + 1):  # insertion chain
                if cur[j - 1] + 1 < cur[j]:
                    cur[j] = cur[j - 1] + 1
            prev = cur
        return int(prev[-1])

    ge, gt = 0, 0
    for _ in range(6):
        reads = S.tr_region(rng, 2, 700, 0.1, extra_frac=0.5, n_rate=0)
        x, y = sorted(reads, key=len, reverse=True)
        h = P.wfo_edit_score(S.b(x), len(x), S.b(y), len(y), 0, 0, 0, 0, 0, None)
        e = exact(x, y)
        assert h >= e
        ge += 1
        gt += h > e
    assert ge == 6


def test_prune_closed_form_equals_literal_loops():
    """wf_prune in the kernel replaces two scans of wavefront_compute_edit.c:244-267 with
    closed forms (best(k) = |k - ak| does not depend on the offset); check them against the
    literal loops."""
    rng = np.random.default_rng(11)
    for _ in range(20000):
        lo = int(rng.integers(-3000, 100))
        hi = lo + int(rng.integers(0, 3000))
        ak = int(rng.integers(-4000, 1000))
        w = int(rng.integers(0, 3000))
        # literal
        lo_r = lo
        for k in range(lo, hi + 1):
            if abs(k - ak) <= w:
                break
            lo_r += 1
        hi_r = hi
        for k in range(hi, lo_r, -1):
            if abs(k - ak) <= w:
                break
            hi_r -= 1
        # closed form (otter_b200/csrc/wfa_edit.cuh: wf_prune)
        nlo = max(lo, ak - w)
        if nlo > min(hi, ak + w):
            assert lo_r == hi + 1
            continue
        nhi = min(hi, ak + w)
        if nhi <= nlo:
            nhi = nlo
        assert (nlo, nhi) == (lo_r, hi_r), (lo, hi, ak, w)


@pytest.mark.skipif(not S.have_ref(), reason="reference library only exists in the build container")
def test_port_equals_reference_random():
    P, R = S.port(), S.ref()
    rng = np.random.default_rng(99)
    n = 0
    for tl, e, xf in [(200, 0.01, 0.2), (400, 0.08, 0.5), (1500, 0.05, 0.3), (800, 0.25, 0.3),
                      (3000, 0.005, 0.2), (3000, 0.05, 0.2)]:
        reads = S.tr_region(rng, 6, tl, e, extra_frac=xf, n_rate=0.2)
        for i, j in itertools.combinations(range(len(reads)), 2):
            x, y = S.b(reads[i]), S.b(reads[j])
            if len(x) < len(y):
                x, y = y, x
            assert P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None) == \
                R.ref_edit_e2e(x, len(x), y, len(y))
            d = len(x) - len(y)
            for pb, pe in ((0, d), (d, 0), (d // 2, d // 2)):
                assert P.wfo_edit_score(x, len(x), y, len(y), 1, pb, pe, 0, 0, None) == \
                    R.ref_edit_endsfree(x, len(x), pb, pe, y, len(y), 0, 0)
            n += 4
    assert n > 300


@pytest.mark.skipif(not S.have_ref(), reason="reference library only exists in the build container")
def test_port_fill_dist_matrix_equals_reference():
    P, R = S.port(), S.ref()
    rng = np.random.default_rng(5)
    for tl, e, n in [(300, 0.01, 9), (900, 0.05, 7)]:
        reads = S.tr_region(rng, n, tl, e, extra_frac=0.4)
        bases, off, lens = S.pack_reads(reads)
        idx = np.arange(n, dtype=np.int32)
        a = np.zeros(n * (n - 1) // 2)
        b = np.zeros_like(a)
        P.wfo_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), None, S.ptr(idx), n, S.ptr(a))
        R.ref_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), None, n, S.ptr(idx), n, S.ptr(b))
        assert np.array_equal(a, b)
