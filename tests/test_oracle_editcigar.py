"""CPU tests of the edit-distance + CIGAR oracle (oracle/wfa_oracle_edit_cigar.c), the aligner
`otter compare` uses (src/compare.cpp:50-66,95): against the committed golden vectors produced by
the unmodified reference and, in the build container, against the reference on fresh inputs."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def egold():
    with open(os.path.join(ROOT, "tests", "golden", "editcigar_golden.json")) as f:
        return json.load(f)


def test_port_matches_golden_edit_cigars(egold):
    assert len(egold["cigars"]) >= 60
    for g in egold["cigars"]:
        got = S.port_edit_cigar(g["p"].encode(), g["t"].encode(), g["ef"], g["pb"], g["pe"], g["tb"], g["te"])
        assert got == (g["score"], g["cigar"]), (len(g["p"]), len(g["t"]), g["ef"])


def test_golden_edit_cigars_are_consistent(egold):
    """A CIGAR consumes exactly pattern and text; end-to-end, its non-match operations are the score."""
    for g in egold["cigars"]:
        c = g["cigar"]
        assert c.count("M") + c.count("X") + c.count("D") == len(g["p"])
        assert c.count("M") + c.count("X") + c.count("I") == len(g["t"])
        if not g["ef"]:
            assert len(c) - c.count("M") == g["score"]


def test_score_equals_score_only_aligner(egold):
    """WFAlignerEdit(Alignment) and WFAlignerEdit(Score) walk the same wavefronts."""
    P = S.port()
    for g in egold["cigars"]:
        p, t = g["p"].encode(), g["t"].encode()
        sc = P.wfo_edit_score(p, len(p), t, len(t), g["ef"], g["pb"], g["pe"], g["tb"], g["te"], None)
        assert sc == g["score"]


@pytest.mark.skipif(not S.have_ref(), reason="reference library only exists in the build container")
def test_port_equals_reference_random_edit_cigar():
    R = S.ref()
    rng = np.random.default_rng(4242)
    n = 0
    for it in range(40):
        L = int(rng.choice([1, 40, 400, 2500]))
        e = float(rng.choice([0.005, 0.05, 0.2]))
        reg = S.tr_region(rng, 2, L + 150, e, extra_frac=float(rng.choice([0, 0.3])), n_rate=0.2)
        x, y = S.b(reg[0]), S.b(reg[1])
        if len(x) < len(y):
            x, y = y, x
        d = len(x) - len(y)
        for f in [(0, 0, 0, 0, 0), (1, 0, d, 0, 0), (1, d // 2, d // 2, 0, 0)]:
            buf = C.create_string_buffer(len(x) + len(y) + 8)
            sc, st = C.c_int(), C.c_int()
            m = R.ref_cigar(1, f[0], x, len(x), f[1], f[2], y, len(y), f[3], f[4], C.byref(sc), C.byref(st),
                            buf, len(buf))
            assert S.port_edit_cigar(x, y, *f) == (sc.value, buf.raw[:m].decode())
            n += 1
    assert n == 120


def py_get_distances(subjs, querys):
    """get_distances (src/compare.cpp:50-66) on top of the port's aligner: the host rules in Python."""
    rows = []
    for i, s in enumerate(subjs):
        for j, q in enumerate(querys):
            if s == q or (s == "N" and q == "NDNNN") or (q == "N" and s == "NDNNN"):
                rows.append([i, j, 0, len(q)])
            elif "N" in (s, q) or "NDNNN" in (s, q):
                rows.append([i, j, len(q) - 1, len(q)])
            else:
                p, t = (s, q) if len(s) > len(q) else (q, s)
                sc, cg = S.port_edit_cigar(p.encode(), t.encode())
                rows.append([i, j, sc, len(cg)])
    return rows


def test_get_distances_rules_match_golden(egold):
    assert len(egold["get_distances"]) == 3
    for g in egold["get_distances"]:
        assert py_get_distances(g["subjs"], g["querys"]) == [[int(x) for x in r] for r in g["rows"]]
