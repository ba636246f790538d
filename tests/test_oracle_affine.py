"""CPU tests of the gap-affine oracle (oracle/wfa_oracle_affine.c): against the committed golden
vectors (CIGARs produced by the unmodified reference, incl. the CIGARs behind the reference's own
PPOA known answers, test/ppoa_test.cpp:39-105) and, in the build container, against the
reference on fresh inputs."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def agold():
    with open(os.path.join(ROOT, "tests", "golden", "affine_golden.json")) as f:
        return json.load(f)


def port_cigar(p, t, ef, pb, pe, tb, te):
    P = S.port()
    buf = C.create_string_buffer(len(p) + len(t) + 8)
    sc = C.c_int()
    n = P.wfo_affine_cigar(p, len(p), t, len(t), ef, pb, pe, tb, te, 4, 6, 2, C.byref(sc), buf, len(buf))
    return sc.value, buf.raw[:max(n, 0)].decode()


def test_port_matches_golden_cigars(agold):
    for g in agold["cigars"]:
        sc, cg = port_cigar(g["p"].encode(), g["t"].encode(), g["ef"], g["pb"], g["pe"], g["tb"], g["te"])
        assert (sc, cg) == (g["score"], g["cigar"]), (len(g["p"]), len(g["t"]), g["ef"])


def test_port_matches_reference_ppoa_kat_cigars(agold):
    """The CIGARs that feed the reference's own consensus known answers."""
    assert len(agold["ppoa_kat"]) == 4
    for kat in agold["ppoa_kat"]:
        for s, cg in zip(kat["seqs"], kat["cigars"]):
            assert port_cigar(kat["seqs"][0].encode(), s.encode(), 0, 0, 0, 0, 0)[1] == cg


def test_golden_cigars_are_consistent(agold):
    """Structural sanity of the fixture: a CIGAR consumes exactly pattern and text."""
    forms = set()
    for g in agold["cigars"]:
        c = g["cigar"]
        assert c.count("M") + c.count("X") + c.count("D") == len(g["p"])
        assert c.count("M") + c.count("X") + c.count("I") == len(g["t"])
        forms.add((g["ef"], g["pb"] > 0, g["pe"] > 0, g["tb"] > 0, g["te"] > 0))
    assert len(forms) >= 5  # end-to-end + the ends-free forms rapid_consensus issues


@pytest.mark.skipif(not S.have_ref(), reason="reference library only exists in the build container")
def test_port_equals_reference_random_affine():
    R = S.ref()
    rng = np.random.default_rng(77)
    n = 0
    for tl, e, xf in [(150, 0.03, 0.3), (500, 0.08, 0.5), (1200, 0.05, 0.3), (2500, 0.01, 0.2)]:
        reads = S.tr_region(rng, 5, tl, e, extra_frac=xf, n_rate=0.2)
        for i in range(4):
            x, y = S.b(reads[i]), S.b(reads[i + 1])
            d = len(x) - len(y)
            forms = [(0, 0, 0, 0, 0)]
            forms += [(1, 0, d, 0, 0), (1, d, 0, 0, 0), (1, d // 2, d // 2, 0, 0)] if d >= 0 else \
                [(1, 0, 0, 0, -d), (1, 0, 0, -d, 0)]
            for f in forms:
                buf = C.create_string_buffer(len(x) + len(y) + 8)
                sc, st = C.c_int(), C.c_int()
                k = R.ref_cigar(0, f[0], x, len(x), f[1], f[2], y, len(y), f[3], f[4], C.byref(sc),
                                C.byref(st), buf, len(buf))
                assert port_cigar(x, y, *f) == (sc.value, buf.raw[:k].decode())
                n += 1
    assert n > 50


@pytest.mark.skipif(not S.have_ref(), reason="reference library only exists in the build container")
def test_reference_reproduces_its_ppoa_known_answers(agold):
    """Pins the harness: reference aligner2 + PPOA through our shim gives test/ppoa_test.cpp's
    expected consensus strings."""
    R = S.ref()
    for kat in agold["ppoa_kat"]:
        reads = [np.frombuffer(s.encode(), dtype=np.uint8) for s in kat["seqs"]]
        bases, off, lens = S.pack_reads(reads)
        buf = C.create_string_buffer(1000)
        n = R.ref_ppoa_consensus(S.ptr(bases), S.ptr(off), S.ptr(lens), len(reads), buf, 1000)
        assert buf.raw[:n].decode() == kat["expected"]


def test_golden_local_realignment_replayed_with_port(agold):
    """local_realignment (src/analignments.cpp:11-60) replayed in Python on the port's CIGARs
    reproduces what the reference did to the reads — pins the flank strings of the fixture
    (faidx fetch is end-inclusive: flank + 1 bases)."""
    for g in agold["realign"]:
        assert len(g["ref_left"]) == g["flank"] + 1 and len(g["ref_right"]) == g["flank"] + 1
        for seq, sp, (c0, c1), exp, exp_sp in zip(g["reads"], g["span"], g["ccoords"], g["out_reads"],
                                                  g["out_span"]):
            out, out_sp = seq, sp
            if sp in (1, 2):
                left = sp == 2 and c0 >= g["flank"]
                right = sp == 1 and len(seq) - c1 >= g["flank"]
                sub = seq[:c0] if left else (seq[c1:] if right else "")
                if sub:
                    cg = port_cigar(sub.encode(), (g["ref_left"] if left else g["ref_right"]).encode(),
                                    0, 0, 0, 0, 0)[1]
                    scores, j = [0] * len(sub), 0
                    for op in cg:
                        if op == "I":
                            continue
                        if op == "M":
                            scores[j] = 1 if j == 0 else scores[j - 1] + 1
                        elif j > 0 and scores[j - 1] > 0:
                            scores[j] = scores[j - 1] - 1
                        j += 1
                    best = max(range(len(scores)), key=lambda i: (scores[i], -i))
                    start_i = best
                    while start_i > 0 and scores[start_i] > 0:
                        start_i -= 1
                    if scores[best] / float(g["flank"]) >= g["min_sim"]:
                        out = seq[best:] if left else seq[:c1 + start_i]
                        out_sp = 3
            assert (out, out_sp) == (exp, exp_sp)
