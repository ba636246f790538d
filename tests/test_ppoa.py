"""Partial-order consensus (otter_b200/host/ppoa_b200.hpp) against the
reference's PPOA (src/anppoa.hpp): committed golden vectors everywhere, and live against the
unmodified reference where oracle/_ref is built.  Host code only: no device needed."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    l.othost_ppoa_cigars.restype = C.c_int
    l.othost_ppoa_cigars.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_char_p, C.c_void_p, C.c_float,
                                                       C.c_float, C.c_char_p, C.c_int]
    return l


def run(fn, seqs, cigars, span, c, t):
    bases, off, lens = S.pack_reads([np.frombuffer(s.encode(), dtype=np.uint8) for s in seqs])
    cg = "".join(cigars).encode()
    coff = np.zeros(len(cigars) + 1, dtype=np.uint64)
    coff[1:] = np.cumsum([len(x) for x in cigars])
    sp = np.array(span, dtype=np.uint8)
    cap = int(lens.sum()) + 16
    out = C.create_string_buffer(cap)
    n = fn(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), len(seqs), cg, S.ptr(coff), c, t, out, cap)
    return out.raw[:n].decode()


def test_ppoa_golden(built, host):
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ppoa_golden.json")))
    assert len(gold["cases"]) >= 70
    differ = 0
    for g in gold["cases"]:
        got = run(host.othost_ppoa_cigars, g["seqs"], g["cigars"], g["span"], g["c"], g["t"])
        assert got == g["expected"]
        differ += got != g["seqs"][0]
    assert differ > 10  # the consensus is not just the backbone


def test_ppoa_known_answers(built, host):
    """test/ppoa_test.cpp's expected strings (CIGARs from the committed reference alignments)."""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ppoa_golden.json")))
    kat = json.load(open(os.path.join(ROOT, "tests", "golden", "affine_golden.json")))["ppoa_kat"]
    for k, g in zip(kat, gold["cases"]):
        assert g["seqs"][1:] == k["seqs"]
        assert run(host.othost_ppoa_cigars, g["seqs"], g["cigars"], g["span"], g["c"], g["t"]) == k["expected"]


@pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref is built in the build container only")
def test_ppoa_live_vs_reference(built, host):
    """Bigger graphs than the fixtures hold: kb-long reads, 30-60 per allele, reference CIGARs."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden_ppoa as G
    R = S.ref()
    R.ref_ppoa_cigars.restype = C.c_int
    R.ref_ppoa_cigars.argtypes = host.othost_ppoa_cigars.argtypes
    rng = np.random.default_rng(31337)
    for tl, e, n in [(1500, 0.01, 30), (1000, 0.05, 40), (2500, 0.05, 12)]:
        reads = [S.b(r).decode() for r in S.tr_region(rng, n, tl, e, extra_frac=0.0, n_rate=0.02)]
        span = [3] * n
        for i in range(1, n, 5):
            cut = int(rng.integers(5, 60))
            reads[i], span[i] = (reads[i][:-cut], 1) if i % 2 else (reads[i][cut:], 2)
        cig = [""] + [G.consensus_cigar(R, reads[0].encode(), reads[i].encode(), span[i]) for i in range(1, n)]
        c, t = float(np.float32(n) * np.float32(0.4)), 0.3
        a = run(R.ref_ppoa_cigars, reads, cig, span, c, t)
        b = run(host.othost_ppoa_cigars, reads, cig, span, c, t)
        assert a == b and len(a) > 0
