#!/usr/bin/env python
"""Generate tests/golden/affine_golden.json from the UNMODIFIED reference (oracle/_ref):
gap-affine (4,6,2) CIGARs, the reference's own partial-order-consensus known answers
(test/ppoa_test.cpp:39-105), local_realignment outcomes and whole-region assemble results
(labels + alleles).  Run in the build container:  python tests/golden/make_golden_affine.py"""
import ctypes as C
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import support as S  # noqa: E402
from ppoa_kat import PPOA_KAT  # noqa: E402


def u8(s):
    return np.frombuffer(s.encode(), dtype=np.uint8)


def ref_cigar(R, p, t, ef, pb, pe, tb, te):
    buf = C.create_string_buffer(len(p) + len(t) + 8)
    sc, st = C.c_int(), C.c_int()
    n = R.ref_cigar(0, ef, p, len(p), pb, pe, t, len(t), tb, te, C.byref(sc), C.byref(st), buf, len(buf))
    return sc.value, buf.raw[:n].decode()


def gen_cigars(R, rng, out):
    # the reference's own known answers (consensus) + the CIGARs behind them
    for seqs, expected in PPOA_KAT:
        reads = [u8(s) for s in seqs]
        bases, off, lens = S.pack_reads(reads)
        buf = C.create_string_buffer(1000)
        n = R.ref_ppoa_consensus(S.ptr(bases), S.ptr(off), S.ptr(lens), len(seqs), buf, 1000)
        got = buf.raw[:n].decode()
        assert got == expected, (got, expected)  # the reference reproduces its own KAT here
        cigs = [ref_cigar(R, seqs[0].encode(), s.encode(), 0, 0, 0, 0, 0)[1] for s in seqs]
        out["ppoa_kat"].append(dict(seqs=seqs, expected=expected, cigars=cigs))
    # raw aligner2 calls: end-to-end and every ends-free form rapid_consensus issues
    for tl, e, xf in [(120, 0.02, 0.2), (300, 0.05, 0.4), (300, 0.15, 0.5), (1500, 0.005, 0.2),
                      (1500, 0.05, 0.3), (3000, 0.05, 0.2), (800, 0.3, 0.3)]:
        reads = S.tr_region(rng, 4, tl, e, extra_frac=xf, n_rate=0.2)
        for i in range(3):
            x, y = S.b(reads[i]), S.b(reads[i + 1])
            d = len(x) - len(y)
            forms = [(0, 0, 0, 0, 0)]
            forms += [(1, 0, d, 0, 0), (1, d, 0, 0, 0), (1, d // 2, d // 2, 0, 0)] if d >= 0 else \
                [(1, 0, 0, 0, -d), (1, 0, 0, -d, 0)]
            for f in forms:
                sc, cg = ref_cigar(R, x, y, *f)
                out["cigars"].append(dict(p=x.decode(), t=y.decode(), ef=f[0], pb=f[1], pe=f[2],
                                          tb=f[3], te=f[4], score=sc, cigar=cg))
    for a, b_ in (("ACGT", "ACGT"), ("A", ""), ("", "A"), ("ACGTNACGT", "ACGTACGT"), ("AAAA", "TTTT"),
                  ("ACGTACGTACGT", "ACGTTTTTACGTACGT")):
        sc, cg = ref_cigar(R, a.encode(), b_.encode(), 0, 0, 0, 0, 0)
        out["cigars"].append(dict(p=a, t=b_, ef=0, pb=0, pe=0, tb=0, te=0, score=sc, cigar=cg))
