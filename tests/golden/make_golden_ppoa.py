#!/usr/bin/env python
"""Generate tests/golden/ppoa_golden.json from the UNMODIFIED reference (oracle/_ref): partial-order
consensus (src/anppoa.hpp) of read sets threaded along CIGARs, as rapid_consensus builds it
(src/analignments.cpp:258-292).  CIGARs come from the reference aligner2 (gap-affine 4/6/2) in the
three forms rapid_consensus uses, plus random well-formed CIGARs that reach the corner cases of
insert_alignment (leading gaps, first-node mismatches / insertions, non-spanning reads).  Run in
the build container:   python tests/golden/make_golden_ppoa.py"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402


def ref_cigar(R, p, t, ef, pb, pe, tb, te):
    buf = C.create_string_buffer(len(p) + len(t) + 8)
    sc, st = C.c_int(), C.c_int()
    n = R.ref_cigar(0, ef, p, len(p), pb, pe, t, len(t), tb, te, C.byref(sc), C.byref(st), buf, len(buf))
    return buf.raw[:n].decode()


def consensus_cigar(R, rep, read, span):
    """The alignment rapid_consensus asks for (src/analignments.cpp:262-281)."""
    d = len(rep) - len(read)
    sl, sr = span & 1, span & 2
    if (sl and sr) or d < 0:
        if d >= 0:
            return ref_cigar(R, rep, read, 0, 0, 0, 0, 0)
        if sl:
            return ref_cigar(R, rep, read, 1, 0, 0, 0, -d)
        if sr:
            return ref_cigar(R, rep, read, 1, 0, 0, -d, 0)
        return ref_cigar(R, rep, read, 0, 0, 0, 0, 0)  # (the reference re-uses the last CIGAR here)
    if sl:
        return ref_cigar(R, rep, read, 1, 0, d, 0, 0)
    if sr:
        return ref_cigar(R, rep, read, 1, d, 0, 0, 0)
    return ref_cigar(R, rep, read, 1, d // 2, d // 2, 0, 0)


def random_cigar(rng, blen, rlen):
    """A well-formed CIGAR over (backbone blen, read rlen) with random operations."""
    ops, b, r = [], 0, 0
    while b < blen or r < rlen:
        choices = []
        if b < blen and r < rlen:
            choices += ["M"] * 6 + ["X"] * 2
        if b < blen:
            choices.append("D")
        if r < rlen:
            choices.append("I")
        op = choices[int(rng.integers(0, len(choices)))]
        ops.append(op)
        b += op in "MXD"
        r += op in "MXI"
    return "".join(ops)


def run_ref(R, seqs, cigars, span, c, t):
    bases, off, lens = S.pack_reads([np.frombuffer(s.encode(), dtype=np.uint8) for s in seqs])
    cg = "".join(cigars).encode()
    coff = np.zeros(len(cigars) + 1, dtype=np.uint64)
    coff[1:] = np.cumsum([len(x) for x in cigars])
    sp = np.array(span, dtype=np.uint8)
    cap = int(lens.sum()) + 16
    out = C.create_string_buffer(cap)
    R.ref_ppoa_cigars.restype = C.c_int
    R.ref_ppoa_cigars.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_char_p, C.c_void_p, C.c_float,
                                                    C.c_float, C.c_char_p, C.c_int]
    n = R.ref_ppoa_cigars(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), len(seqs), cg, S.ptr(coff),
                          c, t, out, cap)
    return out.raw[:n].decode()


def main():
    R = S.ref()
    rng = np.random.default_rng(0x990A)
    cases = []

    def add(seqs, cigars, span):
        n = len(seqs)
        c = 1.0 if n < 4 else float(np.float32(n) * np.float32(0.4))  # :285-287
        t = 0.3
        cases.append(dict(seqs=seqs, cigars=cigars, span=span, c=c, t=t,
                          expected=run_ref(R, seqs, cigars, span, c, t)))

    # read sets of test/ppoa_test.cpp (every read incl. the first aligned to the first)
    kat = json.load(open(os.path.join(HERE, "affine_golden.json")))["ppoa_kat"]
    for k in kat:
        seqs = [k["seqs"][0]] + list(k["seqs"])
        add(seqs, [""] + [ref_cigar(R, seqs[0].encode(), s.encode(), 0, 0, 0, 0, 0) for s in seqs[1:]],
            [3] * len(seqs))
    # tandem-repeat alleles, errors from HiFi-like to noisy, partial reads
    for tl, e, n in [(120, 0.005, 5), (300, 0.02, 9), (300, 0.05, 14), (500, 0.1, 12), (700, 0.02, 20),
                     (200, 0.15, 8), (400, 0.05, 3), (250, 0.3, 6)]:
        reads = [S.b(r).decode() for r in S.tr_region(rng, n, tl, e, extra_frac=0.0, n_rate=0.05)]
        span = [3] * n
        for i in range(1, n):
            if rng.random() < 0.25 and len(reads[i]) > 60:
                cut = int(rng.integers(5, 40))
                how = int(rng.integers(0, 3))
                if how == 0:
                    reads[i], span[i] = reads[i][:-cut], 1
                elif how == 1:
                    reads[i], span[i] = reads[i][cut:], 2
                else:
                    reads[i], span[i] = reads[i][cut // 2:-cut // 2 - 1], 0
        cig = [""] + [consensus_cigar(R, reads[0].encode(), reads[i].encode(), span[i]) for i in range(1, n)]
        add(reads, cig, span)
    # random CIGARs on short reads: every branch of insert_alignment
    for _ in range(60):
        blen = int(rng.integers(1, 40))
        n = int(rng.integers(2, 9))
        seqs = [S.b(S.rand_seq(rng, blen)).decode()]
        cig, span = [""], [3]
        for i in range(1, n):
            rlen = int(rng.integers(1, 45))
            seqs.append(S.b(S.rand_seq(rng, rlen)).decode())
            cig.append(random_cigar(rng, blen, rlen))
            span.append(int(rng.integers(0, 4)))
        add(seqs, cig, span)
    path = os.path.join(HERE, "ppoa_golden.json")
    with open(path, "w") as f:
        json.dump(dict(source="oracle/_ref (unmodified reference PPOA, src/anppoa.hpp)", cases=cases), f)
    print(path, len(cases), "cases", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
