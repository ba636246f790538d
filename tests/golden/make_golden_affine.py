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


def gen_realign(R, rng, out):
    """local_realignment against an indexed FASTA written to a temp dir."""
    with tempfile.TemporaryDirectory() as td:
        ref = S.rand_seq(rng, 3000)
        motif = S.rand_seq(rng, 3)
        start, end, flank = 1000, 1000 + 3 * 80, 100
        ref[start:end] = np.tile(motif, 80)
        fa = os.path.join(td, "ref.fa")
        with open(fa, "w") as f:
            f.write(">chr1\n")
            s = S.b(ref).decode()
            for i in range(0, len(s), 60):
                f.write(s[i:i + 60] + "\n")
        for trial in range(3):
            reads, span, cc = [], [], []
            for r in range(8):
                kind = r % 4  # 0 spanning, 1 left side open, 2 right side open, 3 short clip
                body = S.mutate(rng, ref[start - 150:end + 150].copy(), 0.02)
                if kind == 0:
                    reads.append(body); span.append(3); cc.append((150, len(body) - 150))
                elif kind == 1:
                    reads.append(body); span.append(2)
                    cc.append((140 + int(rng.integers(0, 20)), len(body)))
                elif kind == 2:
                    reads.append(body); span.append(1)
                    cc.append((0, len(body) - 140 - int(rng.integers(0, 20))))
                else:
                    reads.append(body[100:]); span.append(2); cc.append((50, len(body) - 100))
            bases, off, lens = S.pack_reads(reads)
            sp = np.array(span, dtype=np.uint8)
            ccv = np.array(cc, dtype=np.int32).reshape(-1)
            cap = int(lens.sum()) + 64
            ob_ = np.zeros(cap, dtype=np.uint8)
            ol = np.zeros(len(reads), dtype=np.uint32)
            os_ = np.zeros(len(reads), dtype=np.uint8)
            R.ref_local_realignment(fa.encode(), b"chr1", start, end, flank, C.c_double(0.9),
                                    S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), S.ptr(ccv),
                                    len(reads), S.ptr(ob_), cap, S.ptr(ol), S.ptr(os_))
            new, pos = [], 0
            for n in ol:
                new.append(ob_[pos:pos + int(n)].tobytes().decode())
                pos += int(n)
            out["realign"].append(dict(flank=flank, min_sim=0.9,
                                       # faidx_fetch_seq takes an INCLUSIVE end (src/faidx.c), so
                                       # FaidxInstance::fetch(chr, a, b) returns b - a + 1 bases
                                       ref_left=S.b(ref[start - flank:start + 1]).decode(),
                                       ref_right=S.b(ref[end:end + flank + 1]).decode(),
                                       reads=[S.b(r).decode() for r in reads], span=span,
                                       ccoords=[list(c) for c in cc], out_reads=new,
                                       out_span=os_.tolist()))


def gen_assemble(R, rng, out):
    """Whole regions through assemble_process' body (src/assemble.cpp:103-142)."""
    for n_reads, e, copies in [(12, 0.01, (60, 90)), (20, 0.03, (40, 48)), (9, 0.005, (50, 50)),
                               (30, 0.05, (100, 130)), (3, 0.01, (30, 40))]:
        motif = S.rand_seq(rng, int(rng.integers(2, 6)))
        fl, fr = S.rand_seq(rng, 80), S.rand_seq(rng, 80)
        al = [np.concatenate([fl, np.tile(motif, c), fr]) for c in copies]
        reads = [S.mutate(rng, al[int(rng.integers(0, 2))], e) for _ in range(n_reads)]
        span = [3] * n_reads
        for a in (0, 1):
            r = S.mutate(rng, al[a], e)
            cut = int(rng.integers(20, 60))
            reads.append(r[:-cut]); span.append(1)
            reads.append(r[cut:]); span.append(2)
        res = S.assemble_region(R, reads, span)
        out["assemble"].append(dict(reads=[S.b(r).decode() for r in reads], span=span, result=res))


def main():
    R = S.ref()
    rng = np.random.default_rng(20261105)
    out = {"source": "holstegelab/otter @ 2573c48e + WFA2-lib v2.2: WFAlignerGapAffine(4,6,2,"
                     "Alignment,MemoryMed); PPOA src/anppoa.hpp; pipeline src/assemble.cpp:103-142",
           "cigars": [], "ppoa_kat": [], "realign": [], "assemble": []}
    gen_cigars(R, rng, out)
    gen_realign(R, rng, out)
    gen_assemble(R, rng, out)
    path = os.path.join(HERE, "affine_golden.json")
    json.dump(out, open(path, "w"))
    chg = sum(a != b for g in out["realign"] for a, b in zip(g["reads"], g["out_reads"]))
    print(path, os.path.getsize(path), "bytes;", len(out["cigars"]), "cigars,", len(out["ppoa_kat"]),
          "ppoa KATs,", len(out["realign"]), "realign cases (", chg, "reads trimmed ),",
          len(out["assemble"]), "regions:", [g["result"]["n_alleles"] for g in out["assemble"]])


if __name__ == "__main__":
    main()
