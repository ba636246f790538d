#!/usr/bin/env python
"""Generate tests/golden/editcigar_golden.json from the UNMODIFIED reference (oracle/_ref):
edit-distance alignments WITH CIGAR as `otter compare` runs them
(wfa::WFAlignerEdit(Alignment, MemoryMed), src/compare.cpp:95; get_distances :50-66 uses the
score and the CIGAR length), end-to-end and ends-free, incl. N/empty/equal reads and the
get_distances special cases.  Run in the build container:
    python tests/golden/make_golden_editcigar.py"""
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
    n = R.ref_cigar(1, ef, p, len(p), pb, pe, t, len(t), tb, te, C.byref(sc), C.byref(st), buf, len(buf))
    return sc.value, buf.raw[:n].decode()


def main():
    R = S.ref()
    rng = np.random.default_rng(0xED17)
    out = dict(source="oracle/_ref (unmodified reference), WFAlignerEdit(Alignment, MemoryMed)", cigars=[])

    def add(x, y, f):
        sc, cg = ref_cigar(R, x, y, *f)
        out["cigars"].append(dict(p=x.decode(), t=y.decode(), ef=f[0], pb=f[1], pe=f[2], tb=f[3],
                                  te=f[4], score=sc, cigar=cg))

    for tl, e, xf in [(60, 0.05, 0.3), (300, 0.02, 0.4), (300, 0.15, 0.5), (1500, 0.005, 0.2),
                      (1500, 0.05, 0.3), (3000, 0.05, 0.2), (800, 0.3, 0.3), (5000, 0.05, 0.2)]:
        reads = S.tr_region(rng, 4, tl, e, extra_frac=xf, n_rate=0.2)
        for i in range(3):
            x, y = S.b(reads[i]), S.b(reads[i + 1])
            if len(x) < len(y):  # get_distances aligns (longer, shorter), src/compare.cpp:60
                x, y = y, x
            d = len(x) - len(y)
            add(x, y, (0, 0, 0, 0, 0))
            if i == 0:
                for f in [(1, 0, d, 0, 0), (1, d, 0, 0, 0), (1, d // 2, d // 2, 0, 0),
                          (1, min(3, len(x)), min(2, len(x)), min(4, len(y)), min(1, len(y)))]:
                    add(x, y, f)
    for a, b_ in (("ACGT", "ACGT"), ("A", ""), ("", "A"), ("", ""), ("ACGTNACGT", "ACGTACGT"),
                  ("AAAA", "TTTT"), ("ACGTACGTACGT", "ACGTTTTTACGTACGT"), ("N", "ACGT"),
                  ("NDNNN", "ACGTAC"), ("acgtACGT", "ACGTACGT")):
        add(a.encode(), b_.encode(), (0, 0, 0, 0, 0))
    # whole get_distances calls (src/compare.cpp:50-66): subjects x queries -> {i, j, edit, ops}
    out["get_distances"] = []
    R.ref_get_distances.restype = None
    R.ref_get_distances.argtypes = [C.c_void_p] * 3 + [C.c_uint32, C.c_uint32, C.c_void_p]
    for tl, e in [(200, 0.02), (900, 0.05), (2500, 0.01)]:
        reads = [S.b(r).decode() for r in S.tr_region(rng, 6, tl, e, extra_frac=0.3, n_rate=0.0)]
        subj = [reads[0], reads[1]]
        query = [reads[2], reads[3], reads[0], "N", "NDNNN", reads[4][: len(reads[0])]]
        if tl == 900:
            subj[1] = "N"
        if tl == 2500:
            subj[0] = "NDNNN"
        seqs = [np.frombuffer(x.encode(), dtype=np.uint8) for x in subj + query]
        bases, off, lens = S.pack_reads(seqs)
        rows = np.zeros((len(subj) * len(query), 4), dtype=np.float64)
        R.ref_get_distances(S.ptr(bases), S.ptr(off), S.ptr(lens), len(subj), len(query), S.ptr(rows))
        out["get_distances"].append(dict(subjs=subj, querys=query, rows=rows.tolist()))
    path = os.path.join(HERE, "editcigar_golden.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print(path, len(out["cigars"]), "alignments", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
