#!/usr/bin/env python
"""Generate tests/golden/edit_golden.json from the UNMODIFIED reference (oracle/_ref).

Run in the build container (needs /root/reference compiled by `make -C oracle ref`):
    python tests/golden/make_golden.py
The fixture stores the literal sequences next to what the reference returned, so it does not
depend on any RNG being reproducible elsewhere.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402


def s(a):
    return S.b(a).decode()


def main():
    R = S.ref()
    rng = np.random.default_rng(20261104)
    out = {"source": "holstegelab/otter @ 2573c48e + WFA2-lib v2.2 (52e8ac1d), "
                     "WFAlignerEdit(Score, MemoryMed), default WF-adaptive(10,50,1)",
           "pairs": [], "anreads": [], "regions": [], "reassign": []}

    # ---- raw aligner calls: end-to-end and the three ends-free forms otter uses ----
    shapes = [(120, 0.01, 0.2), (300, 0.05, 0.3), (600, 0.15, 0.5), (1500, 0.005, 0.2),
              (1500, 0.05, 0.4), (2500, 0.05, 0.6), (900, 0.3, 0.3)]
    for tl, e, xf in shapes:
        reads = S.tr_region(rng, 4, tl, e, extra_frac=xf, n_rate=0.25)
        for i in range(3):
            x, y = reads[i], reads[i + 1]
            if len(x) < len(y):
                x, y = y, x
            px, py = S.b(x), S.b(y)
            d = len(px) - len(py)
            out["pairs"].append(dict(p=s(x), t=s(y), ef=0, pb=0, pe=0, tb=0, te=0,
                                     score=R.ref_edit_e2e(px, len(px), py, len(py))))
            for pb, pe in ((0, d), (d, 0), (d // 2, d // 2)):
                out["pairs"].append(dict(p=s(x), t=s(y), ef=1, pb=pb, pe=pe, tb=0, te=0,
                                         score=R.ref_edit_endsfree(px, len(px), pb, pe, py, len(py), 0, 0)))
    # text free ends too (rapid_consensus uses them with the other aligner; the edit aligner
    # accepts them all the same)
    reads = S.tr_region(rng, 3, 500, 0.05, extra_frac=0.4)
    x, y = sorted(reads[:2], key=len)
    px, py = S.b(x), S.b(y)
    d = len(py) - len(px)
    for tb, te in ((0, d), (d, 0), (d // 2, d // 2)):
        out["pairs"].append(dict(p=s(x), t=s(y), ef=1, pb=0, pe=0, tb=tb, te=te,
                                 score=R.ref_edit_endsfree(px, len(px), 0, 0, py, len(py), tb, te)))
    # edge cases
    for a, b_ in (("N", "N"), ("ACGT", "N"), ("A", ""), ("", ""), ("ACGTACGT", "ACGT"),
                  ("ACGTNACGT", "ACGTAACGT"), ("A" * 70, "A" * 33), ("ACGT" * 40, "ACGT" * 40),
                  ("acgtacgtacgt", "ACGTACGTACGT"), ("RYKMACGTACGT", "RYKMACGTTCGT")):
        pa, pb_ = a.encode(), b_.encode()
        out["pairs"].append(dict(p=a, t=b_, ef=0, pb=0, pe=0, tb=0, te=0,
                                 score=R.ref_edit_e2e(pa, len(pa), pb_, len(pb_))))
    # a pair whose wavefront outgrows the shared-memory ring and reaches the >=1000 exact prune
    # (width >= 1000 switches the kernel to separate compute / prune / extend passes; the prune's
    # own trigger condition is not met by any input we could construct under the heuristic)
    rng2 = np.random.default_rng(1)
    motif = S.rand_seq(rng2, 2)
    fl = S.rand_seq(rng2, 100)
    a0 = np.concatenate([fl, np.tile(motif, 5000)])
    a1 = np.concatenate([fl, np.tile(motif, 7000)])
    x, y = S.mutate(rng2, a1, 0.08), S.mutate(rng2, a0, 0.08)
    px, py = S.b(x), S.b(y)
    ctr = S.Counters()
    S.port().wfo_edit_score(px, len(px), py, len(py), 0, 0, 0, 0, 0, ctr)
    out["pairs"].append(dict(p=s(x), t=s(y), ef=0, pb=0, pe=0, tb=0, te=0,
                             score=R.ref_edit_e2e(px, len(px), py, len(py)),
                             note=f"max_width={ctr.max_width} prunes={ctr.prunes}"))

    # ---- align_anreads: every branch (spanning flags x length order) ----
    reads = S.tr_region(rng, 6, 400, 0.03, extra_frac=0.5)
    for i in range(5):
        for xs in (0, 1, 2, 3):
            for ys in (3, 1):
                for x, y in ((reads[i], reads[i + 1]), (reads[i + 1], reads[i])):
                    d = R.ref_align_anreads(S.b(x), len(x), xs, S.b(y), len(y), ys)
                    out["anreads"].append(dict(x=s(x), xspan=xs, y=s(y), yspan=ys, dist=float(d).hex()))
    out["anreads"].append(dict(x=s(reads[0]), xspan=3, y=s(reads[0]), yspan=3,
                               dist=float(R.ref_align_anreads(S.b(reads[0]), len(reads[0]), 3,
                                                              S.b(reads[0]), len(reads[0]), 3)).hex()))

    # ---- fill_dist_matrix on whole regions ----
    for tl, e, n in ((300, 0.005, 8), (500, 0.05, 10), (1200, 0.05, 7), (200, 0.1, 2), (200, 0.1, 1)):
        reads = S.tr_region(rng, n, tl, e, extra_frac=0.4, n_rate=0.1)
        bases, off, lens = S.pack_reads(reads)
        idx = np.arange(n, dtype=np.int32)
        dm = np.zeros(n * (n - 1) // 2, dtype=np.float64)
        R.ref_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), None, n, S.ptr(idx), n, S.ptr(dm))
        out["regions"].append(dict(reads=[s(r) for r in reads], span=[3] * n,
                                   indeces=idx.tolist(), dist=[float(v).hex() for v in dm]))
    # a region whose valid set is a subset and contains half-spanning reads
    reads = S.tr_region(rng, 9, 400, 0.03, extra_frac=0.4)
    reads[2] = reads[2][: len(reads[2]) // 2]
    reads[5] = reads[5][len(reads[5]) // 3:]
    span = np.array([3, 3, 1, 3, 3, 2, 3, 0, 3], dtype=np.uint8)
    bases, off, lens = S.pack_reads(reads)
    idx = np.array([0, 1, 2, 3, 5, 6, 8], dtype=np.int32)
    dm = np.zeros(len(idx) * (len(idx) - 1) // 2, dtype=np.float64)
    R.ref_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(span), len(reads), S.ptr(idx),
                           len(idx), S.ptr(dm))
    out["regions"].append(dict(reads=[s(r) for r in reads], span=span.tolist(), indeces=idx.tolist(),
                               dist=[float(v).hex() for v in dm]))

    # ---- invalid_reassignment ----
    for trial in range(4):
        motif = S.rand_seq(rng, 4)
        fl, fr = S.rand_seq(rng, 60), S.rand_seq(rng, 60)
        alle = [np.concatenate([fl, np.tile(motif, c), fr]) for c in (40, 70)]
        reads, labels, span = [], [], []
        for a in (0, 1):
            for _ in range(4):
                reads.append(S.mutate(rng, alle[a], 0.01))
                labels.append(a)
                span.append(3)
        for a in (0, 1, 0, 1):  # clipped, unlabeled reads
            r = S.mutate(rng, alle[a], 0.01)
            side = int(rng.integers(0, 3))
            cut = int(rng.integers(20, 80))
            if side == 0:
                r, sp = r[:-cut], 1
            elif side == 1:
                r, sp = r[cut:], 2
            else:
                r, sp = r[cut // 2: -cut // 2], 0
            reads.append(r)
            labels.append(-1)
            span.append(sp)
        if trial == 3:  # a spanning but unlabeled read (the --haps situation)
            reads.append(S.mutate(rng, alle[1], 0.01))
            labels.append(-1)
            span.append(3)
        bases, off, lens = S.pack_reads(reads)
        lab = np.array(labels, dtype=np.int32)
        sp = np.array(span, dtype=np.uint8)
        R.ref_invalid_reassignment(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), len(reads),
                                   C.c_double(0.9), C.c_double(0.01), 2, S.ptr(lab))
        out["reassign"].append(dict(reads=[s(r) for r in reads], span=span, labels_in=labels,
                                    labels_out=lab.tolist(), min_sim=0.9, max_error=0.01,
                                    total_alleles=2))
    path = os.path.join(HERE, "edit_golden.json")
    json.dump(out, open(path, "w"))
    print(path, os.path.getsize(path), "bytes;", len(out["pairs"]), "pairs,", len(out["anreads"]),
          "anreads,", len(out["regions"]), "regions,", len(out["reassign"]), "reassign;",
          out["pairs"][-1].get("note"))


if __name__ == "__main__":
    main()
