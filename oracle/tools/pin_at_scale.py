#!/usr/bin/env python
"""oracle/tools/pin_at_scale.py — TEST INFRASTRUCTURE ONLY.

The C restatements (oracle/wfa_oracle*.c) against the UNMODIFIED reference (oracle/_ref) on many
more random inputs than the unit tests take: the same loops as tests/test_oracle.py
(test_port_equals_reference_random), tests/test_oracle_affine.py and tests/test_oracle_editcigar.py,
over <seeds> seeds each.  Needs oracle/_ref (built where /root/reference exists).

    python oracle/tools/pin_at_scale.py [seeds=40] [threads]
"""
import ctypes as C
import itertools
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import support as S  # noqa: E402
from test_oracle_affine import port_cigar  # noqa: E402

P, R = S.port(), S.ref()


def edit_scores(seed):
    rng = np.random.default_rng(seed)
    n = bad = 0
    for tl, e, xf in [(200, 0.01, 0.2), (400, 0.08, 0.5), (1500, 0.05, 0.3), (800, 0.25, 0.3),
                      (3000, 0.005, 0.2), (3000, 0.05, 0.2), (6000, 0.1, 0.4)]:
        reads = S.tr_region(rng, 6, tl, e, extra_frac=xf, n_rate=0.2)
        for i, j in itertools.combinations(range(len(reads)), 2):
            x, y = S.b(reads[i]), S.b(reads[j])
            if len(x) < len(y):
                x, y = y, x
            bad += P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None) != R.ref_edit_e2e(x, len(x), y, len(y))
            d = len(x) - len(y)
            for pb, pe in ((0, d), (d, 0), (d // 2, d // 2)):
                bad += P.wfo_edit_score(x, len(x), y, len(y), 1, pb, pe, 0, 0, None) != \
                    R.ref_edit_endsfree(x, len(x), pb, pe, y, len(y), 0, 0)
            n += 4
    return n, bad


def cigars(seed, edit):
    rng = np.random.default_rng(seed)
    n = bad = 0
    for tl, e, xf in [(150, 0.03, 0.3), (500, 0.08, 0.5), (1200, 0.05, 0.3), (2500, 0.01, 0.2), (4000, 0.12, 0.3)]:
        reads = S.tr_region(rng, 5, tl, e, extra_frac=xf, n_rate=0.2)
        for i in range(4):
            x, y = S.b(reads[i]), S.b(reads[i + 1])
            if edit and len(x) < len(y):
                x, y = y, x
            d = len(x) - len(y)
            forms = [(0, 0, 0, 0, 0)]
            forms += [(1, 0, d, 0, 0), (1, d, 0, 0, 0), (1, d // 2, d // 2, 0, 0)] if d >= 0 else \
                [(1, 0, 0, 0, -d), (1, 0, 0, -d, 0)]
            for f in forms:
                buf = C.create_string_buffer(len(x) + len(y) + 8)
                sc, st = C.c_int(), C.c_int()
                k = R.ref_cigar(1 if edit else 0, f[0], x, len(x), f[1], f[2], y, len(y), f[3], f[4], C.byref(sc),
                                C.byref(st), buf, len(buf))
                got = S.port_edit_cigar(x, y, *f) if edit else port_cigar(x, y, *f)
                bad += got != (sc.value, buf.raw[:k].decode())
                n += 1
    return n, bad


def main():
    seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    threads = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    with ThreadPoolExecutor(threads) as ex:
        for name, fn in (("edit distance scores (end-to-end + three ends-free forms)", edit_scores),
                         ("gap-affine (4,6,2) scores + CIGARs", lambda s: cigars(s, False)),
                         ("edit distance scores + CIGARs", lambda s: cigars(s, True))):
            res = list(ex.map(fn, range(5000, 5000 + seeds)))
            print(f"{name}: {sum(r[0] for r in res)} alignments over {seeds} seeds, "
                  f"{sum(r[1] for r in res)} differ from the unmodified reference", flush=True)


if __name__ == "__main__":
    main()
