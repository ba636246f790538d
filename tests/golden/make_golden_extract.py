#!/usr/bin/env python
"""Generate tests/golden/extract_golden.json from the UNMODIFIED reference (oracle/_ref):
parse_alignment (src/anseqs.cpp:412-436, get_breakpoints :286-391) on synthetic BAM records —
position, CIGAR, query — against a region [start, end].  Records cover: spanning reads, reads
ending inside the region on either side, soft/hard clipped ends (the re-adjustment branch),
deletions over the region and over its borders, insertions at the borders, records that miss the
region, IUPAC/N bases.  Run in the build container:   python tests/golden/make_golden_extract.py"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402

OPS = "MIDNSHP=X"


def cigar_words(cig):
    return np.array([(l << 4) | OPS.index(o) for l, o in cig], dtype=np.uint32)


def qlen(cig):
    return sum(l for l, o in cig if o in "MIS=X")


def ref_parse(R, pos, cig, seq, start, end):
    w = cigar_words(cig)
    out = C.create_string_buffer(len(seq) + 8)
    ok, span = C.c_int(), C.c_int()
    coords = (C.c_int * 2)()
    R.ref_parse_alignment.restype = C.c_int
    R.ref_parse_alignment.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_char_p, C.c_int, C.c_int,
                                      C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p,
                                      C.c_char_p, C.c_int]
    n = R.ref_parse_alignment(pos, w.ctypes.data_as(C.c_void_p), len(w), seq.encode(), len(seq), start,
                              end, C.byref(ok), C.byref(span), coords, out, len(seq) + 8)
    return dict(ok=ok.value, span=span.value, coords=[coords[0], coords[1]], out=out.raw[:n].decode())


def random_record(rng, start, end):
    """A record somewhere around [start, end] with a random but well-formed CIGAR."""
    kind = int(rng.integers(0, 8))
    width = end - start
    pos = start - int(rng.integers(-width // 2, 400)) if kind != 7 else end + int(rng.integers(1, 50))
    pos = max(pos, 0)
    cig = []
    if rng.random() < 0.3:
        cig.append((int(rng.integers(1, 60)), "S" if rng.random() < 0.8 else "H"))
    target = int(rng.integers(50, width + 900))
    rlen = 0
    while rlen < target:
        r = rng.random()
        if r < 0.62:
            l = int(rng.integers(1, 120))
            cig.append((l, "M=X"[int(rng.integers(0, 3))] if rng.random() < 0.3 else "M"))
            rlen += l
        elif r < 0.8:
            cig.append((int(rng.integers(1, 30)), "I"))
        elif r < 0.97:
            l = int(rng.integers(1, 60 if rng.random() < 0.9 else width + 100))
            cig.append((l, "D"))
            rlen += l
        else:
            cig.append((int(rng.integers(1, 200)), "N"))
    while cig and cig[-1][1] in "DIN":
        cig.pop()
    if rng.random() < 0.3:
        cig.append((int(rng.integers(1, 60)), "S" if rng.random() < 0.8 else "H"))
    # merge nothing, keep adjacent ops as generated (BAM allows it)
    n = qlen(cig)
    alphabet = "ACGT" if rng.random() < 0.8 else "ACGTNRYKM"
    seq = "".join(alphabet[int(i)] for i in rng.integers(0, len(alphabet), n))
    return pos, cig, seq


def main():
    R = S.ref()
    rng = np.random.default_rng(0xBA4)
    cases = []
    # hand-made corner cases (region [1000, 1100])
    hand = [
        (900, [(400, "M")]), (1000, [(101, "M")]), (1001, [(99, "M")]), (950, [(100, "M")]),
        (1050, [(200, "M")]), (900, [(50, "M"), (300, "D"), (50, "M")]),      # region deleted
        (900, [(120, "M"), (60, "D"), (120, "M")]),                          # deletion over the end
        (1020, [(30, "S"), (200, "M")]), (1020, [(30, "H"), (200, "M")]),    # clipped left inside
        (900, [(150, "M"), (40, "S")]), (900, [(150, "M"), (40, "H")]),      # clipped right inside
        (1020, [(30, "S"), (10, "I"), (200, "M")]), (1020, [(30, "S"), (5, "D"), (200, "M")]),
        (2000, [(100, "M")]), (100, [(100, "M")]), (1100, [(10, "M")]), (990, [(11, "M")]),
        (990, [(10, "M")]), (1101, [(10, "M")]), (1000, [(1, "M")]), (1100, [(1, "M")]),
        (900, [(100, "M"), (20, "I"), (200, "M")]), (900, [(200, "M"), (20, "I"), (200, "M")]),
        (900, [(101, "M"), (7, "I"), (99, "M"), (9, "I"), (100, "M")]),
        (900, [(10, "S"), (90, "="), (10, "X"), (300, "M"), (10, "S")]),
        (1000, [(5, "S")]), (1000, []), (1050, [(0, "M"), (10, "M")]),
    ]
    for pos, cig in hand:
        n = qlen(cig)
        seq = "".join("ACGT"[int(i)] for i in rng.integers(0, 4, n))
        cases.append(dict(pos=pos, cigar=cig, seq=seq, start=1000, end=1100,
                          **ref_parse(R, pos, cig, seq, 1000, 1100)))
    for _ in range(260):
        start = int(rng.integers(500, 5000))
        end = start + int(rng.integers(1, 600))
        pos, cig, seq = random_record(rng, start, end)
        cases.append(dict(pos=pos, cigar=cig, seq=seq, start=start, end=end,
                          **ref_parse(R, pos, cig, seq, start, end)))
    path = os.path.join(HERE, "extract_golden.json")
    with open(path, "w") as f:
        json.dump(dict(source="oracle/_ref (unmodified reference parse_alignment, src/anseqs.cpp:412)",
                       cases=cases), f)
    import collections
    print(path, len(cases), "cases", os.path.getsize(path), "bytes",
          collections.Counter((c["ok"], c["span"]) for c in cases))


if __name__ == "__main__":
    main()
