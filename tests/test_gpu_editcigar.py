"""GPU parity tests of edit distance WITH CIGAR — `otter compare`'s aligner and get_distances
(reference src/compare.cpp:50-66,95): C ABI, WFAligner facade and the
reference-shaped get_distances, against the golden vectors of the unmodified reference, the C
port on fresh inputs and (where the reference library travelled) the reference itself."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def egold():
    with open(os.path.join(ROOT, "tests", "golden", "editcigar_golden.json")) as f:
        return json.load(f)


def u8(s):
    return np.frombuffer(s.encode() if isinstance(s, str) else s, dtype=np.uint8)


def PAIRS(rows):
    from otter_b200.capi import PAIR_DTYPE
    return np.array(rows, dtype=PAIR_DTYPE)


def host_lib():
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    l.othost_edit_cigar.restype = C.c_int
    l.othost_edit_cigar.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                    C.c_char_p, C.c_int]
    l.othost_get_distances.restype = None
    l.othost_get_distances.argtypes = [vp, vp, vp, C.c_uint32, C.c_uint32, vp]
    return l


def test_golden_edit_cigars_through_c_abi(ctx, egold):
    reads, rows = [], []
    for g in egold["cigars"]:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, g["pb"] if g["ef"] else -1, g["pe"], g["tb"], g["te"]))
    bases, off, lens = S.pack_reads(reads)
    scores, cigs = ctx.edit_cigars(bases, off, lens, PAIRS(rows))
    for g, sc, cg in zip(egold["cigars"], scores, cigs):
        assert isinstance(cg, bytes), cg
        assert (int(sc), cg.decode()) == (g["score"], g["cigar"]), (len(g["p"]), len(g["t"]), g["ef"])


def test_golden_edit_cigars_through_facade(built, egold):
    H = host_lib()
    for g in egold["cigars"][::2]:
        p, t = g["p"].encode(), g["t"].encode()
        buf = C.create_string_buffer(len(p) + len(t) + 8)
        sc, st = C.c_int(), C.c_int()
        n = H.othost_edit_cigar(p, len(p), t, len(t), g["ef"], g["pb"], g["pe"], g["tb"], g["te"],
                                C.byref(sc), C.byref(st), buf, len(buf))
        assert st.value == 0
        assert (sc.value, buf.raw[:n].decode()) == (g["score"], g["cigar"])


def test_edit_cigars_vs_port_random(ctx):
    """Fresh seeded pairs (HiFi-like, ONT-like, divergent, N-carrying, long): CIGAR bytes and
    scores equal the C port's; the score also equals the score-only kernel's."""
    rng = np.random.default_rng(20261104)
    reads, rows = [], []
    for L, e, xf in [(60, 0.05, 0.3), (500, 0.005, 0.2), (500, 0.15, 0.5), (3000, 0.05, 0.2),
                     (3000, 0.005, 0.3), (9000, 0.05, 0.1), (15000, 0.02, 0.05)]:
        reg = S.tr_region(rng, 5, L, e, extra_frac=xf, n_rate=0.3)
        base = len(reads)
        reads += reg
        for i in range(4):
            a, b = base + i, base + i + 1
            if len(reads[a]) < len(reads[b]):
                a, b = b, a
            d = len(reads[a]) - len(reads[b])
            rows.append((a, b, -1, 0, 0, 0))
            if i % 2 == 0:
                rows.append((a, b, d // 2, d - d // 2, 0, 0))
    bases, off, lens = S.pack_reads(reads)
    pairs = PAIRS(rows)
    scores, cigs = ctx.edit_cigars(bases, off, lens, pairs)
    so = ctx.edit_scores(bases, off, lens, pairs)
    sc_only, len_only = ctx.edit_cigar_lengths(bases, off, lens, pairs)  # no CIGAR bytes requested
    assert np.array_equal(sc_only, scores)
    assert len_only.tolist() == [len(c) for c in cigs]
    for r, sc, cg, s2 in zip(rows, scores, cigs, so):
        p, t = S.b(reads[r[0]]), S.b(reads[r[1]])
        ef = r[2] >= 0
        exp = S.port_edit_cigar(p, t, int(ef), max(r[2], 0), r[3], r[4], r[5])
        assert isinstance(cg, bytes), cg
        assert (int(sc), cg.decode()) == exp, (len(p), len(t), r)
        assert int(s2) == exp[0]


def test_cigar_roundtrip_property(ctx):
    """Size-independent property at read lengths of C3: replaying the CIGAR over the pattern
    reproduces the text, and (end-to-end) the non-match operations number the score."""
    rng = np.random.default_rng(77)
    reg = S.tr_region(rng, 6, 16000, 0.05, flank=500, extra_frac=0.1, n_rate=0.0)
    rows = []
    for i in range(5):
        a, b = (i, i + 1) if len(reg[i]) >= len(reg[i + 1]) else (i + 1, i)
        rows.append((a, b, -1, 0, 0, 0))
    bases, off, lens = S.pack_reads(reg)
    scores, cigs = ctx.edit_cigars(bases, off, lens, PAIRS(rows))
    for r, sc, cg in zip(rows, scores, cigs):
        p, t = S.b(reg[r[0]]), S.b(reg[r[1]])
        ops = np.frombuffer(cg, dtype=np.uint8)
        pc = np.cumsum((ops != ord("I"))) - (ops != ord("I"))  # pattern position of each op
        keep = (ops == ord("M")) | (ops == ord("X"))
        tc = np.cumsum((ops != ord("D"))) - (ops != ord("D"))
        pa, ta = np.frombuffer(p, np.uint8), np.frombuffer(t, np.uint8)
        assert (ops != ord("I")).sum() == len(p) and (ops != ord("D")).sum() == len(t)
        m = ops == ord("M")
        assert np.array_equal(pa[pc[m]], ta[tc[m]])          # every M is a real match
        x = ops == ord("X")
        assert not np.any(pa[pc[x]] == ta[tc[x]])            # every X a real mismatch
        assert int(sc) == int((~m).sum()) and keep.sum() <= min(len(p), len(t))


def test_get_distances_matches_golden(built, egold):
    H = host_lib()
    for g in egold["get_distances"]:
        seqs = [u8(x) for x in g["subjs"] + g["querys"]]
        bases, off, lens = S.pack_reads(seqs)
        rows = np.zeros((len(g["subjs"]) * len(g["querys"]), 4), dtype=np.float64)
        H.othost_get_distances(S.ptr(bases), S.ptr(off), S.ptr(lens), len(g["subjs"]), len(g["querys"]),
                               S.ptr(rows))
        assert rows.tolist() == g["rows"]


@pytest.mark.skipif(not S.have_ref(), reason="reference library is prebuilt in the build container")
def test_get_distances_equals_reference_random(built):
    """C5-shaped allele lists (two subjects x a cohort of query alleles) through the reference's own
    get_distances and ours."""
    H, R = host_lib(), S.ref()
    R.ref_get_distances.restype = None
    R.ref_get_distances.argtypes = [C.c_void_p] * 3 + [C.c_uint32, C.c_uint32, C.c_void_p]
    rng = np.random.default_rng(5005)
    for tl, e, nq in [(300, 0.01, 40), (2500, 0.005, 24), (1200, 0.05, 16)]:
        reads = S.tr_region(rng, nq + 2, tl, e, extra_frac=0.3, n_rate=0.05)
        reads[5] = u8("N")
        reads[6] = reads[0].copy()
        bases, off, lens = S.pack_reads(reads)
        a = np.zeros((2 * nq, 4), dtype=np.float64)
        b = np.zeros_like(a)
        R.ref_get_distances(S.ptr(bases), S.ptr(off), S.ptr(lens), 2, nq, S.ptr(a))
        H.othost_get_distances(S.ptr(bases), S.ptr(off), S.ptr(lens), 2, nq, S.ptr(b))
        assert np.array_equal(a, b)
