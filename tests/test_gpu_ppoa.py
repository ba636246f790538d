"""Partial-order consensus on the device (otter_b200/csrc/ppoa.cuh, otgpu_ppoa_consensus; SURVEY.md
§8f N3; reference src/anppoa.hpp:63-369): byte-identical to the golden strings of the unmodified
reference (tests/golden/ppoa_golden.json, the known answers of test/ppoa_test.cpp among them) and,
on kb-long alleles, to the host restatement (ppoa_b200.hpp), all graphs of a batch in ONE call."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GRAPH = np.dtype([("backbone", "<u4"), ("first_member", "<u4"), ("n_members", "<u4"), ("c", "<f4"), ("t", "<f4")])
MEMBER = np.dtype([("read", "<u4"), ("flags", "<u4"), ("cigar_off", "<u8"), ("cigar_len", "<u4"), ("reserved", "<u4")])


def device_consensus(ctx, cases):
    """cases: list of dict(seqs, cigars, span, c, t) -> list of consensus strings, one call."""
    reads, graphs, members, cig = [], [], [], []
    coff = 0
    for g in cases:
        base = len(reads)
        reads += [np.frombuffer(s.encode(), dtype=np.uint8) for s in g["seqs"]]
        first = len(members)
        for i in range(1, len(g["seqs"])):
            c = g["cigars"][i]
            members.append((base + i, g["span"][i], coff, len(c), 0))
            cig.append(c)
            coff += len(c)
        graphs.append((base, first, len(g["seqs"]) - 1, g["c"], g["t"]))
    bases, off, lens = S.pack_reads(reads)
    G = np.array(graphs, dtype=GRAPH)
    M = np.array(members, dtype=MEMBER) if members else np.zeros(0, dtype=MEMBER)
    cg = np.frombuffer("".join(cig).encode() or b"\0", dtype=np.uint8)
    out_len = np.zeros(len(G), dtype=np.int32)
    cons, begin = C.c_void_p(), C.c_void_p()
    vp = C.c_void_p
    st = ctx._l.otgpu_ppoa_consensus(ctx._h, bases.ctypes.data_as(vp), C.c_uint64(bases.size), off.ctypes.data_as(vp),
                                     lens.ctypes.data_as(vp), C.c_uint32(len(lens)), G.ctypes.data_as(vp),
                                     C.c_uint32(len(G)), M.ctypes.data_as(vp), C.c_uint32(len(M)),
                                     cg.ctypes.data_as(vp), C.c_uint64(coff), out_len.ctypes.data_as(vp),
                                     C.byref(cons), C.byref(begin))
    assert st == 0, (st, ctx._l.otgpu_last_cuda_error(ctx._h))
    b = np.ctypeslib.as_array(C.cast(begin, C.POINTER(C.c_uint64)), shape=(len(G) + 1,))
    total = int(b[len(G)])
    raw = C.string_at(cons, total) if total else b""
    return [raw[int(b[i]):int(b[i]) + int(out_len[i])].decode() for i in range(len(G))]


def host_consensus(g):
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    l.othost_ppoa_cigars.restype = C.c_int
    l.othost_ppoa_cigars.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_char_p, C.c_void_p, C.c_float,
                                                       C.c_float, C.c_char_p, C.c_int]
    bases, off, lens = S.pack_reads([np.frombuffer(s.encode(), dtype=np.uint8) for s in g["seqs"]])
    cg = "".join(g["cigars"]).encode()
    coff = np.zeros(len(g["cigars"]) + 1, dtype=np.uint64)
    coff[1:] = np.cumsum([len(x) for x in g["cigars"]])
    sp = np.array(g["span"], dtype=np.uint8)
    cap = int(lens.sum()) + 16
    out = C.create_string_buffer(cap)
    n = l.othost_ppoa_cigars(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), len(g["seqs"]), cg, S.ptr(coff),
                             g["c"], g["t"], out, cap)
    return out.raw[:n].decode()


def test_device_ppoa_golden_in_one_call(ctx):
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ppoa_golden.json")))["cases"]
    got = device_consensus(ctx, gold)
    bad = [i for i, (a, g) in enumerate(zip(got, gold)) if a != g["expected"]]
    assert not bad, (bad[:5], got[bad[0]][:80], gold[bad[0]]["expected"][:80])
    assert sum(a != g["seqs"][0] for a, g in zip(got, gold)) > 10   # not just the backbones


def test_device_ppoa_known_answers(ctx):
    """test/ppoa_test.cpp's expected strings."""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ppoa_golden.json")))["cases"]
    kat = json.load(open(os.path.join(ROOT, "tests", "golden", "affine_golden.json")))["ppoa_kat"]
    got = device_consensus(ctx, gold[:len(kat)])
    for k, a in zip(kat, got):
        assert a == k["expected"]


def test_device_ppoa_live_alleles_equal_host(ctx, built):
    """kb-long alleles with clipped members (both spanning flags off on some), CIGARs from the
    device's own gap-affine aligner; 14 graphs in one call; and degenerate graphs."""
    from otter_b200.capi import PAIR_DTYPE
    rng = np.random.default_rng(4242)
    cases = []
    for tl, e, n in [(1500, 0.01, 30), (1000, 0.05, 40), (2500, 0.05, 12), (3500, 0.05, 30)] * 3:
        reads = S.tr_region(rng, n, tl, e, extra_frac=0.0, n_rate=0.02)
        span = [3] * n
        for i in range(1, n, 5):
            cut = int(rng.integers(5, 60))
            reads[i], span[i] = (reads[i][:-cut], 1) if i % 2 else (reads[i][cut:], 2)
        bases, off, lens = S.pack_reads(reads)
        rows = []
        for i in range(1, n):   # rapid_consensus' forms (src/analignments.cpp:262-279)
            d = int(lens[0]) - int(lens[i])
            if span[i] == 3 or d < 0:
                rows.append((0, i, -1, 0, 0, 0) if d >= 0 else (0, i, 0, 0, 0, -d))
            elif span[i] == 1:
                rows.append((0, i, 0, d, 0, 0))
            else:
                rows.append((0, i, d, 0, 0, 0))
        sc, cigs = ctx.affine_cigars(bases, off, lens, np.array(rows, dtype=PAIR_DTYPE))
        c = float(np.float32(n * 0.4)) if n >= 4 else 1.0
        cases.append(dict(seqs=[S.b(r).decode() for r in reads], cigars=[""] + [x.decode() for x in cigs],
                          span=span, c=c, t=float(np.float32(0.3))))
    # degenerate: a backbone alone; a one-base backbone with members; a member made of insertions only
    cases.append(dict(seqs=["ACGTACGTAC"], cigars=[""], span=[3], c=1.0, t=float(np.float32(0.3))))
    cases.append(dict(seqs=["A", "AC", "AC", "AC"], cigars=["", "MI", "MI", "MI"], span=[3] * 4, c=1.0,
                      t=float(np.float32(0.3))))
    cases.append(dict(seqs=["ACGTACGTACGTACG", "TTT", "ACGTACGTACGTACG", "ACGTACGTACGTACG"],
                      cigars=["", "III" + "D" * 15, "M" * 15, "M" * 15], span=[3] * 4, c=1.0,
                      t=float(np.float32(0.3))))
    got = device_consensus(ctx, cases)
    for i, (a, g) in enumerate(zip(got, cases)):
        assert a == host_consensus(g), (i, len(g["seqs"]), a[:60])
    assert sum(len(a) > 500 for a in got) >= 12
