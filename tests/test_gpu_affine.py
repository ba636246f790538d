"""GPU parity tests of the gap-affine + CIGAR path:
C ABI, WFAligner facade, local_realignment, and whole regions through the reference's own
pipeline (otterclust + hclust-cpp + anppoa) running on the GPU hot path — labels and alleles
must equal the unmodified reference's."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def agold():
    with open(os.path.join(ROOT, "tests", "golden", "affine_golden.json")) as f:
        return json.load(f)


def u8(s):
    return np.frombuffer(s.encode() if isinstance(s, str) else s, dtype=np.uint8)


def PAIRS(rows):
    from otter_b200.capi import PAIR_DTYPE
    return np.array(rows, dtype=PAIR_DTYPE)


def host_lib():
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    l.othost_affine_cigar.restype = C.c_int
    l.othost_affine_cigar.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                      C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                      C.c_char_p, C.c_int]
    l.othost_local_realignment.restype = C.c_uint64
    l.othost_local_realignment.argtypes = [C.c_int, C.c_double, C.c_char_p, C.c_uint32, C.c_char_p,
                                           C.c_uint32, vp, vp, vp, vp, vp, C.c_uint32, vp, C.c_uint64,
                                           vp, vp]
    return l


def test_golden_cigars_through_c_abi(ctx, agold):
    reads, rows = [], []
    for g in agold["cigars"]:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, g["pb"] if g["ef"] else -1, g["pe"], g["tb"], g["te"]))
    bases, off, lens = S.pack_reads(reads)
    scores, cigs = ctx.affine_cigars(bases, off, lens, PAIRS(rows))
    for g, sc, cg in zip(agold["cigars"], scores, cigs):
        assert isinstance(cg, bytes), cg
        assert (int(sc), cg.decode()) == (g["score"], g["cigar"]), (len(g["p"]), len(g["t"]), g["ef"])


def test_packed_cigar_output_equals_capacity_layout(ctx, agold):
    """otgpu_*_cigars_packed: the same scores, lengths and bytes, end to end in pinned memory."""
    import json
    import os
    reads, rows = [], []
    for g in agold["cigars"]:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, g["pb"] if g["ef"] else -1, g["pe"], g["tb"], g["te"]))
    bases, off, lens = S.pack_reads(reads)
    sc, ln, begin, packed = ctx.cigars_packed(bases, off, lens, PAIRS(rows))
    assert int(begin[len(rows)]) == int(np.maximum(ln, 0).sum()) == len(packed)
    for i, g in enumerate(agold["cigars"]):
        got = packed[int(begin[i]):int(begin[i]) + int(ln[i])].tobytes().decode()
        assert (int(sc[i]), got) == (g["score"], g["cigar"])
    eg = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "editcigar_golden.json")))["cigars"]
    reads, rows = [], []
    for g in eg:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, g["pb"] if g["ef"] else -1, g["pe"], g["tb"], g["te"]))
    bases, off, lens = S.pack_reads(reads)
    sc, ln, begin, packed = ctx.cigars_packed(bases, off, lens, PAIRS(rows), edit=True)
    for i, g in enumerate(eg):
        got = packed[int(begin[i]):int(begin[i]) + int(ln[i])].tobytes().decode()
        assert (int(sc[i]), got) == (g["score"], g["cigar"])
    # empty batch
    sc, ln, begin, packed = ctx.cigars_packed(bases, off, lens, PAIRS([]))
    assert len(sc) == 0 and len(packed) == 0


def test_ppoa_kat_cigars_through_c_abi(ctx, agold):
    for kat in agold["ppoa_kat"]:
        reads = [u8(s) for s in kat["seqs"]]
        bases, off, lens = S.pack_reads(reads)
        rows = [(0, i, -1, 0, 0, 0) for i in range(len(reads))]
        _, cigs = ctx.affine_cigars(bases, off, lens, PAIRS(rows))
        assert [c.decode() for c in cigs] == kat["cigars"]


def test_golden_cigars_through_facade(built, agold):
    H = host_lib()
    for g in agold["cigars"][::3]:
        p, t = g["p"].encode(), g["t"].encode()
        buf = C.create_string_buffer(len(p) + len(t) + 8)
        sc, st = C.c_int(), C.c_int()
        n = H.othost_affine_cigar(p, len(p), t, len(t), g["ef"], g["pb"], g["pe"], g["tb"], g["te"],
                                  C.byref(sc), C.byref(st), buf, len(buf))
        assert st.value == 0
        assert (sc.value, buf.raw[:n].decode()) == (g["score"], g["cigar"])


def test_affine_vs_port_random(ctx):
    P = S.port()
    rng = np.random.default_rng(31)
    reads, rows = [], []
    for tl, e, xf in [(200, 0.03, 0.3), (700, 0.1, 0.5), (2000, 0.05, 0.3), (5000, 0.02, 0.2),
                      (12000, 0.05, 0.2)]:
        rr = S.tr_region(rng, 4, tl, e, extra_frac=xf, n_rate=0.25)
        base = len(reads)
        reads += rr
        for i in range(3):
            a, b = base + i, base + i + 1
            d = len(reads[a]) - len(reads[b])
            rows.append((a, b, -1, 0, 0, 0))
            rows += [(a, b, 0, d, 0, 0), (a, b, d, 0, 0, 0), (a, b, d // 2, d // 2, 0, 0)] if d >= 0 else \
                [(a, b, 0, 0, 0, -d), (a, b, 0, 0, -d, 0)]
    bases, off, lens = S.pack_reads(reads)
    pairs = PAIRS(rows)
    scores, cigs = ctx.affine_cigars(bases, off, lens, pairs)
    for i, (p, t, pb, pe, tb, te) in enumerate(pairs.tolist()):
        x, y = S.b(reads[p]), S.b(reads[t])
        buf = C.create_string_buffer(len(x) + len(y) + 8)
        sc = C.c_int()
        n = P.wfo_affine_cigar(x, len(x), y, len(y), int(pb >= 0), max(pb, 0), pe, tb, te, 4, 6, 2,
                               C.byref(sc), buf, len(buf))
        assert (int(scores[i]), cigs[i]) == (sc.value, buf.raw[:n]), (i, len(x), len(y), pb, pe, tb, te)
    # other penalties go through the same kernel
    s2, c2 = ctx.affine_cigars(bases, off, lens, pairs[:8], x=2, o=3, e=1)
    for i, (p, t, pb, pe, tb, te) in enumerate(pairs[:8].tolist()):
        x, y = S.b(reads[p]), S.b(reads[t])
        buf = C.create_string_buffer(len(x) + len(y) + 8)
        sc = C.c_int()
        n = P.wfo_affine_cigar(x, len(x), y, len(y), int(pb >= 0), max(pb, 0), pe, tb, te, 2, 3, 1,
                               C.byref(sc), buf, len(buf))
        assert (int(s2[i]), c2[i]) == (sc.value, buf.raw[:n])


def test_golden_local_realignment(built, agold):
    H = host_lib()
    trimmed = 0
    for g in agold["realign"]:
        reads = [u8(r) for r in g["reads"]]
        bases, off, lens = S.pack_reads(reads)
        sp = np.array(g["span"], dtype=np.uint8)
        cc = np.array(g["ccoords"], dtype=np.int32).reshape(-1)
        cap = int(lens.sum()) + 64
        ob = np.zeros(cap, dtype=np.uint8)
        ol = np.zeros(len(reads), dtype=np.uint32)
        os_ = np.zeros(len(reads), dtype=np.uint8)
        L, R = g["ref_left"].encode(), g["ref_right"].encode()
        H.othost_local_realignment(g["flank"], g["min_sim"], L, len(L), R, len(R), S.ptr(bases), S.ptr(off),
                                   S.ptr(lens), S.ptr(sp), S.ptr(cc), len(reads), S.ptr(ob), cap,
                                   S.ptr(ol), S.ptr(os_))
        new, pos = [], 0
        for n in ol:
            new.append(ob[pos:pos + int(n)].tobytes().decode())
            pos += int(n)
        assert new == g["out_reads"]
        assert os_.tolist() == g["out_span"]
        trimmed += sum(a != b for a, b in zip(g["reads"], new))
    assert trimmed > 0


@pytest.mark.skipif(not S.have_intree(), reason="integration/_build/libotter_intree.so is prebuilt in the build container")
def test_whole_regions_equal_reference_golden(built, agold):
    """fill_dist_matrix -> otter_hclust -> invalid_reassignment -> rapid_consensus with the
    reference's own clustering/consensus code on top of the GPU hot path: labels, alleles,
    coverage and SE identical to what the unmodified reference produced."""
    T = S.intree()
    for g in agold["assemble"]:
        res = S.assemble_region(T, [u8(r) for r in g["reads"]], g["span"])
        assert res == g["result"]


@pytest.mark.skipif(not S.have_intree(), reason="integration/_build/libotter_intree.so is prebuilt in the build container")
def test_ppoa_known_answers_on_gpu_cigars(built, agold):
    """test/ppoa_test.cpp's expected consensus strings, with the CIGARs coming from the GPU."""
    T = S.intree()
    for kat in agold["ppoa_kat"]:
        reads = [u8(s) for s in kat["seqs"]]
        bases, off, lens = S.pack_reads(reads)
        buf = C.create_string_buffer(1000)
        n = T.ref_ppoa_consensus(S.ptr(bases), S.ptr(off), S.ptr(lens), len(reads), buf, 1000)
        assert buf.raw[:n].decode() == kat["expected"]


@pytest.mark.skipif(not (S.have_intree() and S.have_ref()), reason="needs both prebuilt pipeline libraries")
def test_whole_regions_equal_reference_live(built):
    """Same comparison on fresh regions: reference on the CPU vs reference-on-GPU-hot-path."""
    T, R = S.intree(), S.ref()
    rng = np.random.default_rng(404)
    n_alleles = []
    for trial in range(12):
        motif = S.rand_seq(rng, int(rng.integers(2, 7)))
        fl, fr = S.rand_seq(rng, 100), S.rand_seq(rng, 100)
        c0 = int(rng.integers(30, 200))
        al = [np.concatenate([fl, np.tile(motif, c), fr]) for c in (c0, c0 + int(rng.integers(0, 40)))]
        e = float(rng.choice([0.005, 0.02, 0.05]))
        n = int(rng.integers(4, 40))
        reads = [S.mutate(rng, al[int(rng.integers(0, 2))], e, n_rate=0.05) for _ in range(n)]
        span = [3] * n
        for a in (0, 1):
            r = S.mutate(rng, al[a], e)
            cut = int(rng.integers(10, 80))
            reads += [r[:-cut], r[cut:], r[cut // 2:-cut // 2]]
            span += [1, 2, 0]
        a = S.assemble_region(R, reads, span)
        b = S.assemble_region(T, reads, span)
        assert a == b, trial
        n_alleles.append(a["n_alleles"])
    assert max(n_alleles) == 2


@pytest.mark.skipif(not (S.have_intree() and S.have_ref()), reason="needs both prebuilt pipeline libraries")
def test_wave_of_regions_equals_reference(built):
    """assemble_process's loop body over many regions at once: the in-tree build runs every
    region's distance matrix, re-assignment and consensus alignments as ONE batch each on the GPU
    (clustering and partial-order consensus on host threads); labels, allele counts and allele
    digests must equal the reference run region by region, and region by region through the
    single-region entry point."""
    T, R = S.intree(), S.ref()
    rng = np.random.default_rng(808)
    regions, spans = [], []
    for trial in range(24):
        motif = S.rand_seq(rng, int(rng.integers(2, 7)))
        fl, fr = S.rand_seq(rng, 100), S.rand_seq(rng, 100)
        c0 = int(rng.integers(30, 160))
        al = [np.concatenate([fl, np.tile(motif, c), fr]) for c in (c0, c0 + int(rng.integers(0, 40)))]
        e = float(rng.choice([0.005, 0.02, 0.05]))
        n = int(rng.integers(1, 30))
        reads = [S.mutate(rng, al[int(rng.integers(0, 2))], e, n_rate=0.05) for _ in range(n)]
        span = [3] * n
        if trial % 3 == 0:  # partial reads: re-assignment + ends-free consensus alignments
            for a in (0, 1):
                r = S.mutate(rng, al[a], e)
                cut = int(rng.integers(10, 80))
                reads += [r[:-cut], r[cut:], r[cut // 2:-cut // 2]]
                span += [1, 2, 0]
        if trial == 5:  # a region without any spanning read
            span = [1] * len(span)
        regions.append(reads)
        spans.append(span)
    la, na, da = S.assemble_regions(R, regions, spans, threads=4)
    lb, nb, db = S.assemble_regions(T, regions, spans, threads=4)
    assert na == nb and la == lb and da == db
    assert na[5] == 0 and max(na) == 2
    one = S.assemble_region(R, regions[0], spans[0])
    assert one["labels"] == la[0] and one["n_alleles"] == na[0]
    # the same in five overlapped waves, device stages from one issuing thread and from two
    # (wave_b200.hpp, device_turn.hpp): results depend on neither
    for issuers in ("1", "2"):
        os.environ["OTREF_WAVE"], os.environ["OTREF_ISSUERS"] = "5", issuers
        try:
            lc, nc, dc = S.assemble_regions(T, regions, spans, threads=4)
        finally:
            del os.environ["OTREF_WAVE"], os.environ["OTREF_ISSUERS"]
        assert nc == na and lc == la and dc == da, issuers


def test_skipped_consensus_call_replays_the_last_cigar_deterministically(built):
    """src/analignments.cpp:262-281: a member longer than the representative and spanning on
    neither side gets NO aligner call, and the reference inserts whatever CIGAR its thread-local
    aligner holds.  Inside a batch that is the previous member's CIGAR (reproduced); for a skipped
    call that comes first it is schedule-dependent in the reference, and deliberately the facade's
    own last single-call CIGAR here (empty in a fresh thread) - realign_b200.cpp."""
    import ctypes as C
    H = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    H.othost_consensus_cigars.restype = C.c_uint64
    H.othost_consensus_cigars.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, C.c_int, C.c_char_p, C.c_uint64]
    rng = np.random.default_rng(12)
    base = S.rand_seq(rng, 300)
    longer = np.concatenate([base, S.rand_seq(rng, 40)])      # longer than any representative
    spanning = [S.mutate(rng, base, 0.01) for _ in range(4)]

    def run(order):
        reads, span = [], []
        for kind in order:
            if kind == "L":
                reads.append(longer); span.append(0)          # spanning on neither side
            else:
                reads.append(spanning[int(kind)]); span.append(3)
        bases, off, lens = S.pack_reads(reads)
        labels = np.zeros(len(reads), dtype=np.int32)
        sp = np.array(span, dtype=np.uint8)
        buf = C.create_string_buffer(1 << 16)
        n = H.othost_consensus_cigars(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(sp), len(reads),
                                      S.ptr(labels), 1, buf, len(buf))
        return buf.raw[:n].decode().split("\n")[:-1]

    first = run(["L", "0", "1", "2", "3"])      # the skipped member is read 0: first in plan order
    assert len(first) == 4 and first[0] == ""   # nothing aligned before it in this fresh aligner
    later = run(["0", "1", "L", "2", "3"])      # skipped member after an aligned one
    assert len(later) == 4
    k = [i for i, c in enumerate(later) if i > 0 and c == later[i - 1]]
    assert k, later                              # it repeats its predecessor's CIGAR
