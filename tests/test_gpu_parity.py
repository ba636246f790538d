"""GPU parity tests (run on the B200 box): the CUDA hot path, called through the C ABI
(include/otter_b200.h) and through the C++ host layer, against the golden vectors of the
unmodified reference and against the oracle port on fresh inputs.  Bit-exact or fail."""
import ctypes as C
import itertools
import os

import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def PAIRS(rows):
    from otter_b200.capi import PAIR_DTYPE
    return np.array(rows, dtype=PAIR_DTYPE)


def u8(s):
    return np.frombuffer(s.encode() if isinstance(s, str) else s, dtype=np.uint8)


def port_scores(reads, pairs):
    P = S.port()
    out = []
    for p, t, pb, pe, tb, te in pairs.tolist():
        x, y = S.b(reads[p]), S.b(reads[t])
        out.append(P.wfo_edit_score(x, len(x), y, len(y), int(pb >= 0), max(pb, 0), pe, tb, te, None))
    return np.array(out, dtype=np.int32)


def host_lib():
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    l.othost_align_anreads.restype = C.c_double
    l.othost_align_anreads.argtypes = [C.c_char_p, C.c_uint32, C.c_int, C.c_char_p, C.c_uint32, C.c_int]
    l.othost_fill_dist_matrix.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, C.c_uint32, vp]
    l.othost_fill_dist_matrices.argtypes = [vp, vp, vp, vp, vp, C.c_uint32, vp, vp]
    l.othost_invalid_reassignment.argtypes = [vp, vp, vp, vp, C.c_uint32, C.c_double, C.c_double,
                                              C.c_int, vp]
    l.othost_fill_dist_matrices_multi.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, C.c_int, vp, vp, vp]
    return l


def test_host_layer_multi_device_equals_single(built):
    """fill_dist_matrices_multi (one process, one aligner per device, regions split by work, no
    collective) fills the same matrices as one aligner.  With one GPU the aligners share device 0
    (two contexts, two host threads); with more, every device of the box takes a share."""
    import otter_b200
    H = host_lib()
    rng = np.random.default_rng(81)
    regions = [S.tr_region(rng, int(n), int(l), e, extra_frac=0.3)
               for n, l, e in zip(rng.integers(1, 14, 40), rng.integers(200, 1500, 40),
                                  rng.choice([0.005, 0.05], 40))]
    B = S.region_batch(regions)
    one = np.zeros(B["n_scores"])
    H.othost_fill_dist_matrices(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]), None,
                                S.ptr(B["region_start"]), len(regions), S.ptr(one),
                                S.ptr(B["score_off"]))
    n_dev = otter_b200.lib().otgpu_device_count()
    layouts = [[0, 0], [0, 0, 0]] + ([list(range(n_dev))] if n_dev > 1 else [])
    for devices in layouts:
        dev = np.array(devices, np.int32)
        got = np.full(B["n_scores"], -7.0)
        part = np.full(len(regions), -1, np.int32)
        H.othost_fill_dist_matrices_multi(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]),
                                          S.ptr(B["region_start"]), len(regions), S.ptr(dev),
                                          len(devices), S.ptr(got), S.ptr(B["score_off"]), S.ptr(part))
        assert np.array_equal(got, one), devices
        assert set(part.tolist()) == set(range(len(devices)))  # every aligner took a share


# ---- C ABI vs golden ---------------------------------------------------------------------------
def test_golden_pairs_through_c_abi(ctx, golden):
    reads, rows = [], []
    for g in golden["pairs"]:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, g["pb"] if g["ef"] else -1, g["pe"], g["tb"], g["te"]))
    bases, off, lens = S.pack_reads(reads)
    got = ctx.edit_scores(bases, off, lens, PAIRS(rows))
    exp = np.array([g["score"] for g in golden["pairs"]], dtype=np.int32)
    assert np.array_equal(got, exp), np.nonzero(got != exp)
    st = ctx.stats()
    assert st["n_pairs_8bit"] > 0  # the raw-byte variant was exercised
    # the batch holds a 14 kb read, so it ran on the 2048-slot ring; pin the 1024-slot ring to
    # send the widest pair (1,167 diagonals) through the HBM-ring kernel as well
    import otter_b200
    os.environ["OTGPU_RING_SLOTS"] = "1024"
    try:
        with otter_b200.Context(0) as c2:
            got2 = c2.edit_scores(bases, off, lens, PAIRS(rows))
            assert np.array_equal(got2, exp)
            assert c2.stats()["n_pairs_wide"] > 0
    finally:
        del os.environ["OTGPU_RING_SLOTS"]


def test_long_read_tier_finishes_wide_pairs_inside_the_kernel(built):
    """Batches of long reads (2048-slot tier): a pair whose wavefront outgrows the shared ring is
    started again on the warp's HBM ring inside k_align<2048>; OTGPU_WIDE_TWO_PASS=1 keeps the
    older second kernel (k_align<0>).  Same scores either way, and equal to the oracle.
    The two 45 kb regions hold pairs whose wavefronts reach ~2,500 diagonals (instrumented port)."""
    import otter_b200
    regions = [S.tr_region(np.random.default_rng(seed), 4, 45000, 0.05, extra_frac=0.1)
               for seed in (26, 27)]
    B = S.region_batch(regions)
    out = {}
    for two_pass in ("0", "1"):
        os.environ["OTGPU_WIDE_TWO_PASS"] = two_pass
        try:
            with otter_b200.Context(0) as c:
                out[two_pass] = c.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"],
                                                     B["score_off"], B["n_scores"])
                assert c.stats()["n_pairs_wide"] > 0, "workload no longer reaches the HBM ring"
        finally:
            del os.environ["OTGPU_WIDE_TWO_PASS"]
    assert np.array_equal(out["0"], out["1"])
    P = S.port()
    q = 0
    for reads in regions:
        for i in range(len(reads)):
            for j in range(i + 1, len(reads)):
                x, y = S.b(reads[i]), S.b(reads[j])
                if len(x) < len(y):
                    x, y = y, x
                assert out["0"][q] == P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None), (q, i, j)
                q += 1


def test_golden_regions_through_region_abi(ctx, golden):
    regs = [g for g in golden["regions"] if all(s == 3 for s in g["span"])]
    B = S.region_batch([[u8(r) for r in g["reads"]] for g in regs])
    got = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                                 B["n_scores"])
    q = 0
    for g in regs:
        n = len(g["reads"])
        ls = [len(r) for r in g["reads"]]
        for i, j in itertools.combinations(range(n), 2):
            d = got[q] / float(max(ls[i], ls[j]))  # src/analignments.cpp:68-72
            assert d == float.fromhex(g["dist"][q - int(B["score_off"][regs.index(g)])])
            q += 1
    assert q == B["n_scores"]


# ---- C++ host layer (reference call surface) vs golden -----------------------------------------
def test_golden_align_anreads_through_host_layer(built, golden):
    H = host_lib()
    for g in golden["anreads"]:
        x, y = g["x"].encode(), g["y"].encode()
        d = H.othost_align_anreads(x, len(x), g["xspan"], y, len(y), g["yspan"])
        assert d == float.fromhex(g["dist"]), (g["xspan"], g["yspan"], len(x), len(y))


def test_golden_fill_dist_matrix_through_host_layer(built, golden):
    H = host_lib()
    for g in golden["regions"]:
        reads = [u8(r) for r in g["reads"]]
        bases, off, lens = S.pack_reads(reads)
        span = np.array(g["span"], dtype=np.uint8)
        idx = np.array(g["indeces"], dtype=np.int32)
        out = np.full(max(len(g["dist"]), 1), -7.0)
        H.othost_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(span), len(reads),
                                  S.ptr(idx), len(idx), S.ptr(out))
        exp = np.array([float.fromhex(h) for h in g["dist"]])
        assert np.array_equal(out[:len(exp)], exp)


def test_golden_invalid_reassignment_through_host_layer(built, golden):
    H = host_lib()
    changed = 0
    for g in golden["reassign"]:
        reads = [u8(r) for r in g["reads"]]
        bases, off, lens = S.pack_reads(reads)
        span = np.array(g["span"], dtype=np.uint8)
        lab = np.array(g["labels_in"], dtype=np.int32)
        H.othost_invalid_reassignment(S.ptr(bases), S.ptr(off), S.ptr(lens), S.ptr(span), len(reads),
                                      g["min_sim"], g["max_error"], g["total_alleles"], S.ptr(lab))
        assert lab.tolist() == g["labels_out"]
        changed += sum(a != b for a, b in zip(g["labels_in"], g["labels_out"]))
    assert changed > 0


# ---- C ABI vs the oracle port on fresh inputs ---------------------------------------------------
@pytest.mark.parametrize("tl,e,xf,n", [(150, 0.01, 0.2, 10), (400, 0.08, 0.5, 10), (2000, 0.005, 0.2, 8),
                                       (2000, 0.05, 0.2, 8), (900, 0.3, 0.3, 6), (6000, 0.05, 0.3, 5)])
def test_regions_vs_port(ctx, tl, e, xf, n):
    rng = np.random.default_rng(hash((tl, n)) % 2**32)
    regions = [S.tr_region(rng, n, tl, e, extra_frac=xf, n_rate=0.1) for _ in range(3)]
    B = S.region_batch(regions)
    got = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                                 B["n_scores"])
    P = S.port()
    q = 0
    for reg in regions:
        for i, j in itertools.combinations(range(len(reg)), 2):
            x, y = S.b(reg[i]), S.b(reg[j])
            if len(x) < len(y):
                x, y = y, x  # pattern = longer, tie -> read i
            assert got[q] == P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None), (q, i, j)
            q += 1


def test_ends_free_forms_vs_port(ctx):
    rng = np.random.default_rng(17)
    reads = [r for tl, e in ((300, 0.02), (1200, 0.05), (2500, 0.01))
             for r in S.tr_region(rng, 6, tl, e, extra_frac=0.5, n_rate=0.1)]
    bases, off, lens = S.pack_reads(reads)
    rows = []
    for i in range(len(reads) - 1):
        p, t = (i, i + 1) if lens[i] >= lens[i + 1] else (i + 1, i)
        d = int(lens[p]) - int(lens[t])
        rows += [(p, t, 0, d, 0, 0), (p, t, d, 0, 0, 0), (p, t, d // 2, d // 2, 0, 0),
                 (t, p, 0, 0, 0, d), (t, p, 0, 0, d, 0), (t, p, 0, 0, d // 2, d // 2),
                 (p, t, 0, 0, 0, 0), (p, t, -1, 0, 0, 0)]
    pairs = PAIRS(rows)
    got = ctx.edit_scores(bases, off, lens, pairs)
    assert np.array_equal(got, port_scores(reads, pairs))


def test_edge_cases_vs_port(ctx):
    seqs = [b"", b"A", b"N", b"ACGT", b"ACGTN", b"A" * 15, b"A" * 16, b"A" * 17, b"A" * 31,
            b"A" * 32, b"A" * 33, b"A" * 64, b"A" * 65, b"ACGT" * 300, b"ACGT" * 300 + b"A",
            b"acgt" * 20, b"ACGT" * 20, b"\x00\xff" * 10, b"T" + b"ACGT" * 300]
    reads = [np.frombuffer(s, dtype=np.uint8) for s in seqs]
    bases, off, lens = S.pack_reads(reads)
    rows = []
    for i, j in itertools.product(range(len(reads)), repeat=2):
        p, t = (i, j) if lens[i] >= lens[j] else (j, i)
        rows.append((p, t, -1, 0, 0, 0))
    pairs = PAIRS(rows)
    got = ctx.edit_scores(bases, off, lens, pairs)
    assert np.array_equal(got, port_scores(reads, pairs))
    assert all(got[i * len(reads) + i] == 0 for i in range(len(reads)))  # identical -> 0 (:64)


def test_unaligned_offsets_and_shared_reads(ctx):
    """Reads may sit at any byte offset of the caller's buffer and be used by many pairs."""
    rng = np.random.default_rng(23)
    reads = S.tr_region(rng, 6, 700, 0.05, extra_frac=0.3, n_rate=0.5)
    pad = [rng.integers(0, 7) for _ in reads]
    chunks, off, pos = [], [], 0
    for r, p in zip(reads, pad):
        chunks.append(np.full(p, ord("#"), dtype=np.uint8))
        pos += p
        off.append(pos)
        chunks.append(r)
        pos += len(r)
    bases = np.concatenate(chunks)
    off = np.array(off, dtype=np.uint64)
    lens = np.array([len(r) for r in reads], dtype=np.uint32)
    rows = [((i, j) if lens[i] >= lens[j] else (j, i)) + (-1, 0, 0, 0)
            for i, j in itertools.combinations(range(len(reads)), 2)]
    pairs = PAIRS(rows)
    got = ctx.edit_scores(bases, off, lens, pairs)
    assert np.array_equal(got, port_scores(reads, pairs))


def test_empty_and_degenerate_batches(ctx):
    z8, z64, z32 = np.zeros(0, np.uint8), np.zeros(0, np.uint64), np.zeros(0, np.uint32)
    assert len(ctx.edit_scores(z8, z64, z32, PAIRS([]))) == 0
    # regions with 0, 1 and 2 reads (src/otterclust.cpp:123-148 handles n=1,2 downstream)
    rng = np.random.default_rng(1)
    regs = [[], S.tr_region(rng, 1, 200, 0.01), S.tr_region(rng, 2, 200, 0.01), []]
    B = S.region_batch(regs)
    got = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                                 B["n_scores"])
    assert len(got) == 1


def test_bad_arguments_are_errors(ctx):
    import otter_b200
    reads = [u8("ACGT"), u8("ACGA")]
    bases, off, lens = S.pack_reads(reads)
    with pytest.raises(otter_b200.OtgpuError):
        ctx.edit_scores(bases, off, lens, PAIRS([(0, 2, -1, 0, 0, 0)]))  # read index out of range
    with pytest.raises(otter_b200.OtgpuError):
        ctx.edit_scores(bases, off, lens, PAIRS([(0, 1, 5, 0, 0, 0)]))  # free end > length
    bad_off = off.copy()
    bad_off[1] = 100
    with pytest.raises(otter_b200.OtgpuError):
        ctx.edit_scores(bases, bad_off, lens, PAIRS([(0, 1, -1, 0, 0, 0)]))
    # the context is still usable afterwards
    assert ctx.edit_scores(bases, off, lens, PAIRS([(0, 1, -1, 0, 0, 0)]))[0] == 1


def test_maximum_read_length(ctx):
    """Reads up to the documented limit (2^20 bases; offsets, ring nulls and buffer padding are
    sized for it) against the port; one base more is an argument error, not a wrong answer."""
    import otter_b200
    rng = np.random.default_rng(77)
    top = 1 << 20
    a = S.rand_seq(rng, top)
    b = S.mutate(rng, a, 0.0003)[:top]            # a few hundred edits over a megabase
    c = S.mutate(rng, a[:300000], 0.001)
    d = a[:300000].copy()
    d[1000] = ord("N")                            # 8-bit fallback on a long read
    reads = [a, b, c, d, a[:300000]]
    bases, off, lens = S.pack_reads(reads)
    rows = [(0, 1, -1, 0, 0, 0), (4, 2, -1, 0, 0, 0) if lens[4] >= lens[2] else (2, 4, -1, 0, 0, 0),
            (3, 4, -1, 0, 0, 0), (0, 4, 0, top - 300000, 0, 0)]
    pairs = PAIRS(rows)
    got = ctx.edit_scores(bases, off, lens, pairs)
    assert np.array_equal(got, port_scores(reads, pairs))
    assert got[2] == 1
    too_long = [np.concatenate([a, a[:1]]), a[:10]]
    bases, off, lens = S.pack_reads(too_long)
    with pytest.raises(otter_b200.OtgpuError):
        ctx.edit_scores(bases, off, lens, PAIRS([(0, 1, -1, 0, 0, 0)]))


# ---- full-size, size-independent properties ----------------------------------------------------
def test_full_size_properties(ctx):
    """BASELINE config C1 at 1/10 scale plus structural properties that need no oracle:
    determinism, region ABI == explicit-pair ABI, identical reads -> 0, score bounds,
    |len difference| <= score <= max(len)."""
    regions = S.make_regions("C1", seed=3, n_regions=100)
    # duplicate one read inside a region so the equal-string case is present
    regions[0][1] = regions[0][0].copy()
    B = S.region_batch(regions)
    a = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                               B["n_scores"])
    b = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                               B["n_scores"])
    assert np.array_equal(a, b)
    x, y = S.pair_lengths(B)
    assert (a >= np.abs(x - y)).all() and (a <= np.maximum(x, y)).all()
    assert a[0] == 0
    rows = []
    rs = B["region_start"]
    for g in range(len(regions)):
        for i, j in itertools.combinations(range(rs[g], rs[g + 1]), 2):
            rows.append(((i, j) if B["lens"][i] >= B["lens"][j] else (j, i)) + (-1, 0, 0, 0))
    c = ctx.edit_scores(B["bases"], B["off"], B["lens"], PAIRS(rows))
    assert np.array_equal(a, c)
    # spot-check 200 random pairs against the port
    rng = np.random.default_rng(0)
    pick = rng.choice(len(rows), 200, replace=False)
    sel = PAIRS([rows[i] for i in pick])
    assert np.array_equal(a[pick], port_scores(B["reads"], sel))


def test_host_layer_batched_equals_per_region(built):
    H = host_lib()
    rng = np.random.default_rng(8)
    regions = [S.tr_region(rng, n, 600, 0.05, extra_frac=0.3) for n in (5, 1, 8, 2, 6)]
    B = S.region_batch(regions)
    out = np.zeros(B["n_scores"])
    H.othost_fill_dist_matrices(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]), None,
                                S.ptr(B["region_start"]), len(regions), S.ptr(out),
                                S.ptr(B["score_off"]))
    P = S.port()
    exp = np.zeros_like(out)
    for g, reg in enumerate(regions):
        bases, off, lens = S.pack_reads(reg)
        idx = np.arange(len(reg), dtype=np.int32)
        n = len(reg) * (len(reg) - 1) // 2
        if n:
            P.wfo_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), None, S.ptr(idx), len(reg),
                                   S.ptr(exp[int(B["score_off"][g]):]))
    assert np.array_equal(out, exp)


def test_smoke_entry(built):
    import __graft_entry__ as g
    g.smoke()


def test_streamed_chunks_equal_single_batch(ctx):
    """otgpu_region_edit_scores streams large batches in chunks over two buffer sets; with the
    chunk size forced down to 1 MB a 3 MB batch must take several chunks and give the same scores."""
    import otter_b200
    rng = np.random.default_rng(404)
    regions = [S.tr_region(rng, 12, int(rng.integers(1500, 4000)), 0.03, extra_frac=0.2) for _ in range(90)]
    regions.insert(7, [])            # an empty region and a single-read region inside a chunk
    regions.insert(30, regions[3][:1])
    B = S.region_batch(regions)
    assert B["bases"].size > 2_500_000
    exp = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"], B["n_scores"])
    assert ctx.last_chunk_count() == 1
    os.environ["OTGPU_CHUNK_MB"] = "1"
    try:
        with otter_b200.Context(0) as c2:
            for _ in range(2):  # second call reuses both buffer sets
                got = c2.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"],
                                            B["score_off"], B["n_scores"])
                assert c2.last_chunk_count() >= 3
                assert np.array_equal(got, exp)
                assert c2.stats()["n_pairs"] == B["n_scores"]
    finally:
        del os.environ["OTGPU_CHUNK_MB"]


def test_first_window_follows_the_previous_batch_and_never_changes_scores(built, monkeypatch):
    """k_align<1024> has two flavours (64-bit first windows / 32-bit first windows + 128-bit
    continuations).  A context starts with the first, counts the low-error pairs of every call
    and takes the second for the next call when they were >= 30 %.  Scores never depend on it."""
    import otter_b200
    rng = np.random.default_rng(77)
    hifi = S.region_batch([S.tr_region(rng, 10, 1500, 0.005, extra_frac=0.2, n_rate=0.1) for _ in range(6)])
    noisy = S.region_batch([S.tr_region(rng, 10, 1500, 0.06, extra_frac=0.2, n_rate=0.1) for _ in range(6)])
    run = lambda c, B: c.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"],
                                            B["score_off"], B["n_scores"])
    monkeypatch.delenv("OTGPU_FIRST_WINDOW", raising=False)
    with otter_b200.Context(0) as c:
        assert c.first_window() == 0
        h0 = run(c, hifi)                     # 64-bit windows; counted as a low-error batch
        assert c.first_window() == 1
        h1 = run(c, hifi)                     # 32-bit windows
        assert c.first_window() == 1
        n1 = run(c, noisy)                    # still 32-bit windows (slower, same scores)
        assert c.first_window() == 0
        n0 = run(c, noisy)
        assert c.first_window() == 0
    assert np.array_equal(h0, h1) and np.array_equal(n0, n1)
    for force in ("0", "1"):
        monkeypatch.setenv("OTGPU_FIRST_WINDOW", force)
        with otter_b200.Context(0) as c:
            assert c.first_window() == int(force)
            assert np.array_equal(run(c, hifi), h0) and np.array_equal(run(c, noisy), n0)
            assert c.first_window() == int(force)
    # and against the port, pair by pair
    P = S.port()
    for B, got in ((hifi, h0), (noisy, n0)):
        rs, lens, off = B["region_start"], B["lens"], B["off"]
        q = 0
        for g in range(len(rs) - 1):
            for i, j in itertools.combinations(range(int(rs[g]), int(rs[g + 1])), 2):
                x = B["bases"][int(off[i]):int(off[i]) + int(lens[i])].tobytes()
                y = B["bases"][int(off[j]):int(off[j]) + int(lens[j])].tobytes()
                if len(x) < len(y):
                    x, y = y, x
                assert got[q] == P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None), (g, i, j)
                q += 1


def test_streamed_batch_with_trailing_empty_regions(built):
    """A streamed batch whose last regions hold no reads (region_start[g] == n_reads): they are
    folded into the last chunk instead of starting a chunk that would address read n_reads."""
    import otter_b200
    rng = np.random.default_rng(405)
    regions = [S.tr_region(rng, 10, int(rng.integers(1500, 3000)), 0.03, extra_frac=0.2) for _ in range(70)]
    regions += [[], [], []]
    B = S.region_batch(regions)
    with otter_b200.Context(0) as c1:
        exp = c1.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                                    B["n_scores"])
    os.environ["OTGPU_CHUNK_MB"] = "1"
    try:
        with otter_b200.Context(0) as c2:
            got = c2.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"],
                                        B["score_off"], B["n_scores"])
            assert c2.last_chunk_count() >= 2
    finally:
        del os.environ["OTGPU_CHUNK_MB"]
    assert np.array_equal(got, exp)


def test_hooked_call_waits_for_bases_and_hands_scores_back_in_order(built):
    """otgpu_region_edit_scores_hooked: the caller fills `bases` while the call runs.  Here the
    buffer starts as garbage and wait_bases copies the real bytes in only as far as it is asked to;
    scores_ready must cover the score array once, in order, and every range must already hold
    its final values when the hook runs."""
    import otter_b200
    rng = np.random.default_rng(406)
    regions = [S.tr_region(rng, 9, int(rng.integers(1500, 3500)), 0.04, extra_frac=0.3) for _ in range(200)]
    B = S.region_batch(regions)
    with otter_b200.Context(0) as c1:
        exp = c1.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"],
                                    B["n_scores"])
    os.environ["OTGPU_CHUNK_MB"] = "1"
    try:
        with otter_b200.Context(0) as c2:
            late = np.full_like(B["bases"], ord("T"))
            filled, asked, ready = [0], [], []

            def wait_bases(upto):
                asked.append(upto)
                late[filled[0]:upto] = B["bases"][filled[0]:upto]
                filled[0] = max(filled[0], upto)

            def scores_ready(first, n, out):
                ready.append((first, n))
                assert np.array_equal(out[first:first + n], exp[first:first + n])

            got = c2.region_edit_scores_hooked(late, B["off"], B["lens"], B["region_start"], B["score_off"],
                                               B["n_scores"], wait_bases, scores_ready)
            chunks = c2.last_chunk_count()
    finally:
        del os.environ["OTGPU_CHUNK_MB"]
    assert np.array_equal(got, exp)
    assert chunks >= 3 and len(asked) == chunks and len(ready) == chunks
    assert asked == sorted(asked) and asked[-1] == B["bases"].size and asked[0] < B["bases"].size
    at = 0
    for first, n in ready:
        assert first == at
        at += n
    assert at == B["n_scores"]


def test_host_layer_overlapped_staging_equals_single_batch(built):
    """fill_dist_matrices feeds the hooked call from worker threads (staging copy of chunk i+1 and
    score conversion of chunk i-1 under the kernels of chunk i): same matrices as one batch, also
    with regions that have no pairs at either end."""
    H = host_lib()
    rng = np.random.default_rng(407)
    regions = [[]] + [S.tr_region(rng, int(n), int(l), 0.03, extra_frac=0.3)
                      for n, l in zip(rng.integers(1, 12, 1300), rng.integers(800, 2500, 1300))] + [[], []]
    B = S.region_batch(regions)
    assert B["bases"].size > (10 << 20)   # large enough for the helper threads (> 8 MiB)
    one = np.zeros(B["n_scores"])
    H.othost_fill_dist_matrices(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]), None,
                                S.ptr(B["region_start"]), len(regions), S.ptr(one), S.ptr(B["score_off"]))
    os.environ["OTGPU_CHUNK_MB"] = "1"
    try:
        import subprocess, sys, textwrap
        # a fresh process: the host layer's aligner (and its context's chunk size) is made once per thread
        code = textwrap.dedent("""
            import sys, numpy as np
            sys.path.insert(0, %r); sys.path.insert(0, %r)
            import support as S, test_gpu_parity as T
            Z = np.load(sys.argv[1])
            B = {k: Z[k] for k in Z.files}   # (NpzFile makes a fresh array per access)
            H = T.host_lib()
            out = np.zeros(int(B["n_scores"]))
            H.othost_fill_dist_matrices(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(B["lens"]), None,
                                        S.ptr(B["region_start"]), len(B["region_start"]) - 1, S.ptr(out),
                                        S.ptr(B["score_off"]))
            np.save(sys.argv[2], out)
        """ % (ROOT, os.path.join(ROOT, "tests")))
        import tempfile
        d = tempfile.mkdtemp()
        np.savez(os.path.join(d, "b.npz"), bases=B["bases"], off=B["off"], lens=B["lens"],
                 region_start=B["region_start"], score_off=B["score_off"], n_scores=B["n_scores"])
        subprocess.check_call([sys.executable, "-c", code, os.path.join(d, "b.npz"), os.path.join(d, "o.npy")])
        got = np.load(os.path.join(d, "o.npy"))
    finally:
        del os.environ["OTGPU_CHUNK_MB"]
    assert np.array_equal(got, one)
    P = S.port()
    g = 5
    reg = regions[g]
    bases, off, lens = S.pack_reads(reg)
    n = len(reg) * (len(reg) - 1) // 2
    if n:
        exp = np.zeros(n)
        P.wfo_fill_dist_matrix(S.ptr(bases), S.ptr(off), S.ptr(lens), None,
                               S.ptr(np.arange(len(reg), dtype=np.int32)), len(reg), S.ptr(exp))
        assert np.array_equal(one[int(B["score_off"][g]):int(B["score_off"][g]) + n], exp)
