"""Batched read extraction (otgpu_extract_reads; reference get_breakpoints + parse_alignment,
src/anseqs.cpp:286-436) against golden vectors generated from the unmodified reference, live
against the reference where oracle/_ref is built, and chained into the distance-matrix path."""
import ctypes as C
import json
import os
import sys

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(built):
    import otter_b200
    c = otter_b200.Context(0)
    yield c
    c.close()


def batch(cases):
    """Lay records out as BAM holds them: CIGAR words and 4-bit sequences back to back."""
    from otter_b200.capi import BAM_READ_DTYPE, encode_seq4
    import make_golden_extract as G
    rec = np.zeros(len(cases), dtype=BAM_READ_DTYPE)
    cig, seq4, co, so = [], [], 0, 0
    for i, c in enumerate(cases):
        w = G.cigar_words([tuple(x) for x in c["cigar"]])
        s4 = encode_seq4(c["seq"])
        rec[i] = (c["pos"], len(c["seq"]), len(w), 0, co, so, c["start"], c["end"])
        cig.append(w)
        seq4.append(s4)
        co += len(w)
        so += len(s4)
    return (rec, np.concatenate(cig) if co else np.zeros(0, np.uint32),
            np.concatenate(seq4) if so else np.zeros(0, np.uint8))


def check(ctx, cases):
    rec, cig, seq4 = batch(cases)
    status, coords, lens, begin, bases = ctx.extract_reads(rec, cig, seq4)
    flagged = 0
    for i, c in enumerate(cases):
        st = int(status[i])
        assert (st & 1) == c["ok"], (i, c["pos"], c["cigar"][:4])
        assert ((st >> 1) & 3) == c["span"], i
        if st & 8:  # the reference exits or reads outside the record here: nothing to compare
            flagged += 1
            continue
        assert coords[i].tolist() == c["coords"], (i, c["cigar"][:6], c["start"], c["end"], c["pos"])
        got = bases[int(begin[i]):int(begin[i]) + int(lens[i])].tobytes().decode() if st & 1 else ""
        assert got == c["out"], i
    return flagged


def test_extract_golden(ctx):
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "extract_golden.json")))["cases"]
    assert len(gold) >= 250
    assert check(ctx, gold) <= len(gold) // 20
    outs = [c["out"] for c in gold if c["ok"]]
    assert sum(o == "N" for o in outs) >= 3 and sum(len(o) > 50 for o in outs) > 100


def test_extract_empty_and_bad_arguments(ctx):
    import otter_b200
    from otter_b200.capi import BAM_READ_DTYPE
    st, co, ln, bg, bs = ctx.extract_reads(np.zeros(0, BAM_READ_DTYPE), np.zeros(0, np.uint32), np.zeros(0, np.uint8))
    assert len(st) == 0 and len(bs) == 0
    rec = np.zeros(1, dtype=BAM_READ_DTYPE)
    rec[0] = (10, 8, 1, 0, 5, 0, 10, 12)  # CIGAR offset outside the array
    with pytest.raises(otter_b200.OtgpuError):
        ctx.extract_reads(rec, np.array([8 << 4], np.uint32), np.zeros(4, np.uint8))
    rec[0] = (10, 80, 1, 0, 0, 0, 10, 12)  # sequence longer than the bytes given
    with pytest.raises(otter_b200.OtgpuError):
        ctx.extract_reads(rec, np.array([80 << 4], np.uint32), np.zeros(4, np.uint8))


@pytest.mark.skipif(not S.have_ref(), reason="oracle/_ref is built in the build container only")
def test_extract_live_vs_reference(ctx):
    import make_golden_extract as G
    R = S.ref()
    rng = np.random.default_rng(4242)
    cases = []
    for _ in range(1500):
        start = int(rng.integers(500, 50000))
        end = start + int(rng.integers(1, 2500))
        pos, cig, seq = G.random_record(rng, start, end)
        cases.append(dict(pos=pos, cigar=cig, seq=seq, start=start, end=end,
                          **G.ref_parse(R, pos, cig, seq, start, end)))
    assert check(ctx, cases) <= len(cases) // 20


def test_extracted_reads_feed_the_distance_matrix(ctx):
    """A region's records -> otgpu_extract_reads -> otgpu_region_edit_scores, the pointers handed
    over as they come back; same scores as aligning the expected substrings."""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "extract_golden.json")))["cases"]
    use = [c for c in gold if c["ok"] and c["span"] == 3 and len(c["out"]) > 20][:24]
    rec, cig, seq4 = batch(use)
    status, coords, lens, begin, bases = ctx.extract_reads(rec, cig, seq4)
    n = len(use)
    rs = np.array([0, n], dtype=np.uint32)
    so = np.array([0], dtype=np.uint64)
    got = ctx.region_edit_scores(bases, begin[:n].copy(), lens, rs, so, n * (n - 1) // 2)
    reads = [np.frombuffer(c["out"].encode(), dtype=np.uint8) for c in use]
    B = S.region_batch([reads])
    want = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"], B["n_scores"])
    assert np.array_equal(got, want)


def test_extracted_reads_stay_on_the_device(built):
    """otgpu_extract_reads leaves its sequences on the device; otgpu_region_edit_scores_extracted
    builds the alignment batch from them device to device: a subset of the records, in a different
    order, over several regions (one empty), gives the scores of the same reads uploaded from the
    host.  A CIGAR call in between must not disturb the extraction result; a context without one
    refuses."""
    import otter_b200
    from otter_b200.capi import OtgpuError, PAIR_DTYPE
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "extract_golden.json")))["cases"]
    rec, cig, seq4 = batch(gold)
    with otter_b200.Context(0) as c:
        with pytest.raises(OtgpuError):
            c.region_edit_scores_extracted(np.zeros(2, np.uint32), np.array([0, 2], np.uint32),
                                           np.array([0], np.uint64), 1)
        status, coords, lens, begin, bases = c.extract_reads(rec, cig, seq4)
        ok = [i for i, g in enumerate(gold) if g["ok"] and len(g["out"]) > 5]
        rng = np.random.default_rng(5)
        pick = rng.permutation(ok)[:60]
        sizes = [9, 0, 14, 1, 20, 16]
        rs = np.zeros(len(sizes) + 1, dtype=np.uint32)
        rs[1:] = np.cumsum(sizes)
        npairs = np.array([n * (n - 1) // 2 for n in sizes], dtype=np.uint64)
        so = np.zeros(len(sizes), dtype=np.uint64)
        so[1:] = np.cumsum(npairs[:-1])
        # something else uses the context's CIGAR buffers in between
        two = S.pack_reads([np.frombuffer(b"ACGTACGTAC", np.uint8), np.frombuffer(b"ACGTTCGTAC", np.uint8)])
        c.cigars_packed(two[0], two[1], two[2], np.array([(0, 1, -1, 0, 0, 0)], dtype=PAIR_DTYPE))
        got = c.region_edit_scores_extracted(pick[:int(rs[-1])], rs, so, int(npairs.sum()))
        reads = [np.frombuffer(gold[i]["out"].encode(), dtype=np.uint8) for i in pick[:int(rs[-1])]]
        hb, ho, hl = S.pack_reads(reads)
        want = c.region_edit_scores(hb, ho, hl, rs, so, int(npairs.sum()))
        assert np.array_equal(got, want)
        with pytest.raises(OtgpuError):  # an index past the extracted records
            c.region_edit_scores_extracted(np.array([len(gold) + 3, 0], np.uint32), np.array([0, 2], np.uint32),
                                           np.array([0], np.uint64), 1)


@pytest.mark.parametrize("omit", [0, 1])
def test_read_extractor_host_layer(built, omit):
    """ReadExtractor (otter_b200/host/anseqs_b200.hpp): what parse_anreads keeps per region —
    successful records, optionally only spanning ones — as ANREADs (sequence, flags, ccoords)."""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "extract_golden.json")))["cases"]
    rec, cig, seq4 = batch(gold)
    n_regions = 7
    flat = np.zeros((len(gold), 8), dtype=np.int64)
    for i, r in enumerate(rec):
        flat[i] = (r["pos"], r["l_qseq"], r["n_cigar"], r["cigar_off"], r["seq_off"], i % n_regions,
                   r["start"], r["end"])
    H = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    H.othost_extract_reads.restype = C.c_uint32
    H.othost_extract_reads.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_uint32,
                                       C.c_void_p, C.c_void_p, C.c_uint64]
    meta = np.zeros((len(gold), 5), dtype=np.int32)
    cap = sum(len(c["seq"]) for c in gold) + len(gold)
    out = np.zeros(cap, dtype=np.uint8)
    # records the reference cannot parse (it exits) are left out of this batch
    keep_idx = [i for i, c in enumerate(gold)]
    kept = H.othost_extract_reads(S.ptr(flat), len(gold), S.ptr(cig), S.ptr(seq4), omit, n_regions,
                                  S.ptr(meta), S.ptr(out), cap)
    want = []
    for g in range(n_regions):
        for i in keep_idx:
            c = gold[i]
            if i % n_regions == g and c["ok"] and (not omit or c["span"] == 3):
                want.append((g, c["span"], c["coords"][0], c["coords"][1], c["out"]))
    assert kept == len(want)
    pos = 0
    for k, w in enumerate(want):
        m = meta[k]
        seq = out[pos:pos + int(m[4])].tobytes().decode()
        pos += int(m[4])
        assert (int(m[0]), int(m[1]), int(m[2]), int(m[3]), seq) == w, k
