"""The exact prune (WFA2-lib wavefront/wavefront_compute_edit.c:214-269, call site :341-347).

tests/golden/prune_golden.json holds pairs on which the UNMODIFIED reference's prune actually
removes diagonals under WF-adaptive(10,50,1) (found by oracle/tools/prune_search.c, expected
scores from oracle/_ref).  CPU half: the C restatement reproduces them and the kernel's gate (the
necessary condition under which a step is split around the prune) was open on every firing step.
GPU half: every ring tier of the edit kernel and the edit-CIGAR kernel give the reference's
numbers."""
import hashlib
import json
import os

import numpy as np
import pytest

import support as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def pgold():
    with open(os.path.join(ROOT, "tests", "golden", "prune_golden.json")) as f:
        return json.load(f)["pairs"]


def u8(s):
    return np.frombuffer(s.encode(), dtype=np.uint8)


def test_fixture_covers_both_orientations_and_every_tier(pgold):
    assert len(pgold) >= 16
    assert all(g["prunes"] > 0 and g["max_width"] >= 990 for g in pgold)  # width at extend time, i.e. after the prune
    assert any(len(g["t"]) > len(g["p"]) for g in pgold)  # text longer (facade use; ak > 0)
    assert any(g["max_width"] > 1912 for g in pgold)       # outgrows the 2048-slot shared ring too
    assert any(g["max_width"] <= 1912 for g in pgold)      # stays inside it


def test_port_reproduces_the_reference_where_the_prune_fires(pgold):
    P = S.port()
    for g in pgold:
        p, t = g["p"].encode(), g["t"].encode()
        ctr = S.Counters()
        assert P.wfo_edit_score(p, len(p), t, len(t), 0, 0, 0, 0, 0, ctr) == g["score"]
        assert ctr.prunes == g["prunes"]
        # "prune fired => gate open" (a closed gate on a firing step is recorded as a negative count)
        assert ctr.prune_gate >= ctr.prunes


def test_port_edit_cigar_reproduces_the_reference_where_the_prune_fires(pgold):
    for g in pgold[::3]:
        sc, cg = S.port_edit_cigar(g["p"].encode(), g["t"].encode())
        assert sc == g["cigar_score"] and len(cg) == g["cigar_len"]
        assert hashlib.md5(cg.encode()).hexdigest() == g["cigar_md5"]


def test_gate_is_necessary_on_fresh_adversarial_pairs():
    """The kernel splits a step around the prune only when max(|lo-1-ak|, |hi+1-ak|) >= dm - 1
    (dm = minimum cut-off distance after the last extend).  Fresh unrelated / low-complexity pairs:
    the prune never fires with the gate closed."""
    P = S.port()
    rng = np.random.default_rng(404)
    fired = 0
    for i in range(24):
        if i % 2:
            x = rng.integers(0, 4, int(rng.integers(1500, 2600)))
            y = rng.integers(0, 4, int(rng.integers(600, len(x))))
        else:  # blocks of short motifs
            def blocks(n):
                out = []
                while sum(map(len, out)) < n:
                    m = rng.integers(0, 4, int(rng.integers(1, 4)))
                    out.append(np.tile(m, int(rng.integers(20, 700))))
                return np.concatenate(out)
            x, y = blocks(2500), blocks(1200)
            if len(x) < len(y):
                x, y = y, x
        px, py = S.b(S.ALPHA[x]), S.b(S.ALPHA[y])
        ctr = S.Counters()
        P.wfo_edit_score(px, len(px), py, len(py), 0, 0, 0, 0, 0, ctr)
        assert ctr.prune_gate >= ctr.prunes
        fired += ctr.prunes > 0
    assert fired > 0


# ---- GPU ------------------------------------------------------------------------------------------
def _batch(pgold):
    from otter_b200.capi import PAIR_DTYPE
    reads, rows = [], []
    for g in pgold:
        reads += [u8(g["p"]), u8(g["t"])]
        rows.append((len(reads) - 2, len(reads) - 1, -1, 0, 0, 0))
    return S.pack_reads(reads), np.array(rows, dtype=PAIR_DTYPE)


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{}, {"OTGPU_RING_SLOTS": "2048"},
                                 {"OTGPU_RING_SLOTS": "2048", "OTGPU_WIDE_TWO_PASS": "1"},
                                 {"OTGPU_RING_SLOTS": "1024"}])
def test_gpu_scores_where_the_prune_fires(built, pgold, env):
    """Default tier choice; the 2048-slot shared ring (prune inside k_align<2048>, wider pairs
    continued on the warp's HBM ring); the same with the second kernel; the 1024-slot ring (every
    firing pair overflows into k_align<0>)."""
    import otter_b200
    (bases, off, lens), pairs = _batch(pgold)
    exp = np.array([g["score"] for g in pgold], dtype=np.int32)
    os.environ.update(env)
    try:
        with otter_b200.Context(0) as c:
            got = c.edit_scores(bases, off, lens, pairs)
            wide = c.stats()["n_pairs_wide"]
    finally:
        for k in env:
            del os.environ[k]
    assert np.array_equal(got, exp), np.nonzero(got != exp)
    if env.get("OTGPU_RING_SLOTS") == "1024":
        assert wide == len(pgold)
    if env.get("OTGPU_RING_SLOTS") == "2048":
        assert 0 < wide < len(pgold)  # some pairs ran their prune on the shared ring


@pytest.mark.gpu
def test_gpu_edit_cigars_where_the_prune_fires(ctx, pgold):
    sub = pgold[::3]
    (bases, off, lens), pairs = _batch(sub)
    scores, cigs = ctx.edit_cigars(bases, off, lens, pairs)
    for g, sc, cg in zip(sub, scores, cigs):
        assert int(sc) == g["cigar_score"] and len(cg) == g["cigar_len"]
        assert hashlib.md5(cg).hexdigest() == g["cigar_md5"]
