"""otgpu_edit_scores_static — the entry point that allocates nothing (no counterpart in the
reference): its scores against the existing
otgpu_edit_scores on the same inputs, through the C ABI, on two calls in a row that share the
static scratch array.  Collected last on purpose; passes alone in
a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import support as S

pytestmark = pytest.mark.gpu


def batch(seed, n_regions, n_reads, tl, e, n_rate):
    """reads of a few regions + every pair inside a region (end-to-end, longer read = pattern) and
    a few ends-free requests."""
    from otter_b200.capi import PAIR_DTYPE
    rng = np.random.default_rng(seed)
    reads, rows = [], []
    for _ in range(n_regions):
        reg = S.tr_region(rng, n_reads, tl, e, extra_frac=0.3, n_rate=n_rate)
        base = len(reads)
        reads += reg
        for i in range(len(reg)):
            for j in range(i + 1, len(reg)):
                a, b = (base + i, base + j) if len(reg[i]) >= len(reg[j]) else (base + j, base + i)
                d = len(reads[a]) - len(reads[b])
                rows.append((a, b, -1, 0, 0, 0) if (i + j) % 5 else (a, b, d // 2, d - d // 2, 0, 0))
    bases, off, lens = S.pack_reads(reads)
    return bases, off, lens, np.array(rows, dtype=PAIR_DTYPE)


def static_scores(lib, bases, off, lens, pairs, device=0):
    vp = C.c_void_p
    out = np.full(len(pairs), -12345, dtype=np.int32)
    st = lib.otgpu_edit_scores_static(device, None, bases.ctypes.data_as(vp), C.c_uint64(bases.size),
                                      off.ctypes.data_as(vp), lens.ctypes.data_as(vp), C.c_uint32(len(lens)),
                                      pairs.ctypes.data_as(vp), C.c_uint64(len(pairs)), out.ctypes.data_as(vp))
    return st, out


def test_static_entry_point_equals_the_existing_one_on_two_calls_in_a_row(built):
    import otter_b200
    lib = otter_b200.lib()
    # call 1: ONT-like reads, some with an N (8-bit path, read flags set in the shared array);
    # call 2: fewer, shorter, cleaner reads - whatever call 1 left behind must not matter;
    # call 3: unrelated reads whose wavefronts outgrow the 1024-slot ring (HBM rings in the array)
    rng = np.random.default_rng(3)
    wide_reads = [S.ALPHA[rng.integers(0, 4, n)] for n in (5200, 3100, 4700, 2500)]
    from otter_b200.capi import PAIR_DTYPE
    wb, wo, wl = S.pack_reads(wide_reads)
    wide_pairs = np.array([(0, 1, -1, 0, 0, 0), (0, 2, -1, 0, 0, 0), (2, 3, -1, 0, 0, 0), (0, 3, -1, 0, 0, 0)],
                          dtype=PAIR_DTYPE)
    calls = [batch(11, 6, 14, 2500, 0.05, 0.5), batch(12, 3, 9, 700, 0.005, 0.0), (wb, wo, wl, wide_pairs)]
    with otter_b200.Context(0) as ctx:
        for k, (bases, off, lens, pairs) in enumerate(calls):
            want = ctx.edit_scores(bases, off, lens, pairs)
            st, got = static_scores(lib, bases, off, lens, pairs)
            assert st == 0, lib.otgpu_strerror(st)
            assert np.array_equal(got, want), (k, np.nonzero(got != want)[0][:5])   # integer scores: bit-exact
            if k == 2:
                assert ctx.stats()["n_pairs_wide"] > 0   # the wide tier was exercised
        # and once more with the first batch: a repeat after other contents gives the same scores
        bases, off, lens, pairs = calls[0]
        st, again = static_scores(lib, bases, off, lens, pairs)
        assert st == 0 and np.array_equal(again, ctx.edit_scores(bases, off, lens, pairs))


def test_static_entry_point_rejects_what_exceeds_its_limits(built):
    import otter_b200
    lib = otter_b200.lib()
    from otter_b200.capi import PAIR_DTYPE
    one = np.array([(0, 1, -1, 0, 0, 0)], dtype=PAIR_DTYPE)
    # a read longer than OTGPU_STATIC_MAX_READ_LEN (8,192)
    reads = [S.ALPHA[np.zeros(9000, dtype=np.int64)], S.ALPHA[np.zeros(100, dtype=np.int64)]]
    bases, off, lens = S.pack_reads(reads)
    st, out = static_scores(lib, bases, off, lens, one)
    assert st == 6 and b"fixed limits" in lib.otgpu_strerror(st)
    assert (out == -12345).all()                               # nothing was written
    # more than OTGPU_STATIC_MAX_BASES (4 MiB) of reads
    reads = [S.ALPHA[np.zeros(8000, dtype=np.int64)]] * 600
    bases, off, lens = S.pack_reads(reads)
    assert static_scores(lib, bases, off, lens, one)[0] == 6
    # more than OTGPU_STATIC_MAX_PAIRS (131,072) pairs
    reads = [S.ALPHA[np.zeros(50, dtype=np.int64)]] * 2
    bases, off, lens = S.pack_reads(reads)
    many = np.repeat(one, 131073)
    assert static_scores(lib, bases, off, lens, many)[0] == 6
    # ordinary argument errors keep their own code; an empty pair list is fine
    bad = np.array([(0, 7, -1, 0, 0, 0)], dtype=PAIR_DTYPE)
    assert static_scores(lib, bases, off, lens, bad)[0] == 1
    assert static_scores(lib, bases, off, lens, one[:0])[0] == 0
