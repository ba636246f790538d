"""Shared test support: seeded tandem-repeat read generators and ctypes handles.

Library roles:
  product  otter_b200/lib/libotter_b200.so      CUDA hot path behind include/otter_b200.h
  port     oracle/libwfa_oracle.so              our CPU restatement (checker only)
  ref      oracle/_ref/libotter_ref.so          the unmodified reference + shim (checker only;
                                                prebuilt in the build container, may be absent)
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT_SO = os.path.join(ROOT, "oracle", "libwfa_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so")

c_char_p, c_int, c_u32, c_u64, c_dbl = C.c_char_p, C.c_int, C.c_uint32, C.c_uint64, C.c_double
vp = C.c_void_p


class Counters(C.Structure):
    _fields_ = [("steps", C.c_int64), ("width_sum", C.c_int64), ("extend_bases", C.c_int64),
                ("extend_words", C.c_int64), ("max_width", C.c_int32), ("cutoffs", C.c_int32),
                ("prunes", C.c_int32)]


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            import subprocess
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
        l = C.CDLL(PORT_SO)
        l.wfo_edit_score.restype = c_int
        l.wfo_edit_score.argtypes = [c_char_p, c_int, c_char_p, c_int, c_int, c_int, c_int, c_int,
                                     c_int, C.POINTER(Counters)]
        l.wfo_align_anreads.restype = c_dbl
        l.wfo_align_anreads.argtypes = [c_char_p, c_u32, c_int, c_char_p, c_u32, c_int,
                                        C.POINTER(c_int)]
        l.wfo_fill_dist_matrix.restype = None
        l.wfo_fill_dist_matrix.argtypes = [vp, vp, vp, vp, vp, c_u32, vp]
        if hasattr(l, "wfo_affine_cigar"):
            l.wfo_affine_cigar.restype = c_int
            l.wfo_affine_cigar.argtypes = [c_char_p, c_int, c_char_p, c_int, c_int, c_int, c_int,
                                           c_int, c_int, c_int, c_int, c_int, C.POINTER(c_int),
                                           c_char_p, c_int]
        _port = l
    return _port


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        l = C.CDLL(REF_SO)
        l.ref_edit_e2e.restype = c_int
        l.ref_edit_e2e.argtypes = [c_char_p, c_int, c_char_p, c_int]
        l.ref_edit_endsfree.restype = c_int
        l.ref_edit_endsfree.argtypes = [c_char_p, c_int, c_int, c_int, c_char_p, c_int, c_int, c_int]
        l.ref_cigar.restype = c_int
        l.ref_cigar.argtypes = [c_int, c_int, c_char_p, c_int, c_int, c_int, c_char_p, c_int, c_int,
                                c_int, C.POINTER(c_int), C.POINTER(c_int), c_char_p, c_int]
        l.ref_align_anreads.restype = c_dbl
        l.ref_align_anreads.argtypes = [c_char_p, c_u32, c_int, c_char_p, c_u32, c_int]
        l.ref_fill_dist_matrix.restype = None
        l.ref_fill_dist_matrix.argtypes = [vp, vp, vp, vp, c_u32, vp, c_u32, vp]
        l.ref_fill_dist_matrix_regions.restype = c_dbl
        l.ref_fill_dist_matrix_regions.argtypes = [vp, vp, vp, vp, c_u32, vp, vp, c_int]
        l.ref_invalid_reassignment.restype = None
        l.ref_invalid_reassignment.argtypes = [vp, vp, vp, vp, c_u32, c_dbl, c_dbl, c_int, vp]
        _ref = l
    return _ref


# --------------------------------------------------------------------------------------------
# Synthetic tandem-repeat reads: flank + motif x copies + flank, two alleles,
# per-base sub/ins/del errors at rate e.
ALPHA = np.frombuffer(b"ACGT", dtype=np.uint8)


def rand_seq(rng, n):
    return ALPHA[rng.integers(0, 4, n)]


def mutate(rng, seq, e, n_rate=0.0):
    """Per-base errors at rate e split 1:1:1 sub/ins/del; optionally one 'N'."""
    n = len(seq)
    if n == 0 or e <= 0:
        out = seq.copy()
    else:
        r = rng.random(n)
        kind = np.zeros(n, dtype=np.int8)  # 0 keep, 1 sub, 2 ins (after base), 3 del
        kind[r < e] = rng.integers(1, 4, int((r < e).sum()))
        sub = kind == 1
        s2 = seq.copy()
        s2[sub] = ALPHA[(np.searchsorted(ALPHA, seq[sub]) + rng.integers(1, 4, int(sub.sum()))) % 4]
        reps = np.ones(n, dtype=np.int64)
        reps[kind == 3] = 0
        reps[kind == 2] = 2
        out = np.repeat(s2, reps)
        # inserted copy -> random base
        pos = np.cumsum(reps) - 1
        ins_pos = pos[kind == 2]
        out[ins_pos] = ALPHA[rng.integers(0, 4, len(ins_pos))]
    if n_rate > 0 and len(out) and rng.random() < n_rate:
        out[rng.integers(0, len(out))] = ord("N")
    return out


def tr_region(rng, n_reads, target_len, e, flank=100, extra_frac=0.2, n_rate=0.01, vntr_p=0.1):
    """One region: reads drawn from two alleles (allele1 = allele0 + up to extra_frac copies)."""
    mlen = int(rng.integers(20, 61)) if rng.random() < vntr_p else int(rng.integers(2, 7))
    motif = rand_seq(rng, mlen)
    copies0 = max(1, (target_len - 2 * flank) // mlen)
    copies1 = copies0 + int(rng.integers(0, int(copies0 * extra_frac) + 1))
    fl, fr = rand_seq(rng, flank), rand_seq(rng, flank)
    alleles = [np.concatenate([fl, np.tile(motif, c), fr]) for c in (copies0, copies1)]
    return [mutate(rng, alleles[int(rng.integers(0, 2))], e, n_rate) for _ in range(n_reads)]


def pack_reads(reads):
    """list of uint8 arrays -> (bases, off u64, len u32)."""
    lens = np.array([len(r) for r in reads], dtype=np.uint32)
    off = np.zeros(len(reads), dtype=np.uint64)
    if len(reads):
        off[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    bases = np.concatenate(reads).astype(np.uint8) if len(reads) and lens.sum() else np.zeros(0, np.uint8)
    return bases, off, lens


def region_batch(regions):
    """list of regions (each a list of reads) -> dict of flat arrays for the region ABI."""
    reads = [r for reg in regions for r in reg]
    bases, off, lens = pack_reads(reads)
    rs = np.zeros(len(regions) + 1, dtype=np.uint32)
    rs[1:] = np.cumsum([len(r) for r in regions])
    n = np.diff(rs).astype(np.uint64)
    npairs = n * (n - (n > 0)) // 2
    so = np.zeros(len(regions), dtype=np.uint64)
    if len(regions):
        so[1:] = np.cumsum(npairs[:-1])
    return dict(bases=bases, off=off, lens=lens, region_start=rs, score_off=so,
                n_scores=int(npairs.sum()), reads=reads)


def b(arr):
    return arr.tobytes() if isinstance(arr, np.ndarray) else bytes(arr)


def ptr(a):
    return a.ctypes.data_as(vp) if a is not None else None
