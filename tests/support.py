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
import sys
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PORT_SO = os.path.join(ROOT, "oracle", "libwfa_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so")
INTREE_SO = os.path.join(ROOT, "integration", "_build", "libotter_intree.so")

c_char_p, c_int, c_u32, c_u64, c_dbl = C.c_char_p, C.c_int, C.c_uint32, C.c_uint64, C.c_double
vp = C.c_void_p


class Counters(C.Structure):
    _fields_ = [("steps", C.c_int64), ("width_sum", C.c_int64), ("extend_bases", C.c_int64),
                ("extend_words", C.c_int64), ("max_width", C.c_int32), ("cutoffs", C.c_int32),
                ("prunes", C.c_int32), ("prune_gate", C.c_int32), ("width_hist", C.c_int64 * 8),
                ("run_hist", C.c_int64 * 8)]


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
        l.wfo_edit_cigar.restype = c_int
        l.wfo_edit_cigar.argtypes = [c_char_p, c_int, c_char_p, c_int, c_int, c_int, c_int, c_int,
                                     c_int, C.POINTER(c_int), c_char_p, c_int]
        _port = l
    return _port


def port_edit_cigar(p, t, ef=0, pb=0, pe=0, tb=0, te=0):
    """(score, CIGAR) of the C port's edit-distance aligner with CIGAR."""
    buf = C.create_string_buffer(len(p) + len(t) + 8)
    sc = c_int()
    n = port().wfo_edit_cigar(p, len(p), t, len(t), ef, pb, pe, tb, te, C.byref(sc), buf, len(buf))
    return sc.value, buf.raw[:max(n, 0)].decode()


def have_ref():
    return os.path.exists(REF_SO)


def _bind_ref(l):
    if True:
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
        l.ref_ppoa_consensus.restype = c_int
        l.ref_ppoa_consensus.argtypes = [vp, vp, vp, c_u32, c_char_p, c_int]
        l.ref_local_realignment.restype = c_u64
        l.ref_local_realignment.argtypes = [c_char_p, c_char_p, c_int, c_int, c_int, c_dbl, vp, vp, vp,
                                            vp, vp, c_u32, vp, c_u64, vp, vp]
        l.ref_assemble_region.restype = c_int
        l.ref_assemble_region.argtypes = [vp, vp, vp, vp, c_u32, c_int, vp, vp, c_u64, vp, vp, vp, vp]
    return l


def ref():
    """The unmodified reference (CPU)."""
    global _ref
    if _ref is None:
        _ref = _bind_ref(C.CDLL(REF_SO))
    return _ref


_intree = None


def have_intree():
    return os.path.exists(INTREE_SO)


def intree():
    """The reference pipeline (otterclust, anppoa, ...) running on the GPU hot path."""
    global _intree
    if _intree is None:
        _intree = _bind_ref(C.CDLL(INTREE_SO))
    return _intree


def assemble_region(lib, reads, span, max_alleles=2):
    """Run one region through src/assemble.cpp:103-142 (via integration/otter_shim.cpp) -> dict."""
    bases, off, lens = pack_reads(reads)
    n = len(reads)
    sp = np.array(span, dtype=np.uint8)
    labels = np.zeros(n, dtype=np.int32)
    cap = int(lens.sum()) * 4 + 1024
    seq = np.zeros(cap, dtype=np.uint8)
    seq_len = np.zeros(max(n, 4), dtype=np.uint32)
    cov = np.zeros(3 * max(n, 4), dtype=np.int32)
    se = np.zeros(max(n, 4), dtype=np.float32)
    ic = c_int(0)
    k = lib.ref_assemble_region(ptr(bases), ptr(off), ptr(lens), ptr(sp), n, max_alleles, ptr(labels),
                                ptr(seq), cap, ptr(seq_len), ptr(cov), ptr(se), C.byref(ic))
    alleles, pos = [], 0
    for a in range(k):
        alleles.append(seq[pos:pos + int(seq_len[a])].tobytes().decode())
        pos += int(seq_len[a])
    return dict(n_alleles=k, labels=labels.tolist(), alleles=alleles, cov=cov[:3 * k].tolist(),
                se=[float(x).hex() for x in se[:k]], ic=ic.value)


def assemble_regions(lib, regions, spans, max_alleles=2, threads=4):
    """Many regions through ref_assemble_regions (reference: per region on host threads; in-tree:
    one wave with the alignments batched on the GPU) -> (labels per region, n_alleles, digests)."""
    reads = [r for reg in regions for r in reg]
    bases, off, lens = pack_reads(reads)
    sp = np.array([x for s in spans for x in s], dtype=np.uint8)
    rs = np.zeros(len(regions) + 1, dtype=np.uint32)
    rs[1:] = np.cumsum([len(r) for r in regions])
    labels = np.zeros(len(reads), dtype=np.int32)
    n_al = np.zeros(len(regions), dtype=np.int32)
    dig = np.zeros(len(regions), dtype=np.uint64)
    lib.ref_assemble_regions.restype = c_dbl
    lib.ref_assemble_regions.argtypes = [vp, vp, vp, vp, vp, c_u32, c_int, c_int, vp, vp, vp]
    lib.ref_assemble_regions(ptr(bases), ptr(off), ptr(lens), ptr(sp), ptr(rs), len(regions), max_alleles,
                             threads, ptr(labels), ptr(n_al), ptr(dig))
    return [labels[rs[g]:rs[g + 1]].tolist() for g in range(len(regions))], n_al.tolist(), dig.tolist()


from harness.synth import (ALPHA, rand_seq, mutate, tr_region, pack_reads,  # noqa: E402,F401
                               region_batch, make_regions, pair_lengths)


def b(arr):
    return arr.tobytes() if isinstance(arr, np.ndarray) else bytes(arr)


def ptr(a):
    return a.ctypes.data_as(vp) if a is not None else None
