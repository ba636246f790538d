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


from otter_b200.synth import (ALPHA, rand_seq, mutate, tr_region, pack_reads,  # noqa: E402,F401
                               region_batch, make_regions, pair_lengths)


def b(arr):
    return arr.tobytes() if isinstance(arr, np.ndarray) else bytes(arr)


def ptr(a):
    return a.ctypes.data_as(vp) if a is not None else None
