"""CPU tests of the boundary: the C-ABI library loads, exports every symbol the header declares,
and fails loudly (no fallback) when there is no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "otter_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(otgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    import otter_b200
    from otter_b200 import capi
    l = otter_b200.lib()
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(l, s), f"{s} declared in include/otter_b200.h but not exported"
    assert sorted(capi.SYMBOLS) == syms, "otter_b200/capi.py binding list out of sync with the header"
    assert l.otgpu_abi_version() == 1


def test_library_is_built_for_sm_100a(built):
    import subprocess
    import otter_b200
    out = subprocess.run(["cuobjdump", "-lelf", otter_b200.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_host_layer_exports(built):
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    for s in ("othost_fill_dist_matrix", "othost_fill_dist_matrices", "othost_invalid_reassignment",
              "othost_align_anreads", "othost_have_device", "othost_fill_dist_matrices_multi",
              "othost_partition_regions"):
        assert hasattr(l, s)


def test_partition_regions_by_work(built):
    """fill_dist_matrices_multi's split of regions over devices: every region goes
    to exactly one share, the greedy heaviest-first rule is reproduced, shares are balanced."""
    import numpy as np
    l = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
    vp = C.c_void_p
    l.othost_partition_regions.argtypes = [vp, vp, C.c_uint32, C.c_int, vp, vp]
    rng = np.random.default_rng(5)
    n_reads = rng.integers(0, 40, 300).astype(np.uint32)
    rs = np.zeros(301, np.uint32)
    rs[1:] = np.cumsum(n_reads)
    lens = rng.integers(1, 6000, int(rs[-1])).astype(np.uint32)
    work = np.zeros(300)
    for g in range(300):
        x = lens[rs[g]:rs[g + 1]].astype(np.float64)
        work[g] = (x.sum() ** 2 - (x * x).sum()) / 2
    for parts in (1, 2, 8):
        part = np.full(300, -1, np.int32)
        loads = np.zeros(parts)
        l.othost_partition_regions(lens.ctypes.data, rs.ctypes.data, 300, parts, part.ctypes.data,
                                   loads.ctypes.data)
        exp_part = np.zeros(300, np.int32)
        exp_load = np.zeros(parts)
        if parts > 1:
            for g in sorted(range(300), key=lambda g: -work[g]):  # stable: ties keep region order
                p = int(np.argmin(exp_load))
                exp_part[g] = p
                exp_load[p] += work[g]
            assert np.allclose(loads, exp_load)
            assert loads.max() <= 1.05 * loads.mean()
        assert np.array_equal(part, exp_part)


def test_no_device_means_error_not_fallback(built):
    import otter_b200
    l = otter_b200.lib()
    if l.otgpu_device_count() > 0:
        pytest.skip("a device is present")
    h = C.c_void_p()
    assert l.otgpu_create(0, C.byref(h)) == 2  # OTGPU_ERR_NO_DEVICE
    assert not h.value
    assert b"no CPU fallback" in l.otgpu_strerror(2)
    with pytest.raises(otter_b200.OtgpuError):
        otter_b200.Context(0)
    p = C.c_void_p()
    assert l.otgpu_host_alloc(64, C.byref(p)) == 2
    # NULL context is an argument error everywhere, never a crash
    assert l.otgpu_batch_align(None) == 1
    assert l.otgpu_edit_scores(None, None, 0, None, None, 0, None, 0, None) == 1


def test_missing_library_fails_loudly(monkeypatch):
    from otter_b200 import capi
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setattr(capi, "LIB_PATH", "/nonexistent/libotter_b200.so")
    with pytest.raises(capi.OtgpuError, match="no CPU fallback"):
        capi.lib()


def test_product_never_touches_the_oracle():
    """The product path must not import, link or load anything under oracle/."""
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "otter_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"wfa_oracle|libotter_ref|oracle/|wfo_|ref_shim", txt):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad
    for lib in ("libotter_b200.so", "libotter_b200_host.so"):
        p = os.path.join(ROOT, "otter_b200", "lib", lib)
        if os.path.exists(p):
            data = open(p, "rb").read()
            assert b"wfo_edit_score" not in data and b"ref_edit_e2e" not in data


def test_distmatrix_index_matches_reference_layout():
    """src/andistmat.cpp:20 index == row-major condensed order used by the region ABI."""
    for n in (2, 3, 7, 30, 200):
        q = 0
        for i in range(n):
            for j in range(i + 1, n):
                assert ((2 * n - 3 - i) * i >> 1) + j - 1 == q
                assert i * (2 * n - i - 1) // 2 + (j - i - 1) == q  # k_expand_regions' formula
                q += 1


def test_static_entry_point_checks_its_limits_before_any_device_work(built):
    """otgpu_edit_scores_static rejects a batch beyond its fixed limits with OTGPU_ERR_LIMIT (6) and a
    message that names them, before it looks for a device - so this holds on a CPU-only box too."""
    import numpy as np
    import otter_b200
    from otter_b200.capi import PAIR_DTYPE
    l = otter_b200.lib()
    vp = C.c_void_p
    bases = np.full(9000 + 100, ord("A"), dtype=np.uint8)
    off = np.array([0, 9000], dtype=np.uint64)
    lens = np.array([9000, 100], dtype=np.uint32)      # first read longer than 8,192
    pairs = np.array([(0, 1, -1, 0, 0, 0)], dtype=PAIR_DTYPE)
    out = np.zeros(1, dtype=np.int32)
    st = l.otgpu_edit_scores_static(0, None, bases.ctypes.data_as(vp), C.c_uint64(bases.size),
                                    off.ctypes.data_as(vp), lens.ctypes.data_as(vp), C.c_uint32(2),
                                    pairs.ctypes.data_as(vp), C.c_uint64(1), out.ctypes.data_as(vp))
    assert st == 6
    msg = l.otgpu_strerror(st)
    assert b"fixed limits" in msg and b"8,192" in msg
