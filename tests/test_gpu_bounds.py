"""The alignment kernels issue their sequence-window loads without guards; the batch buffers carry
guard regions instead.  compute-sanitizer is not available on the GPU pool,
so the claim is checked from the inside: libotter_b200_check.so is the same library compiled with
-DOTGPU_CHECK_BOUNDS, in which every window address is compared with the two buffers it may fall
into.  A batch of adversarial shapes (one-base reads, N-only reads, 0-30 % error, ends-free pairs,
45 kb reads on the HBM ring) must produce the oracle's scores and not one out-of-bounds address."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECK = os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_check.so")

SCRIPT = r"""
import ctypes as C, json, os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
import support as S
import otter_b200 as ob
from otter_b200.capi import PAIR_DTYPE

rng = np.random.default_rng(77)
regions = []
for _ in range(120):
    tl = int(rng.choice([1, 20, 80, 300, 1000, 2500, 5000]))
    e = float(rng.choice([0.0, 0.01, 0.05, 0.12, 0.3]))
    n = int(rng.integers(2, 12))
    reads = S.tr_region(rng, n, max(tl, 8), e, extra_frac=float(rng.choice([0.0, 0.3, 0.6])),
                        n_rate=float(rng.choice([0, 0.2, 0.5])))
    if rng.random() < 0.2:
        reads[0] = np.frombuffer(b"N", dtype=np.uint8)
    if rng.random() < 0.2:
        reads[1] = reads[-1][:1].copy()
    regions.append(reads)
SHORT = os.environ.get("OTGPU_FIRST_WINDOW") == "1"  # the low-error flavour of k_align<1024>
if not SHORT:  # (one long read sends the whole batch to the 2,048-slot tier, which has one flavour)
    regions.append(S.tr_region(np.random.default_rng(26), 4, 45000, 0.05, extra_frac=0.1))  # HBM ring
else:
    regions.append(regions[3])
B = S.region_batch(regions)
lib = ob.lib()
out = {}
with ob.Context(0) as ctx:
    got = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"], B["n_scores"])
    out["wide"] = int(ctx.stats()["n_pairs_wide"]); out["bytes8"] = int(ctx.stats()["n_pairs_8bit"])
    # ends-free pairs through the pair ABI (invalid_reassignment's forms)
    reads = regions[5] + regions[40]
    bases, off, lens = S.pack_reads(reads)
    rows = []
    for i in range(len(reads)):
        for j in range(len(reads)):
            if i != j and len(reads[i]) >= len(reads[j]) and len(reads[j]) > 0:
                d = len(reads[i]) - len(reads[j])
                rows.append((i, j, d // 2, d - d // 2, 0, 0))
    pairs = np.zeros(len(rows), dtype=PAIR_DTYPE)
    for q, r in enumerate(rows):
        pairs[q] = r
    ef = ctx.edit_scores(bases, off, lens, pairs)
    loads, oob = C.c_ulonglong(0), C.c_ulonglong(0)
    lib.otgpu_check_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    assert lib.otgpu_check_counts(ctx._h, C.byref(loads), C.byref(oob)) == 0
    out["loads"], out["oob"] = loads.value, oob.value
P = S.port()
bad = 0
q = 0
for reg in regions[:-1]:
    for i in range(len(reg)):
        for j in range(i + 1, len(reg)):
            x, y = S.b(reg[i]), S.b(reg[j])
            if len(x) < len(y):
                x, y = y, x
            bad += int(got[q] != P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None))
            q += 1
for (i, j, pb, pe, tb, te), s in zip(rows, ef):
    x, y = S.b(reads[i]), S.b(reads[j])
    bad += int(s != P.wfo_edit_score(x, len(x), y, len(y), 1, pb, pe, tb, te, None))
out["bad"] = bad
out["pairs"] = int(q + len(rows))
print("RESULT " + json.dumps(out))
"""


@pytest.mark.parametrize("first_window", ["0", "1"])
def test_no_window_load_leaves_its_buffers(first_window):
    if not os.path.exists(CHECK):
        pytest.skip("checking library not built")
    env = dict(os.environ, OTTER_B200_LIB=CHECK, OTGPU_FIRST_WINDOW=first_window)
    r = subprocess.run([sys.executable, "-c", SCRIPT], cwd=ROOT, env=env, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    assert res["bad"] == 0, res
    assert res["loads"] > 1000000, res     # the check was compiled in and ran
    assert res["oob"] == 0, res
    assert res["bytes8"] > 0 and (res["wide"] > 0 or first_window == "1"), res
