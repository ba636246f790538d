"""Harness code the bench and the GPU tests rely on (CPU only): the 100 k-region description with
per-region seeds, its partition by estimated work, and the synthetic BAM writer (run through the
UNMODIFIED `otter` CLI where the integration binaries are built)."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness import sharding, synth  # noqa: E402

BUILD = os.path.join(ROOT, "integration", "_build")


def test_c4s_regions_are_reproducible_in_any_order_and_from_any_process():
    a = synth.c4s_batch([5, 17, 3], procs=1)
    b = synth.c4s_batch([3, 5, 17, 40, 41, 42, 43, 44], procs=2)   # different order, worker processes
    assert a["n_regions"] == 3 and a["n_scores"] == 3 * 435
    ra = {g: a["bases"][int(a["off"][30 * i]):int(a["off"][30 * i]) + int(a["lens"][30 * i:30 * i + 30].sum())]
          for i, g in enumerate([5, 17, 3])}
    rb = {g: b["bases"][int(b["off"][30 * i]):int(b["off"][30 * i]) + int(b["lens"][30 * i:30 * i + 30].sum())]
          for i, g in enumerate([3, 5, 17, 40, 41, 42, 43, 44])}
    for g in (3, 5, 17):
        assert np.array_equal(ra[g], rb[g])
    # the header alone gives the target length the region was generated with
    reg = synth.c4s_region(17)
    tl = synth.c4s_target_len(17)
    assert 2000 <= tl <= 5000 and abs(np.median([len(r) for r in reg]) - tl) < 0.25 * tl


def test_c4s_partition_by_header_work_covers_and_balances():
    work = synth.c4s_work(4000)
    for world in (1, 2, 8):
        shards = sharding.partition_regions(work, world)
        assert np.array_equal(np.sort(np.concatenate(shards)), np.arange(4000))
        loads = np.array([work[s].sum() for s in shards])
        assert loads.max() <= 1.001 * loads.mean() + work.max()


@pytest.mark.skipif(not all(os.path.exists(os.path.join(BUILD, b)) for b in ("otter_stock", "sam2bam")),
                    reason="integration/_build binaries are built in the build container")
def test_synthetic_bam_runs_through_the_unmodified_cli():
    from harness import synth_bam
    d = tempfile.mkdtemp(prefix="otb200_bam_")
    data = synth_bam.write_dataset(d, 6, 20, 200, 500, 0.005, seed=9)
    assert data["n_records"] == 120 and os.path.exists(data["bam"] + ".bai")
    p = subprocess.run([os.path.join(BUILD, "otter_stock"), "assemble", "-b", data["bed"], "-R", "s", "--fasta",
                        "-t", "2", data["bam"]], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr[-400:]
    names = [l for l in p.stdout.split("\n") if l.startswith(">")]
    regions = {n.split("#")[1] for n in names}
    assert len(regions) == 6 and 6 <= len(names) <= 12   # one or two alleles per region
