"""The boundary, proven with the real binary: `otter assemble` built twice from the reference —
as it is (otter_stock), and with integration/assemble_waves.patch + assemble_b200.cpp linked on
otter_b200/host and libotter_b200.so instead of src/analignments.cpp + WFA2-lib (otter_b200) —
run on the same synthetic BAM (written with the reference's own src/sam.c); allele records must be
identical per region name.  Binaries are prebuilt in the build container (integration/Makefile,
target `binaries`) and travel with the tree."""
import os
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "integration"))
BUILD = os.path.join(ROOT, "integration", "_build")
have = all(os.path.exists(os.path.join(BUILD, b)) for b in ("otter_stock", "otter_b200", "sam2bam"))
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not have, reason="integration/_build binaries are prebuilt in the build container")]


@pytest.fixture(scope="module")
def dataset():
    from harness import synth_bam
    d = tempfile.mkdtemp(prefix="otb200_bam_")
    return synth_bam.write_dataset(d, 220, 30, 300, 1500, 0.005, seed=3)


@pytest.fixture(scope="module")
def noisy_dataset():
    from harness import synth_bam
    d = tempfile.mkdtemp(prefix="otb200_bam_")
    return synth_bam.write_dataset(d, 60, 40, 300, 1200, 0.04, seed=4, clipped_frac=0.2)


def test_assemble_fasta_identical(dataset):
    import run_binaries as RB
    r = RB.compare(dataset, threads=8, fasta=True)
    assert r["records"] >= 220 and r["identical"], r
    assert r["same_warnings"], r


def test_assemble_sam_with_reference_identical(dataset):
    """-r ref.fa: local_realignment runs first (batched per wave on the GPU)."""
    import run_binaries as RB
    r = RB.compare(dataset, threads=8, fasta=False, reference=True)
    assert r["records"] >= 220 and r["identical"], r


def test_assemble_noisy_reads_and_small_waves_identical(noisy_dataset):
    import run_binaries as RB
    os.environ["OTTER_B200_WAVE"] = "16"   # several overlapped waves
    try:
        r = RB.compare(noisy_dataset, threads=4, fasta=True, reference=True)
    finally:
        del os.environ["OTTER_B200_WAVE"]
    assert r["records"] >= 60 and r["identical"], r


def test_reads_only_identical(noisy_dataset):
    import run_binaries as RB
    r = RB.compare(noisy_dataset, threads=4, fasta=True, reference=True, extra=("--reads-only",))
    assert r["records"] > 1000 and r["identical"], r


def test_max_alleles_one_identical(dataset):
    import run_binaries as RB
    r = RB.compare(dataset, threads=8, fasta=True, extra=("-a", "1"))
    assert r["identical"], r
