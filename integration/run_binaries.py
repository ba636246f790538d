#!/usr/bin/env python
"""Run `otter assemble` twice on one synthetic BAM — the unmodified CLI (integration/_build/otter_stock)
and the same CLI on the B200 hot path (otter_b200: assemble_waves.patch + assemble_b200.cpp +
otter_b200/host) — and compare the allele records per region name.  Prints one JSON line.

  python integration/run_binaries.py [--regions 1000] [--reads 30] [--len 2000 5000] [--err 0.005]
                                     [--threads N] [--reference] [--sam]
Needs a GPU for the second run (there is no CPU fallback) and the prebuilt binaries
(`make -C integration binaries` in the build container; they travel with the tree)."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from harness import synth_bam  # noqa: E402

BUILD = os.path.join(ROOT, "integration", "_build")


def records(text, fasta):
    """name -> record body (everything but the name), for order-independent comparison."""
    out = {}
    if fasta:
        name = None
        for line in text.split("\n"):
            if line.startswith(">"):
                name = line
                out[name] = ""
            elif name is not None:
                out[name] += line
    else:
        for line in text.split("\n"):
            if line and not line.startswith("@"):
                f = line.split("\t")
                out[f[0]] = "\t".join(f[1:])
    return out


def run(binary, data, threads, fasta, reference, extra=()):
    cmd = [os.path.join(BUILD, binary), "assemble", "-b", data["bed"], "-R", "smp", "-t", str(threads)]
    if fasta:
        cmd.append("--fasta")
    if reference:
        cmd += ["-r", data["fasta"]]
    cmd += list(extra) + [data["bam"]]
    t0 = time.perf_counter()
    p = subprocess.run(cmd, capture_output=True, text=True)
    dt = time.perf_counter() - t0
    if p.returncode != 0:
        raise RuntimeError(f"{binary} failed ({p.returncode}): {p.stderr[-800:]}")
    return dt, p.stdout, p.stderr


def compare(data, threads, fasta=True, reference=False, extra=()):
    ts, so, se = run("otter_stock", data, threads, fasta, reference, extra)
    tb, bo, be = run("otter_b200", data, threads, fasta, reference, extra)
    rs, rb = records(so, fasta), records(bo, fasta)
    differing = sorted(k for k in set(rs) | set(rb) if rs.get(k) != rb.get(k))
    warn = lambda e: sorted(l.split("): ", 1)[1] for l in e.split("\n") if "[WARNING]" in l)  # noqa: E731
    return dict(stock_s=ts, b200_s=tb, records=len(rs), records_b200=len(rb), differing=differing[:5],
                n_differing=len(differing), identical=not differing and len(rs) == len(rb),
                same_warnings=warn(se) == warn(be))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--regions", type=int, default=1000)
    ap.add_argument("--reads", type=int, default=30)
    ap.add_argument("--len", type=int, nargs=2, default=[2000, 5000])
    ap.add_argument("--err", type=float, default=0.005)
    ap.add_argument("--threads", type=int, default=min(32, os.cpu_count() or 1))
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--sam", action="store_true")
    a = ap.parse_args()
    d = tempfile.mkdtemp(prefix="otb200_bam_")
    t0 = time.perf_counter()
    data = synth_bam.write_dataset(d, a.regions, a.reads, a.len[0], a.len[1], a.err, seed=0)
    res = compare(data, a.threads, fasta=not a.sam, reference=a.reference)
    res.update(workload=f"{a.regions} TR regions x {a.reads} reads, repeat of {a.len[0]}-{a.len[1]} bp + 1 kb flanks, "
                        f"error {a.err}; otter assemble -t {a.threads}" + (" -r ref.fa" if a.reference else ""),
               bam_records=data["n_records"], dataset_s=time.perf_counter() - t0 - res["stock_s"] - res["b200_s"],
               regions_per_s_stock=a.regions / res["stock_s"], regions_per_s_b200=a.regions / res["b200_s"],
               threads=a.threads)
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
