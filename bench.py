#!/usr/bin/env python
"""bench.py — read-pair alignments/s of the B200 hot path (fill_dist_matrix over many regions).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: every region's all-pairs edit-distance
alignment (WFA, WF-adaptive heuristic, bit-exact to the reference) for the named workload.
  value     pairs/s with the raw reads already resident in HBM (pack + pair expansion + alignment
            kernels), CUDA events on the library's stream, max over ranks.
  e2e       pairs/s through the C-ABI call a host would make (otgpu_region_edit_scores) with
            PINNED HOST buffers: H2D of reads + kernels + D2H of scores inside the timed region.
  e2e_host_layer  the same through the C++ seam otter itself calls (fill_dist_matrices from
            per-read std::strings): staging copy and score conversion included.
  roofline  dominant kernel k_align<1024> (k_align<2048> for long-read batches): algorithmic bytes
            (2-bit reads once per pair + 4 B score) / its CUDA-event duration, against
            MEASURED_PEAKS.json hbm_gbs; issue_slots only from an ncu capture of THIS csrc tree.
  cpu_baseline  the reference's own fill_dist_matrix (oracle/_ref) on all host threads over >= 20
            regions per thread, and identical_to_reference: the GPU scores of those regions,
            as dissimilarities, against its doubles bit for bit.
  per_config    (N=1) the same measurement for C1, C3 and one eighth of the 100 k-region description.
  c4_strong     north_star's target: ONE 100 k-region HiFi description partitioned by estimated
            work over the ranks, each rank's pass + a host-side gather on rank 0; with N>1 rank 0
            also runs the whole description alone, so speedup_vs_n1 comes from one run.
  single_process  (N>1, rank 0) one host process feeding all N devices (fill_dist_matrices_multi).
N>1 headline: one process per GPU (torchrun), each rank aligns its own seeded batch of the same
shape (weak scaling; regions are independent, no data-path collective); value = pairs of all
ranks / max-over-ranks time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from harness import synth  # noqa: E402

METRIC = "read-pair alignments/sec"
UNIT = "pairs/s"


def load_peaks():
    """HBM peak for the roofline: the driver-measured STREAM-style copy of MEASURED_PEAKS.json, else
    the fallback /opt/skills/guides/B200_PROFILING.md states (6.65 TB/s, an earlier measurement on
    this pool)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                 "100", "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def mark(self):
        return len(self.rows)

    def stop(self, start=0):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = self.rows[start:] or self.rows[-3:]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def workload_arrays(name, seed, n_regions=None):
    regions = synth.make_regions(name, seed, n_regions)
    B = synth.region_batch(regions)
    x, y = synth.pair_lengths(B)
    B["alg_bytes"] = int(((x + 3) // 4 + (y + 3) // 4 + 4).sum())
    B["cells"] = int((x * y).sum())
    del B["reads"]
    return B


# ------------------------------------------------------------------------------------------------
def cpu_reference(B, budget_s, threads=None):
    """Reference fill_dist_matrix on host threads over the first R regions (R sized to budget)."""
    import ctypes as C
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so")
    port_so = os.path.join(ROOT, "oracle", "libwfa_oracle.so")
    threads = threads or os.cpu_count() or 1
    rs, so = B["region_start"], B["score_off"]
    vp = C.c_void_p

    def p(a):
        return a.ctypes.data_as(vp)

    if os.path.exists(ref_so):
        kind = "reference"
        lib = C.CDLL(ref_so)
        lib.ref_fill_dist_matrix_regions.restype = C.c_double
        lib.ref_fill_dist_matrix_regions.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, vp, C.c_int]

        def run(R):
            n_sc = int(so[R - 1] + (rs[R] - rs[R - 1]) * (rs[R] - rs[R - 1] - 1) // 2)
            out = np.empty(n_sc, dtype=np.float64)
            t = lib.ref_fill_dist_matrix_regions(p(B["bases"]), p(B["off"]), p(B["lens"]), p(rs), R,
                                                 p(out), p(so), threads)
            return n_sc, t
    else:
        kind = "port"
        if not os.path.exists(port_so):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
        lib = C.CDLL(port_so)
        lib.wfo_fill_dist_matrix.restype = None
        lib.wfo_fill_dist_matrix.argtypes = [vp, vp, vp, vp, vp, C.c_uint32, vp]

        def run(R):
            from concurrent.futures import ThreadPoolExecutor
            n_sc = int(so[R - 1] + (rs[R] - rs[R - 1]) * (rs[R] - rs[R - 1] - 1) // 2)
            out = np.empty(n_sc, dtype=np.float64)

            def one(g):
                idx = np.arange(rs[g], rs[g + 1], dtype=np.int32)
                lib.wfo_fill_dist_matrix(p(B["bases"]), p(B["off"]), p(B["lens"]), None, p(idx),
                                         len(idx), p(out[int(so[g]):]))
            t0 = time.perf_counter()
            with ThreadPoolExecutor(threads) as ex:  # ctypes releases the GIL
                list(ex.map(one, range(R)))
            return n_sc, time.perf_counter() - t0

    n_regions = len(rs) - 1
    R = min(n_regions, max(threads, 8))
    n_sc, t = run(R)
    if t < budget_s * 0.6 and R < n_regions:
        R = int(min(n_regions, max(R + 1, R * budget_s / max(t, 1e-3))))
        R = max(threads, R - R % threads) if R >= threads else R
        R = min(R, n_regions)
        n_sc, t = run(R)
    return dict(value=n_sc / t, unit=UNIT, cores=threads, kind=kind, pairs=n_sc, seconds=t,
                sample=f"first {R} of {n_regions} regions ({n_sc} pairs) in {t:.2f} s; "
                       f"{'unmodified reference fill_dist_matrix, WFA2 -O3 -march=x86-64-v3' if kind == 'reference' else 'C port of the reference algorithm (oracle/wfa_oracle.c, -O2)'}, "
                       f"static contiguous region blocks over {threads} threads"), run, R


def work_counts(B, n_regions=2):
    """Algorithmic operations per pair (secondary): offset updates W and 32-base
    word comparisons E, counted by the instrumented C restatement (oracle/wfa_oracle.c) over the
    first regions of the workload.  Part of the cpu_baseline leg (the oracle is never on the
    measured path)."""
    import ctypes as C
    port_so = os.path.join(ROOT, "oracle", "libwfa_oracle.so")
    if not os.path.exists(port_so):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    lib = C.CDLL(port_so)

    class Counters(C.Structure):
        _fields_ = [("steps", C.c_int64), ("width_sum", C.c_int64), ("extend_bases", C.c_int64),
                    ("extend_words", C.c_int64), ("max_width", C.c_int32), ("cutoffs", C.c_int32),
                    ("prunes", C.c_int32), ("prune_gate", C.c_int32), ("width_hist", C.c_int64 * 8),
                ("run_hist", C.c_int64 * 8)]
    lib.wfo_edit_score.restype = C.c_int
    lib.wfo_edit_score.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int] + [C.c_int] * 5 + [C.c_void_p]
    rs, off, lens = B["region_start"], B["off"], B["lens"]
    raw = B["bases"].tobytes() if B["bases"].size < (1 << 28) else None
    tot = dict(pairs=0, steps=0, W=0, E=0, max_width=0)
    hist = [0] * 8
    runs = [0] * 8
    for g in range(min(n_regions, len(rs) - 1)):
        reads = []
        for r in range(int(rs[g]), int(rs[g + 1])):
            o, n = int(off[r]), int(lens[r])
            reads.append(raw[o:o + n] if raw is not None else B["bases"][o:o + n].tobytes())
        for i in range(len(reads)):
            for j in range(i + 1, len(reads)):
                x, y = reads[i], reads[j]
                if x == y:
                    tot["pairs"] += 1
                    continue
                if len(x) < len(y):  # pattern = the longer read, tie -> x (analignments.cpp:67-70)
                    x, y = y, x
                c = Counters()
                lib.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, C.byref(c))
                tot["pairs"] += 1
                tot["steps"] += c.steps
                tot["W"] += c.width_sum
                tot["E"] += c.extend_words
                tot["max_width"] = max(tot["max_width"], c.max_width)
                for b in range(8):
                    hist[b] += c.width_hist[b]
                    runs[b] += c.run_hist[b]
    n = max(tot["pairs"], 1)
    return {"sample": f"first {min(n_regions, len(rs) - 1)} regions ({tot['pairs']} pairs), instrumented C restatement",
            "score_steps_per_pair": tot["steps"] / n, "offset_updates_per_pair": tot["W"] / n,
            "word_compares_per_pair": tot["E"] / n, "max_width": tot["max_width"],
            "steps_by_width": dict(zip(("le30", "le61", "le125", "le189", "le253", "le509", "le1021",
                                        "wider"), hist)),
            "extensions_by_matched_bases": dict(zip(("lt31", "lt62", "lt93", "lt124", "lt248", "lt1116",
                                                     "lt4000", "more"), runs))}


def consensus_requests(B):
    """rapid_consensus-shaped requests (src/analignments.cpp:262-268): every read of a region
    against the region's first read (stand-in for the medoid); spanning reads, so end-to-end when
    the read is not longer, else ends-free on the text side."""
    rs, lens = B["region_start"], B["lens"].astype(np.int64)
    rows = []
    for g in range(len(rs) - 1):
        rep = int(rs[g])
        for r in range(rep + 1, int(rs[g + 1])):
            d = int(lens[rep] - lens[r])
            rows.append((rep, r, -1, 0, 0, 0) if d >= 0 else (rep, r, 0, 0, 0, -d))
    return np.array(rows, dtype=np.int32).reshape(-1, 6)


def run_consensus(args, rank, world, local_rank):
    """Secondary measurement: gap-affine (4,6,2) alignments with CIGAR, GPU e2e vs the reference."""
    import ctypes as C
    B = workload_arrays(args.workload, rank, args.regions or 200)
    req = consensus_requests(B)
    n = len(req)
    line = {"metric": "gap-affine CIGAR alignments/sec", "unit": "alignments/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "dtype": "int32",
            "data": "synthetic", "config": {"workload": f"{args.workload} reads vs region "
                                            f"representative ({n} alignments)",
                                            "regions": B["n_regions"]}}
    if args.impl == "reference":
        if rank != 0:
            return
        lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so"))
        lib.ref_affine_cigars_batch.restype = C.c_double
        threads = os.cpu_count() or 1
        clen = np.zeros(n, dtype=np.int32)
        vp = C.c_void_p
        t = 0.0
        for _ in range(args.steps):
            t += lib.ref_affine_cigars_batch(B["bases"].ctypes.data_as(vp), B["off"].ctypes.data_as(vp),
                                             B["lens"].ctypes.data_as(vp), req.ctypes.data_as(vp),
                                             C.c_uint64(n), threads, clen.ctypes.data_as(vp))
        line.update(impl="reference", value=n * args.steps / t, ms_per_step=1e3 * t / args.steps,
                    cpu_baseline={"kind": "reference", "cores": threads, "value": n * args.steps / t,
                                  "unit": "alignments/s", "sample": f"all {n} alignments"})
        print(json.dumps(line), flush=True)
        return
    import otter_b200
    from otter_b200.capi import PAIR_DTYPE
    pairs = np.zeros(n, dtype=PAIR_DTYPE)
    for i, f in enumerate(("pattern", "text", "pbf", "pef", "tbf", "tef")):
        pairs[f] = req[:, i]
    ctx = otter_b200.Context(local_rank)
    for _ in range(max(args.warmup, 1)):
        scores, clen, begin, packed = ctx.cigars_packed(B["bases"], B["off"], B["lens"], pairs)
    l0 = ctx.stats()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        scores, clen, begin, packed = ctx.cigars_packed(B["bases"], B["off"], B["lens"], pairs)
    t = time.perf_counter() - t0
    line.update(value=n * args.steps / t, ms_per_step=1e3 * t / args.steps,
                gpu_launches=int(ctx.stats()["kernel_launches"] - l0),
                e2e={"value": n * args.steps / t, "unit": "alignments/s",
                     "h2d_bytes_per_step": int(B["bases"].nbytes + pairs.nbytes),
                     "d2h_bytes_per_step": int(len(packed) + begin.nbytes + scores.nbytes + clen.nbytes)},
                retried=int(ctx.stats()["n_pairs_wide"]),
                note="host buffers -> C ABI (otgpu_affine_cigars_packed) -> CIGARs packed in pinned "
                     "host memory")
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()


def run_compare(args, rank, world, local_rank):
    """Secondary measurement (BASELINE.json configs[4]): all-pairs allele x allele
    edit distance with CIGAR length, as `otter compare`'s get_distances computes it
    (src/compare.cpp:50-66).  Per region a cohort of assembled alleles (two true alleles, consensus-
    grade error 0.1 %); every pair (longer, shorter) end-to-end."""
    import ctypes as C
    n_regions = args.regions or 16
    n_alleles = 200
    rng = np.random.default_rng(0xC5 + rank)
    reads, rows = [], []
    for _ in range(n_regions):
        tl = int(rng.integers(2000, 5001))
        reg = synth.tr_region(rng, n_alleles, tl, 0.001, n_rate=0.0)
        base = len(reads)
        reads += reg
        ln = np.array([len(r) for r in reg])
        i, j = np.triu_indices(n_alleles, 1)
        swap = ln[i] < ln[j]
        a = np.where(swap, j, i) + base
        b = np.where(swap, i, j) + base
        rows.append(np.stack([a, b, np.full_like(a, -1), 0 * a, 0 * a, 0 * a], axis=1))
    req = np.concatenate(rows).astype(np.int32)
    bases, off, lens = synth.pack_reads(reads)
    n = len(req)
    line = {"metric": "allele-pair edit-distance+CIGAR-length alignments/sec", "unit": "alignments/s",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"C5-shaped: {n_regions} regions x {n_alleles} alleles of 2-5 kb, "
                                   f"all pairs per region ({n} alignments)", "regions": n_regions}}
    if args.impl == "reference":
        if rank != 0:
            return
        lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so"))
        lib.ref_edit_cigars_batch.restype = C.c_double
        threads = os.cpu_count() or 1
        vp = C.c_void_p
        m = min(n, 40000)  # bounded sample
        clen = np.zeros(m, dtype=np.int32)
        t = 0.0
        for _ in range(args.steps):
            t += lib.ref_edit_cigars_batch(bases.ctypes.data_as(vp), off.ctypes.data_as(vp),
                                           lens.ctypes.data_as(vp), req[:m].ctypes.data_as(vp),
                                           C.c_uint64(m), threads, clen.ctypes.data_as(vp))
        line.update(impl="reference", value=m * args.steps / t, ms_per_step=1e3 * t / args.steps,
                    cpu_baseline={"kind": "reference", "cores": threads, "value": m * args.steps / t,
                                  "unit": "alignments/s", "sample": f"first {m} of {n} alignments"})
        print(json.dumps(line), flush=True)
        return
    import otter_b200
    from otter_b200.capi import PAIR_DTYPE
    pairs = np.zeros(n, dtype=PAIR_DTYPE)
    for i, f in enumerate(("pattern", "text", "pbf", "pef", "tbf", "tef")):
        pairs[f] = req[:, i]
    ctx = otter_b200.Context(local_rank)
    for _ in range(max(args.warmup, 1)):
        scores, clen = ctx.edit_cigar_lengths(bases, off, lens, pairs)
    l0 = ctx.stats()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        scores, clen = ctx.edit_cigar_lengths(bases, off, lens, pairs)
    t = time.perf_counter() - t0
    line.update(value=n * args.steps / t, ms_per_step=1e3 * t / args.steps,
                gpu_launches=int(ctx.stats()["kernel_launches"] - l0),
                e2e={"value": n * args.steps / t, "unit": "alignments/s",
                     "h2d_bytes_per_step": int(bases.nbytes + pairs.nbytes),
                     "d2h_bytes_per_step": int(scores.nbytes + clen.nbytes)},
                retried=int(ctx.stats()["n_pairs_wide"]),
                checksum=[int(scores.astype(np.int64).sum()), int(clen.astype(np.int64).sum())],
                note="host buffers -> otgpu_edit_cigars (scores + CIGAR lengths only) -> host, wall clock")
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()


# ------------------------------------------------------------------------------------------------
def extract_workload(n_regions, reads_per_region, seed):
    """Synthetic BAM records: per region (2-5 kb wide) long reads (region + 0.5-5 kb either side,
    10 % of them ending inside the region) with ONT-like CIGARs (an edit every ~20 bases)."""
    from otter_b200.capi import BAM_READ_DTYPE
    rng = np.random.default_rng(seed)
    n = n_regions * reads_per_region
    rec = np.zeros(n, dtype=BAM_READ_DTYPE)
    cig_parts, co, so, i = [], 0, 0, 0
    seq_lens = []
    for g in range(n_regions):
        start = 10000 + int(rng.integers(0, 1000000))
        end = start + int(rng.integers(2000, 5001))
        for _ in range(reads_per_region):
            pos = start - int(rng.integers(500, 5000))
            stop = end + int(rng.integers(500, 5000))
            if rng.random() < 0.1:
                stop = start + int(rng.integers(100, end - start))
            nseg = max(1, (stop - pos) // 20)
            m = rng.integers(1, 40, nseg).astype(np.uint32)
            kind = rng.integers(0, 3, nseg)           # 0: M only, 1: M + I, 2: M + D
            gap = rng.integers(1, 4, nseg).astype(np.uint32)
            words = np.empty(2 * nseg, dtype=np.uint32)
            words[0::2] = m << 4
            words[1::2] = np.where(kind == 1, gap << 4 | 1, np.where(kind == 2, gap << 4 | 2, 0 << 4 | 1))
            words = words[(words >> 4) > 0]
            if words[-1] & 0xf:
                words = words[:-1]
            ops, lens = words & 0xf, words >> 4
            l_qseq = int(lens[(ops == 0) | (ops == 1)].sum())
            rec[i] = (pos, l_qseq, len(words), 0, co, so, start, end)
            cig_parts.append(words)
            co += len(words)
            so += (l_qseq + 1) // 2
            seq_lens.append(l_qseq)
            i += 1
    codes = np.array([1, 2, 4, 8], dtype=np.uint8)
    seq4 = (codes[rng.integers(0, 4, so)] << 4 | codes[rng.integers(0, 4, so)]).astype(np.uint8)
    return rec, np.concatenate(cig_parts), seq4


def run_extract(args, rank, world, local_rank):
    """Secondary measurement: get_breakpoints + parse_alignment per BAM record
    (src/anseqs.cpp:286-436), host records in, extracted reads out."""
    import ctypes as C
    if rank != 0:
        return
    n_regions = args.regions or 200
    rec, cig, seq4 = extract_workload(n_regions, 60, 0xE7)
    n = len(rec)
    line = {"metric": "BAM records extracted/sec", "unit": "records/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"{n_regions} regions x 60 long reads (region + 0.5-5 kb flanks), "
                                   f"ONT-like CIGARs", "records": n, "cigar_ops": int(len(cig)),
                       "query_bases": int(rec["l_qseq"].sum())}}
    flat = np.zeros((n, 8), dtype=np.int64)
    for k, f in enumerate(("pos", "l_qseq", "n_cigar", "cigar_off", "seq_off")):
        flat[:, k] = rec[f]
    flat[:, 6], flat[:, 7] = rec["start"], rec["end"]
    vp = C.c_void_p

    def reference(threads):
        lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libotter_ref.so"))
        lib.ref_parse_alignments_batch.restype = C.c_double
        lib.ref_parse_alignments_batch.argtypes = [vp, C.c_uint32, vp, vp, C.c_int, vp]
        out = np.zeros(n, dtype=np.int32)
        t = lib.ref_parse_alignments_batch(flat.ctypes.data_as(vp), n, cig.ctypes.data_as(vp),
                                           seq4.ctypes.data_as(vp), threads, out.ctypes.data_as(vp))
        return t, out

    threads = os.cpu_count() or 1
    if args.impl == "reference":
        t = sum(reference(threads)[0] for _ in range(args.steps))
        line.update(impl="reference", value=n * args.steps / t, ms_per_step=1e3 * t / args.steps,
                    cpu_baseline={"kind": "reference", "cores": threads, "value": n * args.steps / t,
                                  "unit": "records/s", "sample": f"all {n} records"})
        print(json.dumps(line), flush=True)
        return
    import otter_b200
    from otter_b200.capi import PinnedArray
    ctx = otter_b200.Context(local_rank)
    # pinned host staging: the buffers a host integration fills while iterating the BAM
    # (ReadExtractor::add copies each record's CIGAR and 4-bit bases out of htslib's buffer anyway)
    pins = []
    for a in (rec, cig, seq4):
        pa = PinnedArray(a.shape, a.dtype)
        pa.array[...] = a
        pins.append(pa)
    rec, cig, seq4 = (pa.array for pa in pins)
    for _ in range(max(args.warmup, 1)):
        status, coords, lens, begin, bases = ctx.extract_reads(rec, cig, seq4)
    l0 = ctx.stats()["kernel_launches"]
    t0 = time.perf_counter()
    for _ in range(args.steps):
        status, coords, lens, begin, bases = ctx.extract_reads(rec, cig, seq4)
    t = time.perf_counter() - t0
    line.update(value=n * args.steps / t, ms_per_step=1e3 * t / args.steps,
                gpu_launches=int(ctx.stats()["kernel_launches"] - l0),
                extracted_bases=int(begin[n]),
                e2e={"value": n * args.steps / t, "unit": "records/s",
                     "h2d_bytes_per_step": int(rec.nbytes + cig.nbytes + seq4.nbytes),
                     "d2h_bytes_per_step": int(len(bases) + begin.nbytes + status.nbytes + coords.nbytes + lens.nbytes)})
    # ---- the hand-off to the alignment: every successfully extracted record of a region against
    # every other one, (a) through the host (the extracted characters come back and go up again),
    # (b) device to device (otgpu_region_edit_scores_extracted: only the tables cross PCIe) ----
    okr = np.nonzero(status & 1)[0].astype(np.uint32)
    per_region = np.bincount(okr // 60, minlength=n_regions).astype(np.uint64)
    rs = np.zeros(n_regions + 1, dtype=np.uint32)
    rs[1:] = np.cumsum(per_region)
    npairs = per_region * (per_region - (per_region > 0)) // 2
    so = np.zeros(n_regions, dtype=np.uint64)
    so[1:] = np.cumsum(npairs[:-1])
    n_sc = int(npairs.sum())
    off_sel, len_sel = begin[:n][okr].copy(), lens[okr].copy()

    def via_host():
        st_, co_, le_, be_, ba_ = ctx.extract_reads(rec, cig, seq4)
        return ctx.region_edit_scores(ba_, off_sel, len_sel, rs, so, n_sc)

    def on_device():
        ctx.extract_reads(rec, cig, seq4)
        return ctx.region_edit_scores_extracted(okr, rs, so, n_sc)

    chain = {}
    for name, fn in (("via_host", via_host), ("device_resident", on_device)):
        sc = fn()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sc = fn()
        chain[name] = {"ms_per_step": 1e3 * (time.perf_counter() - t0) / args.steps, "checksum": int(sc.astype(np.int64).sum())}
    chain["identical"] = chain["via_host"]["checksum"] == chain["device_resident"]["checksum"]
    chain["pairs"] = n_sc
    chain["note"] = "extract + all-pairs edit distance per region; device_resident skips the H2D of the extracted bases"
    line["extract_then_align"] = chain
    if not args.no_cpu_baseline:
        tr, out = reference(threads)
        mine = np.where(status & 1, lens.astype(np.int64), -1)
        line["cpu_baseline"] = {"kind": "reference", "cores": threads, "value": n / tr, "unit": "records/s",
                                "sample": f"all {n} records, parse_alignment only"}
        line["identical_to_reference"] = bool(np.array_equal(mine, out))
    print(json.dumps(line), flush=True)
    ctx.close()


def run_pipeline(args, rank, world, local_rank):
    """Integration measurement: assemble_process's loop body (src/assemble.cpp:103-142: distance
    matrix -> otter_hclust -> re-assignment -> rapid_consensus) over a wave of regions, host reads
    in, alleles out.  `ours` = the unmodified reference clustering / partial-order-consensus
    objects linked on the product's host layer (integration/_build/libotter_intree.so, built by
    integration/Makefile from /root/reference + otter_b200/host): every alignment runs batched on the
    GPU, everything else on the host threads.  `reference` = the unmodified reference
    (libotter_ref.so), one region at a time per host thread.  Both return per-region labels and
    allele digests; the ours arm checks them against the reference on a sample."""
    import ctypes as C
    if rank != 0:
        return
    B = workload_arrays(args.workload, 0, args.regions or 200)
    n_regions = B["n_regions"]
    threads = os.cpu_count() or 1
    vp = C.c_void_p

    def bind(name):
        where = ("integration", "_build") if "intree" in name else ("oracle", "_ref")
        lib = C.CDLL(os.path.join(ROOT, *where, name))
        lib.ref_assemble_regions.restype = C.c_double
        lib.ref_assemble_regions.argtypes = [vp, vp, vp, vp, vp, C.c_uint32, C.c_int, C.c_int, vp, vp, vp]
        return lib

    def run(lib, R):
        labels = np.zeros(int(B["region_start"][R]), dtype=np.int32)
        n_al = np.zeros(R, dtype=np.int32)
        dig = np.zeros(R, dtype=np.uint64)
        p = lambda a: a.ctypes.data_as(vp)  # noqa: E731
        t = lib.ref_assemble_regions(p(B["bases"]), p(B["off"]), p(B["lens"]), None, p(B["region_start"]),
                                     R, 2, threads, p(labels), p(n_al), p(dig))
        return t, labels, n_al, dig

    line = {"metric": "regions assembled/sec", "unit": "regions/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: distance matrix + clustering + re-assignment + "
                                   f"consensus per region, max_alleles 2", "regions": n_regions,
                       "host_threads": threads}}
    if args.impl == "reference":
        lib = bind("libotter_ref.so")
        R = min(n_regions, max(threads * 2, 32))
        t = sum(run(lib, R)[0] for _ in range(args.steps))
        line.update(impl="reference", value=R * args.steps / t, ms_per_step=1e3 * t / args.steps,
                    cpu_baseline={"kind": "reference", "cores": threads, "value": R * args.steps / t,
                                  "unit": "regions/s", "sample": f"first {R} of {n_regions} regions"})
        print(json.dumps(line), flush=True)
        return
    lib = bind("libotter_intree.so")
    for _ in range(max(args.warmup, 1)):
        run(lib, n_regions)
    t, res = 0.0, None
    for _ in range(args.steps):
        r = run(lib, n_regions)
        t += r[0]
        res = r[1:]
    line.update(value=n_regions * args.steps / t, ms_per_step=1e3 * t / args.steps,
                e2e={"value": n_regions * args.steps / t, "unit": "regions/s",
                     "h2d_bytes_per_step": int(B["bases"].nbytes), "d2h_bytes_per_step": None},
                note="host reads -> alleles; GPU: all alignments (batched per wave); host threads: the "
                     "clustering logic + hclust-cpp (densities from the GPU), partial-order consensus (ppoa_b200.hpp)")
    if not args.no_cpu_baseline:
        R = min(n_regions, max(threads * 2, 32))
        tr, labels, n_al, dig = run(bind("libotter_ref.so"), R)
        same = bool(np.array_equal(labels, res[0][:len(labels)]) and np.array_equal(n_al, res[1][:R])
                    and np.array_equal(dig, res[2][:R]))
        line["cpu_baseline"] = {"kind": "reference", "cores": threads, "value": R / tr, "unit": "regions/s",
                                "sample": f"first {R} of {n_regions} regions in {tr:.2f} s"}
        line["identical_to_reference"] = same
    print(json.dumps(line), flush=True)


def csrc_sha():
    """Hash of the kernel sources: ncu-derived constants (traffic, warp instructions) are only
    quoted when they were captured on this very tree."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "otter_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".inc")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def load_capture(workload, n_regions):
    """The committed ncu capture of the dominant kernel for this workload (profiles/*_traffic.json:
    DRAM bytes and warp instructions per launch), or (None, reason)."""
    best = None
    for f in sorted(os.listdir(os.path.join(ROOT, "profiles"))):
        if f.endswith("_traffic.json"):
            try:
                j = json.load(open(os.path.join(ROOT, "profiles", f)))
            except Exception:
                continue
            if j.get("workload") == workload and j.get("n_regions") == n_regions:
                best = (f, j)
    if best is None:
        return None, "no capture of this workload"
    f, j = best
    if j.get("csrc_sha") != csrc_sha():
        return None, f"stale capture (profiles/{f} was taken on csrc {j.get('csrc_sha')}, this tree is {csrc_sha()})"
    j["file"] = f
    return j, None


def reference_doubles(B, R, threads=None, so="libotter_ref.so"):
    """The unmodified reference's dissimilarities of the first R regions (oracle/_ref), or None."""
    import ctypes as C
    ref_so = os.path.join(ROOT, "oracle", "_ref", so)
    if not os.path.exists(ref_so) or R < 1:
        return None, 0.0
    lib = C.CDLL(ref_so)
    lib.ref_fill_dist_matrix_regions.restype = C.c_double
    vp = C.c_void_p
    lib.ref_fill_dist_matrix_regions.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, vp, C.c_int]
    rs, so = B["region_start"], B["score_off"]
    n_sc = int(so[R - 1] + (rs[R] - rs[R - 1]) * (rs[R] - rs[R - 1] - 1) // 2)
    out = np.empty(n_sc, dtype=np.float64)
    p = lambda a: a.ctypes.data_as(vp)  # noqa: E731
    t = lib.ref_fill_dist_matrix_regions(p(B["bases"]), p(B["off"]), p(B["lens"]), p(rs), R, p(out), p(so),
                                         threads or os.cpu_count() or 1)
    return out, t


def scores_to_doubles(B, scores, n_sc):
    """score / (double)largest, as fill_dist_matrix forms it (src/analignments.cpp:68-72)."""
    lens = B["lens"].astype(np.float64)
    rs = B["region_start"]
    out = np.empty(n_sc, dtype=np.float64)
    q = 0
    g = 0
    while q < n_sc:
        L = lens[rs[g]:rs[g + 1]]
        i, j = np.triu_indices(len(L), 1)
        k = len(i)
        out[q:q + k] = scores[q:q + k] / np.maximum(L[i], L[j])
        q += k
        g += 1
    return out


def cpu_variants(B, R, threads, want):
    """The same reference sources under the other two builds BASELINE.md lists, on the first R
    regions: WFA2-lib as otter's README builds it (-Wall -g, no -O), and -O3 -march=x86-64-v4 where
    the host has AVX-512 (what -march=native could add over the -march=x86-64-v3 library that
    travels).  Results must equal the main build's."""
    out = {}
    try:
        flags = open("/proc/cpuinfo").read()
    except OSError:
        flags = ""
    avx512 = all(f in flags for f in ("avx512f", "avx512bw", "avx512vl", "avx512dq"))
    for key, so, ok in (("readme_build_-g", "libotter_ref_readme.so", True),
                        ("O3_march_x86-64-v4", "libotter_ref_v4.so", avx512)):
        if not ok:
            out[key] = "host has no AVX-512"
            continue
        got, t = reference_doubles(B, R, threads, so)
        if got is None:
            out[key] = "library not built"
            continue
        out[key] = {"value": len(got) / t, "unit": UNIT, "regions": R, "seconds": t,
                    "same_results": bool(np.array_equal(got.view(np.uint64), want[:len(got)].view(np.uint64)))}
    g3, t = reference_doubles(B, R, threads)
    out["O3_march_x86-64-v3_same_sample"] = {"value": len(g3) / t, "unit": UNIT, "regions": R, "seconds": t}
    return out


def cpu_baseline_leg(B, scores, regions_per_thread=20, min_seconds=8.0, with_variants=False):
    """The reference's fill_dist_matrix on all host threads over the first R regions, R >= 20 per
    thread (static contiguous blocks: fewer and the slowest block sets the number), and the check
    of the GPU scores of those regions against its doubles."""
    threads = os.cpu_count() or 1
    n_regions = B["n_regions"]
    base, run, _ = cpu_reference(B, 2.0, threads)           # short probe: rate of this workload
    R = min(n_regions, regions_per_thread * threads)
    pairs_per_region = max(B["n_scores"] / max(n_regions, 1), 1.0)
    R = int(min(n_regions, max(R, base["value"] * min_seconds / pairs_per_region)))
    R = max(1, R - R % threads) if R >= threads else R
    want, t = reference_doubles(B, R, threads)
    if want is not None and t < 4.0:  # a short sample is noisy: the reference keeps its better pass
        t = min(t, reference_doubles(B, R, threads)[1])
    if want is None:                                          # reference library absent: the C port
        n_sc, t = run(R)
        return dict(value=n_sc / t, unit=UNIT, cores=threads, kind=base["kind"],
                    sample=f"first {R} of {n_regions} regions ({n_sc} pairs) in {t:.2f} s; " + base["sample"].split("; ", 1)[1]), None
    n_sc = len(want)
    same = None
    if scores is not None:
        mine = scores_to_doubles(B, scores, n_sc)
        same = bool(np.array_equal(mine.view(np.uint64), want.view(np.uint64)))
    variants = None
    if with_variants:
        variants = cpu_variants(B, max(threads, min(R, 8 * threads)), threads, want)
    return dict(value=n_sc / t, unit=UNIT, cores=threads, kind="reference", other_builds=variants,
                sample=f"first {R} of {n_regions} regions = {R // threads if R >= threads else 0} per thread "
                       f"({n_sc} pairs) in {t:.2f} s; unmodified reference fill_dist_matrix, WFA2 -O3 "
                       f"-march=x86-64-v3 (built in the CPU-only container: no AVX-512, which -march=native "
                       f"on this host might add), static contiguous region blocks over {threads} threads"), same


def run_reference(args, rank, world):
    if rank != 0:
        return
    B = workload_arrays(args.workload, 0, args.regions)
    threads = os.cpu_count() or 1
    base, run, _ = cpu_reference(B, 2.0, threads)
    pairs_per_region = max(B["n_scores"] / max(B["n_regions"], 1), 1.0)
    R = int(min(B["n_regions"], max(20 * threads, base["value"] * 8.0 / pairs_per_region)))
    R = max(1, R - R % threads) if R >= threads else R
    for _ in range(max(args.warmup, 0)):
        run(min(R, max(threads, R // 4)))
    tot_p, tot_t = 0, 0.0
    for _ in range(args.steps):
        n_sc, t = run(R)
        tot_p += n_sc
        tot_t += t
    v = tot_p / tot_t
    base.update(value=v, sample=f"first {R} of {B['n_regions']} regions = {R // threads if R >= threads else 0} per "
                                f"thread, {args.steps} passes; " + base["sample"].split("; ", 1)[1])
    base.pop("pairs"), base.pop("seconds")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": config_dict(args.workload, B, world),
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def config_dict(workload, B, world):
    c = synth.CONFIGS[workload]
    return {"workload": f"{workload}: {c['desc']}", "regions_per_gpu": B["n_regions"],
            "reads_per_region": c["n_reads"], "pairs_per_gpu": B["n_scores"],
            "bases_per_gpu": int(B["bases"].size), "parallelism": f"region-sharded x{world}",
            "l2_policy": "inputs_exceed_l2" if B["bases"].size > 126e6 else "l2_flushed_between_steps"}


class Bench:
    """One rank's device, context and timing helpers."""

    def __init__(self, rank, world, local_rank):
        import torch
        import torch.distributed as dist
        import otter_b200
        self.torch, self.dist, self.rank, self.world, self.local_rank = torch, dist, rank, world, local_rank
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: otter_b200 has no CPU fallback")
        torch.cuda.set_device(local_rank)
        self.gloo = None
        if world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.gloo = dist.new_group(backend="gloo")  # host-side gathers only
        self.ctx = otter_b200.Context(local_rank)
        self.ext = torch.cuda.ExternalStream(self.ctx.stream, device=local_rank)
        self.flush = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, vals):
        """(max over ranks, sum over ranks) of a list of floats."""
        torch = self.torch
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if self.world == 1:
            return t.tolist(), t.tolist()
        tmax, tsum = t.clone(), t.clone()
        self.dist.all_reduce(tmax, op=self.dist.ReduceOp.MAX)
        self.dist.all_reduce(tsum, op=self.dist.ReduceOp.SUM)
        return tmax.tolist(), tsum.tolist()

    def pin(self, B):
        from otter_b200.capi import PinnedArray
        pin = {}
        for k in ("bases", "off", "lens", "region_start", "score_off"):
            pa = PinnedArray(B[k].shape, B[k].dtype)
            pa.array[...] = B[k]
            pin[k] = pa
        pin["out"] = PinnedArray((B["n_scores"],), np.int32)
        return pin

    @staticmethod
    def unpin(pin):
        for pa in pin.values():
            pa.free()

    def e2e_step(self, B, pin):
        c = self.ctx
        c._ck(c._l.otgpu_region_edit_scores(
            c._h, pin["bases"].ptr, B["bases"].size, pin["off"].ptr, pin["lens"].ptr, len(B["lens"]),
            pin["region_start"].ptr, B["n_regions"], pin["score_off"].ptr, B["n_scores"], pin["out"].ptr))

    def measure(self, B, steps, warmup, sampler=None):
        """value (device-resident, CUDA events on the library's stream) and e2e (pinned host buffers
        -> C ABI -> host scores) of one workload on this rank; max / sum over ranks."""
        torch = self.torch
        ctx, ext = self.ctx, self.ext
        pin = self.pin(B)
        small = B["bases"].size <= 126e6
        if small and self.flush is None:
            self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        flush = self.flush if small else None
        for _ in range(warmup):
            if flush is not None:  # loads torch's fill kernel outside the timed region
                with torch.cuda.stream(ext):
                    flush.fill_(1)
            self.e2e_step(B, pin)
        first = pin["out"].array.copy()
        launches0 = ctx.stats()["kernel_launches"]
        mark = sampler.mark() if sampler else 0
        # ---- value: inputs resident in HBM, device time ----
        ctx.upload_regions(pin["bases"].array, pin["off"].array, pin["lens"].array,
                           pin["region_start"].array, pin["score_off"].array, B["n_scores"])
        ctx.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stage = dict(pack=0.0, expand=0.0, align=0.0, wide=0.0)
        self.barrier()
        ev0.record(ext)
        for _ in range(steps):
            if flush is not None:
                with torch.cuda.stream(ext):
                    flush.fill_(1)
            ctx.align()
            ms = ctx.last_align_ms()  # waits for this step; per-stage CUDA events on the same stream
            for k in stage:
                stage[k] += ms[k]
        ev1.record(ext)
        self.barrier()
        dev_ms = ev0.elapsed_time(ev1)
        ctx.download(pin["out"].array)
        same = bool(np.array_equal(first, pin["out"].array))
        launches = ctx.stats()["kernel_launches"] - launches0
        # ---- e2e: host buffers -> C ABI -> host scores, wall clock bracketed by sync ----
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            self.e2e_step(B, pin)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        self.barrier()
        clocks = sampler.stop(mark) if sampler else None
        stats = ctx.stats()
        scores = pin["out"].array.copy()
        self.unpin(pin)
        (dev_ms, e2e_ms), _ = self.reduce([dev_ms, e2e_s * 1e3])
        _, (tot_pairs, tot_cells) = self.reduce([float(B["n_scores"]), float(B["cells"])])
        k_ms = stage["align"] / steps
        peak, peak_src = load_peaks()
        achieved = B["alg_bytes"] / (k_ms * 1e-3) / 1e9
        h2d = int(sum(B[k].nbytes for k in ("bases", "off", "lens", "region_start", "score_off")))
        long_ring = B["lens"].max() > 8000
        res = {
            "value": tot_pairs * steps / (dev_ms * 1e-3), "ms_per_step": dev_ms / steps,
            "gcups_equiv": tot_cells * steps / (dev_ms * 1e-3) / 1e9,
            "e2e": {"value": tot_pairs * steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": int(B["n_scores"] * 4), "ms_per_step": e2e_ms / steps,
                    "gcups_equiv": tot_cells * steps / (e2e_ms * 1e-3) / 1e9},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm",
                         "kernel": "k_align<2048>" if long_ring else
                                   ("k_align<1024, 1>" if ctx.first_window() else "k_align<1024, 0>"),
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None,
                         "alg_bytes_per_launch": B["alg_bytes"], "kernel_ms": k_ms,
                         "note": "latency/issue-bound integer DP: HBM fraction is expected to be "
                                 "tiny; see profiles/ for issue-slot and stall evidence"},
            "stage_ms_per_step": {k: v / steps for k, v in stage.items()},
            "pairs_8bit": stats["n_pairs_8bit"], "pairs_wide": stats["n_pairs_wide"],
            "deterministic": same,
        }
        if clocks is not None:
            res["clocks"] = clocks
        return res, scores

    # ------------------------------------------------------------------------------------------
    def host_layer(self, B, devices, warmup, reps):
        """The seam otter calls: fill_dist_matrices(_multi) from per-read std::strings that exist
        before the clock starts (othost_bench_fill); staging copy, device call and score conversion
        inside the timed region."""
        import ctypes as C
        H = C.CDLL(os.path.join(ROOT, "otter_b200", "lib", "libotter_b200_host.so"))
        vp = C.c_void_p
        H.othost_bench_fill.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp]
        H.othost_bench_fill.restype = None
        dev = np.array(devices, dtype=np.int32)
        out = np.zeros(B["n_scores"], dtype=np.float64)
        times = np.zeros((reps, 6), dtype=np.float64)
        p = lambda a: a.ctypes.data_as(vp)  # noqa: E731
        H.othost_bench_fill(p(B["bases"]), p(B["off"]), p(B["lens"]), p(B["region_start"]), B["n_regions"],
                            p(dev), len(devices), warmup, reps, p(out), p(B["score_off"]), p(times))
        m = times.mean(axis=0)
        return {"value": B["n_scores"] / m[0], "unit": UNIT, "ms_per_call": 1e3 * m[0],
                "stage_ms": {"share_total_max": 1e3 * m[1], "device_call_max": 1e3 * m[2],
                             "wait_for_staging_copy_max": 1e3 * m[3], "score_conversion_max": 1e3 * m[4]},
                "chunks": int(m[5]), "devices": list(map(int, devices)), "calls": reps,
                "note": "fill_dist_matrices from std::string reads: pinned staging copy by helper threads "
                        "and score -> dissimilarity conversion overlap the kernels chunk by chunk "
                        "(otgpu_region_edit_scores_hooked)"}, out


def c4_strong(bench, args):
    """north_star's target configuration: ONE 100 k-region HiFi description (harness/synth.py C4S),
    partitioned region-wise by estimated work over the ranks, each rank generates and aligns its
    shard end to end from pinned host buffers, rank 0 gathers the scores on the host (gloo).  Timed
    region = every rank's pass + the gather, max over ranks.  With more than one rank, rank 0 then
    runs the SAME description on its one GPU so that the speed-up is measured inside one run."""
    from harness import sharding
    torch, dist = bench.torch, bench.dist
    rank, world = bench.rank, bench.world
    n_regions = args.c4_regions
    procs = max(1, (os.cpu_count() or 1) // world)
    t0 = time.perf_counter()
    work = synth.c4s_work(n_regions)
    shards = sharding.partition_regions(work, world)
    mine = shards[rank]
    B = synth.c4s_batch(mine, procs)
    gen_s = time.perf_counter() - t0
    pin = bench.pin(B)
    reps = max(1, min(args.steps, 3))
    n_total = 435 * n_regions

    # Host-side gather (one box): a shared-memory array of all the scores in description order;
    # every rank places its shard's rows, a gloo barrier closes the step.  No device collective.
    # Where /dev/shm is too small for the array (a SIGBUS otherwise), rank 0 collects the shards with
    # a gloo gather instead.
    shm = None
    use_shm = world > 1
    if world > 1:
        try:
            vfs = os.statvfs("/dev/shm")
            use_shm = vfs.f_bavail * vfs.f_frsize > 4 * n_total + (64 << 20) and \
                not os.environ.get("OTB200_NO_SHM")  # (switch: exercise the fallback)
        except OSError:
            use_shm = False
        flag = torch.tensor([1 if use_shm else 0], dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=bench.gloo)  # every rank must agree
        use_shm = bool(flag.item())
    if not use_shm:
        full = np.empty(n_total, dtype=np.int32) if rank == 0 else None
    else:
        from multiprocessing import shared_memory
        name = f"otb200_c4_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            try:
                shared_memory.SharedMemory(name=name).unlink()  # left over from a crashed run
            except FileNotFoundError:
                pass
            shm = shared_memory.SharedMemory(name=name, create=True, size=4 * n_total)
        dist.barrier(group=bench.gloo)
        if rank != 0:
            shm = shared_memory.SharedMemory(name=name)
            try:  # rank 0 owns the segment: keep this process's resource tracker out of it
                from multiprocessing import resource_tracker
                resource_tracker.unregister(shm._name, "shared_memory")
            except Exception:
                pass
        full = np.ndarray((n_total,), dtype=np.int32, buffer=shm.buf)
    rows = full.reshape(n_regions, 435) if full is not None else None

    def gather(scores):
        if world == 1 or use_shm:
            rows[mine] = scores.reshape(-1, 435)
            if world > 1:
                dist.barrier(group=bench.gloo)
            return
        t = torch.from_numpy(scores)
        if rank == 0:
            bufs = [torch.empty(435 * len(sh), dtype=torch.int32) for sh in shards]
            dist.gather(t, bufs, dst=0, group=bench.gloo)
            for sh, buf in zip(shards, bufs):
                rows[sh] = buf.numpy().reshape(-1, 435)
        else:
            dist.gather(t, None, dst=0, group=bench.gloo)

    bench.e2e_step(B, pin)  # warm-up (first touch of the device buffers)
    gather(pin["out"].array)
    pass_ms, gather_ms, total_ms = [], [], []
    for _ in range(reps):
        bench.barrier()
        t0 = time.perf_counter()
        bench.e2e_step(B, pin)
        t1 = time.perf_counter()
        gather(pin["out"].array)
        t2 = time.perf_counter()
        (pm, tm), _ = bench.reduce([(t1 - t0) * 1e3, (t2 - t0) * 1e3])
        pass_ms.append(pm)
        total_ms.append(tm)
        gather_ms.append(tm - pm)
    chunks = bench.ctx.last_chunk_count()
    bench.unpin(pin)
    (load_max,), (load_sum,) = bench.reduce([float(work[mine].sum())])
    rec = None
    if rank == 0:
        sec = float(np.mean(total_ms)) * 1e-3
        rec = {"workload": f"C4S: ONE description of {n_regions} TR regions x 30 HiFi reads (per-region seeds), "
                           f"partitioned by estimated work over {world} rank(s), generated per shard",
               "regions": n_regions, "pairs": n_total, "n_gpus": world,
               "seconds_per_pass": sec, "value": n_total / sec, "unit": UNIT,
               "stage_ms": {"shard_pass_max": float(np.mean(pass_ms)), "host_gather": float(np.mean(gather_ms))},
               "h2d_bytes_per_rank": int(B["bases"].nbytes), "gathered_bytes": int(4 * n_total),
               "chunks_per_rank": int(chunks), "passes": reps, "generation_s": gen_s,
               "partition_imbalance": load_max / (load_sum / world) if load_sum else 1.0,
               "checksum": int(full.astype(np.int64).sum())}
    if world > 1:
        # the same description on rank 0's GPU alone (the other ranks wait)
        if rank == 0:
            t0 = time.perf_counter()
            one = synth.c4s_batch(np.arange(n_regions), os.cpu_count() or 1)
            gen1 = time.perf_counter() - t0
            pin1 = bench.pin(one)
            bench.e2e_step(one, pin1)
            ts = []
            for _ in range(max(1, min(reps, 2))):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                bench.e2e_step(one, pin1)
                ts.append(time.perf_counter() - t0)
            same = bool(np.array_equal(pin1["out"].array, full))
            bench.unpin(pin1)
            rec["n1_seconds_per_pass"] = float(np.mean(ts))
            rec["speedup_vs_n1"] = rec["n1_seconds_per_pass"] / rec["seconds_per_pass"]
            rec["identical"] = same  # sharded + gathered == one GPU, score for score
            rec["n1_generation_s"] = gen1
            lim = "shard pass (device + H2D)" if rec["stage_ms"]["shard_pass_max"] > 4 * rec["stage_ms"]["host_gather"] else "host gather"
            rec["limiter"] = lim
        dist.barrier(group=bench.gloo)
    if rank == 0:
        rec["gather"] = ("shared-memory array on the host, every rank places its shard's rows; gloo barrier"
                         if use_shm else "gloo gather to rank 0 (/dev/shm too small for the array)") \
            if world > 1 else "local placement"
    if rank == 0 and not args.no_cpu_baseline:
        # spot check against the unmodified reference: the first regions of the description
        sub = synth.c4s_batch(np.arange(min(n_regions, 64)), 1)
        want, _ = reference_doubles(sub, sub["n_regions"])
        if want is not None:
            got = scores_to_doubles(sub, full[:len(want)], len(want))
            rec["identical_to_reference"] = bool(np.array_equal(got.view(np.uint64), want.view(np.uint64)))
    if shm is not None:
        del rows, full
        dist.barrier(group=bench.gloo)
        shm.close()
        if rank == 0:
            shm.unlink()
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="C2", choices=list(synth.CONFIGS))
    ap.add_argument("--regions", type=int, default=None, help="override regions per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--headline-only", action="store_true",
                    help="skip per_config / c4_strong / single_process / e2e_host_layer (kernel experiments)")
    ap.add_argument("--c4-regions", type=int, default=synth.C4S_REGIONS,
                    help="regions of the C4S description (c4_strong record)")
    ap.add_argument("--path", default="distmatrix", choices=["distmatrix", "consensus", "compare", "pipeline", "extract"],
                    help="distmatrix = fill_dist_matrix (the headline); consensus = the gap-affine "
                         "+ CIGAR alignments rapid_consensus issues (secondary row)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.path == "consensus":
        run_consensus(args, rank, world, local_rank)
        return
    if args.path == "pipeline":
        return run_pipeline(args, rank, world, local_rank)
    if args.path == "extract":
        return run_extract(args, rank, world, local_rank)
    if args.path == "compare":
        run_compare(args, rank, world, local_rank)
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    bench = Bench(rank, world, local_rank)
    torch = bench.torch
    extras = not args.headline_only and args.regions is None

    # ---- headline: the workload BASELINE.json's metric is quoted on (C2), weak scaling ----
    B = workload_arrays(args.workload, rank, args.regions)
    sampler = ClockSampler(local_rank)
    time.sleep(0.3)
    res, scores = bench.measure(B, args.steps, args.warmup, sampler)
    line = None
    if rank == 0:
        line = {"metric": METRIC, "value": res.pop("value"), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": res.pop("ms_per_step"), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": config_dict(args.workload, B, world)}
        line.update(res)
        cap, why = load_capture(args.workload, B["n_regions"])
        k_ms = line["roofline"]["kernel_ms"]
        if cap:
            line["roofline"]["traffic"] = cap.get("traffic_bytes_per_launch")
            clk = (line.get("clocks") or {}).get("sm_mhz")
            if cap.get("warp_instructions") and clk:
                # what actually bounds the kernel: one warp instruction per sub-partition per cycle
                n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
                peak_i = n_sm * 4 * clk * 1e6
                wi = cap["warp_instructions"]
                line["roofline"]["issue_slots"] = {
                    "warp_instructions_per_launch": wi, "source": f"profiles/{cap['file']} (ncu, this csrc, same workload)",
                    "achieved_per_s": wi / (k_ms * 1e-3), "peak_per_s": peak_i, "frac": wi / (k_ms * 1e-3) / peak_i}
        else:
            line["roofline"]["issue_slots"] = None
            line["roofline"]["issue_slots_reason"] = why
        if world == 1 and not args.no_cpu_baseline:
            base, same = cpu_baseline_leg(B, scores, with_variants=extras)
            line["cpu_baseline"] = base
            line["identical_to_reference"] = same
            w = work_counts(B)
            w["offset_updates_per_s"] = w["offset_updates_per_pair"] * line["value"]
            w["word_compares_per_s"] = w["word_compares_per_pair"] * line["value"]
            line["work"] = w

    # ---- the seam otter calls, same workload (N=1): fill_dist_matrices from std::string reads ----
    if extras and world == 1:
        hl, doubles = bench.host_layer(B, [local_rank], 1, max(2, min(args.steps, 3)))
        hl["identical_to_c_abi"] = bool(np.array_equal(
            doubles.view(np.uint64), scores_to_doubles(B, scores, B["n_scores"]).view(np.uint64)))
        hl["frac_of_e2e"] = hl["value"] / line["e2e"]["value"]
        line["e2e_host_layer"] = hl
    del B, scores

    # ---- the other configurations north_star names, same measurement each (N=1) ----
    if extras and world == 1:
        per = {}
        for name, nreg in (("C1", None), ("C3", None)):
            Bc = workload_arrays(name, 0, nreg)
            r, sc = bench.measure(Bc, max(2, min(args.steps, 3)), 3)
            r["config"] = config_dict(name, Bc, 1)
            capc, _ = load_capture(name, Bc["n_regions"])
            clkc = (r.get("clocks") or line.get("clocks") or {}).get("sm_mhz")
            if capc and capc.get("warp_instructions") and clkc:
                n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
                wi, k_ms_c = capc["warp_instructions"], r["roofline"]["kernel_ms"]
                r["roofline"]["traffic"] = capc.get("traffic_bytes_per_launch")
                r["roofline"]["issue_slots"] = {
                    "warp_instructions_per_launch": wi, "source": f"profiles/{capc['file']} (ncu, this csrc, same workload)",
                    "frac": wi / (k_ms_c * 1e-3) / (n_sm * 4 * clkc * 1e6)}
            if not args.no_cpu_baseline:
                r["cpu_baseline"], r["identical_to_reference"] = cpu_baseline_leg(Bc, sc, min_seconds=5.0)
            per[name] = r
            del Bc, sc
        # one eighth of the 100 k-region description: the shard a rank of 8 aligns
        from harness import sharding
        n_desc = args.c4_regions
        shard0 = sharding.partition_regions(synth.c4s_work(n_desc), 8)[0]
        Bc = synth.c4s_batch(shard0, os.cpu_count() or 1)
        x, y = synth.pair_lengths(Bc)
        Bc["alg_bytes"] = int(((x + 3) // 4 + (y + 3) // 4 + 4).sum())
        Bc["cells"] = int((x * y).sum())
        r, sc = bench.measure(Bc, 2, 3)
        r["config"] = {"workload": f"C4 shard: share 0 of 8 of the {n_desc}-region HiFi description "
                                   f"({Bc['n_regions']} regions x 30 reads), streamed in chunks",
                       "regions_per_gpu": Bc["n_regions"], "pairs_per_gpu": Bc["n_scores"],
                       "bases_per_gpu": int(Bc["bases"].size), "l2_policy": "inputs_exceed_l2"}
        if not args.no_cpu_baseline:
            r["cpu_baseline"], r["identical_to_reference"] = cpu_baseline_leg(Bc, sc, min_seconds=5.0)
        hl, _ = bench.host_layer(Bc, [local_rank], 1, 2)
        hl["frac_of_e2e"] = hl["value"] / r["e2e"]["value"]
        r["e2e_host_layer"] = hl
        per["C4_shard"] = r
        del Bc, sc
        line["per_config"] = per

    # ---- the target configuration: the 100 k-region description, strong scaling ----
    if extras:
        rec = c4_strong(bench, args)
        if rank == 0:
            line["c4_strong"] = rec

    # ---- one host process feeding all N devices (rank 0 only; the other ranks wait) ----
    if extras and world > 1:
        if rank == 0:
            n_desc = args.c4_regions
            ids = np.arange(min(n_desc, 1500 * world))
            Bm = synth.c4s_batch(ids, os.cpu_count() or 1)
            multi, dm = bench.host_layer(Bm, list(range(world)), 1, 2)
            single, d1 = bench.host_layer(Bm, [0], 1, 1)
            multi["multi_identical"] = bool(np.array_equal(dm.view(np.uint64), d1.view(np.uint64)))
            multi["single_device_ms_per_call"] = single["ms_per_call"]
            multi["speedup_vs_one_device"] = single["ms_per_call"] / multi["ms_per_call"]
            multi["workload"] = f"first {len(ids)} regions of the C4S description ({Bm['n_scores']} pairs, " \
                                f"{Bm['bases'].size >> 20} MB of reads) from ONE host process"
            multi["gather"] = "none: every region's DistMatrix is filled in place by the thread that owns its device"
            line["single_process"] = multi
        bench.dist.barrier(group=bench.gloo)

    if rank == 0:
        print(json.dumps(line), flush=True)
    bench.ctx.close()
    if world > 1:
        bench.dist.destroy_process_group()


if __name__ == "__main__":
    main()
