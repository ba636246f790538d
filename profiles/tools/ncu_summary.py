#!/usr/bin/env python
"""ncu_summary.py <report.ncu-rep> <workload> <n_regions> <out-prefix>: the metrics the docs quote
(one launch of the dominant kernel), as <out-prefix>_summary.txt, and the two per-launch constants
bench.py may quote (DRAM bytes, warp instructions) as <out-prefix>_traffic.json, stamped with the
hash of the kernel sources they were captured on."""
import csv, io, json, os, subprocess, sys
rep, workload, n_regions, prefix = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, vals = rows[0], rows[1], rows[-1]
m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg.per_second"]
lines = []
for k in want:
    if k in m:
        lines.append(f"{k:70s} {m[k][0]} {m[k][1]}")
stalls = sorted(((float(v[0].replace(",", "")), k) for k, v in m.items()
                 if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v[0] not in ("", "n/a")),
                reverse=True)
lines.append("warp stalls (average warps stalled per issue-active cycle), top 8:")
for v, k in stalls[:8]:
    lines.append(f"   {k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):30s} {v:.3f}")
open(prefix + "_summary.txt", "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
def num(k):
    v, u = m[k]
    x = float(v.replace(",", ""))
    return x * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1}.get(u, 1)
import bench
json.dump({"workload": workload, "n_regions": n_regions, "kernel": m["Kernel Name"][0][:60],
           "traffic_bytes_per_launch": num("dram__bytes_read.sum") + num("dram__bytes_write.sum"),
           "dram_read_bytes": num("dram__bytes_read.sum"), "dram_write_bytes": num("dram__bytes_write.sum"),
           "warp_instructions": int(float(m["smsp__inst_executed.sum"][0].replace(",", ""))),
           "kernel_ms_under_ncu": float(m["gpu__time_duration.sum"][0].replace(",", "")),
           "csrc_sha": bench.csrc_sha(), "source": os.path.basename(rep)},
          open(prefix + "_traffic.json", "w"), indent=1)
