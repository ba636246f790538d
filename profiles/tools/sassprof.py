#!/usr/bin/env python
"""sassprof.py <report.ncu-rep> <lib.so> [kernel-substr]: executed warp instructions of a kernel by
function of wfa_edit.cuh.  Joins ncu's per-SASS-instruction counts with nvdisasm line info of the
same library; leaf helpers (window64, first_diff, ring, intrinsics) are attributed to the last
top-level function seen in address order."""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, lib = sys.argv[1], sys.argv[2]
kern = sys.argv[3] if len(sys.argv) > 3 else "k_alignILi1024"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
# function extents of the kernel headers
FILES = ["wfa_edit.cuh", "wfa_affine.cuh", "wfa_edit_cigar.cuh"]
funcs = {}
for fn in FILES:
    src = open(os.path.join(ROOT, "otter_b200/csrc", fn)).read().split("\n")
    fl = []
    for i, l in enumerate(src, 1):
        m = re.match(r"^(?:__device__|template|struct)?.*?\b([A-Za-z_0-9]+)\s*\((?:const|int|uint|unsigned|typename|Ring|Grp|otgpu|AffWork|AffSlot|EditJob|char)", l)
        if l.startswith("__device__") and m:
            fl.append((i, m.group(1)))
    funcs[fn] = fl
def func_of(fn, line):
    name = "?"
    for a, n in funcs[fn]:
        if a <= line: name = n
        else: break
    return name
LEAF = {"window64", "first_diff", "match_window", "lds32", "sts32", "bitsel", "opaque", "wf_dist", "?"}
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l and l.endswith(":")][0]
tags, cur = [], ("", 0)
for l in dis[start:]:
    if l.startswith("//-----"): break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s*/\*[0-9a-f]+\*/\s+\S", l): tags.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; ie = hdr.index("Instructions Executed"); sm = hdr.index("# Samples")
body = [r for r in rows[hi + 1:] if r and r[0].startswith("0x")]
assert len(body) == len(tags), (len(body), len(tags))
agg, smp, leafagg = collections.Counter(), collections.Counter(), collections.Counter()
owner = "kernel"
tot = ts = 0
for r, (f, ln) in zip(body, tags):
    n, s = int(r[ie]), int(r[sm])
    if f in funcs:
        fn = func_of(f, ln)
        if fn not in LEAF: owner = fn
        leaf = fn
    elif f == "otgpu.cu": owner = leaf = "kernel"
    else: leaf = f
    agg[owner] += n; smp[owner] += s; leafagg[(owner, leaf)] += n; tot += n; ts += s
print(f"total warp instructions {tot:,}  samples {ts}")
for k, v in agg.most_common():
    print(f"{100*v/tot:5.1f}%i {100*smp[k]/max(ts,1):5.1f}%s  {k}")
    for (o, lf), c in sorted(leafagg.items(), key=lambda x: -x[1]):
        if o == k and lf != k and c > 0.004 * tot: print(f"          {100*c/tot:5.1f}%  {lf}")
