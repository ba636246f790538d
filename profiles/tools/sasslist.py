#!/usr/bin/env python
"""sasslist.py <report.ncu-rep> <lib.so> <out.txt> [kernel-substr]: per-instruction listing: exec count, samples, SASS, source tag."""
import csv, io, os, re, subprocess, sys, tempfile
rep, lib, outp = sys.argv[1], sys.argv[2], sys.argv[3]
kern = sys.argv[4] if len(sys.argv) > 4 else "k_alignILi1024"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l and l.endswith(":")][0]
items, cur = [], ""
for l in dis[start:]:
    if l.startswith("//-----"): break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur = f"{m.group(1).split('/')[-1]}:{m.group(2)}"; continue
    m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*);", l)
    if m: items.append(("I", m.group(1), m.group(2).strip(), cur))
    elif l.startswith(".L_"): items.append(("L", l, "", ""))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; ie = hdr.index("Instructions Executed"); sm = hdr.index("# Samples")
body = [r for r in rows[hi + 1:] if r and r[0].startswith("0x")]
bi = 0
with open(outp, "w") as f:
    for kind, a, b, c in items:
        if kind == "L": f.write(f"{a}\n"); continue
        r = body[bi]; bi += 1
        f.write(f"{int(r[ie]):>12,d} {int(r[sm]):>6d}  {a:6s} {b:64s} {c}\n")
