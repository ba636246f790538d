#!/usr/bin/env python
"""lineprof.py <report.ncu-rep> <lib.so> <kernel-substr> <srcroot> [file]: samples and executed warp
instructions per source line of one kernel (ncu SASS page joined with nvdisasm line info)."""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, lib, kern, root = sys.argv[1:5]
only = sys.argv[5] if len(sys.argv) > 5 else "wfa_affine.cuh"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=tmp, capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l and l.endswith(":")][0]
tags, cur = [], ("", 0)
for l in dis[start:]:
    if l.startswith("//-----"): break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        # inlined: keep the outermost call site inside `only` when the leaf is elsewhere
        cur = (m.group(1).split("/")[-1], int(m.group(2)), m.group(3)); continue
    if re.match(r"\s*/\*[0-9a-f]+\*/\s+\S", l): tags.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]; ie = hdr.index("Instructions Executed"); sm = hdr.index("# Samples")
st = {n: hdr.index(n) for n in ("stall_long_sb", "stall_wait", "stall_short_sb", "stall_not_selected", "stall_math", "stall_branch_resolving", "stall_lg")}
body = [r for r in rows[hi + 1:] if r and r[0].startswith("0x")]
assert len(body) == len(tags), (len(body), len(tags))
agg = collections.defaultdict(lambda: collections.Counter())
tot = ts = 0
for r, tag in zip(body, tags):
    f, ln = tag[0], tag[1]
    key = (f, ln)
    a = agg[key]
    a["i"] += int(r[ie]); a["s"] += int(r[sm])
    for n, ix in st.items(): a[n] += int(r[ix])
    tot += int(r[ie]); ts += int(r[sm])
src = {}
print(f"total instr {tot:,} samples {ts}")
byfile = collections.Counter()
for (f, ln), a in agg.items(): byfile[f] += a["s"]
print({k: round(100 * v / ts, 1) for k, v in byfile.most_common(8)})
for (f, ln), a in sorted(agg.items()):
    if f != only or a["s"] < 0.002 * ts: continue
    if f not in src:
        src[f] = open(os.path.join(root, "otter_b200/csrc", f)).read().split("\n")
    code = src[f][ln - 1].strip()[:70]
    print(f"{ln:4d} {100*a['i']/tot:5.1f}%i {100*a['s']/ts:5.1f}%s  lsb {100*a['stall_long_sb']/ts:4.1f} wait {100*a['stall_wait']/ts:4.1f} ssb {100*a['stall_short_sb']/ts:4.1f} | {code}")
