# group the ncu source page (cuda view) of k_align by function of wfa_edit.cuh
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
groups = [("window64/first_diff", 121, 140), ("match_window", 141, 156), ("lcp_warp", 164, 216), ("lcp_uniform", 234, 262),
          ("first_diff", 264, 279), ("lcp_more", 281, 309), ("wf_pass", 345, 426), ("wf_trim", 428, 450),
          ("wf_scan", 497, 520), ("wf_edges", 522, 580), ("wf_step_narrow", 582, 644), ("wf_prune", 646, 675),
          ("run loop", 677, 733), ("ring", 48, 101)]
hdr = None; cur_file = None; agg = collections.Counter(); tot = 0; smp = collections.Counter(); ts = 0
for r in rows:
    if not r: continue
    if r[0] in ('File Path','File Name'): cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; ie = hdr.index('Instructions Executed'); sm = hdr.index('# Samples'); continue
    if hdr is None or not r[0].isdigit(): continue
    f = lambda x: int(x) if x.isdigit() else 0
    n = f(r[ie]); s = f(r[sm]); ln = int(r[0]); tot += n; ts += s
    key = cur_file
    if cur_file == 'wfa_edit.cuh':
        for g, a, b in groups:
            if a <= ln <= b: key = g; break
        else: key = 'wfa_edit.cuh:other'
    agg[key] += n; smp[key] += s
print("total warp instructions", tot, "samples", ts)
for k, v in agg.most_common(): print(f"{100*v/tot:5.1f}%i {100*smp[k]/ts:5.1f}%s  {k}")
