#!/usr/bin/env python
"""Generate tests/golden/prune_golden.json: pairs on which the reference's exact prune
(WFA2-lib wavefront/wavefront_compute_edit.c:214-269, called from :341-347) actually changes the
wavefront under otter's WF-adaptive(10,50,1) edit aligner.

Found by oracle/tools/prune_search.c (instrumented C restatement over adversarial families:
unrelated reads, shifted overlaps, low-complexity blocks, mosaics, ...).  Expected values come
from the UNMODIFIED reference (oracle/_ref): WFAlignerEdit(Score) score, and for the smaller
pairs WFAlignerEdit(Alignment) score + CIGAR digest (the `compare` aligner takes the same prune).
Run in the build container:
    python tests/golden/make_golden_prune.py
"""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402

PER_FAMILY = 4          # smallest firing pairs kept per family
SEEDS = [(7, 3000, 2600), (11, 3000, 3500), (1, 2000, 7000)]  # (seed, pairs, longest base sequence)


def search():
    tmp = tempfile.mkdtemp()
    exe = os.path.join(tmp, "prune_search")
    tools = os.path.join(ROOT, "oracle", "tools")
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-o", exe, os.path.join(tools, "prune_search.c"),
                           os.path.join(ROOT, "oracle", "wfa_oracle.c")])
    found = []
    for seed, n, lmax in SEEDS:
        out = os.path.join(tmp, f"fires_{seed}.txt")
        log = subprocess.run([exe, str(seed), str(n), out, "1", str(lmax)], capture_output=True, text=True)
        print(log.stdout.strip().splitlines()[-1])
        rows = open(out).read().split("\n")
        for i in range(0, len(rows) - 2, 3):
            head = rows[i].split()
            found.append(dict(fam=int(head[0][4:]), p=rows[i + 1], t=rows[i + 2],
                              origin=f"prune_search seed {seed} {head[1]} {head[2]}"))
    return found


def main():
    R = S.ref()
    P = S.port()
    found = search()
    keep = []
    for fam in sorted({f["fam"] for f in found}):
        rows = sorted((f for f in found if f["fam"] == fam), key=lambda f: len(f["p"]) + len(f["t"]))
        # smallest ones, but at least one with the text longer than the pattern if there is any
        pick = rows[:PER_FAMILY]
        swapped = [f for f in rows if len(f["t"]) > len(f["p"])]
        if swapped and not any(len(f["t"]) > len(f["p"]) for f in pick):
            pick[-1] = swapped[0]
        keep += pick
    # two pairs whose wavefront also outgrows the 2048-slot shared ring (prune on the HBM ring)
    very_wide = []
    for f in sorted(found, key=lambda f: len(f["p"]) + len(f["t"])):
        if f in keep:
            continue
        ctr = S.Counters()
        p, t = f["p"].encode(), f["t"].encode()
        P.wfo_edit_score(p, len(p), t, len(t), 0, 0, 0, 0, 0, ctr)
        if ctr.max_width > 1912:
            very_wide.append(f)
            if len(very_wide) == 2:
                break
    keep += very_wide
    out = dict(source="holstegelab/otter @ 2573c48e + WFA2-lib v2.2 (52e8ac1d), WFAlignerEdit, default "
                      "WF-adaptive(10,50,1); pairs found by oracle/tools/prune_search.c", pairs=[])
    for f in keep:
        p, t = f["p"].encode(), f["t"].encode()
        ctr = S.Counters()
        sp = P.wfo_edit_score(p, len(p), t, len(t), 0, 0, 0, 0, 0, ctr)
        sr = R.ref_edit_e2e(p, len(p), t, len(t))
        assert sp == sr and ctr.prunes > 0 and ctr.prune_gate >= ctr.prunes, (sp, sr, ctr.prunes)
        row = dict(p=f["p"], t=f["t"], score=sr, prunes=ctr.prunes, gate_steps=ctr.prune_gate,
                   max_width=ctr.max_width, fam=f["fam"], origin=f["origin"])
        buf = C.create_string_buffer(len(p) + len(t) + 8)
        sc, st = C.c_int(), C.c_int()
        n = R.ref_cigar(1, 0, p, len(p), 0, 0, t, len(t), 0, 0, C.byref(sc), C.byref(st), buf, len(buf))
        row["cigar_score"] = sc.value
        row["cigar_len"] = n
        row["cigar_md5"] = hashlib.md5(buf.raw[:n]).hexdigest()
        out["pairs"].append(row)
    path = os.path.join(HERE, "prune_golden.json")
    json.dump(out, open(path, "w"))
    print(path, os.path.getsize(path), "bytes;", len(out["pairs"]), "pairs; families",
          sorted({r["fam"] for r in out["pairs"]}), "; text-longer pairs",
          sum(len(r["t"]) > len(r["p"]) for r in out["pairs"]))


if __name__ == "__main__":
    main()
