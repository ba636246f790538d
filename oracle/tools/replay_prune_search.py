#!/usr/bin/env python
"""oracle/tools/replay_prune_search.py — TEST INFRASTRUCTURE ONLY.

Replays every pair on which prune_search.c saw the exact prune fire (its out files) through the
UNMODIFIED reference (oracle/_ref/libotter_ref.so, WFAlignerEdit end-to-end) and compares the score
with the one the C restatement printed in the record header.  Needs /root/reference to have been
built (oracle/Makefile); run where the search ran:

    python oracle/tools/replay_prune_search.py <dir with out*.txt and log*.txt> [threads]
"""
import glob
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import support as S  # noqa: E402


def records(path):
    rows = open(path).read().split("\n")
    for i in range(0, len(rows) - 2, 3):
        head = rows[i].split()
        yield dict(fam=int(head[0][4:]), score=int(head[4]), prunes=int(head[6]), p=rows[i + 1], t=rows[i + 2])


def main():
    d = sys.argv[1]
    threads = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    R = S.ref()
    recs = [r for f in sorted(glob.glob(os.path.join(d, "out*.txt"))) for r in records(f)]

    def one(r):
        p, t = r["p"].encode(), r["t"].encode()
        return R.ref_edit_e2e(p, len(p), t, len(t)) == r["score"]  # ctypes releases the GIL

    with ThreadPoolExecutor(threads) as ex:
        same = list(ex.map(one, recs))
    tot = dict(pairs=0, wide=0, gate=0, fired=0, viol=0)
    for f in sorted(glob.glob(os.path.join(d, "log*.txt"))):
        last = [l for l in open(f).read().splitlines() if l.startswith("TOTAL")][-1].split()
        tot["pairs"] += int(last[4]); tot["wide"] += int(last[6]); tot["gate"] += int(last[9])
        tot["fired"] += int(last[13]); tot["viol"] += int(last[16])
    print(f"searched pairs {tot['pairs']}, wavefront >= 1000 diagonals on {tot['wide']}, gate-open steps {tot['gate']}, "
          f"pairs on which the prune fired {tot['fired']}, firings with the gate closed {tot['viol']}")
    print(f"firing pairs replayed through the unmodified reference: {len(recs)}, "
          f"score equal to the restatement's: {sum(same)}, different: {len(recs) - sum(same)}")
    by = {}
    for r, ok in zip(recs, same):
        a = by.setdefault(r["fam"], [0, 0, 0])
        a[0] += 1; a[1] += ok; a[2] = max(a[2], r["prunes"])
    for fam in sorted(by):
        print(f"  family {fam}: {by[fam][0]} firing pairs, {by[fam][1]} equal, up to {by[fam][2]} firings in one pair")


if __name__ == "__main__":
    main()
