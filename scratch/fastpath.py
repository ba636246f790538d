"""Fast path (slow-path branches skipped) of the 2-bit end-to-end wf_pass round loop in
k_align<false>, with a pipe tag per instruction."""
import re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else "otter_b200/lib/libotter_b200.so"
show = len(sys.argv) > 2
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fn = None; L = []
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m: fn = m.group(1); continue
    if fn and "k_alignILi1024" in fn:
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m: L.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: i for i, (a, _) in enumerate(L)}
ALU = ("ISETP","VIADD","LOP3","VIMNMX","VIADDMNMX","SEL","LEA","SHF.","PLOP3","IADD3","PRMT","IABS")
def pipe(op):
    if op.startswith(ALU): return "ALU"
    if op.startswith(("IMAD","FFMA","HFMA","FMUL")): return "FMA"
    if op.startswith(("LDG","LDS","STS","STG","LDC","SHFL","ATOM","RED","LDL","STL")): return "LSU"
    if op.startswith(("BREV","FLO","POPC","MUFU")): return "XU"
    return "CTL"
loops = []
for i, (a, t) in enumerate(L):
    mm = re.search(r"BRA.*0x([0-9a-f]+)", t)
    if mm:
        tgt = int(mm.group(1), 16)
        if tgt < a and tgt in addr:
            body = L[addr[tgt]:i + 1]
            txts = [x[1] for x in body]
            if sum("LDS.64" in x for x in txts) == 1 and any("0x1c" in x for x in txts):
                loops.append(body)
for body in loops:
    amap = {a: i for i, (a, _) in enumerate(body)}
    i = 0; out = []; vote = False
    while i < len(body):
        a, t = body[i]; out.append((a, t))
        if "VOTE" in t: vote = True
        mm = re.search(r"BRA.*0x([0-9a-f]+)", t)
        if mm and vote and t.startswith("@"):
            tgt = int(mm.group(1), 16)
            if tgt > a and tgt in amap: i = amap[tgt]; vote = False; continue
        i += 1
    cnt = {}
    for a, t in out:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        p = pipe(op); cnt[p] = cnt.get(p, 0) + 1
        if show: print(f"{a:#06x} {p:4s} {t[:64]}")
    print(hex(body[0][0]), "fast path", len(out), cnt)
