"""Print the static SASS of the wf_pass round loops of k_align<false> (2-bit instances) with a
pipe tag per instruction, and ALU/FMA/other counts."""
import re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else "otter_b200/lib/libotter_b200.so"
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fn = None; L = []
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m: fn = m.group(1); continue
    if fn and "k_alignILb0" in fn:
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m: L.append((int(m.group(1), 16), m.group(2).strip()))
addr = {a: i for i, (a, _) in enumerate(L)}
ALU = ("ISETP","VIADD","LOP3","VIMNMX","VIADDMNMX","SEL","LEA","SHF","PLOP3","IADD3","PRMT","IABS","VIMNMX3")
def pipe(op):
    if op.startswith(ALU): return "ALU"
    if op.startswith(("IMAD","FFMA","HFMA","FMUL")): return "FMA"
    if op.startswith(("LDG","LDS","STS","STG","LDC","SHFL","ATOM","RED")): return "LSU"
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
            if sum("SHFL.UP" in x for x in txts) == 1 and any("0x1c" in x for x in txts):
                loops.append(body)
print("candidate loops:", [(hex(b[0][0]), len(b)) for b in loops])
for n, body in enumerate(loops):
    if which >= 0 and n != which: continue
    cnt = {}
    for a, t in body:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        p = pipe(op); cnt[p] = cnt.get(p, 0) + 1
        if which >= 0: print(f"{a:#06x} {p:4s} {t[:70]}")
    print(hex(body[0][0]), len(body), cnt)
