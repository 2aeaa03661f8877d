"""Dump the SASS of one kernel from libb200vr.so and summarise its innermost loops (backward branches): instruction count
and mnemonic histogram per loop body. Usage: python tools/sass_loop.py <kernel-substring> [min_len]"""
import collections
import re
import subprocess
import sys

so = "open-volume-renderer_b200/libb200vr.so"
pat = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
fn = [b for b in blocks if pat in b.split("\n")[0]]
assert fn, "kernel not found"
body = fn[0]
ins = []
for line in body.split("\n"):
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
print(f"{fn[0].splitlines()[0][:90]}: {len(ins)} instructions")
loops = []
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA(?:\.\w+)*\s+(?:!?U?P\d,\s*)*(?:`\()?0x([0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr_index:
            loops.append((addr_index[tgt], i))
minlen = int(sys.argv[2]) if len(sys.argv) > 2 else 8
for lo, hi in sorted(loops, key=lambda x: x[1] - x[0]):
    n = hi - lo + 1
    if n < minlen:
        continue
    hist = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in ins[lo:hi + 1])
    print(f"loop {ins[lo][0]:#x}..{ins[hi][0]:#x}: {n} instr | " + " ".join(f"{k}:{v}" for k, v in hist.most_common(14)))
if "--dump" in sys.argv:
    lo, hi = [int(x, 16) for x in sys.argv[sys.argv.index("--dump") + 1].split("-")]
    for a, t in ins:
        if lo <= a <= hi:
            print(f"{a:#06x}  {t}")
