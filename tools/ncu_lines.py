"""Per-CUDA-line totals (instructions executed, stall samples) from an ncu report captured with --import-source on.
Usage: python tools/ncu_lines.py <report.ncu-rep> [top_n]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
SORT = 0 if "--by-inst" in sys.argv else 1
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
cur, hdr = None, None
agg, src = collections.defaultdict(lambda: [0, 0, 0]), {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Line No":
        hdr = r
        ie, iss, it = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
        continue
    if hdr is None or not r[0].strip() or len(r) <= it or r[2] != "-":
        continue
    try:
        e, s, t = int(r[ie]), int(r[iss]), int(r[it])
    except ValueError:
        continue
    k = (cur, int(r[0]))
    agg[k][0] += e
    agg[k][1] += s
    agg[k][2] += t
    src[k] = r[1]
tot_e, tot_s = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
print(f"total warp instructions {tot_e}, stall samples {tot_s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][SORT])[:top]:
    print(f"{k[0]:18s}:{k[1]:4d} inst={100 * v[0] / tot_e:5.1f}% smp={100 * v[1] / tot_s:5.1f}% thr={v[2] / max(v[0], 1):4.1f} | {src[k].strip()[:100]}")
