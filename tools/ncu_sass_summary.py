#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv --print-source sass` for the first kernel in the report:
opcode mix (warp instructions executed), lane efficiency, top stall sites.  Usage: ncu_sass_summary.py X.ncu-rep [N]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
# split into kernels
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        kernels.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
        cur["rows"].append(r)
k = kernels[0]
h = k["hdr"]
ci = {n: i for i, n in enumerate(h)}
print(k["name"][:120])
ops = collections.Counter(); thr = collections.Counter(); samples = collections.Counter()
tot_i = tot_t = tot_s = 0
for r in k["rows"]:
    src = r[ci["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    op = op.split(".")[0]
    ie = float(r[ci["Instructions Executed"]] or 0); te = float(r[ci["Thread Instructions Executed"]] or 0)
    s = float(r[ci["# Samples"]] or 0)
    ops[op] += ie; thr[op] += te; samples[op] += s
    tot_i += ie; tot_t += te; tot_s += s
print(f"warp insts {tot_i:.0f}  thread insts {tot_t:.0f}  lane eff {tot_t / tot_i / 32:.3f}  sass lines {len(k['rows'])}")
print("opcode mix (warp insts %, lane eff, stall-sample %):")
for op, v in ops.most_common(topn):
    print(f"  {op:12s} {100 * v / tot_i:6.2f}%  eff {thr[op] / max(v, 1) / 32:.2f}  samples {100 * samples[op] / max(tot_s, 1):5.1f}%")
print("top stall sites:")
srt = sorted(k["rows"], key=lambda r: -float(r[ci["# Samples"]] or 0))[:topn]
stall_cols = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
for r in srt:
    st = sorted(((float(r[ci[n]] or 0), n) for n in stall_cols), reverse=True)[:2]
    print(f"  {100 * float(r[ci['# Samples']]) / tot_s:5.2f}%  {r[ci['Address']][-5:]}  {r[ci['Source']].strip()[:70]:70s} {st[0][1]}:{st[0][0]:.0f} {st[1][1]}:{st[1][0]:.0f}")
# basic blocks: runs of equal (instructions executed)
print("basic blocks (share of warp insts, n inst, executions, lane eff, head):")
blocks = []
cur = None
for r in k["rows"]:
    ie = float(r[ci["Instructions Executed"]] or 0); te = float(r[ci["Thread Instructions Executed"]] or 0)
    src = r[ci["Source"]].strip()
    if cur is None or cur["exec"] != ie:
        cur = {"exec": ie, "n": 0, "thr": 0.0, "addr": r[ci["Address"]][-5:], "ops": []}
        blocks.append(cur)
    cur["n"] += 1; cur["thr"] += te; cur["ops"].append(src.split("(")[0].split()[1 if src.startswith("@") else 0] if src else "?")
for b in blocks:
    share = 100 * b["exec"] * b["n"] / tot_i
    if share >= float(sys.argv[3]) if len(sys.argv) > 3 else share >= 0.5:
        print(f"  {share:5.2f}%  {b['addr']}  n={b['n']:4d}  x{b['exec']:.0f}  eff {b['thr'] / max(b['exec'] * b['n'], 1) / 32:.2f}  {' '.join(b['ops'][:12])}")
