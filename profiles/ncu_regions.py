#!/usr/bin/env python
"""Per-region / per-opcode breakdown of one kernel in an .ncu-rep (source page).
usage: ncu_regions.py file.ncu-rep   -- groups SASS instructions by their execution count (a proxy for
'which loop nest'), then by opcode: warp-instructions executed and stall samples."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
data = []
for r in rows[2:]:
    try:
        data.append((int(r[isamp]), int(r[iex]), r[ia]))
    except Exception:
        pass
tot_s = sum(d[0] for d in data) or 1
tot_e = sum(d[1] for d in data) or 1
def opc(t):
    p = t.split()
    o = p[1] if p[0].startswith('@') else p[0]
    return o.split('.')[0]
# regions: bucket exec counts to 2 significant digits
def bucket(e):
    if e == 0: return 0
    s = f"{e:.1e}"
    return float(s)
reg = collections.defaultdict(lambda: [0, 0, 0, collections.defaultdict(lambda: [0, 0, 0])])
for s, e, t in data:
    b = bucket(e)
    R = reg[b]
    R[0] += 1; R[1] += e; R[2] += s
    o = R[3][opc(t)]
    o[0] += 1; o[1] += e; o[2] += s
print(f"total: {len(data)} SASS, {tot_e} warp-instr, {tot_s} samples")
for b, R in sorted(reg.items(), key=lambda kv: -kv[1][1])[:8]:
    print(f"\nregion exec~{b:.3g}: {R[0]} SASS, {100*R[1]/tot_e:.1f}% of executed, {100*R[2]/tot_s:.1f}% of samples")
    for o, v in sorted(R[3].items(), key=lambda kv: -kv[1][1])[:14]:
        print(f"    {o:10s} n={v[0]:4d} exec%={100*v[1]/tot_e:5.1f} samp%={100*v[2]/tot_s:5.1f}")
