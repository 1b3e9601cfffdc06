#!/usr/bin/env python
"""Summarise an .ncu-rep: key metrics + top stall instructions.  usage: ncu_summary.py file.ncu-rep [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 14
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, v = rows[0], rows[1], rows[-1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio']
for w in want:
    if w in h:
        i = h.index(w)
        print(f'{w:72s} {u[i]:>16s} {v[i]}')
st = {n.replace('smsp__pcsamp_warps_issue_stalled_', ''): int(v[i]) for i, n in enumerate(h)
      if 'pcsamp_warps_issue_stalled' in n and 'not_issued' not in n and v[i].isdigit()}
tot = sum(st.values()) or 1
print('stalls:', ', '.join(f'{k} {100*x/tot:.0f}%' for k, x in sorted(st.items(), key=lambda kv: -kv[1])[:8]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
data = []
for r in rows[2:]:
    try:
        data.append((int(r[isamp]), int(r[iex]), r[ia], r))
    except Exception:
        pass
tots = sum(d[0] for d in data) or 1
print(f'total samples {tots}, SASS instructions {len(data)}, warp-instr executed {sum(d[1] for d in data)}')
ops = {}
for s_, e, text, r in data:
    t = text.split()
    op = (t[1] if t and t[0].startswith('@') and len(t) > 1 else (t[0] if t else '?')).split('.')[0]
    ops[op] = ops.get(op, 0) + e
te = sum(ops.values()) or 1
print('dynamic opcode mix:', ', '.join(f'{k} {100*x/te:.1f}%' for k, x in sorted(ops.items(), key=lambda kv: -kv[1])[:14]))
for s, e, text, r in sorted(data, key=lambda d: -d[0])[:topn]:
    sts = {hdr[k].replace('stall_', ''): r[k] for k in range(len(hdr)) if hdr[k].startswith('stall_') and '(' not in hdr[k] and r[k] not in ('0', '')}
    top = sorted(sts.items(), key=lambda kv: -int(kv[1]))[:2]
    print(f'{100*s/tots:5.1f}%  exec={e:9d}  {text[:70]:70s} {top}')
