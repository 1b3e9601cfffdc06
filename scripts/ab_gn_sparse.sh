#!/bin/bash
# usage: ab_gn_sparse.sh "name:VAR=value VAR2=value" ...   (GN leg only; prints the sparse-front shape too)
for spec in "$@"; do
  IFS=: read name kv <<< "$spec"
  env $kv timeout 300 python bench.py --pairs 2 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/g_$name.json 2> gpurun_out/g_$name.err
  python -c "import sys,json; d=json.loads(open('gpurun_out/g_$name.json').read().strip().splitlines()[-1])['gn']; print('$name', 'iter', d['ms_per_iter'], 'acc', d['accumulate_ms'], 'solve', d['solve_update_ms'], d['status']['last_dx_norm'], d['sparse_front'])" || tail -3 gpurun_out/g_$name.err
done
