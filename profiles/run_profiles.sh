#!/bin/bash
# Profiling recipe (run under gpurun on one B200):  bash profiles/run_profiles.sh <tag>
# 1. plain run of the reduced bench command (must exit 0), 2. ncu launch list of the same command
# (cold-cache, serialised: compare SHARES), 3. one `--set full` capture per hot kernel.
# The reduced command = bench.py with 8 pairs instead of 64 (same per-pair grid geometry, 6144 CTAs),
# 1 timed step, full 200-KF Gauss-Newton leg with 1 iteration, no CPU / e2e legs.
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --pairs 8 --steps 1 --warmup 3 --no-cpu --no-e2e --gn-iters 1"
KREGEX='regex:iter_proj|refine_|gn_|chol_|sp_|prep_|match_tail'
$CMD > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -20 $OUT/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREGEX" -c 1500 --csv \
    --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_list_$TAG.log 2>&1
for spec in "iter_proj_tile_kernel:4" "refine_tiled_kernel:3" "gn_accum_stream_kernel:2"; do
  k=${spec%%:*}; s=${spec##*:}
  ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c 1 -f \
      -o $OUT/prof_${k}_$TAG $CMD > $OUT/ncu_${k}_$TAG.log 2>&1
done
ls -la $OUT
