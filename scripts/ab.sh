#!/bin/bash
# usage: ab.sh name:lib:impl ...
for spec in "$@"; do
  IFS=: read name lib impl <<< "$spec"
  M3S_LIB=$PWD/mast3r-slam_b200/$lib M3S_ITER_PROJ_IMPL=$impl python bench.py --no-cpu --no-e2e --no-gn > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err
  python -c "import sys,json; d=json.loads(open('gpurun_out/b_$name.json').read().strip().splitlines()[-1]); k=d['roofline']['kernels']; print('$name', d['ms_per_step'], 'iter_proj', k['iter_proj_tile_kernel']['ms_per_launch'], 'refine', k['refine_tiled_kernel']['ms_per_launch'])" || tail -3 gpurun_out/b_$name.err
done
