#!/bin/bash
# The evidence run kept under profiles/ (run on the GPU box: gpurun -- 'bash tools/profile_round.sh r1l'):
#   1. the full default bench line (never under a profiler),
#   2. the ncu launch list of a short bench command (gpu__time_duration.sum only: shares, not absolutes),
#   3. one `ncu --set full` capture of the traversal launch at a steady-state iteration.
# Summaries: tools/ncu_summary.py launches|kernel, tools/ncu_lines.py.
TAG=${1:-rX}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || exit 1
CMD="python bench.py --steps 2 --warmup 3 --no-raycast --no-cpu-baseline --no-reference-opencl --no-configs --no-headless --no-strong-scaling --no-lbvh"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_ext_shadow" --launch-skip 100 -c 1 -o gpurun_out/${TAG}_k_ext_shadow $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log
#   4. the shading stages (logic + the fused material launch) at the same steady-state iteration,
#   5. the cfg5 ray-cast kernel (10.1 M triangles, 2^24 incoherent rays): the HBM-facing case.
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_logic|k_materials" --launch-skip 200 -c 2 -o gpurun_out/${TAG}_shade $CMD > gpurun_out/${TAG}_ncu_shade.log 2>&1
tail -1 gpurun_out/${TAG}_ncu_shade.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_batch" --launch-skip 4 -c 1 -o gpurun_out/${TAG}_k_trace_batch python tools/raycast_only.py 2 > gpurun_out/${TAG}_ncu_raycast.log 2>&1
tail -1 gpurun_out/${TAG}_ncu_raycast.log
cut -c1-400 gpurun_out/${TAG}_bench.json
