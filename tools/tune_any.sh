#!/bin/bash
# Sweep the any-hit (shadow-ray) vote thresholds of the fused traversal launch (run on the GPU box).
# usage: tools/tune_any.sh "R L SL SI" ...   (FLX_VOTE_RETIRE_ANY FLX_VOTE_LEAF_ANY FLX_STAY_LEAF_ANY FLX_STAY_INNER_ANY)
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  set -- $cfg
  touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_microkernel.cu
  make -s -j4 -C fluctus_b200/csrc EXTRA="-DFLX_VOTE_RETIRE_ANY=$1 -DFLX_VOTE_LEAF_ANY=$2 -DFLX_STAY_LEAF_ANY=$3 -DFLX_STAY_INNER_ANY=$4" || exit 1
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-raycast 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']
print('any $cfg: value %.1f  trace %.2f logic %.2f mat %.2f' % (d['value'], s['trace'], s['logic'], s['materials']))"
done
touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_microkernel.cu; make -s -j4 -C fluctus_b200/csrc
