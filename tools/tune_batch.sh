#!/bin/bash
# Sweep the vote thresholds of the ray-batch kernels on cfg5 (run on the GPU box).
# usage: tools/tune_batch.sh "R L SL SI" ...   (FLX_VOTE_RETIRE_BATCH FLX_VOTE_LEAF_BATCH FLX_STAY_LEAF_BATCH FLX_STAY_INNER_BATCH)
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  set -- $cfg
  touch fluctus_b200/csrc/flx_trace.cu
  make -s -j4 -C fluctus_b200/csrc EXTRA="-DFLX_VOTE_RETIRE_BATCH=$1 -DFLX_VOTE_LEAF_BATCH=$2 -DFLX_STAY_LEAF_BATCH=$3 -DFLX_STAY_INNER_BATCH=$4" || exit 1
  python tools/raycast_only.py 3 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('batch $cfg: %.0f Mrays/s kernel %.2f ms frac %.3f e2e %.0f' % (r['value'], r['roofline']['kernel_ms'], r['roofline']['frac'], r['e2e']['value']))"
done
touch fluctus_b200/csrc/flx_trace.cu; make -s -j4 -C fluctus_b200/csrc
