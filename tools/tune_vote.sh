#!/bin/bash
# Rebuild the traversal TU with different vote thresholds and print the stage times (run on the GPU box).
# usage: tools/tune_vote.sh "R L SL SI" ...   (FLX_VOTE_RETIRE FLX_VOTE_LEAF FLX_STAY_LEAF FLX_STAY_INNER)
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  set -- $cfg
  touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_logic.cu fluctus_b200/csrc/flx_repack.cu
  make -s -C fluctus_b200/csrc EXTRA="-DFLX_VOTE_RETIRE=$1 -DFLX_VOTE_LEAF=$2 -DFLX_STAY_LEAF=$3 -DFLX_STAY_INNER=$4 $EXTRA_DEFS" || exit 1
  python bench.py --steps 2 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; r=d['raycast']
print('cfg $cfg $EXTRA_DEFS: value %.1f  trace %.2f logic %.2f mat %.2f raycast %.0f Mrays/s (kernel %.2f ms) parity %s' % (d['value'], s['trace'], s['logic'], s['materials'], r['value'], r['roofline']['kernel_ms'], r.get('parity',{}).get('hit_i_equal')))"
done
