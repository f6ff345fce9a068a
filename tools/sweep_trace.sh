#!/bin/bash
# Rebuild the traversal TUs with extra -D flags and print kitchen + cfg5 numbers for each set (run on the GPU box).
# usage: tools/sweep_trace.sh "-DFLX_VOTE_RETIRE=8" "-DFLX_RETIRE_RING=0" ...     ("" = the defaults)
cd "$(dirname "$0")/.."
for defs in "$@"; do
  touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_noalloc.cu
  make -s -C fluctus_b200/csrc EXTRA="$defs" 2>&1 | grep -i "error" && exit 1
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-opencl --no-configs --no-headless --no-strong-scaling --no-lbvh 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; r=d['raycast']
print('[$defs] value %.1f  trace %.2f logic %.2f mat %.2f | cfg5 %.0f Mrays/s (kernel %.3f ms) parity %s' % (d['value'], s['trace'], s['logic'], s['materials'], r['value'], r['roofline']['kernel_ms'], r.get('parity',{}).get('hit_i_equal')))"
done
touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_noalloc.cu; make -s -C fluctus_b200/csrc
