#!/bin/bash
# Rebuild the traversal TUs with extra -D flags and print render value, stage times and the cfg5 ray-cast figures
# for each set (run on the GPU box).  usage: tools/tune_trace.sh "" "-DFLX_PF_PUSH=1" "-DFLX_PF_PUSH=1 -DFLX_PF_CUR=1" ...
cd "$(dirname "$0")/.."
TUS="$(ls fluctus_b200/csrc/*.cu)"
for defs in "$@"; do
  touch $TUS
  make -s -j4 -C fluctus_b200/csrc EXTRA="$defs" || exit 1
  python bench.py --steps ${STEPS:-3} --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; r=d['raycast']
print('[$defs] value %.1f e2e %.1f trace %.2f logic %.2f mat %.2f | raycast %.0f Mrays/s (kernel %.2f ms, frac %.3f) e2e %.0f parity %s' % (d['value'], d['e2e']['value'], s['trace'], s['logic'], s['materials'], r['value'], r['roofline']['kernel_ms'], r['roofline']['frac'], r['e2e']['value'], r.get('parity',{}).get('hit_i_equal')))"
done
touch $TUS; make -s -j4 -C fluctus_b200/csrc
