#!/bin/bash
# Rebuild the library with extra -D flags and print value + stage times for each set (run on the GPU box).
# usage: tools/tune_defs.sh "-DFLX_MAT_MIN_BLOCKS=3" "-DFLX_LOGIC_MIN_BLOCKS=5 -DFLX_MAT_MIN_BLOCKS=2" ...
cd "$(dirname "$0")/.."
for defs in "$@"; do
  touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_logic.cu fluctus_b200/csrc/flx_wavefront.cu fluctus_b200/csrc/flx_microkernel.cu fluctus_b200/csrc/flx_repack.cu
  make -s -C fluctus_b200/csrc EXTRA="$defs" || exit 1
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-raycast 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']
print('[$defs] value %.1f  trace %.2f logic %.2f mat %.2f' % (d['value'], s['trace'], s['logic'], s['materials']))"
done
touch fluctus_b200/csrc/flx_trace.cu fluctus_b200/csrc/flx_logic.cu fluctus_b200/csrc/flx_wavefront.cu fluctus_b200/csrc/flx_microkernel.cu fluctus_b200/csrc/flx_repack.cu; make -s -C fluctus_b200/csrc
