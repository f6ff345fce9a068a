CMD="python bench.py --steps 2 --warmup 3 --no-raycast --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_logic|k_materials" --launch-skip 200 -c 2 -o gpurun_out/r1h_shade $CMD > gpurun_out/r1h_ncu_full.log 2>&1
tail -2 gpurun_out/r1h_ncu_full.log
