CMD="python bench.py --steps 3 --warmup 3 --no-raycast --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/r1e_bench_short.json 2> gpurun_out/r1e_bench_short.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_logic|k_material" --launch-skip 245 -c 8 -o gpurun_out/r1e_shade $CMD > gpurun_out/r1e_ncu_full.log 2>&1
tail -2 gpurun_out/r1e_ncu_full.log
