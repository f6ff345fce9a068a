CMD="python bench.py --steps 3 --warmup 3 --no-raycast --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/r1c_bench_short.json 2> gpurun_out/r1c_bench_short.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1c_launches.csv $CMD > gpurun_out/r1c_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_ext|k_shadow" --launch-skip 60 -c 2 -o gpurun_out/r1c_k_ext $CMD > gpurun_out/r1c_ncu_full.log 2>&1
tail -2 gpurun_out/r1c_ncu_full.log
