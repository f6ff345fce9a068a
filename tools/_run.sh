# r1f: final-state evidence — full bench line, ncu launch list of the same command, one full capture of the traversal launch
python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err || exit 1
CMD="python bench.py --steps 2 --warmup 3 --no-raycast --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r1f_launches.csv $CMD > gpurun_out/r1f_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_ext_shadow" --launch-skip 100 -c 1 -o gpurun_out/r1f_k_ext_shadow $CMD > gpurun_out/r1f_ncu_full.log 2>&1
tail -2 gpurun_out/r1f_ncu_full.log
cut -c1-400 gpurun_out/r1f_bench.json
