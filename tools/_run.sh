python tools/raycast_only.py 3 > gpurun_out/r1g_raycast.json 2> gpurun_out/r1g_raycast.err || { tail -5 gpurun_out/r1g_raycast.err; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_batch" --launch-skip 4 -c 1 -o gpurun_out/r1g_k_trace_batch python tools/raycast_only.py 2 > gpurun_out/r1g_ncu_full.log 2>&1
tail -2 gpurun_out/r1g_ncu_full.log; cut -c1-300 gpurun_out/r1g_raycast.json
