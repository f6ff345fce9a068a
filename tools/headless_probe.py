"""Run on a multi-GPU box: the C++ host's rate for several budgets and GPU counts (fixed cost vs rate)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
teapot, scene, env, settings = bench.load_workload()
for g in [int(x) for x in sys.argv[1].split(",")]:
    for spp in (12, 48):
        r = bench.run_headless_cpp(scene, env, settings, g, spp_per_gpu=spp)
        print("gpus %d spp/gpu %d: %s" % (g, spp, {k: r.get(k) for k in ("msamples_per_s", "seconds", "samples", "error")}), flush=True)
