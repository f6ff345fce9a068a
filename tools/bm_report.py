"""Evidence run of `fluctus_headless --benchmark-file` (Tracer::runBenchmarkFromFile, reference src/tracer.cpp:567-870):
writes the benchmark's scenes as .flxs containers, a .bm.json in the reference's format, runs the binary and keeps the
reference-format outputs (per-scene .csv / .txt, summary lines) under gpurun_out/bm_report/.  Run on the GPU box:

    python tools/bm_report.py [seconds-per-scene]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fluctus_b200 import scene as fs  # noqa: E402

seconds = int(sys.argv[1]) if len(sys.argv) > 1 else 5
out = os.path.join(ROOT, "gpurun_out", "bm_report")
os.makedirs(out, exist_ok=True)
teapot, kitchen, env, settings = bench.load_workload()
fs.save_flxs(os.path.join(out, "teapot.flxs"), teapot, env)
fs.save_flxs(os.path.join(out, "kitchen_standin.flxs"), kitchen, env)
cam = {"pos": [float(x) for x in settings["cam_pos"]], "dir": [float(x) for x in settings["cam_dir"]], "fov": float(settings["fov"])}
light = {"pos": [float(x) for x in settings["light_pos"]], "N": [float(x) for x in settings["light_n"]], "E": [float(x) for x in settings["light_e"]],
         "size": [float(x) for x in settings["light_size"]]}
view = {"camera": cam, "areaLight": light}
job = {"settings": {"windowWidth": 1024, "windowHeight": 1024, "useWavefront": True, "wfBufferSize": 1 << 22, "maxPathDepth": 8, "useEnvMap": True,
                    "useAreaLight": True, "useSeparateQueues": True, "maxRenderTime": seconds, "maxSpp": 0, "tonemap": 2},
       "outputFolder": "results",
       "scenes": [{"file": "teapot.flxs", "outputFile": "teapot", "settings": {"useAreaLight": False, "maxPathDepth": 4}},
                  {"file": "kitchen_standin.flxs", "outputFile": "kitchen_standin", "settings": view},
                  {"file": "kitchen_standin.flxs", "outputFile": "kitchen_standin_mk_64spp", "settings": dict({"useWavefront": False, "maxRenderTime": 0, "maxSpp": 64}, **view)}]}
bm = os.path.join(out, "b200.bm.json")
json.dump(job, open(bm, "w"), indent=1)
r = subprocess.run([os.path.join(ROOT, "fluctus_b200", "fluctus_headless"), "--benchmark-file", bm], capture_output=True, text=True)
open(os.path.join(out, "stdout.txt"), "w").write(r.stdout)
open(os.path.join(out, "stderr.txt"), "w").write(r.stderr[-4000:])
print(r.stdout)
print("exit", r.returncode)
for f in ("teapot.flxs", "kitchen_standin.flxs"):
    os.remove(os.path.join(out, f))
for f in os.listdir(os.path.join(out, "results")):
    if f.endswith(".pfm") or f.endswith(".ppm"):
        os.remove(os.path.join(out, "results", f))
