#!/usr/bin/env python
"""bench.py — the headline measurement of the wavefront path-tracing hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json metric "Msamples/s (1080p kitchen, 8 bounces)", configs[3]): the kitchen
stand-in scene (Country-Kitchen.obj is stripped from the reference snapshot, SURVEY.md §7.3.7;
fluctus_b200/standin.py rebuilds a room with the kitchen's material mix from the committed teapot
fixture), 1920x1080, 8 bounces, environment-map MIS + area light, NUM_TASKS = 2^22 paths per GPU.

One STEP = one frame: reset, ITERS wavefront iterations (logic+raygen -> materials -> ext+shadow+
bookkeeping: the reference's benchmark inner loop, src/tracer.cpp:737-778, in three launches), one NCCL
reduce of the float accumulation buffer into rank 0, tonemap there.  Per-GPU work is fixed as N
grows (each rank traces its own N paths with rank-offset seeds): weak scaling.
  value : Msamples/s, device-resident, CUDA events around K graph-replayed steps, max over ranks
          (the K steps are then repeated launch by launch with events around every stage: stage times, roofline)
  e2e   : the same frame through the public API with host buffers: RenderParams uploaded and
          the tonemapped frame + counters read back to pinned host memory inside the timed region
  roofline : the traversal launch (k_ext_shadow: extension closest-hit + shadow any-hit rays of an iteration),
          algorithmic bytes of SURVEY.md §8(d) / DESIGN.md
  cpu_baseline / --impl reference : the reference's own OpenCL kernels compiled as C++ (oracle/_ref, see
          oracle/refcl.py), OpenMP over work-items, on the host cores
  reference_opencl : the reference's untouched kernels on the SAME B200 through NVIDIA's OpenCL platform
          (oracle/clrun.py): the benchmark frame and the cfg5 ray batch, beside our numbers
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
WIDTH, HEIGHT = 1920, 1080
# wfBufferSize ("the amount of paths in flight simultaneously, limited by VRAM", clcontext.hpp:93; default 2^20 in
# settings.cpp:26).  On a 180 GB B200 the pool is 2^22 paths (0.87 GB of state + 0.15 GB of queues): the persistent
# traversal launch ends with a tail in which a few long rays keep the GPU waiting, and a 4x larger pool amortises
# it (measured: 2^20 145, 2^21 153, 2^22 157, 2^23 158 Msamples/s).  FLX_BENCH_TASKS overrides it for experiments.
NUM_TASKS = int(os.environ.get("FLX_BENCH_TASKS", 1 << 22))
BITSTACK = os.environ.get("FLX_BENCH_BITSTACK", "0") == "1"  # experiment switch: Settings::clUseBitstack traversal
MAX_BOUNCES = 8
ITERS_PER_STEP = 48          # wavefront iterations per frame
CPU_TASKS = 1 << 18          # path pool of the CPU arm (throughput per path does not depend on it)
CPU_ITERS_PER_STEP = 2
CPU_RAMP_ITERS = 14          # untimed iterations that bring the CPU pool to steady state
METRIC = "Msamples/s (1080p kitchen, 8 bounces)"
# algorithmic bytes of one extension-ray traversal
B_RAY_IN, B_RAY_OUT, B_INNER, B_TRI, B_HIT = 28, 56, 56, 40, 64
B_SHADOW_STATE = 36  # `shadow` state per ray (queue index, origin + length, direction, blocked flag)


def load_workload():
    from fluctus_b200 import scene as fs
    from fluctus_b200.standin import kitchen_standin
    teapot = fs.load_scene(os.path.join(GOLDEN, "teapot_sbvh.npz"))
    env = fs.load_envmap(os.path.join(GOLDEN, "night_env.npz"))
    scene, view = kitchen_standin(teapot)
    settings = dict(max_bounces=MAX_BOUNCES, use_env_map=True, use_area_light=True, sample_implicit=True,
                    sample_explicit=True, use_roulette=False, separate_queues=os.environ.get("FLX_BENCH_SEPARATE", "1") != "0", tonemap=2, **view)
    return teapot, scene, env, settings


def config_dict(n_gpus, scene, impl):
    return {"workload": "kitchen stand-in (%d tris, kitchen material mix, textured floor), 1920x1080, 8 bounces, "
                        "env-map MIS + area light" % len(scene["tris"]),
            "num_tasks_per_gpu": NUM_TASKS if impl == "ours" else CPU_TASKS,
            "iterations_per_step": ITERS_PER_STEP if impl == "ours" else CPU_ITERS_PER_STEP,
            "step": "one frame: reset + iterations + reduce to rank 0 + tonemap" if impl == "ours"
                    else "steady-state wavefront iterations on the CPU path pool",
            "parallelism": "sample-sharded x%d, one NCCL reduce of the accumulation buffer into rank 0 per frame" % n_gpus,
            "l2": "inputs larger than L2: %d MB path state + 33 MB frame per pass (L2 126 MB), no flush needed"
                  % ((NUM_TASKS if impl == "ours" else CPU_TASKS) * 212 // (1 << 20)),
            "stand_in_for": "assets/country_kitchen/Country-Kitchen.obj (stripped from the reference snapshot)"}


# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def window(self, t0, t1):
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        sm, mx, reasons, power = [], [], set(), []
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.15:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)), "samples": len(sm)}

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own kernels (src/*.cl compiled as C++ into oracle/_ref, oracle/refcl.py), work-items spread
# over the host cores with OpenMP; the oracle port stands in only where those objects were never built.
class CpuWavefront:
    def __init__(self, scene, env, settings):
        from oracle import pyoracle, refcl  # allowed here: cpu_baseline / --impl reference legs only
        from fluctus_b200 import abi, scene as fs
        # all the host cores this process may use (torchrun exports OMP_NUM_THREADS=1, which is not the machine)
        self.threads = max(pyoracle.max_threads(), len(os.sched_getaffinity(0)))
        try:
            refcl.program("wf_logic", refcl.Options(use_area_light=True, use_env_map=True, separate_queues=settings.get("separate_queues", True)))
            self.kind = "reference"
            self.wf = refcl.RefWavefront(CPU_TASKS, WIDTH, HEIGHT)
            self.wf.set_threads(self.threads)
        except RuntimeError:
            self.kind = "port"
            pyoracle.set_threads(self.threads)
            self.wf = pyoracle.OracleWavefront(CPU_TASKS, WIDTH, HEIGHT)
        self.wf.set_scene(scene)
        self.wf.set_envmap(env)
        p = abi.make_render_params(WIDTH, HEIGHT, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), **settings)
        self.wf.set_params(p.tobytes())
        self.wf.reset(); self.wf.clear_queues(); self.wf.set_pixel_index(0)
        self.wf.raygen(); self.wf.ext()
        self.first = True

    def iterate(self, n):
        """n passes of the benchmark inner loop (tracer.cpp:737-778); returns (samples, ext rays, shadow rays)."""
        wf, s, e, sh = self.wf, 0, 0, 0
        for _ in range(n):
            wf.logic(False); wf.raygen(); wf.materials()
            c = wf.counters().copy()
            wf.ext(); wf.shadow(); wf.clear_queues()
            wf.update_pixel_index(int(c["raygenQueue"]))
            if not self.first:
                s += int(c["raygenQueue"])
            self.first = False
            e += int(c["extensionQueue"]); sh += int(c["shadowQueue"])
        return s, e, sh


def cpu_baseline(scene, env, settings, budget_s=12.0):
    cw = CpuWavefront(scene, env, settings)
    cw.iterate(CPU_RAMP_ITERS)
    t0 = time.perf_counter()
    samples = rays = iters = 0
    while True:
        s, e, sh = cw.iterate(1)
        samples += s; rays += e + sh; iters += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or iters >= 64:
            break
    return {"value": samples / dt / 1e6, "unit": "Msamples/s", "cores": cw.threads, "kind": cw.kind,
            "mrays_per_s": rays / dt / 1e6,
            "sample": "%d steady-state wavefront iterations of a %d-path pool (after %d untimed ramp-up iterations), "
                      "same scene / resolution / params; %.1f s" % (iters, CPU_TASKS, CPU_RAMP_ITERS, dt)}


def run_reference_arm(args):
    """--impl reference: the reference's own kernels on the host cores (oracle/_ref: src/*.cl compiled as C++; the
    timing of the same kernels on the B200 through NVIDIA's OpenCL platform is the `reference_opencl` block of
    the other arm's line).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    teapot, scene, env, settings = load_workload()
    cw = CpuWavefront(scene, env, settings)
    cw.iterate(CPU_RAMP_ITERS)
    for _ in range(args.warmup):
        cw.iterate(CPU_ITERS_PER_STEP)
    t0 = time.perf_counter()
    samples = 0
    for _ in range(args.steps):
        samples += cw.iterate(CPU_ITERS_PER_STEP)[0]
    dt = time.perf_counter() - t0
    val = samples / dt / 1e6
    sample = ("%d steps x %d steady-state iterations of a %d-path pool after %d ramp-up iterations"
              % (args.steps, CPU_ITERS_PER_STEP, CPU_TASKS, CPU_RAMP_ITERS))
    emit({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.gpus, scene, "reference"),
        "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": cw.threads, "kind": cw.kind, "sample": sample},
        "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0})


def reference_opencl(scene, env, settings, ours_hit_i=None, rays=None, grid=None):
    """The reference's untouched kernels on this GPU through its OpenCL platform (oracle/clrun.py), after the timed
    regions: (1) the benchmark frame — reset + ITERS_PER_STEP iterations + tonemap + frame read-back, host loop as in
    Tracer::runBenchmarkFromFile (one blocking counter read per iteration) — with the benchmark's pool and queues and
    with the reference's defaults (wfBufferSize 2^20, single material queue); (2) the cfg5 ray batch through
    traceExtension, compared ray by ray with our result.  None where the box has no OpenCL device."""
    try:
        from oracle import clrun
        ok, what = clrun.available()
        if not ok:
            return {"unavailable": what}
        from fluctus_b200 import abi, scene as fs
        out = {"device": what, "build_options": " ".join(o for o in clrun.global_build_options("reference") if not o.startswith("/")).replace("-I ", "-I<reference>/src "),
               "frames": {}}
        for tag, n_tasks, sepq in (("benchmark settings (2^%d paths, separate queues)" % int(np.log2(NUM_TASKS)), NUM_TASKS, True),
                                   ("reference defaults (2^20 paths, single material queue)", 1 << 20, False)):
            st = dict(settings); st["separate_queues"] = sepq
            p = abi.make_render_params(WIDTH, HEIGHT, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), **st)
            cl = clrun.CLWavefront(n_tasks, WIDTH, HEIGHT, "reference")
            cl.set_scene(scene); cl.set_envmap(env); cl.set_params(p.tobytes())
            cl.render(4)
            best = None
            for _ in range(3):
                t0 = time.perf_counter()
                it, tot = cl.render(ITERS_PER_STEP)
                cl.postprocess()
                dt = time.perf_counter() - t0
                if best is None or dt < best[0]:
                    best = (dt, tot)
            cl.timing = True; cl.kernel_ms.clear()
            cl.render(ITERS_PER_STEP)
            km = {k: round(v[0], 2) for k, v in cl.kernel_ms.items()}
            cl.close()
            out["frames"][tag] = {"msamples_per_s": best[1]["samples"] / best[0] / 1e6, "frame_ms": best[0] * 1e3,
                                  "mrays_per_s": (best[1]["extension"] + best[1]["shadow"]) / best[0] / 1e6,
                                  "kernel_ms_per_frame": km, "kernel_ms_sum": round(sum(km.values()), 1)}
        if rays is not None:
            r = clrun.trace_closest(grid, rays, mode="reference")
            d = r["hit_i"] != ours_hit_i
            out["raycast"] = {"mrays_per_s": len(rays) / r["ms"] / 1e3, "kernel_ms": r["ms"], "rays": int(len(rays)),
                              "hit_miss_differs_from_ours": int(((r["hit_i"] >= 0) != (ours_hit_i >= 0)).sum()),
                              "hit_index_differs_from_ours": int(d.sum()),
                              "note": "the reference's bits are device-defined (fast-relaxed-math, FMA contraction, rcp.approx): "
                                      "tests/test_opencl_reference_gpu.py re-traces every differing ray with the IEEE build of the reference source"}
        return out
    except Exception as e:  # never let the bonus block cost the bench line
        return {"error": str(e)[-400:]}


# ---------------------------------------------------------------------------------------------------
# one B200 row per config.  After the timed regions, rank 0, one GPU.
def _block_relmse(img, ref, W, H, bs):
    def means(px):
        px = np.asarray(px, np.float32).reshape(H, W, 4)
        c = px[..., :3] / np.maximum(px[..., 3:], 1.0)
        return c[:H // bs * bs, :W // bs * bs].reshape(H // bs, bs, W // bs, bs, 3).mean((1, 3))
    a, b = means(img), means(ref)
    return float((((a - b) ** 2) / (b ** 2 + 1e-3)).mean())


def config_scenes(teapot):
    """cfg1 as BASELINE.json states it; cfg2 / cfg3 on stand-ins (luxball.obj and conference.obj are stripped from the
    reference snapshot) built from the reference's own assets in the scene cache, when it is there."""
    from fluctus_b200 import abi, scene as fs
    out = {"cfg1": dict(scene=teapot, W=512, H=512, N=262144, spp=16, oracle_spp=16, bs=4, stand_in_for=None,
                        what="assets/teapot.ply + env_maps/night.hdr, 512x512, 4 bounces, 16 spp, wfBufferSize 262144 (as stated)",
                        kw=dict(max_bounces=4, use_env_map=True, use_area_light=False))}
    cache = os.path.join(ROOT, "scenes", "_cache")
    try:
        base = fs.load_scene(os.path.join(cache, "psor_cube_sbvh.npz"))
        mats = np.tile(base["materials"][:1], 6)
        types = [abi.BXDF_DIFFUSE, abi.BXDF_GLOSSY, abi.BXDF_GGX_ROUGH_REFLECTION, abi.BXDF_IDEAL_REFLECTION,
                 abi.BXDF_GGX_ROUGH_DIELECTRIC, abi.BXDF_IDEAL_DIELECTRIC]
        for k, t in enumerate(types):
            mats[k]["type"] = t
            mats[k]["Kd"][:3] = (0.7, 0.45 + 0.05 * k, 0.3); mats[k]["Ks"][:3] = (0.9, 0.85, 0.8)
            mats[k]["Ns"], mats[k]["Ni"] = 0.12 + 0.04 * k, 1.5
        tris = base["tris"].copy()
        tris["matId"] = (np.arange(len(tris)) // 256) % 6
        scene = dict(base); scene.update(materials=mats, tris=tris)
        b = base["bounds"]; c = 0.5 * (b[0] + b[1])
        out["cfg2"] = dict(scene=scene, W=1024, H=1024, N=1 << 20, spp=64, oracle_spp=2, bs=16,
                           stand_in_for="assets/luxball/luxball.obj (stripped)",
                           what="psor-cube geometry (18 221 tris) with GGX rough reflection / rough dielectric, ideal dielectric, mirror, "
                                "glossy and diffuse materials, 1024x1024, 8 bounces, 64 spp, env map + area light",
                           kw=dict(max_bounces=8, use_env_map=True, use_area_light=True, separate_queues=True,
                                   cam_pos=(float(c[0]), float(c[1]), float(b[1][2] + 1.2 * (b[1][2] - b[0][2]))),
                                   light_pos=(float(c[0]) + 2.0, float(b[1][1]) + 1.0, float(c[2]))))
    except OSError:
        pass
    try:
        from fluctus_b200.standin import conference_standin
        room, view = conference_standin(fs.load_scene(os.path.join(cache, "egyptcat_sbvh.npz")))
        out["cfg3"] = dict(scene=room, W=1920, H=1080, N=1 << 22, spp=16, oracle_spp=1, bs=40,
                           stand_in_for="assets/conference/conference.obj (stripped)",
                           what="egyptcat.obj instanced 7x1x3 in a closed room (%d tris, textured glossy + diffuse), 1920x1080, 8 bounces, "
                                "area light only, 16 spp" % len(room["tris"]),
                           kw=dict(max_bounces=8, use_env_map=False, use_area_light=True, separate_queues=True, **view))
    except OSError:
        pass
    return out


def run_configs(dev, stream, teapot, env):
    """GPU: the config's full sample budget (CHECK_SPP stop rule of Tracer::runBenchmarkFromFile), timed with CUDA events
    after one warm-up render.  Oracle (checker): the same loop on the host cores with a bounded budget (oracle_spp), block
    means compared.  relMSE is dominated by the Monte-Carlo noise of the lower-spp side."""
    import torch
    import fluctus_b200 as fx
    from oracle import pyoracle
    threads = max(pyoracle.max_threads(), len(os.sched_getaffinity(0)))
    pyoracle.set_threads(threads)
    res = {}
    for name, cfg in config_scenes(teapot).items():
        W, H, N = cfg["W"], cfg["H"], cfg["N"]
        kw = dict(cfg["kw"], max_spp=cfg["spp"])
        tr = fx.Tracer(W, H, num_tasks=N, device=dev.index, stream=stream.cuda_stream, **kw)
        tr.init(cfg["scene"], env if kw.get("use_env_map") else None)
        tr.render(spp=cfg["spp"])                                     # warm-up (graph capture, caches)
        runs = []  # the budget is polled once per batch of iterations: whether a render needs n or n + 1 batches moves its time by
        for _ in range(3):  # 10-20 %, so three renders are timed and the median is reported (all three are kept in `runs_ms`)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            t_run = tr.render(spp=cfg["spp"])
            tr.clctx.enqueuePostprocessKernel()
            e1.record(stream)
            torch.cuda.synchronize(dev)
            runs.append((e0.elapsed_time(e1), t_run))
        ms, tot = sorted(runs, key=lambda r: r[0])[1]
        img = tr.pixels().reshape(-1, 4).copy()
        kw_o = dict(cfg["kw"], max_spp=cfg["oracle_spp"])
        from fluctus_b200 import abi, scene as fs
        p = abi.make_render_params(W, H, n_tris=len(cfg["scene"]["tris"]), world_radius=fs.world_radius(cfg["scene"]), **kw_o)
        wf = pyoracle.OracleWavefront(min(N, 1 << 19), W, H)
        wf.set_scene(cfg["scene"])
        if kw.get("use_env_map"):
            wf.set_envmap(env)
        wf.set_params(p.tobytes())
        t0, first, iters = time.perf_counter(), True, 0
        while iters < 4000:
            it, _ = wf.render(8, first)
            first = False; iters += it
            if it < 8 or (wf.spp() >= cfg["oracle_spp"]).all():
                break
        t_or = time.perf_counter() - t0
        ref = wf.pixels().copy()
        res[name] = {"workload": cfg["what"], "stand_in_for": cfg["stand_in_for"], "triangles": int(len(cfg["scene"]["tris"])),
                     "msamples_per_s": tot["splatted"] / (ms * 1e-3) / 1e6, "seconds": ms * 1e-3, "samples": int(tot["splatted"]),
                     "runs_ms": [round(r[0], 3) for r in runs], "iterations": int(tot["iterations"]), "mrays_per_s": (tot["extension"] + tot["shadow"]) / (ms * 1e-3) / 1e6,
                     "oracle": {"spp": cfg["oracle_spp"], "seconds": t_or, "cores": threads,
                                "msamples_per_s": float(ref[:, 3].sum()) / t_or / 1e6},
                     "relmse_vs_oracle": _block_relmse(img, ref, W, H, cfg["bs"]), "relmse_blocks": "%dx%d pixels" % (cfg["bs"], cfg["bs"])}
        tr.close()
    return res


def run_headless_cpp(scene, env, settings, n_gpus, spp_per_gpu=12):
    """The C++ host (fluctus_headless -> HeadlessTracer -> CudaContext -> the same C-ABI): the benchmark frame rendered by a
    subprocess of rank 0, sample-sharded over n_gpus devices inside that one process, frames combined by
    flx_reduce_frames over NVLink peer memory.  Reports the rate the binary prints (render + combine, without scene load)."""
    import subprocess as sp
    import tempfile
    from fluctus_b200 import abi, scene as fs
    exe = os.path.join(ROOT, "fluctus_b200", "fluctus_headless")
    if not os.path.exists(exe):
        return {"unavailable": "fluctus_headless not built"}
    tmp = tempfile.mkdtemp(prefix="flx_bench_")
    try:
        flxs, cfg, state, out = (os.path.join(tmp, n) for n in ("kitchen.flxs", "settings.json", "state.dat", "frame.pfm"))
        fs.save_flxs(flxs, scene, env)
        json.dump({"release": {"wfBufferSize": NUM_TASKS, "useAreaLight": True, "useEnvMap": True, "maxPathDepth": MAX_BOUNCES,
                               "useSeparateQueues": True, "tonemap": 2}}, open(cfg, "w"))
        p = abi.make_render_params(WIDTH, HEIGHT, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), **settings)
        abi.state_record(p).tofile(state)
        spp = spp_per_gpu * n_gpus  # 12 spp per GPU ~ the samples of one bench frame
        best = None
        for _ in range(2):
            r = sp.run([exe, "-x", str(WIDTH), "-y", str(HEIGHT), "-s", str(spp), "--settings", cfg, "--load-state", state,
                        "--gpus", str(n_gpus), "--out", out, flxs], capture_output=True, text=True, timeout=300)
            if r.returncode != 0:
                return {"error": (r.stderr or r.stdout)[-300:]}
            st = json.loads(r.stdout.strip().splitlines()[-1])
            if best is None or st["msamples_per_s"] > best["msamples_per_s"]:
                best = st
        res = {"msamples_per_s": best["msamples_per_s"], "seconds": best["seconds"], "samples": best["samples"], "gpus": best["gpus"],
               "spp": spp, "iterations": best["iterations"], "kernel_launches": best["kernel_launches"],
               "what": "fluctus_headless -x 1920 -y 1080 -s %d --gpus %d on the kitchen stand-in (.flxs), best of 2 processes; "
                       "time = render + frame combine + tonemap, scene load excluded" % (spp, n_gpus)}
        if n_gpus > 1:
            # the same job with ONE PROCESS PER GPU and NCCL from the C++ host (flx_comm_init / flx_reduce_pixels, no Python):
            # n_gpus copies of the binary, rank r on device r; rate = all ranks' samples / the slowest rank's time
            best = None
            for attempt in range(2):
                comm = os.path.join(tmp, "comm_%d.id" % attempt)
                procs = [sp.Popen([exe, "-x", str(WIDTH), "-y", str(HEIGHT), "-s", str(spp), "--settings", cfg, "--load-state", state,
                                   "--rank", str(r), "--world", str(n_gpus), "--comm-file", comm, "--out", out, flxs],
                                  stdout=sp.PIPE, stderr=sp.PIPE, text=True) for r in range(n_gpus)]
                lines, err = [], None
                for pr in procs:
                    try:
                        o, e = pr.communicate(timeout=300)
                    except sp.TimeoutExpired:
                        pr.kill(); o, e = pr.communicate(); err = "timeout"
                    if pr.returncode != 0:
                        err = (e or o)[-300:]
                    else:
                        lines.append(json.loads(o.strip().splitlines()[-1]))
                if err:
                    res["one_process_per_gpu"] = {"error": err}
                    break
                sec = max(l["seconds"] for l in lines)
                tot = sum(l["samples"] for l in lines)
                cand = {"msamples_per_s": round(tot / sec * 1e-6, 3), "seconds": sec, "samples": tot, "world": n_gpus, "nccl": lines[0]["nccl"],
                        "what": "%d x fluctus_headless --rank r --world %d: ncclReduce of the accumulation buffers to rank 0 on the render "
                                "stream (flx_reduce_pixels), best of 2" % (n_gpus, n_gpus)}
                if best is None or cand["msamples_per_s"] > best["msamples_per_s"]:
                    best = cand
            if best is not None and "one_process_per_gpu" not in res:
                res["one_process_per_gpu"] = best
        return res
    except Exception as e:
        return {"error": str(e)[-300:]}
    finally:
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)


# ---------------------------------------------------------------------------------------------------
# GPU arm
TRAVERSAL_SOURCES = ("flx_trace.cu", "flx_bvh.cuh", "flx_math.cuh", "flx_internal.h")


def kernel_source_sha1():
    """Hash of the sources the traversal kernels are built from: profiles/traffic.json records it when a capture is
    taken (tools/profile_round.sh), and a capture whose hash differs from the tree's is refused (.git does not travel)."""
    import hashlib
    h = hashlib.sha1()
    for name in TRAVERSAL_SOURCES:
        with open(os.path.join(ROOT, "fluctus_b200", "csrc", name), "rb") as f:
            h.update(f.read())
    with open(os.path.join(ROOT, "fluctus_b200", "csrc", "Makefile")) as f:  # the compiler flags, not the file list
        h.update("".join(l for l in f if l.startswith("NVFLAGS") or l.startswith("  -Xcompiler")).encode())
    return h.hexdigest()[:12]


def recorded_traffic(which=None):
    """DRAM bytes per launch from the committed `ncu --set full` capture (profiles/traffic.json): the traversal launch of
    the render (which=None) or the cfg5 ray-cast kernel (which="raycast").  Returns (GB or None, dict of provenance);
    the figure is RECORDED, not measured in this run, and is dropped when the kernels have changed since the capture."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except (OSError, ValueError):
        return None, {"traffic_state": "no capture"}
    rec = t.get("raycast") if which == "raycast" else t
    if not rec:
        return None, {"traffic_state": "no capture"}
    info = {"traffic_source": rec.get("source"), "traffic_kernel_sha1": t.get("kernel_source_sha1"), "traffic_kind": "recorded"}
    if t.get("kernel_source_sha1") != kernel_source_sha1():
        info["traffic_state"] = "stale: the traversal sources changed since the capture (now %s)" % kernel_source_sha1()
        return None, info
    if which is None and int(t.get("num_tasks", 0)) != NUM_TASKS:
        info["traffic_state"] = "capture taken with another pool size"
        return None, info
    info["traffic_state"] = "current"
    info["ncu"] = {k: rec[k] for k in ("issue_slots_busy_pct", "lanes_per_instruction", "l1_hit_pct", "l2_hit_pct") if k in rec}
    return rec["dram_bytes_per_launch"] / 1e9, info


def traversal_bytes(cnt, rays=None, any_hit=False):
    """Algorithmic bytes of SURVEY.md §8(d) for `rays` rays with the per-ray visit statistics of cnt.
    any_hit: shadow rays — 36 B of path state per ray (queue index, origin + length, direction, result word), no
    shading fetch."""
    r = max(int(cnt["rays"]), 1)
    if any_hit:
        per_ray = B_SHADOW_STATE + (cnt["inner"] * B_INNER + cnt["tri"] * B_TRI) / r
    else:
        per_ray = (B_RAY_IN + B_RAY_OUT) + (cnt["inner"] * B_INNER + cnt["tri"] * B_TRI + cnt["hits"] * B_HIT) / r
    return per_ray * (r if rays is None else rays), per_ray


def run_raycast(dev, rank, world, teapot, peak_gbs, steps, warmup, check):
    """cfg5: 2^24 incoherent rays against the 15x15x14 replicated-teapot SBVH (10.1 M triangles)."""
    import torch
    from fluctus_b200 import CudaContext, scene as fs
    from fluctus_b200.context import STAGE_TRACE_BATCH
    from fluctus_b200.multi_gpu import split_rays
    t_build = time.perf_counter()
    grid = fs.replicate_grid(teapot, (15, 15, 14), (2.2, 1.1, 1.4))
    n_total = 1 << 24
    first, n = split_rays(n_total, rank, world)
    rays = fs.bench_rays(n, grid["bounds"], first=first)
    stream = torch.cuda.current_stream(dev)
    ctx = CudaContext(num_tasks=1024, device=dev.index, stream=stream.cuda_stream)
    ctx.uploadSceneData(grid)
    ctx.setTraversal(BITSTACK)
    t_build = time.perf_counter() - t_build
    d_rays = torch.from_numpy(rays).to(dev)
    d_i = torch.empty(n, dtype=torch.int32, device=dev)
    d_t = torch.empty(n, dtype=torch.float32, device=dev)
    h_rays = torch.from_numpy(rays).pin_memory()
    h_i = torch.empty(n, dtype=torch.int32).pin_memory()
    h_t = torch.empty(n, dtype=torch.float32).pin_memory()

    def device_pass():
        ctx.traceClosestDevice(d_rays.data_ptr(), n, d_i.data_ptr(), d_t.data_ptr())

    def host_pass():  # public host-buffer entry: H2D rays, trace, D2H hit index + distance, synchronises
        ctx._check(ctx.L.flx_trace_closest(ctx.h, h_rays.data_ptr(), n, h_i.data_ptr(), h_t.data_ptr(), None))

    out = {}
    for name, fn in (("device", device_pass), ("host", host_pass)):
        for _ in range(max(warmup, 3)):
            fn()
        torch.cuda.synchronize(dev)
        if world > 1:
            torch.distributed.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if name == "device":
            ctx.enableTiming(True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        out[name] = float(ms.item()) / steps
        if name == "device":
            k_ms, k_n = ctx.kernelTime(STAGE_TRACE_BATCH)
            ctx.enableTiming(False)
            out["kernel_ms"] = k_ms / max(k_n, 1)
    # visit statistics from the instrumented kernel variant (untimed)
    ctx.setTraceCounting(True)
    device_pass()
    cnt = ctx.traceCounters()["batch_closest"]
    ctx.setTraceCounting(False)
    # microbench variant of the §8(d) formula: 32 B in, 8 B out per ray
    per_ray = 40 + (cnt["inner"] * B_INNER + cnt["tri"] * B_TRI) / max(cnt["rays"], 1)
    achieved = per_ray * n / (out["kernel_ms"] * 1e-3) / 1e9
    res = {"metric": "incoherent Mrays/s (2^24 rays vs replicated-teapot SBVH, closest hit)",
           "value": n_total / (out["device"] * 1e-3) / 1e6, "unit": "Mrays/s",
           "e2e": {"value": n_total / (out["host"] * 1e-3) / 1e6, "unit": "Mrays/s",
                   "h2d_bytes_per_step": 32 * n, "d2h_bytes_per_step": 8 * n},
           "rays": n_total, "triangles": int(len(grid["tris"])), "nodes": int(len(grid["nodes"])),
           "scene_build_upload_s": round(t_build, 1),
           "per_ray": {"inner_nodes": cnt["inner"] / max(cnt["rays"], 1), "triangle_tests": cnt["tri"] / max(cnt["rays"], 1),
                       "hit_fraction": cnt["hits"] / max(cnt["rays"], 1), "algorithmic_bytes": per_ray},
           "roofline": {"bound": "hbm", "kernel": "k_trace_batch", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                        "frac": achieved / peak_gbs, "traffic": None, "kernel_ms": out["kernel_ms"]}}
    rc_gb, rc_info = recorded_traffic("raycast")
    res["roofline"].update({"traffic": rc_gb, "traffic_unit": "GB per launch (DRAM read + write)"}, **rc_info)
    if rc_gb:
        res["roofline"]["dram_frac"] = rc_gb / (out["kernel_ms"] * 1e-3) / peak_gbs
    keep = None
    if check and rank == 0:  # parity sample against the oracle (checker only; tests/ hold the full 2^24 comparison) + CPU rate
        from oracle import pyoracle, refcl
        m = 1 << 20
        threads = max(pyoracle.max_threads(), len(os.sched_getaffinity(0)))
        pyoracle.set_threads(threads)
        ref = pyoracle.trace_closest(grid, rays[:m], 0)
        gi, gt = d_i.cpu().numpy(), d_t.cpu().numpy()
        res["parity"] = {"rays_checked": m, "hit_i_equal": bool(np.array_equal(gi[:m], ref["hit_i"])),
                         "hit_t_bits_equal": bool(np.array_equal(gt[:m].view(np.uint32), ref["hit_t"].view(np.uint32)))}
        try:  # the reference's bvh_intersect on all host cores, beside the GPU figure
            t0 = time.perf_counter()
            refcl.trace_closest(grid, rays[:m], threads=threads)
            dt = time.perf_counter() - t0
            res["cpu_baseline"] = {"value": m / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference",
                                   "sample": "the first 2^20 rays of the batch through the reference's bvh_intersect (src/bvh.cl compiled as C++), %.1f s" % dt}
        except RuntimeError as e:
            res["cpu_baseline"] = {"unavailable": str(e)[:200]}
        keep = (grid, rays, gi)
    ctx.close()
    return res, keep


def run_lbvh(dev, stream, teapot, scene, env, settings, steps=3):
    """The hierarchy built on the GPU (flx_build_lbvh: Morton order + radix tree, emitted in the reference's Node[] format)
    in place of the reference's SBVH: build time, the benchmark frame and the cfg5 ray batch on it.  Same kernels, same
    timing method as the headline numbers; only the tree differs (rank 0, N=1)."""
    import torch
    import fluctus_b200 as fx
    from fluctus_b200 import CudaContext, scene as fs
    out = {"max_leaf": 4}
    # ---- the benchmark frame on the kitchen stand-in
    fs.build_lbvh(scene, max_leaf=4, device=dev.index)                      # warm-up: context, cub temp sizes
    t0 = time.perf_counter()
    lb = fs.build_lbvh(scene, max_leaf=4, device=dev.index)
    out["kitchen"] = {"triangles": int(len(scene["tris"])), "build_s": round(time.perf_counter() - t0, 4),
                      "nodes": int(len(lb["nodes"])), "depth": lb["lbvh_depth"], "sbvh_nodes": int(len(scene["nodes"]))}
    tr = fx.Tracer(WIDTH, HEIGHT, num_tasks=NUM_TASKS, device=dev.index, stream=stream.cuda_stream, **settings)
    tr.init(lb, env)
    tr.clctx.setTraversal(BITSTACK)
    c = tr.clctx

    def frame():
        c.setPixelIndex(0)
        c.render(0, reset=True)
        c.render(ITERS_PER_STEP)
        c.enqueuePostprocessKernel()
        return c.getRenderTotals()
    for _ in range(2):
        frame()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    e0.record(stream)
    samples = 0
    for _ in range(steps):
        samples += frame()["samples"]
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    out["kitchen"].update({"msamples_per_s": samples / ms / 1e3, "ms_per_step": ms / steps, "steps": steps})
    tr.close()
    # ---- cfg5: 2^24 incoherent rays, 10.1 M triangles
    grid = fs.replicate_grid(teapot, (15, 15, 14), (2.2, 1.1, 1.4))
    t0 = time.perf_counter()
    glb = fs.build_lbvh(grid, max_leaf=4, device=dev.index)
    t_build = time.perf_counter() - t0
    n = 1 << 24
    rays = fs.bench_rays(n, grid["bounds"])
    ctx = CudaContext(num_tasks=1024, device=dev.index, stream=stream.cuda_stream)
    ctx.uploadSceneData(glb)
    ctx.setTraversal(BITSTACK and glb["lbvh_depth"] < 31)
    d_rays = torch.from_numpy(rays).to(dev)
    d_i = torch.empty(n, dtype=torch.int32, device=dev)
    d_t = torch.empty(n, dtype=torch.float32, device=dev)
    for _ in range(3):
        ctx.traceClosestDevice(d_rays.data_ptr(), n, d_i.data_ptr(), d_t.data_ptr())
    torch.cuda.synchronize(dev)
    e0.record(stream)
    for _ in range(steps):
        ctx.traceClosestDevice(d_rays.data_ptr(), n, d_i.data_ptr(), d_t.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    ctx.setTraceCounting(True)
    ctx.traceClosestDevice(d_rays.data_ptr(), n, d_i.data_ptr(), d_t.data_ptr())
    cnt = ctx.traceCounters()["batch_closest"]
    ctx.setTraceCounting(False)
    out["cfg5"] = {"triangles": int(len(grid["tris"])), "build_s": round(t_build, 3), "nodes": int(len(glb["nodes"])),
                   "depth": glb["lbvh_depth"], "sbvh_nodes": int(len(grid["nodes"])), "mrays_per_s": n / ms / 1e3, "ms": ms,
                   "per_ray": {"inner_nodes": cnt["inner"] / max(cnt["rays"], 1), "triangle_tests": cnt["tri"] / max(cnt["rays"], 1),
                               "hit_fraction": cnt["hits"] / max(cnt["rays"], 1)}}
    ctx.close()
    out["what"] = ("flx_build_lbvh (max_leaf 4) replaces the scene's SBVH; build_s = host call incl. H2D of the triangles and D2H of "
                   "the tree; frame and ray batch timed as the headline numbers (CUDA events, device-resident inputs)")
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    import fluctus_b200 as fx
    from fluctus_b200.context import BUF_PREVIEW, STAGE_EXT, STAGE_SHADOW, STAGE_LOGIC, STAGE_RAYGEN, STAGE_MATERIALS
    from fluctus_b200.multi_gpu import pixels_tensor, reduce_pixels

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this backend has no CPU fallback (use --impl reference for the CPU arm)")
    fx.load()  # fails loudly when libfluctus_cuda.so is missing
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)  # the context, torch events and NCCL all order on this stream

    teapot, scene, env, settings = load_workload()
    tr = fx.Tracer(WIDTH, HEIGHT, num_tasks=NUM_TASKS, device=local, rank=rank, world_size=world,
                   stream=stream.cuda_stream, **settings)
    tr.init(scene, env)
    tr.clctx.setTraversal(BITSTACK)
    c = tr.clctx
    pix = pixels_tensor(c, dev)
    # The frame combine is the library's own collective (flx_comm_init / flx_reduce_pixels: ncclReduce from the C-ABI on
    # the context's stream); torch.distributed carries the communicator id and the timing scalars.  If the library cannot
    # set its communicator up the torch collective over the zero-copy view takes over, and the line says so.
    collective = "none (1 GPU)"
    if world > 1:
        try:
            uid = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                uid.copy_(torch.frombuffer(bytearray(fx.CudaContext.commUniqueId()), dtype=torch.uint8))
            dist.broadcast(uid, src=0)
            c.commInit(rank, world, bytes(uid.cpu().numpy().tobytes()))
            collective = "flx_reduce_pixels (ncclReduce from the C-ABI, NCCL %d)" % c.commInfo()["nccl_version"]
        except Exception as e:  # noqa: BLE001
            collective = "torch.distributed.reduce (flx_comm_init failed: %s)" % str(e)[:120]
    own_comm = collective.startswith("flx_")

    def combine():
        if own_comm:
            c.reducePixels(0)
        else:
            reduce_pixels(pix, dst=0)
    cursor0 = tr._cursor_start()
    n_px = WIDTH * HEIGHT
    h_frame = torch.empty((n_px, 4), dtype=torch.float32).pin_memory()
    h_frame_np = h_frame.numpy()

    def frame(host_io):
        if host_io:
            c.updateParams(tr.params)              # the frame's inputs: 256-byte RenderParams from the host
        c.setPixelIndex(0)
        c.render(0, reset=True)                    # reset + clear + priming raygen + ext
        c.setPixelIndex(cursor0)
        c.render(ITERS_PER_STEP)
        tot = None
        if world > 1:
            tot = c.getRenderTotals()              # this rank's totals (synchronises; the reduce below moves pixels only)
            combine()                              # ONE NCCL reduce of the accumulation buffer per frame, into the root
        if rank == 0:                              # the frame exists on the root only: it alone tonemaps and reads back,
            c.enqueuePostprocessKernel()           # the other ranks go straight on to the next frame's reset
            if host_io:
                c.readPixels(BUF_PREVIEW, out=h_frame_np)  # tonemapped frame -> pinned host memory (synchronises)
        return tot if tot is not None else c.getRenderTotals()

    def timed(host_io, steps, with_events):
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        if with_events:
            c.enableTiming(True)
        l0 = c.launchCount()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record(stream)
        tot = dict(samples=0, extension=0, shadow=0, primary=0)
        for _ in range(steps):
            t = frame(host_io)
            for k in tot:
                tot[k] += t[k]
        e1.record(stream)
        torch.cuda.synchronize(dev)
        t1 = time.time()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        agg = torch.tensor([tot["samples"], tot["extension"], tot["shadow"], tot["primary"]], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        stage = None
        if with_events:
            stage = {n: c.kernelTime(s) for n, s in (("ext", STAGE_EXT), ("shadow", STAGE_SHADOW), ("logic", STAGE_LOGIC),
                                                     ("raygen", STAGE_RAYGEN), ("materials", STAGE_MATERIALS))}
            c.enableTiming(False)
        return dict(ms=float(ms.item()), samples=float(agg[0]), ext=float(agg[1]), shadow=float(agg[2]), primary=float(agg[3]),
                    local=tot, launches=c.launchCount() - l0, stage=stage, t0=t0, t1=t1)

    def strong_scaling(total_spp=256):
        """cfg4 as a FIXED budget: 256 spp at 1080p split spp/G over the ranks (shard_plan; the reference's stop rule
        counts regenerated paths, src/tracer.cpp:713,773), one reduce into rank 0, tonemap.  Time to completion, max
        over ranks; work still in flight when a rank reaches its share is the cost of the split."""
        from fluctus_b200.multi_gpu import shard_plan
        share = shard_plan(rank, world, NUM_TASKS, n_px, spp=total_spp)["spp"]
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        tot = tr.render(spp=share, batch=8) if share > 0 else c.getRenderTotals()
        if world > 1:
            combine()
        if rank == 0:
            c.enqueuePostprocessKernel()
        e1.record(stream)
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        done = torch.tensor([float(tot["samples"])], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(done, op=dist.ReduceOp.SUM)
        sec = float(ms.item()) * 1e-3
        return {"spp": total_spp, "spp_per_gpu": share, "seconds": sec, "budget_samples": total_spp * n_px,
                "msamples_per_s": total_spp * n_px / sec / 1e6, "samples_traced": float(done.item()),
                "scaling": "strong", "what": "time to %d spp at 1920x1080 with the budget split spp/G, one reduce, tonemap on rank 0; "
                                             "max over ranks" % total_spp}

    clocks = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 1)):
        frame(False)
    v = timed(False, args.steps, False)         # THE timed region: device-resident frames, CUDA-graph replay
    r = timed(False, args.steps, True)          # same K frames again, launch by launch, CUDA events around every stage
    for _ in range(2):
        frame(True)
    h = timed(True, args.steps, False)          # end to end through the host-buffer API (graph replay)
    clk = clocks.window(v["t0"], r["t1"]) if clocks else None
    if clocks:
        clocks.stop()

    # visit statistics of the extension-ray kernel from the instrumented variant: one untimed frame
    c.setTraceCounting(True)
    frame(False)
    cnt = c.traceCounters()
    c.setTraceCounting(False)
    peak, peak_src = measured_peak_gbs()
    # the render entry traces the extension and the shadow rays of an iteration in ONE persistent launch
    # (k_ext_shadow); its events are filed under the extension stage
    ext_ms, ext_n = r["stage"]["ext"]
    ext_bytes, ext_per_ray = traversal_bytes(cnt["ext"], rays=r["local"]["extension"])
    shw_bytes, shw_per_ray = traversal_bytes(cnt["shadow"], rays=r["local"]["shadow"], any_hit=True)
    trace_achieved = (ext_bytes + shw_bytes) / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
    stage_ms = {("trace" if k == "ext" else k): s[0] for k, s in r["stage"].items() if k not in ("shadow", "raygen") or s[0] > 0}
    stage_sum = sum(stage_ms.values())
    traffic_gb, traffic_info = recorded_traffic()
    # whole-step roofline (SURVEY.md §8d "Msamples/s roofline fraction = sum of stage bytes / t / peak"), per rank, from the
    # device totals of the timed region: logic reads 100 B of every path every iteration, + 76 B when a NEE sample
    # resolves (~ one per shadow ray), + 56 B on terminate, + 292 B on continue; raygen 188 B per restarted path;
    # materials 272 B per shaded path (every extension ray that is not a restart); traversal as above; reset, tonemap.
    lt = v["local"]
    shaded = max(lt["extension"] - lt["primary"], 0)
    ext_b, _ = traversal_bytes(cnt["ext"], rays=lt["extension"])
    shw_b, _ = traversal_bytes(cnt["shadow"], rays=lt["shadow"], any_hit=True)
    stage_bytes = {
        "logic": 100.0 * NUM_TASKS * ITERS_PER_STEP * args.steps + 76.0 * lt["shadow"] + 56.0 * lt["primary"] + 292.0 * shaded,
        "raygen": 188.0 * lt["primary"], "materials": 272.0 * shaded, "trace": ext_b + shw_b,
        "reset_tonemap": args.steps * (152.0 * NUM_TASKS + (52.0 + 32.0) * n_px)}
    step_bytes = sum(stage_bytes.values())
    step_gbs = step_bytes / (v["ms"] * 1e-3) / 1e9

    strong = strong_scaling() if not args.no_strong_scaling else None
    raycast = None
    if not args.no_raycast:
        raycast, keep = run_raycast(dev, rank, world, teapot, peak, max(args.steps // 2, 3), args.warmup, check=(world == 1))
    else:
        keep = None
    cpu = ref_cl = configs = headless = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(scene, env, settings)
    if rank == 0 and world == 1 and not args.no_configs:
        try:
            configs = run_configs(dev, stream, teapot, env)
        except Exception as e:  # a bonus block never costs the bench line
            configs = {"error": str(e)[-300:]}
    lbvh = None
    if rank == 0 and world == 1 and not args.no_lbvh:
        try:
            lbvh = run_lbvh(dev, stream, teapot, scene, env, settings)
        except Exception as e:
            lbvh = {"error": str(e)[-300:]}
    if world > 1 and not args.no_headless:
        # Every rank's timed work is over before rank 0 hands all GPUs to the C++ host — and the other ranks wait on the
        # CPU (gloo): an NCCL barrier is a kernel spinning on their GPUs, which the subprocess would have to share
        # (measured: 158 instead of 353 Msamples/s on 2 GPUs).
        torch.cuda.synchronize(dev)
        cpu_group = dist.new_group(backend="gloo")
        dist.barrier(group=cpu_group)
    if rank == 0 and not args.no_headless:
        headless = run_headless_cpp(scene, env, settings, world)
    if world > 1 and not args.no_headless:
        dist.barrier(group=cpu_group)
    if rank == 0 and world == 1 and not args.no_reference_opencl:
        tr.close()  # free our pools first: the reference's 2^22-path state is another 1.1 GB + the cfg5 tree in its own layout
        tr = None
        if keep is not None:
            ref_cl = reference_opencl(scene, env, settings, ours_hit_i=keep[2], rays=keep[1], grid=keep[0])
        else:
            ref_cl = reference_opencl(scene, env, settings)

    if rank == 0:
        line = {
            "metric": METRIC, "value": v["samples"] / (v["ms"] * 1e-3) / 1e6, "unit": "Msamples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": v["ms"] / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(config_dict(world, scene, "ours"), collective=collective),
            "mrays_per_s": {"extension": v["ext"] / (v["ms"] * 1e-3) / 1e6, "shadow": v["shadow"] / (v["ms"] * 1e-3) / 1e6,
                            "total": (v["ext"] + v["shadow"]) / (v["ms"] * 1e-3) / 1e6},
            "e2e": {"value": h["samples"] / (h["ms"] * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": h["ms"] / args.steps,
                    "h2d_bytes_per_step": 256, "d2h_bytes_per_step": n_px * 16 + 48},
            "gpu_launches": int(v["launches"]),
            "roofline": dict({
                "bound": "issue", "bound_per_contract": "hbm",
                "kernel": "k_ext_shadow (extension closest-hit + shadow any-hit traversal of one iteration, one persistent launch)",
                "achieved": trace_achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": trace_achieved / peak,
                "frac_is": "ALGORITHMIC bytes (SURVEY.md 8d: every node / triangle fetch counted, cache hits not subtracted) / kernel time / HBM "
                           "peak.  This scene's tree (37 MB) is served from L1/L2, so the figure can exceed 1 and is NOT an HBM utilisation: "
                           "dram_frac is.  The kernel is bound by instruction issue at partial lane occupancy (ncu block); the HBM-facing "
                           "traversal figure of the metric is raycast.roofline",
                "traffic": traffic_gb, "traffic_unit": "GB per launch (DRAM read + write)",
                "dram_frac": (traffic_gb / (ext_ms / max(ext_n, 1) * 1e-3) / peak) if traffic_gb and ext_ms > 0 else None,
                "algorithmic_gb_per_launch": (ext_bytes + shw_bytes) / max(ext_n, 1) / 1e9,
                "launches": int(ext_n), "avg_launch_ms": ext_ms / max(ext_n, 1),
                "algorithmic_bytes_per_ray": {"extension": ext_per_ray, "shadow": shw_per_ray},
                "rays_per_launch": {"extension": r["local"]["extension"] / max(ext_n, 1), "shadow": r["local"]["shadow"] / max(ext_n, 1)},
                "share_of_step": ext_ms / stage_sum if stage_sum else None,
                "timed_in": "the same K frames repeated launch by launch right after the graph-replayed region, "
                            "CUDA events around every stage launch on the context's stream"}, **traffic_info),
            "step_roofline": {"bound": "hbm", "achieved": step_gbs, "peak": peak, "unit": "GB/s", "frac": step_gbs / peak,
                              "what": "sum of the stages' algorithmic bytes (SURVEY.md 8d) of the timed region on one rank / its time",
                              "gb_per_step": {k: b / args.steps / 1e9 for k, b in stage_bytes.items()}},
            "stage_ms_per_step": {k: t / args.steps for k, t in stage_ms.items()},
            "events_pass_ms_per_step": r["ms"] / args.steps,
            "strong_scaling": strong, "configs": configs, "lbvh": lbvh, "headless_cpp": headless,
            "clocks": clk, "cpu_baseline": cpu, "raycast": raycast, "reference_opencl": ref_cl,
        }
        emit(line)
    if tr is not None:
        tr.close()
    if world > 1:
        dist.destroy_process_group()


_RESULT_FD = None


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--no-raycast", action="store_true", help="skip the cfg5 ray-cast microbench")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reference-opencl", action="store_true", help="skip timing the reference's kernels through OpenCL")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg1/cfg2/cfg3 rows (BASELINE.md 4)")
    ap.add_argument("--no-headless", action="store_true", help="skip the C++ host (fluctus_headless) leg")
    ap.add_argument("--no-strong-scaling", action="store_true", help="skip the fixed 256-spp budget record")
    ap.add_argument("--no-lbvh", action="store_true", help="skip the GPU-built hierarchy (flx_build_lbvh) rates")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON result: libraries that write to fd 1 on their own (NCCL prints its
    # version banner there when NCCL_DEBUG is set) are sent to stderr for the duration of the run.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
