"""Run on the GPU box: what the reference's untouched OpenCL kernels do on the same B200, next to the CUDA library.
  A  cfg5 at full size (2^24 rays, 10.1 M triangles): hit index / hit-miss / distance, CUDA vs reference kernel; Mrays/s
  B  the benchmark frame (kitchen stand-in, 1080p, 8 bounces) through the reference's kernels: Msamples/s, per-kernel ms
  C  stage-by-stage state, CUDA vs reference kernels, both fed the oracle's pre-stage state: distribution of differences
Writes gpurun_out/cl_evidence.json (+ .txt).  Evidence run; assertions live in tests/test_opencl_reference_gpu.py."""
import json
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
txt = open(os.path.join(ROOT, "gpurun_out", "cl_evidence.txt"), "w")
RES = {}


def say(*a):
    s = " ".join(str(x) for x in a)
    print(s); txt.write(s + "\n"); txt.flush()


def step(name):
    def deco(fn):
        say("==", name)
        t = time.time()
        try:
            fn()
        except Exception:
            say("FAILED:", traceback.format_exc()[-3000:])
        say("   (%.1f s)" % (time.time() - t))
        json.dump(RES, open(os.path.join(ROOT, "gpurun_out", "cl_evidence.json"), "w"), indent=1)
    return deco


from oracle import clrun, pyoracle
from fluctus_b200 import scene as fs, abi
import fluctus_b200 as fx

ok, what = clrun.available()
say("OpenCL:", ok, what)
RES["opencl"] = what
if not ok:
    sys.exit(0)
WHICH = set(sys.argv[1:]) or {"A", "B", "C"}
teapot = fs.load_scene(os.path.join(ROOT, "tests", "golden", "teapot_sbvh.npz"))
env = fs.load_envmap(os.path.join(ROOT, "tests", "golden", "night_env.npz"))


if "A" in WHICH:
    @step("A: cfg5, 2^24 incoherent rays vs the 10.1 M-triangle replicated-teapot SBVH")
    def _():
        grid = fs.replicate_grid(teapot, (15, 15, 14), (2.2, 1.1, 1.4))
        n = 1 << 24
        rays = fs.bench_rays(n, grid["bounds"])
        ctx = fx.CudaContext(num_tasks=1024)
        ctx.uploadSceneData(grid)
        ci, ct = ctx.traceClosest(rays)
        ctx.close()
        out = {"rays": n, "triangles": int(len(grid["tris"]))}
        for mode in ("reference", "strict"):
            r = clrun.trace_closest(grid, rays, mode=mode)
            di = r["hit_i"] != ci
            hit = (ci >= 0) & ~di
            rel = np.abs(r["hit_t"][hit].astype(np.float64) - ct[hit]) / np.abs(ct[hit])
            o = {"hit_i_differs": int(di.sum()), "hit_miss_differs": int(((r["hit_i"] >= 0) != (ci >= 0)).sum()),
                 "hit_t_bits_differ": int((r["hit_t"].view(np.uint32) != ct.view(np.uint32)).sum()),
                 "hit_t_max_rel": float(rel.max()), "hit_t_rel_above_1e-5": int((rel > 1e-5).sum()),
                 "kernel_ms": r["ms"], "mrays_per_s": n / r["ms"] / 1e3, "hit_fraction": float((ci >= 0).mean())}
            if di.any():
                # a differing index must be a near-tie: the triangle the reference kernel reports lies (within its own
                # arithmetic's error) at the distance of the triangle CUDA reports
                k = np.nonzero(di)[0]
                both = (r["hit_i"][k] >= 0) & (ci[k] >= 0)
                gap = np.abs(r["hit_t"][k][both].astype(np.float64) - ct[k][both]) / np.abs(ct[k][both])
                o["differing_rays_rel_gap_max"] = float(gap.max()) if len(gap) else None
                o["differing_rays_one_side_miss"] = int((~both).sum())
                o["first_differing"] = [int(x) for x in k[:8]]
            out[mode] = o
            say("  mode %-10s" % mode, o)
        RES["cfg5"] = out


def kitchen():
    from fluctus_b200.standin import kitchen_standin
    scene, view = kitchen_standin(teapot)
    settings = dict(max_bounces=8, use_env_map=True, use_area_light=True, sample_implicit=True, sample_explicit=True,
                    use_roulette=False, tonemap=2, **view)
    return scene, settings


if "B" in WHICH:
    @step("B: the benchmark frame through the reference's OpenCL kernels")
    def _():
        scene, settings = kitchen()
        W, H = 1920, 1080
        out = {}
        for N, sepq in ((1 << 20, False), (1 << 20, True), (1 << 22, True)):
            p = abi.make_render_params(W, H, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), separate_queues=sepq, **settings)
            cl = clrun.CLWavefront(N, W, H, "reference")
            cl.set_scene(scene); cl.set_envmap(env); cl.set_params(p.tobytes())
            cl.render(8)                                   # builds every kernel, warms up
            cl.rt.finish()
            t0 = time.perf_counter()
            it, st = cl.render(48)                         # one frame of bench.py: reset + 48 iterations ...
            cl.postprocess()                               # ... + tonemap + read-back of the frame
            dt = time.perf_counter() - t0
            cl.timing = True
            cl.kernel_ms.clear()
            cl.render(48)
            km = {k: round(v[0], 3) for k, v in cl.kernel_ms.items()}
            cl.timing = False
            key = "N=2^%d %s" % (int(np.log2(N)), "separate queues" if sepq else "single queue (reference default)")
            out[key] = {"msamples_per_s": st["samples"] / dt / 1e6, "frame_ms": dt * 1e3, "samples": st["samples"],
                        "mrays_per_s": (st["extension"] + st["shadow"]) / dt / 1e6, "kernel_ms_per_frame": km,
                        "kernel_ms_sum": round(sum(km.values()), 2)}
            say("  ", key, out[key])
            cl.close()
        RES["kitchen_opencl"] = out


if "C" in WHICH:
    @step("C: per-stage state, CUDA vs the reference's kernels (same pre-stage state from the oracle)")
    def _():
        from fluctus_b200.standin import kitchen_standin
        from test_stages_gpu import _push_oracle_to_gpu, ALL_FIELDS
        scene, view = kitchen_standin(teapot, grid=(3, 2, 2))
        N, W, H = 1 << 15, 192, 108
        kw = dict(max_bounces=8, use_env_map=True, use_area_light=True, separate_queues=True, **view)
        p = abi.make_render_params(W, H, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), **kw)
        wf = pyoracle.OracleWavefront(N, W, H)
        wf.set_scene(scene); wf.set_envmap(env); wf.set_params(p.tobytes())
        ctx = fx.CudaContext(num_tasks=N)
        ctx.uploadSceneData(scene); ctx.createEnvMap(env); ctx.updateParams(p)
        cls = {}
        for mode in ("reference", "strict"):
            c = clrun.CLWavefront(N, W, H, mode)
            c.set_scene(scene); c.set_envmap(env); c.set_params(p.tobytes())
            cls[mode] = c
        pyoracle.set_threads(1)
        wf.reset(); wf.set_pixel_index(0); wf.raygen(); wf.ext(); wf.clear_queues()
        stats = {}

        def push_cl(c):
            c.write_state(wf.state())
            cnt = wf.counters()
            for qi, name in enumerate(abi.QUEUES):
                n = int(cnt[name + "Queue"])
                if n:
                    c.write_queue(qi, np.concatenate([wf.queue(qi)[:n], np.zeros(N - n, np.uint32)]))
            c.write_counters(cnt)
            c.write_pixels(wf.pixels()); c.write_spp(wf.spp()); c.set_pixel_index(wf.get_pixel_index())

        def compare(tag, a64, b64):
            for name in ALL_FIELDS:
                g, r = abi.state_field(a64, name), abi.state_field(b64, name)
                s = stats.setdefault(tag, {}).setdefault(name, [0, 0, 0, 0, 0])
                s[0] += N
                if name in abi.INT_SLOTS:
                    d = (g != r) if g.ndim == 1 else (g != r).any(1)
                    s[1] += int(d.sum()); s[2] += int(d.sum()); s[3] += int(d.sum())
                    continue
                g64, r64 = g.astype(np.float64), r.astype(np.float64)
                with np.errstate(invalid="ignore"):
                    e = np.abs(g64 - r64) / np.maximum(np.abs(r64), 1e-3)
                    e = np.where(np.isnan(g64) & np.isnan(r64), 0.0, e)
                    e = np.where(np.isinf(g64) & (g64 == r64), 0.0, e)
                    e = np.nan_to_num(e, nan=np.inf)
                e = e if e.ndim == 1 else e.max(1)
                s[1] += int((e > 1e-5).sum()); s[2] += int((e > 1e-4).sum()); s[3] += int((e > 1e-3).sum())
                s[4] = max(s[4], float(np.minimum(e, 1e9).max()))

        for it in range(6):
            for stage, fo, fg, fc in (("logic", lambda: wf.logic(False), lambda: ctx.enqueueWfLogicKernel(firstIteration=False), lambda c: c.logic(False)),
                                      ("raygen", wf.raygen, ctx.enqueueWfRaygenKernel, lambda c: c.raygen()),
                                      ("materials", wf.materials, ctx.enqueueWfMaterialKernels, lambda c: c.materials()),
                                      ("ext", wf.ext, ctx.enqueueWfExtRayKernel, lambda c: c.ext()),
                                      ("shadow", wf.shadow, ctx.enqueueWfShadowRayKernel, lambda c: c.shadow())):
                _push_oracle_to_gpu(wf, ctx)
                for c in cls.values():
                    push_cl(c)
                if stage == "ext":
                    cnt = wf.counters().copy()
                fo(); fg()
                g = ctx.exportState()
                compare("cuda_vs_oracle/" + stage, g, wf.state())
                for mode, c in cls.items():
                    fc(c)
                    compare("cuda_vs_opencl_%s/%s" % (mode, stage), g, c.state())
            wf.clear_queues(); wf.update_pixel_index(int(cnt["raygenQueue"]))
        pyoracle.set_threads(pyoracle.max_threads())
        out = {}
        for tag, fields in stats.items():
            worst = {k: v for k, v in fields.items() if v[1]}
            out[tag] = {k: {"paths": v[0], ">1e-5": v[1], ">1e-4": v[2], ">1e-3": v[3], "max_rel": v[4]} for k, v in worst.items()}
            say("  ", tag, {k: (v[1], v[2], v[3]) for k, v in worst.items()})
        RES["stages"] = {"scene": "kitchen stand-in 3x2x2, N=2^15, 192x108, 6 iterations", "fields": out}
        ctx.close()
        for c in cls.values():
            c.close()

txt.close()
