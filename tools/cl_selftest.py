"""Run on the GPU box: the reference's untouched OpenCL kernels (oracle/clrun.py) against the oracle and the CUDA
library on the teapot fixture — ray batches in both build modes, a short wavefront render, every kernel's build.
Writes gpurun_out/cl_selftest.txt.  Evidence run; the assertions live in tests/test_opencl_reference_gpu.py."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
out = open(os.path.join(ROOT, "gpurun_out", "cl_selftest.txt"), "w")


def say(*a):
    s = " ".join(str(x) for x in a)
    print(s); out.write(s + "\n"); out.flush()


def step(name):
    def deco(fn):
        say("==", name)
        t = time.time()
        try:
            fn()
        except Exception:
            say("FAILED:", traceback.format_exc()[-3000:])
        say("   (%.1f s)" % (time.time() - t))
    return deco


from oracle import clrun, pyoracle
from fluctus_b200 import scene as fs, abi
import fluctus_b200 as fx

ok, what = clrun.available()
say("OpenCL available:", ok, what)
if not ok:
    sys.exit(0)
teapot = fs.load_scene(os.path.join(ROOT, "tests", "golden", "teapot_sbvh.npz"))
env = fs.load_envmap(os.path.join(ROOT, "tests", "golden", "night_env.npz"))
rays = fs.bench_rays(1 << 18, teapot["bounds"])
ref = pyoracle.trace_closest(teapot, rays, 0, want_hits=True)


@step("ray batch through traceExtension, all build modes x traversal variants")
def _():
    ctx = fx.CudaContext(num_tasks=4096)
    ctx.uploadSceneData(teapot)
    ci, ct = ctx.traceClosest(rays)
    say("  CUDA vs oracle: hit_i equal", np.array_equal(ci, ref["hit_i"]), "hit_t bits equal", np.array_equal(ct.view(np.uint32), ref["hit_t"].view(np.uint32)))
    ctx.close()
    for mode in ("reference", "strict", "strict_noopt"):
        for bs in (False, True):
            try:
                r = clrun.trace_closest(teapot, rays, mode=mode, bitstack=bs, want_hits=True)
            except Exception as e:
                say("  mode", mode, "bitstack", bs, "FAILED:", str(e)[-1500:])
                continue
            di = int((r["hit_i"] != ref["hit_i"]).sum())
            dt = int((r["hit_t"].view(np.uint32) != ref["hit_t"].view(np.uint32)).sum())
            hit = ref["hit_i"] >= 0
            rel = np.abs(r["hit_t"][hit].astype(np.float64) - ref["hit_t"][hit]) / np.abs(ref["hit_t"][hit])
            dP = int((r["hits"]["P"].view(np.uint32) != ref["hits"]["P"].view(np.uint32)).any(1).sum())
            dN = int((r["hits"]["N"].view(np.uint32) != ref["hits"]["N"].view(np.uint32)).any(1).sum())
            say("  mode %-12s bitstack %d: hit_i differs %d / %d, hit_t bits differ %d (max rel %.2e), P differs %d, N differs %d, kernel %.3f ms = %.0f Mrays/s"
                % (mode, bs, di, len(rays), dt, rel.max() if len(rel) else 0, dP, dN, r["ms"], len(rays) / r["ms"] / 1e3))


@step("any-hit through traceShadow")
def _():
    any_rays = rays[: 1 << 16].copy()
    t = ref["hit_t"][: len(any_rays)]
    with np.errstate(over="ignore"):
        any_rays[:, 3] = np.where(ref["hit_i"][: len(any_rays)] >= 0, t * np.random.default_rng(3).uniform(0.5, 1.5, len(any_rays)).astype(np.float32), 10.0)
    o = pyoracle.trace_any(teapot, any_rays, 0)
    for mode in ("reference", "strict"):
        occ, ms = clrun.trace_any(teapot, any_rays, mode=mode)
        say("  mode %-10s occluded flags differ %d / %d (%.3f ms)" % (mode, int((occ != o).sum()), len(o), ms))


@step("every wavefront kernel builds with the reference's options; short render vs oracle")
def _():
    W = H = 64
    N = 4096
    for kw in (dict(max_bounces=4, use_env_map=True, use_area_light=False),
               dict(max_bounces=6, use_env_map=True, use_area_light=True, separate_queues=True)):
        p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), **kw)
        cl = clrun.CLWavefront(N, W, H, "reference")
        cl.set_scene(teapot); cl.set_envmap(env); cl.set_params(p.tobytes())
        t = time.time()
        it, st = cl.render(200)
        dt = time.time() - t
        img = cl.pixels().reshape(H, W, 4)
        wf = pyoracle.OracleWavefront(N, W, H)
        wf.set_scene(teapot); wf.set_envmap(env); wf.set_params(p.tobytes())
        _, so = wf.render(200)
        ref_img = wf.pixels().reshape(H, W, 4)
        a = (img[..., :3] / np.maximum(img[..., 3:], 1)).reshape(8, 8, 8, 8, 3).mean((1, 3))
        b = (ref_img[..., :3] / np.maximum(ref_img[..., 3:], 1)).reshape(8, 8, 8, 8, 3).mean((1, 3))
        relmse = float((((a - b) ** 2) / (b ** 2 + 1e-3)).mean())
        say("  %s: OpenCL stats %s | oracle stats %s | relMSE(8x8 blocks) %.5f | build %.1f s, loop %.2f s" % (kw, st, so, relmse, cl.build_seconds, dt))
        cl.close()

out.close()
