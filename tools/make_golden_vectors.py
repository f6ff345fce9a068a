"""Generate the committed golden vectors under tests/golden/ from the CPU oracle.

The reference ships no known-answer vectors and its OpenCL runtime is absent, so
these files do NOT pin the oracle to the reference: they freeze the oracle's own outputs (after its
self-consistency checks: stack == bit-stack == brute force) so that any later change to the oracle
or to the CUDA kernels is detected on a box that has neither /root/reference nor a compiler.

    python tools/make_golden_vectors.py [mk]   # rewrites tests/golden/trace_golden.npz, wf_golden.npz, mk_golden.npz
                                               # ("mk": only the microkernel file, the others stay byte-identical)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from fluctus_b200 import abi, scene as fs  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def special_rays(bounds):
    """Axis-aligned directions (inf reciprocals, 0*inf = NaN in the slab test), rays along box faces,
    zero-length direction components, finite tmax — the edge cases of SURVEY.md §7.3.2."""
    b0, b1 = np.asarray(bounds[0], np.float32), np.asarray(bounds[1], np.float32)
    c = 0.5 * (b0 + b1)
    rows = []
    for axis in range(3):
        for sign in (-1.0, 1.0):
            d = np.zeros(3, np.float32); d[axis] = sign
            for off in ((0, 0, 0), (0.1, 0.05, 0.02), (0.3, 0.2, 0.1)):
                o = c + np.asarray(off, np.float32); o[axis] = c[axis] - sign * 3.0
                rows.append(np.concatenate([o, [3.402823466e+38], d, [0]]))
            o = b0.copy(); o[axis] = c[axis] - sign * 3.0   # grazing a box face exactly
            rows.append(np.concatenate([o, [3.402823466e+38], d, [0]]))
    for tmax in (0.5, 2.0, 3.2):  # finite tmax
        rows.append(np.concatenate([c + np.float32([0, 0, 3.0]), [tmax], [0, 0, -1], [0]]))
    rows.append(np.concatenate([c, [3.402823466e+38], [0, 0, 0], [0]]))  # degenerate direction
    return np.asarray(rows, np.float32)


def microkernel(teapot, env):
    # microkernel renderer: Tracer::renderSingle, 3 spp at 48x32, both lights, serial
    W, H, SPP = 48, 32, 3
    p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=3,
                               use_env_map=True, use_area_light=True, cam_pos=(0.0, 0.2, 1.3))  # the teapot fills the frame
    mk = pyoracle.OracleWavefront(W * H, W, H)
    mk.set_scene(teapot); mk.set_envmap(env); mk.set_params(p.tobytes())
    mk.mk_render_single(SPP)
    fields = ("orig", "dir", "T", "Ei", "P", "N", "uvTex", "phase", "lastPdfW", "pathLen", "seed", "lastSpecular", "t", "i", "matId")
    rows = sorted(set(r for f in fields for r in range(abi.SLOTS[f], abi.SLOTS[f] + abi.SLOT_WIDTH.get(f, 1))))
    np.savez_compressed(os.path.join(GOLD, "mk_golden.npz"), params=p.view(np.uint8), spp=np.uint32(SPP), rows=np.array(rows, np.int32),
                        state_rows=mk.state()[rows].copy(), pixels=mk.pixels().copy(), stats=mk.mk_stats().copy())
    print("mk_golden:", mk.mk_stats(), "mean", mk.pixels()[:, :3].sum(0) / mk.pixels()[:, 3].sum())


def main():
    teapot = fs.load_scene(os.path.join(GOLD, "teapot_sbvh.npz"))
    env = fs.load_envmap(os.path.join(GOLD, "night_env.npz"))
    pyoracle.set_threads(1)
    if sys.argv[1:] == ["mk"]:
        microkernel(teapot, env)
        pyoracle.set_threads(pyoracle.max_threads())
        return

    rays = np.concatenate([fs.bench_rays(4096, teapot["bounds"]), special_rays(teapot["bounds"])])
    out = {}
    for mode, tag in ((0, "stack"), (1, "bitstack"), (2, "brute")):
        r = pyoracle.trace_closest(teapot, rays, mode, want_hits=True)
        out[tag] = r
    for tag in ("bitstack", "brute"):
        assert np.array_equal(out["stack"]["hit_i"], out[tag]["hit_i"]), tag
        assert np.array_equal(out["stack"]["hit_t"].view(np.uint32), out[tag]["hit_t"].view(np.uint32)), tag
    any_rays = rays.copy()
    any_rays[:2048, 3] = 1.5
    occ = pyoracle.trace_any(teapot, any_rays, 0)
    assert np.array_equal(occ, pyoracle.trace_any(teapot, any_rays, 1))
    h = out["stack"]["hits"]
    np.savez_compressed(os.path.join(GOLD, "trace_golden.npz"), rays=rays, hit_i=out["stack"]["hit_i"], hit_t=out["stack"]["hit_t"],
                        P=h["P"], N=h["N"], uv=h["uv"], matId=h["matId"], any_rays=any_rays, occluded=occ)
    print("trace_golden: %d rays, %d hits, %d occluded" % (len(rays), int((out["stack"]["hit_i"] >= 0).sum()), int(occ.sum())))

    # wavefront: state / queues / pixels after 6 iterations of the benchmark loop, N = 1024, 48x32, serial
    N, W, H = 1024, 48, 32
    p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=4,
                               use_env_map=True, use_area_light=True, separate_queues=True)
    wf = pyoracle.OracleWavefront(N, W, H)
    wf.set_scene(teapot); wf.set_envmap(env); wf.set_params(p.tobytes())
    it, st = wf.render(6, True)
    np.savez_compressed(os.path.join(GOLD, "wf_golden.npz"), params=p.view(np.uint8), state=wf.state().copy(),
                        pixels=wf.pixels().copy(), preview=wf.postprocess(), stats=np.array([st[k] for k in ("primary", "extension", "shadow", "samples")], np.uint64),
                        pixel_index=np.uint32(wf.get_pixel_index()))
    print("wf_golden:", st, "pixel index", wf.get_pixel_index())
    microkernel(teapot, env)
    pyoracle.set_threads(pyoracle.max_threads())


if __name__ == "__main__":
    main()
