"""The GPU hierarchy builder (flx_build_lbvh): its output is a hierarchy in the REFERENCE's format
(Node[] in depth-first order + index list), so it is checked three ways —
  * structure: the invariants the reference's traversal relies on (src/bvh.cl:234-310: left child at index + 1,
    rightChild, leaves index the list, every triangle exactly once, boxes contain what they claim);
  * traversal: CUDA kernels and CPU oracle walk the SAME tree -> every hit field bit-exact;
  * geometry: against brute force (no tree at all) -> t bit-exact, triangle equal except exact-t ties."""
import time

import numpy as np
import pytest

from conftest import bits, cached_scene
from test_trace_gpu import _check_closest, _rays_inside

pytestmark = pytest.mark.gpu


def _validate(scene, max_leaf):
    nodes, idx, tris = scene["nodes"], scene["indices"], scene["tris"]
    n, m = len(tris), len(nodes)
    assert sorted(idx.tolist()) == list(range(n)), "every triangle exactly once in the index list"
    leaf = nodes["nPrims"] > 0
    assert nodes["nPrims"].max() <= max_leaf
    assert nodes["parent"][0] == -1
    inner = np.nonzero(~leaf)[0]
    right = nodes["iStartOrRight"][inner].astype(np.int64)
    left = inner + 1
    assert np.all(right > left) and np.all(right < m)
    assert np.all(nodes["parent"][left] == inner) and np.all(nodes["parent"][right] == inner)
    assert leaf.sum() + len(inner) == m and len(inner) == leaf.sum() - 1        # full binary tree
    # leaves tile the index list in depth-first order
    ls = np.nonzero(leaf)[0]
    start, cnt = nodes["iStartOrRight"][ls].astype(np.int64), nodes["nPrims"][ls].astype(np.int64)
    assert start[0] == 0 and np.all(start[1:] == start[:-1] + cnt[:-1]) and start[-1] + cnt[-1] == n
    # boxes: a leaf's box is exactly the box of its triangles; a parent's box is exactly the union of its children
    v = tris["v"][:, :, 0, :3]                                                                  # (n, vertex, xyz)
    tlo, thi = v.min(1)[idx], v.max(1)[idx]
    lo = np.minimum.reduceat(tlo, start); hi = np.maximum.reduceat(thi, start)
    assert np.array_equal(nodes["bmin"][ls, :3], lo) and np.array_equal(nodes["bmax"][ls, :3], hi)
    assert np.array_equal(nodes["bmin"][inner, :3], np.minimum(nodes["bmin"][left, :3], nodes["bmin"][right, :3]))
    assert np.array_equal(nodes["bmax"][inner, :3], np.maximum(nodes["bmax"][left, :3], nodes["bmax"][right, :3]))
    # depth reported == depth found
    depth = np.zeros(m, np.int64)
    for i in range(1, m):                       # parents precede children in depth-first order
        depth[i] = depth[nodes["parent"][i]] + 1
    assert depth.max() == scene["lbvh_depth"] <= 64
    return depth.max()


@pytest.mark.parametrize("name,max_leaf", [("teapot", 1), ("teapot", 4), ("egyptcat", 4), ("psor_cube", 8)])
def test_lbvh_structure_and_traversal_parity(oracle, teapot, name, max_leaf):
    from fluctus_b200 import CudaContext, scene as fs
    base = teapot if name == "teapot" else cached_scene(name)
    lb = fs.build_lbvh(base, max_leaf=max_leaf)
    _validate(lb, max_leaf)
    ctx = CudaContext(num_tasks=1024)
    rays = np.concatenate([fs.bench_rays(6000, lb["bounds"]), _rays_inside(lb, 3000, 7)])
    ref = _check_closest(ctx, oracle, lb, rays)                   # same tree on both sides: everything bit-exact
    assert (ref["hit_i"] >= 0).sum() > 500
    brute = oracle.trace_closest(lb, rays, 2)                     # no tree at all
    assert np.array_equal(bits(brute["hit_t"]), bits(ref["hit_t"]))
    differ = brute["hit_i"] != ref["hit_i"]
    assert differ.mean() <= 2e-3                                   # exact-t ties only (shared edges)
    # and against the reference builder's SBVH of the same triangles
    sb = oracle.trace_closest(base, rays, 0)
    assert np.array_equal(bits(sb["hit_t"]), bits(ref["hit_t"])) and (sb["hit_i"] != ref["hit_i"]).mean() <= 2e-3
    # any hit
    any_rays = rays.copy(); any_rays[::2, 3] = 1.0
    ctx.uploadSceneData(lb)
    assert np.array_equal(ctx.traceAny(any_rays), oracle.trace_any(lb, any_rays, 0))
    ctx.close()


def test_lbvh_degenerate_inputs(oracle, teapot):
    """one triangle; two identical triangles (equal Morton codes); max_leaf larger than the scene; bad arguments"""
    from fluctus_b200 import CudaContext, FluctusError, scene as fs
    ctx = CudaContext(num_tasks=64)
    for pick, max_leaf in ((slice(0, 1), 4), (np.array([5, 5]), 1), (slice(0, 7), 255), (np.array([3, 3, 3, 9, 9, 200]), 2)):
        s = dict(teapot); s["tris"] = teapot["tris"][pick].copy()
        lb = fs.build_lbvh(s, max_leaf=max_leaf)
        _validate(lb, max_leaf)
        rays = _rays_inside(lb, 256, 3)
        c = 0.5 * (lb["bounds"][0] + lb["bounds"][1])
        rays[:, :3] = c + np.float32([0, 0, 2.0]); rays[:, 4:7] = np.float32([0, 0, -1]) + 0.05 * rays[:, 4:7]
        _check_closest(ctx, oracle, lb, rays)
    with pytest.raises(FluctusError):
        fs.build_lbvh(teapot, max_leaf=0)
    with pytest.raises(FluctusError):
        fs.build_lbvh(teapot, max_leaf=256)
    ctx.close()


def test_lbvh_too_deep_for_the_traversal_stack_is_refused_at_build_time(oracle, teapot):
    """A chain of ~60 Morton levels (one triangle per top bit) ending in 200 coincident triangles (index tie-break, 8
    more levels) is deeper than the 64-entry stack: flx_build_lbvh says so itself instead of handing out a hierarchy
    that flx_upload_scene refuses; a larger max_leaf collapses the coincident run and the build goes through."""
    from fluctus_b200 import CudaContext, FluctusError, scene as fs
    t0 = teapot["tris"][:1].copy()
    shape = t0["v"][0, :, 0, :3] - t0["v"][0, :, 0, :3].mean(0)
    shape *= np.float32(0.05 / np.abs(shape).max())
    pos = [np.zeros(3, np.float32)] * 200
    for j in range(63):
        c = np.zeros(3, np.float32); c[j % 3] = np.float32(2.0 ** (j // 3))
        pos.append(c)
    tris = np.repeat(t0, len(pos))
    tris["v"][:, :, 0, :3] = shape[None] + np.stack(pos)[:, None, :]
    s = dict(teapot); s["tris"] = tris
    with pytest.raises(FluctusError) as e:
        fs.build_lbvh(s, max_leaf=1)
    assert "levels deep" in str(e.value) and "max_leaf" in str(e.value)
    lb = fs.build_lbvh(s, max_leaf=255)
    assert lb["lbvh_depth"] + 1 <= 64
    ctx = CudaContext(num_tasks=64)
    rays = _rays_inside(lb, 512, 5)
    _check_closest(ctx, oracle, lb, rays)
    ctx.close()


def test_upload_rejects_misaligned_texture_offsets(teapot):
    """readTexel loads RGBA8 texels as words (flx_shade.cuh): an offset that is not a multiple of 4 is refused at upload
    instead of faulting in the shading kernels."""
    from fluctus_b200 import CudaContext, FluctusError
    from fluctus_b200.abi import TEX_DT
    s = dict(teapot)
    s["textures"] = np.array([(2, 4, 4)], TEX_DT)
    s["texels"] = np.zeros(2 + 4 * 4 * 4, np.uint8)
    ctx = CudaContext(num_tasks=64)
    with pytest.raises(FluctusError) as e:
        ctx.uploadSceneData(s)
    assert "multiple of 4" in str(e.value)
    s["textures"] = np.array([(4, 4, 4)], TEX_DT)
    s["texels"] = np.zeros(4 + 4 * 4 * 4, np.uint8)
    ctx.uploadSceneData(s)
    ctx.close()


def test_lbvh_ten_million_triangles(oracle, teapot):
    """cfg5-size input (15x15x14 teapots, 10 098 900 triangles): builds in well under a second of GPU work where the
    reference's single-threaded builder needs minutes; rays agree with the replicated-SBVH hierarchy on t bit for bit."""
    from fluctus_b200 import CudaContext, scene as fs
    grid = fs.replicate_grid(teapot, (15, 15, 14), (2.2, 1.1, 1.4))
    fs.build_lbvh(dict(tris=grid["tris"][:1000], materials=grid["materials"]))      # warm-up (context creation)
    t0 = time.perf_counter()
    lb = fs.build_lbvh(grid, max_leaf=4)
    dt = time.perf_counter() - t0
    print("LBVH build of %d triangles incl. 1.6 GB upload + download: %.2f s, %d nodes, depth %d" % (len(grid["tris"]), dt, len(lb["nodes"]), lb["lbvh_depth"]))
    assert dt < 20.0 and lb["lbvh_depth"] <= 64
    nodes = lb["nodes"]
    assert sorted(np.unique(lb["indices"][::1000]).tolist()) == sorted(lb["indices"][::1000].tolist())
    assert (nodes["nPrims"] > 0).sum() == (nodes["nPrims"] == 0).sum() + 1
    rays = fs.bench_rays(1 << 18, grid["bounds"])
    ctx = CudaContext(num_tasks=1024)
    ctx.uploadSceneData(grid)
    i0, t0_ = ctx.traceClosest(rays)
    ctx.uploadSceneData(lb)
    i1, t1_ = ctx.traceClosest(rays)
    assert np.array_equal(bits(t0_), bits(t1_)) and (i0 != i1).mean() <= 2e-3 and (i0 >= 0).mean() > 0.2
    # a sample against the oracle walking the LBVH itself
    ref = oracle.trace_closest(lb, rays[:4096], 0)
    assert np.array_equal(ref["hit_i"], i1[:4096]) and np.array_equal(bits(ref["hit_t"]), bits(t1_[:4096]))
    ctx.close()
