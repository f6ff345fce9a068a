"""Closest-hit / any-hit parity of the CUDA traversal kernels against the CPU oracle, through the
C-ABI (flx_trace_closest / flx_trace_any).  Contract (BASELINE.json north_star): hit index and
hit/miss bit-exact; we also require hit.t, P, N, uv bit-exact because the arithmetic is op-for-op
the reference's (no FMA, IEEE div/sqrt)."""
import os

import numpy as np
import pytest

from conftest import bits, cached_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from fluctus_b200 import CudaContext
    c = CudaContext(num_tasks=4096)
    yield c
    c.close()


def _rays_inside(scene, n, seed):
    """Rays that start inside the scene box (many hits, deep traversals), with finite tmax for some."""
    rng = np.random.default_rng(seed)
    b = scene["bounds"]
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = rng.uniform(b[0], b[1], (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 4:7] = d
    rays[:, 3] = np.float32(3.402823466e+38)
    return rays


def _check_closest(ctx, oracle, scene, rays):
    ctx.uploadSceneData(scene)
    ref = oracle.trace_closest(scene, rays, 0, want_hits=True)
    hit_i, hit_t, hits = ctx.traceClosest(rays, want_hits=True)
    assert np.array_equal(hit_i, ref["hit_i"])
    assert np.array_equal(bits(hit_t), bits(ref["hit_t"]))
    for f in ("P", "N", "uv", "t"):
        assert np.array_equal(bits(hits[f]), bits(ref["hits"][f])), f
    for f in ("i", "matId", "areaLightHit"):
        assert np.array_equal(hits[f], ref["hits"][f]), f
    # the index-only path (no attribute fetch) must agree too
    hit_i2, hit_t2 = ctx.traceClosest(rays)
    assert np.array_equal(hit_i2, hit_i) and np.array_equal(bits(hit_t2), bits(hit_t))
    return ref


def test_closest_teapot_incoherent(ctx, oracle, teapot):
    from fluctus_b200 import scene as fs
    rays = fs.bench_rays(1 << 17, teapot["bounds"])
    ref = _check_closest(ctx, oracle, teapot, rays)
    assert 0.1 < (ref["hit_i"] >= 0).mean() < 0.4


def test_closest_teapot_inside(ctx, oracle, teapot):
    _check_closest(ctx, oracle, teapot, _rays_inside(teapot, 1 << 16, 1))


def test_closest_axis_aligned_and_degenerate_rays(ctx, oracle, teapot):
    """Zero direction components give +-inf reciprocals and 0*inf = NaN in the slab test
    (src/intersect.cl:43-50): fmin/fmax must drop the NaN exactly as OpenCL does."""
    rays = _rays_inside(teapot, 6 * 512, 2)
    for k, axis in enumerate(("x+", "x-", "y+", "y-", "z+", "z-")):
        d = np.zeros(3, np.float32)
        d[k // 2] = 1.0 if k % 2 == 0 else -1.0
        rays[k * 512:(k + 1) * 512, 4:7] = d
    rays[:64, 4:7] = 0.0                    # null direction (what an emissive hit produces)
    rays[64:128, 3] = 0.05                  # short tmax
    rays[128:192, 3] = 0.0                  # tmax = 0: nothing can be closer
    rays[192:200, 0] = np.nan               # NaN origin must come back as a miss
    _check_closest(ctx, oracle, teapot, rays)


def test_any_hit_matches_oracle(ctx, oracle, teapot):
    rays = _rays_inside(teapot, 1 << 15, 3)
    rays[:, 3] = np.random.default_rng(4).uniform(0.01, 1.5, len(rays)).astype(np.float32)
    ctx.uploadSceneData(teapot)
    assert np.array_equal(ctx.traceAny(rays), oracle.trace_any(teapot, rays, 0))


def test_replicated_grid(ctx, oracle, teapot):
    """Instanced teapots: exact-t ties between distinct triangles never occur across instances, SBVH
    duplicates inside one do; the visit order must resolve them like the reference."""
    from fluctus_b200 import scene as fs
    grid = fs.replicate_grid(teapot, (3, 2, 2), (2.2, 1.1, 1.4))
    assert len(grid["tris"]) == 12 * len(teapot["tris"])
    rays = fs.bench_rays(1 << 16, grid["bounds"])
    ref = _check_closest(ctx, oracle, grid, rays)
    assert (ref["hit_i"] >= 0).mean() > 0.2
    # bit-stack traversal of the oracle walks the same order (SURVEY §4 consistency pair 1)
    alt = oracle.trace_closest(grid, rays, 1)
    assert np.array_equal(alt["hit_i"], ref["hit_i"])


@pytest.mark.parametrize("name", ["egyptcat", "psor_cube"])
def test_closest_reference_assets(ctx, oracle, name):
    scene = cached_scene(name)
    rays = _rays_inside(scene, 1 << 15, 5)
    _check_closest(ctx, oracle, scene, rays)


def test_single_leaf_scene(ctx, oracle, teapot):
    """A hierarchy whose root is a leaf (<= 8 triangles): the root reference itself is a leaf reference."""
    from fluctus_b200.abi import NODE_DT
    tris = teapot["tris"][:5].copy()
    nodes = np.zeros(1, NODE_DT)
    p = tris["v"][:, :, 0, :3].reshape(-1, 3)
    nodes["bmin"][0, :3], nodes["bmax"][0, :3] = p.min(0), p.max(0)
    nodes["parent"], nodes["iStartOrRight"], nodes["nPrims"] = -1, 0, 5
    scene = dict(teapot, tris=tris, nodes=nodes, indices=np.arange(5, dtype=np.uint32))
    scene["bounds"] = np.stack([p.min(0), p.max(0)])
    rays = _rays_inside(scene, 4096, 6)
    rays[:, :3] += np.float32(0.3) * np.random.default_rng(7).normal(size=(4096, 3)).astype(np.float32)
    _check_closest(ctx, oracle, scene, rays)


def test_empty_batch_and_errors(ctx, teapot):
    from fluctus_b200 import FluctusError
    ctx.uploadSceneData(teapot)
    hit_i, hit_t = ctx.traceClosest(np.zeros((0, 8), np.float32))
    assert len(hit_i) == 0
    bad = dict(teapot, indices=teapot["indices"].copy())
    bad["indices"][0] = len(teapot["tris"]) + 7
    with pytest.raises(FluctusError):
        ctx.uploadSceneData(bad)
    ctx.uploadSceneData(teapot)  # context stays usable after a rejected upload


def _chain_scene(teapot, levels):
    """Degenerate hierarchy: every inner node has a one-triangle leaf on the left and the rest of the chain on
    the right — depth = levels + 1.  Obeys the reference's Node[] invariants."""
    from fluctus_b200.abi import NODE_DT
    tris = teapot["tris"][:levels + 1].copy()
    P = tris["v"][:, :, 0, :3]
    nodes = np.zeros(2 * levels + 1, NODE_DT)
    lo = np.minimum.accumulate(P.min(1)[::-1])[::-1]   # box of triangles k..end
    hi = np.maximum.accumulate(P.max(1)[::-1])[::-1]
    for k in range(levels):
        inner, leaf = 2 * k, 2 * k + 1
        nodes[inner]["bmin"][:3], nodes[inner]["bmax"][:3] = lo[k], hi[k]
        nodes[inner]["parent"], nodes[inner]["iStartOrRight"], nodes[inner]["nPrims"] = (2 * (k - 1) if k else -1), 2 * k + 2, 0
        nodes[leaf]["bmin"][:3], nodes[leaf]["bmax"][:3] = P[k].min(0), P[k].max(0)
        nodes[leaf]["parent"], nodes[leaf]["iStartOrRight"], nodes[leaf]["nPrims"] = inner, k, 1
    last = 2 * levels
    nodes[last]["bmin"][:3], nodes[last]["bmax"][:3] = P[levels].min(0), P[levels].max(0)
    nodes[last]["parent"], nodes[last]["iStartOrRight"], nodes[last]["nPrims"] = 2 * (levels - 1), levels, 1
    scene = dict(teapot, tris=tris, nodes=nodes, indices=np.arange(levels + 1, dtype=np.uint32))
    scene["bounds"] = np.stack([lo[0], hi[0]])
    return scene


def test_deep_chain_hierarchy(ctx, oracle, teapot):
    """Depth 61 (within the 64-entry stack, deeper than the shared-memory top levels are wide) traverses
    bit-exactly; depth 71 is refused at upload instead of overflowing the stack (reference: unchecked)."""
    from fluctus_b200 import FluctusError
    scene = _chain_scene(teapot, 60)
    rays = _rays_inside(scene, 8192, 9)
    _check_closest(ctx, oracle, scene, rays)
    with pytest.raises(FluctusError) as e:
        ctx.uploadSceneData(_chain_scene(teapot, 70))
    assert "deeper" in str(e.value)
    ctx.uploadSceneData(teapot)


def test_bitstack_traversal_matches_stack_and_oracle(oracle, teapot):
    """Settings::clUseBitstack / -DUSE_BITSTACK (src/bvh.cl:12-230): the stackless bit-trail traversal gives the same
    closest hit (index, distance, attributes) and the same occlusion flags as the stack traversal and as the oracle's
    restatement of BOTH halves of bvh.cl — on the teapot, two reference assets, a replicated grid (depth 25), rays
    starting inside, axis-parallel rays, finite tmax, a single-leaf scene and a depth-31 chain; deeper hierarchies are
    refused while the switch is on."""
    from fluctus_b200 import CudaContext, FluctusError, scene as fs
    c = CudaContext(num_tasks=4096)
    try:
        c.setTraversal(True)
        scenes = [("teapot", teapot), ("egyptcat", cached_scene("egyptcat")), ("psor_cube", cached_scene("psor_cube")),
                  ("grid", fs.replicate_grid(teapot, (4, 3, 3), (2.2, 1.1, 1.4))), ("chain30", _chain_scene(teapot, 30))]
        from fluctus_b200.abi import NODE_DT
        tris = teapot["tris"][:5].copy()
        node = np.zeros(1, NODE_DT)
        pts = tris["v"][:, :, 0, :3].reshape(-1, 3)
        node["bmin"][0, :3], node["bmax"][0, :3] = pts.min(0), pts.max(0)
        node["parent"], node["iStartOrRight"], node["nPrims"] = -1, 0, 5
        single = dict(teapot, tris=tris, nodes=node, indices=np.arange(5, dtype=np.uint32))
        single["bounds"] = np.stack([pts.min(0), pts.max(0)])
        scenes.append(("single_leaf", single))
        for name, scene in scenes:
            rays = np.concatenate([fs.bench_rays(1 << 15, scene["bounds"]), _rays_inside(scene, 1 << 14, 3)])
            ax = _rays_inside(scene, 3072, 5)
            ax[:1024, 4:7] = (1, 0, 0); ax[1024:2048, 4:7] = (0, -1, 0); ax[2048:, 4:7] = (0, 0, 1)
            rays = np.concatenate([rays, ax])
            rays[::7, 3] = np.float32(1.5)  # finite tmax on some
            c.setTraversal(True)
            c.uploadSceneData(scene)
            bi, bt, bh = c.traceClosest(rays, want_hits=True)
            bo = c.traceAny(rays)
            c.setTraversal(False)
            si, st_, sh = c.traceClosest(rays, want_hits=True)
            so = c.traceAny(rays)
            assert np.array_equal(bi, si) and np.array_equal(bits(bt), bits(st_)), name
            assert bh.tobytes() == sh.tobytes() and np.array_equal(bo, so), name
            for mode in (0, 1):  # the oracle's stack and bit-stack restatements
                ref = oracle.trace_closest(scene, rays, mode)
                assert np.array_equal(bi, ref["hit_i"]) and np.array_equal(bits(bt), bits(ref["hit_t"])), (name, mode)
                assert np.array_equal(bo, oracle.trace_any(scene, rays, mode)), (name, mode)
        # depth 32 = one level more than the trails can hold
        c.setTraversal(True)
        with pytest.raises(FluctusError) as e:
            c.uploadSceneData(_chain_scene(teapot, 31))
        assert "31 levels" in str(e.value)
        c.setTraversal(False)
        c.uploadSceneData(_chain_scene(teapot, 31))
        with pytest.raises(FluctusError):
            c.setTraversal(True)
    finally:
        c.close()


def test_bitstack_render_agrees_with_stack_render(teapot, night_env):
    """The wavefront render through flx_render with the bit-stack traversal: same ray / sample statistics and the same
    image (up to the order of the float atomics) as with the stack."""
    from fluctus_b200 import Tracer
    out = []
    for bitstack in (False, True):
        tr = Tracer(96, 64, num_tasks=8192, max_bounces=4, use_env_map=True, use_area_light=True)
        tr.init(teapot, night_env)
        tr.clctx.setTraversal(bitstack)
        tot = tr.render(iterations=120)
        px = tr.pixels()
        out.append((tot, px[..., :3].sum((0, 1)) / max(px[..., 3].sum(), 1)))
        tr.close()
    for k in ("primary", "extension", "shadow", "samples"):
        # the same hits for every ray; which pixel a restarted path takes depends on the order of the queue atomics
        assert abs(out[0][0][k] - out[1][0][k]) <= 0.02 * out[0][0][k], k
    assert np.allclose(out[0][1], out[1][1], rtol=0.03)


def _twin_scene(teapot):
    """Two coincident copies of the teapot under one root whose children have IDENTICAL boxes.  The left child (visited
    first on equal entry distance, src/bvh.cl:293-300) owns the triangles [n, 2n) and the second half of the index
    list, the right child the triangles [0, n): every hit has an exact-`t` twin, the traversal must report the left
    copy (strictly-closer wins, so the twin met later does not replace it) while a loop over the triangle array in
    index order reports the right copy."""
    from fluctus_b200.abi import NODE_DT
    A, B = teapot["nodes"].copy(), teapot["nodes"].copy()
    nA, nT, nI = len(A), len(teapot["tris"]), len(teapot["indices"])
    for half, node_off, idx_off in ((A, 1, nI), (B, 1 + nA, 0)):
        inner = half["nPrims"] == 0
        half["iStartOrRight"][inner] += node_off
        half["iStartOrRight"][~inner] += idx_off
        half["parent"][1:] += node_off
        half["parent"][0] = 0
    root = np.zeros(1, NODE_DT)
    root["bmin"], root["bmax"] = teapot["nodes"]["bmin"][0], teapot["nodes"]["bmax"][0]
    root["parent"], root["iStartOrRight"], root["nPrims"] = -1, 1 + nA, 0
    scene = dict(teapot, nodes=np.concatenate([root, A, B]), tris=np.concatenate([teapot["tris"], teapot["tris"]]),
                 indices=np.concatenate([teapot["indices"], teapot["indices"] + np.uint32(nT)]))
    return scene


def test_exact_ties_follow_the_reference_visit_order(oracle, teapot):
    """Exact-`t` ties (coincident triangles in different subtrees): the reported index is decided by the visit order
    and the strictly-closer rule, and must be the oracle's for every ray — with the stack and with the bit-stack."""
    from fluctus_b200 import CudaContext, scene as fs
    scene = _twin_scene(teapot)
    n = len(teapot["tris"])
    rays = np.concatenate([fs.bench_rays(1 << 16, scene["bounds"]), _rays_inside(scene, 1 << 15, 2)])
    ref = oracle.trace_closest(scene, rays, 0)
    hit = ref["hit_i"] >= 0
    assert hit.mean() > 0.2 and (ref["hit_i"][hit] >= n).all()              # the left copy wins every tie
    brute = oracle.trace_closest(scene, rays[:8192], 2)
    assert np.array_equal(bits(brute["hit_t"]), bits(ref["hit_t"][:8192]))
    assert (brute["hit_i"][hit[:8192]] < n).all()                           # index order would report the other copy
    c = CudaContext(num_tasks=64)
    try:
        c.uploadSceneData(scene)
        for bitstack in (False, True):
            c.setTraversal(bitstack)
            hit_i, hit_t, hits = c.traceClosest(rays, want_hits=True)
            assert np.array_equal(hit_i, ref["hit_i"]) and np.array_equal(bits(hit_t), bits(ref["hit_t"])), bitstack
            assert np.array_equal(hits["i"][hit], ref["hit_i"][hit])
            assert np.array_equal(c.traceAny(rays) != 0, hit)
    finally:
        c.close()


def test_host_batch_pipeline_chunk_boundaries(ctx, oracle, teapot):
    """flx_trace_closest / flx_trace_any with host buffers run as an upload / traverse / download pipeline over chunks of
    2^20 rays: a batch that ends 777 rays into its third chunk (and pageable, not pinned, memory) must give what one
    device-resident launch gives, and the rays either side of every chunk boundary must match the oracle."""
    import torch
    from fluctus_b200 import scene as fs
    ctx.uploadSceneData(teapot)
    n = (2 << 20) + 777
    rays = fs.bench_rays(n, teapot["bounds"])
    hit_i, hit_t, hits = ctx.traceClosest(rays, want_hits=True)
    d_rays = torch.from_numpy(rays).cuda()
    d_i = torch.empty(n, dtype=torch.int32, device="cuda"); d_t = torch.empty(n, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    ctx.traceClosestDevice(d_rays.data_ptr(), n, d_i.data_ptr(), d_t.data_ptr())
    ctx.finishQueue()
    assert np.array_equal(hit_i, d_i.cpu().numpy()) and np.array_equal(bits(hit_t), bits(d_t.cpu().numpy()))
    h = hit_i >= 0
    assert np.array_equal(hits["i"][h], hit_i[h]) and np.array_equal(bits(hits["t"][h]), bits(hit_t[h]))
    sel = np.concatenate([np.arange(0, 512), np.arange((1 << 20) - 512, (1 << 20) + 512),
                          np.arange((2 << 20) - 512, (2 << 20) + 512), np.arange(n - 1024, n)])
    ref = oracle.trace_closest(teapot, rays[sel], 0)
    assert np.array_equal(hit_i[sel], ref["hit_i"]) and np.array_equal(bits(hit_t[sel]), bits(ref["hit_t"]))
    occ = ctx.traceAny(rays)
    assert np.array_equal(occ != 0, hit_i >= 0)  # tmax = FLT_MAX: occluded <=> some hit
    assert np.array_equal(occ[sel], oracle.trace_any(teapot, rays[sel], 0))


def test_cfg5_full_size_microbench_parity(oracle, teapot):
    """BASELINE config 5 at full size: 15x15x14 replicated teapots (10 098 900 triangles), 2^24 incoherent rays.
    Oracle comparison (bit-exact index + distance) on ALL 2^24 rays (the oracle is OpenMP-parallel: seconds on the
    box's host cores); size-independent properties besides: idempotence, any-hit <=> closest hit below tmax, hit
    distances invariant under batch order."""
    from fluctus_b200 import CudaContext, scene as fs
    grid = fs.replicate_grid(teapot, (15, 15, 14), (2.2, 1.1, 1.4))
    assert len(grid["tris"]) == 10098900
    n = 1 << 24
    rays = fs.bench_rays(n, grid["bounds"])
    c = CudaContext(num_tasks=1024)
    try:
        c.uploadSceneData(grid)
        hit_i, hit_t = c.traceClosest(rays)
        oracle.set_threads(max(oracle.max_threads(), len(os.sched_getaffinity(0))))
        ref = oracle.trace_closest(grid, rays, 0)
        assert np.array_equal(hit_i, ref["hit_i"]), "%d of 2^24 hit indices differ" % int((hit_i != ref["hit_i"]).sum())
        assert np.array_equal(bits(hit_t), bits(ref["hit_t"]))
        frac = float((hit_i >= 0).mean())
        assert 0.3 < frac < 0.55, frac
        assert hit_i.max() < len(grid["tris"]) and hit_i.min() == -1
        # idempotence: the persistent scheduler hands rays to lanes in a different order every launch
        hit_i2, hit_t2 = c.traceClosest(rays)
        assert np.array_equal(hit_i2, hit_i) and np.array_equal(bits(hit_t2), bits(hit_t))
        # batch order does not matter: reversed batch gives the reversed result
        k = 1 << 22
        ri, rt = c.traceClosest(np.ascontiguousarray(rays[:k][::-1]))
        assert np.array_equal(ri[::-1], hit_i[:k]) and np.array_equal(bits(rt[::-1]), bits(hit_t[:k]))
        # any hit with tmax just beyond / just before the closest hit
        sel = np.nonzero(hit_i[:k] >= 0)[0]
        far, near = rays[sel].copy(), rays[sel].copy()
        # (a margin, not one ulp: the slab test of a flat leaf box may place the box an ulp behind the triangle's
        # own t, and bvh_occluded culls boxes against tmax — reference behaviour, reproduced bit for bit below)
        far[:, 3] = hit_t[sel] * np.float32(1.001)
        near[:, 3] = hit_t[sel] * np.float32(0.999)
        occ_far, occ_near = c.traceAny(far), c.traceAny(near)
        assert occ_far.all()
        assert not occ_near.any()
        edge = rays[sel[:1 << 17]].copy()
        edge[:, 3] = np.nextafter(hit_t[sel[:1 << 17]], np.float32(np.inf))
        assert np.array_equal(c.traceAny(edge), oracle.trace_any(grid, edge, 0))
        miss = rays[:k][hit_i[:k] < 0]
        assert not c.traceAny(miss).any()
    finally:
        c.close()


def test_pick_single_matches_oracle(oracle, teapot):
    """CLContext::pickSingle / kernel_pick.cl: the Hit record under NDC points, bit-exact, including the light quad
    reported as an implicit hit (i = 0, matId = 0, areaLightHit = 1) when sampleImpl && useAreaLight."""
    from fluctus_b200 import CudaContext, abi, scene as fs
    W, H = 640, 360
    for impl in (True, False):
        p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), use_env_map=False,
                                   use_area_light=True, sample_implicit=impl, cam_pos=(0.0, 0.2, 1.6),
                                   light_pos=(0.0, 0.9, 0.0), light_n=(0.0, 0.0, 1.0), light_right=(1.0, 0.0, 0.0), light_up=(0.0, 1.0, 0.0),
                                   light_size=(0.6, 0.3))
        wf = oracle.OracleWavefront(64, W, H)
        wf.set_scene(teapot); wf.set_params(p.tobytes())
        ctx = CudaContext(num_tasks=64)
        ctx.uploadSceneData(teapot); ctx.updateParams(p)
        kinds = set()
        for y in np.linspace(0.02, 0.98, 13):
            for x in np.linspace(0.02, 0.98, 17):
                g, r = ctx.pickSingle(float(x), float(y)), wf.pick(float(x), float(y))
                assert g.tobytes() == r.tobytes(), (x, y, g, r)
                kinds.add("light" if r["areaLightHit"] else ("tri" if r["i"] >= 0 else "miss"))
        assert kinds == ({"light", "tri", "miss"} if impl else {"tri", "miss"}), kinds
        ctx.close()
