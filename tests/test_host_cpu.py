"""CPU-only tests of the host-side logic: RenderParams builder, scene containers, instancing and the
stand-in scene builders (the Node[] invariants every traversal relies on), sharding plan, and the
world_size-2 gloo path of the per-frame accumulation reduce."""
import os
import socket

import numpy as np
import pytest

from conftest import ROOT


def _check_node_invariants(scene):
    """Reference Node[] contract (src/bvhnode.hpp:50-59, bvh.cpp flattening): DFS order, left child = i+1,
    right child > i+1, leaf iff nPrims != 0, leaf ranges tile the index list, parents consistent, child
    boxes inside the parent's."""
    nodes, idx = scene["nodes"], scene["indices"]
    n = len(nodes)
    leaf = nodes["nPrims"] != 0
    inner = np.nonzero(~leaf)[0]
    right = nodes["iStartOrRight"][inner].astype(np.int64)
    assert (right > inner + 1).all() and (right < n).all()
    assert (nodes["parent"][inner + 1] == inner).all()
    assert (nodes["parent"][right] == inner).all()
    assert nodes["parent"][0] == -1
    s = nodes["iStartOrRight"][leaf].astype(np.int64)
    e = s + nodes["nPrims"][leaf]
    order = np.argsort(s)
    assert s[order][0] == 0 and e[order][-1] == len(idx) and (s[order][1:] == e[order][:-1]).all()
    assert idx.max() < len(scene["tris"])
    for child in (inner + 1, right):
        assert (nodes["bmin"][child, :3] >= nodes["bmin"][inner, :3] - 1e-5).all()
        assert (nodes["bmax"][child, :3] <= nodes["bmax"][inner, :3] + 1e-5).all()
    # every triangle lies inside its leaf's box
    P = scene["tris"]["v"][:, :, 0, :3]
    li = np.nonzero(leaf)[0]
    for k in li[:: max(1, len(li) // 200)]:
        t = idx[nodes["iStartOrRight"][k]: nodes["iStartOrRight"][k] + nodes["nPrims"][k]]
        # SBVH leaves hold clipped references: the triangle must at least overlap the box
        assert (P[t].min(1) <= nodes["bmax"][k, :3] + 1e-4).all() and (P[t].max(1) >= nodes["bmin"][k, :3] - 1e-4).all()


def test_teapot_fixture_is_a_valid_reference_hierarchy(teapot):
    assert len(teapot["tris"]) == 3206 and len(teapot["nodes"]) == 2065 and len(teapot["indices"]) == 3290
    _check_node_invariants(teapot)
    assert (teapot["nodes"]["nPrims"] <= 8).all()  # SBVH leaf size limit (src/sbvh.hpp)


def test_replicate_grid_invariants(teapot):
    from fluctus_b200 import scene as fs
    g = fs.replicate_grid(teapot, (3, 2, 2), (2.2, 1.1, 1.4))
    assert len(g["tris"]) == 12 * 3206 and len(g["nodes"]) == 11 + 12 * 2065
    _check_node_invariants(g)
    # instance 0 is the untouched model
    assert np.array_equal(g["tris"]["v"][:3206], teapot["tris"]["v"])


def test_kitchen_standin_invariants(teapot):
    from fluctus_b200 import abi
    from fluctus_b200.standin import kitchen_standin
    scene, view = kitchen_standin(teapot, grid=(3, 2, 2))
    _check_node_invariants(scene)
    types = set(int(t) for t in scene["materials"]["type"])
    assert {abi.BXDF_DIFFUSE, abi.BXDF_GLOSSY, abi.BXDF_IDEAL_REFLECTION, abi.BXDF_GGX_ROUGH_REFLECTION,
            abi.BXDF_IDEAL_DIELECTRIC} <= types
    assert scene["tris"]["matId"].max() < len(scene["materials"])
    assert len(scene["textures"]) == 1 and scene["texels"].size == 256 * 256 * 4
    # deterministic: same bytes every time (the CPU and GPU arms of bench.py build it independently)
    again, _ = kitchen_standin(teapot, grid=(3, 2, 2))
    assert np.array_equal(scene["nodes"].view(np.uint8), again["nodes"].view(np.uint8))
    assert np.array_equal(scene["materials"].view(np.uint8), again["materials"].view(np.uint8))


def test_render_params_follow_reference_defaults():
    """src/settings.cpp:17-59 defaults and src/tracer.cpp:1127-1168 derived fields."""
    from fluctus_b200 import abi
    p = abi.make_render_params(1920, 1080)[0]
    assert p["width"] == 1920 and p["height"] == 1080
    assert p["width1"] == np.float32(1.0) / np.float32(1920) and p["height1"] == np.float32(1.0) / np.float32(1080)
    assert abs(float(p["camera"]["fovSCALE"]) - np.tan(np.radians(30.0))) < 1e-6
    assert tuple(p["camera"]["pos"][:3]) == (0.0, 1.0, 3.5) and tuple(p["camera"]["dir"][:3]) == (0.0, 0.0, -1.0)
    assert tuple(p["areaLight"]["E"][:3]) == (100.0, 100.0, 100.0) and tuple(p["areaLight"]["size"]) == (0.5, 0.5)
    assert p["sampleImpl"] == 1 and p["sampleExpl"] == 1 and p["useRoulette"] == 0 and p["wfSeparateQueues"] == 0
    assert p["maxBounces"] == 10 and p["tmOperator"] == 2 and p["maxSpp"] == 0


def test_world_radius_is_half_the_diagonal(teapot):
    from fluctus_b200 import scene as fs
    b = teapot["bounds"]
    assert abs(fs.world_radius(teapot) - 0.5 * float(np.linalg.norm(b[1] - b[0]))) < 1e-6


def test_shard_plan_and_ray_split():
    from fluctus_b200.multi_gpu import shard_plan, split_rays
    N, px = 1 << 20, 1920 * 1080
    plans = [shard_plan(r, 8, N, px, spp=256 + 3) for r in range(8)]
    assert [p["seed_offset"] for p in plans] == [r * N for r in range(8)]           # seeds of one 8N-task run
    assert [p["cursor_start"] for p in plans] == [r * px // 8 for r in range(8)]
    assert sum(p["spp"] for p in plans) == 259 and max(p["spp"] for p in plans) - min(p["spp"] for p in plans) == 1
    assert shard_plan(0, 1, N, px) == dict(seed_offset=0, cursor_start=0, spp=None)  # 1 GPU == the reference
    with pytest.raises(ValueError):
        shard_plan(2, 2, N, px)
    chunks = [split_rays((1 << 24) + 5, r, 8) for r in range(8)]
    assert chunks[0][0] == 0 and sum(c[1] for c in chunks) == (1 << 24) + 5
    assert all(chunks[i][0] + chunks[i][1] == chunks[i + 1][0] for i in range(7))


# ---- world_size 2 over gloo: rank-sharded oracle renders + one all-reduce ------------------------------
def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _rank_frame(rank, world, N, W, H, iters):
    """One rank's share of a frame, rendered by the CPU oracle with the rank's shard plan."""
    from oracle import pyoracle
    from fluctus_b200 import abi, scene as fs
    from fluctus_b200.multi_gpu import shard_plan
    teapot = fs.load_scene(os.path.join(ROOT, "tests", "golden", "teapot_sbvh.npz"))
    env = fs.load_envmap(os.path.join(ROOT, "tests", "golden", "night_env.npz"))
    plan = shard_plan(rank, world, N, W * H)
    p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=3,
                               use_env_map=True, use_area_light=False)
    pyoracle.set_threads(1)
    wf = pyoracle.OracleWavefront(N, W, H)
    wf.set_scene(teapot); wf.set_envmap(env); wf.set_params(p.tobytes())
    wf.set_options(seed_offset=plan["seed_offset"])
    wf.reset(); wf.clear_queues(); wf.set_pixel_index(0); wf.raygen(); wf.ext()
    wf.set_pixel_index(plan["cursor_start"])
    wf.render(iters, False)
    return wf.pixels().copy()


def _worker(rank, world, port, N, W, H, iters, out_dir):
    import torch.distributed as dist
    from fluctus_b200.multi_gpu import allreduce_pixels
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    px = _rank_frame(rank, world, N, W, H, iters)
    np.save(os.path.join(out_dir, "local%d.npy" % rank), px)
    allreduce_pixels(px)          # numpy path -> CPU tensor -> gloo
    np.save(os.path.join(out_dir, "sum%d.npy" % rank), px)
    dist.destroy_process_group()


def test_two_rank_frame_reduce_gloo(tmp_path):
    import torch.multiprocessing as mp
    N, W, H, iters = 512, 32, 24, 12
    port = _free_port()
    mp.spawn(_worker, args=(2, port, N, W, H, iters, str(tmp_path)), nprocs=2, join=True)
    l0, l1 = np.load(tmp_path / "local0.npy"), np.load(tmp_path / "local1.npy")
    s0, s1 = np.load(tmp_path / "sum0.npy"), np.load(tmp_path / "sum1.npy")
    assert np.array_equal(s0, s1)                       # every rank ends with the same frame
    assert np.array_equal(s0, l0 + l1)                  # two-operand float sum is order-independent
    assert s0[:, 3].sum() == l0[:, 3].sum() + l1[:, 3].sum() > 0   # .w carries the total sample count
    assert not np.array_equal(l0, l1)                   # ranks are decorrelated (seed offset + cursor start)
    # and each rank's share is exactly what a stand-alone run with that plan produces
    assert np.array_equal(l1, _rank_frame(1, 2, N, W, H, iters))
