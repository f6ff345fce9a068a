"""Lock-step parity of the kernels the BENCHMARK runs: flx_render's iteration = k_logic<REGEN> (logic + in-place restart
of terminated paths) -> k_materials_separate -> k_ext_shadow (extension + shadow rays + the iteration's bookkeeping),
replayed from a CUDA graph — against the oracle's logic -> raygen -> materials -> ext -> shadow -> clear
(reference: src/tracer.cpp:737-778, wf_logic.cl:171-205 + wf_raygen.cl:4-100, clcontext.cpp:903-921).

Every iteration starts from the oracle's state (pushed to the GPU), runs ONE graph iteration, and compares all fields
of all paths.  A restarted path receives the pixel `cursor + position in the raygen queue`, and queue positions come
from atomics (schedule-dependent in the reference as well): the GPU's assignment must be a
permutation of the oracle's pixel range, and the oracle then consumes the GPU's order, after which restarted paths are
compared field by field like all others.  Also here: Tracer::update's restart preview against an oracle `update()`.
"""
import numpy as np
import pytest

from conftest import cached_scene
from test_stages_gpu import ALL_FIELDS, MAX_COUNT_FRACTION, _compare_state, _push_oracle_to_gpu, _setup

pytestmark = pytest.mark.gpu


def _fused_lockstep(oracle, scene, env, N, W, H, iterations, warm=3, bitstack=False, **kw):
    from fluctus_b200 import abi
    from fluctus_b200.context import BUF_PIXELS, BUF_SPP
    wf, ctx, p = _setup(oracle, scene, env, N, W, H, **kw)
    ctx.setTraversal(bitstack)
    wf.set_options(bitstack=bitstack)
    npx = W * H
    RAYGEN = abi.QUEUES.index("raygen")
    pix_slot = abi.SLOTS["pixelIndex"]
    oracle.set_threads(1)
    try:
        # Tracer::runBenchmarkFromFile start (tracer.cpp:633-645, 725-729), then `warm` iterations on the oracle alone
        wf.reset(); wf.clear_queues(); wf.set_pixel_index(0); wf.raygen(); wf.ext()
        for _ in range(warm):
            wf.logic(False); wf.raygen(); wf.materials()
            n = int(wf.counters()["raygenQueue"])
            wf.ext(); wf.shadow(); wf.clear_queues(); wf.update_pixel_index(n)
        ctx.enqueueWfResetKernel()  # zeroes the device totals
        restarted_total = 0
        for it in range(iterations):
            tag = "fused iteration %d " % it
            _push_oracle_to_gpu(wf, ctx)
            cursor = wf.get_pixel_index()
            t0 = ctx.getRenderTotals()
            ctx.render(1)
            g = ctx.exportState()
            t1 = ctx.getRenderTotals()

            wf.logic(False)
            cnt = wf.counters().copy()
            n = int(cnt["raygenQueue"])
            T = wf.queue(RAYGEN)[:n].copy()
            # the pixels the GPU handed to the restarted paths: as a multiset, the oracle's (cursor + k) % (w*h), k < n
            # (a batch larger than the image wraps and gives some pixels two paths)
            rel = (g[pix_slot][T].astype(np.int64) - cursor) % npx
            assert np.array_equal(np.sort(rel), np.sort(np.arange(n) % npx)), tag + "restarted paths do not cover the pixel range"
            # the oracle consumes the GPU's queue order: the j-th path the GPU put on relative pixel r sits in slot r + j * w*h
            order = np.argsort(rel, kind="stable")
            r_sorted = rel[order]
            first_of_run = np.r_[0, np.nonzero(np.diff(r_sorted))[0] + 1]
            run_start = np.repeat(first_of_run, np.diff(np.r_[first_of_run, n]))
            slot = r_sorted + (np.arange(n) - run_start) * npx
            assert np.array_equal(np.sort(slot), np.arange(n))
            q = np.empty(n, np.uint32); q[slot] = T[order]
            wf.queue(RAYGEN)[:n] = q
            wf.raygen(); wf.materials()
            cnt2 = wf.counters().copy()
            wf.ext(); wf.shadow()
            bad = _compare_state(g, wf.state(), ALL_FIELDS, tag + "state", vector_norm=True)
            # restarted paths: integer fields (pixel, seed after the jitter draws, reset flags) exact whatever the tolerance story
            for f in ("pixelIndex", "seed", "pathLen", "lastSpecular", "firstDiffuseHit"):
                a, b = abi.state_field(g, f)[T], abi.state_field(wf.state(), f)[T]
                assert (a != b).sum() <= max(2, MAX_COUNT_FRACTION * N), tag + f
            # device bookkeeping == the host's (tracer.cpp:761-765): totals advance by the oracle's queue lengths
            tol = max(2, MAX_COUNT_FRACTION * N)
            assert t1["primary"] - t0["primary"] == n, tag + "primary"
            assert abs((t1["extension"] - t0["extension"]) - int(cnt2["extensionQueue"])) <= tol, tag + "extension"
            assert abs((t1["shadow"] - t0["shadow"]) - int(cnt2["shadowQueue"])) <= tol, tag + "shadow"
            assert t1["samples"] - t0["samples"] == (n if it > 0 else 0), tag + "samples"
            assert t1["iterations"] - t0["iterations"] == 1
            assert all(int(v) == 0 for v in ctx.getCounters().tolist()), tag + "queue counters not cleared by the last CTA"
            # splat: same terminated paths, same Ei
            px_g, px_r = ctx.readPixels(BUF_PIXELS), wf.pixels()
            assert np.allclose(px_g, px_r, rtol=1e-4, atol=1e-4), tag + "splat"
            assert float(px_g[:, 3].sum()) == float(px_r[:, 3].sum()), tag + "splat count"
            if int(p[0]["maxSpp"]) > 0:
                assert np.abs(ctx.readPixels(BUF_SPP).astype(np.int64) - wf.spp().astype(np.int64)).max() <= 1
            wf.clear_queues(); wf.update_pixel_index(n)
            restarted_total += n
        assert restarted_total > 0
        return wf, ctx
    finally:
        oracle.set_threads(oracle.max_threads())


def test_fused_iteration_teapot_envmap(oracle, teapot, night_env):
    wf, ctx = _fused_lockstep(oracle, teapot, night_env, 8192, 64, 64, 8, max_bounces=4, use_env_map=True, use_area_light=False)
    ctx.close()


def test_fused_iteration_teapot_both_lights_roulette_dof(oracle, teapot, night_env):
    wf, ctx = _fused_lockstep(oracle, teapot, night_env, 8192, 80, 48, 8, warm=7, max_bounces=0, use_roulette=True,
                              use_env_map=True, use_area_light=True, aperture=0.02, focal_dist=3.0)
    ctx.close()


@pytest.mark.parametrize("bitstack", [False, True])
def test_fused_iteration_kitchen_small(oracle, teapot, night_env, bitstack):
    """The benchmark's scene family and parameter set (kitchen stand-in, 8 bounces, MIS, both lights), small instance;
    more paths than pixels, so the cursor wraps inside one restart batch."""
    from fluctus_b200.standin import kitchen_standin
    scene, view = kitchen_standin(teapot, grid=(3, 2, 2))
    wf, ctx = _fused_lockstep(oracle, scene, night_env, 16384, 96, 54, 6, warm=5, bitstack=bitstack, max_bounces=8,
                              use_env_map=True, use_area_light=True, separate_queues=True, **view)
    ctx.close()


def test_fused_iteration_material_mix(oracle, night_env):
    from test_render_gpu import _material_mix_scene
    scene, view = _material_mix_scene()
    wf, ctx = _fused_lockstep(oracle, scene, night_env, 8192, 64, 64, 6, warm=4, max_bounces=6, use_env_map=True,
                              use_area_light=True, separate_queues=True, **view)
    ctx.close()


# ---- Tracer::update: restart preview (tracer.cpp:250-294) -----------------------------------------------------------------
def _oracle_update(wf, params, state):
    """One Tracer::update on the oracle: 2-bounce cap + firstIteration limit + three priming passes when iteration == 0."""
    from fluctus_b200 import abi
    p = params.copy()
    max_bounces = int(p[0]["maxBounces"])
    n_pass = 1
    first = state["iteration"] == 0
    if first:
        p[0]["maxBounces"] = min(2, max_bounces)
        wf.set_params(p.tobytes())
        n_pass = 3
        wf.set_pixel_index(0)
        wf.reset(); wf.raygen(); wf.ext(); wf.clear_queues()
    cnt = None
    for _ in range(n_pass):
        wf.logic(first); wf.raygen(); wf.materials()
        cnt = wf.counters().copy()
        wf.ext(); wf.shadow(); wf.clear_queues()
    if first:
        wf.set_params(params.tobytes())
    wf.update_pixel_index(int(cnt["raygenQueue"]))
    state["iteration"] += 1
    return cnt


@pytest.mark.parametrize("N", [2048, 6000])  # fewer / more paths than the 64 x 48 pixels: firstIteration caps logic at min(w*h, N)
def test_update_loop_matches_oracle_update(oracle, teapot, night_env, N):
    from fluctus_b200 import Tracer
    W, H, calls = 64, 48, 60
    kw = dict(max_bounces=5, use_env_map=True, use_area_light=True)
    tr = Tracer(W, H, num_tasks=N, **kw)
    tr.init(teapot, night_env)
    wf = oracle.OracleWavefront(N, W, H)
    wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(tr.params.tobytes())
    st = {"iteration": 0}
    tot_g = np.zeros(3, np.int64); tot_o = np.zeros(3, np.int64)
    for k in range(calls):
        cg = tr.update()
        co = _oracle_update(wf, tr.params, st)
        g = np.array([int(cg[q]) for q in ("raygenQueue", "extensionQueue", "shadowQueue")])
        o = np.array([int(co[q]) for q in ("raygenQueue", "extensionQueue", "shadowQueue")])
        if k == 0:
            # the three priming passes start from identical state, but the paths restarted in passes 1 and 2 get their
            # pixels in the order of the GPU's atomics: from then on the two sides trace different (equally distributed)
            # paths, and the lengths of the raygen / shadow queues of the last pass are two draws of the same binomial —
            # 6 sigma of their difference.  The extension queue holds every live path: exact.
            assert g[1] == o[1], (g, o)
            assert np.all(np.abs(g - o) <= 6.0 * np.sqrt(2.0 * np.maximum(o, 1)) + 4), (g, o)
        tot_g += g; tot_o += o
    # two independent Monte-Carlo runs (the GPU's path <-> pixel assignment comes from atomics): the totals are sums of
    # ~Poisson counts, so their difference has a standard deviation of about sqrt(2 * total) — 6 sigma + 1 %
    assert np.all(np.abs(tot_g - tot_o) <= 6.0 * np.sqrt(2.0 * tot_o) + 0.01 * tot_o + 64), (tot_g, tot_o)
    assert tr.iteration == st["iteration"] == calls
    img, ref = tr.pixels(), wf.pixels().reshape(H, W, 4)
    ws = float(ref[..., 3].sum())
    assert abs(float(img[..., 3].sum()) - ws) <= 6.0 * np.sqrt(2.0 * ws) + 0.01 * ws + 64
    a = (img[..., :3] / np.maximum(img[..., 3:], 1)).reshape(H // 8, 8, W // 8, 8, 3).mean((1, 3))
    b = (ref[..., :3] / np.maximum(ref[..., 3:], 1)).reshape(H // 8, 8, W // 8, 8, 3).mean((1, 3))
    relmse = float((((a - b) ** 2) / (b ** 2 + 1e-3)).mean())
    # 8 x 8 blocks hold ~64 * calls * min(N, w*h) * P(terminate) / (w*h) samples each (a few hundred at N = 2048): the
    # difference of two independent block means is noise of that size, observed 0.005 - 0.02 with a tail — a wrong update
    # sequence (no priming, no bounce cap, wrong first-iteration limit) shows up as > 0.3
    assert relmse < 0.1, relmse
    # preview restart: a parameter edit sets iteration back to 0 and the next update() clears the frame
    tr.setParams(exposure=1.5)
    assert tr.iteration == 0
    tr.update()
    assert float(tr.pixels()[..., 3].sum()) <= 3.0 * min(N, W * H) + 1
    tr.close()
