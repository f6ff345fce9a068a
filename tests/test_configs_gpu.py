"""BASELINE.json's configurations rendered AS STATED (not at toy size) against the oracle.

cfg1 is the one configuration whose assets are all in the reference snapshot: assets/teapot.ply + env_maps/night.hdr,
512 x 512, 4 bounces, 16 spp, wfBufferSize 262144 (defaults of src/settings.cpp:17-59).  It is rendered
 (i) by the microkernel renderer — Tracer::renderSingle, what the reference's own `-b -s 16` batch mode runs
     (src/tracer.cpp:108-182) — compared PER PIXEL with the oracle's renderSingle (deterministic on both sides), and
 (ii) by the wavefront renderer with maxSpp = 16 (CHECK_SPP; Tracer::runBenchmarkFromFile's stop rule,
     src/tracer.cpp:713,730,773) — compared on 4 x 4 pixel blocks (its pixel <-> path mapping comes from atomics).
The kitchen stand-in (cfg4's scene, Country-Kitchen.obj is stripped) is rendered at the metric's 1920 x 1080, one
sample per pixel, and compared on 40 x 40 blocks.  The oracle runs on all host cores; ~1 minute for the file.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CFG1 = dict(max_bounces=4, use_env_map=True, use_area_light=False, sample_implicit=True, sample_explicit=True,
            use_roulette=False, tonemap=2)
CFG1_W = CFG1_H = 512
CFG1_TASKS = 262144
CFG1_SPP = 16


def _blocks(px, H, W, bs):
    px = px.reshape(H, W, 4)
    c = px[..., :3] / np.maximum(px[..., 3:], 1.0)
    return c.reshape(H // bs, bs, W // bs, bs, 3).mean((1, 3))


def _relmse(a, b):
    return float((((a - b) ** 2) / (b ** 2 + 1e-3)).mean())


def _all_cores(oracle):
    oracle.set_threads(max(oracle.max_threads(), len(os.sched_getaffinity(0))))


def _oracle_budget_render(wf, max_spp, max_iterations=4000):
    """Tracer::runBenchmarkFromFile's loop on the oracle until the CHECK_SPP budget is used up: the reference's stop rule
    (splatted >= w*h*maxSpp) is unreachable without racing extra samples, because its first logic pass saturates
    samplesPerPixel[0] without splatting (quirk 13) — so stop when every pixel is saturated."""
    done, first = 0, True
    while done < max_iterations:
        it, _ = wf.render(8, first)
        first = False
        done += it
        if it < 8 or (wf.spp() >= max_spp).all():
            break
    assert (wf.spp() >= max_spp).all(), "oracle did not use up its budget in %d iterations" % done
    return done


def test_cfg1_microkernel_render_per_pixel(oracle, teapot, night_env):
    from fluctus_b200 import Tracer
    tr = Tracer(CFG1_W, CFG1_H, num_tasks=CFG1_TASKS, **CFG1)
    tr.init(teapot, night_env)
    stats = tr.renderSingle(CFG1_SPP)
    img = tr.pixels().reshape(-1, 4)
    _all_cores(oracle)
    mk = oracle.OracleWavefront(CFG1_TASKS, CFG1_W, CFG1_H)
    mk.set_scene(teapot); mk.set_envmap(night_env); mk.set_params(tr.params.tobytes())
    mk.mk_render_single(CFG1_SPP)
    ref = mk.pixels()
    assert np.all(img[:, 3] == CFG1_SPP) and np.all(ref[:, 3] == CFG1_SPP)        # exactly 16 samples in every pixel
    assert stats["samples"] == int(mk.mk_stats()[3]) == CFG1_SPP * CFG1_W * CFG1_H
    for k, name in ((0, "primaryRays"), (1, "extensionRays"), (2, "shadowRays")):
        assert abs(stats[name] - int(mk.mk_stats()[k])) <= 2e-3 * int(mk.mk_stats()[k]), name
    # per pixel: identical arithmetic per task; libm ulps (sincos / atan2 / acos / pow) reach a few pixels through the
    # bilinear env lookup or a flipped `rand() < F`
    err = np.abs(img[:, :3].astype(np.float64) - ref[:, :3]) / np.maximum(np.abs(ref[:, :3]), 1e-2)
    off = (err > 1e-4).any(1)
    print("cfg1 MK: %.4f %% of pixels beyond 1e-4, %.4f %% beyond 1e-2" % (100 * off.mean(), 100 * (err > 1e-2).any(1).mean()))
    assert off.mean() < 0.01, off.mean()
    assert _relmse(_blocks(img, CFG1_H, CFG1_W, 4), _blocks(ref, CFG1_H, CFG1_W, 4)) < 1e-4
    tr.close()


def test_cfg1_wavefront_render_16spp(oracle, teapot, night_env):
    from fluctus_b200 import Tracer
    kw = dict(CFG1, max_spp=CFG1_SPP)
    tr = Tracer(CFG1_W, CFG1_H, num_tasks=CFG1_TASKS, **kw)
    tr.init(teapot, night_env)
    tot = tr.render(spp=CFG1_SPP)
    img = tr.pixels().reshape(-1, 4)
    spp = tr.clctx.readPixels(4)  # FLX_BUF_SAMPLES_PER_PIXEL
    _all_cores(oracle)
    wf = oracle.OracleWavefront(CFG1_TASKS, CFG1_W, CFG1_H)
    wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(tr.params.tobytes())
    it = _oracle_budget_render(wf, CFG1_SPP)
    ref = wf.pixels()
    target = CFG1_SPP * CFG1_W * CFG1_H
    assert tot["splatted"] >= target - CFG1_SPP - 64 and float(ref[:, 3].sum()) >= target - CFG1_SPP  # pixel 0 starves (quirk 13)
    # CHECK_SPP: every pixel ends with at least maxSpp samples; `<=` admits maxSpp + 1 (quirk 3) and the reference's
    # non-atomic clamp of the counter (wf_logic.cl:186-189, kept in flx_logic.cu) lets a path that had already passed the
    # first check slip in after it: with as many paths as pixels a few pixels reach maxSpp + 2.  Pixel 0 starves (quirk 13).
    w = img[1:, 3]
    assert w.min() >= CFG1_SPP and w.max() <= CFG1_SPP + 3, (w.min(), w.max())
    print("cfg1 WF: pixels with maxSpp+1 samples %.3f %%, with more %.4f %%" % (100 * (w == CFG1_SPP + 1).mean(), 100 * (w > CFG1_SPP + 1).mean()))
    assert (w > CFG1_SPP + 1).mean() < 2e-3
    assert np.all(spp.reshape(-1)[1:] == CFG1_SPP)
    r4 = _relmse(_blocks(img, CFG1_H, CFG1_W, 4), _blocks(ref, CFG1_H, CFG1_W, 4))
    r16 = _relmse(_blocks(img, CFG1_H, CFG1_W, 16), _blocks(ref, CFG1_H, CFG1_W, 16))
    print("cfg1 WF: relMSE 4x4 blocks %.5f, 16x16 blocks %.5f, GPU iterations %d, oracle iterations %d" % (r4, r16, tot["iterations"], it))
    # two independent 16-spp estimates: the difference of 4 x 4 block means (256 samples each) is Monte-Carlo noise.
    # VERDICT r1 asked for < 0.01 on 4 x 4 blocks; observed 0.00026 (and 0.00001 on 16 x 16): limits at ~4x observed
    assert r4 < 0.001, r4
    assert r16 < 1e-4, r16
    mg = img[:, :3].sum(0) / img[:, 3].sum(); mr = ref[:, :3].sum(0) / ref[:, 3].sum()
    assert np.allclose(mg, mr, rtol=0.01), (mg, mr)
    tr.close()


def test_kitchen_standin_full_resolution_one_spp(oracle, teapot, night_env):
    from fluctus_b200 import Tracer
    from fluctus_b200.standin import kitchen_standin
    scene, view = kitchen_standin(teapot)
    W, H = 1920, 1080
    kw = dict(max_bounces=8, use_env_map=True, use_area_light=True, separate_queues=True, max_spp=1, **view)
    tr = Tracer(W, H, num_tasks=1 << 20, **kw)
    tr.init(scene, night_env)
    tot = tr.render(spp=1)
    img = tr.pixels().reshape(-1, 4)
    _all_cores(oracle)
    wf = oracle.OracleWavefront(1 << 19, W, H)
    wf.set_scene(scene); wf.set_envmap(night_env); wf.set_params(tr.params.tobytes())
    it = _oracle_budget_render(wf, 1)
    ref = wf.pixels()
    assert tot["splatted"] >= W * H - 64 and float(ref[:, 3].sum()) >= W * H - 1
    assert img[1:, 3].min() >= 1.0 and img[1:, 3].max() <= 4.0 and (img[1:, 3] > 2.0).mean() < 2e-3
    assert set(np.unique(ref[1:, 3]).tolist()) <= {1.0, 2.0}
    r = _relmse(_blocks(img, H, W, 40), _blocks(ref, H, W, 40))
    print("kitchen stand-in 1080p, 1 spp: relMSE on 40x40 blocks %.5f (GPU %d iterations, oracle %d)" % (r, tot["iterations"], it))
    assert r < 0.04, r  # observed 0.0227: 1 600 one-sample pixels per block, two independent estimates
    mg = img[:, :3].sum(0) / img[:, 3].sum(); mr = ref[:, :3].sum(0) / ref[:, 3].sum()
    assert np.allclose(mg, mr, rtol=0.02), (mg, mr)
    tr.close()
