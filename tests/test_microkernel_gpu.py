"""The microkernel ("MK") renderer on CUDA against the CPU restatement of src/mk_*.cl (oracle/o_microkernel.cpp).

MK keeps one task per pixel and never lets two tasks touch the same pixel, so — unlike the wavefront renderer,
whose splat order comes from atomics — an MK image is a deterministic function of (scene, params, spp).  That
gives the strict image-level checks SURVEY.md §8(f)-2 asks for: per-stage state in lock step, per-pixel image
parity after Tracer::renderSingle, exact sample counts, and bit-identical images from run to run."""
import numpy as np
import pytest

from conftest import cached_scene
from test_stages_gpu import _compare_state, _setup, MAX_COUNT_FRACTION

pytestmark = pytest.mark.gpu

# fields the MK kernels own (the NEE slots are scratch on the CUDA side: shadow rays are parked there between
# the sub-steps of sampleBsdf, the reference traces them inline and never stores them)
MK_FIELDS = ["orig", "dir", "T", "Ei", "P", "N", "uvTex", "phase", "lastPdfW", "pathLen", "seed", "lastSpecular",
             "firstDiffuseHit", "t", "i", "areaLightHit", "matId"]


def _push(wf, ctx):
    from fluctus_b200.context import BUF_PIXELS, BUF_SPP, BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL
    ctx.importState(wf.state())
    ctx.writePixels(BUF_PIXELS, wf.pixels())
    ctx.writePixels(BUF_SPP, wf.spp())
    ctx.writePixels(BUF_DENOISER_ALBEDO, wf.denoiser_albedo())
    ctx.writePixels(BUF_DENOISER_NORMAL, wf.denoiser_normal())


def _lockstep(oracle, scene, env, N, W, H, samples, aov=False, **kw):
    """Tracer::renderSingle's schedule, stage by stage: the oracle's pre-stage state goes to the GPU, both run the
    stage, the post-stage state is compared.  Stats are compared as per-stage increments."""
    from fluctus_b200.context import BUF_PIXELS, BUF_SPP
    wf, ctx, p = _setup(oracle, scene, env, N, W, H, **kw)
    if aov:
        wf.set_options(denoiser_aov=True); ctx.setOptions(denoiser_aov=True)
    bounces = int(p[0]["maxBounces"])
    oracle.set_threads(1)
    try:
        wf.mk_reset(); ctx.enqueueResetKernel()
        _compare_state(ctx.exportState(), wf.state(), MK_FIELDS, "mk reset", allow_flips=False)
        assert np.array_equal(ctx.readPixels(BUF_PIXELS), wf.pixels())

        def stage(name, fo, fg, exact=False):
            _push(wf, ctx)
            s0, g0 = wf.mk_stats().astype(np.int64).copy(), ctx.getStats()
            fo(); fg()
            _compare_state(ctx.exportState(), wf.state(), MK_FIELDS, name, allow_flips=not exact)
            s1, g1 = wf.mk_stats().astype(np.int64), ctx.getStats()
            dg = np.array([g1[k] - g0[k] for k in ("primaryRays", "extensionRays", "shadowRays", "samples")])
            assert np.abs(dg - (s1 - s0)).max() <= max(2, MAX_COUNT_FRACTION * N), (name, dg, s1 - s0)

        for s in range(samples):
            stage("s%d raygen" % s, wf.mk_raygen, ctx.enqueueRayGenKernel)
            for b in range(bounces + 1):
                stage("s%d b%d nextVertex" % (s, b), wf.mk_next_vertex, ctx.enqueueNextVertexKernel)
                stage("s%d b%d sampleBsdf" % (s, b), wf.mk_sample_bsdf, ctx.enqueueBsdfSampleKernel)
            stage("s%d splat" % s, wf.mk_splat, ctx.enqueueSplatKernel, exact=True)
            px_g, px_r = ctx.readPixels(BUF_PIXELS), wf.pixels()
            assert np.array_equal(px_g, px_r), "splat of identical Ei must be bit-identical"
            assert np.array_equal(ctx.readPixels(BUF_SPP), wf.spp())
        return wf, ctx
    finally:
        oracle.set_threads(oracle.max_threads())


def test_mk_stages_teapot_env_and_area(oracle, teapot, night_env):
    """both lights: every vertex samples the environment AND the quad (mk_sample_bsdf.cl:75-140); NUM_TASKS > w*h"""
    wf, ctx = _lockstep(oracle, teapot, night_env, 64 * 48 + 100, 64, 48, 2, max_bounces=3, use_env_map=True, use_area_light=True)
    assert np.all(wf.pixels()[:, 3] == 2)
    ctx.close()


def test_mk_stages_env_only_check_spp_dof(oracle, teapot, night_env):
    """environment only, thin lens, CHECK_SPP branch of mk_splat.cl (maxSpp + 1 samples, then MK_DONE)"""
    wf, ctx = _lockstep(oracle, teapot, night_env, 48 * 32, 48, 32, 4, max_bounces=2, max_spp=2, use_env_map=True,
                        use_area_light=False, aperture=0.01, focal_dist=3.0)
    assert np.all(wf.pixels()[:, 3] == 3) and np.all(wf.spp() == 2)
    ctx.close()


def test_mk_stages_area_only_explicit_only(oracle, teapot):
    """area light only, implicit sampling off (no light quad in nextVertex, no MIS term)"""
    wf, ctx = _lockstep(oracle, teapot, None, 40 * 40, 40, 40, 2, max_bounces=3, use_env_map=False, use_area_light=True,
                        sample_implicit=False, light_pos=(1.5, 1.5, 0.5))
    ctx.close()


def test_mk_stages_no_nee(oracle, teapot, night_env):
    """explicit sampling off: sampleBsdf is a single launch"""
    wf, ctx = _lockstep(oracle, teapot, night_env, 32 * 32, 32, 32, 1, max_bounces=4, use_env_map=True, use_area_light=True,
                        sample_explicit=False)
    ctx.close()


def test_mk_stages_textured_aov(oracle, egyptcat, night_env):
    """egyptcat (textured glossy + diffuse, normal handling) with the denoiser AOVs on"""
    from fluctus_b200.context import BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL
    b = egyptcat["bounds"]
    c = 0.5 * (b[0] + b[1])
    cam = (float(c[0]), float(c[1]), float(b[1][2] + 1.5 * (b[1][2] - b[0][2])))
    wf, ctx = _lockstep(oracle, egyptcat, night_env, 48 * 48, 48, 48, 2, aov=True, max_bounces=3, use_env_map=True,
                        use_area_light=False, cam_pos=cam)
    assert np.allclose(ctx.readPixels(BUF_DENOISER_ALBEDO), wf.denoiser_albedo(), rtol=1e-5, atol=1e-6)
    assert np.allclose(ctx.readPixels(BUF_DENOISER_NORMAL), wf.denoiser_normal(), rtol=1e-5, atol=1e-6)
    ctx.close()


def _material_mix_scene():
    from fluctus_b200 import abi
    base = cached_scene("psor_cube")
    scene = dict(base)
    mats = np.tile(base["materials"][:1], 8)
    types = [abi.BXDF_DIFFUSE, abi.BXDF_GLOSSY, abi.BXDF_GGX_ROUGH_REFLECTION, abi.BXDF_IDEAL_REFLECTION,
             abi.BXDF_GGX_ROUGH_DIELECTRIC, abi.BXDF_IDEAL_DIELECTRIC, abi.BXDF_EMISSIVE, abi.BXDF_MIXED]
    for k, t in enumerate(types):
        mats[k]["type"] = t
        mats[k]["Kd"][:3] = (0.7, 0.5 + 0.05 * k, 0.3)
        mats[k]["Ks"][:3] = (0.9, 0.8, 0.7) if k != 1 else (0.0, 0.0, 0.0)
        mats[k]["Ke"][:3] = (2.0, 2.0, 2.0) if t == abi.BXDF_EMISSIVE else (0.0, 0.0, 0.0)
        mats[k]["Ns"] = 0.15 + 0.05 * k
        mats[k]["Ni"] = 1.5 if k != 1 else 0.0
    tris = base["tris"].copy()
    tris["matId"] = (np.arange(len(tris)) // 64) % 8
    scene.update(materials=mats, tris=tris)
    b = base["bounds"]
    c = 0.5 * (b[0] + b[1])
    cam = (float(c[0]), float(c[1]), float(b[1][2] + 1.2 * (b[1][2] - b[0][2])))
    light = (float(c[0]) + 2.0, float(b[1][1]) + 1.0, float(c[2]))
    return scene, cam, light


def test_mk_stages_material_mix(oracle, night_env):
    """all BxDF classes (incl. emissive with its unset pdf, and MIXED) through nextVertex / sampleBsdf"""
    scene, cam, light = _material_mix_scene()
    wf, ctx = _lockstep(oracle, scene, night_env, 56 * 56, 56, 56, 2, max_bounces=5, use_env_map=True, use_area_light=True,
                        cam_pos=cam, light_pos=light)
    ctx.close()


@pytest.mark.parametrize("which", ["teapot", "mix"])
def test_mk_render_single_image_parity(oracle, teapot, night_env, which):
    """Tracer::renderSingle end to end (graph replay on the GPU, plain loop in the oracle): per-PIXEL comparison
    of the accumulated radiance.  Differences can only come from libm ulps (sinf/cosf/atan2f/acosf/powf) — mostly
    below 1e-5, occasionally amplified by the bilinear env lookup, and in rare cases flipping a `rand() < F`
    branch of one sample; such pixels are counted and bounded."""
    from fluctus_b200 import Tracer
    from fluctus_b200.context import BUF_PIXELS
    W, H, SPP = 96, 64, 16
    if which == "teapot":
        scene, kw = teapot, dict(max_bounces=4, use_env_map=True, use_area_light=True)
    else:
        scene, cam, light = _material_mix_scene()
        kw = dict(max_bounces=5, use_env_map=True, use_area_light=True, cam_pos=cam, light_pos=light)
    tr = Tracer(W, H, num_tasks=W * H, **kw)
    tr.init(scene, night_env)
    stats = tr.renderSingle(SPP)
    g = tr.clctx.readPixels(BUF_PIXELS).astype(np.float64)

    wf = oracle.OracleWavefront(W * H, W, H)
    wf.set_scene(scene); wf.set_envmap(night_env); wf.set_params(tr.params.tobytes())
    wf.mk_render_single(SPP)
    r = wf.pixels().astype(np.float64)
    so = wf.mk_stats()

    assert np.all(g[:, 3] == SPP) and np.all(r[:, 3] == SPP)          # exact sample count per pixel
    assert stats["samples"] == SPP * W * H == int(so[3]) and stats["primaryRays"] == int(so[0])
    assert abs(stats["extensionRays"] - int(so[1])) <= 2e-3 * so[1] and abs(stats["shadowRays"] - int(so[2])) <= 2e-3 * so[2]
    with np.errstate(invalid="ignore"):
        rel = np.abs(g[:, :3] - r[:, :3]).max(1) / np.maximum(np.abs(r[:, :3]).max(1), 1e-3)
    assert np.isfinite(g).all() == np.isfinite(r).all()
    rel = np.nan_to_num(rel, nan=0.0)
    assert (rel > 1e-4).mean() <= 0.01, "pixels off by more than 1e-4: %d of %d" % ((rel > 1e-4).sum(), W * H)
    assert (rel > 1e-5).mean() <= 0.05
    fin = np.isfinite(g[:, :3]).all(1) & np.isfinite(r[:, :3]).all(1)
    relmse = np.mean((g[fin, :3] - r[fin, :3]) ** 2 / (r[fin, :3] ** 2 + 1e-3))
    assert relmse < 1e-4, relmse
    tr.close()


def test_mk_images_are_bit_reproducible(teapot, night_env):
    """Same scene, params and spp => the same bits, whatever order the warps ran in (two fresh contexts, one of
    them driven stage by stage instead of through the graph)."""
    from fluctus_b200 import Tracer
    from fluctus_b200.context import BUF_PIXELS
    W, H, SPP = 128, 96, 6
    kw = dict(max_bounces=3, use_env_map=True, use_area_light=True)
    imgs = []
    for mode in ("graph", "stages"):
        tr = Tracer(W, H, num_tasks=W * H + 5000, **kw)
        tr.init(teapot, night_env)
        c = tr.clctx
        if mode == "graph":
            tr.renderSingle(SPP)
        else:
            c.enqueueResetKernel()
            for _ in range(SPP):
                c.enqueueRayGenKernel()
                for _ in range(4):
                    c.enqueueNextVertexKernel(); c.enqueueBsdfSampleKernel()
                c.enqueueSplatKernel()
        imgs.append((c.readPixels(BUF_PIXELS).copy(), c.getStats()))
        tr.close()
    assert np.array_equal(imgs[0][0].view(np.uint32), imgs[1][0].view(np.uint32))
    assert imgs[0][1] == imgs[1][1]


def test_mk_exact_spp_full_resolution_and_preview(teapot, night_env):
    """BASELINE size (1920x1080, NUM_TASKS 2^21 >= w*h): size-independent properties — every pixel has exactly spp
    samples, RenderStats add up, the MK image agrees with the wavefront renderer's in the mean, and
    splatPreview overwrites with weight 0 (mk_splat_preview.cl)."""
    from fluctus_b200 import Tracer
    from fluctus_b200.context import BUF_PIXELS
    W, H, SPP = 1920, 1080, 4
    kw = dict(max_bounces=4, use_env_map=True, use_area_light=False)
    tr = Tracer(W, H, num_tasks=1 << 21, **kw)
    tr.init(teapot, night_env)
    st = tr.renderSingle(SPP)
    px = tr.clctx.readPixels(BUF_PIXELS)
    assert np.all(px[:, 3] == SPP) and np.isfinite(px).all()
    assert st["samples"] == SPP * W * H and st["primaryRays"] == SPP * W * H and st["extensionRays"] > 0 and st["shadowRays"] > 0
    mean_mk = (px[:, :3] / px[:, 3:4]).mean(0)

    # wavefront renderer, same context, same scene.  Its sample counts per pixel are uneven at any cut-off (pixels
    # whose paths end early are revisited sooner), so the comparison is between per-pixel normalised images.
    tr.render(spp=12)
    q = tr.clctx.readPixels(BUF_PIXELS)
    mean_wf = (q[:, :3] / np.maximum(q[:, 3:4], 1.0)).mean(0)
    assert np.allclose(mean_mk, mean_wf, rtol=0.03), (mean_mk, mean_wf)

    c = tr.clctx
    c.enqueueResetKernel(); c.enqueueRayGenKernel(); c.enqueueNextVertexKernel(); c.enqueueBsdfSampleKernel()
    c.enqueueNextVertexKernel(); c.enqueueBsdfSampleKernel(); c.enqueueSplatPreviewKernel()
    pv = c.readPixels(BUF_PIXELS)
    assert np.all(pv[:, 3] == 0) and pv[:, :3].sum() > 0
    assert np.all(c.exportState()[46][:W * H] == 5)
    tr.close()


def test_mk_against_committed_golden_image(teapot, night_env):
    """The committed fixture (tests/golden/mk_golden.npz, frozen oracle output): RenderStats and phases exact, image
    per pixel within the float tolerance."""
    import os
    from conftest import GOLDEN
    from fluctus_b200 import CudaContext, abi
    from fluctus_b200.context import BUF_PIXELS
    g = np.load(os.path.join(GOLDEN, "mk_golden.npz"))
    W, H = 48, 32
    ctx = CudaContext(num_tasks=W * H)
    ctx.uploadSceneData(teapot); ctx.createEnvMap(night_env)
    ctx.updateParams(np.frombuffer(g["params"].tobytes(), abi.PARAMS_DT))
    ctx.renderSingle(int(g["spp"]))
    st = ctx.getStats()
    assert st["samples"] == int(g["stats"][3]) and st["primaryRays"] == int(g["stats"][0])
    assert abs(st["extensionRays"] - int(g["stats"][1])) <= 4 and abs(st["shadowRays"] - int(g["stats"][2])) <= 8
    px, ref = ctx.readPixels(BUF_PIXELS).astype(np.float64), g["pixels"].astype(np.float64)
    assert np.array_equal(px[:, 3], ref[:, 3])
    rel = np.abs(px[:, :3] - ref[:, :3]).max(1) / np.maximum(np.abs(ref[:, :3]).max(1), 1e-3)
    assert (rel > 1e-4).mean() <= 0.01 and (rel > 1e-5).mean() <= 0.05, ((rel > 1e-4).sum(), (rel > 1e-5).sum())
    rows = list(g["rows"])
    assert np.array_equal(ctx.exportState()[46], g["state_rows"][rows.index(46)])       # phase
    assert (ctx.exportState()[49] != g["state_rows"][rows.index(49)]).mean() <= 2e-3    # seed: same number of RNG draws
    ctx.close()
