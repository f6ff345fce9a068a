"""CPU-only tests of the oracle (test infrastructure): committed golden vectors, the self-consistency
pairs of SURVEY.md §4 (stack == bit-stack == brute force; alias table reproduces the pdf; sampling
modes converge to the same image), and the reference quirks of SURVEY.md §8(g)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits


@pytest.fixture(scope="module")
def trace_golden():
    return np.load(os.path.join(GOLDEN, "trace_golden.npz"))


def test_trace_golden_vectors(oracle, teapot, trace_golden):
    g = trace_golden
    for mode in (0, 1, 2):  # stack, bit-stack (src/bvh.cl:12-230), brute force
        r = oracle.trace_closest(teapot, g["rays"], mode, want_hits=True)
        assert np.array_equal(r["hit_i"], g["hit_i"]), mode
        assert np.array_equal(bits(r["hit_t"]), bits(g["hit_t"])), mode
        if mode == 2:
            continue  # the brute-force loop reports index and distance only
        assert np.array_equal(bits(r["hits"]["P"]), bits(g["P"])), mode
        assert np.array_equal(bits(r["hits"]["N"]), bits(g["N"])), mode
        assert np.array_equal(bits(r["hits"]["uv"]), bits(g["uv"])), mode
        assert np.array_equal(r["hits"]["matId"], g["matId"]), mode
    for mode in (0, 1):
        assert np.array_equal(oracle.trace_any(teapot, g["any_rays"], mode), g["occluded"]), mode


def test_any_hit_consistent_with_closest(oracle, teapot, trace_golden):
    """occluded(tmax) <=> closest hit exists with 0 < t < tmax."""
    g = trace_golden
    rays = g["any_rays"]
    r = oracle.trace_closest(teapot, rays, 0)
    expect = (r["hit_i"] >= 0) & (r["hit_t"] < rays[:, 3])
    assert np.array_equal(g["occluded"].astype(bool), expect)


def test_traversal_variants_agree_on_replicated_grid(oracle, teapot):
    """Instanced hierarchy (exact-t ties between distinct triangles cannot occur across instances, but
    equal-entry-distance child ordering does): stack and bit-stack visit order give the same winner."""
    from fluctus_b200 import scene as fs
    grid = fs.replicate_grid(teapot, (3, 2, 2), (2.2, 1.1, 1.4))
    rays = fs.bench_rays(20000, grid["bounds"])
    a = oracle.trace_closest(grid, rays, 0, want_counters=True)
    b = oracle.trace_closest(grid, rays, 1)
    assert np.array_equal(a["hit_i"], b["hit_i"]) and np.array_equal(bits(a["hit_t"]), bits(b["hit_t"]))
    c = oracle.trace_closest(grid, rays[:1500], 2)
    assert np.array_equal(a["hit_i"][:1500], c["hit_i"])
    assert a["counters"]["max_stack"] < 64  # reference: uint stack[64] (src/bvh.cl:240)
    assert a["counters"]["inner"] > 0 and a["counters"]["tri"] > 0


def test_rng_matches_reference_hash(oracle):
    """Wang hash chain of src/random.cl:7-22, cross-checked against the numpy restatement that generates
    the benchmark rays (fluctus_b200.scene.bench_rays)."""
    seed = np.arange(16, dtype=np.uint32)
    s = seed.copy()
    s = (s ^ np.uint32(61)) ^ (s >> np.uint32(16)); s = s * np.uint32(9); s = s ^ (s >> np.uint32(4))
    s = s * np.uint32(0x27d4eb2d); s = s ^ (s >> np.uint32(15))
    from fluctus_b200 import scene as fs
    rays = fs.bench_rays(16, np.array([[0, 0, 0], [1, 1, 1]], np.float32))
    r0 = s.astype(np.float32) * np.float32(2.3283064365386963e-10)
    assert np.allclose(rays[:, 0], -0.25 + r0 * 1.5)
    # quirk 2: rand can return exactly 1.0f
    assert np.float32(np.uint32(0xffffff80)) * np.float32(2.3283064365386963e-10) == np.float32(1.0)


def test_wavefront_golden_state(oracle, teapot, night_env):
    g = np.load(os.path.join(GOLDEN, "wf_golden.npz"))
    N, W, H = g["state"].shape[1], 48, 32
    oracle.set_threads(1)
    try:
        wf = oracle.OracleWavefront(N, W, H)
        wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(g["params"].tobytes())
        it, st = wf.render(6, True)
        assert it == 6
        assert [st[k] for k in ("primary", "extension", "shadow", "samples")] == [int(x) for x in g["stats"]]
        assert wf.get_pixel_index() == int(g["pixel_index"])
        assert np.array_equal(wf.state(), g["state"])
        assert np.array_equal(bits(wf.pixels()), bits(g["pixels"]))
        assert np.array_equal(bits(wf.postprocess()), bits(g["preview"]))
    finally:
        oracle.set_threads(oracle.max_threads())


def test_threaded_oracle_same_image_statistics(oracle, teapot, night_env):
    """The OpenMP path (timing baseline) and the serial path (parity) accumulate the same samples."""
    from fluctus_b200 import abi, scene as fs
    N, W, H = 2048, 32, 32
    p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=3,
                               use_env_map=True, use_area_light=False)
    imgs = []
    for threads in (1, max(2, oracle.max_threads())):
        oracle.set_threads(threads)
        wf = oracle.OracleWavefront(N, W, H)
        wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(p.tobytes())
        wf.render(30, True)
        imgs.append(wf.pixels().copy())
    oracle.set_threads(oracle.max_threads())
    a, b = imgs
    assert abs(a[:, 3].sum() - b[:, 3].sum()) <= 0.02 * a[:, 3].sum()
    ma, mb = a[:, :3].sum(0) / a[:, 3].sum(), b[:, :3].sum(0) / b[:, 3].sum()
    assert np.allclose(ma, mb, rtol=0.05)


def test_alias_table_reproduces_pdf(night_env):
    """SURVEY.md §4 pair (5): alias-method tables (reference src/envmap.cpp:31-116) sample texel k with
    probability proportional to the pdf table's luminance*sin(theta) weights."""
    prob, alias, pdf = night_env["prob"], night_env["alias"], night_env["pdf"]
    n = len(prob)
    assert alias.min() >= 0 and alias.max() < n and prob.min() >= 0 and prob.max() <= 1.0 + 1e-6
    # exact sampling probability of texel k: (prob[k] + sum_{j: alias[j]==k} (1 - prob[j])) / n
    q = prob.astype(np.float64).copy()
    np.add.at(q, alias, 1.0 - prob.astype(np.float64))
    q /= n
    assert abs(q.sum() - 1.0) < 1e-6
    w = pdf.astype(np.float64)
    w = w / w.sum()
    # tables are built in float: allow small absolute error relative to the largest bin
    assert np.abs(q - w).max() <= 2e-3 * w.max() + 1e-9


def test_sampling_modes_converge(oracle, teapot, night_env):
    """SURVEY.md §4 pair (4): explicit-only, implicit-only and MIS renders agree in the mean."""
    from fluctus_b200 import abi, scene as fs
    N, W, H = 4096, 24, 16
    means = []
    for impl, expl in ((True, True), (True, False), (False, True)):
        p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=3,
                                   use_env_map=True, use_area_light=False, sample_implicit=impl, sample_explicit=expl)
        wf = oracle.OracleWavefront(N, W, H)
        wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(p.tobytes())
        wf.render(160, True)
        px = wf.pixels()
        means.append(px[:, :3].sum(0) / px[:, 3].sum())
    ref = means[0]
    # explicit-only omits directly visible environment at path length > 1 only through NEE, and the
    # primary-ray background (len == 1) is added in every mode (wf_logic.cl:106-107)
    assert np.allclose(means[1], ref, rtol=0.08), (means[1], ref)
    assert np.allclose(means[2], ref, rtol=0.08), (means[2], ref)


def test_benchmark_loop_pixel0_quirk(oracle, teapot, night_env):
    """Quirk 13: reset -> clear -> logic leaves every path at pixelIndex 0 / pathLen 0, so with
    CHECK_SPP samplesPerPixel[0] is saturated before pixel 0 receives a sample (wf_logic.cl:180-189)."""
    from fluctus_b200 import abi, scene as fs
    N, W, H, spp = 1024, 32, 32, 2
    p = abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), max_bounces=2,
                               max_spp=spp, use_env_map=True, use_area_light=False)
    oracle.set_threads(1)
    try:
        wf = oracle.OracleWavefront(N, W, H)
        wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(p.tobytes())
        wf.render(60, True)
        w = wf.pixels()[:, 3]
        assert w[0] == 0.0 and wf.spp()[0] == spp
        assert w[1:].min() >= spp and w.max() <= spp + 1  # `<=` admits maxSpp + 1 (quirk 3)
    finally:
        oracle.set_threads(oracle.max_threads())


def _mk_params(teapot, W, H, **kw):
    from fluctus_b200 import abi, scene as fs
    return abi.make_render_params(W, H, n_tris=len(teapot["tris"]), world_radius=fs.world_radius(teapot), **kw)


def test_microkernel_exact_spp_and_matches_wavefront(oracle, teapot, night_env):
    """SURVEY.md §4 pair (3): the MK renderer (src/mk_*.cl, Tracer::renderSingle) gives every pixel exactly
    `spp` samples, and its image agrees with the wavefront renderer's in the mean."""
    W, H, SPP = 24, 16, 48
    p = _mk_params(teapot, W, H, max_bounces=3, use_env_map=True, use_area_light=True)
    mk = oracle.OracleWavefront(W * H, W, H)
    mk.set_scene(teapot); mk.set_envmap(night_env); mk.set_params(p.tobytes())
    mk.mk_render_single(SPP)
    px = mk.pixels()
    assert np.all(px[:, 3] == SPP)
    st = mk.mk_stats()
    assert st[3] == SPP * W * H and st[0] == SPP * W * H      # samples, primary rays
    assert st[1] > 0 and st[2] > 0
    assert np.all(mk.state()[46] == 5)                        # every task back in MK_GENERATE_CAMERA_RAY
    assert np.isfinite(px).all()

    wf = oracle.OracleWavefront(4096, W, H)
    wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(p.tobytes())
    wf.render(200, True)
    q = wf.pixels()
    m_mk, m_wf = px[:, :3].sum(0) / px[:, 3].sum(), q[:, :3].sum(0) / q[:, 3].sum()
    # MK samples BOTH lights at every vertex, WF picks one with probability 1/2 and divides by it: same mean
    assert np.allclose(m_mk, m_wf, rtol=0.08), (m_mk, m_wf)


def test_microkernel_phases_and_check_spp(oracle, teapot, night_env):
    """Stage-by-stage phase machine (geom.h:205-216) and the CHECK_SPP branch of mk_splat.cl:21-32."""
    W, H = 16, 8
    p = _mk_params(teapot, W, H, max_bounces=2, use_env_map=True, use_area_light=False, max_spp=2)
    mk = oracle.OracleWavefront(W * H + 40, W, H)               # NUM_TASKS > w*h: the extra tasks stay idle
    mk.set_scene(teapot); mk.set_envmap(night_env); mk.set_params(p.tobytes())
    mk.mk_reset()
    ph = mk.state()[46]
    assert np.all(ph[:W * H] == 5) and np.all(mk.state()[49][:W * H] == np.arange(W * H))
    mk.mk_raygen()
    assert np.all(ph[:W * H] == 0)
    mk.mk_next_vertex()
    assert set(np.unique(ph[:W * H])) <= {1, 4} and np.all(mk.state()[48][:W * H] == 1)
    for s in range(5):
        if s:
            mk.mk_raygen()
        for b in range(3):
            mk.mk_next_vertex() if (s or b) else None
            mk.mk_sample_bsdf()
        assert set(np.unique(ph[:W * H])) <= {4, 6}
        mk.mk_splat()
    # `<=` on the pre-increment counter: maxSpp + 1 samples, then MK_DONE and the counter pinned to maxSpp
    assert np.all(mk.pixels()[:, 3] == 3) and np.all(mk.spp() == 2) and np.all(ph[:W * H] == 6)
    assert np.all(ph[W * H:] == 0) and np.all(mk.state()[48][W * H:] == 0)   # idle tasks untouched


def test_microkernel_golden_image(oracle, teapot, night_env):
    """tests/golden/mk_golden.npz (tools/make_golden_vectors.py mk): renderSingle state, image and RenderStats."""
    g = np.load(os.path.join(GOLDEN, "mk_golden.npz"))
    W, H = 48, 32
    oracle.set_threads(1)
    try:
        mk = oracle.OracleWavefront(W * H, W, H)
        mk.set_scene(teapot); mk.set_envmap(night_env); mk.set_params(g["params"].tobytes())
        mk.mk_render_single(int(g["spp"]))
        assert np.array_equal(mk.mk_stats(), g["stats"])
        assert np.array_equal(mk.state()[g["rows"]], g["state_rows"])
        assert np.array_equal(bits(mk.pixels()), bits(g["pixels"]))
    finally:
        oracle.set_threads(oracle.max_threads())
