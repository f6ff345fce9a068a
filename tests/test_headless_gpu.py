"""The C++ host (fluctus_headless -> HeadlessTracer -> CudaContext -> C-ABI) renders the same image as the
Python mirror of the same loop, and both agree with the CPU oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "fluctus_b200", "fluctus_headless")


def _read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(x) for x in f.readline().split())
        assert float(f.readline()) < 0
        return np.frombuffer(f.read(), np.float32).reshape(h, w, 3)


def _binned(img, by=8):
    h, w = img.shape[:2]
    return img.reshape(h // by, by, w // by, by, 3).mean((1, 3))


def test_headless_render_matches_python_tracer_and_oracle(tmp_path, oracle, teapot, night_env):
    assert os.path.exists(BIN), "fluctus_headless not built"
    from fluctus_b200 import Tracer, scene as fs
    W, H, spp, bounces, N = 96, 64, 24, 4, 8192
    flxs, out, cfg = str(tmp_path / "teapot.flxs"), str(tmp_path / "img.pfm"), str(tmp_path / "settings.json")
    fs.save_flxs(flxs, teapot, night_env)
    json.dump({"release": {"wfBufferSize": N, "useAreaLight": False, "maxPathDepth": bounces, "tonemap": 2}}, open(cfg, "w"))
    r = subprocess.run([BIN, "-x", str(W), "-y", str(H), "-s", str(spp), "-b", "--settings", cfg, "--out", out, flxs],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    assert stats["samples"] >= spp * W * H and stats["kernel_launches"] > 0 and stats["triangles"] == 3206
    a = _read_pfm(out)

    tr = Tracer(W, H, num_tasks=N, max_bounces=bounces, use_env_map=True, use_area_light=False)
    tr.init(teapot, night_env)
    tr.render(spp=spp, batch=16)
    px = tr.pixels()
    b = px[..., :3] / np.maximum(px[..., 3:], 1)
    tr.close()

    wf = oracle.OracleWavefront(N, W, H)
    wf.set_scene(teapot); wf.set_envmap(night_env); wf.set_params(tr.params.tobytes())
    wf.render(stats["iterations"], True)
    po = wf.pixels().reshape(H, W, 4)
    c = po[..., :3] / np.maximum(po[..., 3:], 1)

    def relmse(x, y):
        x, y = _binned(x), _binned(y)
        return float((((x - y) ** 2) / (y ** 2 + 1e-3)).mean())
    assert relmse(a, b) < 0.03, relmse(a, b)   # C++ host vs Python host: same kernels, different atomics order
    assert relmse(a, c) < 0.05, relmse(a, c)   # vs the CPU oracle
    # the tonemapped PPM path works too
    ppm = str(tmp_path / "img.ppm")
    r = subprocess.run([BIN, "-x", str(W), "-y", str(H), "-s", "2", "--settings", cfg, "--out", ppm, flxs], capture_output=True, text=True)
    assert r.returncode == 0 and os.path.getsize(ppm) == len("P6\n%d %d\n255\n" % (W, H)) + W * H * 3


def test_headless_microkernel_mode_is_bit_identical_to_python_host(tmp_path, teapot, night_env):
    """`--renderer mk` = Tracer::renderSingle (the reference's own batch mode): exact spp, deterministic — so the C++
    host and the Python host must produce the same bits through the same C-ABI."""
    from fluctus_b200 import Tracer, scene as fs
    W, H, spp, bounces = 80, 48, 5, 3
    flxs, out, cfg = str(tmp_path / "teapot.flxs"), str(tmp_path / "mk.pfm"), str(tmp_path / "settings.json")
    fs.save_flxs(flxs, teapot, night_env)
    json.dump({"release": {"wfBufferSize": 1024, "useAreaLight": True, "maxPathDepth": bounces, "useRussianRoulette": True}}, open(cfg, "w"))
    r = subprocess.run([BIN, "-x", str(W), "-y", str(H), "-s", str(spp), "-b", "--renderer", "mk", "--settings", cfg, "--out", out, flxs],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Turning off russian roulette" in r.stdout       # tracer.cpp:118-122
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    assert stats["samples"] == spp * W * H == stats["primary"] and stats["iterations"] == spp
    a = _read_pfm(out)

    tr = Tracer(W, H, num_tasks=W * H, max_bounces=bounces, use_env_map=True, use_area_light=True, use_roulette=True)
    tr.init(teapot, night_env)
    st = tr.renderSingle(spp)
    px = tr.pixels()
    assert np.all(px[..., 3] == spp)
    b = px[..., :3] / np.maximum(px[..., 3:], 1)
    tr.close()
    assert st["extensionRays"] == stats["extension"] and st["shadowRays"] == stats["shadow"]
    assert np.array_equal(a.view(np.uint32), b.astype(np.float32).view(np.uint32))


def test_headless_gpu_built_hierarchy_renders_the_same_image(tmp_path, teapot, night_env):
    """`--lbvh N` swaps the scene's SBVH for a hierarchy built on the GPU (flx_build_lbvh).  The deterministic MK
    renderer then yields the same image except where a ray meets two triangles at exactly the same distance."""
    from fluctus_b200 import scene as fs
    W, H, spp = 96, 64, 4
    flxs, cfg = str(tmp_path / "teapot.flxs"), str(tmp_path / "settings.json")
    fs.save_flxs(flxs, teapot, night_env)
    json.dump({"release": {"wfBufferSize": W * H, "useAreaLight": True, "maxPathDepth": 3}}, open(cfg, "w"))
    imgs = []
    for extra in ([], ["--lbvh", "4"]):
        out = str(tmp_path / ("img%d.pfm" % len(imgs)))
        r = subprocess.run([BIN, "-x", str(W), "-y", str(H), "-s", str(spp), "--renderer", "mk", "--settings", cfg, "--out", out] + extra + [flxs],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        if extra:
            assert "rebuilt hierarchy on the GPU" in r.stderr
        imgs.append(_read_pfm(out))
    same = (imgs[0].view(np.uint32) == imgs[1].view(np.uint32)).all(-1)
    assert same.mean() >= 0.995, same.mean()
