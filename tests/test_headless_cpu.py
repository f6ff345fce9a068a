"""CPU-only tests of the C++ host side (fluctus_b200/host): the headless binary builds and links the C-ABI
library, its scene preparation through the reference's own loaders reproduces the committed fixtures, the
.flxs container round-trips between C++ and Python, and without a GPU it fails loudly."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "fluctus_b200", "fluctus_headless")
REF = os.environ.get("FLUCTUS_REF", "/root/reference")


def _need_bin():
    if not os.path.exists(BIN):
        pytest.skip("fluctus_headless not built (python -c 'import __graft_entry__ as g; g.build()')")


def test_headless_links_the_c_abi_library():
    _need_bin()
    out = subprocess.check_output(["ldd", BIN], text=True)
    assert "libfluctus_cuda.so" in out and "oracle" not in out
    r = subprocess.run([BIN, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "usage: fluctus_headless" in r.stderr
    r = subprocess.run([BIN], capture_output=True, text=True)
    assert r.returncode == 2 and "no scene given" in r.stderr
    r = subprocess.run([BIN, "-x", "0", "scene.flxs"], capture_output=True, text=True)
    assert r.returncode == 2


def test_headless_one_process_per_gpu_flags():
    """--rank / --world / --comm-file (NCCL from the C++ host, flx_comm.cu): argument errors are caught before any work."""
    _need_bin()
    for args, msg in ((["--world", "2", "s.flxs"], "--comm-file"),
                      (["--world", "2", "--rank", "2", "--comm-file", "/tmp/x", "s.flxs"], "0 <= R < W"),
                      (["--world", "2", "--rank", "1", "--comm-file", "/tmp/x", "--gpus", "2", "s.flxs"], "excludes")):
        r = subprocess.run([BIN] + args, capture_output=True, text=True)
        assert r.returncode == 2 and msg in r.stderr, (args, r.stderr)
    # NCCL is not a link-time dependency of the library or the binary (dlopen at the first flx_comm_* call)
    lib = os.path.join(ROOT, "fluctus_b200", "libfluctus_cuda.so")
    assert "nccl" not in subprocess.check_output(["ldd", BIN], text=True) + subprocess.check_output(["ldd", lib], text=True)


def test_settings_and_benchmark_file_are_parsed_by_the_reference_settings_class(tmp_path):
    """settings.json / .bm.json go through the reference's own Settings::import (src/settings.cpp compiled in place):
    nested camera / areaLight objects, lookAt -> dir / up / right, scalar E, per-scene overrides on top of the file's
    base settings, envMap and outputFolder relative to the benchmark file (Tracer::runBenchmarkFromFile, tracer.cpp:567-630)."""
    import json
    _need_bin()
    cfg = tmp_path / "s.json"
    cfg.write_text(json.dumps({"release": {"maxPathDepth": 7, "wfBufferSize": 4096, "camera": {"pos": [1, 2, 3], "lookAt": [0, 0, 0], "fov": 40}}}))
    r = subprocess.run([BIN, "--settings", str(cfg), "--dump-settings"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["maxPathDepth"] == 7 and d["wfBufferSize"] == 4096 and d["tonemap"] == 2 and d["useAreaLight"] is True   # rest: Settings::init defaults
    if d["settings_by"] != "reference settings.cpp":
        pytest.skip("binary built without the reference tree: scalar keys only")
    want = -np.array([1, 2, 3.0]) / np.linalg.norm([1, 2, 3.0])
    assert np.allclose(d["camera"]["dir"], want, atol=1e-6) and d["camera"]["fov"] == 40 and d["camera"]["pos"] == [1, 2, 3]
    assert abs(np.dot(d["camera"]["dir"], d["camera"]["up"])) < 1e-6 and abs(np.dot(d["camera"]["dir"], d["camera"]["right"])) < 1e-6
    bm = tmp_path / "job.bm.json"
    bm.write_text(json.dumps({
        "settings": {"windowWidth": 320, "windowHeight": 200, "useWavefront": True, "maxSpp": 4, "maxPathDepth": 5, "envMap": "env/night.hdr",
                     "areaLight": {"pos": [0, 3, 0], "N": [0, -1, 0], "E": [50], "size": [1, 2]}},
        "outputFolder": "out",
        "scenes": [{"file": "a.ply", "outputFile": "first"},
                   {"file": "sub/b.obj", "outputFile": "second", "settings": {"useWavefront": False, "maxPathDepth": 3, "tonemap": 1}},
                   {"file": "c.ply", "outputFile": "third"}]}))
    r = subprocess.run([BIN, "--benchmark-file", str(bm), "--dump-settings"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    a, b, c = (json.loads(l) for l in r.stdout.strip().splitlines()[-3:])
    base = str(tmp_path) + "/"
    assert (a["scene"], b["scene"], c["scene"]) == (base + "a.ply", base + "sub/b.obj", base + "c.ply")
    assert a["output"] == base + "out/first" and (tmp_path / "out").is_dir() and a["envMap"] == base + "env/night.hdr"
    assert (a["windowWidth"], a["windowHeight"], a["maxSpp"], a["maxPathDepth"], a["useWavefront"]) == (320, 200, 4, 5, True)
    assert a["areaLight"]["E"] == [50, 50, 50] and a["areaLight"]["size"] == [1, 2] and a["areaLight"]["N"] == [0, -1, 0]
    assert (b["useWavefront"], b["maxPathDepth"], b["tonemap"]) == (False, 3, 1)
    # the base settings are imported again before every scene; a key only a scene set stays (one singleton, as in the reference)
    assert (c["useWavefront"], c["maxPathDepth"], c["tonemap"]) == (True, 5, 1)
    r = subprocess.run([BIN, "--benchmark-file", str(bm), "scene.flxs"], capture_output=True, text=True)
    assert r.returncode == 2 and "names its own scenes" in r.stderr


def test_flxs_roundtrip_python(tmp_path, teapot, night_env):
    from fluctus_b200 import scene as fs
    p = str(tmp_path / "t.flxs")
    fs.save_flxs(p, teapot, night_env)
    s, e = fs.load_flxs(p)
    for k in ("tris", "nodes", "materials"):
        assert np.array_equal(s[k].view(np.uint8), teapot[k].view(np.uint8)), k
    assert np.array_equal(s["indices"], teapot["indices"])
    for k in ("rgb", "prob", "alias", "pdf"):
        assert np.array_equal(e[k], night_env[k]), k
    with open(p, "r+b") as f:
        f.write(b"XXXX")
    with pytest.raises(ValueError):
        fs.load_flxs(p)


def test_headless_reference_loader_reproduces_fixture(tmp_path, teapot, night_env):
    """The C++ host loads teapot.ply + night.hdr with the reference's Scene / SBVH / EnvironmentMap code and
    writes a .flxs; it must hold the same arrays as the committed fixture (made through oracle/_ref from the
    same reference sources).  Then the render step must fail loudly here: there is no GPU and no CPU fallback."""
    _need_bin()
    if not os.path.isdir(os.path.join(REF, "assets")):
        pytest.skip("reference tree not present")
    from fluctus_b200 import scene as fs
    out = str(tmp_path / "teapot.flxs")
    r = subprocess.run([BIN, "--save-scene", out, "--env", os.path.join(REF, "assets/env_maps/night.hdr"),
                        os.path.join(REF, "assets/teapot.ply")], capture_output=True, text=True)
    import torch
    if not torch.cuda.is_available():
        assert r.returncode == 1 and "no CUDA device" in r.stderr
    s, e = fs.load_flxs(out)
    assert np.array_equal(s["tris"]["v"], teapot["tris"]["v"]) and np.array_equal(s["tris"]["matId"], teapot["tris"]["matId"])
    assert np.array_equal(s["indices"], teapot["indices"])
    for f in ("bmin", "bmax", "parent", "iStartOrRight", "nPrims"):
        assert np.array_equal(s["nodes"][f], teapot["nodes"][f]), f
    assert np.array_equal(s["materials"]["type"], teapot["materials"]["type"])
    for k in ("rgb", "prob", "alias", "pdf"):
        assert np.array_equal(e[k], night_env[k]), k


def test_headless_rejects_bad_scene_files(tmp_path):
    _need_bin()
    bad = tmp_path / "bad.flxs"
    bad.write_bytes(b"FLXS0001" + b"\0" * 64)
    r = subprocess.run([BIN, str(bad)], capture_output=True, text=True)
    assert r.returncode == 1 and "empty scene" in r.stderr
    r = subprocess.run([BIN, str(tmp_path / "missing.flxs")], capture_output=True, text=True)
    assert r.returncode == 1 and "cannot open" in r.stderr
