"""NCCL from the C-ABI — flx_comm_unique_id / flx_comm_init / flx_allreduce_pixels / flx_reduce_pixels (flx_comm.cu):
the per-frame sum of the accumulation buffers with ONE PROCESS PER GPU, without Python in the loop (SURVEY.md §8b/§8e;
the reference is single-device and has no collective).

One GPU is enough for the single-rank communicator (sum over one rank = identity, bit for bit) and the error paths;
the two-process tests need `gpurun --gpus 2` and compare `fluctus_headless --rank r --world 2` (two processes, ncclReduce)
with `fluctus_headless --gpus 2` (one process, peer-memory combine): same shards, same seeds, a two-term sum -> the
microkernel renderer's frame must be BIT-IDENTICAL, the wavefront renderer's sample counts must add up.
"""
import json
import subprocess

import numpy as np
import pytest

from test_headless_gpu import BIN, _binned, _gpu_count, _read_pfm, _run, _teapot_job

pytestmark = pytest.mark.gpu


def test_single_rank_communicator_sum_is_identity(teapot, night_env):
    from fluctus_b200 import CudaContext, Tracer
    from fluctus_b200.context import BUF_PIXELS
    tr = Tracer(128, 96, num_tasks=16384, max_bounces=4, use_env_map=True, use_area_light=False)
    tr.init(teapot, night_env)
    tr.render(spp=4, batch=8)
    ctx = tr.clctx
    before = ctx.readPixels(BUF_PIXELS).copy()
    assert before[:, 3].sum() >= 4 * 128 * 96
    assert ctx.commInfo()["world"] == 0
    with pytest.raises(Exception) as e:
        ctx.allreducePixels()
    assert "flx_comm_init" in str(e.value)
    uid = CudaContext.commUniqueId()
    assert len(uid) == 128 and any(uid)
    with pytest.raises(Exception):
        ctx.commInit(1, 1, uid)                       # rank outside the communicator
    ctx.commInit(0, 1, uid)
    info = ctx.commInfo()
    assert info["rank"] == 0 and info["world"] == 1 and info["nccl_version"] >= 22000, info
    with pytest.raises(Exception) as e:
        ctx.commInit(0, 1, uid)                       # one communicator per context
    assert "already" in str(e.value)
    ctx.allreducePixels(); ctx.finishQueue()
    assert np.array_equal(ctx.readPixels(BUF_PIXELS).view(np.uint32), before.view(np.uint32))
    ctx.reducePixels(0); ctx.finishQueue()
    assert np.array_equal(ctx.readPixels(BUF_PIXELS).view(np.uint32), before.view(np.uint32))
    with pytest.raises(Exception):
        ctx.reducePixels(1)
    # the collective is a stage on the render stream: render -> allreduce -> render keeps accumulating
    ctx.render(8)
    ctx.allreducePixels(); ctx.finishQueue()
    assert ctx.readPixels(BUF_PIXELS)[:, 3].sum() > before[:, 3].sum()
    ctx.commDestroy()
    assert ctx.commInfo()["world"] == 0
    ctx.commDestroy()                                 # idempotent
    tr.close()


def test_denoiser_aov_sums_travel_with_the_frame(teapot, night_env):
    from fluctus_b200 import CudaContext, Tracer
    from fluctus_b200.context import BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL, BUF_PIXELS
    tr = Tracer(64, 64, num_tasks=8192, max_bounces=3, use_env_map=True, use_area_light=False)
    tr.init(teapot, night_env)
    tr.clctx.setOptions(denoiser_aov=True, seed_offset=0)
    tr.render(spp=2, batch=8)
    ctx = tr.clctx
    bufs = [ctx.readPixels(w).copy() for w in (BUF_PIXELS, BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL)]
    ctx.commInit(0, 1, CudaContext.commUniqueId())
    ctx.allreducePixels(); ctx.finishQueue()
    for w, b in zip((BUF_PIXELS, BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL), bufs):
        assert np.array_equal(ctx.readPixels(w).view(np.uint32), b.view(np.uint32))
    tr.close()                                        # flx_destroy tears the communicator down


def _run_ranks(args, world, tmp_path, tag):
    comm = str(tmp_path / ("comm_%s.id" % tag))
    procs = [subprocess.Popen([BIN] + [str(a) for a in args] + ["--rank", str(r), "--world", str(world), "--comm-file", comm],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = []
    for r, p in enumerate(procs):
        try:
            o, e = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0, "rank %d: %s" % (r, e)
        outs.append(json.loads(o.strip().splitlines()[-1]))
    return outs


def test_two_processes_nccl_reduce_equals_in_process_combine(tmp_path, teapot, night_env):
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    flxs, cfg = _teapot_job(tmp_path, teapot, night_env)
    # microkernel renderer: deterministic per shard -> NCCL's two-term sum == the peer-memory kernel's, bit for bit
    mk = ["-x", 96, "-y", 64, "-s", 5, "--settings", cfg, "--renderer", "mk"]
    _run(mk + ["--gpus", 2, "--save-accum", tmp_path / "a.flxa", "--out", tmp_path / "a.pfm", flxs])
    stats = _run_ranks(mk + ["--save-accum", tmp_path / "b.flxa", "--out", tmp_path / "b.pfm", flxs], 2, tmp_path, "mk")
    a = np.fromfile(tmp_path / "a.flxa", np.uint8)[12:].view(np.uint32)
    b = np.fromfile(tmp_path / "b.flxa", np.uint8)[12:].view(np.uint32)
    assert np.array_equal(a, b)
    assert np.all(b.view(np.float32).reshape(64, 96, 4)[..., 3] == 5.0)     # 3 samples from rank 0 + 2 from rank 1
    assert np.array_equal(_read_pfm(str(tmp_path / "a.pfm")), _read_pfm(str(tmp_path / "b.pfm")))
    assert [s["rank"] for s in stats] == [0, 1] and all(s["world"] == 2 and s["nccl"] >= 22000 for s in stats)
    assert stats[0]["samples"] == 3 * 96 * 64 and stats[1]["samples"] == 2 * 96 * 64
    assert stats[0]["image"] and not stats[1]["image"]
    # wavefront renderer: every rank's samples are in rank 0's frame
    wf = ["-x", 96, "-y", 64, "-s", 32, "--settings", cfg]
    stats = _run_ranks(wf + ["--save-accum", tmp_path / "w.flxa", "--out", tmp_path / "w.pfm", flxs], 2, tmp_path, "wf")
    total = sum(s["samples"] for s in stats)
    acc = np.fromfile(tmp_path / "w.flxa", np.uint8)[12:].view(np.float32).reshape(64, 96, 4)
    assert total >= 32 * 96 * 64 and abs(acc[..., 3].sum() - total) <= 0.02 * total
    r1 = _run(wf + ["--out", tmp_path / "one.pfm", flxs])
    x, y = _binned(_read_pfm(str(tmp_path / "w.pfm"))), _binned(_read_pfm(str(tmp_path / "one.pfm")))
    assert float((((x - y) ** 2) / (y ** 2 + 1e-3)).mean()) < 0.05
