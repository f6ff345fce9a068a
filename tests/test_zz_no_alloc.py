"""flx_trace_closest_static — the ray cast that allocates nothing (staging in a static device array; no reference
counterpart) — against the existing entry point flx_trace_closest,
through the C-ABI.  Tolerance: the one the ray-cast tests use everywhere, i.e. none — hit index equal, hit distance
bit-identical.

Collected last on purpose: a failure here must not hide other tests.  Self-contained:
it passes alone in a fresh process and after the rest of the suite (the static array is shared by every context of
the process and keeps the previous call's contents, which is exactly what the consecutive calls below exercise).
"""
import ctypes as C

import numpy as np
import pytest

from conftest import bits

pytestmark = pytest.mark.gpu

MAX_RAYS = 262144  # FLX_STATIC_MAX_RAYS, include/fluctus_cuda.h


def _ctx(scene, **kw):
    from fluctus_b200 import CudaContext
    ctx = CudaContext(num_tasks=1024, **kw)
    ctx.uploadSceneData(scene)
    return ctx


def test_static_scratch_ray_cast_matches_the_allocating_entry_point(teapot):
    from fluctus_b200 import scene as fs
    ctx = _ctx(teapot)
    # call 1: a full-size batch; call 2, straight after it: a smaller batch of DIFFERENT rays (the tail of the array
    # still holds call 1's rays and results, the work cursor holds call 1's end value); call 3: call 1 again
    batches = [fs.bench_rays(MAX_RAYS, teapot["bounds"]),
               fs.bench_rays(70001, teapot["bounds"], first=1 << 20),
               fs.bench_rays(MAX_RAYS, teapot["bounds"])]
    batches[1][:64, 4:7] = 0.0
    batches[1][:64, 4 + (np.arange(64) % 3)] = np.where(np.arange(64) % 2, 1.0, -1.0)  # axis-parallel rays (0 * inf in the slab test)
    batches[1][64:128, 3] = 0.75                                                        # finite tmax
    got = [ctx.traceClosestStatic(r) for r in batches]          # consecutive calls of the new entry point
    want = [ctx.traceClosest(r) for r in batches]               # the existing entry point
    for k, ((gi, gt), (wi, wt)) in enumerate(zip(got, want)):
        assert np.array_equal(gi, wi), "call %d: hit index differs for %d rays" % (k + 1, int((gi != wi).sum()))
        assert np.array_equal(bits(gt), bits(wt)), "call %d: hit distance bits differ" % (k + 1)
    assert 0.05 < (want[0][0] >= 0).mean() < 0.95 and 0.05 < (want[1][0] >= 0).mean() < 0.95  # hits and misses both present
    assert np.array_equal(got[0][0], got[2][0]) and np.array_equal(bits(got[0][1]), bits(got[2][1]))
    ctx.close()


def test_static_scratch_is_shared_between_contexts_and_scenes(teapot):
    """Two contexts on the device (different scenes) take turns on the one array."""
    from fluctus_b200 import scene as fs
    grid = fs.replicate_grid(teapot, (3, 2, 2), (2.2, 1.1, 1.4))
    a, b = _ctx(teapot), _ctx(grid)
    ra, rb = fs.bench_rays(50000, teapot["bounds"]), fs.bench_rays(120000, grid["bounds"])
    for _ in range(2):
        for ctx, rays in ((a, ra), (b, rb)):
            gi, gt = ctx.traceClosestStatic(rays)
            wi, wt = ctx.traceClosest(rays)
            assert np.array_equal(gi, wi) and np.array_equal(bits(gt), bits(wt))
    a.close(); b.close()


def test_static_scratch_limits_and_errors(teapot):
    import fluctus_b200 as fx
    from fluctus_b200 import CudaContext, scene as fs
    ctx = _ctx(teapot)
    hit_i, hit_t = ctx.traceClosestStatic(np.zeros((0, 8), np.float32))          # empty batch: nothing to do
    assert hit_i.size == 0 and hit_t.size == 0
    too_many = fs.bench_rays(MAX_RAYS + 1, teapot["bounds"])
    with pytest.raises(Exception) as e:                                             # beyond the maximum: a clear error, no work
        ctx.traceClosestStatic(too_many)
    assert "262144" in str(e.value) and "maximum" in str(e.value)
    L = fx.load()
    rays = fs.bench_rays(16, teapot["bounds"])
    out_i = np.empty(16, np.int32)
    rc = L.flx_trace_closest_static(ctx.h, rays.ctypes.data_as(C.c_void_p), 16, out_i.ctypes.data_as(C.c_void_p), None)
    assert rc != 0                                                                  # outputs are required
    gi, gt = ctx.traceClosestStatic(too_many[:MAX_RAYS])                            # the context stays usable after the errors
    wi, wt = ctx.traceClosest(too_many[:MAX_RAYS])
    assert np.array_equal(gi, wi) and np.array_equal(bits(gt), bits(wt))
    ctx.close()
    empty = CudaContext(num_tasks=1024)                                             # no scene uploaded
    with pytest.raises(Exception):
        empty.traceClosestStatic(rays)
    empty.close()
