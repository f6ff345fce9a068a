"""CudaContext — host-side mirror of the reference's `class CLContext` (src/clcontext.hpp:26-215)
on top of the C-ABI.  Method names, argument meaning and error behaviour follow the reference so
that the render loop in tracer.py reads like Tracer::update / runBenchmarkFromFile."""
import ctypes as C

import numpy as np

from . import _lib
from .abi import COUNTERS_DT, HIT_DT, PARAMS_DT, QUEUES, TEX_DT
from ._lib import FluctusError, V

BUF_PIXELS, BUF_PREVIEW, BUF_DENOISER_ALBEDO, BUF_DENOISER_NORMAL, BUF_SPP, BUF_ALBEDO_OUT, BUF_NORMAL_OUT = range(7)
STAGE_EXT, STAGE_SHADOW, STAGE_LOGIC, STAGE_RAYGEN, STAGE_MATERIALS, STAGE_TRACE_BATCH = range(6)


def _p(a):
    return a.ctypes.data_as(V) if a is not None else None


class CudaContext:
    def __init__(self, num_tasks=1 << 20, device=0, stream=None):
        self.L = _lib.load()
        h = V()
        rc = self.L.flx_create(C.byref(h), int(device), int(num_tasks), V(stream) if stream else None)
        if rc != 0:
            raise FluctusError(self.L.flx_last_error(None).decode())
        self.h = h
        self.NUM_TASKS = int(num_tasks)
        self.width = self.height = 0
        self._pending = []  # keeps counter arrays alive until finishQueue

    # ---- plumbing
    def _check(self, rc):
        if rc != 0:
            raise FluctusError(self.L.flx_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.flx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def getNumTasks(self):
        return int(self.L.flx_num_tasks(self.h))

    @property
    def stream(self):
        return self.L.flx_stream(self.h)

    # ---- uploads
    def uploadSceneData(self, scene):
        """scene: dict with tris (TRI_DT), indices (u32), nodes (NODE_DT), materials (MAT_DT) and optionally
        textures (TEX_DT) + texels (u8) — what CLContext::uploadSceneData takes from BVH* and Scene*."""
        tris = np.ascontiguousarray(scene["tris"])
        idx = np.ascontiguousarray(scene["indices"], np.uint32)
        nodes = np.ascontiguousarray(scene["nodes"])
        mats = np.ascontiguousarray(scene["materials"])
        tex = np.ascontiguousarray(scene.get("textures", np.zeros(0, TEX_DT)))
        texels = np.ascontiguousarray(scene.get("texels", np.zeros(0, np.uint8)), np.uint8)
        assert tris.dtype.itemsize == 160 and nodes.dtype.itemsize == 48 and mats.dtype.itemsize == 96
        self._check(self.L.flx_upload_scene(self.h, _p(tris), len(tris), _p(idx), len(idx), _p(nodes), len(nodes),
                                            _p(mats), len(mats), _p(tex) if len(tex) else None, len(tex),
                                            _p(texels) if len(texels) else None, texels.nbytes))

    def createEnvMap(self, env):
        rgb = np.ascontiguousarray(env["rgb"], np.float32)
        h, w = rgb.shape[:2]
        prob = np.ascontiguousarray(env["prob"], np.float32)
        alias = np.ascontiguousarray(env["alias"], np.int32)
        pdf = np.ascontiguousarray(env["pdf"], np.float32)
        self._check(self.L.flx_upload_envmap(self.h, _p(rgb), w, h, _p(prob), _p(alias), _p(pdf)))

    def setupPixelStorage(self, width, height):
        self._check(self.L.flx_resize(self.h, int(width), int(height)))
        self.width, self.height = int(width), int(height)

    def updateParams(self, params):
        p = np.ascontiguousarray(params).view(np.uint8).reshape(-1)
        assert p.size == PARAMS_DT.itemsize
        self._check(self.L.flx_update_params(self.h, _p(p)))
        pr = p.view(PARAMS_DT)[0]
        self.width, self.height = int(pr["width"]), int(pr["height"])

    def setOptions(self, denoiser_aov=False, seed_offset=0):
        self._check(self.L.flx_set_options(self.h, int(denoiser_aov), int(seed_offset)))

    def setTraversal(self, use_bitstack):
        """Settings::clUseBitstack: stackless bit-trail traversal (src/bvh.cl:12-230) instead of the 64-entry stack."""
        self._check(self.L.flx_set_traversal(self.h, int(bool(use_bitstack))))

    def setPixelSharding(self, rank, world):
        """Preview sharding: this context traces only the image rows y % world == rank (world=1: off)."""
        self._check(self.L.flx_set_pixel_sharding(self.h, int(rank), int(world)))

    def recompileKernels(self, setArgs=False):
        """No-op: every specialisation is compiled ahead of time and picked by updateParams."""

    # ---- pixel cursor
    def resetPixelIndex(self): self._check(self.L.flx_reset_pixel_index(self.h))
    def setPixelIndex(self, v): self._check(self.L.flx_set_pixel_index(self.h, int(v)))
    def updatePixelIndex(self, numPixels, numNewPaths): self._check(self.L.flx_update_pixel_index(self.h, int(numPixels), int(numNewPaths)))

    # ---- stages (the params argument is kept for signature parity; the device copy is the one set by updateParams)
    def enqueueWfResetKernel(self, params=None): self._check(self.L.flx_wf_reset(self.h))
    def enqueueWfRaygenKernel(self, params=None): self._check(self.L.flx_wf_raygen(self.h))
    def enqueueWfLogicKernel(self, params=None, firstIteration=False): self._check(self.L.flx_wf_logic(self.h, int(bool(firstIteration))))
    def enqueueWfMaterialKernels(self, params=None): self._check(self.L.flx_wf_materials(self.h))
    def enqueueWfExtRayKernel(self, params=None): self._check(self.L.flx_wf_ext(self.h))
    def enqueueWfShadowRayKernel(self, params=None): self._check(self.L.flx_wf_shadow(self.h))
    def enqueueClearWfQueues(self): self._check(self.L.flx_wf_clear_queues(self.h))
    def enqueuePostprocessKernel(self, params=None): self._check(self.L.flx_postprocess(self.h))

    def enqueueGetCounters(self):
        """Asynchronous like the reference: the returned record is valid after finishQueue()."""
        cnt = np.zeros(1, COUNTERS_DT)
        self._pending.append(cnt)
        self._check(self.L.flx_get_counters(self.h, _p(cnt)))
        return cnt[0]

    def finishQueue(self):
        self._check(self.L.flx_finish(self.h))
        self._pending.clear()

    # ---- headless render entry
    def render(self, iterations, reset=False):
        self._check(self.L.flx_render(self.h, int(iterations), int(bool(reset))))

    def getRenderTotals(self):
        t = np.zeros(6, np.uint64)
        self._check(self.L.flx_get_render_totals(self.h, _p(t)))
        return dict(zip(("primary", "extension", "shadow", "samples", "splatted", "iterations"), (int(x) for x in t)))

    # ---- microkernel renderer (clcontext.cpp:727-768 -> src/mk_*.cl): one task per pixel, exact spp
    def enqueueResetKernel(self, params=None): self._check(self.L.flx_mk_reset(self.h))
    def enqueueRayGenKernel(self, params=None): self._check(self.L.flx_mk_raygen(self.h))
    def enqueueNextVertexKernel(self, params=None): self._check(self.L.flx_mk_next_vertex(self.h))
    def enqueueBsdfSampleKernel(self, params=None): self._check(self.L.flx_mk_sample_bsdf(self.h))
    def enqueueSplatKernel(self, params=None): self._check(self.L.flx_mk_splat(self.h))
    def enqueueSplatPreviewKernel(self, params=None): self._check(self.L.flx_mk_splat_preview(self.h))

    def renderSingle(self, spp, reset=True):
        """Tracer::renderSingle's loop (tracer.cpp:130-159) replayed from a CUDA graph: exactly `spp` samples per pixel."""
        self._check(self.L.flx_mk_render(self.h, int(spp), int(bool(reset))))

    def getStats(self):
        """Device RenderStats maintained by the MK kernels (synchronises)."""
        st = np.zeros(4, np.uint32)
        self._check(self.L.flx_mk_get_stats(self.h, _p(st)))
        return dict(zip(("primaryRays", "extensionRays", "shadowRays", "samples"), (int(x) for x in st)))

    # ---- results
    def readPixels(self, which=BUF_PIXELS, out=None):
        n = self.width * self.height
        if out is None:
            out = np.empty(n, np.uint32) if which == BUF_SPP else np.empty((n, 4), np.float32)
        self._check(self.L.flx_read_pixels(self.h, int(which), _p(out), out.nbytes))
        return out

    def writePixels(self, which, data):
        data = np.ascontiguousarray(data)
        self._check(self.L.flx_write_pixels(self.h, int(which), _p(data), data.nbytes))

    def devicePtr(self, which=BUF_PIXELS):
        return self.L.flx_device_ptr(self.h, int(which))

    def pickSingle(self, NDCx, NDCy):
        hit = np.zeros(1, HIT_DT)
        self._check(self.L.flx_pick(self.h, float(NDCx), float(NDCy), _p(hit)))
        return hit[0]

    # ---- ray casting
    def traceClosest(self, rays, want_hits=False):
        rays = np.ascontiguousarray(rays, np.float32)
        n = rays.shape[0]
        hit_i, hit_t = np.empty(n, np.int32), np.empty(n, np.float32)
        hits = np.zeros(n, HIT_DT) if want_hits else None
        self._check(self.L.flx_trace_closest(self.h, _p(rays), n, _p(hit_i), _p(hit_t), _p(hits)))
        return (hit_i, hit_t, hits) if want_hits else (hit_i, hit_t)

    # ---- one process per GPU: NCCL from the C-ABI (flx_comm.cu; the reference is single-device)
    @staticmethod
    def _prefer_bundled_nccl():
        """The library dlopen()s `libnccl.so.2` at the first flx_comm_* call and takes the copy the process already
        holds.  In a Python process that copy must be the one torch ships (site-packages/nvidia/nccl): were the
        system's older libnccl.so.2 loaded first, a later `import torch` would be handed it under the same soname and
        fail on missing symbols.  So load the bundled one now, unless torch (hence its NCCL) is already in."""
        import ctypes as C
        import importlib.util
        import os
        import sys
        if "torch" in sys.modules or os.environ.get("FLX_NCCL_LIB"):
            return
        spec = importlib.util.find_spec("nvidia")
        for root in (list(spec.submodule_search_locations) if spec and spec.submodule_search_locations else []):
            path = os.path.join(root, "nccl", "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return

    @staticmethod
    def commUniqueId():
        """128 bytes drawn by rank 0 and handed to every rank's commInit."""
        import ctypes as C
        from ._lib import load
        CudaContext._prefer_bundled_nccl()
        buf = C.create_string_buffer(128)
        if load().flx_comm_unique_id(buf) != 0:
            raise RuntimeError("flx_comm_unique_id failed: libnccl.so.2 not loadable")
        return buf.raw

    def commInit(self, rank, world, unique_id):
        assert len(unique_id) == 128
        self._prefer_bundled_nccl()
        self._check(self.L.flx_comm_init(self.h, int(rank), int(world), unique_id))

    def commDestroy(self):
        self._check(self.L.flx_comm_destroy(self.h))

    def commInfo(self):
        import ctypes as C
        r, w, v = C.c_int(0), C.c_int(0), C.c_int(0)
        self._check(self.L.flx_comm_info(self.h, C.byref(r), C.byref(w), C.byref(v)))
        return {"rank": r.value, "world": w.value, "nccl_version": v.value}

    def allreducePixels(self):
        self._check(self.L.flx_allreduce_pixels(self.h))

    def reducePixels(self, root=0):
        self._check(self.L.flx_reduce_pixels(self.h, int(root)))

    def traceClosestStatic(self, rays):
        """flx_trace_closest_static: no allocation on the call path (staging in a static device array, at most
        FLX_STATIC_MAX_RAYS = 262144 rays per call).  No reference counterpart."""
        rays = np.ascontiguousarray(rays, np.float32)
        n = rays.shape[0]
        hit_i, hit_t = np.empty(n, np.int32), np.empty(n, np.float32)
        self._check(self.L.flx_trace_closest_static(self.h, _p(rays), n, _p(hit_i), _p(hit_t)))
        return hit_i, hit_t

    def traceAny(self, rays):
        rays = np.ascontiguousarray(rays, np.float32)
        occ = np.empty(rays.shape[0], np.uint8)
        self._check(self.L.flx_trace_any(self.h, _p(rays), rays.shape[0], _p(occ)))
        return occ

    def traceClosestDevice(self, d_rays, n, d_hit_i=None, d_hit_t=None, d_hits=None):
        self._check(self.L.flx_trace_closest_device(self.h, V(d_rays), int(n), V(d_hit_i) if d_hit_i else None,
                                                    V(d_hit_t) if d_hit_t else None, V(d_hits) if d_hits else None))

    def traceAnyDevice(self, d_rays, n, d_occ):
        self._check(self.L.flx_trace_any_device(self.h, V(d_rays), int(n), V(d_occ)))

    # ---- state inspection (stage parity)
    def exportState(self):
        s = np.empty((64, self.NUM_TASKS), np.uint32)
        self._check(self.L.flx_export_state(self.h, _p(s)))
        return s

    def importState(self, state64):
        s = np.ascontiguousarray(state64, np.uint32)
        assert s.shape == (64, self.NUM_TASKS)
        self._check(self.L.flx_import_state(self.h, _p(s)))

    def readQueue(self, q, n):
        q = QUEUES.index(q) if isinstance(q, str) else int(q)
        out = np.empty(int(n), np.uint32)
        self._check(self.L.flx_read_queue(self.h, q, _p(out), int(n)))
        return out

    def writeQueue(self, q, data):
        q = QUEUES.index(q) if isinstance(q, str) else int(q)
        data = np.ascontiguousarray(data, np.uint32)
        self._check(self.L.flx_write_queue(self.h, q, _p(data), len(data)))

    def setCounters(self, cnt):
        c = np.ascontiguousarray(cnt).view(np.uint8).reshape(-1)
        assert c.size == 40
        self._check(self.L.flx_set_counters(self.h, _p(c)))

    def getCounters(self):
        """Synchronous convenience: enqueueGetCounters + finishQueue."""
        c = self.enqueueGetCounters()
        self.finishQueue()
        return c.copy()

    # ---- instrumentation
    def launchCount(self): return int(self.L.flx_launch_count(self.h))
    def enableTiming(self, on=True): self._check(self.L.flx_enable_timing(self.h, int(bool(on))))

    def setTraceCounting(self, on=True): self._check(self.L.flx_set_trace_counting(self.h, int(bool(on))))

    def traceCounters(self):
        """dict group -> {inner, leaf, tri, hits, rays}; groups ext, shadow, batch_closest, batch_any."""
        c = np.zeros(20, np.uint64)
        self._check(self.L.flx_get_trace_counters(self.h, _p(c)))
        names = ("inner", "leaf", "tri", "hits", "rays")
        return {g: dict(zip(names, (int(x) for x in c[5 * k:5 * k + 5])))
                for k, g in enumerate(("ext", "shadow", "batch_closest", "batch_any"))}

    def kernelTime(self, stage):
        ms, n = C.c_double(), C.c_uint64()
        self._check(self.L.flx_get_kernel_time(self.h, int(stage), C.byref(ms), C.byref(n)))
        return ms.value, n.value
