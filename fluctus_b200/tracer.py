"""Headless Tracer — the reference's render driver (src/tracer.cpp) without the window.

`update()` issues the stage sequence of Tracer::update (tracer.cpp:250-294, :333-336) call by
call; `render()` is the fixed-budget loop of Tracer::runBenchmarkFromFile (tracer.cpp:633-645,
:725-778) with the per-iteration host work (counter read-back, statistics, pixel cursor) moved
onto the device, so the host only polls once per batch of iterations.  Samples are sharded over
ranks by seed offset and pixel-cursor start; the float accumulation buffers are
summed by the caller with one all-reduce per frame (see bench.py / multi_gpu.py)."""
import numpy as np

from .abi import PARAMS_DT, make_render_params
from .context import BUF_PIXELS, BUF_PREVIEW, CudaContext
from .scene import world_radius


class Tracer:
    def __init__(self, width, height, num_tasks=1 << 20, device=0, rank=0, world_size=1, stream=None, **settings):
        self.rank, self.world_size = int(rank), int(world_size)
        self.clctx = CudaContext(num_tasks, device, stream)  # name kept from the reference (tracer.hpp:131)
        self.settings = dict(settings)
        # CHECK_SPP's per-pixel budget (samplesPerPixel) stays per context — the frame combine sums pixels only — so a
        # rank of a sample-sharded render takes its share of maxSpp, as fluctus_headless does (ceil(maxSpp / world))
        if self.world_size > 1 and int(self.settings.get("max_spp", 0)) > 0:
            self.settings["max_spp"] = -(-int(self.settings["max_spp"]) // self.world_size)
        self.params = make_render_params(width, height, **self.settings)
        self.iteration = 0
        self.stats = dict(primaryRays=0, extensionRays=0, shadowRays=0, samples=0)  # CLContext::statsAsync
        self.scene = None

    # Tracer::init (tracer.cpp:68-93): scene -> hierarchy -> worldRadius -> upload
    def init(self, scene, envmap=None):
        self.scene = scene
        p = self.params[0]
        p["n_tris"] = len(scene["tris"])
        p["worldRadius"] = world_radius(scene)
        if envmap is not None:
            self.clctx.createEnvMap(envmap)
        else:
            p["useEnvMap"] = 0
        self.clctx.uploadSceneData(scene)
        # decorrelate ranks: seeds of a single world_size*N-task run, cursor start spread over the image
        self.clctx.setOptions(denoiser_aov=bool(self.settings.get("denoiser_aov", False)),
                              seed_offset=self.rank * self.clctx.NUM_TASKS)
        self.clctx.updateParams(self.params)
        self.iteration = 0

    def setParams(self, **kw):
        for k, v in kw.items():
            self.params[0][k] = v
        self.clctx.updateParams(self.params)
        self.iteration = 0

    def _cursor_start(self):
        n = int(self.params[0]["width"]) * int(self.params[0]["height"])
        return (self.rank * n) // self.world_size

    # Tracer::update, wavefront branch (tracer.cpp:250-294 + :333-336, :361-368)
    def update(self):
        c, params = self.clctx, self.params
        maxBounces = int(params[0]["maxBounces"])
        N = 1
        if self.iteration == 0:
            params[0]["maxBounces"] = min(2, maxBounces)  # 2-bounce preview
            c.updateParams(params)
            N = 3
            c.setPixelIndex(self._cursor_start())
            c.enqueueWfResetKernel(params)
            c.enqueueWfRaygenKernel(params)
            c.enqueueWfExtRayKernel(params)
            c.enqueueClearWfQueues()
        cnt = None
        for _ in range(N):
            c.enqueueWfLogicKernel(params, self.iteration == 0)
            c.enqueueWfRaygenKernel(params)
            c.enqueueWfMaterialKernels(params)
            cnt = c.enqueueGetCounters()
            c.enqueueWfExtRayKernel(params)
            c.enqueueWfShadowRayKernel(params)
            c.enqueueClearWfQueues()
        if self.iteration == 0:
            params[0]["maxBounces"] = maxBounces
            c.updateParams(params)
        c.enqueuePostprocessKernel(params)
        c.finishQueue()
        c.updatePixelIndex(int(params[0]["width"]) * int(params[0]["height"]), int(cnt["raygenQueue"]))
        self.stats["extensionRays"] += int(cnt["extensionQueue"])
        self.stats["shadowRays"] += int(cnt["shadowQueue"])
        self.stats["primaryRays"] += int(cnt["raygenQueue"])
        self.stats["samples"] += int(cnt["raygenQueue"]) if self.iteration > 0 else 0
        self.iteration += 1
        return cnt

    # Tracer::runBenchmarkFromFile inner loop with a sample or iteration budget
    def render(self, spp=None, iterations=None, batch=16, reset=True, max_iterations=1 << 20):
        """Render until this rank has splatted spp * w * h samples (spp=None: exactly `iterations` passes).
        With spp, maxSpp must be 0 (unbounded per pixel) or >= spp.  Returns the device totals."""
        c = self.clctx
        if reset:
            c.setPixelIndex(0)
            c.render(0, reset=True)           # reset + clear + priming raygen + ext
            c.setPixelIndex(self._cursor_start())
        if spp is None:
            c.render(int(iterations))
            return c.getRenderTotals()
        target = int(spp) * int(self.params[0]["width"]) * int(self.params[0]["height"])
        key = "splatted" if int(self.params[0]["maxSpp"]) > 0 else "samples"
        done, last = 0, None
        while done < max_iterations:
            c.render(batch)
            done += batch
            t = c.getRenderTotals()
            if t[key] >= target:
                break
            # CHECK_SPP budget exhausted (every pixel saturated, a whole batch without a splat): the reference's stop rule
            # can then never be met — its first logic pass saturates samplesPerPixel[0] without splatting (DESIGN.md §7,
            # quirk 13), so the total tops out at target - maxSpp unless racing paths slip extra samples in
            if key == "splatted" and t[key] == last:
                break
            last = t[key]
        return c.getRenderTotals()

    # Tracer::renderSingle (tracer.cpp:108-182): final frame with a predefined spp on the microkernel renderer —
    # the only mode that guarantees the sample count of every pixel; images are deterministic
    def renderSingle(self, spp):
        """Exactly `spp` samples per pixel (this rank's share when world_size > 1: spp // world, remainder to low
        ranks).  Needs num_tasks >= width * height.  Returns the device RenderStats."""
        c, p = self.clctx, self.params[0]
        if c.NUM_TASKS < int(p["width"]) * int(p["height"]):
            raise ValueError("renderSingle: NUM_TASKS must cover the image (mk_raygen.cl:8-9)")
        if int(p["useRoulette"]):
            p["useRoulette"] = 0  # "Turning off russian roulette" (tracer.cpp:118-122)
            c.updateParams(self.params)
        share = int(spp) // self.world_size + (1 if self.rank < int(spp) % self.world_size else 0)
        c.renderSingle(share, reset=True)
        return c.getStats()

    def postprocess(self):
        self.clctx.enqueuePostprocessKernel(self.params)

    def pixels(self):
        """Raw accumulation buffer (h, w, 4): sum of radiance in rgb, sample count in a."""
        w, h = int(self.params[0]["width"]), int(self.params[0]["height"])
        return self.clctx.readPixels(BUF_PIXELS).reshape(h, w, 4)

    def preview(self):
        w, h = int(self.params[0]["width"]), int(self.params[0]["height"])
        self.clctx.enqueuePostprocessKernel(self.params)
        return self.clctx.readPixels(BUF_PREVIEW).reshape(h, w, 4)

    # CLContext::saveImage (clcontext.cpp:404-483): .png from the tonemapped preview (clamp, *255, truncate),
    # .hdr-like float dump from the raw buffer divided by the sample count
    def saveImage(self, filename):
        if filename.lower().endswith((".npy", ".hdr")):
            px = self.pixels()
            rgb = px[..., :3] / np.maximum(px[..., 3:4], 1.0)
            np.save(filename if filename.endswith(".npy") else filename + ".npy", rgb.astype(np.float32))
            return
        img = np.clip(self.preview()[..., :3], 0.0, 1.0)
        rgb8 = (img * 255.0).astype(np.uint8)[::-1]  # the reference's image origin is bottom-left
        from PIL import Image
        Image.fromarray(rgb8, "RGB").save(filename)

    def close(self):
        self.clctx.close()
