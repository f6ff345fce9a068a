// C-ABI, second half: stage dispatch, the graph-replayed render loop, read-back, ray-cast entry
// points, state inspection.  See include/fluctus_cuda.h for the reference method each one replaces.
#include "flx_internal.h"
#include <algorithm>
#include <cmath>
#include <cstring>

#define FLX_CHECK_CTX(ctx) do { if (!(ctx)) return FLX_ERR_ARG; } while (0)
#define FLX_TRY(expr) do { int rc__ = (expr); if (rc__ != FLX_OK) return rc__; } while (0)

namespace {

struct TimerScope // CUDA-event pair around one stage launch, only when flx_enable_timing is on
{
    flx_ctx *ctx; int stage; cudaEvent_t e0 = nullptr, e1 = nullptr;
    TimerScope(flx_ctx *c, int s) : ctx(c), stage(s)
    {
        if (!ctx->timing) return;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0, ctx->stream);
    }
    ~TimerScope()
    {
        if (!e0) return;
        cudaEventRecord(e1, ctx->stream);
        ctx->timers[stage].pending.push_back(std::make_pair(e0, e1));
    }
};

int ensureScratch(flx_ctx *ctx, size_t bytes)
{
    if (ctx->scratchBytes >= bytes) return FLX_OK;
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->dScratch); ctx->dScratch = nullptr; ctx->scratchBytes = 0;
    FLX_CUDA(ctx, cudaMalloc(&ctx->dScratch, bytes));
    ctx->scratchBytes = bytes;
    return FLX_OK;
}

int stageLogic(flx_ctx *ctx, int first) { TimerScope t(ctx, 2); return flx_launch_logic(ctx, first); }
int stageLogicRegen(flx_ctx *ctx) { TimerScope t(ctx, 2); return flx_launch_logic_regen(ctx); }
int stageRaygen(flx_ctx *ctx) { TimerScope t(ctx, 3); return flx_launch_raygen(ctx); }
int stageMaterials(flx_ctx *ctx, bool renderEntry = false) { TimerScope t(ctx, 4); return flx_launch_materials(ctx, renderEntry); }
int stageExt(flx_ctx *ctx) { TimerScope t(ctx, 0); return flx_launch_ext(ctx); }
int stageShadow(flx_ctx *ctx) { TimerScope t(ctx, 1); return flx_launch_shadow(ctx); }
int stageExtShadow(flx_ctx *ctx) { TimerScope t(ctx, 0); return flx_launch_ext_shadow(ctx); } // timed under the extension stage

// one pass of the reference's benchmark inner loop (tracer.cpp:737-778)
int enqueueIteration(flx_ctx *ctx)
{
    FLX_TRY(stageLogicRegen(ctx)); // logic + raygen: terminated paths restart inside the logic kernel
    FLX_TRY(stageMaterials(ctx, true));
    if (ctx->counting) // the visit counters are kept per stage
    {
        FLX_TRY(stageExt(ctx));
        FLX_TRY(stageShadow(ctx));
        FLX_TRY(flx_launch_bookkeeping(ctx)); // getCounters + clear + updatePixelIndex, on the device
    }
    else FLX_TRY(stageExtShadow(ctx)); // both ray stages + the bookkeeping (last CTA) in one persistent launch
    return FLX_OK;
}

void *bufferPtr(flx_ctx *ctx, int which, size_t *bytes)
{
    const size_t np = ctx->fb.numPixels;
    switch (which)
    {
    case FLX_BUF_PIXELS: *bytes = np * 16; return ctx->fb.pixels;
    case FLX_BUF_PREVIEW: *bytes = np * 16; return ctx->fb.preview;
    case FLX_BUF_DENOISER_ALBEDO: *bytes = np * 16; return ctx->fb.denoiserAlbedo;
    case FLX_BUF_DENOISER_NORMAL: *bytes = np * 16; return ctx->fb.denoiserNormal;
    case FLX_BUF_SAMPLES_PER_PIXEL: *bytes = np * 4; return ctx->fb.spp;
    case FLX_BUF_DENOISER_ALBEDO_OUT: *bytes = np * 16; return ctx->fb.denoiserAlbedoOut;
    case FLX_BUF_DENOISER_NORMAL_OUT: *bytes = np * 16; return ctx->fb.denoiserNormalOut;
    }
    *bytes = 0;
    return nullptr;
}

} // namespace

extern "C" {

int flx_wf_reset(flx_ctx *ctx)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, false));
    FLX_CUDA(ctx, cudaMemsetAsync(ctx->dTotals, 0, sizeof(RenderTotals), ctx->stream));
    return flx_launch_reset(ctx);
}
int flx_wf_raygen(flx_ctx *ctx) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, false)); return stageRaygen(ctx); }
int flx_wf_logic(flx_ctx *ctx, int first) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, true)); return stageLogic(ctx, first); }
int flx_wf_materials(flx_ctx *ctx) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, true)); return stageMaterials(ctx); }
int flx_wf_ext(flx_ctx *ctx) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, true)); return stageExt(ctx); }
int flx_wf_shadow(flx_ctx *ctx) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, true)); return stageShadow(ctx); }
int flx_postprocess(flx_ctx *ctx) { FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, false)); return flx_launch_postprocess(ctx); }

int flx_wf_clear_queues(flx_ctx *ctx)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    FLX_CUDA(ctx, cudaMemsetAsync(ctx->qs.counters, 0, sizeof(flx_queue_counters), ctx->stream));
    return FLX_OK;
}

int flx_get_counters(flx_ctx *ctx, flx_queue_counters *out)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!out) return flx_fail(ctx, FLX_ERR_ARG, "flx_get_counters: null output");
    if (ctx->pendingCounterReads.size() >= 64) FLX_TRY(flx_finish(ctx)); // staging ring is full: drain it
    flx_queue_counters *slot = ctx->hCounters + ctx->pendingCounterReads.size();
    FLX_CUDA(ctx, cudaMemcpyAsync(slot, ctx->qs.counters, sizeof(flx_queue_counters), cudaMemcpyDeviceToHost, ctx->stream));
    ctx->pendingCounterReads.push_back(out);
    return FLX_OK;
}

int flx_finish(flx_ctx *ctx)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < ctx->pendingCounterReads.size(); i++) *ctx->pendingCounterReads[i] = ctx->hCounters[i];
    ctx->pendingCounterReads.clear();
    return FLX_OK;
}

int flx_render(flx_ctx *ctx, uint32_t iterations, int do_reset)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device)); FLX_TRY(flx_ready(ctx, true));
    if (do_reset) // resetRenderer + priming segment (tracer.cpp:633-645, :725-729)
    {
        FLX_TRY(flx_wf_reset(ctx));
        FLX_TRY(flx_wf_clear_queues(ctx));
        FLX_TRY(flx_reset_pixel_index(ctx));
        FLX_TRY(stageRaygen(ctx));
        FLX_TRY(stageExt(ctx));
    }
    if (iterations == 0) return FLX_OK;
    FLX_TRY(flx_reset_trace_cursors(ctx)); // k_ext_shadow expects (and leaves) them zero; a stage call may have used them
    if (ctx->timing) // per-stage events need individual launches
    {
        for (uint32_t i = 0; i < iterations; i++) FLX_TRY(enqueueIteration(ctx));
        return FLX_OK;
    }
    if (!ctx->graphValid)
    {
        cudaGraph_t graph = nullptr;
        FLX_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
        const uint64_t before = ctx->launches;
        const int rc = enqueueIteration(ctx);
        ctx->graphKernels = ctx->launches - before; // kernels in one iteration
        ctx->launches = before;                     // capture enqueues nothing
        cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
        if (rc != FLX_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) return flx_fail(ctx, FLX_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
        e = cudaGraphInstantiate(&ctx->graphExec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) return flx_fail(ctx, FLX_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
        ctx->graphValid = true;
    }
    for (uint32_t i = 0; i < iterations; i++)
    {
        FLX_CUDA(ctx, cudaGraphLaunch(ctx->graphExec, ctx->stream));
        ctx->launches += ctx->graphKernels;
    }
    return FLX_OK;
}

// ---- microkernel renderer (CLContext::enqueueResetKernel ... enqueueSplatPreviewKernel, clcontext.cpp:727-768) ----
#define FLX_MK_ENTRY(ctx, needScene) FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice((ctx)->device)); FLX_TRY(flx_ready(ctx, needScene))
int flx_mk_reset(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, false); return flx_launch_mk_reset(ctx); }
int flx_mk_raygen(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, false); return flx_launch_mk_raygen(ctx); }
int flx_mk_next_vertex(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, true); return flx_launch_mk_next_vertex(ctx); }
int flx_mk_sample_bsdf(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, true); return flx_launch_mk_sample_bsdf(ctx); }
int flx_mk_splat(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, false); return flx_launch_mk_splat(ctx); }
int flx_mk_splat_preview(flx_ctx *ctx) { FLX_MK_ENTRY(ctx, false); return flx_launch_mk_splat_preview(ctx); }

int flx_mk_get_stats(flx_ctx *ctx, flx_render_stats *out)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!out) return flx_fail(ctx, FLX_ERR_ARG, "flx_mk_get_stats: null output");
    FLX_CUDA(ctx, cudaMemcpyAsync(out, ctx->dStats, sizeof(*out), cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

// Tracer::renderSingle (tracer.cpp:108-182): every pixel gets exactly `spp` more samples
int flx_mk_render(flx_ctx *ctx, uint32_t spp, int do_reset)
{
    FLX_MK_ENTRY(ctx, true);
    if (ctx->params.maxBounces == 0u) return flx_fail(ctx, FLX_ERR_ARG, "flx_mk_render: maxBounces = 0 (unbounded paths) has no fixed schedule");
    if (ctx->params.useRoulette) return flx_fail(ctx, FLX_ERR_ARG, "flx_mk_render: switch Russian roulette off first (tracer.cpp:118-122)");
    if (do_reset) FLX_TRY(flx_launch_mk_reset(ctx));
    if (spp == 0) return FLX_OK;
    if (!ctx->mkGraphValid)
    {
        cudaGraph_t graph = nullptr;
        FLX_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
        const uint64_t before = ctx->launches;
        int rc = flx_launch_mk_raygen(ctx);
        for (uint32_t b = 0; rc == FLX_OK && b < ctx->params.maxBounces + 1u; b++)
        {
            rc = flx_launch_mk_next_vertex(ctx);
            if (rc == FLX_OK) rc = flx_launch_mk_sample_bsdf(ctx);
        }
        if (rc == FLX_OK) rc = flx_launch_mk_splat(ctx);
        ctx->mkGraphKernels = ctx->launches - before;
        ctx->launches = before;
        cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
        if (rc != FLX_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) return flx_fail(ctx, FLX_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
        e = cudaGraphInstantiate(&ctx->mkGraphExec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) return flx_fail(ctx, FLX_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
        ctx->mkGraphValid = true;
    }
    for (uint32_t s = 0; s < spp; s++)
    {
        FLX_CUDA(ctx, cudaGraphLaunch(ctx->mkGraphExec, ctx->stream));
        ctx->launches += ctx->mkGraphKernels;
    }
    return FLX_OK;
}

int flx_get_render_totals(flx_ctx *ctx, uint64_t totals[6])
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    RenderTotals t;
    FLX_CUDA(ctx, cudaMemcpyAsync(&t, ctx->dTotals, sizeof(t), cudaMemcpyDeviceToHost, ctx->stream));
    FLX_TRY(flx_finish(ctx));
    totals[0] = t.primary; totals[1] = t.extension; totals[2] = t.shadow; totals[3] = t.samples; totals[4] = t.splatted; totals[5] = t.iterations;
    return FLX_OK;
}

int flx_read_pixels(flx_ctx *ctx, int which, void *host_out, size_t bytes)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->fbBlock) return flx_fail(ctx, FLX_ERR_STATE, "flx_read_pixels: no frame buffers (flx_resize)");
    size_t have = 0;
    void *src = bufferPtr(ctx, which, &have);
    if (!src || !host_out || bytes > have) return flx_fail(ctx, FLX_ERR_ARG, "flx_read_pixels: bad buffer id or size");
    FLX_CUDA(ctx, cudaMemcpyAsync(host_out, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

int flx_write_pixels(flx_ctx *ctx, int which, const void *host_in, size_t bytes)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->fbBlock) return flx_fail(ctx, FLX_ERR_STATE, "flx_write_pixels: no frame buffers (flx_resize)");
    size_t have = 0;
    void *dst = bufferPtr(ctx, which, &have);
    if (!dst || !host_in || bytes > have) return flx_fail(ctx, FLX_ERR_ARG, "flx_write_pixels: bad buffer id or size");
    FLX_CUDA(ctx, cudaMemcpyAsync(dst, host_in, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return flx_finish(ctx);
}

void *flx_device_ptr(flx_ctx *ctx, int which)
{
    if (!ctx || !ctx->fbBlock) return nullptr;
    size_t b = 0;
    return bufferPtr(ctx, which, &b);
}

// ---- ray casting ------------------------------------------------------------------------------------
int flx_trace_closest_device(flx_ctx *ctx, const void *d_rays, size_t n, void *d_hit_i, void *d_hit_t, void *d_hits)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    TimerScope t(ctx, 5);
    return flx_launch_trace_batch(ctx, (const float4 *)d_rays, n, (int32_t *)d_hit_i, (float *)d_hit_t, (flx_hit *)d_hits);
}

int flx_trace_any_device(flx_ctx *ctx, const void *d_rays, size_t n, void *d_occluded)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    if (!d_occluded) return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_any: null output");
    TimerScope t(ctx, 5);
    return flx_launch_any_batch(ctx, (const float4 *)d_rays, n, (uint8_t *)d_occluded);
}

// Host-buffer ray casts are PCIe-bound: 32 B up and 8 B (+64 B with full hit records) down per ray against ~0.4 ns of
// traversal.  The batch is cut into chunks of FLX_RAY_CHUNK rays: chunk c+1 is uploaded on `copyIn` while chunk c is
// traversed on the context's stream and chunk c-1 is downloaded on `copyOut` (both copy engines and the SMs busy at
// the same time), so the call costs about max(upload, traversal, download) instead of their sum.
#ifndef FLX_RAY_CHUNK
#define FLX_RAY_CHUNK (1u << 20)
#endif
namespace {
struct RayPipe
{
    flx_ctx *ctx; size_t nChunks;
    int init(flx_ctx *c, size_t n)
    {
        ctx = c; nChunks = (n + FLX_RAY_CHUNK - 1) / FLX_RAY_CHUNK;
        if (!ctx->copyIn) FLX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copyIn, cudaStreamNonBlocking));
        if (!ctx->copyOut) FLX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copyOut, cudaStreamNonBlocking));
        while (ctx->pipeEvents.size() < 2 * nChunks + 1)
        {
            cudaEvent_t e; FLX_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ctx->pipeEvents.push_back(e);
        }
        // the copy streams start behind whatever is already queued on the context's stream (scratch reuse)
        cudaEvent_t start = ctx->pipeEvents[2 * nChunks];
        FLX_CUDA(ctx, cudaEventRecord(start, ctx->stream));
        FLX_CUDA(ctx, cudaStreamWaitEvent(ctx->copyIn, start, 0));
        FLX_CUDA(ctx, cudaStreamWaitEvent(ctx->copyOut, start, 0));
        return FLX_OK;
    }
    int uploaded(size_t c) // chunk c is on the device: the context's stream may traverse it
    {
        FLX_CUDA(ctx, cudaEventRecord(ctx->pipeEvents[2 * c], ctx->copyIn));
        FLX_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEvents[2 * c], 0));
        return FLX_OK;
    }
    int traced(size_t c) // results of chunk c are ready: the download stream may fetch them
    {
        FLX_CUDA(ctx, cudaEventRecord(ctx->pipeEvents[2 * c + 1], ctx->stream));
        FLX_CUDA(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEvents[2 * c + 1], 0));
        return FLX_OK;
    }
    int finish()
    {
        FLX_CUDA(ctx, cudaStreamSynchronize(ctx->copyOut));
        return flx_finish(ctx);
    }
};
} // namespace

int flx_trace_closest(flx_ctx *ctx, const float *rays, size_t n, int32_t *hit_i, float *hit_t, flx_hit *hits)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return FLX_OK;
    if (!rays) return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_closest: null rays");
    if (!ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    // sub-buffers on 256-byte boundaries: the hit records are written as float4 whatever the parity of n
    const size_t rayBytes = n * 32, iBytes = (n * 4 + 255) & ~(size_t)255, hBytes = hits ? n * sizeof(flx_hit) : 0;
    FLX_TRY(ensureScratch(ctx, rayBytes + 2 * iBytes + hBytes));
    uint8_t *base = (uint8_t *)ctx->dScratch;
    float4 *dRays = (float4 *)base; int32_t *dI = (int32_t *)(base + rayBytes); float *dT = (float *)(base + rayBytes + iBytes);
    flx_hit *dH = hits ? (flx_hit *)(base + rayBytes + 2 * iBytes) : nullptr;
    RayPipe pipe; FLX_TRY(pipe.init(ctx, n));
    for (size_t c = 0; c < pipe.nChunks; c++)
    {
        const size_t k0 = c * (size_t)FLX_RAY_CHUNK, m = std::min((size_t)FLX_RAY_CHUNK, n - k0);
        FLX_CUDA(ctx, cudaMemcpyAsync(dRays + 2 * k0, rays + 8 * k0, m * 32, cudaMemcpyHostToDevice, ctx->copyIn));
        FLX_TRY(pipe.uploaded(c));
        FLX_TRY(flx_trace_closest_device(ctx, dRays + 2 * k0, m, dI + k0, dT + k0, dH ? dH + k0 : nullptr));
        FLX_TRY(pipe.traced(c));
        if (hit_i) FLX_CUDA(ctx, cudaMemcpyAsync(hit_i + k0, dI + k0, m * 4, cudaMemcpyDeviceToHost, ctx->copyOut));
        if (hit_t) FLX_CUDA(ctx, cudaMemcpyAsync(hit_t + k0, dT + k0, m * 4, cudaMemcpyDeviceToHost, ctx->copyOut));
        if (hits) FLX_CUDA(ctx, cudaMemcpyAsync(hits + k0, dH + k0, m * sizeof(flx_hit), cudaMemcpyDeviceToHost, ctx->copyOut));
    }
    return pipe.finish();
}

int flx_trace_any(flx_ctx *ctx, const float *rays, size_t n, uint8_t *occluded)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return FLX_OK;
    if (!rays || !occluded) return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_any: null argument");
    if (!ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    const size_t rayBytes = n * 32;
    FLX_TRY(ensureScratch(ctx, rayBytes + n));
    uint8_t *base = (uint8_t *)ctx->dScratch, *dOcc = base + rayBytes;
    RayPipe pipe; FLX_TRY(pipe.init(ctx, n));
    for (size_t c = 0; c < pipe.nChunks; c++)
    {
        const size_t k0 = c * (size_t)FLX_RAY_CHUNK, m = std::min((size_t)FLX_RAY_CHUNK, n - k0);
        FLX_CUDA(ctx, cudaMemcpyAsync(base + 32 * k0, rays + 8 * k0, m * 32, cudaMemcpyHostToDevice, ctx->copyIn));
        FLX_TRY(pipe.uploaded(c));
        FLX_TRY(flx_trace_any_device(ctx, base + 32 * k0, m, dOcc + k0));
        FLX_TRY(pipe.traced(c));
        FLX_CUDA(ctx, cudaMemcpyAsync(occluded + k0, dOcc + k0, m, cudaMemcpyDeviceToHost, ctx->copyOut));
    }
    return pipe.finish();
}

// CLContext::pickSingle (clcontext.cpp:934-949) + kernel_pick.cl: one primary ray through the
// given NDC point.  The ray set-up (kernel_pick.cl:12-24) is a dozen flops, done here on the host.
int flx_pick(flx_ctx *ctx, float ndc_x, float ndc_y, flx_hit *out)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!out) return flx_fail(ctx, FLX_ERR_ARG, "flx_pick: null output");
    FLX_TRY(flx_ready(ctx, true));
    FLX_TRY(ensureScratch(ctx, sizeof(flx_hit)));
    FLX_TRY(flx_launch_pick(ctx, ndc_x, ndc_y, (flx_hit *)ctx->dScratch)); // ray set-up, trace and light test on the device
    FLX_CUDA(ctx, cudaMemcpyAsync(out, ctx->dScratch, sizeof(flx_hit), cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

// ---- state inspection ---------------------------------------------------------------------------------
int flx_export_state(flx_ctx *ctx, uint32_t *slots)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!slots) return flx_fail(ctx, FLX_ERR_ARG, "flx_export_state: null output");
    const size_t bytes = (size_t)FLX_NUM_SLOTS * ctx->N * 4;
    FLX_TRY(ensureScratch(ctx, bytes));
    FLX_CUDA(ctx, cudaMemsetAsync(ctx->dScratch, 0, bytes, ctx->stream));
    FLX_TRY(flx_launch_export_state(ctx, (uint32_t *)ctx->dScratch));
    FLX_CUDA(ctx, cudaMemcpyAsync(slots, ctx->dScratch, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

int flx_import_state(flx_ctx *ctx, const uint32_t *slots)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!slots) return flx_fail(ctx, FLX_ERR_ARG, "flx_import_state: null input");
    const size_t bytes = (size_t)FLX_NUM_SLOTS * ctx->N * 4;
    FLX_TRY(ensureScratch(ctx, bytes));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dScratch, slots, bytes, cudaMemcpyHostToDevice, ctx->stream));
    FLX_TRY(flx_launch_import_state(ctx, (const uint32_t *)ctx->dScratch));
    return flx_finish(ctx);
}

int flx_read_queue(flx_ctx *ctx, int queue, uint32_t *out, size_t n)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (queue < 0 || queue >= FLX_NUM_QUEUES || !out || n > ctx->N) return flx_fail(ctx, FLX_ERR_ARG, "flx_read_queue: bad arguments");
    if (n) FLX_CUDA(ctx, cudaMemcpyAsync(out, ctx->qs.q[queue], n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

int flx_write_queue(flx_ctx *ctx, int queue, const uint32_t *in, size_t n)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (queue < 0 || queue >= FLX_NUM_QUEUES || !in || n > ctx->N) return flx_fail(ctx, FLX_ERR_ARG, "flx_write_queue: bad arguments");
    if (n) FLX_CUDA(ctx, cudaMemcpyAsync(ctx->qs.q[queue], in, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    return flx_finish(ctx);
}

int flx_set_counters(flx_ctx *ctx, const flx_queue_counters *in)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!in) return flx_fail(ctx, FLX_ERR_ARG, "flx_set_counters: null input");
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->qs.counters, in, sizeof(*in), cudaMemcpyHostToDevice, ctx->stream));
    return flx_finish(ctx);
}

// ---- instrumentation ------------------------------------------------------------------------------------
int flx_enable_timing(flx_ctx *ctx, int enable)
{
    FLX_CHECK_CTX(ctx);
    flx_drain_timers(ctx);
    if (enable) for (int s = 0; s < 6; s++) { ctx->timers[s].ms = 0.0; ctx->timers[s].launches = 0; }
    ctx->timing = enable != 0;
    return FLX_OK;
}

int flx_set_trace_counting(flx_ctx *ctx, int enable)
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if ((enable != 0) != ctx->counting) flx_invalidate_graph(ctx);
    ctx->counting = enable != 0;
    if (enable) FLX_CUDA(ctx, cudaMemsetAsync(ctx->dTraceStats, 0, 20 * sizeof(unsigned long long), ctx->stream));
    return FLX_OK;
}

int flx_get_trace_counters(flx_ctx *ctx, uint64_t out[20])
{
    FLX_CHECK_CTX(ctx); FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!out) return flx_fail(ctx, FLX_ERR_ARG, "flx_get_trace_counters: null output");
    FLX_CUDA(ctx, cudaMemcpyAsync(out, ctx->dTraceStats, 20 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    return flx_finish(ctx);
}

int flx_get_kernel_time(flx_ctx *ctx, int stage, double *ms, uint64_t *launches)
{
    FLX_CHECK_CTX(ctx);
    if (stage < 0 || stage >= 6) return flx_fail(ctx, FLX_ERR_ARG, "flx_get_kernel_time: bad stage id");
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    flx_drain_timers(ctx);
    if (ms) *ms = ctx->timers[stage].ms;
    if (launches) *launches = ctx->timers[stage].launches;
    return FLX_OK;
}

} // extern "C"
