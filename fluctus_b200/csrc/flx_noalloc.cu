// flx_trace_closest_static — the host-buffer closest-hit ray cast for callers that must not allocate.
//
// No counterpart in the reference.  Some callers cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  flx_trace_closest
// (flx_api.cu) stages the rays and the results in a scratch block that it (re)allocates with cudaMalloc whenever a
// batch is larger than the last one.  This variant keeps that scratch in a file-scope `static __device__` array of
// fixed size instead, so the call path contains no cudaMalloc / cudaFree / allocating library routine:
//
//     cudaMemsetAsync (work cursor) -> cudaMemcpyAsync H2D (rays) -> k_trace_batch_static -> cudaMemcpyAsync D2H -> sync
//
// all on the context's stream.  Cost: every GPU that loads the library holds FLX_STATIC_SCRATCH_BYTES (10.0 MiB)
// until the process exits, and a batch is limited to FLX_STATIC_MAX_RAYS (2^18) rays.
//
// The array is ONE per device and process: every context on a device shares it and it keeps the previous call's
// contents.  What must start at zero (the work cursor of the persistent kernel) is cleared by every call; ray and
// result lanes beyond `n` are never read.  Calls that overlap on the GPU (two contexts with different streams on
// the same device, or two host threads) race on it: callers serialise them.  The entry point itself synchronises
// the context's stream before it returns, so back-to-back calls from one thread are safe.
//
// The traversal is traceLoop of flx_bvh.cuh — the same code k_trace_batch runs — with the stack trail (the
// reference's default, src/bvh.cl:234-310): hit index and distance are bit-identical to flx_trace_closest.
#include "flx_internal.h"
#include "../../include/fluctus_cuda.h"
#include <mutex>

#define FLX_STATIC_HEADER_BYTES 256u // [0, 256): word 0 = work cursor of the persistent kernel, rest unused
#define FLX_STATIC_SCRATCH_BYTES (FLX_STATIC_HEADER_BYTES + (size_t)FLX_STATIC_MAX_RAYS * (32u + 4u + 4u))
#define FLX_STATIC_THREADS 640
#define FLX_STATIC_CTAS_PER_SM 2
#define FLX_STATIC_MAX_DEVICES 64

// In HOST code this name is a host-side stand-in, not a device address: handing it to a kernel faults (or, where the
// GPU can reach host memory, silently uses the wrong memory).  Host code below only ever passes the pointer that
// cudaGetSymbolAddress returned for the current device.
static __device__ __align__(256) unsigned char g_staticScratch[FLX_STATIC_SCRATCH_BYTES];

namespace {

struct StaticLayout
{
    unsigned int *cursor; const float4 *rays; int32_t *hitI; float *hitT;
};
__host__ __device__ inline StaticLayout layoutOf(unsigned char *base)
{
    StaticLayout l;
    l.cursor = (unsigned int *)base;
    l.rays = (const float4 *)(base + FLX_STATIC_HEADER_BYTES);
    l.hitI = (int32_t *)(base + FLX_STATIC_HEADER_BYTES + (size_t)FLX_STATIC_MAX_RAYS * 32u);
    l.hitT = (float *)(base + FLX_STATIC_HEADER_BYTES + (size_t)FLX_STATIC_MAX_RAYS * 36u);
    return l;
}

// rays as 2 x float4 {o.xyz, tmax} {d.xyz, -}; results: triangle index (-1 = miss) and distance
struct StaticBatchJob
{
    static constexpr bool BATCH = true;
    const DevBVH &bvh; const float4 *rays; int32_t *hitI; float *hitT;
    unsigned int k;
    FLX_DEV bool load(unsigned int kk, f3 &o, f3 &d, float &tmax)
    {
        k = kk;
        const float4 r0 = rays[2 * (size_t)kk], r1 = rays[2 * (size_t)kk + 1];
        o = mk3(r0); d = mk3(r1); tmax = r0.w;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3, f3)
    {
        hitI[k] = h.slot >= 0 ? __float_as_int(__ldg(bvh.slots + 3 * (size_t)h.slot).w) : -1;
        hitT[k] = h.t;
    }
};

__global__ void __launch_bounds__(FLX_STATIC_THREADS, FLX_STATIC_CTAS_PER_SM)
k_trace_batch_static(DevBVH bvh, unsigned char *scratch, unsigned int n)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    const StaticLayout l = layoutOf(scratch);
    StaticBatchJob job = {bvh, l.rays, l.hitI, l.hitT, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<false, false, StackTrail>(bvh, topNodes, job, l.cursor, n, cnt);
}

// Set-up, once per device, with that device current: the device address of the array.
std::mutex g_addrLock;
unsigned char *g_addr[FLX_STATIC_MAX_DEVICES] = {};

int scratchAddress(flx_ctx *ctx, unsigned char **out)
{
    if (ctx->device < 0 || ctx->device >= FLX_STATIC_MAX_DEVICES)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_closest_static: device index out of range");
    std::lock_guard<std::mutex> guard(g_addrLock);
    if (!g_addr[ctx->device])
    {
        void *p = nullptr;
        FLX_CUDA(ctx, cudaGetSymbolAddress(&p, g_staticScratch)); // the current device is ctx->device (set by the caller)
        if (!p) return flx_fail(ctx, FLX_ERR_CUDA, "flx_trace_closest_static: cudaGetSymbolAddress returned a null address");
        g_addr[ctx->device] = (unsigned char *)p;
    }
    *out = g_addr[ctx->device];
    return FLX_OK;
}

} // namespace

extern "C" int flx_trace_closest_static(flx_ctx *ctx, const float *rays, size_t n, int32_t *hit_i, float *hit_t)
{
    if (!ctx) return FLX_ERR_ARG;
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) return FLX_OK;
    if (!rays || !hit_i || !hit_t) return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_closest_static: null argument");
    if (n > (size_t)FLX_STATIC_MAX_RAYS)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_trace_closest_static: " + std::to_string(n) + " rays exceed the static scratch's maximum of " +
                                              std::to_string(FLX_STATIC_MAX_RAYS) + " per call (FLX_STATIC_MAX_RAYS); split the batch or use flx_trace_closest");
    if (!ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    unsigned char *scratch = nullptr;
    { const int rc = scratchAddress(ctx, &scratch); if (rc != FLX_OK) return rc; }
    const StaticLayout l = layoutOf(scratch);
    // the array keeps the last call's contents: the cursor must start at zero, everything else is overwritten or unread
    FLX_CUDA(ctx, cudaMemsetAsync(scratch, 0, FLX_STATIC_HEADER_BYTES, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync((void *)l.rays, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream));
    k_trace_batch_static<<<ctx->numSMs * FLX_STATIC_CTAS_PER_SM, FLX_STATIC_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, scratch, (unsigned int)n);
    ctx->launches++;
    FLX_CUDA(ctx, cudaGetLastError());
    FLX_CUDA(ctx, cudaMemcpyAsync(hit_i, l.hitI, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync(hit_t, l.hitT, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FLX_OK;
}
