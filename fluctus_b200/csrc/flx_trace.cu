// Ray-traversal kernels (closest hit / any hit), both as wavefront stages (reference
// src/wf_extrays.cl, src/wf_shadowrays.cl) and as plain ray-batch entry points.
//
// Persistent warps: the grid is a fixed multiple of the SM count and every warp pulls the next 32
// queue entries with one atomic (lane 0) + shuffle broadcast until the work is gone.
// COUNT=true instantiations additionally accumulate visit counters (inner pops, leaf visits,
// triangle tests, hits, rays) for the algorithmic-bytes figure; they are never the timed ones.
#include "flx_internal.h"

#define FLX_TRACE_THREADS 128

FLX_DEV bool fetchWork(unsigned int *cursor, unsigned int total, unsigned int &idx)
{
    unsigned int base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    idx = base + (threadIdx.x & 31);
    return base < total;
}

FLX_DEV void flushCounts(unsigned long long *stats, const TravCount &c, unsigned int hits, unsigned int rays)
{
    unsigned int v[5] = {c.inner, c.leaf, c.tri, hits, rays};
#pragma unroll
    for (int k = 0; k < 5; k++)
    {
        unsigned int x = v[k];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if ((threadIdx.x & 31) == 0 && x) atomicAdd(stats + k, (unsigned long long)x);
    }
}

FLX_DEV AreaLightGeom makeLightGeom(const flx_area_light &al)
{
    const f3 pos = mk3(al.pos.x, al.pos.y, al.pos.z), right = mk3(al.right.x, al.right.y, al.right.z), up = mk3(al.up.x, al.up.y, al.up.z);
    const float sx = al.size.x, sy = al.size.y;
    AreaLightGeom g;
    g.tl = pos + sx * right + sy * up;
    g.tr = pos - sx * right + sy * up;
    g.bl = pos + sx * right - sy * up;
    g.br = pos - sx * right - sy * up;
    g.N = mk3(al.N.x, al.N.y, al.N.z);
    return g;
}

// wf_extrays.cl:5-36
template <bool COUNT>
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_ext(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
      unsigned int *cursor, unsigned long long *stats, const __grid_constant__ flx_render_params params)
{
    const unsigned int total = counters->extensionQueue;
    const bool lightOn = params.sampleImpl && params.useAreaLight;
    TravCount cnt = {0u, 0u, 0u};
    unsigned int nHits = 0, nRays = 0;
    unsigned int k;
    while (fetchWork(cursor, total, k))
    {
        if (k >= total) continue;
        const uint32_t gid = queue[k];
        const float4 r0 = st.ray0[gid], r1 = st.ray1[gid];
        const f3 o = mk3(r0), d = mk3(r1);
        const RayHit h = traverse<false, COUNT>(bvh, o, d, FLX_FLT_MAX, &cnt);

        f3 P = mk3(0.0f), N = mk3(0.0f);
        float uvx = 0.0f, uvy = 0.0f, t = FLX_FLT_MAX;
        int hi = -1, matId = -1, al = 0;
        if (h.slot >= 0)
        {
            t = h.t;
            fetchHitAttribs(bvh, h, o, d, hi, matId, P, N, uvx, uvy);
        }
        if (COUNT) { nRays++; nHits += (h.slot >= 0); }
        if (lightOn)
        {
            const AreaLightGeom g = makeLightGeom(params.areaLight);
            if (intersectLight(g, o, d, t)) { al = 1; P = o + t * d; N = g.N; hi = 0; matId = 0; }
        }
        st.hit0[gid] = mk4(P, t);
        st.hit1[gid] = mk4i(N, hi);
        st.hit2[gid] = make_float4(uvx, uvy, __int_as_float(matId), __int_as_float(al));
        st.ray1[gid].w = __uint_as_float(__float_as_uint(r1.w) + 1u); // pathLen += 1
    }
    if (COUNT) flushCounts(stats, cnt, nHits, nRays);
}

// wf_shadowrays.cl:6-39
template <bool COUNT>
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_shadow(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
         unsigned int *cursor, unsigned long long *stats, const __grid_constant__ flx_render_params params)
{
    const unsigned int total = counters->shadowQueue;
    TravCount cnt = {0u, 0u, 0u};
    unsigned int nHits = 0, nRays = 0;
    unsigned int k;
    while (fetchWork(cursor, total, k))
    {
        if (k >= total) continue;
        const uint32_t gid = queue[k];
        const float4 s0 = st.shw0[gid], s1 = st.shw1[gid];
        const f3 o = mk3(s0), d = mk3(s1);
        const float lenL = s0.w;
        bool occluded = false;
        if (params.useAreaLight)
        {
            float t = lenL;
            occluded = intersectLight(makeLightGeom(params.areaLight), o, d, t);
        }
        if (!occluded)
        {
            occluded = traverse<true, COUNT>(bvh, o, d, lenL, &cnt).slot >= 0;
            if (COUNT) { nRays++; nHits += occluded; }
        }
        st.shadowBlocked[gid] = occluded ? 1u : 0u;
    }
    if (COUNT) flushCounts(stats, cnt, nHits, nRays);
}

// ray-batch closest hit: rays as 2 x float4 {o.xyz, tmax} {d.xyz, -}
template <bool COUNT>
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_trace_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor, unsigned long long *stats,
              int32_t *__restrict__ hitI, float *__restrict__ hitT, float4 *__restrict__ hits)
{
    TravCount cnt = {0u, 0u, 0u};
    unsigned int nHits = 0, nRays = 0;
    unsigned int k;
    while (fetchWork(cursor, n, k))
    {
        if (k >= n) continue;
        const float4 r0 = __ldg(rays + 2 * (size_t)k), r1 = __ldg(rays + 2 * (size_t)k + 1);
        const f3 o = mk3(r0), d = mk3(r1);
        const RayHit h = traverse<false, COUNT>(bvh, o, d, r0.w, &cnt);
        if (COUNT) { nRays++; nHits += (h.slot >= 0); }
        int hi = -1, matId = -1;
        f3 P = mk3(0.0f), N = mk3(0.0f);
        float uvx = 0.0f, uvy = 0.0f;
        if (h.slot >= 0)
        {
            if (hits) fetchHitAttribs(bvh, h, o, d, hi, matId, P, N, uvx, uvy);
            else hi = __float_as_int(__ldg(bvh.slots + 3 * (size_t)h.slot).w);
        }
        if (hitI) hitI[k] = hi;
        if (hitT) hitT[k] = h.t;
        if (hits)
        {
            float4 *o4 = hits + 4 * (size_t)k; // flx_hit, 64 B
            o4[0] = mk4(P, 0.0f);
            o4[1] = mk4(N, 0.0f);
            o4[2] = make_float4(uvx, uvy, h.t, __int_as_float(hi));
            o4[3] = make_float4(__int_as_float(0), __int_as_float(matId), 0.0f, 0.0f);
        }
    }
    if (COUNT) flushCounts(stats, cnt, nHits, nRays);
}

template <bool COUNT>
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_any_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor, unsigned long long *stats,
            uint8_t *__restrict__ occ)
{
    TravCount cnt = {0u, 0u, 0u};
    unsigned int nHits = 0, nRays = 0;
    unsigned int k;
    while (fetchWork(cursor, n, k))
    {
        if (k >= n) continue;
        const float4 r0 = __ldg(rays + 2 * (size_t)k), r1 = __ldg(rays + 2 * (size_t)k + 1);
        const bool hit = traverse<true, COUNT>(bvh, mk3(r0), mk3(r1), r0.w, &cnt).slot >= 0;
        if (COUNT) { nRays++; nHits += hit; }
        occ[k] = hit ? 1 : 0;
    }
    if (COUNT) flushCounts(stats, cnt, nHits, nRays);
}

// ---------------------------------------------------------------------------------------------
static int traceGrid(const flx_ctx *ctx) { return ctx->numSMs * 8; }

// cursors live right behind the pixel cursor: [1] ext, [2] shadow / batch
static unsigned int *workCursor(flx_ctx *ctx, int which) { return (unsigned int *)(ctx->dPixelIdx + 1 + which); }

#define FLX_LAUNCHED(ctx) do { (ctx)->launches++; FLX_CUDA(ctx, cudaGetLastError()); } while (0)

int flx_launch_ext(flx_ctx *ctx)
{
    unsigned int *cur = workCursor(ctx, 0);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_ext<true><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, ctx->dTraceStats, ctx->params);
    else
        k_ext<false><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, nullptr, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_shadow(flx_ctx *ctx)
{
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_shadow<true><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, ctx->dTraceStats + 5, ctx->params);
    else
        k_shadow<false><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, nullptr, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_trace_batch(flx_ctx *ctx, const float4 *dRays, size_t n, int32_t *dHitI, float *dHitT, flx_hit *dHits)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_trace_batch<true><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, ctx->dTraceStats + 10, dHitI, dHitT, (float4 *)dHits);
    else
        k_trace_batch<false><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dHitI, dHitT, (float4 *)dHits);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_any_batch(flx_ctx *ctx, const float4 *dRays, size_t n, uint8_t *dOcc)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_any_batch<true><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, ctx->dTraceStats + 15, dOcc);
    else
        k_any_batch<false><<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dOcc);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}
