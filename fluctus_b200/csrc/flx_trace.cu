// Ray-traversal kernels (closest hit / any hit), both as wavefront stages (reference
// src/wf_extrays.cl, src/wf_shadowrays.cl) and as plain ray-batch entry points.
//
// Persistent warps: the grid is a fixed multiple of the SM count; lanes that run out of work pull the
// next queue entries with one atomic per warp (see traceLoop in flx_bvh.cuh) until the work is gone.
// COUNT=true instantiations additionally accumulate visit counters (inner pops, leaf visits,
// triangle tests, hits, rays) for the algorithmic-bytes figure; they are never the timed ones.
#include "flx_internal.h"

// Two CTAs of 640 threads per SM (40 warps at 48 registers) instead of one of 1024 (32 warps at 64): traversal time
// follows the number of resident warps (768 / 896 / 1024 threads: 126 / 115 / 108 ms per kitchen frame), and 1280
// threads x 48 registers is the largest population whose loop bodies stay (almost) free of spills — 2 x 672 and
// 2 x 704 drop to 40 registers and spill 200 bytes.  Each CTA stages its own copy of the top of the tree, so the
// copy is halved to 440 nodes (flx_bvh.cuh), the plateau of that parameter.  cfg5: 2 731 -> 2 965 Mrays/s, kitchen
// traversal 108.3 -> 106.8 ms per frame (tools/tune_trace.sh).
#ifndef FLX_TRACE_THREADS
#define FLX_TRACE_THREADS 640
#endif
#ifndef FLX_TRACE_MIN_BLOCKS
#define FLX_TRACE_MIN_BLOCKS 2
#endif

template <bool BITS> struct TrailOf { typedef StackTrail type; };
template <> struct TrailOf<true> { typedef BitTrail type; };

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

// ---- jobs: where the rays of each kernel come from and where the results go ------------------------------
// Path state is touched once per ray: with FLX_STREAM_STATE its loads and stores are marked evict-first (ld.global.cs /
// st.global.cs) so that they do not push the hierarchy out of L1, which serves 61 % of the traversal's sectors.
#ifndef FLX_STREAM_STATE
#define FLX_STREAM_STATE 1
#endif
#if FLX_STREAM_STATE
#define FLX_LD_STATE(ref) __ldcs(&(ref))
#define FLX_LD_ONCE(ptr) __ldcs(ptr)
#define FLX_ST_STATE(ref, v) __stcs(&(ref), (v))
#else
#define FLX_LD_STATE(ref) (ref)
#define FLX_LD_ONCE(ptr) __ldg(ptr)
#define FLX_ST_STATE(ref, v) ((ref) = (v))
#endif

// wf_extrays.cl:5-36
struct ExtJob
{
    static constexpr bool BATCH = false;
    const DevBVH &bvh; const PathState &st; const uint32_t *queue; const flx_render_params &params;
    uint32_t gid; unsigned int nHits, nRays;
    uint32_t len; // pathLen, which travels in the .w lane of the direction: kept so that the store needs no read
    FLX_DEV bool load(unsigned int k, f3 &o, f3 &d, float &tmax)
    {
        gid = FLX_LD_ONCE(queue + k);
        const float4 r1 = FLX_LD_STATE(st.ray1[gid]);
        o = mk3(FLX_LD_STATE(st.ray0[gid])); d = mk3(r1); len = __float_as_uint(r1.w); tmax = FLX_FLT_MAX;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3 o, f3 d)
    {
        f3 P = mk3(0.0f), N = mk3(0.0f);
        float uvx = 0.0f, uvy = 0.0f, t = FLX_FLT_MAX;
        int hi = -1, matId = -1, al = 0;
        if (h.slot >= 0)
        {
            t = h.t;
            fetchHitAttribs(bvh, h, o, d, hi, matId, P, N, uvx, uvy);
        }
        nRays++; nHits += (h.slot >= 0);
        if (params.sampleImpl && params.useAreaLight) // wf_extrays.cl:27-28
        {
            const AreaLightGeom g = makeLightGeom(params.areaLight);
            if (intersectLight(g, o, d, t)) { al = 1; P = o + t * d; N = g.N; hi = 0; matId = 0; }
        }
        FLX_ST_STATE(st.hit0[gid], mk4(P, t));
        FLX_ST_STATE(st.hit1[gid], mk4i(N, hi));
        FLX_ST_STATE(st.hit2[gid], make_float4(uvx, uvy, __int_as_float(matId), __int_as_float(al)));
        FLX_ST_STATE(st.ray1[gid].w, __uint_as_float(len + 1u)); // pathLen += 1
    }
};

// wf_shadowrays.cl:6-39
struct ShadowJob
{
    static constexpr bool BATCH = false;
    const PathState &st; const uint32_t *queue; const flx_render_params &params;
    uint32_t gid; unsigned int nHits, nRays;
    FLX_DEV bool load(unsigned int k, f3 &o, f3 &d, float &tmax)
    {
        gid = FLX_LD_ONCE(queue + k);
        nRays++;
        const float4 s0 = FLX_LD_STATE(st.shw0[gid]);
        o = mk3(s0); d = mk3(FLX_LD_STATE(st.shw1[gid])); tmax = s0.w;
        if (params.useAreaLight) // the light quad occludes from its front side (wf_shadowrays.cl:31-33)
        {
            float t = tmax;
            if (intersectLight(makeLightGeom(params.areaLight), o, d, t)) { st.shadowBlocked[gid] = 1u; return false; }
        }
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3, f3)
    {
        nHits += (h.slot >= 0);
        st.shadowBlocked[gid] = h.slot >= 0 ? 1u : 0u;
    }
};

// Microkernel renderer (src/mk_next_vertex.cl:33-41): the closest-hit part of `nextVertex` for every task in phase
// MK_RT_NEXT_VERTEX.  No queue: task k is pixel k, tasks in other phases are skipped at load time.  The result
// handling is ExtJob's (hit record + pathLen); the shading half of the stage runs in k_mk_vertex_shade.
struct MkVertexJob
{
    static constexpr bool BATCH = false;
    ExtJob ext;
    FLX_DEV bool load(unsigned int k, f3 &o, f3 &d, float &tmax)
    {
        if (ext.st.phase[k] != FLX_MK_RT_NEXT_VERTEX) return false;
        ext.gid = k;
        const float4 r1 = ext.st.ray1[k];
        o = mk3(ext.st.ray0[k]); d = mk3(r1); ext.len = __float_as_uint(r1.w); tmax = FLX_FLT_MAX;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3 o, f3 d) { ext.store(h, o, d); }
};

// The inline shadow rays of src/mk_sample_bsdf.cl:90-96 (environment sample: the light quad occludes too, LIGHT =
// true) and :121-126 (area-light sample: hierarchy only).  A task takes part when its NEE slot says "pending".
template <bool LIGHT>
struct MkShadowJob
{
    static constexpr bool BATCH = false;
    const PathState &st; const flx_render_params &params;
    uint32_t gid;
    FLX_DEV bool load(unsigned int k, f3 &o, f3 &d, float &tmax)
    {
        if (st.shadowBlocked[k] != FLX_MK_NEE_PENDING) return false;
        gid = k;
        const float4 s0 = st.shw0[k];
        o = mk3(s0); d = mk3(st.shw1[k]); tmax = s0.w;
        if (LIGHT && params.useAreaLight)
        {
            float t = tmax;
            if (intersectLight(makeLightGeom(params.areaLight), o, d, t)) { st.shadowBlocked[k] = 1u; return false; }
        }
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3, f3) { st.shadowBlocked[gid] = h.slot >= 0 ? 1u : 0u; }
};

// kernel_pick.cl:6-41: the ray through NDC point (x, y) from the camera position (no lens sample), closest hit,
// the light quad as an implicit hit when sampleImpl && useAreaLight; result as a Hit record (geom.h:152-163)
struct PickJob
{
    static constexpr bool BATCH = false;
    const DevBVH &bvh; const flx_render_params &params; float ndcX, ndcY; float4 *out;
    FLX_DEV bool load(unsigned int, f3 &o, f3 &d, float &tmax)
    {
        float SCRx = ndcX + ndcX - 1.0f;
        float SCRy = ndcY + ndcY - 1.0f;
        SCRx *= (float)params.width * params.height1;
        SCRx *= params.camera.fovSCALE;
        SCRy *= params.camera.fovSCALE;
        const f3 pos = mk3(params.camera.pos.x, params.camera.pos.y, params.camera.pos.z);
        const f3 right = mk3(params.camera.right.x, params.camera.right.y, params.camera.right.z);
        const f3 up = mk3(params.camera.up.x, params.camera.up.y, params.camera.up.z);
        const f3 dir = mk3(params.camera.dir.x, params.camera.dir.y, params.camera.dir.z);
        const f3 rayTarget = pos + right * SCRx + up * SCRy + dir;
        o = pos; d = normalize(rayTarget - pos); tmax = FLX_FLT_MAX;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3 o, f3 d)
    {
        f3 P = mk3(0.0f), N = mk3(0.0f);
        float uvx = 0.0f, uvy = 0.0f, t = FLX_FLT_MAX;
        int hi = -1, matId = -1, al = 0;
        if (h.slot >= 0) { t = h.t; fetchHitAttribs(bvh, h, o, d, hi, matId, P, N, uvx, uvy); }
        if (params.sampleImpl && params.useAreaLight)
        {
            const AreaLightGeom g = makeLightGeom(params.areaLight);
            if (intersectLight(g, o, d, t)) { al = 1; P = o + t * d; N = g.N; hi = 0; matId = 0; }
        }
        out[0] = mk4(P, 0.0f);
        out[1] = mk4(N, 0.0f);
        out[2] = make_float4(uvx, uvy, t, __int_as_float(hi));
        out[3] = make_float4(__int_as_float(al), __int_as_float(matId), 0.0f, 0.0f);
    }
};

// ray batch, closest hit: rays as 2 x float4 {o.xyz, tmax} {d.xyz, -}
struct BatchJob
{
    static constexpr bool BATCH = true;
    const DevBVH &bvh; const float4 *rays; int32_t *hitI; float *hitT; float4 *hits;
    unsigned int k; unsigned int nHits, nRays;
    FLX_DEV bool load(unsigned int kk, f3 &o, f3 &d, float &tmax)
    {
        k = kk;
        const float4 r0 = FLX_LD_ONCE(rays + 2 * (size_t)kk), r1 = FLX_LD_ONCE(rays + 2 * (size_t)kk + 1);
        o = mk3(r0); d = mk3(r1); tmax = r0.w;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3 o, f3 d)
    {
        nRays++; nHits += (h.slot >= 0);
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
};

struct AnyBatchJob
{
    static constexpr bool BATCH = true;
    const float4 *rays; uint8_t *occ;
    unsigned int k; unsigned int nHits, nRays;
    FLX_DEV bool load(unsigned int kk, f3 &o, f3 &d, float &tmax)
    {
        k = kk;
        const float4 r0 = FLX_LD_ONCE(rays + 2 * (size_t)kk), r1 = FLX_LD_ONCE(rays + 2 * (size_t)kk + 1);
        o = mk3(r0); d = mk3(r1); tmax = r0.w;
        return true;
    }
    FLX_DEV void store(const RayHit &h, f3, f3)
    {
        nRays++; nHits += (h.slot >= 0);
        occ[k] = h.slot >= 0 ? 1 : 0;
    }
};

// BITS selects the trail (flx_bvh.cuh): false = the reference's stack, true = its -DUSE_BITSTACK traversal
template <bool COUNT, bool BITS>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_ext(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
      unsigned int *cursor, unsigned long long *stats, const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    const unsigned int total = counters->extensionQueue;
    if (total == 0u) return;
    stageTopNodes(bvh, topNodes);
    ExtJob job = {bvh, st, queue, params, 0u, 0u, 0u, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<false, COUNT, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursor, total, cnt);
    if (COUNT) flushCounts(stats, cnt, job.nHits, job.nRays);
}

template <bool COUNT, bool BITS>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_shadow(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
         unsigned int *cursor, unsigned long long *stats, const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    const unsigned int total = counters->shadowQueue;
    if (total == 0u) return;
    stageTopNodes(bvh, topNodes);
    ShadowJob job = {st, queue, params, 0u, 0u, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<true, COUNT, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursor, total, cnt);
    if (COUNT) flushCounts(stats, cnt, job.nHits, job.nRays);
}

// Extension and shadow rays of one iteration in ONE persistent launch (render entry).  The two stages are
// independent (extension rays read ray0/ray1 and write the hit record, shadow rays read shw0/shw1 and write
// shadowBlocked), so a warp that finds the extension pool empty moves straight on to the shadow pool instead of
// idling until the slowest extension ray of the grid has finished; the top of the tree is staged once.
template <bool BITS>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_ext_shadow(DevBVH bvh, PathState st, const uint32_t *__restrict__ extQueue, const uint32_t *__restrict__ shadowQueue,
             flx_queue_counters *counters, unsigned int *cursors, RenderTotals *totals, uint32_t *pixelIdx, uint32_t ownedPixels,
             const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    const unsigned int nExt = counters->extensionQueue, nShadow = counters->shadowQueue;
    if (nExt != 0u || nShadow != 0u)
    {
        stageTopNodes(bvh, topNodes);
        TravCount cnt = {0u, 0u, 0u};
        if (nExt)
        {
            ExtJob job = {bvh, st, extQueue, params, 0u, 0u, 0u, 0u};
            traceLoop<false, false, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursors, nExt, cnt);
        }
        if (nShadow)
        {
            ShadowJob job = {st, shadowQueue, params, 0u, 0u, 0u};
            traceLoop<true, false, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursors + 1, nShadow, cnt);
        }
    }
    // The CTA that finishes last closes the iteration: what the reference's host does between iterations
    // (statistics src/tracer.cpp:761-765, updatePixelIndex + enqueueClearWfQueues src/clcontext.cpp:903-921) and
    // the reset of the two work cursors.  Every CTA has read the counters before it arrives here.
    __syncthreads();
    if (threadIdx.x == 0)
    {
        __threadfence();
        if (atomicAdd(cursors + 2, 1u) == gridDim.x - 1u)
        {
            closeIteration(counters, totals, pixelIdx, ownedPixels);
            cursors[0] = 0u; cursors[1] = 0u; cursors[2] = 0u;
        }
    }
}

__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_mk_vertex(DevBVH bvh, PathState st, unsigned int total, unsigned int *cursor, const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    MkVertexJob job = {{bvh, st, nullptr, params, 0u, 0u, 0u, 0u}};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<false, false, StackTrail>(bvh, topNodes, job, cursor, total, cnt);
}

template <bool LIGHT>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_mk_shadow(DevBVH bvh, PathState st, unsigned int total, unsigned int *cursor, const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    MkShadowJob<LIGHT> job = {st, params, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<true, false, StackTrail>(bvh, topNodes, job, cursor, total, cnt);
}

__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_pick(DevBVH bvh, float ndcX, float ndcY, float4 *out, unsigned int *cursor, const __grid_constant__ flx_render_params params)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    PickJob job = {bvh, params, ndcX, ndcY, out};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<false, false, StackTrail>(bvh, topNodes, job, cursor, 1u, cnt);
}

template <bool COUNT, bool BITS>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_trace_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor, unsigned long long *stats,
              int32_t *__restrict__ hitI, float *__restrict__ hitT, float4 *__restrict__ hits)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    BatchJob job = {bvh, rays, hitI, hitT, hits, 0u, 0u, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<false, COUNT, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursor, n, cnt);
    if (COUNT) flushCounts(stats, cnt, job.nHits, job.nRays);
}

template <bool COUNT, bool BITS>
__global__ void __launch_bounds__(FLX_TRACE_THREADS, FLX_TRACE_MIN_BLOCKS)
k_any_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor, unsigned long long *stats,
            uint8_t *__restrict__ occ)
{
    extern __shared__ float4 topNodes[];
    stageTopNodes(bvh, topNodes);
    AnyBatchJob job = {rays, occ, 0u, 0u, 0u};
    TravCount cnt = {0u, 0u, 0u};
    traceLoop<true, COUNT, typename TrailOf<BITS>::type>(bvh, topNodes, job, cursor, n, cnt);
    if (COUNT) flushCounts(stats, cnt, job.nHits, job.nRays);
}

// ---------------------------------------------------------------------------------------------
// Persistent grid: FLX_TRACE_MIN_BLOCKS CTAs per SM, each with its own copy of the top of the tree (see the top of this
// file).  Many small CTAs do not pay: eight 28 KB copies evict the L1 that the deeper nodes and the triangles live in
// (-8 % on the kitchen render, -17 % on incoherent rays; tools/tune_vote.sh).
#ifndef FLX_TRACE_CTAS_PER_SM
#define FLX_TRACE_CTAS_PER_SM ((1024 / FLX_TRACE_THREADS) > FLX_TRACE_MIN_BLOCKS ? (1024 / FLX_TRACE_THREADS) : FLX_TRACE_MIN_BLOCKS)
#endif
static int traceGrid(const flx_ctx *ctx) { return ctx->numSMs * FLX_TRACE_CTAS_PER_SM; }

// dynamic shared memory above 48 KB has to be opted into, per kernel and per device
static cudaError_t traceKernelSetup(flx_ctx *ctx)
{
    if (ctx->traceSetupDone) return cudaSuccess;
    const void *kernels[] = {(const void *)k_ext<false, false>, (const void *)k_ext<true, false>, (const void *)k_ext<false, true>,
                             (const void *)k_shadow<false, false>, (const void *)k_shadow<true, false>, (const void *)k_shadow<false, true>,
                             (const void *)k_trace_batch<false, false>, (const void *)k_trace_batch<true, false>, (const void *)k_trace_batch<false, true>,
                             (const void *)k_any_batch<false, false>, (const void *)k_any_batch<true, false>, (const void *)k_any_batch<false, true>,
                             (const void *)k_ext_shadow<false>, (const void *)k_ext_shadow<true>,
                             (const void *)k_pick, (const void *)k_mk_vertex, (const void *)k_mk_shadow<false>, (const void *)k_mk_shadow<true>};
    for (const void *k : kernels)
    {
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, FLX_TOP_SMEM_BYTES);
        if (e != cudaSuccess) return e;
    }
    ctx->traceSetupDone = true;
    return cudaSuccess;
}

// cursors live right behind the pixel cursor: [1] ext, [2] shadow / batch, [3] CTA arrival counter of k_ext_shadow
static unsigned int *workCursor(flx_ctx *ctx, int which) { return (unsigned int *)(ctx->dPixelIdx + 1 + which); }

#define FLX_LAUNCHED(ctx) do { (ctx)->launches++; FLX_CUDA(ctx, cudaGetLastError()); } while (0)

int flx_launch_ext(flx_ctx *ctx)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 0);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    // the counting variants (visit statistics under the reference's default order) always walk with the stack
    if (ctx->counting)
        k_ext<true, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, ctx->dTraceStats, ctx->params);
    else if (ctx->useBitstack)
        k_ext<false, true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, nullptr, ctx->params);
    else
        k_ext<false, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, nullptr, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

// zero the work cursors and the arrival counter of k_ext_shadow (which leaves them zero again when it ends)
int flx_reset_trace_cursors(flx_ctx *ctx)
{
    FLX_CUDA(ctx, cudaMemsetAsync(workCursor(ctx, 0), 0, 3 * sizeof(unsigned int), ctx->stream));
    return FLX_OK;
}

int flx_launch_ext_shadow(flx_ctx *ctx)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    if (ctx->useBitstack)
        k_ext_shadow<true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.q[FLX_Q_SHADOW],
            ctx->qs.counters, workCursor(ctx, 0), ctx->dTotals, ctx->dPixelIdx, flx_owned_pixels(ctx), ctx->params);
    else
        k_ext_shadow<false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.q[FLX_Q_SHADOW],
            ctx->qs.counters, workCursor(ctx, 0), ctx->dTotals, ctx->dPixelIdx, flx_owned_pixels(ctx), ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

// microkernel renderer: traversal halves of nextVertex / sampleBsdf over tasks [0, total)
int flx_launch_mk_vertex_trace(flx_ctx *ctx, uint32_t total)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 0);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_mk_vertex<<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, total, cur, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_mk_shadow_trace(flx_ctx *ctx, uint32_t total, int lightOccludes)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (lightOccludes) k_mk_shadow<true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, total, cur, ctx->params);
    else k_mk_shadow<false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, total, cur, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_shadow(flx_ctx *ctx)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_shadow<true, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, ctx->dTraceStats + 5, ctx->params);
    else if (ctx->useBitstack)
        k_shadow<false, true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, nullptr, ctx->params);
    else
        k_shadow<false, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, nullptr, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_trace_batch(flx_ctx *ctx, const float4 *dRays, size_t n, int32_t *dHitI, float *dHitT, flx_hit *dHits)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_trace_batch<true, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, ctx->dTraceStats + 10, dHitI, dHitT, (float4 *)dHits);
    else if (ctx->useBitstack)
        k_trace_batch<false, true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dHitI, dHitT, (float4 *)dHits);
    else
        k_trace_batch<false, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dHitI, dHitT, (float4 *)dHits);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

// CLContext::pickSingle (clcontext.cpp:934-949): one ray, one warp
int flx_launch_pick(flx_ctx *ctx, float ndcX, float ndcY, flx_hit *dHit)
{
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_pick<<<1, 32, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, ndcX, ndcY, (float4 *)dHit, cur, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_any_batch(flx_ctx *ctx, const float4 *dRays, size_t n, uint8_t *dOcc)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    FLX_CUDA(ctx, traceKernelSetup(ctx));
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    if (ctx->counting)
        k_any_batch<true, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, ctx->dTraceStats + 15, dOcc);
    else if (ctx->useBitstack)
        k_any_batch<false, true><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dOcc);
    else
        k_any_batch<false, false><<<traceGrid(ctx), FLX_TRACE_THREADS, FLX_TOP_SMEM_BYTES, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, nullptr, dOcc);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}
