// Scene re-packing and the ray-traversal kernels (closest hit / any hit), both as wavefront
// stages (reference src/wf_extrays.cl, src/wf_shadowrays.cl) and as plain ray-batch entry points.
#include "flx_internal.h"
#include <cstring>

// ---------------------------------------------------------------------------------------------
// Re-pack the reference's Node[48 B] / RTTriangle[160 B] / indices into the streams described in
// flx_bvh.cuh.  Runs once per scene on the host (O(n)), then three bulk uploads.
// ---------------------------------------------------------------------------------------------
int flx_repack_scene(flx_ctx *ctx, const flx_triangle *tris, size_t n_tris, const uint32_t *indices, size_t n_indices,
                     const flx_node *nodes, size_t n_nodes)
{
    if (!tris || !indices || !nodes || n_tris == 0 || n_nodes == 0 || n_indices == 0)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: empty geometry");
    if (n_indices >= 0x7fffffffu || n_nodes >= 0x7fffffffu)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: scene too large for 31-bit references");

    // compact numbering of inner nodes, in DFS order like the input
    std::vector<int32_t> innerIndex(n_nodes, -1);
    size_t nInner = 0;
    for (size_t i = 0; i < n_nodes; i++)
        if (nodes[i].nPrims == 0) innerIndex[i] = (int32_t)nInner++;

    auto refOf = [&](size_t child) -> int32_t {
        const flx_node &c = nodes[child];
        return c.nPrims != 0 ? ~(int32_t)c.iStartOrRight : innerIndex[child];
    };

    std::vector<float4> hNodes(4 * (nInner ? nInner : 1));
    std::vector<float4> hSlots(3 * n_indices);
    std::vector<uint8_t> covered(n_indices, 0);

    for (size_t i = 0; i < n_nodes; i++)
    {
        const flx_node &n = nodes[i];
        if (n.nPrims == 0)
        {
            const size_t li = i + 1, ri = n.iStartOrRight;
            if (li >= n_nodes || ri >= n_nodes) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: child index out of range");
            const flx_node &L = nodes[li], &R = nodes[ri];
            float4 *o = &hNodes[4 * (size_t)innerIndex[i]];
            o[0] = make_float4(L.bmin.x, L.bmin.y, L.bmin.z, L.bmax.x);
            o[1] = make_float4(L.bmax.y, L.bmax.z, R.bmin.x, R.bmin.y);
            o[2] = make_float4(R.bmin.z, R.bmax.x, R.bmax.y, R.bmax.z);
            int32_t refs[4] = {refOf(li), refOf(ri), 0, 0};
            memcpy(&o[3], refs, 16);
        }
        else
        {
            const size_t s = n.iStartOrRight, e = s + n.nPrims;
            if (e > n_indices) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: leaf range outside index list");
            for (size_t k = s; k < e; k++)
            {
                if (covered[k]) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: overlapping leaf ranges");
                covered[k] = 1;
                const uint32_t ti = indices[k];
                if (ti >= n_tris) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: triangle index out of range");
                const flx_triangle &t = tris[ti];
                const int32_t last = (k + 1 == e) ? 1 : 0;
                float4 *o = &hSlots[3 * k];
                o[0] = make_float4(t.v0.p.x, t.v0.p.y, t.v0.p.z, 0.f); memcpy(&o[0].w, &ti, 4);
                o[1] = make_float4(t.v1.p.x, t.v1.p.y, t.v1.p.z, 0.f); memcpy(&o[1].w, &last, 4);
                o[2] = make_float4(t.v2.p.x, t.v2.p.y, t.v2.p.z, 0.f); memcpy(&o[2].w, &t.matId, 4);
            }
        }
    }
    // uncovered slots (never referenced) still get a terminator so a stray read cannot run away
    for (size_t k = 0; k < n_indices; k++)
        if (!covered[k]) { const int32_t one = 1; hSlots[3 * k] = hSlots[3 * k + 1] = hSlots[3 * k + 2] = make_float4(0, 0, 0, 0); memcpy(&hSlots[3 * k + 1].w, &one, 4); }

    std::vector<float4> hShade(4 * n_tris);
    for (size_t i = 0; i < n_tris; i++)
    {
        const flx_triangle &t = tris[i];
        float4 *o = &hShade[4 * i];
        o[0] = make_float4(t.v0.n.x, t.v0.n.y, t.v0.n.z, t.v1.n.x);
        o[1] = make_float4(t.v1.n.y, t.v1.n.z, t.v2.n.x, t.v2.n.y);
        o[2] = make_float4(t.v2.n.z, t.v0.t.x, t.v0.t.y, t.v1.t.x);
        o[3] = make_float4(t.v1.t.y, t.v2.t.x, t.v2.t.y, 0.f); memcpy(&o[3].w, &t.matId, 4);
    }

    cudaFree(ctx->dNodes); cudaFree(ctx->dSlots); cudaFree(ctx->dShade);
    ctx->dNodes = ctx->dSlots = ctx->dShade = nullptr;
    FLX_CUDA(ctx, cudaMalloc(&ctx->dNodes, hNodes.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMalloc(&ctx->dSlots, hSlots.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMalloc(&ctx->dShade, hShade.size() * sizeof(float4)));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dNodes, hNodes.data(), hNodes.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dSlots, hSlots.data(), hSlots.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaMemcpyAsync(ctx->dShade, hShade.data(), hShade.size() * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // host vectors die at return
    ctx->bvh.nodes = ctx->dNodes; ctx->bvh.slots = ctx->dSlots; ctx->bvh.shade = ctx->dShade;
    ctx->bvh.rootRef = refOf(0);
    ctx->nInner = nInner; ctx->nSlots = n_indices; ctx->nTris = n_tris;
    return FLX_OK;
}

// ---------------------------------------------------------------------------------------------
// Kernels.  Persistent warps: the grid is a fixed multiple of the SM count and every warp pulls
// the next 32 queue entries with one atomic (lane 0) + shuffle broadcast until the work is gone.
// ---------------------------------------------------------------------------------------------
#define FLX_TRACE_THREADS 128

FLX_DEV bool fetchWork(unsigned int *cursor, unsigned int total, unsigned int &idx)
{
    unsigned int base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    idx = base + (threadIdx.x & 31);
    return base < total;
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
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_ext(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
      unsigned int *cursor, const __grid_constant__ flx_render_params params)
{
    const unsigned int total = counters->extensionQueue;
    const bool lightOn = params.sampleImpl && params.useAreaLight;
    unsigned int k;
    while (fetchWork(cursor, total, k))
    {
        if (k >= total) continue;
        const uint32_t gid = queue[k];
        const float4 r0 = st.ray0[gid], r1 = st.ray1[gid];
        const f3 o = mk3(r0), d = mk3(r1);
        const RayHit h = traverse<false>(bvh, o, d, FLX_FLT_MAX);

        f3 P = mk3(0.0f), N = mk3(0.0f);
        float uvx = 0.0f, uvy = 0.0f, t = FLX_FLT_MAX;
        int hi = -1, matId = -1, al = 0;
        if (h.slot >= 0)
        {
            t = h.t;
            fetchHitAttribs(bvh, h, o, d, hi, matId, P, N, uvx, uvy);
        }
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
}

// wf_shadowrays.cl:6-39
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_shadow(DevBVH bvh, PathState st, const uint32_t *__restrict__ queue, const flx_queue_counters *__restrict__ counters,
         unsigned int *cursor, const __grid_constant__ flx_render_params params)
{
    const unsigned int total = counters->shadowQueue;
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
        if (!occluded) occluded = traverse<true>(bvh, o, d, lenL).slot >= 0;
        st.shadowBlocked[gid] = occluded ? 1u : 0u;
    }
}

// ray-batch closest hit: rays as 2 x float4 {o.xyz, tmax} {d.xyz, -}
__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_trace_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor,
              int32_t *__restrict__ hitI, float *__restrict__ hitT, float4 *__restrict__ hits)
{
    unsigned int k;
    while (fetchWork(cursor, n, k))
    {
        if (k >= n) continue;
        const float4 r0 = __ldg(rays + 2 * (size_t)k), r1 = __ldg(rays + 2 * (size_t)k + 1);
        const f3 o = mk3(r0), d = mk3(r1);
        const RayHit h = traverse<false>(bvh, o, d, r0.w);
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
}

__global__ void __launch_bounds__(FLX_TRACE_THREADS)
k_any_batch(DevBVH bvh, const float4 *__restrict__ rays, unsigned int n, unsigned int *cursor, uint8_t *__restrict__ occ)
{
    unsigned int k;
    while (fetchWork(cursor, n, k))
    {
        if (k >= n) continue;
        const float4 r0 = __ldg(rays + 2 * (size_t)k), r1 = __ldg(rays + 2 * (size_t)k + 1);
        occ[k] = traverse<true>(bvh, mk3(r0), mk3(r1), r0.w).slot >= 0 ? 1 : 0;
    }
}

// ---------------------------------------------------------------------------------------------
static int traceGrid(const flx_ctx *ctx) { return ctx->numSMs * 8; }

static unsigned int *workCursor(flx_ctx *ctx, int which)
{
    // two cursors (ext, shadow/batch) live right behind the pixel cursor allocation
    return (unsigned int *)(ctx->dPixelIdx + 1 + which);
}

int flx_launch_ext(flx_ctx *ctx)
{
    unsigned int *cur = workCursor(ctx, 0);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_ext<<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_EXTENSION], ctx->qs.counters, cur, ctx->params);
    ctx->launches++;
    FLX_CUDA(ctx, cudaGetLastError());
    return FLX_OK;
}

int flx_launch_shadow(flx_ctx *ctx)
{
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_shadow<<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, ctx->st, ctx->qs.q[FLX_Q_SHADOW], ctx->qs.counters, cur, ctx->params);
    ctx->launches++;
    FLX_CUDA(ctx, cudaGetLastError());
    return FLX_OK;
}

int flx_launch_trace_batch(flx_ctx *ctx, const float4 *dRays, size_t n, int32_t *dHitI, float *dHitT, flx_hit *dHits)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_trace_batch<<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, dHitI, dHitT, (float4 *)dHits);
    ctx->launches++;
    FLX_CUDA(ctx, cudaGetLastError());
    return FLX_OK;
}

int flx_launch_any_batch(flx_ctx *ctx, const float4 *dRays, size_t n, uint8_t *dOcc)
{
    if (n == 0) return FLX_OK;
    if (n > 0xfffffff0u) return flx_fail(ctx, FLX_ERR_ARG, "ray batch too large");
    unsigned int *cur = workCursor(ctx, 1);
    FLX_CUDA(ctx, cudaMemsetAsync(cur, 0, sizeof(unsigned int), ctx->stream));
    k_any_batch<<<traceGrid(ctx), FLX_TRACE_THREADS, 0, ctx->stream>>>(ctx->bvh, dRays, (unsigned int)n, cur, dOcc);
    ctx->launches++;
    FLX_CUDA(ctx, cudaGetLastError());
    return FLX_OK;
}
