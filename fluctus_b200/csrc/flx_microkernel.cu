// The microkernel ("MK") renderer: one task per pixel, a phase word per task, exact sample counts.
// Behavioural contract: reference src/mk_reset.cl, mk_raygen.cl, mk_next_vertex.cl, mk_sample_bsdf.cl,
// mk_splat.cl, mk_splat_preview.cl, driven by Tracer::renderSingle (src/tracer.cpp:108-182) and the MK branch
// of the benchmark loops (:478-481, :749-752).
//
// What is different from the reference's kernels: `nextVertex` and `sampleBsdf` trace their rays INLINE, one
// thread per pixel, which is the pattern the reference itself marks "BAD! Collect all shadow ray casts
// together" (mk_sample_bsdf.cl:92).  Here every stage is split at its ray casts; the casts run on the same
// persistent traversal engine as the wavefront stages (flx_trace.cu: MkVertexJob / MkShadowJob, tasks filtered
// by phase at load time, no queue) and the shading halves are plain grid-stride kernels:
//     nextVertex = k_mk_vertex (closest hit + hit record + pathLen)  ->  k_mk_vertex_shade
//     sampleBsdf = k_mk_bsdf<NONE, ENV> -> k_mk_shadow<light> -> k_mk_bsdf<ENV, AREA> -> k_mk_shadow<no light>
//                  -> k_mk_bsdf<AREA, FINISH>           (steps for lights that are switched off are skipped)
// Per task the arithmetic and the order of RNG draws and of additions into Ei are the reference's, and no
// task ever touches another task's state or pixel: images are bit-reproducible from run to run.
#include "flx_bxdf.cuh"
#include "flx_raygen.cuh"

#define FLX_MK_THREADS 256
#define FLX_ALL_TYPES (FLX_BXDF_DIFFUSE | FLX_BXDF_GLOSSY | FLX_BXDF_GGX_ROUGH_REFLECTION | FLX_BXDF_IDEAL_REFLECTION | \
                       FLX_BXDF_GGX_ROUGH_DIELECTRIC | FLX_BXDF_IDEAL_DIELECTRIC | FLX_BXDF_EMISSIVE)

// RenderStats counters (geom.h:279-285): one atomic per warp, like the -DNVIDIA ballot path of mk_next_vertex.cl:46-52
FLX_DEV void statAdd(uint32_t *counter, uint32_t mine)
{
    const uint32_t total = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31u) == 0u && total) atomicAdd(counter, total);
}

// mk_reset.cl:4-45
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_reset(PathState st, FrameBuffers fb, uint32_t limit, uint32_t seedOffset)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < limit; gid += gridDim.x * blockDim.x)
    {
        fb.pixels[gid] = make_float4(0.f, 0.f, 0.f, 0.f);
        fb.denoiserNormal[gid] = make_float4(0.f, 0.f, 0.f, 0.f);
        fb.denoiserAlbedo[gid] = make_float4(0.1f, 0.1f, 0.1f, 0.f);
        fb.spp[gid] = 0u;
        st.phase[gid] = FLX_MK_GENERATE_CAMERA_RAY;
        st.acc[gid] = mk4(mk3(0.0f), st.acc[gid].w);          // Ei = 0
        st.thr[gid] = mk4u(mk3(1.0f), gid + seedOffset);      // T = 1, seed = gid (+ rank offset)
        st.ray1[gid].w = __uint_as_float(0u);                 // pathLen
        st.ray0[gid].w = 1.0f;                                // lastPdfW
        st.lastSpecular[gid] = 1u;
        st.logicFlags[gid] &= ~2u;                            // firstDiffuseHit = 0
    }
}

// mk_raygen.cl:5-67: task gid looks through pixel gid
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_raygen(PathState st, uint32_t limit, const __grid_constant__ flx_render_params params)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < limit; gid += gridDim.x * blockDim.x)
    {
        if (st.phase[gid] != FLX_MK_GENERATE_CAMERA_RAY) continue;
        const float4 thr = st.thr[gid];
        uint32_t seed = __float_as_uint(thr.w);
        f3 o, d;
        cameraRay(gid, seed, params, o, d);
        st.ray0[gid] = mk4(o, st.ray0[gid].w);
        st.ray1[gid] = mk4(d, st.ray1[gid].w);
        st.thr[gid] = mk4u(mk3(thr), seed);
        st.phase[gid] = FLX_MK_RT_NEXT_VERTEX;
    }
}

// mk_next_vertex.cl:43-122: everything after the trace (the hit record and pathLen += 1 were written by k_mk_vertex)
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_vertex_shade(PathState st, FrameBuffers fb, EnvMapDev env, flx_render_stats *stats, uint32_t limit, int denoiserAOV,
                  const __grid_constant__ flx_render_params params)
{
    const uint32_t rounded = (limit + 31u) & ~31u;
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < rounded; gid += gridDim.x * blockDim.x)
    {
        const bool mine = gid < limit && st.phase[gid] == FLX_MK_RT_NEXT_VERTEX;
        uint32_t primary = 0u, extension = 0u;
        if (mine)
        {
            const float4 r0 = st.ray0[gid], r1 = st.ray1[gid], h0 = st.hit0[gid], h1 = st.hit1[gid], h2 = st.hit2[gid];
            const uint32_t len = __float_as_uint(r1.w); // already incremented
            const f3 rayOrig = mk3(r0), rayDir = mk3(r1), hitN = mk3(h1);
            const int hitI = __float_as_int(h1.w), areaLightHit = __float_as_int(h2.w);
            if (len == 1u) primary = 1u; else extension = 1u;

            if (denoiserAOV && len == 1u) // :59-69 (one task per pixel: no atomics needed)
            {
                const f3 c1 = ld3(params.camera.right), c2 = ld3(params.camera.up), c3 = -ld3(params.camera.dir);
                float4 a = fb.denoiserNormal[gid];
                a.x += dot(c1, hitN); a.y += dot(c2, hitN); a.z += dot(c3, hitN); a.w += 1.0f;
                fb.denoiserNormal[gid] = a;
            }

            if (hitI < 0) // implicit environment sample (:72-93)
            {
                f3 bg = mk3(0.0f);
                if (params.useEnvMap && (len == 1u || params.sampleImpl)) bg = evalEnvMapDir(env, rayDir) * params.envMapStrength;
                float weight = 1.0f;
                if (params.sampleImpl && params.sampleExpl && params.useEnvMap && len > 1u && !st.lastSpecular[gid])
                {
                    const float lightPickProb = 1.0f;
                    const float directPdfW = envMapPdf(env, rayDir);
                    const float actualPdfW = r0.w;
                    weight = (actualPdfW * lightPickProb) / (actualPdfW * lightPickProb + directPdfW);
                }
                const f3 T = mk3(st.thr[gid]);
                const float4 acc = st.acc[gid];
                st.acc[gid] = mk4(mk3(acc) + weight * T * bg, acc.w);
                st.phase[gid] = FLX_MK_SPLAT_SAMPLE;
            }
            else if (areaLightHit) // implicit area-light sample (:95-116)
            {
                float misWeight = 1.0f;
                if (params.sampleExpl && len > 1u && !st.lastSpecular[gid])
                {
                    const float directPdfA = recip(4.0f * params.areaLight.size.x * params.areaLight.size.y);
                    const float directPdfW = pdfAtoW(directPdfA, length(mk3(h0) - rayOrig), -dot(rayDir, hitN));
                    const float lightPickProb = 1.0f;
                    const float lastPdfW = r0.w;
                    misWeight = lastPdfW / (lastPdfW + directPdfW * lightPickProb);
                }
                const f3 T = mk3(st.thr[gid]);
                const float4 acc = st.acc[gid];
                st.acc[gid] = mk4(mk3(acc) + T * misWeight * ld3(params.areaLight.E), acc.w);
                st.phase[gid] = FLX_MK_SPLAT_SAMPLE;
            }
            else st.phase[gid] = FLX_MK_SAMPLE_BSDF;
        }
        statAdd(&stats->primaryRays, primary);
        statAdd(&stats->extensionRays, extension);
    }
}

// One step of mk_sample_bsdf.cl:11-188 for the tasks in phase MK_SAMPLE_BSDF.
//   RESOLVE: which light's shadow ray has just come back (its NEE contribution is added now)
//   NEXT:    which light is sampled now (its shadow ray is traced by the next launch), or FINISH
enum { MK_NONE = 0, MK_ENV = 1, MK_AREA = 2, MK_FINISH = 3 };
template <int RESOLVE, int NEXT>
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_bsdf(PathState st, FrameBuffers fb, MaterialsDev md, EnvMapDev env, flx_render_stats *stats, uint32_t limit, int denoiserAOV,
          const __grid_constant__ flx_render_params params)
{
    const uint32_t rounded = (limit + 31u) & ~31u;
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < rounded; gid += gridDim.x * blockDim.x)
    {
        const bool mine = gid < limit && st.phase[gid] == FLX_MK_SAMPLE_BSDF;
        uint32_t shadowRays = 0u;
        if (mine)
        {
            const float4 thr = st.thr[gid], r1 = st.ray1[gid], h0 = st.hit0[gid], h1 = st.hit1[gid], h2 = st.hit2[gid];
            uint32_t seed = __float_as_uint(thr.w);
            const f3 T = mk3(thr), rayDir = mk3(r1), hitP = mk3(h0);
            const int hitI = __float_as_int(h1.w);
            const flx_material *mp = md.materials + __float_as_int(h2.z);
            const Mat mat = loadMat(mp);
            const int mapN = __float_as_int(__ldg(((const float4 *)mp) + 5).y);

            // shading frame (:46-53), recomputed by every step instead of being written back: the reference keeps
            // the geometric normal in the state
            SurfHit hit; hit.N = mk3(h1); hit.uvx = h2.x; hit.uvy = h2.y;
            if (mapN != -1) hit.N = tangentSpaceNormal(hit.N, hit.uvx, hit.uvy, hitI, mapN, md);
            const bool backface = dot(hit.N, rayDir) > 0.0f;
            if (backface) hit.N = -hit.N;
            const f3 orig = hitP - 1e-3f * rayDir;
            const bool nee = params.sampleExpl && !FLX_BXDF_IS_SINGULAR(mat.type);

            if (RESOLVE == MK_NONE && denoiserAOV) // :56-67
            {
                const uint32_t flags = st.logicFlags[gid];
                if (!FLX_BXDF_IS_SINGULAR(mat.type) && !(flags & 2u))
                {
                    st.logicFlags[gid] = flags | 2u;
                    const f3 albedo = matGetFloat3(mk3(__ldg((const float4 *)mp)), hit.uvx, hit.uvy, mat.map_Kd, md); // raw Kd: no gamma on the AOV
                    float4 a = fb.denoiserAlbedo[gid];
                    a.x += albedo.x; a.y += albedo.y; a.z += albedo.z; a.w += 1.0f;
                    fb.denoiserAlbedo[gid] = a;
                }
            }

            // ---- contribution of the light sampled by the previous step (:99-107, :130-139) -------------------
            if (RESOLVE != MK_NONE && nee && st.shadowBlocked[gid] == 0u)
            {
                const float lightPickProb = 1.0f;
                const float4 s0 = st.shw0[gid];
                const f3 L = mk3(st.shw1[gid]);
                const float pdfSampled = st.nee0[gid].w;
                f3 Li = mk3(0.0f); float directPdfW = 0.0f; bool ok;
                if (RESOLVE == MK_ENV)
                {
                    directPdfW = pdfSampled;
                    ok = directPdfW != 0.0f;
                    if (ok) Li = evalEnvMapDir(env, L) * params.envMapStrength;
                }
                else
                {
                    const float cosLight = fmaxf(dot(ld3(params.areaLight.N), -L), 0.0f);
                    ok = cosLight > 0.0f;
                    if (ok) { directPdfW = pdfAtoW(pdfSampled, s0.w, cosLight); Li = ld3(params.areaLight.E); }
                }
                if (ok)
                {
                    const f3 brdf = bxdfEval<FLX_ALL_TYPES>(hit, mat, backface, md, rayDir, L);
                    const float cosTh = fmaxf(0.0f, dot(L, hit.N));
                    const float bsdfPdfW = fmaxf(0.0f, bxdfPdf<FLX_ALL_TYPES>(hit, mat, backface, md, rayDir, L));
                    const f3 contrib = brdf * T * Li * cosTh / (lightPickProb * directPdfW + (float)(params.sampleImpl) * bsdfPdfW);
                    const float4 acc = st.acc[gid];
                    st.acc[gid] = mk4(mk3(acc) + contrib, acc.w);
                }
            }

            // ---- sample the next light and park its shadow ray (:75-91, :111-123) ------------------------------
            if (NEXT == MK_ENV || NEXT == MK_AREA)
            {
                uint32_t slot = FLX_MK_NEE_NONE;
                if (nee)
                {
                    if (NEXT == MK_ENV)
                    {
                        f3 L; float directPdfW = 0.0f;
                        sampleEnvMapAlias(rnd(seed), L, directPdfW, env);
                        st.shw0[gid] = mk4(orig, params.worldRadius + params.worldRadius);
                        st.shw1[gid] = mk4(L, 0.0f);
                        st.nee0[gid].w = directPdfW;
                    }
                    else
                    {
                        float directPdfA; f3 posL;
                        sampleAreaLight(params.areaLight, directPdfA, posL, seed);
                        f3 L = posL - orig;
                        const float lenL = length(L);
                        L /= lenL;
                        st.shw0[gid] = mk4(orig, lenL);
                        st.shw1[gid] = mk4(L, 0.0f);
                        st.nee0[gid].w = directPdfA;
                    }
                    slot = FLX_MK_NEE_PENDING;
                    shadowRays = 1u;
                    st.thr[gid] = mk4u(T, seed);
                }
                st.shadowBlocked[gid] = slot;
            }

            // ---- termination, continuation ray, state update (:143-187) ---------------------------------------
            if (NEXT == MK_FINISH)
            {
                float contProb = 1.0f;
                const uint32_t len = __float_as_uint(r1.w);
                bool terminate = params.maxBounces > 0u && (len >= params.maxBounces + 1u);
                if (!terminate && params.useRoulette && len > 5u)
                {
                    contProb = clampf(luminance(T), 0.01f, 0.5f);
                    terminate = (rnd(seed) > contProb);
                }
                float pdfW = 0.0f;     // left unset by the reference for emissive / rejected glossy samples;
                f3 newDir = mk3(0.0f); // pinned to 0 like the oracle
                const f3 bsdf = bxdfSample<FLX_ALL_TYPES>(hit, mat, backface, md, rayDir, newDir, pdfW, seed);
                pdfW *= contProb;
                if (pdfW == 0.0f || isZero(bsdf)) terminate = true;
                const f3 newT = T * bsdf * dot(hit.N, newDir) / pdfW;
                const f3 newOrig = hitP + 1e-4f * newDir;
                st.thr[gid] = mk4u(newT, seed);
                st.ray0[gid] = mk4(newOrig, pdfW);
                st.ray1[gid] = make_float4(newDir.x, newDir.y, newDir.z, r1.w);
                st.lastSpecular[gid] = FLX_BXDF_IS_SINGULAR(mat.type) ? 1u : 0u;
                st.phase[gid] = terminate ? FLX_MK_SPLAT_SAMPLE : FLX_MK_RT_NEXT_VERTEX;
            }
        }
        if (NEXT == MK_ENV || NEXT == MK_AREA) statAdd(&stats->shadowRays, shadowRays);
    }
}

// mk_splat.cl:5-57
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_splat(PathState st, FrameBuffers fb, flx_render_stats *stats, uint32_t limit, uint32_t maxSpp)
{
    const uint32_t rounded = (limit + 31u) & ~31u;
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < rounded; gid += gridDim.x * blockDim.x)
    {
        uint32_t splatted = 0u;
        if (gid < limit && st.phase[gid] == FLX_MK_SPLAT_SAMPLE)
        {
            bool shouldSplat = true;
            if (maxSpp > 0u) // -DCHECK_SPP: `<=` on the old value admits maxSpp + 1 samples (quirk kept)
            {
                const uint32_t old = fb.spp[gid];
                fb.spp[gid] = old + 1u;
                shouldSplat = old <= maxSpp;
                if (!shouldSplat) { st.phase[gid] = FLX_MK_DONE; fb.spp[gid] = maxSpp; }
            }
            if (shouldSplat)
            {
                const float4 acc = st.acc[gid];
                float4 color = make_float4(acc.x, acc.y, acc.z, 1.0f);
                const float4 prev = fb.pixels[gid];
                if (prev.w > 0.0f) { color.x += prev.x; color.y += prev.y; color.z += prev.z; color.w += prev.w; }
                fb.pixels[gid] = color;
                splatted = 1u;
                st.acc[gid] = mk4(mk3(0.0f), acc.w);
                st.thr[gid] = mk4(mk3(1.0f), st.thr[gid].w);
                st.ray1[gid].w = __uint_as_float(0u);
                st.logicFlags[gid] &= ~2u;
                st.phase[gid] = FLX_MK_GENERATE_CAMERA_RAY;
            }
        }
        statAdd(&stats->samples, splatted);
    }
}

// mk_splat_preview.cl:5-25
__global__ void __launch_bounds__(FLX_MK_THREADS)
k_mk_splat_preview(PathState st, FrameBuffers fb, uint32_t limit)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < limit; gid += gridDim.x * blockDim.x)
    {
        const float4 acc = st.acc[gid];
        fb.pixels[gid] = make_float4(acc.x, acc.y, acc.z, 0.0f); // weight 0: overwritten by the next splat
        st.acc[gid] = mk4(mk3(0.0f), acc.w);
        st.thr[gid] = mk4(mk3(1.0f), st.thr[gid].w);
        st.ray1[gid].w = __uint_as_float(0u);
        st.phase[gid] = FLX_MK_GENERATE_CAMERA_RAY;
    }
}

// ---- launchers -----------------------------------------------------------------------------------
static inline uint32_t mkLimit(const flx_ctx *ctx) { return ctx->fb.numPixels < ctx->N ? ctx->fb.numPixels : ctx->N; }
static inline int mkGrid(const flx_ctx *ctx, uint64_t items)
{
    const uint64_t blocks = (items + FLX_MK_THREADS - 1) / FLX_MK_THREADS;
    const uint64_t cap = (uint64_t)ctx->numSMs * 8;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}
#define FLX_LAUNCHED(ctx) do { (ctx)->launches++; FLX_CUDA(ctx, cudaGetLastError()); } while (0)
#define FLX_TRY(expr) do { int rc__ = (expr); if (rc__ != FLX_OK) return rc__; } while (0)
#define MK_ARGS(ctx) (ctx)->st, (ctx)->fb, (ctx)->mats, (ctx)->env, (ctx)->dStats, limit, (ctx)->denoiserAOV, (ctx)->params

int flx_launch_mk_reset(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    FLX_CUDA(ctx, cudaMemsetAsync(ctx->dStats, 0, sizeof(flx_render_stats), ctx->stream));
    k_mk_reset<<<mkGrid(ctx, limit), FLX_MK_THREADS, 0, ctx->stream>>>(ctx->st, ctx->fb, limit, ctx->seedOffset);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_mk_raygen(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    k_mk_raygen<<<mkGrid(ctx, limit), FLX_MK_THREADS, 0, ctx->stream>>>(ctx->st, limit, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_mk_next_vertex(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    FLX_TRY(flx_launch_mk_vertex_trace(ctx, limit));
    k_mk_vertex_shade<<<mkGrid(ctx, limit), FLX_MK_THREADS, 0, ctx->stream>>>(ctx->st, ctx->fb, ctx->env, ctx->dStats, limit, ctx->denoiserAOV, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_mk_sample_bsdf(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    const int grid = mkGrid(ctx, limit);
    const bool envNEE = ctx->params.sampleExpl && ctx->params.useEnvMap, areaNEE = ctx->params.sampleExpl && ctx->params.useAreaLight;
    if (envNEE && areaNEE)
    {
        k_mk_bsdf<MK_NONE, MK_ENV><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
        FLX_TRY(flx_launch_mk_shadow_trace(ctx, limit, 1));
        k_mk_bsdf<MK_ENV, MK_AREA><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
        FLX_TRY(flx_launch_mk_shadow_trace(ctx, limit, 0));
        k_mk_bsdf<MK_AREA, MK_FINISH><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
    }
    else if (envNEE)
    {
        k_mk_bsdf<MK_NONE, MK_ENV><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
        FLX_TRY(flx_launch_mk_shadow_trace(ctx, limit, 1));
        k_mk_bsdf<MK_ENV, MK_FINISH><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
    }
    else if (areaNEE)
    {
        k_mk_bsdf<MK_NONE, MK_AREA><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
        FLX_TRY(flx_launch_mk_shadow_trace(ctx, limit, 0));
        k_mk_bsdf<MK_AREA, MK_FINISH><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
    }
    else
    {
        k_mk_bsdf<MK_NONE, MK_FINISH><<<grid, FLX_MK_THREADS, 0, ctx->stream>>>(MK_ARGS(ctx)); FLX_LAUNCHED(ctx);
    }
    return FLX_OK;
}

int flx_launch_mk_splat(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    k_mk_splat<<<mkGrid(ctx, limit), FLX_MK_THREADS, 0, ctx->stream>>>(ctx->st, ctx->fb, ctx->dStats, limit, ctx->params.maxSpp);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_mk_splat_preview(flx_ctx *ctx)
{
    const uint32_t limit = mkLimit(ctx);
    k_mk_splat_preview<<<mkGrid(ctx, limit), FLX_MK_THREADS, 0, ctx->stream>>>(ctx->st, ctx->fb, limit);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}
