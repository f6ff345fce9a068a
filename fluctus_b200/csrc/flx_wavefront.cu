// Wavefront stages other than ray traversal: reset, raygen, logic (state machine + splat + NEE +
// queue compaction), material shaders, postprocess, plus the device-side bookkeeping that lets a
// whole iteration run from a CUDA graph without host round trips.
// Behavioural contract: reference src/wf_reset.cl, wf_raygen.cl, wf_logic.cl, wf_mat_*.cl,
// mk_postprocess.cl, tonemap.cl.
#include "flx_bxdf.cuh"
#include "flx_raygen.cuh"

#define FLX_WF_THREADS 256

// Warp-aggregated queue push (the reference's atomicAggInc, src/ptx_asm.cl:92-111, expressed
// with the native intrinsics): one atomicAdd per warp per queue, slot = base + rank among the
// pushing lanes.  Must be called by all 32 lanes of a converged warp.
FLX_DEV uint32_t warpPush(uint32_t *counter, bool pred)
{
    const unsigned mask = __ballot_sync(0xffffffffu, pred);
    if (mask == 0u) return 0u;
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(mask & ((1u << lane) - 1u));
}

// wf_reset.cl:5-68, launched over max(NUM_TASKS, w*h) like clcontext.cpp:783-788
__global__ void __launch_bounds__(FLX_WF_THREADS)
k_reset(PathState st, Queues qs, FrameBuffers fb, uint32_t N, uint32_t seedOffset, float worldRadius)
{
    const uint32_t total = max(N, fb.numPixels);
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < total; gid += gridDim.x * blockDim.x)
    {
        if (gid < fb.numPixels)
        {
            fb.pixels[gid] = make_float4(0.f, 0.f, 0.f, 0.f);
            fb.denoiserNormal[gid] = make_float4(0.f, 0.f, 0.f, 0.f);
            fb.denoiserAlbedo[gid] = make_float4(0.1f, 0.1f, 0.1f, 0.f);
            fb.spp[gid] = 0u;
        }
        if (gid >= N) continue;
        resetPathFields(st, gid, worldRadius);
        st.acc[gid] = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));        // Ei = 0, pixelIndex = 0
        st.thr[gid] = make_float4(1.f, 1.f, 1.f, __uint_as_float(gid + seedOffset)); // T = 1, seed = gid (+rank offset)
        st.ray0[gid].w = 1.0f;                                                 // lastPdfW
        st.ray1[gid].w = __uint_as_float(0u);                                  // pathLen
        qs.q[FLX_Q_RAYGEN][gid] = gid;
        if (gid == 0) qs.counters->raygenQueue = N;
    }
}

// wf_raygen.cl:4-100
__global__ void __launch_bounds__(FLX_WF_THREADS)
k_raygen(PathState st, Queues qs, const uint32_t *__restrict__ currPixelIdx, uint32_t tileRank, uint32_t tileWorld,
         const __grid_constant__ flx_render_params params)
{
    const uint32_t qlen = qs.counters->raygenQueue;
    const uint32_t rounded = (qlen + 31u) & ~31u;
    const uint32_t cursor = *currPixelIdx;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < rounded; k += gridDim.x * blockDim.x)
    {
        const bool valid = k < qlen;
        uint32_t gid = 0;
        if (valid)
        {
            gid = qs.q[FLX_Q_RAYGEN][k];
            generateCameraPath(st, gid, __float_as_uint(st.thr[gid].w), cursor, k, tileRank, tileWorld, params);
        }
        const uint32_t slot = warpPush(&qs.counters->extensionQueue, valid);
        if (valid) qs.q[FLX_Q_EXTENSION][slot] = gid;
    }
}

// mk_postprocess.cl:7-56 + tonemap.cl:8-33
FLX_DEV f3 uc2Func(f3 x)
{
    const float A = 0.22f, B = 0.30f, CB = 0.03f, DE = 0.002f, DF = 0.06f, ExF = 0.033333333333333333333f;
    return ((x * (A * x + mk3(CB)) + mk3(DE)) / (x * (A * x + mk3(B)) + mk3(DF))) - mk3(ExF);
}
__global__ void __launch_bounds__(FLX_WF_THREADS)
k_postprocess(FrameBuffers fb, float exposure, uint32_t tmOperator, int denoiserAOV)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < fb.numPixels; gid += gridDim.x * blockDim.x)
    {
        float4 c = fb.pixels[gid];
        if (c.w > 0.0f) { const float w = c.w; c.x = c.x / w; c.y = c.y / w; c.z = c.z / w; c.w = c.w / w; }
        f3 col = mk3(c);
        col *= exposure;
        if (tmOperator == 1u) col = col / (mk3(1.0f) + col);
        if (tmOperator == 2u) col = uc2Func(2.0f * col) / uc2Func(mk3(11.2f));
        if (tmOperator != 3u) col = pow3x(col, FLX_ONE_2_2);
        fb.preview[gid] = mk4(col, c.w);
        if (denoiserAOV) // mk_postprocess.cl:49-55: normalised copies for an external denoiser
        {
            const float4 n = fb.denoiserNormal[gid], a = fb.denoiserAlbedo[gid];
            fb.denoiserNormalOut[gid] = (n.w > 1.0f) ? make_float4(n.x / n.w, n.y / n.w, n.z / n.w, n.w / n.w) : n;
            fb.denoiserAlbedoOut[gid] = (a.w > 1.0f) ? make_float4(a.x / a.w, a.y / a.w, a.z / a.w, a.w / a.w) : a;
        }
    }
}

// What Tracer does on the host after each iteration (tracer.cpp:761-773, clcontext.cpp:903-921):
// add the queue lengths to the statistics, advance the pixel cursor by the number of regenerated
// paths, zero the counters.  One thread; keeps flx_render free of host synchronisation.
__global__ void k_bookkeeping(flx_queue_counters *counters, RenderTotals *totals, uint32_t *pixelIdx, uint32_t numPixels)
{
    closeIteration(counters, totals, pixelIdx, numPixels);
}

__global__ void k_advance_pixel(uint32_t *pixelIdx, uint32_t numPixels, uint32_t numNew)
{
    *pixelIdx = (*pixelIdx + numNew) % numPixels;
}

// ---- state import/export in the reference's 64-slot SoA layout (fluctus_abi.h) -----------------
FLX_DEV void put3(uint32_t *s, uint32_t N, uint32_t gid, int slot, float4 v)
{
    s[(size_t)slot * N + gid] = __float_as_uint(v.x);
    s[(size_t)(slot + 1) * N + gid] = __float_as_uint(v.y);
    s[(size_t)(slot + 2) * N + gid] = __float_as_uint(v.z);
}
FLX_DEV float4 get3(const uint32_t *s, uint32_t N, uint32_t gid, int slot, int wslot)
{
    return make_float4(__uint_as_float(s[(size_t)slot * N + gid]), __uint_as_float(s[(size_t)(slot + 1) * N + gid]),
                       __uint_as_float(s[(size_t)(slot + 2) * N + gid]), wslot >= 0 ? __uint_as_float(s[(size_t)wslot * N + gid]) : 0.0f);
}
#define PUTW(slot, val) s[(size_t)(slot) * N + gid] = (val)
#define GETW(slot) s[(size_t)(slot) * N + gid]

__global__ void k_export_state(PathState st, uint32_t *s, uint32_t N)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < N; gid += gridDim.x * blockDim.x)
    {
        float4 v;
        v = st.ray0[gid]; put3(s, N, gid, FLX_SLOT_ORIG, v); PUTW(FLX_SLOT_LAST_PDF_W, __float_as_uint(v.w));
        v = st.ray1[gid]; put3(s, N, gid, FLX_SLOT_DIR, v); PUTW(FLX_SLOT_PATH_LEN, __float_as_uint(v.w));
        v = st.hit0[gid]; put3(s, N, gid, FLX_SLOT_P, v); PUTW(FLX_SLOT_HIT_T, __float_as_uint(v.w));
        v = st.hit1[gid]; put3(s, N, gid, FLX_SLOT_N, v); PUTW(FLX_SLOT_HIT_I, __float_as_uint(v.w));
        v = st.hit2[gid]; PUTW(FLX_SLOT_UV, __float_as_uint(v.x)); PUTW(FLX_SLOT_UV + 1, __float_as_uint(v.y));
        PUTW(FLX_SLOT_MAT_ID, __float_as_uint(v.z)); PUTW(FLX_SLOT_AREA_LIGHT_HIT, __float_as_uint(v.w));
        v = st.thr[gid]; put3(s, N, gid, FLX_SLOT_T, v); PUTW(FLX_SLOT_SEED, __float_as_uint(v.w));
        v = st.acc[gid]; put3(s, N, gid, FLX_SLOT_EI, v); PUTW(FLX_SLOT_PIXEL_INDEX, __float_as_uint(v.w));
        v = st.shw0[gid]; put3(s, N, gid, FLX_SLOT_SHADOW_ORIG, v); PUTW(FLX_SLOT_SHADOW_LEN, __float_as_uint(v.w));
        v = st.shw1[gid]; put3(s, N, gid, FLX_SLOT_SHADOW_DIR, v); PUTW(FLX_SLOT_LAST_COS_TH, __float_as_uint(v.w));
        v = st.nee0[gid]; put3(s, N, gid, FLX_SLOT_LAST_EMISSION, v); PUTW(FLX_SLOT_LAST_PDF_DIRECT, __float_as_uint(v.w));
        v = st.nee1[gid]; put3(s, N, gid, FLX_SLOT_LAST_BSDF, v); PUTW(FLX_SLOT_LAST_PDF_IMPLICIT, __float_as_uint(v.w));
        v = st.nee2[gid]; put3(s, N, gid, FLX_SLOT_LAST_T, v);
        PUTW(FLX_SLOT_LAST_SPECULAR, st.lastSpecular[gid]);
        PUTW(FLX_SLOT_SHADOW_BLOCKED, st.shadowBlocked[gid]);
        const uint32_t lf = st.logicFlags[gid];
        PUTW(FLX_SLOT_BACKFACE, lf & 1u); PUTW(FLX_SLOT_FIRST_DIFFUSE, (lf >> 1) & 1u);
        PUTW(FLX_SLOT_LAST_PICK_PROB, __float_as_uint(st.lastPickProb[gid]));
        PUTW(FLX_SLOT_PHASE, (uint32_t)st.phase[gid]);
    }
}

__global__ void k_import_state(PathState st, const uint32_t *s, uint32_t N)
{
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < N; gid += gridDim.x * blockDim.x)
    {
        st.ray0[gid] = get3(s, N, gid, FLX_SLOT_ORIG, FLX_SLOT_LAST_PDF_W);
        st.ray1[gid] = get3(s, N, gid, FLX_SLOT_DIR, FLX_SLOT_PATH_LEN);
        st.hit0[gid] = get3(s, N, gid, FLX_SLOT_P, FLX_SLOT_HIT_T);
        st.hit1[gid] = get3(s, N, gid, FLX_SLOT_N, FLX_SLOT_HIT_I);
        st.hit2[gid] = make_float4(__uint_as_float(GETW(FLX_SLOT_UV)), __uint_as_float(GETW(FLX_SLOT_UV + 1)),
                                   __uint_as_float(GETW(FLX_SLOT_MAT_ID)), __uint_as_float(GETW(FLX_SLOT_AREA_LIGHT_HIT)));
        st.thr[gid] = get3(s, N, gid, FLX_SLOT_T, FLX_SLOT_SEED);
        st.acc[gid] = get3(s, N, gid, FLX_SLOT_EI, FLX_SLOT_PIXEL_INDEX);
        st.shw0[gid] = get3(s, N, gid, FLX_SLOT_SHADOW_ORIG, FLX_SLOT_SHADOW_LEN);
        st.shw1[gid] = get3(s, N, gid, FLX_SLOT_SHADOW_DIR, FLX_SLOT_LAST_COS_TH);
        st.nee0[gid] = get3(s, N, gid, FLX_SLOT_LAST_EMISSION, FLX_SLOT_LAST_PDF_DIRECT);
        st.nee1[gid] = get3(s, N, gid, FLX_SLOT_LAST_BSDF, FLX_SLOT_LAST_PDF_IMPLICIT);
        st.nee2[gid] = get3(s, N, gid, FLX_SLOT_LAST_T, -1);
        st.lastSpecular[gid] = GETW(FLX_SLOT_LAST_SPECULAR);
        st.shadowBlocked[gid] = GETW(FLX_SLOT_SHADOW_BLOCKED);
        st.logicFlags[gid] = (GETW(FLX_SLOT_BACKFACE) & 1u) | ((GETW(FLX_SLOT_FIRST_DIFFUSE) & 1u) << 1);
        st.lastPickProb[gid] = __uint_as_float(GETW(FLX_SLOT_LAST_PICK_PROB));
        st.phase[gid] = (int32_t)GETW(FLX_SLOT_PHASE);
    }
}

// ---- launchers -----------------------------------------------------------------------------------
static inline int wfGrid(const flx_ctx *ctx, uint64_t items)
{
    const uint64_t blocks = (items + FLX_WF_THREADS - 1) / FLX_WF_THREADS;
    const uint64_t cap = (uint64_t)ctx->numSMs * 8; // grid-stride beyond 8 resident blocks per SM
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}
#define FLX_LAUNCHED(ctx) do { (ctx)->launches++; FLX_CUDA(ctx, cudaGetLastError()); } while (0)

int flx_launch_reset(flx_ctx *ctx)
{
    const uint64_t total = ctx->N > ctx->fb.numPixels ? ctx->N : ctx->fb.numPixels;
    k_reset<<<wfGrid(ctx, total), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs, ctx->fb, ctx->N, ctx->seedOffset, ctx->params.worldRadius);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_raygen(flx_ctx *ctx)
{
    k_raygen<<<wfGrid(ctx, ctx->N), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs, ctx->dPixelIdx, ctx->tileRank, ctx->tileWorld, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_postprocess(flx_ctx *ctx)
{
    k_postprocess<<<wfGrid(ctx, ctx->fb.numPixels), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->fb, ctx->params.ppParams.exposure, ctx->params.ppParams.tmOperator, ctx->denoiserAOV);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_bookkeeping(flx_ctx *ctx)
{
    k_bookkeeping<<<1, 1, 0, ctx->stream>>>(ctx->qs.counters, ctx->dTotals, ctx->dPixelIdx, flx_owned_pixels(ctx));
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_advance_pixel(flx_ctx *ctx, uint32_t numPixels, uint32_t numNew)
{
    k_advance_pixel<<<1, 1, 0, ctx->stream>>>(ctx->dPixelIdx, numPixels, numNew);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_export_state(flx_ctx *ctx, uint32_t *dSlots)
{
    k_export_state<<<wfGrid(ctx, ctx->N), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, dSlots, ctx->N);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

int flx_launch_import_state(flx_ctx *ctx, const uint32_t *dSlots)
{
    k_import_state<<<wfGrid(ctx, ctx->N), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, dSlots, ctx->N);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}
