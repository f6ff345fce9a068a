// C-ABI of libfluctus_cuda.so (include/fluctus_cuda.h): context life cycle, uploads, stage
// dispatch, read-back.  Replaces the reference's CLContext (src/clcontext.cpp); there is no CPU
// or library fallback — without an sm_100 device flx_create fails.
#include "flx_internal.h"
#include "flx_libm.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>

static std::string g_createError;

int flx_fail(flx_ctx *ctx, int code, const std::string &msg)
{
    if (ctx) ctx->err = msg; else g_createError = msg;
    return code;
}

#define FLX_CHECK_CTX(ctx) do { if (!(ctx)) return FLX_ERR_ARG; } while (0)
#define FLX_TRY(expr) do { int rc__ = (expr); if (rc__ != FLX_OK) return rc__; } while (0)

static void invalidateGraph(flx_ctx *ctx)
{
    // replays of the old graph may still be in flight on the stream: let them drain before the executable goes away
    if ((ctx->graphExec || ctx->mkGraphExec) && ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->graphExec) { cudaGraphExecDestroy(ctx->graphExec); ctx->graphExec = nullptr; }
    ctx->graphValid = false;
    if (ctx->mkGraphExec) { cudaGraphExecDestroy(ctx->mkGraphExec); ctx->mkGraphExec = nullptr; }
    ctx->mkGraphValid = false;
}

// ---- stage timers ---------------------------------------------------------------------------------
static void drainTimers(flx_ctx *ctx)
{
    for (int s = 0; s < 6; s++)
    {
        StageTimer &t = ctx->timers[s];
        for (auto &p : t.pending)
        {
            float ms = 0.f;
            cudaEventSynchronize(p.second);
            cudaEventElapsedTime(&ms, p.first, p.second);
            t.ms += ms; t.launches++;
            cudaEventDestroy(p.first); cudaEventDestroy(p.second);
        }
        t.pending.clear();
    }
}

extern "C" {

const char *flx_last_error(flx_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }
uint32_t flx_num_tasks(flx_ctx *ctx) { return ctx ? ctx->N : 0; }
uint64_t flx_launch_count(flx_ctx *ctx) { return ctx ? ctx->launches : 0; }
void *flx_stream(flx_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

int flx_create(flx_ctx **out, int device, uint32_t num_tasks, void *cuda_stream)
{
    if (!out || num_tasks == 0) return flx_fail(nullptr, FLX_ERR_ARG, "flx_create: bad arguments");
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return flx_fail(nullptr, FLX_ERR_NO_DEVICE, "flx_create: no CUDA device visible (this backend has no CPU fallback)");
    if (device < 0 || device >= count) return flx_fail(nullptr, FLX_ERR_ARG, "flx_create: device index out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return flx_fail(nullptr, FLX_ERR_CUDA, "flx_create: cudaGetDeviceProperties failed");
    if (prop.major < 10)
        return flx_fail(nullptr, FLX_ERR_NO_DEVICE, std::string("flx_create: device '") + prop.name + "' is not sm_100-class; kernels are built for sm_100a only");

    flx_ctx *ctx = new flx_ctx();
    ctx->device = device; ctx->N = num_tasks; ctx->numSMs = prop.multiProcessorCount;
#define CREATE_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { \
        int rc__ = flx_fail(nullptr, FLX_ERR_CUDA, std::string("flx_create: " #call ": ") + cudaGetErrorString(e__)); flx_destroy(ctx); return rc__; } } while (0)
    CREATE_CUDA(cudaSetDevice(device));
    if (cuda_stream) ctx->stream = (cudaStream_t)cuda_stream;
    else { CREATE_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->ownStream = true; }

    // path state: twelve float4 streams + five scalar streams, zero-initialised like a fresh cl::Buffer
    // is not — the oracle starts from zeros too, so untouched fields compare equal.
    const size_t N = num_tasks;
    const size_t stateBytes = FLX_STATE_AOS ? N * (size_t)FLX_STATE_RECORD_BYTES
                                            : N * (FLX_STATE_VEC_STREAMS * sizeof(float4) + FLX_STATE_SCALAR_STREAMS * 4);
    CREATE_CUDA(cudaMalloc(&ctx->stateBlock, stateBytes));
    CREATE_CUDA(cudaMemsetAsync(ctx->stateBlock, 0, stateBytes, ctx->stream));
    {
        char *b = (char *)ctx->stateBlock;
        PathState &st = ctx->st;
        StateView<float4> *vec[FLX_STATE_VEC_STREAMS] = {&st.ray0, &st.ray1, &st.thr, &st.acc, &st.hit0, &st.hit1, &st.hit2, &st.shw0,
                                                         &st.shw1, &st.nee0, &st.nee1, &st.nee2};
#if FLX_STATE_AOS
        // byte offsets inside the 256-byte record: four 64-byte groups, fields that are gathered together share one
        const int vecOff[FLX_STATE_VEC_STREAMS] = {0, 16, 32, 48, 64, 80, 96, 128, 144, 160, 176, 192};
        for (int i = 0; i < FLX_STATE_VEC_STREAMS; i++) vec[i]->base = b + vecOff[i];
        st.lastSpecular.base = b + 112; st.shadowBlocked.base = b + 116; st.logicFlags.base = b + 120; st.lastPickProb.base = b + 124;
        st.phase.base = b + 208;
#else
        for (int i = 0; i < FLX_STATE_VEC_STREAMS; i++) vec[i]->base = b + (size_t)i * N * sizeof(float4);
        char *s = b + (size_t)FLX_STATE_VEC_STREAMS * N * sizeof(float4);
        st.lastSpecular.base = s; st.shadowBlocked.base = s + 4 * N; st.logicFlags.base = s + 8 * N; st.lastPickProb.base = s + 12 * N;
        st.phase.base = s + 16 * N;
#endif
    }
    // queues + counters + cursors + totals
    const size_t queueBytes = FLX_NUM_QUEUES * N * 4 + 512;
    CREATE_CUDA(cudaMalloc(&ctx->queueBlock, queueBytes));
    CREATE_CUDA(cudaMemsetAsync(ctx->queueBlock, 0, queueBytes, ctx->stream));
    {
        uint32_t *q = (uint32_t *)ctx->queueBlock;
        for (int i = 0; i < FLX_NUM_QUEUES; i++) ctx->qs.q[i] = q + (size_t)i * N;
        uint8_t *tail = (uint8_t *)(q + (size_t)FLX_NUM_QUEUES * N);
        ctx->qs.counters = (flx_queue_counters *)tail;       // 40 B
        ctx->dPixelIdx = (uint32_t *)(tail + 64);            // pixel cursor + 2 work cursors + CTA arrival counter
        ctx->dTotals = (RenderTotals *)(tail + 128);         // 48 B
        ctx->dStats = (flx_render_stats *)(tail + 192);      // 16 B
        ctx->dTraceStats = (unsigned long long *)(tail + 256); // 20 x u64
    }
    CREATE_CUDA(cudaMallocHost(&ctx->hCounters, 64 * sizeof(flx_queue_counters)));
    CREATE_CUDA(cudaStreamSynchronize(ctx->stream));
#undef CREATE_CUDA
    *out = ctx;
    return FLX_OK;
}

void flx_destroy(flx_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->comm) flx_comm_destroy(ctx);
    drainTimers(ctx);
    invalidateGraph(ctx);
    cudaFree(ctx->dNodes); cudaFree(ctx->dSlots); cudaFree(ctx->dShade);
    cudaFree(ctx->dMaterials); cudaFree(ctx->dTextures); cudaFree(ctx->dTexels); cudaFree(ctx->dTris160); cudaFree(ctx->dGammaLut);
    cudaFree(ctx->dEnv); cudaFree(ctx->dProb); cudaFree(ctx->dPdf); cudaFree(ctx->dAlias);
    cudaFree(ctx->stateBlock); cudaFree(ctx->queueBlock); cudaFree(ctx->fbBlock); cudaFree(ctx->dScratch);
    if (ctx->hCounters) cudaFreeHost(ctx->hCounters);
    for (cudaEvent_t e : ctx->pipeEvents) cudaEventDestroy(e);
    if (ctx->copyIn) cudaStreamDestroy(ctx->copyIn);
    if (ctx->copyOut) cudaStreamDestroy(ctx->copyOut);
    if (ctx->ownStream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// see MaterialsDev::gammaLut
__global__ void k_prepare_materials(flx_material *materials, int nMaterials, float *gammaLut)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 256) gammaLut[i] = flx_powf((float)i * FLX_ONE_255, 2.2f);
    if (i < nMaterials)
    {
        const flx_float3 kd = materials[i].Kd;
        flx_float3 g = kd;
        g.x = flx_powf(kd.x, 2.2f); g.y = flx_powf(kd.y, 2.2f); g.z = flx_powf(kd.z, 2.2f);
        materials[i].Kt = g;
    }
}

int flx_upload_scene(flx_ctx *ctx, const flx_triangle *tris, size_t n_tris, const uint32_t *indices, size_t n_indices,
                     const flx_node *nodes, size_t n_nodes, const flx_material *materials, size_t n_materials,
                     const flx_tex_desc *textures, size_t n_textures, const uint8_t *texels, size_t texel_bytes)
{
    FLX_CHECK_CTX(ctx);
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!materials || n_materials == 0) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: at least the default material is required");
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    invalidateGraph(ctx);
    ctx->sceneReady = false;
    FLX_TRY(flx_repack_scene(ctx, tris, n_tris, indices, n_indices, nodes, n_nodes));

    bool needTris160 = false;
    for (size_t i = 0; i < n_materials; i++)
    {
        const flx_material &m = materials[i];
        const int maps[3] = {m.map_Kd, m.map_Ks, m.map_N};
        for (int k = 0; k < 3; k++)
            if (maps[k] != -1 && (maps[k] < 0 || (size_t)maps[k] >= n_textures))
                return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: material references a missing texture");
        if (m.map_N != -1) needTris160 = true;
    }
    for (size_t i = 0; i < n_tris; i++)
        if (tris[i].matId < 0 || (size_t)tris[i].matId >= n_materials)
            return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: triangle references a missing material");
    for (size_t i = 0; i < n_textures; i++)
        if ((size_t)textures[i].offset + (size_t)textures[i].width * textures[i].height * 4 > texel_bytes)
            return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: texture descriptor outside texel data");
    for (size_t i = 0; i < n_textures; i++) // readTexel (flx_shade.cuh) loads a texel as one uchar4; the reference's loader packs RGBA8, offsets are multiples of 4
        if (textures[i].offset & 3u)
            return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_scene: texture offset " + std::to_string(textures[i].offset) + " is not a multiple of 4 (texels are RGBA8 words)");

    cudaFree(ctx->dMaterials); cudaFree(ctx->dTextures); cudaFree(ctx->dTexels); cudaFree(ctx->dTris160);
    ctx->dMaterials = nullptr; ctx->dTextures = nullptr; ctx->dTexels = nullptr; ctx->dTris160 = nullptr;
    FLX_CUDA(ctx, cudaMalloc(&ctx->dMaterials, n_materials * sizeof(flx_material)));
    FLX_CUDA(ctx, cudaMemcpy(ctx->dMaterials, materials, n_materials * sizeof(flx_material), cudaMemcpyHostToDevice));
    if (n_textures)
    {
        FLX_CUDA(ctx, cudaMalloc(&ctx->dTextures, n_textures * sizeof(flx_tex_desc)));
        FLX_CUDA(ctx, cudaMemcpy(ctx->dTextures, textures, n_textures * sizeof(flx_tex_desc), cudaMemcpyHostToDevice));
        FLX_CUDA(ctx, cudaMalloc(&ctx->dTexels, texel_bytes ? texel_bytes : 4));
        if (texel_bytes) FLX_CUDA(ctx, cudaMemcpy(ctx->dTexels, texels, texel_bytes, cudaMemcpyHostToDevice));
    }
    if (needTris160) // positions + uvs by triangle id, read only when a normal map is hit
    {
        FLX_CUDA(ctx, cudaMalloc(&ctx->dTris160, n_tris * sizeof(flx_triangle)));
        FLX_CUDA(ctx, cudaMemcpy(ctx->dTris160, tris, n_tris * sizeof(flx_triangle), cudaMemcpyHostToDevice));
    }
    if (!ctx->dGammaLut) FLX_CUDA(ctx, cudaMalloc(&ctx->dGammaLut, 256 * sizeof(float)));
    {
        const int n = (int)std::max<size_t>(n_materials, 256);
        k_prepare_materials<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->dMaterials, (int)n_materials, ctx->dGammaLut);
        FLX_CUDA(ctx, cudaGetLastError());
        FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    ctx->mats.gammaLut = ctx->dGammaLut;
    ctx->mats.materials = ctx->dMaterials; ctx->mats.nMaterials = (int)n_materials;
    ctx->mats.textures = ctx->dTextures; ctx->mats.nTextures = (int)n_textures;
    ctx->mats.texels = ctx->dTexels; ctx->mats.tris160 = ctx->dTris160;
    ctx->sceneReady = true;
    return FLX_OK;
}

int flx_upload_envmap(flx_ctx *ctx, const float *rgb, int width, int height, const float *prob, const int32_t *alias, const float *pdf)
{
    FLX_CHECK_CTX(ctx);
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!rgb || !prob || !alias || !pdf || width <= 0 || height <= 0) return flx_fail(ctx, FLX_ERR_ARG, "flx_upload_envmap: bad arguments");
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    invalidateGraph(ctx);
    const size_t n = (size_t)width * height;
    std::vector<float4> rgba(n); // RGB -> RGBA with alpha 1, as createEnvMap does (clcontext.cpp:490-505)
    for (size_t i = 0; i < n; i++) rgba[i] = make_float4(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 1.0f);
    cudaFree(ctx->dEnv); cudaFree(ctx->dProb); cudaFree(ctx->dPdf); cudaFree(ctx->dAlias);
    ctx->dEnv = nullptr; ctx->dProb = ctx->dPdf = nullptr; ctx->dAlias = nullptr;
    FLX_CUDA(ctx, cudaMalloc(&ctx->dEnv, n * sizeof(float4)));
    FLX_CUDA(ctx, cudaMalloc(&ctx->dProb, n * 4)); FLX_CUDA(ctx, cudaMalloc(&ctx->dPdf, n * 4)); FLX_CUDA(ctx, cudaMalloc(&ctx->dAlias, n * 4));
    FLX_CUDA(ctx, cudaMemcpy(ctx->dEnv, rgba.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
    FLX_CUDA(ctx, cudaMemcpy(ctx->dProb, prob, n * 4, cudaMemcpyHostToDevice));
    FLX_CUDA(ctx, cudaMemcpy(ctx->dPdf, pdf, n * 4, cudaMemcpyHostToDevice));
    FLX_CUDA(ctx, cudaMemcpy(ctx->dAlias, alias, n * 4, cudaMemcpyHostToDevice));
    ctx->env.rgba = ctx->dEnv; ctx->env.prob = ctx->dProb; ctx->env.pdf = ctx->dPdf; ctx->env.alias = ctx->dAlias;
    ctx->env.w = width; ctx->env.h = height;
    return FLX_OK;
}

int flx_resize(flx_ctx *ctx, uint32_t width, uint32_t height)
{
    FLX_CHECK_CTX(ctx);
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (width == 0 || height == 0) return flx_fail(ctx, FLX_ERR_ARG, "flx_resize: zero-sized frame");
    FLX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    invalidateGraph(ctx);
    const size_t np = (size_t)width * height;
    cudaFree(ctx->fbBlock); ctx->fbBlock = nullptr;
    const size_t bytes = np * (6 * sizeof(float4) + 4);
    FLX_CUDA(ctx, cudaMalloc(&ctx->fbBlock, bytes));
    FLX_CUDA(ctx, cudaMemset(ctx->fbBlock, 0, bytes));
    float4 *v = (float4 *)ctx->fbBlock;
    ctx->fb.pixels = v; ctx->fb.preview = v + np; ctx->fb.denoiserAlbedo = v + 2 * np; ctx->fb.denoiserNormal = v + 3 * np;
    ctx->fb.denoiserAlbedoOut = v + 4 * np; ctx->fb.denoiserNormalOut = v + 5 * np;
    ctx->fb.spp = (uint32_t *)(v + 6 * np);
    ctx->fb.numPixels = (uint32_t)np;
    return FLX_OK;
}

int flx_update_params(flx_ctx *ctx, const flx_render_params *params)
{
    FLX_CHECK_CTX(ctx);
    if (!params) return flx_fail(ctx, FLX_ERR_ARG, "flx_update_params: null params");
    if (params->width == 0 || params->height == 0) return flx_fail(ctx, FLX_ERR_ARG, "flx_update_params: zero-sized frame");
    if (memcmp(&ctx->params, params, sizeof(*params)) != 0) invalidateGraph(ctx);
    ctx->params = *params; // kernels receive it by value (constant bank), so no device copy is needed
    ctx->paramsSet = true;
    if (!ctx->fbBlock || ctx->fb.numPixels != params->width * params->height) return flx_resize(ctx, params->width, params->height);
    return FLX_OK;
}

int flx_set_options(flx_ctx *ctx, int denoiser_aov, uint32_t seed_offset)
{
    FLX_CHECK_CTX(ctx);
    if (ctx->denoiserAOV != denoiser_aov) invalidateGraph(ctx);
    ctx->denoiserAOV = denoiser_aov; ctx->seedOffset = seed_offset;
    return FLX_OK;
}

// Settings::clUseBitstack -> -DUSE_BITSTACK (src/settings.cpp:27, src/clcontext.cpp:145-155): which half of src/bvh.cl walks the tree
int flx_set_traversal(flx_ctx *ctx, int use_bitstack)
{
    FLX_CHECK_CTX(ctx);
    const bool want = use_bitstack != 0;
    if (want && ctx->sceneReady && ctx->treeDepth > FLX_BITSTACK_MAX_DEPTH)
        return flx_fail(ctx, FLX_ERR_ARG, "flx_set_traversal: the uploaded hierarchy is deeper than the 31 levels the bit-stack traversal can track");
    if (want != ctx->useBitstack) invalidateGraph(ctx);
    ctx->useBitstack = want;
    return FLX_OK;
}

int flx_set_pixel_sharding(flx_ctx *ctx, uint32_t rank, uint32_t world)
{
    FLX_CHECK_CTX(ctx);
    if (world == 0 || rank >= world) return flx_fail(ctx, FLX_ERR_ARG, "flx_set_pixel_sharding: rank outside world");
    if (ctx->paramsSet && world > ctx->params.height) return flx_fail(ctx, FLX_ERR_ARG, "flx_set_pixel_sharding: more ranks than image rows");
    if (ctx->tileRank != rank || ctx->tileWorld != world) invalidateGraph(ctx);
    ctx->tileRank = rank; ctx->tileWorld = world;
    return FLX_OK;
}

int flx_set_pixel_index(flx_ctx *ctx, uint32_t value)
{
    FLX_CHECK_CTX(ctx);
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    // 4-byte value travels inside the API call itself (cudaMemsetAsync cannot express it; a
    // one-thread kernel can): advance from zero.
    FLX_CUDA(ctx, cudaMemsetAsync(ctx->dPixelIdx, 0, 4, ctx->stream));
    if (value) return flx_launch_advance_pixel(ctx, 0xffffffffu, value);
    return FLX_OK;
}
int flx_reset_pixel_index(flx_ctx *ctx) { return flx_set_pixel_index(ctx, 0); }
int flx_update_pixel_index(flx_ctx *ctx, uint32_t num_pixels, uint32_t num_new_paths)
{
    FLX_CHECK_CTX(ctx);
    FLX_CUDA(ctx, cudaSetDevice(ctx->device));
    if (num_pixels == 0) return flx_fail(ctx, FLX_ERR_ARG, "flx_update_pixel_index: num_pixels == 0");
    return flx_launch_advance_pixel(ctx, num_pixels, num_new_paths);
}

} // extern "C"

// the rest of the C-ABI (stages, render loop, read-back) lives in flx_api.cu
int flx_ready(flx_ctx *ctx, bool needScene)
{
    if (!ctx->paramsSet || !ctx->fbBlock) return flx_fail(ctx, FLX_ERR_STATE, "render params not set (flx_update_params)");
    if (needScene && !ctx->sceneReady) return flx_fail(ctx, FLX_ERR_STATE, "scene not uploaded (flx_upload_scene)");
    if (needScene && ctx->params.useEnvMap && !ctx->dEnv) return flx_fail(ctx, FLX_ERR_STATE, "useEnvMap set but no environment map uploaded");
    return FLX_OK;
}
void flx_invalidate_graph(flx_ctx *ctx) { invalidateGraph(ctx); }
void flx_drain_timers(flx_ctx *ctx) { drainTimers(ctx); }
