// Internal definitions shared by the .cu translation units of libfluctus_cuda.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/fluctus_cuda.h"
#include "flx_bvh.cuh"

// ---- path state in HBM ------------------------------------------------------------------------
// The reference keeps GPUTaskState as 64 word-slots of SoA (256 B/path, src/geom.h:222-259) of
// which 12 are float3 padding.  Here the 52 live words are grouped by the stage that writes them
// into twelve float4 streams + four scalar streams (208 B/path), so every stage moves its fields
// with 16-byte coalesced accesses and no stage has to read a field just to write it back:
//
//   ray0  orig.xyz       | lastPdfW          written by raygen, material      read by ext, logic
//   ray1  dir.xyz        | pathLen (u32)     written by raygen, material; ext bumps .w
//   hit0  P.xyz          | t                 written by ext (raygen/reset: empty hit)
//   hit1  N.xyz          | hit.i (i32)       written by ext; logic rewrites N (shading normal/flip)
//   hit2  uv.xy | matId (i32) | areaLightHit (i32)                     written by ext
//   thr   T.xyz          | seed (u32)        written by raygen, logic (RR, rng), material
//   acc   Ei.xyz         | pixelIndex (u32)  written by raygen, logic
//   shw0  shadowOrig.xyz | shadowRayLen      written by logic (NEE)            read by shadow
//   shw1  shadowDir.xyz  | lastCosTh         written by logic                  read by shadow, material
//   nee0  lastEmission   | lastPdfDirect     written by logic
//   nee1  lastBsdf       | lastPdfImplicit   written by material
//   nee2  lastT.xyz      | (spare)           written by material
//   scalars: lastSpecular (material), shadowBlocked (shadow, logic), logicFlags = backface |
//            firstDiffuse<<1 (logic), lastPickProb (logic)
// Every field is reached as st.field[gid] through a strided view, so the layout is a build-time choice:
//   FLX_STATE_AOS = 0 (default): one stream per field (stride = sizeof(field)), coalesced for the dense stages;
//   FLX_STATE_AOS = 1 (experiment): one 256-byte record per path — {ray0 ray1 thr acc | hit0 hit1 hit2
//   scalars | shw0 shw1 nee0 nee1 | nee2 phase} — meant to let the stages that GATHER by sparse path id (material queues,
//   ray queues) use whole DRAM atoms; measured slower in every stage (logic +41 %, materials +14 %, traversal +3 %).
#ifndef FLX_STATE_AOS
#define FLX_STATE_AOS 0
#endif
#define FLX_STATE_RECORD_BYTES 256
template <class T>
struct StateView
{
    char *base;
    __host__ __device__ __forceinline__ T &operator[](size_t i) const
    {
        return *(T *)(base + i * (size_t)(FLX_STATE_AOS ? FLX_STATE_RECORD_BYTES : sizeof(T)));
    }
};
struct PathState
{
    StateView<float4> ray0, ray1, hit0, hit1, hit2, thr, acc, shw0, shw1, nee0, nee1, nee2;
    StateView<uint32_t> lastSpecular, shadowBlocked, logicFlags;
    StateView<float> lastPickProb;
    StateView<int32_t> phase; // microkernel renderer only (PathPhase, geom.h:205-216)
};
// PathPhase (geom.h:205-216) and the values the microkernel stages park in shadowBlocked while a shadow ray is out
enum { FLX_MK_RT_NEXT_VERTEX = 0, FLX_MK_SAMPLE_BSDF = 1, FLX_MK_SPLAT_SAMPLE = 4, FLX_MK_GENERATE_CAMERA_RAY = 5, FLX_MK_DONE = 6 };
enum { FLX_MK_NEE_PENDING = 2u, FLX_MK_NEE_NONE = 3u };
#define FLX_STATE_VEC_STREAMS 12
#define FLX_STATE_SCALAR_STREAMS 5

struct Queues
{
    uint32_t *q[FLX_NUM_QUEUES]; // N entries each (reference src/clcontext.cpp:130-138); one allocation: q[i] = q[0] + i * N
    flx_queue_counters *counters; // device copy of QueueCounters
};

struct EnvMapDev
{
    const float4 *rgba; // w*h texels, alpha = 1 (what createEnvMap uploads)
    const float *prob; const int *alias; const float *pdf;
    int w, h;
};

struct MaterialsDev
{
    const flx_material *materials; int nMaterials;
    const flx_tex_desc *textures; int nTextures;
    const uint8_t *texels;
    const flx_triangle *tris160; // original triangles: only read for normal-mapped hits (utils.cl:184-190)
    // albedo gamma (utils.cl:161-166) taken out of the per-path work: gammaLut[k] = powf((float)k * (1/255), 2.2f) for
    // the 256 values an RGBA8 texel can take, and the DEVICE copy of every material carries powf(Kd, 2.2f) in its Kt
    // slot (Kt is not read by any kernel).  Both are computed on the device with the powf the shading kernels would
    // have called on the same input bits, so shaded values are bit-identical to evaluating it per path.
    const float *gammaLut;
};

struct FrameBuffers
{
    float4 *pixels, *preview, *denoiserAlbedo, *denoiserNormal, *denoiserAlbedoOut, *denoiserNormalOut;
    uint32_t *spp;
    uint32_t numPixels;
};

// Device-side bookkeeping that the reference does on the host between iterations
// (tracer.cpp:761-773 statistics, clcontext.cpp:917-921 pixel cursor).
struct RenderTotals
{
    unsigned long long primary, extension, shadow, samples, splatted, iterations;
};

// Closes one iteration on the device: queue lengths -> statistics, pixel cursor advanced by the number of
// regenerated paths, counters cleared.  Called by ONE thread (k_bookkeeping, or the last CTA of k_ext_shadow).
#ifdef __CUDACC__
static __device__ __forceinline__ void closeIteration(flx_queue_counters *counters, RenderTotals *totals, uint32_t *pixelIdx, uint32_t numPixels)
{
    const flx_queue_counters c = *counters;
    totals->primary += c.raygenQueue;
    totals->extension += c.extensionQueue;
    totals->shadow += c.shadowQueue;
    totals->samples += (totals->iterations > 0) ? c.raygenQueue : 0u;
    totals->splatted += c.splattedSamples;
    totals->iterations += 1;
    *pixelIdx = (*pixelIdx + c.raygenQueue) % numPixels;
    flx_queue_counters z = {};
    *counters = z;
}
#endif

struct StageTimer
{
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
    double ms = 0.0; uint64_t launches = 0;
};

struct flx_ctx
{
    int device = 0;
    cudaStream_t stream = nullptr; bool ownStream = false;
    uint32_t N = 0;
    int numSMs = 148;
    std::string err;
    uint64_t launches = 0;

    // scene
    DevBVH bvh{}; float4 *dNodes = nullptr, *dSlots = nullptr, *dShade = nullptr;
    size_t nInner = 0, nSlots = 0, nTris = 0;
    MaterialsDev mats{}; flx_material *dMaterials = nullptr; flx_tex_desc *dTextures = nullptr; uint8_t *dTexels = nullptr;
    flx_triangle *dTris160 = nullptr;
    float *dGammaLut = nullptr;
    bool sceneReady = false;
    bool useBitstack = false; uint32_t treeDepth = 0; // Settings::clUseBitstack (flx_set_traversal); depth of the uploaded hierarchy
    bool traceSetupDone = false; // cudaFuncSetAttribute(MaxDynamicSharedMemorySize) done on this device
    EnvMapDev env{}; float4 *dEnv = nullptr; float *dProb = nullptr, *dPdf = nullptr; int *dAlias = nullptr;

    // wavefront
    void *stateBlock = nullptr; PathState st{};
    void *queueBlock = nullptr; Queues qs{};
    uint32_t *dPixelIdx = nullptr;
    RenderTotals *dTotals = nullptr;
    flx_render_stats *dStats = nullptr; // device RenderStats of the microkernel renderer
    FrameBuffers fb{}; void *fbBlock = nullptr;
    flx_render_params params{}; bool paramsSet = false;
    int denoiserAOV = 0; uint32_t seedOffset = 0;
    uint32_t tileRank = 0, tileWorld = 1; // preview sharding: rows y % tileWorld == tileRank belong to this context
    void *comm = nullptr; int commRank = 0, commWorld = 1; // ncclComm_t of flx_comm_init (flx_comm.cu), one process per GPU
    flx_queue_counters *hCounters = nullptr; // pinned staging for async counter reads
    std::vector<flx_queue_counters *> pendingCounterReads;

    // graph for flx_render
    cudaGraphExec_t graphExec = nullptr; bool graphValid = false; uint64_t graphKernels = 0;
    // graph for flx_mk_render: one sample per pixel (raygen, (nextVertex, sampleBsdf) x (maxBounces + 1), splat)
    cudaGraphExec_t mkGraphExec = nullptr; bool mkGraphValid = false; uint64_t mkGraphKernels = 0;

    // timing
    bool timing = false; StageTimer timers[6];
    // visit counters of the COUNT=true trace kernels: 4 groups (ext, shadow, batch closest, batch any)
    // x {inner, leaf, tri, hits, rays}
    bool counting = false; unsigned long long *dTraceStats = nullptr;

    // scratch for the host-buffer ray-cast entry points; two copy streams + one event pair per chunk so that the
    // upload of chunk c+1, the traversal of chunk c and the download of chunk c-1 overlap (flx_api.cu: RayPipe)
    void *dScratch = nullptr; size_t scratchBytes = 0;
    cudaStream_t copyIn = nullptr, copyOut = nullptr;
    std::vector<cudaEvent_t> pipeEvents;
};

// error helpers -----------------------------------------------------------------------------------
int flx_fail(flx_ctx *ctx, int code, const std::string &msg);
#define FLX_CUDA(ctx, call)                                                                         \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            return flx_fail(ctx, FLX_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

// launch geometry: persistent grid-stride kernels sized to the machine (148 SMs on B200)
static inline int flx_grid(const flx_ctx *ctx, int blocksPerSM) { return ctx->numSMs * blocksPerSM; }

// implemented in flx_trace.cu
int flx_repack_scene(flx_ctx *ctx, const flx_triangle *tris, size_t n_tris, const uint32_t *indices, size_t n_indices,
                     const flx_node *nodes, size_t n_nodes);
int flx_launch_ext(flx_ctx *ctx);
int flx_launch_shadow(flx_ctx *ctx);
int flx_launch_ext_shadow(flx_ctx *ctx); // both in one persistent launch + the iteration's bookkeeping (render entry)
int flx_reset_trace_cursors(flx_ctx *ctx);
int flx_launch_trace_batch(flx_ctx *ctx, const float4 *dRays, size_t n, int32_t *dHitI, float *dHitT, flx_hit *dHits);
int flx_launch_any_batch(flx_ctx *ctx, const float4 *dRays, size_t n, uint8_t *dOcc);
int flx_launch_pick(flx_ctx *ctx, float ndcX, float ndcY, flx_hit *dHit);

int flx_launch_mk_vertex_trace(flx_ctx *ctx, uint32_t total);
int flx_launch_mk_shadow_trace(flx_ctx *ctx, uint32_t total, int lightOccludes);

// implemented in flx_microkernel.cu (src/mk_*.cl)
int flx_launch_mk_reset(flx_ctx *ctx);
int flx_launch_mk_raygen(flx_ctx *ctx);
int flx_launch_mk_next_vertex(flx_ctx *ctx);
int flx_launch_mk_sample_bsdf(flx_ctx *ctx);
int flx_launch_mk_splat(flx_ctx *ctx);
int flx_launch_mk_splat_preview(flx_ctx *ctx);

// implemented in flx_wavefront.cu
int flx_launch_reset(flx_ctx *ctx);
int flx_launch_raygen(flx_ctx *ctx);
int flx_launch_logic(flx_ctx *ctx, int firstIteration);
int flx_launch_logic_regen(flx_ctx *ctx);
int flx_launch_materials(flx_ctx *ctx, bool renderEntry = false); // renderEntry: flx_render's iteration (always per-BxDF queues)
int flx_launch_postprocess(flx_ctx *ctx);
int flx_launch_bookkeeping(flx_ctx *ctx); // counters -> totals, pixel cursor advance, clear
int flx_launch_advance_pixel(flx_ctx *ctx, uint32_t numPixels, uint32_t numNew);
int flx_launch_export_state(flx_ctx *ctx, uint32_t *dSlots64xN);
int flx_launch_import_state(flx_ctx *ctx, const uint32_t *dSlots64xN);

// pixels this context generates paths for (all of them unless preview sharding is on)
static inline uint32_t flx_owned_pixels(const flx_ctx *ctx)
{
    if (ctx->tileWorld <= 1u) return ctx->fb.numPixels;
    return ((ctx->params.height - ctx->tileRank + ctx->tileWorld - 1u) / ctx->tileWorld) * ctx->params.width;
}

// implemented in flx_context.cu
int flx_ready(flx_ctx *ctx, bool needScene);
void flx_invalidate_graph(flx_ctx *ctx);
void flx_drain_timers(flx_ctx *ctx);
