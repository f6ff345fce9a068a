// TEST INFRASTRUCTURE — host launchers for the reference's own OpenCL kernels, compiled as C++.
//
// Each translation unit is this file built with -DTU_<name>: it includes ONE of the reference's
// kernel files (the lexically rewritten copy clpp.py puts under oracle/_ref/cl_gen/, see cl_shim.hpp)
// inside a namespace — an OpenCL program per kernel file, exactly as CLContext builds them
// (src/kernel_impl.hpp) — and exports one C launcher that runs the kernel for every work-item of the
// NDRange (global sizes as in src/clcontext.cpp:727-873).  The -D options are the ones
// CLContext::setKernelBuildSettings (src/clcontext.cpp:145-155) and the kernels'
// getAdditionalBuildOptions (src/kernel_impl.hpp:51-63,277-282,...) would pass; NVIDIA is left
// undefined (no inline PTX on the host), which selects the portable atomic_inc queue pushes.
//
// Work-items run in ascending gid order on one thread by default (deterministic queue order, the
// order the oracle restates); refcl_set_threads(n > 1) runs them under OpenMP for the timed CPU
// baseline — the kernels' atomics are real atomics (cl_shim.hpp).
#include "cl_shim.hpp"
#include "ref_cl_buffers.h"

#ifndef REFCL_NS
#error "build with -DREFCL_NS=<namespace> -DTU_<name>"
#endif

namespace REFCL_NS {
#if defined(TU_wf_reset)
#include "wf_reset.cl"
#elif defined(TU_wf_raygen)
#include "wf_raygen.cl"
#elif defined(TU_wf_logic)
#include "wf_logic.cl"
#elif defined(TU_wf_extrays)
#include "wf_extrays.cl"
#elif defined(TU_wf_shadowrays)
#include "wf_shadowrays.cl"
#elif defined(TU_wf_mat_diffuse)
#include "wf_mat_diffuse.cl"
#elif defined(TU_wf_mat_glossy)
#include "wf_mat_glossy.cl"
#elif defined(TU_wf_mat_ggx_reflection)
#include "wf_mat_ggx_reflection.cl"
#elif defined(TU_wf_mat_ggx_refraction)
#include "wf_mat_ggx_refraction.cl"
#elif defined(TU_wf_mat_delta)
#include "wf_mat_delta.cl"
#elif defined(TU_wf_mat_emissive)
#include "wf_mat_emissive.cl"
#elif defined(TU_wf_mat_all)
#include "wf_mat_all.cl"
#elif defined(TU_mk_reset)
#include "mk_reset.cl"
#elif defined(TU_mk_raygen)
#include "mk_raygen.cl"
#elif defined(TU_mk_next_vertex)
#include "mk_next_vertex.cl"
#elif defined(TU_mk_sample_bsdf)
#include "mk_sample_bsdf.cl"
#elif defined(TU_mk_splat)
#include "mk_splat.cl"
#elif defined(TU_mk_splat_preview)
#include "mk_splat_preview.cl"
#elif defined(TU_mk_postprocess)
#include "mk_postprocess.cl"
#elif defined(TU_kernel_pick)
#include "kernel_pick.cl"
#else
#error "unknown translation unit"
#endif
} // namespace

using namespace REFCL_NS;

// sizes of the shared host/device ABI as the reference's geom.h declares it
static_assert(sizeof(GPUNode) == 48 && sizeof(Triangle) == 160 && sizeof(Material) == 96, "geom.h layout");
static_assert(sizeof(Hit) == 64 && sizeof(RenderParams) == 256 && sizeof(GPUTaskState) == 256, "geom.h layout");
static_assert(sizeof(QueueCounters) == 40 && sizeof(RenderStats) == 16 && sizeof(TexDescriptor) == 12, "geom.h layout");

// NDRange of `n` work-items on `threads` host threads
#define REFCL_RANGE_T(threads, n, ...)                                             \
    do {                                                                           \
        const int64_t n_ = (int64_t)(n);                                           \
        const int t_ = (int)(threads);                                             \
        if (t_ <= 1) { for (int64_t g_ = 0; g_ < n_; g_++) { clshim_gid = (size_t)g_; __VA_ARGS__; } } \
        else {                                                                     \
            _Pragma("omp parallel for schedule(dynamic, 256) num_threads(t_)")      \
            for (int64_t g_ = 0; g_ < n_; g_++) { clshim_gid = (size_t)g_; __VA_ARGS__; }   \
        }                                                                          \
    } while (0)
#define REFCL_RANGE(n, ...) REFCL_RANGE_T(b->threads, n, __VA_ARGS__)
// 2-D NDRange (width x height), dimension 0 fastest: the microkernel reset / splat launches (src/clcontext.cpp:727-768)
#define REFCL_RANGE_2D(w, h, ...)                                                  \
    do {                                                                           \
        const int64_t w_ = (int64_t)(w), h_ = (int64_t)(h);                        \
        const int t_ = (int)(b->threads);                                          \
        if (t_ <= 1) { for (int64_t y_ = 0; y_ < h_; y_++) for (int64_t x_ = 0; x_ < w_; x_++) { clshim_gid = (size_t)x_; clshim_gid1 = (size_t)y_; __VA_ARGS__; } } \
        else {                                                                     \
            _Pragma("omp parallel for schedule(dynamic, 4) num_threads(t_)")        \
            for (int64_t y_ = 0; y_ < h_; y_++) for (int64_t x_ = 0; x_ < w_; x_++) { clshim_gid = (size_t)x_; clshim_gid1 = (size_t)y_; __VA_ARGS__; } \
        }                                                                          \
        clshim_gid1 = 0;                                                           \
    } while (0)

#define TASKS(b) ((GPUTaskState *)(b)->tasks)
#define COUNTERS(b) ((QueueCounters *)(b)->queueCounters)
#define PARAMS(b) ((RenderParams *)(b)->params)
#define TRIS(b) ((Triangle *)(b)->tris)
#define NODES(b) ((GPUNode *)(b)->nodes)
#define MATS(b) ((Material *)(b)->materials)
#define TEXS(b) ((TexDescriptor *)(b)->textures)
#define STATS(b) ((RenderStats *)(b)->stats)
#define Q(b, i) ((b)->queues[i])
enum { Q_RAYGEN = 0, Q_EXT, Q_SHADOW, Q_DIFFUSE, Q_GLOSSY, Q_GGXREFL, Q_GGXREFR, Q_DELTA, Q_EMISSIVE };

static inline uint32_t maxu(uint32_t a, uint32_t b) { return a > b ? a : b; }
static inline uint32_t minu(uint32_t a, uint32_t b) { return a < b ? a : b; }

extern "C" {

#if defined(TU_wf_reset)
// CLContext::enqueueWfResetKernel, src/clcontext.cpp:783-788: global size max(NUM_TASKS, w*h)
void refcl_wf_reset(const refcl_buffers *b)
{
    REFCL_RANGE(maxu(b->numTasks, b->numPixels),
                reset(TASKS(b), b->pixels, b->denoiserAlbedo, b->denoiserNormal, COUNTERS(b), Q(b, Q_RAYGEN), PARAMS(b),
                      b->samplesPerPixel, b->numTasks));
}
#elif defined(TU_wf_raygen)
// enqueueWfRaygenKernel, src/clcontext.cpp:790-794
void refcl_wf_raygen(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, genRays(TASKS(b), PARAMS(b), COUNTERS(b), Q(b, Q_RAYGEN), Q(b, Q_EXT), b->currPixelIdx, b->numTasks));
}
#elif defined(TU_wf_logic)
// enqueueWfLogicKernel, src/clcontext.cpp:808-814: global size rounded up to 32
void refcl_wf_logic(const refcl_buffers *b, uint32_t firstIteration)
{
    const clshim_image2d env = {b->envRGBA, b->envW, b->envH};
    REFCL_RANGE((b->numTasks + 31u) & ~31u,
                logic(TASKS(b), b->pixels, b->denoiserNormal, b->denoiserAlbedo, COUNTERS(b), Q(b, Q_EXT), Q(b, Q_SHADOW),
                      Q(b, Q_RAYGEN), Q(b, Q_DIFFUSE), Q(b, Q_GLOSSY), Q(b, Q_GGXREFL), Q(b, Q_GGXREFR), Q(b, Q_DELTA),
                      Q(b, Q_EMISSIVE), TRIS(b), NODES(b), b->indices, &env, b->probTable, b->aliasTable, b->pdfTable, MATS(b),
                      b->texData, TEXS(b), PARAMS(b), b->samplesPerPixel, b->numTasks, firstIteration));
}
#elif defined(TU_wf_extrays)
// enqueueWfExtRayKernel, src/clcontext.cpp:796-800
void refcl_wf_extrays(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, traceExtension(TASKS(b), COUNTERS(b), Q(b, Q_EXT), TRIS(b), NODES(b), b->indices, PARAMS(b), b->numTasks));
}
// Ray batches through the reference's bvh_intersect (the function traceExtension calls, src/wf_extrays.cl:27;
// the reference has no batch kernel of its own).  rays: n x (ox,oy,oz,tmax, dx,dy,dz,pad).
void refcl_trace_closest(const void *tris, const void *nodes, const uint32_t *indices, const float *rays, uint64_t n,
                         int32_t *hit_i, float *hit_t, void *hits64, int threads)
{
    REFCL_RANGE_T(threads, n, {
        const float *p = rays + 8 * clshim_gid;
        Ray r = {float3(p[0], p[1], p[2]), float3(p[4], p[5], p[6])};
        Hit hit = EMPTY_HIT(p[3]);
        bvh_intersect(&r, &hit, (Triangle *)tris, (GPUNode *)nodes, (uint *)indices);
        if (hit_i) hit_i[clshim_gid] = hit.i;
        if (hit_t) hit_t[clshim_gid] = hit.t;
        if (hits64) ((Hit *)hits64)[clshim_gid] = hit;
    });
}
// The reference's own Wang-hash generator (src/random.cl:7-22): `count` successive rand(&seed) draws.
void refcl_rand_sequence(uint32_t seed, uint32_t count, float *out, uint32_t *seed_out)
{
    for (uint32_t k = 0; k < count; k++) out[k] = rand(&seed);
    if (seed_out) *seed_out = seed;
}
// intersectLight (src/intersect.cl:124-155) on a ray batch: hit.t starts at tmax
void refcl_intersect_light(const void *params, const float *rays, uint64_t n, void *hits64)
{
    REFCL_RANGE_T(1, n, {
        const float *p = rays + 8 * clshim_gid;
        Ray r = {float3(p[0], p[1], p[2]), float3(p[4], p[5], p[6])};
        Hit hit = EMPTY_HIT(p[3]);
        intersectLight(&hit, &r, (RenderParams *)params);
        ((Hit *)hits64)[clshim_gid] = hit;
    });
}
#elif defined(TU_wf_shadowrays)
// enqueueWfShadowRayKernel, src/clcontext.cpp:802-806
void refcl_wf_shadowrays(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, traceShadow(TASKS(b), COUNTERS(b), Q(b, Q_SHADOW), TRIS(b), NODES(b), b->indices, PARAMS(b), b->numTasks));
}
// Ray batches through the reference's bvh_occluded (src/wf_shadowrays.cl:33), tmax in rays[3]
void refcl_trace_any(const void *tris, const void *nodes, const uint32_t *indices, const float *rays, uint64_t n, uint8_t *occluded, int threads)
{
    REFCL_RANGE_T(threads, n, {
        const float *p = rays + 8 * clshim_gid;
        Ray r = {float3(p[0], p[1], p[2]), float3(p[4], p[5], p[6])};
        float maxDist = p[3];
        occluded[clshim_gid] = bvh_occluded(&r, &maxDist, (Triangle *)tris, (GPUNode *)nodes, (uint *)indices) ? 1 : 0;
    });
}
#elif defined(TU_wf_mat_diffuse)
// enqueueWfMaterialKernels, src/clcontext.cpp:816-873: one launch per material queue, global size NUM_TASKS
void refcl_wf_mat_diffuse(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontDiffuse(TASKS(b), COUNTERS(b), Q(b, Q_DIFFUSE), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_glossy)
void refcl_wf_mat_glossy(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontGlossy(TASKS(b), COUNTERS(b), Q(b, Q_GLOSSY), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_ggx_reflection)
void refcl_wf_mat_ggx_reflection(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontGGXReflection(TASKS(b), COUNTERS(b), Q(b, Q_GGXREFL), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_ggx_refraction)
void refcl_wf_mat_ggx_refraction(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontGGXRefraction(TASKS(b), COUNTERS(b), Q(b, Q_GGXREFR), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_delta)
void refcl_wf_mat_delta(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontDelta(TASKS(b), COUNTERS(b), Q(b, Q_DELTA), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_emissive)
void refcl_wf_mat_emissive(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontEmissive(TASKS(b), COUNTERS(b), Q(b, Q_EMISSIVE), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_wf_mat_all)
// single-queue mode: all paths wait in the 'diffuse' queue (src/clcontext.cpp:869-873, kernel_impl.hpp:255-283)
void refcl_wf_mat_all(const refcl_buffers *b)
{
    REFCL_RANGE(b->numTasks, wavefrontAllMaterials(TASKS(b), COUNTERS(b), Q(b, Q_DIFFUSE), Q(b, Q_EXT), MATS(b), b->texData, TEXS(b), PARAMS(b), b->numTasks));
}
#elif defined(TU_mk_reset)
// microkernel renderer launches, src/clcontext.cpp:727-768: reset / splat / splatPreview are 2-D (w x h), the rest 1-D (NUM_TASKS)
void refcl_mk_reset(const refcl_buffers *b)
{
    REFCL_RANGE_2D(PARAMS(b)->width, PARAMS(b)->height,
                   reset(TASKS(b), b->pixels, b->denoiserAlbedo, b->denoiserNormal, PARAMS(b), b->samplesPerPixel, b->numTasks));
}
#elif defined(TU_mk_raygen)
void refcl_mk_raygen(const refcl_buffers *b) { REFCL_RANGE(b->numTasks, genCameraRays(TASKS(b), PARAMS(b), b->numTasks)); }
#elif defined(TU_mk_next_vertex)
void refcl_mk_next_vertex(const refcl_buffers *b)
{
    const clshim_image2d env = {b->envRGBA, b->envW, b->envH};
    REFCL_RANGE(b->numTasks, nextVertex(TASKS(b), MATS(b), b->texData, TEXS(b), b->denoiserNormal, TRIS(b), NODES(b), b->indices,
                                        PARAMS(b), STATS(b), &env, b->pdfTable, b->numTasks));
}
#elif defined(TU_mk_sample_bsdf)
void refcl_mk_sample_bsdf(const refcl_buffers *b)
{
    const clshim_image2d env = {b->envRGBA, b->envW, b->envH};
    REFCL_RANGE(b->numTasks, sampleBsdf(TASKS(b), b->denoiserAlbedo, MATS(b), b->texData, TEXS(b), &env, b->probTable, b->aliasTable,
                                        b->pdfTable, TRIS(b), NODES(b), b->indices, PARAMS(b), STATS(b), b->numTasks));
}
#elif defined(TU_mk_splat)
void refcl_mk_splat(const refcl_buffers *b)
{
    REFCL_RANGE_2D(PARAMS(b)->width, PARAMS(b)->height, splat(TASKS(b), b->pixels, PARAMS(b), b->samplesPerPixel, STATS(b), b->numTasks));
}
#elif defined(TU_mk_splat_preview)
void refcl_mk_splat_preview(const refcl_buffers *b)
{
    REFCL_RANGE_2D(PARAMS(b)->width, PARAMS(b)->height, splatPreview(TASKS(b), b->pixels, PARAMS(b), b->numTasks));
}
#elif defined(TU_mk_postprocess)
// enqueuePostprocessKernel, src/clcontext.cpp:770-781: global size w*h
void refcl_mk_postprocess(const refcl_buffers *b)
{
    REFCL_RANGE(b->numPixels, process(b->pixels, b->denoiserAlbedo, b->denoiserNormal, b->pixelsPreview, b->denoiserAlbedoGL,
                                      b->denoiserNormalGL, PARAMS(b), b->numTasks));
}
#elif defined(TU_kernel_pick)
// CLContext::pickSingle, src/clcontext.cpp:934-949: one work-item
void refcl_kernel_pick(const refcl_buffers *b, float NDCx, float NDCy)
{
    REFCL_RANGE(1, pick(PARAMS(b), TRIS(b), NODES(b), b->indices, (Hit *)b->pickResult, NDCx, NDCy));
}
#endif

} // extern "C"
