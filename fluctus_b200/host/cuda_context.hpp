// CudaContext — C++ host-side drop-in for the reference's `class CLContext`
// (reference src/clcontext.hpp:26-215) on top of the C-ABI of include/fluctus_cuda.h.
//
// Method names, argument meaning and error behaviour follow the reference so that the render loops
// of Tracer (src/tracer.cpp:250-294 update, :459-519 / :730-796 benchmark loops) compile against
// either class: every enqueue* call is asynchronous on one in-order queue, finishQueue() is the only
// wait, counters requested with enqueueGetCounters are valid after the next finishQueue, and any
// failure throws std::runtime_error with the backend's message (CLContext::verify -> clt::check,
// src/clcontext.cpp:952-957).  The GL-coupled pieces (setup(PTWindow*), setupPixelStorage(PTWindow*),
// PBO acquire/release) become plain device buffers; recompileKernels() is a no-op because every
// specialisation is compiled ahead of time.
//
// Header-only, C++14, no CUDA headers needed: link with -lfluctus_cuda.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/fluctus_cuda.h"

namespace flx {

// The reference's struct names for the bytes that cross the boundary (src/geom.h); layouts are identical
// (static_asserts below), so code written against geom.h can pass its own structs by reinterpret_cast.
using RenderParams = flx_render_params;
using QueueCounters = flx_queue_counters;
using RenderStats = flx_render_stats;
using Hit = flx_hit;
static_assert(sizeof(flx_triangle) == 160 && sizeof(flx_node) == 48 && sizeof(flx_material) == 96, "scene ABI");
static_assert(sizeof(flx_render_params) == 256 && sizeof(flx_queue_counters) == 40 && sizeof(flx_hit) == 64, "param ABI");

struct PerfNumbers { float primary, extension, shadow, samples, total; }; // src/clcontext.hpp:19-24 (MRays/s)

class CudaContext
{
public:
    // CLContext::CLContext + setup (src/clcontext.cpp:18-69): NUM_TASKS = settings.wfBufferSize
    explicit CudaContext(uint32_t numTasks = 1u << 20, int device = 0, void *cudaStream = nullptr)
    {
        if (flx_create(&ctx, device, numTasks, cudaStream) != FLX_OK) throw std::runtime_error(flx_last_error(nullptr));
    }
    ~CudaContext() { flx_destroy(ctx); }
    CudaContext(const CudaContext &) = delete;
    CudaContext &operator=(const CudaContext &) = delete;

    uint32_t getNumTasks() const { return flx_num_tasks(ctx); }
    flx_ctx *handle() const { return ctx; }

    // ---- uploads ---------------------------------------------------------------------------------
    // CLContext::uploadSceneData(BVH*, Scene*) (src/clcontext.cpp:540-584) + packTextures (:588-629): the
    // caller passes the arrays those objects hold (RTTriangle[], index list, Node[], Material[], textures).
    void uploadSceneData(const flx_triangle *tris, size_t nTris, const uint32_t *indices, size_t nIndices,
                         const flx_node *nodes, size_t nNodes, const flx_material *materials, size_t nMaterials,
                         const flx_tex_desc *textures = nullptr, size_t nTextures = 0, const uint8_t *texels = nullptr, size_t texelBytes = 0)
    {
        check(flx_upload_scene(ctx, tris, nTris, indices, nIndices, nodes, nNodes, materials, nMaterials, textures, nTextures, texels, texelBytes));
    }
    // CLContext::createEnvMap(EnvironmentMap*) (src/clcontext.cpp:485-529)
    void createEnvMap(const float *rgb, int width, int height, const float *probTable, const int32_t *aliasTable, const float *pdfTable)
    {
        check(flx_upload_envmap(ctx, rgb, width, height, probTable, aliasTable, pdfTable));
    }
    void setupPixelStorage(uint32_t width, uint32_t height) { check(flx_resize(ctx, width, height)); }
    void updateParams(const RenderParams &params) { check(flx_update_params(ctx, &params)); lastParams = params; }
    void recompileKernels(bool /*setArgs*/) {}
    void setPixelSharding(uint32_t rank, uint32_t world) { check(flx_set_pixel_sharding(ctx, rank, world)); }
    void setTraversal(bool useBitstack) { check(flx_set_traversal(ctx, useBitstack ? 1 : 0)); } // Settings::clUseBitstack
    void setOptions(bool denoiserAOV, uint32_t seedOffset) { check(flx_set_options(ctx, denoiserAOV ? 1 : 0, seedOffset)); }

    // ---- pixel cursor (src/clcontext.cpp:917-927) -------------------------------------------------
    void resetPixelIndex() { check(flx_reset_pixel_index(ctx)); }
    void setPixelIndex(uint32_t v) { check(flx_set_pixel_index(ctx, v)); }
    void updatePixelIndex(uint32_t numPixels, uint32_t numNewPaths) { check(flx_update_pixel_index(ctx, numPixels, numNewPaths)); }

    // ---- wavefront stages (src/clcontext.cpp:783-873); params kept for signature parity ------------
    void enqueueWfResetKernel(const RenderParams &) { check(flx_wf_reset(ctx)); }
    void enqueueWfRaygenKernel(const RenderParams &) { check(flx_wf_raygen(ctx)); }
    void enqueueWfExtRayKernel(const RenderParams &) { check(flx_wf_ext(ctx)); }
    void enqueueWfShadowRayKernel(const RenderParams &) { check(flx_wf_shadow(ctx)); }
    void enqueueWfLogicKernel(const RenderParams &, const bool firstIteration) { check(flx_wf_logic(ctx, firstIteration ? 1 : 0)); }
    void enqueueWfMaterialKernels(const RenderParams &) { check(flx_wf_materials(ctx)); }
    void enqueuePostprocessKernel(const RenderParams &) { check(flx_postprocess(ctx)); }
    void enqueueClearWfQueues() { check(flx_wf_clear_queues(ctx)); }
    void enqueueGetCounters(QueueCounters *cnt) { check(flx_get_counters(ctx, cnt)); }
    void finishQueue() { check(flx_finish(ctx)); }

    // ---- microkernel renderer (src/clcontext.cpp:727-768 -> mk_*.cl): one task per pixel, exact spp ---------
    void enqueueResetKernel(const RenderParams &) { check(flx_mk_reset(ctx)); }
    void enqueueRayGenKernel(const RenderParams &) { check(flx_mk_raygen(ctx)); }
    void enqueueNextVertexKernel(const RenderParams &) { check(flx_mk_next_vertex(ctx)); }
    void enqueueBsdfSampleKernel(const RenderParams &) { check(flx_mk_sample_bsdf(ctx)); }
    void enqueueSplatKernel(const RenderParams &) { check(flx_mk_splat(ctx)); }
    void enqueueSplatPreviewKernel(const RenderParams &) { check(flx_mk_splat_preview(ctx)); }
    // Tracer::renderSingle's loop (src/tracer.cpp:130-159) from a CUDA graph
    void renderSingle(uint32_t spp, bool reset = true) { check(flx_mk_render(ctx, spp, reset ? 1 : 0)); }
    // CLContext::fetchStatsAsync (src/clcontext.cpp:660-664): the device RenderStats of the MK kernels
    void fetchStats() { check(flx_mk_get_stats(ctx, &statsAsync)); }

    // ---- headless render entry (new): n passes of the benchmark inner loop from a CUDA graph ---------
    void render(uint32_t iterations, bool reset = false) { check(flx_render(ctx, iterations, reset ? 1 : 0)); }
    struct Totals { uint64_t primary, extension, shadow, samples, splatted, iterations; };
    Totals getRenderTotals()
    {
        uint64_t t[6];
        check(flx_get_render_totals(ctx, t));
        return Totals{t[0], t[1], t[2], t[3], t[4], t[5]};
    }

    // ---- statistics: CLContext::statsAsync / getStats / resetStats / updateRenderPerf (src/clcontext.cpp:631-684)
    RenderStats statsAsync{0, 0, 0, 0}; // written directly by Tracer (src/tracer.cpp:364-367)
    const RenderStats getStats() const { return statsAsync; }
    void resetStats() { statsAsync = RenderStats{0, 0, 0, 0}; }
    void updateRenderPerf(float deltaT)
    {
        const float s = 1e-6f / deltaT;
        perf = PerfNumbers{statsAsync.primaryRays * s, statsAsync.extensionRays * s, statsAsync.shadowRays * s, statsAsync.samples * s,
                           (statsAsync.primaryRays + statsAsync.extensionRays + statsAsync.shadowRays) * s};
        resetStats();
    }
    const PerfNumbers getRenderPerf() const { return perf; }

    // ---- results ---------------------------------------------------------------------------------
    Hit pickSingle(float NDCx, float NDCy) { Hit h; check(flx_pick(ctx, NDCx, NDCy, &h)); return h; }
    // raw accumulation (sum R, sum G, sum B, #samples) or the tonemapped preview, w*h float4 each
    std::vector<float> readPixels(int which = FLX_BUF_PIXELS)
    {
        std::vector<float> px((size_t)lastParams.width * lastParams.height * 4);
        check(flx_read_pixels(ctx, which, px.data(), px.size() * sizeof(float)));
        return px;
    }
    void writePixels(int which, const std::vector<float> &px) { check(flx_write_pixels(ctx, which, px.data(), px.size() * sizeof(float))); }
    void *devicePtr(int which = FLX_BUF_PIXELS) { return flx_device_ptr(ctx, which); }

    // ---- ray casting -------------------------------------------------------------------------------
    void traceClosest(const float *rays8, size_t n, int32_t *hitI, float *hitT, Hit *hits = nullptr) { check(flx_trace_closest(ctx, rays8, n, hitI, hitT, hits)); }
    void traceAny(const float *rays8, size_t n, uint8_t *occluded) { check(flx_trace_any(ctx, rays8, n, occluded)); }

    uint64_t launchCount() const { return flx_launch_count(ctx); }

    // ---- multi-GPU (new; the reference is single-device): contexts on different GPUs rendered sample shards of one
    // frame; afterwards contexts[0] holds the summed accumulation buffer and the others are zero (flx_reduce_frames)
    static void reduceFrames(const std::vector<CudaContext *> &contexts)
    {
        std::vector<flx_ctx *> h;
        for (CudaContext *c : contexts) h.push_back(c->ctx);
        if (flx_reduce_frames(h.data(), (int)h.size()) != FLX_OK)
        {
            std::string msg;
            for (flx_ctx *c : h) { const char *m = flx_last_error(c); if (m && *m) { msg = m; break; } }
            if (msg.empty()) msg = flx_last_error(nullptr);
            throw std::runtime_error(msg);
        }
    }

    // ---- multi-GPU with one process per GPU (new): NCCL over NVLink from the C-ABI (flx_comm.cu) ----------------------
    // rank 0 draws the id and hands its 128 bytes to the others; commInit is collective
    static std::vector<unsigned char> commUniqueId()
    {
        std::vector<unsigned char> id(FLX_COMM_ID_BYTES);
        if (flx_comm_unique_id(id.data()) != FLX_OK) throw std::runtime_error("flx_comm_unique_id failed (is libnccl.so.2 on the library path?)");
        return id;
    }
    void commInit(int rank, int world, const std::vector<unsigned char> &id)
    {
        if (id.size() != FLX_COMM_ID_BYTES) throw std::runtime_error("commInit: the id must have 128 bytes");
        check(flx_comm_init(ctx, rank, world, id.data()));
    }
    void allreducePixels() { check(flx_allreduce_pixels(ctx)); }   // every rank ends with the summed frame
    void reducePixels(int root = 0) { check(flx_reduce_pixels(ctx, root)); } // only `root` does
    int ncclVersion() { int v = 0; check(flx_comm_info(ctx, nullptr, nullptr, &v)); return v; }

    // ---- hierarchy build on the GPU (new): stands in for BVH / SBVH construction (src/bvh.cpp, src/sbvh.cpp) -----
    // Returns Node[] (depth-first, left child at index + 1) and fills `indices`: the arrays uploadSceneData takes.
    static std::vector<flx_node> buildHierarchy(const flx_triangle *tris, size_t nTris, std::vector<uint32_t> &indices,
                                                uint32_t maxLeaf = 4, int device = 0, uint32_t *depth = nullptr)
    {
        std::vector<flx_node> nodes(nTris ? 2 * nTris - 1 : 0);
        indices.assign(nTris, 0u);
        size_t nNodes = 0;
        const int rc = flx_build_lbvh(device, tris, nTris, maxLeaf, nodes.data(), nodes.size(), &nNodes, indices.data(), depth);
        if (rc != FLX_OK) throw std::runtime_error(std::string(flx_last_error(nullptr)) + " (flx_status " + std::to_string(rc) + ")");
        nodes.resize(nNodes);
        return nodes;
    }

private:
    void check(int rc) const
    {
        if (rc != FLX_OK) throw std::runtime_error(std::string(flx_last_error(ctx)) + " (flx_status " + std::to_string(rc) + ")");
    }
    flx_ctx *ctx = nullptr;
    RenderParams lastParams{};
    PerfNumbers perf{0, 0, 0, 0, 0};
};

} // namespace flx
