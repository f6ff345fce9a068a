// TEST INFRASTRUCTURE — CPU oracle.  Wavefront stages restated from reference src/wf_*.cl over
// the reference's own SoA path-state layout (slot * numTasks + gid, src/geom.h:43-67).
#pragma once
#include "o_shade.h"
#include <vector>

namespace orc {

struct Wavefront
{
    uint32_t N = 0;                 // NUM_TASKS
    flx_render_params params;       // device copy of RenderParams
    SceneView geo;
    ShadeScene shade;
    bool useBitstack = false;       // -DUSE_BITSTACK
    bool denoiserAOV = false;       // -DUSE_OPTIX_DENOISER (AOV accumulation only)
    bool dropMixed = false;         // reference-exact handling of BXDF_MIXED with separate queues (wf_logic.cl:399-401)
    uint32_t seedOffset = 0;        // rank * N for multi-GPU decorrelation; 0 = reference
    uint32_t currPixelIdx = 0;      // deviceBuffers.currentPixelIdx
    flx_queue_counters cnt;         // deviceBuffers.queueCounters
    std::vector<uint32_t> state;    // 64 slots x N
    std::vector<uint32_t> queues[FLX_NUM_QUEUES];
    std::vector<float> pixels, denoiserAlbedo, denoiserNormal; // float4 per pixel
    std::vector<uint32_t> spp;
    TraceCounters extCounters, shadowCounters; // accumulated over stages (serial mode only)

    float &F(int slot, uint32_t gid) { return ((float *)state.data())[(size_t)slot * N + gid]; }
    uint32_t &U(int slot, uint32_t gid) { return state[(size_t)slot * N + gid]; }
    int32_t &I(int slot, uint32_t gid) { return ((int32_t *)state.data())[(size_t)slot * N + gid]; }
    f3 F3(int slot, uint32_t gid) { return mk3(F(slot, gid), F(slot + 1, gid), F(slot + 2, gid)); }
    void W3(int slot, uint32_t gid, f3 v) { F(slot, gid) = v.x; F(slot + 1, gid) = v.y; F(slot + 2, gid) = v.z; }

    void allocate(uint32_t numTasks, uint32_t numPixels);
    Hit readHit(uint32_t gid);
    void writeHit(uint32_t gid, const Hit &h);
    uint32_t push(int q, uint32_t *counter, uint32_t gid);

    void reset();                       // wf_reset.cl
    void raygen();                      // wf_raygen.cl
    void logic(bool firstIteration);    // wf_logic.cl
    void materials();                   // wf_mat_*.cl
    void extRays();                     // wf_extrays.cl
    void shadowRays();                  // wf_shadowrays.cl
    void clearQueues();                 // CLContext::enqueueClearWfQueues
    void postprocess(float *outRGBA);   // mk_postprocess.cl
    void updatePixelIndex(uint32_t numNewPaths); // CLContext::updatePixelIndex
    Hit pick(float NDCx, float NDCy);   // kernel_pick.cl

    // microkernel renderer (o_microkernel.cpp): src/mk_*.cl, one task per pixel, phase word in slot 46
    flx_render_stats stats = {0, 0, 0, 0};  // device RenderStats (mk_next_vertex.cl:44-56, mk_sample_bsdf.cl:97,127, mk_splat.cl:39)
    void mkReset();                         // mk_reset.cl
    void mkRaygen();                        // mk_raygen.cl
    void mkNextVertex();                    // mk_next_vertex.cl
    void mkSampleBsdf();                    // mk_sample_bsdf.cl
    void mkSplat();                         // mk_splat.cl
    void mkSplatPreview();                  // mk_splat_preview.cl
    void mkRenderSingle(uint32_t spp);      // Tracer::renderSingle
private:
    void materialQueue(int q, uint32_t len);
};

} // namespace orc
