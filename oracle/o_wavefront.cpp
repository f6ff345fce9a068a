// TEST INFRASTRUCTURE — CPU oracle.  Restates reference src/wf_reset.cl, wf_raygen.cl,
// wf_logic.cl, wf_extrays.cl, wf_shadowrays.cl, wf_mat_*.cl and mk_postprocess.cl.
// Work-items run in ascending gid order when OMP_NUM_THREADS=1 (this defines the queue order
// of the serial oracle); with more threads the same code is the timed CPU baseline.
#include "o_wavefront.h"
#include <cstring>

namespace orc {

void Wavefront::allocate(uint32_t numTasks, uint32_t numPixels)
{
    N = numTasks;
    state.assign((size_t)FLX_NUM_SLOTS * N, 0u);
    for (int q = 0; q < FLX_NUM_QUEUES; q++) queues[q].assign(N, 0u);
    pixels.assign((size_t)numPixels * 4, 0.0f);
    denoiserAlbedo.assign((size_t)numPixels * 4, 0.0f);
    denoiserNormal.assign((size_t)numPixels * 4, 0.0f);
    spp.assign(numPixels, 0u);
    memset(&cnt, 0, sizeof(cnt));
    memset(&extCounters, 0, sizeof(extCounters));
    memset(&shadowCounters, 0, sizeof(shadowCounters));
    currPixelIdx = 0;
}

Hit Wavefront::readHit(uint32_t gid) // utils.cl:238-249
{
    Hit h;
    h.P = F3(FLX_SLOT_P, gid); h.N = F3(FLX_SLOT_N, gid);
    h.uv.x = F(FLX_SLOT_UV, gid); h.uv.y = F(FLX_SLOT_UV + 1, gid);
    h.t = F(FLX_SLOT_HIT_T, gid); h.i = I(FLX_SLOT_HIT_I, gid);
    h.areaLightHit = I(FLX_SLOT_AREA_LIGHT_HIT, gid); h.matId = I(FLX_SLOT_MAT_ID, gid);
    return h;
}
void Wavefront::writeHit(uint32_t gid, const Hit &h) // utils.cl:227-236
{
    W3(FLX_SLOT_P, gid, h.P); W3(FLX_SLOT_N, gid, h.N);
    F(FLX_SLOT_UV, gid) = h.uv.x; F(FLX_SLOT_UV + 1, gid) = h.uv.y;
    F(FLX_SLOT_HIT_T, gid) = h.t; I(FLX_SLOT_HIT_I, gid) = h.i;
    I(FLX_SLOT_AREA_LIGHT_HIT, gid) = h.areaLightHit; I(FLX_SLOT_MAT_ID, gid) = h.matId;
}

uint32_t Wavefront::push(int q, uint32_t *counter, uint32_t gid) // atomic_inc + store
{
    uint32_t idx;
#pragma omp atomic capture
    { idx = *counter; (*counter)++; }
    queues[q][idx] = gid;
    return idx;
}

static inline void atomicAddF(float *p, float v)
{
#pragma omp atomic
    *p += v;
}
static inline void addFloat4(float *p, float x, float y, float z, float w) // utils.cl:298-305, 329-337
{
    atomicAddF(p + 0, x); atomicAddF(p + 1, y); atomicAddF(p + 2, z); atomicAddF(p + 3, w);
}

// Fields both reset and raygen re-initialise (wf_reset.cl:36-58, wf_raygen.cl:79-99)
static inline void resetPathFields(Wavefront &w, uint32_t gid)
{
    w.W3(FLX_SLOT_EI, gid, mk3(0.0f));
    w.W3(FLX_SLOT_T, gid, mk3(1.0f));
    w.U(FLX_SLOT_PATH_LEN, gid) = 0;
    w.U(FLX_SLOT_FIRST_DIFFUSE, gid) = 0;
    w.U(FLX_SLOT_LAST_SPECULAR, gid) = 1;
    w.F(FLX_SLOT_LAST_PDF_W, gid) = 1.0f;
    w.F(FLX_SLOT_LAST_PDF_DIRECT, gid) = 0.0f;
    w.F(FLX_SLOT_LAST_PDF_IMPLICIT, gid) = 0.0f;
    w.F(FLX_SLOT_LAST_COS_TH, gid) = 0.0f;
    w.F(FLX_SLOT_LAST_PICK_PROB, gid) = 1.0f;
    w.F(FLX_SLOT_SHADOW_LEN, gid) = w.params.worldRadius + w.params.worldRadius;
    w.U(FLX_SLOT_BACKFACE, gid) = 0;
    w.U(FLX_SLOT_SHADOW_BLOCKED, gid) = 1;
    w.W3(FLX_SLOT_LAST_EMISSION, gid, mk3(0.0f));
    w.W3(FLX_SLOT_LAST_BSDF, gid, mk3(0.0f));
    w.writeHit(gid, emptyHit(FLT_MAX));
}

// wf_reset.cl:5-68; launched over max(NUM_TASKS, w*h) (clcontext.cpp:783-788)
void Wavefront::reset()
{
    const uint32_t numPixels = params.width * params.height;
    const uint32_t total = N > numPixels ? N : numPixels;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)total; g++)
    {
        const uint32_t gid = (uint32_t)g;
        if (gid < numPixels)
        {
            for (int k = 0; k < 4; k++) { pixels[4 * (size_t)gid + k] = 0.0f; denoiserNormal[4 * (size_t)gid + k] = 0.0f; }
            denoiserAlbedo[4 * (size_t)gid + 0] = 0.1f; denoiserAlbedo[4 * (size_t)gid + 1] = 0.1f;
            denoiserAlbedo[4 * (size_t)gid + 2] = 0.1f; denoiserAlbedo[4 * (size_t)gid + 3] = 0.0f;
            spp[gid] = 0;
        }
        if (gid >= N) continue;
        resetPathFields(*this, gid);
        U(FLX_SLOT_PIXEL_INDEX, gid) = 0;
        U(FLX_SLOT_SEED, gid) = gid + seedOffset;
        queues[FLX_Q_RAYGEN][gid] = gid;
        if (gid == 0) cnt.raygenQueue = N;
    }
}

// wf_raygen.cl:4-100
void Wavefront::raygen()
{
    const uint32_t qlen = cnt.raygenQueue;
    const flx_camera &cam = params.camera;
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < (int64_t)qlen; k++)
    {
        const uint32_t gid_direct = (uint32_t)k;
        const uint32_t gid = queues[FLX_Q_RAYGEN][gid_direct];
        uint32_t seed = U(FLX_SLOT_SEED, gid);

        const uint32_t numPixels = params.width * params.height;
        const uint32_t pixelIdx = (currPixelIdx + gid_direct) % numPixels;
        U(FLX_SLOT_PIXEL_INDEX, gid) = pixelIdx;

        float x = (float)(pixelIdx % params.width);
        float y = (float)(pixelIdx) * params.width1; // not floored
        x += rnd(&seed);
        y += rnd(&seed);
        const float NDCx = x * params.width1;
        const float NDCy = y * params.height1;
        float SCRx = NDCx + NDCx - 1.0f;
        float SCRy = NDCy + NDCy - 1.0f;
        SCRx *= (float)params.width * params.height1;
        SCRx *= cam.fovSCALE;
        SCRy *= cam.fovSCALE;

        f3 rayOrig = ld3(cam.pos);
        const f3 rayTarget = rayOrig + ld3(cam.right) * SCRx + ld3(cam.up) * SCRy + ld3(cam.dir);
        f3 rayDirection = normalize(rayTarget - rayOrig);
        if (cam.apertureSize > 0.0f)
        {
            const f3 fp = ld3(cam.pos) + rayDirection * cam.focalDist;
            const f2 r = uniformSampleDisk(&seed);
            rayOrig += params.worldRadius * cam.apertureSize * (ld3(cam.right) * r.x + ld3(cam.up) * r.y);
            rayDirection = normalize(fp - rayOrig);
        }
        W3(FLX_SLOT_ORIG, gid, rayOrig);
        W3(FLX_SLOT_DIR, gid, rayDirection);
        push(FLX_Q_EXTENSION, &cnt.extensionQueue, gid);
        U(FLX_SLOT_SEED, gid) = seed;
        resetPathFields(*this, gid);
    }
}

// wf_extrays.cl:5-36
void Wavefront::extRays()
{
    const uint32_t qlen = cnt.extensionQueue;
    TraceCounters acc; memset(&acc, 0, sizeof(acc));
#pragma omp parallel
    {
        TraceCounters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 256)
        for (int64_t k = 0; k < (int64_t)qlen; k++)
        {
            const uint32_t gid = queues[FLX_Q_EXTENSION][k];
            Ray r; r.orig = F3(FLX_SLOT_ORIG, gid); r.dir = F3(FLX_SLOT_DIR, gid);
            Hit hit = emptyHit(FLT_MAX);
            if (useBitstack) bvhIntersectBitstack(r, &hit, geo, &c); else bvhIntersectStack(r, &hit, geo, &c);
            if (params.sampleImpl && params.useAreaLight) intersectLight(&hit, r, &params);
            U(FLX_SLOT_PATH_LEN, gid) += 1;
            writeHit(gid, hit);
        }
#pragma omp critical
        {
            acc.inner_visits += c.inner_visits; acc.leaf_visits += c.leaf_visits; acc.tri_tests += c.tri_tests;
            acc.hits += c.hits; if (c.max_stack > acc.max_stack) acc.max_stack = c.max_stack;
        }
    }
    extCounters.inner_visits += acc.inner_visits; extCounters.leaf_visits += acc.leaf_visits;
    extCounters.tri_tests += acc.tri_tests; extCounters.hits += acc.hits;
    if (acc.max_stack > extCounters.max_stack) extCounters.max_stack = acc.max_stack;
}

// wf_shadowrays.cl:6-39
void Wavefront::shadowRays()
{
    const uint32_t qlen = cnt.shadowQueue;
    TraceCounters acc; memset(&acc, 0, sizeof(acc));
#pragma omp parallel
    {
        TraceCounters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 256)
        for (int64_t k = 0; k < (int64_t)qlen; k++)
        {
            const uint32_t gid = queues[FLX_Q_SHADOW][k];
            Ray r; r.orig = F3(FLX_SLOT_SHADOW_ORIG, gid); r.dir = F3(FLX_SLOT_SHADOW_DIR, gid);
            const float lenL = F(FLX_SLOT_SHADOW_LEN, gid);
            Hit hitL = emptyHit(lenL);
            if (params.useAreaLight) intersectLight(&hitL, r, &params);
            const bool occluded = (hitL.i > -1) ||
                (useBitstack ? bvhOccludedBitstack(r, lenL, geo, &c) : bvhOccludedStack(r, lenL, geo, &c));
            U(FLX_SLOT_SHADOW_BLOCKED, gid) = occluded ? 1u : 0u;
        }
#pragma omp critical
        { acc.inner_visits += c.inner_visits; acc.leaf_visits += c.leaf_visits; acc.tri_tests += c.tri_tests; acc.hits += c.hits; }
    }
    shadowCounters.inner_visits += acc.inner_visits; shadowCounters.leaf_visits += acc.leaf_visits;
    shadowCounters.tri_tests += acc.tri_tests; shadowCounters.hits += acc.hits;
}

// kernel_pick.cl:6-41 (CLContext::pickSingle): ray through an NDC point from the camera position, no lens sample
Hit Wavefront::pick(float NDCx, float NDCy)
{
    float SCRx = NDCx + NDCx - 1.0f;
    float SCRy = NDCy + NDCy - 1.0f;
    SCRx *= (float)params.width * params.height1;
    SCRx *= params.camera.fovSCALE;
    SCRy *= params.camera.fovSCALE;
    const f3 pos = ld3(params.camera.pos);
    const f3 rayTarget = pos + ld3(params.camera.right) * SCRx + ld3(params.camera.up) * SCRy + ld3(params.camera.dir);
    Ray r; r.orig = pos; r.dir = normalize(rayTarget - pos);
    Hit hit = emptyHit(FLT_MAX);
    if (useBitstack) bvhIntersectBitstack(r, &hit, geo, nullptr); else bvhIntersectStack(r, &hit, geo, nullptr);
    if (params.sampleImpl && params.useAreaLight) intersectLight(&hit, r, &params);
    return hit;
}

void Wavefront::clearQueues() { memset(&cnt, 0, sizeof(cnt)); } // clcontext.cpp:903-909

void Wavefront::updatePixelIndex(uint32_t numNewPaths) // clcontext.cpp:917-921
{
    currPixelIdx = (currPixelIdx + numNewPaths) % (params.width * params.height);
}

// mk_postprocess.cl:7-56
void Wavefront::postprocess(float *outRGBA)
{
    const uint32_t limit = params.width * params.height;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)limit; g++)
        postprocessPixel(&pixels[4 * (size_t)g], outRGBA + 4 * (size_t)g, params.ppParams.exposure, params.ppParams.tmOperator);
}

} // namespace orc
