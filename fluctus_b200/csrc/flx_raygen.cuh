// Camera-path generation shared by the raygen stage (k_raygen) and the fused logic+raygen kernel of the
// render entry.  Behavioural contract: reference src/wf_raygen.cl:4-100, src/wf_reset.cl:36-58.
#pragma once
#include "flx_shade.cuh"

// Fields re-initialised by both reset and raygen (wf_reset.cl:36-58, wf_raygen.cl:79-99)
FLX_DEV void resetPathFields(const PathState &st, uint32_t gid, float worldRadius)
{
    st.hit0[gid] = make_float4(0.f, 0.f, 0.f, FLX_FLT_MAX);
    st.hit1[gid] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
    st.hit2[gid] = make_float4(0.f, 0.f, __int_as_float(-1), __int_as_float(0));
    st.nee0[gid] = make_float4(0.f, 0.f, 0.f, 0.f);            // lastEmission, lastPdfDirect
    st.nee1[gid] = make_float4(0.f, 0.f, 0.f, 0.f);            // lastBsdf, lastPdfImplicit
    st.shw0[gid].w = worldRadius + worldRadius;                // shadowRayLen
    st.shw1[gid].w = 0.0f;                                     // lastCosTh
    st.lastSpecular[gid] = 1u;
    st.shadowBlocked[gid] = 1u;
    st.logicFlags[gid] = 0u;                                   // backfaceHit, firstDiffuseHit
    st.lastPickProb[gid] = 1.0f;
}

// Jittered pinhole / thin-lens camera ray through pixel `pixelIdx` (wf_raygen.cl:27-70 == mk_raygen.cl:20-58)
FLX_DEV void cameraRay(uint32_t pixelIdx, uint32_t &seed, const flx_render_params &params, f3 &rayOrig, f3 &rayDirection)
{
    const flx_camera &cam = params.camera;
    float x = (float)(pixelIdx % params.width);
    float y = (float)(pixelIdx) * params.width1; // deliberately not floored, as the reference
    x += rnd(seed);
    y += rnd(seed);
    const float NDCx = x * params.width1;
    const float NDCy = y * params.height1;
    float SCRx = NDCx + NDCx - 1.0f;
    float SCRy = NDCy + NDCy - 1.0f;
    SCRx *= (float)params.width * params.height1;
    SCRx *= cam.fovSCALE;
    SCRy *= cam.fovSCALE;

    rayOrig = ld3(cam.pos);
    const f3 rayTarget = rayOrig + ld3(cam.right) * SCRx + ld3(cam.up) * SCRy + ld3(cam.dir);
    rayDirection = normalize(rayTarget - rayOrig);
    if (cam.apertureSize > 0.0f) // thin lens
    {
        const f3 fp = ld3(cam.pos) + rayDirection * cam.focalDist;
        float dx, dy;
        uniformSampleDisk(seed, dx, dy);
        rayOrig += params.worldRadius * cam.apertureSize * (ld3(cam.right) * dx + ld3(cam.up) * dy);
        rayDirection = normalize(fp - rayOrig);
    }
}

// Path `gid` restarts at the pixel that raygen-queue position `k` maps to (wf_raygen.cl:25: (cursor + k) %
// numPixels), with the RNG state `seed` the terminated path left behind.
FLX_DEV void generateCameraPath(const PathState &st, uint32_t gid, uint32_t seed, uint32_t cursor, uint32_t k,
                                uint32_t tileRank, uint32_t tileWorld, const flx_render_params &params)
{
    uint32_t pixelIdx = (cursor + k) % (params.width * params.height);
    if (tileWorld > 1u) // preview sharding: this rank owns the image rows y % tileWorld == tileRank
    {
        const uint32_t ownedRows = (params.height - tileRank + tileWorld - 1u) / tileWorld;
        const uint32_t j = (cursor + k) % (ownedRows * params.width);
        pixelIdx = ((j / params.width) * tileWorld + tileRank) * params.width + (j % params.width);
    }
    f3 rayOrig, rayDirection;
    cameraRay(pixelIdx, seed, params, rayOrig, rayDirection);
    st.ray0[gid] = mk4(rayOrig, 1.0f);                   // orig, lastPdfW = 1
    st.ray1[gid] = mk4u(rayDirection, 0u);               // dir, pathLen = 0
    st.thr[gid] = mk4u(mk3(1.0f), seed);                 // T = 1, seed
    st.acc[gid] = mk4u(mk3(0.0f), pixelIdx);             // Ei = 0, pixelIndex
    resetPathFields(st, gid, params.worldRadius);
}
