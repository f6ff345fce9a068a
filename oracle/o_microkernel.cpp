// TEST INFRASTRUCTURE — CPU oracle.  Restates the reference's microkernel ("MK") renderer: src/mk_reset.cl,
// mk_raygen.cl, mk_next_vertex.cl, mk_sample_bsdf.cl, mk_splat.cl, mk_splat_preview.cl.  One task per pixel
// (gid == pixel index, tasks beyond min(w*h, NUM_TASKS) idle), a phase word per task (geom.h:205-216), shadow
// rays traced inline.  This is the mode Tracer::renderSingle uses for an exact sample count per pixel
// (src/tracer.cpp:108-182): with a fixed pixel<->task map there are no cross-task float atomics, so images
// are deterministic — the strict image-parity check of tests/test_microkernel_gpu.py.
// Same state layout and scene as the wavefront oracle (o_wavefront.h); pinned bit for bit to the reference's mk_*.cl
// compiled as C++ (tests/test_reference_source_cpu.py), like the rest.
#include "o_wavefront.h"
#include <cstring>

namespace orc {

enum { MK_RT_NEXT_VERTEX = 0, MK_SAMPLE_BSDF = 1, MK_SPLAT_SAMPLE = 4, MK_GENERATE_CAMERA_RAY = 5, MK_DONE = 6 }; // geom.h:205-216

static inline uint32_t mkLimit(const Wavefront &w)
{
    const uint32_t np = w.params.width * w.params.height;
    return np < w.N ? np : w.N;
}

// mk_reset.cl:4-45, launched over (width, height) (clcontext.cpp:727-732)
void Wavefront::mkReset()
{
    const uint32_t limit = mkLimit(*this);
    memset(&stats, 0, sizeof(stats));
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        for (int k = 0; k < 4; k++) { pixels[4 * (size_t)gid + k] = 0.0f; denoiserNormal[4 * (size_t)gid + k] = 0.0f; }
        denoiserAlbedo[4 * (size_t)gid + 0] = 0.1f; denoiserAlbedo[4 * (size_t)gid + 1] = 0.1f;
        denoiserAlbedo[4 * (size_t)gid + 2] = 0.1f; denoiserAlbedo[4 * (size_t)gid + 3] = 0.0f;
        spp[gid] = 0;
        I(FLX_SLOT_PHASE, gid) = MK_GENERATE_CAMERA_RAY;
        W3(FLX_SLOT_EI, gid, mk3(0.0f));
        W3(FLX_SLOT_T, gid, mk3(1.0f));
        U(FLX_SLOT_PATH_LEN, gid) = 0;
        U(FLX_SLOT_LAST_SPECULAR, gid) = 1;
        F(FLX_SLOT_LAST_PDF_W, gid) = 1.0f;
        U(FLX_SLOT_FIRST_DIFFUSE, gid) = 0;
        U(FLX_SLOT_SEED, gid) = gid + seedOffset;
    }
}

// mk_raygen.cl:5-67
void Wavefront::mkRaygen()
{
    const uint32_t limit = mkLimit(*this);
    const flx_camera &cam = params.camera;
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        if (I(FLX_SLOT_PHASE, gid) != MK_GENERATE_CAMERA_RAY) continue;
        uint32_t seed = U(FLX_SLOT_SEED, gid);
        float x = (float)(gid % params.width);
        float y = (float)(gid) * params.width1; // not floored
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
            const f2 dsk = uniformSampleDisk(&seed);
            rayOrig += params.worldRadius * cam.apertureSize * (ld3(cam.right) * dsk.x + ld3(cam.up) * dsk.y);
            rayDirection = normalize(fp - rayOrig);
        }
        W3(FLX_SLOT_ORIG, gid, rayOrig);
        W3(FLX_SLOT_DIR, gid, rayDirection);
        U(FLX_SLOT_SEED, gid) = seed;
        I(FLX_SLOT_PHASE, gid) = MK_RT_NEXT_VERTEX;
    }
}

// mk_next_vertex.cl:7-123
void Wavefront::mkNextVertex()
{
    const uint32_t limit = mkLimit(*this);
    uint32_t nPrimary = 0, nExtension = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nPrimary, nExtension)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        if (I(FLX_SLOT_PHASE, gid) != MK_RT_NEXT_VERTEX) continue;
        Ray r; r.orig = F3(FLX_SLOT_ORIG, gid); r.dir = F3(FLX_SLOT_DIR, gid);
        Hit hit = emptyHit(FLT_MAX);
        if (useBitstack) bvhIntersectBitstack(r, &hit, geo, nullptr); else bvhIntersectStack(r, &hit, geo, nullptr);
        if (params.sampleImpl && params.useAreaLight) intersectLight(&hit, r, &params);
        writeHit(gid, hit);

        uint32_t len = U(FLX_SLOT_PATH_LEN, gid);
        if (len == 0) nPrimary++; else nExtension++; // :44-56
        len += 1;
        U(FLX_SLOT_PATH_LEN, gid) = len;

        if (denoiserAOV && len == 1) // :59-69
        {
            const f3 r1 = ld3(params.camera.right), r2 = ld3(params.camera.up), r3 = -ld3(params.camera.dir);
            float *px = &denoiserNormal[4 * (size_t)gid];
            px[0] += dot(r1, hit.N); px[1] += dot(r2, hit.N); px[2] += dot(r3, hit.N); px[3] += 1.0f;
        }

        if (hit.i < 0) // implicit environment sample (:72-93)
        {
            f3 bg = mk3(0.0f);
            if (params.useEnvMap && (len == 1 || params.sampleImpl)) bg = evalEnvMapDir(shade, r.dir) * params.envMapStrength;
            float weight = 1.0f;
            const bool lastSpecular = U(FLX_SLOT_LAST_SPECULAR, gid) != 0;
            if (params.sampleImpl && params.sampleExpl && params.useEnvMap && len > 1 && !lastSpecular)
            {
                const float lightPickProb = 1.0f;
                const float directPdfW = envMapPdf(shade, r.dir);
                const float actualPdfW = F(FLX_SLOT_LAST_PDF_W, gid);
                weight = (actualPdfW * lightPickProb) / (actualPdfW * lightPickProb + directPdfW);
            }
            const f3 T = F3(FLX_SLOT_T, gid);
            W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + weight * T * bg);
            I(FLX_SLOT_PHASE, gid) = MK_SPLAT_SAMPLE;
        }
        else if (hit.areaLightHit) // implicit area-light sample (:95-116)
        {
            float misWeight = 1.0f;
            const bool lastSpecular = U(FLX_SLOT_LAST_SPECULAR, gid) != 0;
            if (params.sampleExpl && len > 1 && !lastSpecular)
            {
                const float directPdfA = recip(4.0f * params.areaLight.size.x * params.areaLight.size.y);
                const float directPdfW = pdfAtoW(directPdfA, length(hit.P - r.orig), -dot(r.dir, hit.N));
                const float lightPickProb = 1.0f;
                const float lastPdfW = F(FLX_SLOT_LAST_PDF_W, gid);
                misWeight = lastPdfW / (lastPdfW + directPdfW * lightPickProb);
            }
            const f3 T = F3(FLX_SLOT_T, gid);
            W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + T * misWeight * ld3(params.areaLight.E));
            I(FLX_SLOT_PHASE, gid) = MK_SPLAT_SAMPLE;
        }
        else I(FLX_SLOT_PHASE, gid) = MK_SAMPLE_BSDF;
    }
    stats.primaryRays += nPrimary; stats.extensionRays += nExtension;
}

// mk_sample_bsdf.cl:11-188
void Wavefront::mkSampleBsdf()
{
    const uint32_t limit = mkLimit(*this);
    uint32_t nShadow = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nShadow)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        if (I(FLX_SLOT_PHASE, gid) != MK_SAMPLE_BSDF) continue;
        uint32_t seed = U(FLX_SLOT_SEED, gid);
        Ray r; r.orig = F3(FLX_SLOT_ORIG, gid); r.dir = F3(FLX_SLOT_DIR, gid);
        Hit hit = readHit(gid);
        const flx_material &mat = shade.materials[hit.matId];
        hit.N = tangentSpaceNormal(hit, mat, shade);
        const bool backface = dot(hit.N, r.dir) > 0.0f;
        if (backface) hit.N = -hit.N;
        f3 orig = hit.P - 1e-3f * r.dir;

        if (denoiserAOV) // :56-67
        {
            const bool isDiffuse = !FLX_BXDF_IS_SINGULAR(mat.type);
            if (isDiffuse && !U(FLX_SLOT_FIRST_DIFFUSE, gid))
            {
                U(FLX_SLOT_FIRST_DIFFUSE, gid) = 1;
                const f3 albedo = matGetFloat3(ld3(mat.Kd), hit.uv, mat.map_Kd, shade);
                float *px = &denoiserAlbedo[4 * (size_t)gid];
                px[0] += albedo.x; px[1] += albedo.y; px[2] += albedo.z; px[3] += 1.0f;
            }
        }

        if (params.sampleExpl && !FLX_BXDF_IS_SINGULAR(mat.type)) // next event estimation (:70-141)
        {
            const float lightPickProb = 1.0f;
            if (params.useEnvMap) // :75-108
            {
                f3 L; float directPdfW = 0.0f;
                sampleEnvMapAlias(rnd(&seed), &L, &directPdfW, shade);
                const float lenL = params.worldRadius + params.worldRadius;
                Ray rLight; rLight.orig = orig; rLight.dir = L;
                Hit hitL = emptyHit(lenL);
                if (params.useAreaLight) intersectLight(&hitL, rLight, &params);
                const bool occluded = (hitL.i > -1) ||
                    (useBitstack ? bvhOccludedBitstack(rLight, lenL, geo, nullptr) : bvhOccludedStack(rLight, lenL, geo, nullptr));
                nShadow++;
                if (!occluded && directPdfW != 0.0f)
                {
                    const f3 brdf = bxdfEval(hit, mat, backface, shade, r.dir, L);
                    const float cosTh = fmaxf(0.0f, dot(L, hit.N));
                    const float bsdfPdfW = fmaxf(0.0f, bxdfPdf(hit, mat, backface, shade, r.dir, L));
                    const f3 T = F3(FLX_SLOT_T, gid);
                    const f3 envMapLi = evalEnvMapDir(shade, L) * params.envMapStrength;
                    const f3 contrib = brdf * T * envMapLi * cosTh / (lightPickProb * directPdfW + (float)(params.sampleImpl) * bsdfPdfW);
                    W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + contrib);
                }
            }
            if (params.useAreaLight) // :111-140
            {
                float directPdfA; f3 posL;
                sampleAreaLight(params.areaLight, &directPdfA, &posL, &seed);
                f3 L = posL - orig;
                const float lenL = length(L);
                L /= lenL;
                Ray rLight; rLight.orig = orig; rLight.dir = L;
                const bool occluded = useBitstack ? bvhOccludedBitstack(rLight, lenL, geo, nullptr) : bvhOccludedStack(rLight, lenL, geo, nullptr);
                nShadow++;
                const float cosLight = fmaxf(dot(ld3(params.areaLight.N), -L), 0.0f);
                if (!occluded && cosLight > 0.0f)
                {
                    const f3 brdf = bxdfEval(hit, mat, backface, shade, r.dir, L);
                    const float cosTh = fmaxf(0.0f, dot(L, hit.N));
                    const float directPdfW = pdfAtoW(directPdfA, lenL, cosLight);
                    const float bsdfPdfW = fmaxf(0.0f, bxdfPdf(hit, mat, backface, shade, r.dir, L));
                    const f3 T = F3(FLX_SLOT_T, gid);
                    const f3 contrib = brdf * T * ld3(params.areaLight.E) * cosTh / (lightPickProb * directPdfW + (float)(params.sampleImpl) * bsdfPdfW);
                    W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + contrib);
                }
            }
        }

        // termination (:143-152)
        float contProb = 1.0f;
        const uint32_t len = U(FLX_SLOT_PATH_LEN, gid);
        bool terminate = params.maxBounces > 0 && (len >= params.maxBounces + 1);
        if (!terminate && params.useRoulette && len > 5 /* MIN_PATH_LENGTH */)
        {
            contProb = clampf(luminance(F3(FLX_SLOT_T, gid)), 0.01f, 0.5f);
            terminate = (rnd(&seed) > contProb);
        }

        // continuation ray (:154-186); pdfW / newDir are left unset by the reference for emissive and rejected
        // glossy samples — pinned to 0 like in the wavefront oracle
        float pdfW = 0.0f;
        f3 newDir = mk3(0.0f);
        const f3 bsdf = bxdfSample(hit, mat, backface, shade, r.dir, &newDir, &pdfW, &seed);
        pdfW *= contProb;
        if (pdfW == 0.0f || isZero(bsdf)) terminate = true;
        const f3 newT = F3(FLX_SLOT_T, gid) * bsdf * dot(hit.N, newDir) / pdfW;
        orig = hit.P + 1e-4f * newDir;

        W3(FLX_SLOT_T, gid, newT);
        W3(FLX_SLOT_ORIG, gid, orig);
        W3(FLX_SLOT_DIR, gid, newDir);
        F(FLX_SLOT_LAST_PDF_W, gid) = pdfW;
        U(FLX_SLOT_SEED, gid) = seed;
        U(FLX_SLOT_LAST_SPECULAR, gid) = FLX_BXDF_IS_SINGULAR(mat.type) ? 1u : 0u;
        I(FLX_SLOT_PHASE, gid) = terminate ? MK_SPLAT_SAMPLE : MK_RT_NEXT_VERTEX;
    }
    stats.shadowRays += nShadow;
}

// mk_splat.cl:5-57, launched over (width, height) (clcontext.cpp:758-762)
void Wavefront::mkSplat()
{
    const uint32_t limit = mkLimit(*this);
    const bool checkSpp = params.maxSpp > 0; // -DCHECK_SPP
    uint32_t nSamples = 0;
#pragma omp parallel for schedule(static) reduction(+ : nSamples)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        if (I(FLX_SLOT_PHASE, gid) != MK_SPLAT_SAMPLE) continue;
        if (checkSpp)
        {
            const bool shouldSplat = spp[gid]++ <= params.maxSpp; // `<=` on the old value: maxSpp + 1 samples (§8g.3)
            if (!shouldSplat)
            {
                I(FLX_SLOT_PHASE, gid) = MK_DONE;
                spp[gid] = params.maxSpp;
                continue;
            }
        }
        float *px = &pixels[4 * (size_t)gid];
        const f3 Ei = F3(FLX_SLOT_EI, gid);
        float c[4] = {Ei.x, Ei.y, Ei.z, 1.0f};
        if (px[3] > 0.0f) { c[0] += px[0]; c[1] += px[1]; c[2] += px[2]; c[3] += px[3]; }
        px[0] = c[0]; px[1] = c[1]; px[2] = c[2]; px[3] = c[3];
        nSamples++;
        W3(FLX_SLOT_EI, gid, mk3(0.0f));
        W3(FLX_SLOT_T, gid, mk3(1.0f));
        U(FLX_SLOT_PATH_LEN, gid) = 0;
        U(FLX_SLOT_FIRST_DIFFUSE, gid) = 0;
        I(FLX_SLOT_PHASE, gid) = MK_GENERATE_CAMERA_RAY;
    }
    stats.samples += nSamples;
}

// mk_splat_preview.cl:5-25: every task overwrites its pixel, weight 0 (the next splat replaces it)
void Wavefront::mkSplatPreview()
{
    const uint32_t limit = mkLimit(*this);
#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)limit; g++)
    {
        const uint32_t gid = (uint32_t)g;
        float *px = &pixels[4 * (size_t)gid];
        const f3 Ei = F3(FLX_SLOT_EI, gid);
        px[0] = Ei.x; px[1] = Ei.y; px[2] = Ei.z; px[3] = 0.0f;
        W3(FLX_SLOT_EI, gid, mk3(0.0f));
        W3(FLX_SLOT_T, gid, mk3(1.0f));
        U(FLX_SLOT_PATH_LEN, gid) = 0;
        I(FLX_SLOT_PHASE, gid) = MK_GENERATE_CAMERA_RAY;
    }
}

// Tracer::renderSingle (src/tracer.cpp:108-182): reset, then per sample raygen -> (nextVertex -> sampleBsdf)
// x (maxBounces + 1) -> splat.  Russian roulette is switched off by the caller, as the reference does.
void Wavefront::mkRenderSingle(uint32_t spp_)
{
    mkReset();
    for (uint32_t s = 0; s < spp_; s++)
    {
        mkRaygen();
        for (uint32_t b = 0; b < params.maxBounces + 1; b++) { mkNextVertex(); mkSampleBsdf(); }
        mkSplat();
    }
}

} // namespace orc
