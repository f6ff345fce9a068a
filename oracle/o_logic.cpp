// TEST INFRASTRUCTURE — CPU oracle.  Restates reference src/wf_logic.cl (state machine, splat,
// NEE sample generation, material-queue routing) and the shared body of src/wf_mat_*.cl.
#include "o_wavefront.h"

namespace orc {

static inline void atomicAddF(float *p, float v)
{
#pragma omp atomic
    *p += v;
}

// wf_logic.cl:352-407 (addToMaterialQueueNaive) incl. -DWF_SINGLE_MAT_QUEUE.  BXDF_MIXED falls
// into `default:` in the reference and the path is dropped: reproduced when
// `dropMixed` is set (tests/test_reference_source_cpu.py does, to stay bit-identical with the
// reference's own code); otherwise it goes to the diffuse queue like the single-queue mode does —
// the documented deviation of the CUDA path.
static inline void addToMaterialQueue(Wavefront &w, uint32_t gid, int type)
{
    if (!w.params.wfSeparateQueues) { w.push(FLX_Q_DIFFUSE, &w.cnt.diffuseQueue, gid); return; }
    switch (type)
    {
    case FLX_BXDF_MIXED: if (w.dropMixed) break; // fall through
    case FLX_BXDF_DIFFUSE: w.push(FLX_Q_DIFFUSE, &w.cnt.diffuseQueue, gid); break;
    case FLX_BXDF_GLOSSY: w.push(FLX_Q_GLOSSY, &w.cnt.glossyQueue, gid); break;
    case FLX_BXDF_GGX_ROUGH_REFLECTION: w.push(FLX_Q_GGX_REFL, &w.cnt.ggxReflQueue, gid); break;
    case FLX_BXDF_GGX_ROUGH_DIELECTRIC: w.push(FLX_Q_GGX_REFR, &w.cnt.ggxRefrQueue, gid); break;
    case FLX_BXDF_IDEAL_REFLECTION:
    case FLX_BXDF_IDEAL_DIELECTRIC: w.push(FLX_Q_DELTA, &w.cnt.deltaQueue, gid); break;
    case FLX_BXDF_EMISSIVE: w.push(FLX_Q_EMISSIVE, &w.cnt.emissiveQueue, gid); break;
    default: break;
    }
}

// wf_logic.cl:14-344
void Wavefront::logic(bool firstIteration)
{
    const uint32_t numPixels = params.width * params.height;
    const uint32_t maxId = firstIteration ? (numPixels < N ? numPixels : N) : N;
    const bool checkSpp = params.maxSpp > 0; // -DCHECK_SPP (kernel_impl.hpp:61)

#pragma omp parallel for schedule(static)
    for (int64_t g = 0; g < (int64_t)maxId; g++)
    {
        const uint32_t gid = (uint32_t)g;
        uint32_t seed = U(FLX_SLOT_SEED, gid);
        const uint32_t len = U(FLX_SLOT_PATH_LEN, gid);
        Hit hit = readHit(gid);
        Ray r; r.orig = F3(FLX_SLOT_ORIG, gid); r.dir = F3(FLX_SLOT_DIR, gid);
        f3 T = F3(FLX_SLOT_T, gid);

        // Russian roulette (:59-74)
        bool terminate = params.maxBounces > 0 && (len >= params.maxBounces + 1);
        if (!terminate && params.useRoulette && len > 5 /* MIN_PATH_LENGTH */)
        {
            const float contProb = clampf(luminance(T), 0.01f, 0.5f);
            terminate = (rnd(&seed) > contProb);
            T /= contProb;
            W3(FLX_SLOT_T, gid, T);
        }

        uint32_t pixIdx = 0; bool maxSamplesReached = false;
        if (checkSpp) // :76-84
        {
            pixIdx = U(FLX_SLOT_PIXEL_INDEX, gid);
            uint32_t cur;
#pragma omp atomic read
            cur = spp[pixIdx];
            if (cur >= params.maxSpp) { terminate = true; maxSamplesReached = true; }
        }

        if (isZero(T) || F(FLX_SLOT_LAST_PDF_W, gid) == 0.0f) terminate = true; // :87-88

        if (hit.i < 0 && !terminate) // implicit environment sample (:99-122)
        {
            float weight = 1.0f;
            const bool lastSpecular = U(FLX_SLOT_LAST_SPECULAR, gid) != 0;
            f3 bg = mk3(0.0f);
            if (params.useEnvMap)
            {
                if (len == 1 || params.sampleImpl) bg = evalEnvMapDir(shade, r.dir) * params.envMapStrength;
                if (params.sampleImpl && params.sampleExpl && len > 1 && !lastSpecular)
                {
                    const float lightPickProb = F(FLX_SLOT_LAST_PICK_PROB, gid);
                    const float directPdfW = envMapPdf(shade, r.dir);
                    const float actualPdfW = F(FLX_SLOT_LAST_PDF_W, gid);
                    weight = (actualPdfW * lightPickProb) / (actualPdfW * lightPickProb + directPdfW);
                }
            }
            W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + weight * T * bg);
            terminate = true;
        }
        else if (params.useAreaLight && hit.areaLightHit && !terminate) // implicit area light (:126-146)
        {
            float misWeight = 1.0f;
            const bool lastSpecular = U(FLX_SLOT_LAST_SPECULAR, gid) != 0;
            if (params.sampleExpl && len > 1 && !lastSpecular)
            {
                const float directPdfA = recip(4.0f * params.areaLight.size.x * params.areaLight.size.y);
                const float directPdfW = pdfAtoW(directPdfA, length(hit.P - r.orig), -dot(r.dir, hit.N));
                const float lightPickProb = F(FLX_SLOT_LAST_PICK_PROB, gid);
                const float lastPdfW = F(FLX_SLOT_LAST_PDF_W, gid);
                misWeight = lastPdfW / (lastPdfW + directPdfW * lightPickProb);
            }
            W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + T * misWeight * ld3(params.areaLight.E));
            terminate = true;
        }

        // resolve the previous NEE sample (:150-168)
        if (U(FLX_SLOT_SHADOW_BLOCKED, gid) == 0)
        {
            const f3 emission = F3(FLX_SLOT_LAST_EMISSION, gid);
            const f3 bsdf = F3(FLX_SLOT_LAST_BSDF, gid);
            const float cosTh = F(FLX_SLOT_LAST_COS_TH, gid);
            const float directPdfW = F(FLX_SLOT_LAST_PDF_DIRECT, gid);
            const float bsdfPdfW = F(FLX_SLOT_LAST_PDF_IMPLICIT, gid);
            const float lightPickProb = F(FLX_SLOT_LAST_PICK_PROB, gid);
            const f3 lastT = F3(FLX_SLOT_LAST_T, gid);
            const f3 contrib = bsdf * lastT * emission * cosTh / (lightPickProb * directPdfW + (float)(params.sampleImpl) * bsdfPdfW);
            W3(FLX_SLOT_EI, gid, F3(FLX_SLOT_EI, gid) + contrib);
        }

        if (terminate) // splat + regenerate (:171-205)
        {
            if (checkSpp)
            {
                bool admit = false;
                if (!maxSamplesReached)
                {
                    uint32_t old;
#pragma omp atomic capture
                    { old = spp[pixIdx]; spp[pixIdx]++; }
                    admit = old <= params.maxSpp && len > 0; // `<=` admits maxSpp+1
                }
                if (admit)
                {
                    const f3 Ei = F3(FLX_SLOT_EI, gid);
                    float *px = &pixels[4 * (size_t)pixIdx];
                    atomicAddF(px + 0, Ei.x); atomicAddF(px + 1, Ei.y); atomicAddF(px + 2, Ei.z); atomicAddF(px + 3, 1.0f);
#pragma omp atomic
                    cnt.splattedSamples++;
                }
                uint32_t cur;
#pragma omp atomic read
                cur = spp[pixIdx];
                if (cur > params.maxSpp)
                {
#pragma omp atomic write
                    spp[pixIdx] = params.maxSpp;
                }
            }
            else if (len > 0)
            {
                const uint32_t pix = U(FLX_SLOT_PIXEL_INDEX, gid);
                const f3 Ei = F3(FLX_SLOT_EI, gid);
                float *px = &pixels[4 * (size_t)pix];
                atomicAddF(px + 0, Ei.x); atomicAddF(px + 1, Ei.y); atomicAddF(px + 2, Ei.z); atomicAddF(px + 3, 1.0f);
            }
            push(FLX_Q_RAYGEN, &cnt.raygenQueue, gid);
            U(FLX_SLOT_SEED, gid) = seed;
            continue;
        }

        // surviving path: material, shading normal, backface flip (:208-241)
        const flx_material &mat = shade.materials[hit.matId];
        hit.N = tangentSpaceNormal(hit, mat, shade);
        const bool backface = dot(hit.N, r.dir) > 0.0f;
        if (backface) hit.N = -hit.N;
        const f3 orig = hit.P - 1e-3f * r.dir;

        if (denoiserAOV) // -DUSE_OPTIX_DENOISER AOVs (:214-237)
        {
            if (len == 1)
            {
                const f3 r1 = ld3(params.camera.right), r2 = ld3(params.camera.up), r3 = -ld3(params.camera.dir);
                const uint32_t pix = U(FLX_SLOT_PIXEL_INDEX, gid);
                float *px = &denoiserNormal[4 * (size_t)pix];
                atomicAddF(px + 0, dot(r1, hit.N)); atomicAddF(px + 1, dot(r2, hit.N)); atomicAddF(px + 2, dot(r3, hit.N)); atomicAddF(px + 3, 1.0f);
            }
            const bool isDiffuse = !FLX_BXDF_IS_SINGULAR(mat.type);
            if (isDiffuse && !U(FLX_SLOT_FIRST_DIFFUSE, gid))
            {
                U(FLX_SLOT_FIRST_DIFFUSE, gid) = 1;
                const uint32_t pix = U(FLX_SLOT_PIXEL_INDEX, gid);
                const f3 albedo = matGetFloat3(ld3(mat.Kd), hit.uv, mat.map_Kd, shade);
                float *px = &denoiserAlbedo[4 * (size_t)pix];
                atomicAddF(px + 0, albedo.x); atomicAddF(px + 1, albedo.y); atomicAddF(px + 2, albedo.z); atomicAddF(px + 3, 1.0f);
            }
        }

        writeHit(gid, hit);
        U(FLX_SLOT_BACKFACE, gid) = backface ? 1u : 0u;

        if (params.sampleExpl) // -DSAMPLE_EXPLICIT (:243-333)
        {
            U(FLX_SLOT_SHADOW_BLOCKED, gid) = 1;
            if (!FLX_BXDF_IS_SINGULAR(mat.type))
            {
                const uint32_t den = params.useEnvMap + params.useAreaLight;
                const float envMapProb = (float)params.useEnvMap / (float)(den > 1u ? den : 1u);
                const bool useEnv = rnd(&seed) < envMapProb;
                const bool useArea = !useEnv && params.useAreaLight;
                if (useEnv && params.useEnvMap)
                {
                    const float lightPickProb = envMapProb;
                    f3 L; float directPdfW = 0.0f;
                    sampleEnvMapAlias(rnd(&seed), &L, &directPdfW, shade);
                    const float lenL = params.worldRadius + params.worldRadius;
                    const float cosTh = fmaxf(0.0f, dot(L, hit.N));
                    const f3 envMapLi = evalEnvMapDir(shade, L) * params.envMapStrength;
                    W3(FLX_SLOT_SHADOW_ORIG, gid, orig);
                    W3(FLX_SLOT_SHADOW_DIR, gid, L);
                    F(FLX_SLOT_SHADOW_LEN, gid) = lenL;
                    F(FLX_SLOT_LAST_PDF_DIRECT, gid) = directPdfW;
                    F(FLX_SLOT_LAST_COS_TH, gid) = cosTh;
                    F(FLX_SLOT_LAST_PICK_PROB, gid) = lightPickProb;
                    W3(FLX_SLOT_LAST_EMISSION, gid, envMapLi);
                    push(FLX_Q_SHADOW, &cnt.shadowQueue, gid);
                }
                if (useArea)
                {
                    const float lightPickProb = 1.0f - envMapProb;
                    float directPdfA; f3 posL;
                    sampleAreaLight(params.areaLight, &directPdfA, &posL, &seed);
                    f3 L = posL - orig;
                    const float cosLight = fmaxf(dot(ld3(params.areaLight.N), -L), 0.0f);
                    if (cosLight > 0.0f)
                    {
                        const float lenL0 = length(L);
                        const float lenL = lenL0 * 0.995f;
                        const float invLEN = recip(lenL0);
                        L *= invLEN;
                        const float directPdfW = pdfAtoW(directPdfA, lenL, cosLight * invLEN);
                        const float cosTh = fmaxf(0.0f, dot(L, hit.N));
                        W3(FLX_SLOT_SHADOW_ORIG, gid, orig);
                        W3(FLX_SLOT_SHADOW_DIR, gid, L);
                        F(FLX_SLOT_SHADOW_LEN, gid) = lenL;
                        F(FLX_SLOT_LAST_PDF_DIRECT, gid) = directPdfW;
                        F(FLX_SLOT_LAST_COS_TH, gid) = cosTh;
                        F(FLX_SLOT_LAST_PICK_PROB, gid) = lightPickProb;
                        W3(FLX_SLOT_LAST_EMISSION, gid, ld3(params.areaLight.E));
                        push(FLX_Q_SHADOW, &cnt.shadowQueue, gid);
                    }
                }
            }
        }
        U(FLX_SLOT_SEED, gid) = seed;
        addToMaterialQueue(*this, gid, mat.type);
    }
}

// shared body of wf_mat_diffuse.cl:7-66 / wf_mat_all.cl:5-66
void Wavefront::materialQueue(int q, uint32_t qlen)
{
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < (int64_t)qlen; k++)
    {
        const uint32_t gid = queues[q][k];
        uint32_t seed = U(FLX_SLOT_SEED, gid);
        const Hit hit = readHit(gid);
        const flx_material &mat = shade.materials[hit.matId];
        const bool backface = U(FLX_SLOT_BACKFACE, gid) != 0;
        const f3 dirIn = F3(FLX_SLOT_DIR, gid);
        const f3 L = F3(FLX_SLOT_SHADOW_DIR, gid);

        const f3 bsdfNEE = bxdfEval(hit, mat, backface, shade, dirIn, L);
        const float bsdfPdfW = fmaxf(0.0f, bxdfPdf(hit, mat, backface, shade, dirIn, L));
        W3(FLX_SLOT_LAST_BSDF, gid, bsdfNEE);
        F(FLX_SLOT_LAST_PDF_IMPLICIT, gid) = bsdfPdfW;

        float pdfW = 0.0f;        // uninitialised in the reference; pinned to 0
        f3 newDir = mk3(0.0f);    // ditto
        const f3 bsdf = bxdfSample(hit, mat, backface, shade, dirIn, &newDir, &pdfW, &seed);

        const f3 oldT = F3(FLX_SLOT_T, gid);
        f3 newT;
        if (pdfW == 0.0f || isZero(bsdf)) newT = mk3(0.0f);
        else newT = oldT * bsdf * dot(hit.N, newDir) / pdfW;
        const f3 orig = hit.P + 1e-4f * newDir;

        W3(FLX_SLOT_LAST_T, gid, oldT);
        W3(FLX_SLOT_T, gid, newT);
        W3(FLX_SLOT_ORIG, gid, orig);
        W3(FLX_SLOT_DIR, gid, newDir);
        F(FLX_SLOT_LAST_PDF_W, gid) = pdfW;
        U(FLX_SLOT_SEED, gid) = seed;
        U(FLX_SLOT_LAST_SPECULAR, gid) = FLX_BXDF_IS_SINGULAR(mat.type) ? 1u : 0u;
        push(FLX_Q_EXTENSION, &cnt.extensionQueue, gid);
    }
}

// CLContext::enqueueWfMaterialKernels (clcontext.cpp:816-831)
void Wavefront::materials()
{
    if (!params.wfSeparateQueues) { materialQueue(FLX_Q_DIFFUSE, cnt.diffuseQueue); return; }
    materialQueue(FLX_Q_DIFFUSE, cnt.diffuseQueue);
    materialQueue(FLX_Q_GLOSSY, cnt.glossyQueue);
    materialQueue(FLX_Q_GGX_REFL, cnt.ggxReflQueue);
    materialQueue(FLX_Q_GGX_REFR, cnt.ggxRefrQueue);
    materialQueue(FLX_Q_DELTA, cnt.deltaQueue);
    materialQueue(FLX_Q_EMISSIVE, cnt.emissiveQueue);
}

} // namespace orc
