// The logic stage (path state machine, splat, next-event-estimation sample generation, queue
// compaction) and the material-queue shaders.
// Behavioural contract: reference src/wf_logic.cl:14-344 and src/wf_mat_*.cl (identical bodies).
#include "flx_bxdf.cuh"
#include "flx_raygen.cuh"

#ifndef FLX_WF_THREADS
#define FLX_WF_THREADS 256
#endif
#ifndef FLX_LOGIC_MIN_BLOCKS
#define FLX_LOGIC_MIN_BLOCKS 4
#endif
#ifndef FLX_WF_GRID_PER_SM
#define FLX_WF_GRID_PER_SM 64 // CTAs per SM in the grid of the logic / material kernels (tools/tune_defs.sh: 8 -> 64 is +2 %: finer-grained balance of divergent blocks)
#endif
#ifndef FLX_MAT_MIN_BLOCKS
#define FLX_MAT_MIN_BLOCKS 4
#endif

FLX_DEV uint32_t warpPush(uint32_t *counter, bool pred) // see flx_wavefront.cu
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

// material type -> queue (wf_logic.cl:352-407); MIXED goes with diffuse
// separate: 0 = single queue, 1 = the reference's separate queues (an unknown type falls into `default:` and the path
// is dropped, wf_logic.cl:399-401), 2 = the render entry's internal per-BxDF compaction of single-queue mode (an
// unknown type is shaded to zero by the diffuse body, exactly what wf_mat_all does with it)
FLX_DEV int materialQueueOf(int type, uint32_t separate)
{
    if (!separate) return FLX_Q_DIFFUSE; // -DWF_SINGLE_MAT_QUEUE
    switch (type)
    {
    case FLX_BXDF_DIFFUSE: case FLX_BXDF_MIXED: return FLX_Q_DIFFUSE;
    case FLX_BXDF_GLOSSY: return FLX_Q_GLOSSY;
    case FLX_BXDF_GGX_ROUGH_REFLECTION: return FLX_Q_GGX_REFL;
    case FLX_BXDF_GGX_ROUGH_DIELECTRIC: return FLX_Q_GGX_REFR;
    case FLX_BXDF_IDEAL_REFLECTION: case FLX_BXDF_IDEAL_DIELECTRIC: return FLX_Q_DELTA;
    case FLX_BXDF_EMISSIVE: return FLX_Q_EMISSIVE;
    }
    return separate == 2u ? FLX_Q_DIFFUSE : -1;
}

static __host__ __device__ __forceinline__ uint32_t *counterOf(flx_queue_counters *c, int q) { return ((uint32_t *)c) + q; } // same order as FLX_Q_*

// REGEN = true is the render entry's variant (flx_render): a terminated path is restarted in place — the body
// of the raygen stage runs on the lane that just splatted, with the seed still in a register, and the path
// goes straight into the extension queue — which saves the raygen launch and one round trip of thr/acc
// through HBM.  The raygen-queue counter is still advanced (bookkeeping derives the sample count and the
// pixel cursor from it) and a path's raygen-queue position still selects its pixel, as in wf_raygen.cl:25.
template <bool REGEN>
__global__ void __launch_bounds__(FLX_WF_THREADS, FLX_LOGIC_MIN_BLOCKS)
k_logic(PathState st, Queues qs, FrameBuffers fb, MaterialsDev md, EnvMapDev env, uint32_t N, uint32_t firstIteration,
        int denoiserAOV, const uint32_t *__restrict__ currPixelIdx, uint32_t tileRank, uint32_t tileWorld,
        const __grid_constant__ flx_render_params params)
{
    const uint32_t numPixels = params.width * params.height;
    const uint32_t maxId = firstIteration ? min(numPixels, N) : N;
    const uint32_t rounded = (maxId + FLX_WF_THREADS - 1u) & ~(uint32_t)(FLX_WF_THREADS - 1u); // whole CTAs: the push below has barriers
    const bool checkSpp = params.maxSpp > 0u; // -DCHECK_SPP
    __shared__ uint32_t sPush[FLX_WF_THREADS / 32][16]; // per warp: slot counts, then slot bases, of the ten counters
    const uint32_t separate = params.wfSeparateQueues; // 0 / 1, or 2 from the render entry (see materialQueueOf)
    const uint32_t cursor = REGEN ? *currPixelIdx : 0u;

    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < rounded; gid += gridDim.x * blockDim.x)
    {
        const bool valid = gid < maxId;
        bool terminate = false, pushShadow = false, splatted = false;
        int matQ = -1;
        uint32_t seed = 0u;

        if (valid)
        {
            const float4 thr = st.thr[gid], r0 = st.ray0[gid], r1 = st.ray1[gid];
            const float4 h0 = st.hit0[gid], h1 = st.hit1[gid], h2 = st.hit2[gid], acc = st.acc[gid];
            // the scalar streams and the pending NEE sample are requested together with the rest of the state, so
            // that the whole record costs one trip to HBM instead of one per use further down
            const uint32_t shadowBlocked = st.shadowBlocked[gid], lastSpecular = st.lastSpecular[gid];
            const float lastPickProb = st.lastPickProb[gid];
            f3 neeContrib = mk3(0.0f);
            if (shadowBlocked == 0u) // the NEE sample whose shadow ray came back unoccluded (:150-168)
            {
                const float4 n0 = st.nee0[gid], n1 = st.nee1[gid], n2 = st.nee2[gid];
                const float cosTh = st.shw1[gid].w;
                neeContrib = mk3(n1) * mk3(n2) * mk3(n0) * cosTh / (lastPickProb * n0.w + (float)(params.sampleImpl) * n1.w);
            }
            f3 T = mk3(thr);
            seed = __float_as_uint(thr.w);
            const uint32_t len = __float_as_uint(r1.w);
            const f3 rayOrig = mk3(r0), rayDir = mk3(r1);
            const float lastPdfW = r0.w;
            const f3 hitP = mk3(h0);
            f3 hitN = mk3(h1);
            const int hitI = __float_as_int(h1.w), matId = __float_as_int(h2.z), areaLightHit = __float_as_int(h2.w);
            f3 Ei = mk3(acc);
            const uint32_t pixIdx = __float_as_uint(acc.w);

            // Russian roulette (:59-74)
            terminate = params.maxBounces > 0u && (len >= params.maxBounces + 1u);
            if (!terminate && params.useRoulette && len > 5u)
            {
                const float contProb = clampf(luminance(T), 0.01f, 0.5f);
                terminate = (rnd(seed) > contProb);
                T /= contProb;
            }
            bool maxSamplesReached = false;
            if (checkSpp && fb.spp[pixIdx] >= params.maxSpp) { terminate = true; maxSamplesReached = true; } // :76-84
            if (isZero(T) || lastPdfW == 0.0f) terminate = true;                                             // :87-88

            const bool escaped = hitI < 0 && !terminate;                                        // :99
            const bool lightHit = !escaped && params.useAreaLight && areaLightHit && !terminate; // :126
            const bool survives = !terminate && !escaped && !lightHit;

            // Environment look-ups.  A lane wants one either for an escaped ray (radiance and / or pdf of the ray's own
            // direction, :99-122) or for the NEE sample of a surviving path (radiance of the sampled direction, :262-280);
            // the two exclude each other per lane.  directionToUV (atan2f + acosf) and the bilinear fetch are the longest
            // straight-line code of this kernel and were 27 % of its instructions when each case evaluated them on its own
            // (2 and 14 of 32 lanes per instruction, profiles/r1j_shade_full.md): both kinds of lanes now go through ONE
            // evaluation.  Per lane the order of RNG draws and of the additions into Ei is unchanged.
            bool envBg = false, envPdf = false, neeEnv = false;
            f3 envDir = rayDir;
            float neePdfW = 0.0f;
            if (escaped && params.useEnvMap)
            {
                envBg = len == 1u || params.sampleImpl;
                envPdf = params.sampleImpl && params.sampleExpl && len > 1u && !lastSpecular;
            }

            if (survives)
            {
                // surviving path: shading normal, backface flip (:208-241)
                const flx_material *mp = md.materials + matId;
                const float4 mtail = __ldg(((const float4 *)mp) + 5); // map_Ks, map_N, type, pad
                const int mapN = __float_as_int(mtail.y), mtype = __float_as_int(mtail.z);
                if (mapN != -1) hitN = tangentSpaceNormal(hitN, h2.x, h2.y, hitI, mapN, md);
                const bool backface = dot(hitN, rayDir) > 0.0f;
                if (backface) hitN = -hitN;
                const f3 orig = hitP - 1e-3f * rayDir;
                uint32_t flags = backface ? 1u : 0u;

                if (denoiserAOV) // -DUSE_OPTIX_DENOISER accumulation (:214-237)
                {
                    if (len == 1u)
                    {
                        const f3 c1 = ld3(params.camera.right), c2 = ld3(params.camera.up), c3 = -ld3(params.camera.dir);
                        atomicAdd(&fb.denoiserNormal[pixIdx], make_float4(dot(c1, hitN), dot(c2, hitN), dot(c3, hitN), 1.0f));
                    }
                    const uint32_t firstDiffuse = (st.logicFlags[gid] >> 1) & 1u;
                    flags |= firstDiffuse << 1;
                    if (!FLX_BXDF_IS_SINGULAR(mtype) && !firstDiffuse)
                    {
                        flags |= 2u;
                        const float4 kd = __ldg((const float4 *)mp), m4 = __ldg(((const float4 *)mp) + 4);
                        const f3 albedo = matGetFloat3(mk3(kd), h2.x, h2.y, __float_as_int(m4.w), md);
                        atomicAdd(&fb.denoiserAlbedo[pixIdx], make_float4(albedo.x, albedo.y, albedo.z, 1.0f));
                    }
                }
                st.hit1[gid] = mk4i(hitN, hitI);
                st.logicFlags[gid] = flags;

                if (params.sampleExpl) // -DSAMPLE_EXPLICIT (:243-333)
                {
                    st.shadowBlocked[gid] = 1u;
                    if (!FLX_BXDF_IS_SINGULAR(mtype))
                    {
                        const uint32_t nLights = params.useEnvMap + params.useAreaLight;
                        const float envMapProb = (float)params.useEnvMap / (float)max(1u, nLights);
                        const bool useEnv = rnd(seed) < envMapProb;
                        const bool useArea = !useEnv && params.useAreaLight;
                        if (useEnv && params.useEnvMap)
                        {
                            f3 L; float directPdfW;
                            sampleEnvMapAlias(rnd(seed), L, directPdfW, env);
                            const float lenL = params.worldRadius + params.worldRadius;
                            const float cosTh = fmaxf(0.0f, dot(L, hitN));
                            st.shw0[gid] = mk4(orig, lenL);
                            st.shw1[gid] = mk4(L, cosTh);
                            st.lastPickProb[gid] = envMapProb;
                            pushShadow = true;
                            neeEnv = true; envDir = L; neePdfW = directPdfW; // radiance of L: shared evaluation below
                        }
                        if (useArea)
                        {
                            float directPdfA; f3 posL;
                            sampleAreaLight(params.areaLight, directPdfA, posL, seed);
                            f3 L = posL - orig;
                            const float cosLight = fmaxf(dot(ld3(params.areaLight.N), -L), 0.0f);
                            if (cosLight > 0.0f) // only samples that see the emitting side
                            {
                                const float lenL0 = length(L);
                                const float lenL = lenL0 * 0.995f;
                                const float invLEN = recip(lenL0);
                                L *= invLEN;
                                const float directPdfW = pdfAtoW(directPdfA, lenL, cosLight * invLEN);
                                const float cosTh = fmaxf(0.0f, dot(L, hitN));
                                st.shw0[gid] = mk4(orig, lenL);
                                st.shw1[gid] = mk4(L, cosTh);
                                st.nee0[gid] = mk4(ld3(params.areaLight.E), directPdfW);
                                st.lastPickProb[gid] = 1.0f - envMapProb;
                                pushShadow = true;
                            }
                        }
                    }
                }
                matQ = materialQueueOf(mtype, separate);
            }

            f3 envLi = mk3(0.0f);
            float envWeight = 1.0f;
            if (envBg || envPdf || neeEnv) // the shared environment evaluation
            {
                float u, v;
                directionToUV(envDir, u, v);
                if (envBg || neeEnv) envLi = envLinear(env, u, v) * params.envMapStrength;
                if (envPdf)
                {
                    const float lightPickProb = lastPickProb;
                    const float directPdfW = envMapPdfUV(env, rayDir, u, v);
                    envWeight = (lastPdfW * lightPickProb) / (lastPdfW * lightPickProb + directPdfW);
                }
                if (neeEnv) st.nee0[gid] = mk4(envLi, neePdfW);
            }
            if (escaped) // implicit environment sample with MIS (:99-122)
            {
                Ei = Ei + envWeight * T * envLi;
                terminate = true;
            }
            else if (lightHit) // implicit area-light hit (:126-146)
            {
                float misWeight = 1.0f;
                if (params.sampleExpl && len > 1u && !lastSpecular)
                {
                    const float directPdfA = recip(4.0f * params.areaLight.size.x * params.areaLight.size.y);
                    const float directPdfW = pdfAtoW(directPdfA, length(hitP - rayOrig), -dot(rayDir, hitN));
                    const float lightPickProb = lastPickProb;
                    misWeight = lastPdfW / (lastPdfW + directPdfW * lightPickProb);
                }
                Ei = Ei + T * misWeight * ld3(params.areaLight.E);
                terminate = true;
            }

            if (shadowBlocked == 0u) Ei = Ei + neeContrib; // resolved after the implicit terms, as the reference orders it

            if (terminate) // splat and hand the slot back to raygen (:171-205)
            {
                bool admit;
                if (checkSpp)
                {
                    admit = false;
                    if (!maxSamplesReached) admit = (atomicAdd(&fb.spp[pixIdx], 1u) <= params.maxSpp) && len > 0u; // `<=`: quirk kept
                    // The reference clamps the counter with a plain read + write (wf_logic.cl:186-189): a benign race, kept.
                    // Paths that passed the first check above before the clamp set the counter back to maxSpp are admitted
                    // again by the `<=`, so a pixel many paths finish on at once can end with maxSpp + 2 (rarely more)
                    // samples instead of maxSpp + 1; splattedSamples counts every one of them, the image stays unbiased.
                    if (fb.spp[pixIdx] > params.maxSpp) fb.spp[pixIdx] = params.maxSpp;
                    splatted = admit;
                }
                else admit = len > 0u;
                if (admit) atomicAdd(&fb.pixels[pixIdx], make_float4(Ei.x, Ei.y, Ei.z, 1.0f));
            }
            if (!(REGEN && terminate)) // a regenerated path overwrites both below
            {
                st.thr[gid] = mk4u(T, seed);
                st.acc[gid] = mk4u(Ei, pixIdx);
            }
        }

        // ---- warp-aggregated compaction into the raygen / shadow / material queues ----------------
        // One ballot per queue, then lane q reserves the slots of queue q: the (up to nine) atomicAdds of a
        // warp are issued by different lanes in ONE instruction, i.e. one round trip to L2 instead of one per
        // queue (the per-queue version spent half of the kernel's stall samples waiting on them one after the
        // other).  Counter q of QueueCounters belongs to queue q (same order), word 9 is splattedSamples.
        {
            const unsigned FULL = 0xffffffffu;
            const unsigned lane = threadIdx.x & 31u;
            const int myQ = (valid && terminate) ? FLX_Q_RAYGEN : matQ; // a path is regenerated or shaded, never both
            unsigned mine = 0u;                                          // ballot of the queue this lane reserves for
#pragma unroll
            for (int q = 0; q < FLX_NUM_QUEUES; q++)
            {
                if (q == FLX_Q_EXTENSION) continue;
                if (!separate && q > FLX_Q_DIFFUSE) continue;
                const unsigned m = __ballot_sync(FULL, q == FLX_Q_SHADOW ? pushShadow : (myQ == q));
                if (lane == (unsigned)q) mine = m;
                if (REGEN && q == FLX_Q_RAYGEN && lane == (unsigned)FLX_Q_EXTENSION) mine = m; // restarted paths
            }
            const unsigned mSplat = checkSpp ? __ballot_sync(FULL, splatted) : 0u;
            if (lane == 9u) mine = mSplat;
            // CTA-level aggregation: the ten counters share one 32-byte sector, and the L2 atomic unit serialises the
            // operations on it — one request per warp (131 k per launch at 2^22 paths) kept that unit busy for most of the
            // kernel.  The eight warps of a CTA pool their counts in shared memory and warp 0 reserves for all of them.
            const unsigned warp = threadIdx.x >> 5;
            if (lane < 10u) sPush[warp][lane] = (uint32_t)__popc(mine);
            __syncthreads();
            if (warp == 0u && lane < 10u)
            {
                uint32_t total = 0u;
#pragma unroll
                for (int w = 0; w < FLX_WF_THREADS / 32; w++) { const uint32_t c = sPush[w][lane]; sPush[w][lane] = total; total += c; }
                const uint32_t b = total ? atomicAdd(((uint32_t *)qs.counters) + lane, total) : 0u;
#pragma unroll
                for (int w = 0; w < FLX_WF_THREADS / 32; w++) sPush[w][lane] += b;
            }
            __syncthreads();
            const uint32_t base = lane < 10u ? sPush[warp][lane] : 0u;
            __syncwarp();
            const unsigned lt = (1u << lane) - 1u;
            // every lane asks the reserving lane of its queue for the base and the mask (two shuffles per push)
            const int srcA = myQ >= 0 ? myQ : 0;
            const uint32_t baseA = __shfl_sync(FULL, base, srcA);
            const unsigned maskA = __shfl_sync(FULL, mine, srcA);
            const uint32_t slotA = baseA + (uint32_t)__popc(maskA & lt);
            if (REGEN)
            {
                const uint32_t baseE = __shfl_sync(FULL, base, FLX_Q_EXTENSION);
                if (myQ == FLX_Q_RAYGEN)
                {
                    generateCameraPath(st, gid, seed, cursor, slotA, tileRank, tileWorld, params);
                    qs.q[FLX_Q_EXTENSION][baseE + (uint32_t)__popc(maskA & lt)] = gid;
                }
                else if (myQ >= 0) (qs.q[0] + (size_t)myQ * N)[slotA] = gid;
            }
            else if (myQ >= 0) (qs.q[0] + (size_t)myQ * N)[slotA] = gid; // queues are contiguous, N apart
            const uint32_t baseS = __shfl_sync(FULL, base, FLX_Q_SHADOW);
            const unsigned maskS = __shfl_sync(FULL, mine, FLX_Q_SHADOW);
            if (pushShadow) qs.q[FLX_Q_SHADOW][baseS + (uint32_t)__popc(maskS & lt)] = gid;
        }
    }
}

// Shared body of the material stages (wf_mat_diffuse.cl:7-66 and its five siblings, wf_mat_all.cl).
// TYPES is the compile-time set of BxDFs this instantiation can meet in its queue.  Called by whole warps:
// lane `k` of the queue (k may run past qlen in the last warp), then one aggregated push of the survivors.
template <int TYPES>
FLX_DEV void materialShade(const PathState &st, uint32_t gid, const MaterialsDev &md)
{
    const float4 thr = st.thr[gid], r1 = st.ray1[gid], h0 = st.hit0[gid], h1 = st.hit1[gid], h2 = st.hit2[gid];
    uint32_t seed = __float_as_uint(thr.w);
    SurfHit hit; hit.N = mk3(h1); hit.uvx = h2.x; hit.uvy = h2.y;
    const Mat mat = loadMat(md.materials + __float_as_int(h2.z));
    const bool backface = (st.logicFlags[gid] & 1u) != 0u;
    const f3 dirIn = mk3(r1);
    const f3 L = mk3(st.shw1[gid]);

    const f3 bsdfNEE = bxdfEval<TYPES>(hit, mat, backface, md, dirIn, L);
    const float bsdfPdfW = fmaxf(0.0f, bxdfPdf<TYPES>(hit, mat, backface, md, dirIn, L));

    float pdfW = 0.0f;     // the reference leaves these unset for emissive / rejected glossy
    f3 newDir = mk3(0.0f); // samples; the oracle pins them to 0 and so do we
    const f3 bsdf = bxdfSample<TYPES>(hit, mat, backface, md, dirIn, newDir, pdfW, seed);

    const f3 oldT = mk3(thr);
    f3 newT;
    if (pdfW == 0.0f || isZero(bsdf)) newT = mk3(0.0f);
    else newT = oldT * bsdf * dot(hit.N, newDir) / pdfW;
    const f3 orig = mk3(h0) + 1e-4f * newDir;

    st.nee1[gid] = mk4(bsdfNEE, bsdfPdfW);
    st.nee2[gid] = mk4(oldT, 0.0f);
    st.thr[gid] = mk4u(newT, seed);
    st.ray0[gid] = mk4(orig, pdfW);
    st.ray1[gid] = make_float4(newDir.x, newDir.y, newDir.z, r1.w);
    st.lastSpecular[gid] = FLX_BXDF_IS_SINGULAR(mat.type) ? 1u : 0u;
}

template <int TYPES>
FLX_DEV void materialBody(const PathState &st, const uint32_t *__restrict__ queue, uint32_t qlen, uint32_t k,
                          uint32_t *__restrict__ extQueue, uint32_t *extCounter, const MaterialsDev &md)
{
    const bool valid = k < qlen;
    uint32_t gid = 0;
    if (valid)
    {
        gid = queue[k];
        materialShade<TYPES>(st, gid, md);
    }
    const uint32_t slot = warpPush(extCounter, valid);
    if (valid) extQueue[slot] = gid;
}

// one queue, one launch: wf_mat_all.cl (no separate queues) and the building block of the tests
template <int TYPES>
__global__ void __launch_bounds__(FLX_WF_THREADS, FLX_MAT_MIN_BLOCKS)
k_material(PathState st, const uint32_t *__restrict__ queue, const uint32_t *__restrict__ queueLen,
           uint32_t *__restrict__ extQueue, uint32_t *extCounter, MaterialsDev md)
{
    const uint32_t qlen = *queueLen;
    const uint32_t rounded = (qlen + 31u) & ~31u;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < rounded; k += gridDim.x * blockDim.x)
        materialBody<TYPES>(st, queue, qlen, k, extQueue, extCounter, md);
}

// The six material queues of CLContext::enqueueWfMaterialKernels (clcontext.cpp:816-831) in ONE launch: the
// queues are laid end to end in units of 32-path chunks and every warp takes chunks round robin, running the
// specialised body of the queue its chunk belongs to.  Most queues are short (a few thousand paths), so six
// launches spend their time ramping up and draining; together they fill the machine once.
// Extension-queue slots without atomics: every shaded path continues with exactly one extension ray
// (wf_mat_diffuse.cl:60-65), so path k of queue i owns slot extBefore + len(queues before i) + k, where extBefore
// is the counter as the logic stage left it.  The counter itself is advanced once, by the CTA that finishes last
// (every other CTA has read extBefore by then) — the reference's one atomic per path, and the one per warp of the
// first version here (15 % of this kernel's stall samples: they all hit one L2 word), are gone.
__global__ void __launch_bounds__(FLX_WF_THREADS, FLX_MAT_MIN_BLOCKS)
k_materials_separate(PathState st, Queues qs, uint32_t N, MaterialsDev md, unsigned int *arrive)
{
    const uint32_t *cnt = (const uint32_t *)qs.counters; // word q = length of queue q
    const uint32_t extBefore = cnt[FLX_Q_EXTENSION];
    uint32_t end[6], pre[6];                              // chunk index one past queue FLX_Q_DIFFUSE + i; paths in the queues before it
    uint32_t acc = 0u, total = 0u;
#pragma unroll
    for (int i = 0; i < 6; i++) { const uint32_t n = cnt[FLX_Q_DIFFUSE + i]; pre[i] = total; total += n; acc += (n + 31u) >> 5; end[i] = acc; }
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = (gridDim.x * blockDim.x) >> 5;
    uint32_t *extQueue = qs.q[FLX_Q_EXTENSION];
    for (uint32_t chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; chunk < acc; chunk += warpsTotal)
    {
        int i = 0;
        uint32_t first = 0u, before = 0u;
#pragma unroll
        for (int j = 0; j < 5; j++) if (chunk >= end[j]) { i = j + 1; first = end[j]; before = pre[j + 1]; }
        const uint32_t k = ((chunk - first) << 5) + lane;
        const uint32_t *queue = qs.q[0] + (size_t)(FLX_Q_DIFFUSE + i) * N; // queues are contiguous, N apart
        if (k < cnt[FLX_Q_DIFFUSE + i])
        {
            const uint32_t gid = queue[k];
            switch (i) // warp-uniform
            {
            case 0: materialShade<FLX_BXDF_DIFFUSE>(st, gid, md); break;
            case 1: materialShade<FLX_BXDF_GLOSSY>(st, gid, md); break;
            case 2: materialShade<FLX_BXDF_GGX_ROUGH_REFLECTION>(st, gid, md); break;
            case 3: materialShade<FLX_BXDF_GGX_ROUGH_DIELECTRIC>(st, gid, md); break;
            case 4: materialShade<FLX_BXDF_IDEAL_REFLECTION | FLX_BXDF_IDEAL_DIELECTRIC>(st, gid, md); break;
            default: materialShade<FLX_BXDF_EMISSIVE>(st, gid, md); break;
            }
            extQueue[extBefore + before + k] = gid;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0u)
    {
        __threadfence();
        if (atomicAdd(arrive, 1u) == gridDim.x - 1u)
        {
            qs.counters->extensionQueue = extBefore + total;
            *arrive = 0u;
        }
    }
}

// ---- launchers -----------------------------------------------------------------------------------
static inline int wfGrid(const flx_ctx *ctx, uint64_t items)
{
    const uint64_t blocks = (items + FLX_WF_THREADS - 1) / FLX_WF_THREADS;
    const uint64_t cap = (uint64_t)ctx->numSMs * FLX_WF_GRID_PER_SM;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}
#define FLX_LAUNCHED(ctx) do { (ctx)->launches++; FLX_CUDA(ctx, cudaGetLastError()); } while (0)

int flx_launch_logic(flx_ctx *ctx, int firstIteration)
{
    k_logic<false><<<wfGrid(ctx, ctx->N), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs, ctx->fb, ctx->mats, ctx->env, ctx->N,
        (uint32_t)(firstIteration != 0), ctx->denoiserAOV, ctx->dPixelIdx, ctx->tileRank, ctx->tileWorld, ctx->params);
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

// logic + raygen in one launch (the render entry's inner loop; the per-stage API keeps them apart).
// The render entry always compacts into the per-BxDF queues, whatever params.wfSeparateQueues says: which queue a
// path waits in between two kernels is invisible from outside flx_render (the host sees totals, not queues), every
// path is shaded by the same arithmetic either way (materialBody<TYPES> is wf_mat_all's body minus the branches a
// queue cannot meet), and the reference's default single queue (useSeparateQueues = false, src/settings.cpp:31) costs
// 4x in the material stage on a GPU because every warp then walks through all BxDF branches.  The per-stage calls
// (flx_wf_logic / flx_wf_materials) honour the flag: the lock-step parity tests compare the queues themselves.
static inline flx_render_params renderEntryParams(const flx_ctx *ctx)
{
    flx_render_params p = ctx->params;
#ifndef FLX_RENDER_SINGLE_QUEUE_AS_IS
    p.wfSeparateQueues = p.wfSeparateQueues ? 1u : 2u;
#endif
    return p;
}

int flx_launch_logic_regen(flx_ctx *ctx)
{
    k_logic<true><<<wfGrid(ctx, ctx->N), FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs, ctx->fb, ctx->mats, ctx->env, ctx->N,
        0u, ctx->denoiserAOV, ctx->dPixelIdx, ctx->tileRank, ctx->tileWorld, renderEntryParams(ctx));
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}

#define FLX_ALL_TYPES (FLX_BXDF_DIFFUSE | FLX_BXDF_GLOSSY | FLX_BXDF_GGX_ROUGH_REFLECTION | FLX_BXDF_IDEAL_REFLECTION | \
                       FLX_BXDF_GGX_ROUGH_DIELECTRIC | FLX_BXDF_IDEAL_DIELECTRIC | FLX_BXDF_EMISSIVE)

// CLContext::enqueueWfMaterialKernels (clcontext.cpp:816-831)
int flx_launch_materials(flx_ctx *ctx, bool renderEntry)
{
    const int grid = wfGrid(ctx, ctx->N);
    if (!(renderEntry ? renderEntryParams(ctx) : ctx->params).wfSeparateQueues)
    {
        k_material<FLX_ALL_TYPES><<<grid, FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs.q[FLX_Q_DIFFUSE], counterOf(ctx->qs.counters, FLX_Q_DIFFUSE), ctx->qs.q[FLX_Q_EXTENSION], &ctx->qs.counters->extensionQueue, ctx->mats);
        FLX_LAUNCHED(ctx);
        return FLX_OK;
    }
    static_assert(FLX_Q_GLOSSY == FLX_Q_DIFFUSE + 1 && FLX_Q_GGX_REFL == FLX_Q_DIFFUSE + 2 && FLX_Q_GGX_REFR == FLX_Q_DIFFUSE + 3 &&
                  FLX_Q_DELTA == FLX_Q_DIFFUSE + 4 && FLX_Q_EMISSIVE == FLX_Q_DIFFUSE + 5, "material queues are consecutive");
    k_materials_separate<<<grid, FLX_WF_THREADS, 0, ctx->stream>>>(ctx->st, ctx->qs, ctx->N, ctx->mats, (unsigned int *)(ctx->dPixelIdx + 5)); // arrival word: zero at rest
    FLX_LAUNCHED(ctx);
    return FLX_OK;
}
