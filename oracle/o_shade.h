// TEST INFRASTRUCTURE — CPU oracle (see o_math.h).  Shading-side helpers restated from the
// reference's src/random.cl, utils.cl, env_map.cl and the BxDF files.
#pragma once
#include "o_trace.h"

namespace orc {

struct ShadeScene
{
    const flx_triangle *tris;
    const flx_material *materials;
    const flx_tex_desc *textures;
    const uint8_t *texData;
    // environment map as RGBA32F rows (what CLContext::createEnvMap uploads, clcontext.cpp:485-507)
    const float *envRGBA; int envW, envH;
    const float *probTable; const int *aliasTable; const float *pdfTable;
};

// src/random.cl:7-22
static inline uint32_t wangHash(uint32_t seed)
{
    seed = (seed ^ 61u) ^ (seed >> 16);
    seed *= 9u;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2du;
    seed = seed ^ (seed >> 15);
    return seed;
}
static inline float rnd(uint32_t *seed)
{
    *seed = wangHash(*seed);
    return (float)(*seed) * kOne2p32; // may round up to exactly 1.0f
}

static inline float luminance(f3 v) { return 0.212671f * v.x + 0.715160f * v.y + 0.072169f * v.z; }
static inline f3 reflect0(f3 dir, f3 n) { return dir - dot(dir, n) * (n + n); }             // utils.cl:30-33
static inline f3 reflectDot(f3 dir, f3 n, float d) { return dir + (d + d) * n; }             // utils.cl:48-51
static inline f3 refract1(f3 wi, f3 n, float eta, float iDotN, float cosT)                    // utils.cl:63-66
{
    return wi * eta + n * (eta * iDotN - cosT);
}
static inline float pdfAtoW(float pdf, float dist, float cosine) { return pdf * (dist * dist) / fabsf(cosine); }

void makeOrthoBasis(f3 N, f3 *a, f3 *b);
f2 uniformSampleDisk(uint32_t *seed);
f3 cosSampleHemisphere(f3 n, uint32_t *seed, float *p);
f3 readTexture(f2 uv, const flx_tex_desc &tex, const uint8_t *data);
f3 matGetAlbedo(f3 fallback, f2 uv, int idx, const ShadeScene &s);
f3 matGetFloat3(f3 fallback, f2 uv, int idx, const ShadeScene &s);
f3 tangentSpaceNormal(const Hit &hit, const flx_material &mat, const ShadeScene &s);
void sampleAreaLight(const flx_area_light &light, float *pdf, f3 *p, uint32_t *seed);

// env_map.cl
f2 directionToUV(f3 dir);
f3 uvToDirection(float u, float v, float *sinPhi);
f3 envLinear(const ShadeScene &s, float u, float v); // read_imagef(samplerFloat)
f3 evalEnvMapDir(const ShadeScene &s, f3 dir);
void sampleEnvMapAlias(float r, f3 *L, float *pdfW, const ShadeScene &s);
float envMapPdf(const ShadeScene &s, f3 direction);

// fresnel.cl
float fresnelDielectric(float cosThI, float etaI, float etaT);
float fresnelDielectric1(float cosThI, float etaI, float etaT, float *cosThetaT);

// bxdf_partial.cl dispatch
f3 bxdfSample(const Hit &hit, const flx_material &mat, bool backface, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed);
f3 bxdfEval(const Hit &hit, const flx_material &mat, bool backface, const ShadeScene &s, f3 dirIn, f3 dirOut);
float bxdfPdf(const Hit &hit, const flx_material &mat, bool backface, const ShadeScene &s, f3 dirIn, f3 dirOut);

// tonemap.cl + mk_postprocess.cl:27-44 applied to one accumulated pixel
void postprocessPixel(const float in[4], float out[4], float exposure, uint32_t tmOperator);

} // namespace orc
