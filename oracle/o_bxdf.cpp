// TEST INFRASTRUCTURE — CPU oracle.  Restates the reference BxDF library: src/diffuse.cl,
// glossy.cl, ggx.cl, ideal_reflection.cl, ideal_dielectric.cl and the dispatch of
// bxdf_partial.cl (with every BXDF_USE_* switch on, as wf_mat_all compiles it for a scene that
// contains all types; a type absent from a scene is never dispatched).
#include "o_shade.h"

namespace orc {

struct Mat // mutable register copy, as glossy.cl:28 does (`Material m = *mat`)
{
    f3 Kd, Ks, Ke, Kt; float Ns, Ni, d; int map_Kd, map_Ks, map_N, type;
};
static inline Mat loadMat(const flx_material &m)
{
    Mat r; r.Kd = ld3(m.Kd); r.Ks = ld3(m.Ks); r.Ke = ld3(m.Ke); r.Kt = ld3(m.Kt);
    r.Ns = m.Ns; r.Ni = m.Ni; r.d = m.d; r.map_Kd = m.map_Kd; r.map_Ks = m.map_Ks; r.map_N = m.map_N; r.type = m.type;
    return r;
}

// ---- diffuse.cl:9-26 ------------------------------------------------------------------------
static f3 sampleDiffuse(const Hit &hit, const Mat &m, const ShadeScene &s, f3 *dirOut, float *pdfW, uint32_t *seed)
{
    *dirOut = cosSampleHemisphere(hit.N, seed, pdfW);
    return matGetAlbedo(m.Kd, hit.uv, m.map_Kd, s) * kInvPI;
}
static f3 evalDiffuse(const Hit &hit, const Mat &m, const ShadeScene &s)
{
    return matGetAlbedo(m.Kd, hit.uv, m.map_Kd, s) * kInvPI;
}
static float pdfDiffuse(const Hit &hit, f3 dirOut) { return dot(hit.N, dirOut) * kInvPI; }

// ---- ggx.cl ---------------------------------------------------------------------------------
static f3 ggxSampleLobe(float alpha, f3 N, uint32_t *seed) // :20-37
{
    f3 X, Y; const f3 Z = N;
    makeOrthoBasis(Z, &X, &Y);
    const float rx = rnd(seed), ry = rnd(seed);
    const float theta = o_atan2f(alpha * sqrtf(rx), sqrtf(1 - rx));
    const float phi = k2PI * ry;
    const float sinTheta = o_sinf(theta), cosTheta = o_cosf(theta);
    const float sinPhi = o_sinf(phi), cosPhi = o_cosf(phi);
    return (X * sinTheta * cosPhi + Y * sinTheta * sinPhi + Z * cosTheta);
}
static float ggxG1(float alpha, f3 v, f3 n, f3 m) // :41-54
{
    const float mDotV = dot(m, v), nDotV = dot(n, v);
    if (nDotV * mDotV <= 0.0f) return 0.0f;
    const float cosThSq = nDotV * nDotV;
    const float tanSq = (cosThSq > 0.0f) ? ((1.0f - cosThSq) / cosThSq) : 0.0f;
    return 2.0f / (1.0f + sqrtf(1.0f + alpha * alpha * tanSq));
}
static float ggxG(float alpha, f3 dirIn, f3 dirOut, f3 n, f3 m) { return ggxG1(alpha, dirIn, n, m) * ggxG1(alpha, dirOut, n, m); }
static float ggxD(float alpha, f3 n, f3 m) // :65-78
{
    const float nDotM = dot(n, m);
    if (nDotM <= 0.0f) return 0.0f;
    const float nDotMSq = nDotM * nDotM;
    const float tanSq = ((1.0f - nDotMSq) / nDotMSq);
    const float aSq = alpha * alpha;
    const float denom = kPI * nDotMSq * nDotMSq * (aSq + tanSq) * (aSq + tanSq);
    return (aSq / denom);
}
static float ggxPdfReflect(float alpha, f3 dirOut, f3 N, f3 H) // :81-89
{
    const float nDotH = dot(N, H), oDotH = dot(dirOut, H);
    return oDotH == 0.0f ? 0.0f : ggxD(alpha, N, H) * fabsf(nDotH * 0.25f / oDotH);
}
static f3 ggxReflTerm(const Hit &hit, const Mat &m, const ShadeScene &s, f3 dirInN, f3 dirOut, f3 H, float alpha)
{
    const float iDotN = dot(dirInN, hit.N), oDotN = dot(dirOut, hit.N);
    const float F = (m.Ni > 1.0f) ? fresnelDielectric(iDotN, 1.0f, m.Ni) : 1.0f;
    const f3 Ks = matGetFloat3(m.Ks, hit.uv, m.map_Ks, s);
    const float D = ggxD(alpha, hit.N, H);
    const float G = ggxG(alpha, dirInN, dirOut, hit.N, H);
    const float den = (iDotN * oDotN);
    return (den != 0.0f) ? (Ks * F * G * D * 0.25f / den) : mk3(0.0f);
}
static f3 sampleGGXReflect(const Hit &hit, const Mat &m, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed) // :91-114
{
    const float alpha = m.Ns;
    const f3 dirInN = -dirIn;
    const f3 H = ggxSampleLobe(alpha, hit.N, seed);
    *dirOut = reflect0(dirIn, H);
    *pdfW = ggxPdfReflect(alpha, *dirOut, hit.N, H);
    return ggxReflTerm(hit, m, s, dirInN, *dirOut, H, alpha);
}
static f3 evalGGXReflect(const Hit &hit, const Mat &m, const ShadeScene &s, f3 dirIn, f3 dirOut) // :116-136
{
    const float alpha = m.Ns;
    const f3 dirInN = -dirIn;
    const f3 H = normalize(dirOut - dirIn);
    return ggxReflTerm(hit, m, s, dirInN, dirOut, H, alpha);
}
static float pdfGGXReflect(const Hit &hit, const Mat &m, f3 dirIn, f3 dirOut) // :138-143
{
    const f3 H = normalize(dirOut - dirIn);
    return ggxPdfReflect(m.Ns, dirOut, hit.N, H);
}
static float ggxPdfRefract(float alpha, float etaI, float etaO, f3 dirIn, f3 dirOut, f3 N, f3 H) // :146-154
{
    const float nDotH = fabsf(dot(N, H)), iDotH = fabsf(dot(dirIn, H)), oDotH = fabsf(dot(dirOut, H));
    const float sqrtJInv = etaI * iDotH + etaO * oDotH;
    return sqrtJInv == 0.0f ? 0.0f : ggxD(alpha, N, H) * nDotH * oDotH * etaO * etaO / (sqrtJInv * sqrtJInv);
}
static f3 sampleGGXRefract(const Hit &hit, const Mat &m, bool backface, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed) // :156-222
{
    const f3 dirInN = -dirIn;
    const float alpha = m.Ns;
    float etaI = 1.0f, etaO = m.Ni;
    if (backface) { const float t = etaI; etaI = etaO; etaO = t; }
    const float iDotN = dot(dirInN, hit.N);
    f3 H = ggxSampleLobe(alpha, hit.N, seed);
    float cosThetaT;
    const float F = fresnelDielectric1(iDotN, etaI, etaO, &cosThetaT);
    if (rnd(seed) < F)
    {
        *dirOut = reflect0(dirIn, H);
        *pdfW = ggxPdfReflect(alpha, *dirOut, hit.N, H);
        const float oDotN = dot(*dirOut, hit.N);
        const float D = ggxD(alpha, hit.N, H);
        const float G = ggxG(alpha, dirInN, *dirOut, hit.N, H);
        const float den = (iDotN * oDotN);
        return mk3((den != 0.0f) ? (F * G * D * 0.25f / den) : 0.0f);
    }
    const float eta = etaI / etaO;
    *dirOut = refract1(dirIn, hit.N, eta, iDotN, cosThetaT);
    H = normalize((dirIn * etaI - *dirOut * etaO));
    const f3 Nn = backface ? -hit.N : hit.N;
    *pdfW = ggxPdfRefract(alpha, etaI, etaO, dirInN, *dirOut, Nn, H);
    f3 bsdf = mk3(eta * eta);
    bsdf *= matGetFloat3(m.Ks, hit.uv, m.map_Ks, s);
    const float iDotH = fabsf(dot(dirIn, H)), oDotH = fabsf(dot(*dirOut, H));
    const float oDotN = dot(*dirOut, hit.N);
    const float focusTermDenom = iDotN * oDotN * (etaI * iDotH + etaO * oDotH) * (etaI * iDotH + etaO * oDotH);
    if (focusTermDenom == 0.0f) return mk3(0.0f);
    const float focusTerm = etaO * etaO * iDotH * oDotH / focusTermDenom;
    const float D = ggxD(alpha, Nn, H);
    const float G = ggxG(alpha, dirInN, *dirOut, Nn, H);
    return (1.0f - F) * bsdf * D * G * focusTerm;
}
static f3 evalGGXRefract(const Hit &hit, const Mat &m, bool backface, const ShadeScene &s, f3 dirIn, f3 dirOut) // :224-275
{
    const f3 dirInN = -dirIn;
    const float alpha = m.Ns;
    float etaI = 1.0f, etaO = m.Ni;
    if (backface) { const float t = etaI; etaI = etaO; etaO = t; }
    const float iDotN = dot(dirInN, hit.N), oDotN = dot(dirOut, hit.N);
    const float F = fresnelDielectric(iDotN, etaI, etaO);
    if (!backface)
    {
        const f3 H = normalize(dirOut - dirIn);
        const float D = ggxD(alpha, hit.N, H);
        const float G = ggxG(alpha, dirInN, dirOut, hit.N, H);
        const float den = (iDotN * oDotN);
        return (den != 0.0f) ? mk3(F * G * D * 0.25f / den) : mk3(0.0f);
    }
    const f3 H = normalize((dirIn * etaI - dirOut * etaO));
    const float eta = etaI / etaO;
    f3 bsdf = mk3(eta * eta);
    bsdf *= matGetFloat3(m.Ks, hit.uv, m.map_Ks, s);
    const float iDotH = fabsf(dot(dirIn, H)), oDotH = fabsf(dot(dirOut, H));
    const float focusTermDenom = iDotN * oDotN * (etaI * iDotH + etaO * oDotH) * (etaI * iDotH + etaO * oDotH);
    if (focusTermDenom == 0.0f) return mk3(0.0f);
    const float focusTerm = etaO * etaO * iDotH * oDotH / focusTermDenom;
    const float D = ggxD(alpha, -hit.N, H);
    const float G = ggxG(alpha, dirInN, dirOut, -hit.N, H);
    return (1.0f - F) * bsdf * D * G * focusTerm;
}
static float pdfGGXRefract(const Hit &hit, const Mat &m, bool backface, f3 dirIn, f3 dirOut) // :277-296
{
    const float alpha = m.Ns;
    float etaI = 1.0f, etaO = m.Ni;
    if (!backface)
    {
        const f3 H = normalize(dirOut - dirIn);
        return ggxPdfReflect(alpha, dirOut, hit.N, H);
    }
    { const float t = etaI; etaI = etaO; etaO = t; }
    const f3 H = normalize((dirIn * etaI - dirOut * etaO));
    return ggxPdfRefract(alpha, etaI, etaO, -dirIn, dirOut, -hit.N, H);
}

// ---- glossy.cl ------------------------------------------------------------------------------
static inline f3 etaToKs(float eta) { const float r = (eta > 0.0f) ? ((eta - 1) / (eta + 1)) : 0.0f; return mk3(r * r); }
static inline float ksToEta(f3 Ks)
{
    const float k = clampf((Ks.x + Ks.y + Ks.z) / 3.0f, 0.0f, 0.99f);
    return (sqrtf(k) + 1) / (1 - sqrtf(k));
}
static f3 sampleGlossy(const Hit &hit, const Mat &mat, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed) // :22-64
{
    Mat m = mat;
    m.Ks = matGetFloat3(mat.Ks, hit.uv, mat.map_Ks, s);
    m.Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(m.Ks);
    if (isZero(m.Ks)) m.Ks = etaToKs(m.Ni);
    const float cosTh = -dot(dirIn, hit.N);
    const float F = fresnelDielectric(cosTh, 1.0f, m.Ni);
    float basePdf, coatingPdf;
    f3 baseBrdf, coatingBrdf;
    if (rnd(seed) < F)
    {
        coatingBrdf = sampleGGXReflect(hit, m, s, dirIn, dirOut, &coatingPdf, seed);
        baseBrdf = evalDiffuse(hit, m, s);
        basePdf = pdfDiffuse(hit, *dirOut);
    }
    else
    {
        baseBrdf = sampleDiffuse(hit, m, s, dirOut, &basePdf, seed);
        coatingBrdf = evalGGXReflect(hit, m, s, dirIn, *dirOut);
        coatingPdf = pdfGGXReflect(hit, m, dirIn, *dirOut);
    }
    if (dot(hit.N, *dirOut) < 1e-5f) return mk3(0.0f); // pdfW left unwritten by the reference here
    *pdfW = (1 - F) * basePdf + F * coatingPdf;
    return (baseBrdf * (1 - F) + coatingBrdf);
}
static f3 evalGlossy(const Hit &hit, const Mat &mat, const ShadeScene &s, f3 dirIn, f3 dirOut) // :66-86
{
    Mat m = mat;
    m.Ks = matGetFloat3(mat.Ks, hit.uv, mat.map_Ks, s);
    m.Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(m.Ks);
    if (length(m.Ks) == 0.0f) m.Ks = etaToKs(m.Ni);
    const f3 baseBrdf = evalDiffuse(hit, m, s);
    const f3 coatingBrdf = evalGGXReflect(hit, m, s, dirIn, dirOut);
    const float cosTh = -dot(dirIn, hit.N);
    const float F = fresnelDielectric(cosTh, 1.0f, m.Ni);
    return (baseBrdf * (1 - F) + coatingBrdf);
}
static float pdfGlossy(const Hit &hit, const Mat &mat, const ShadeScene &s, f3 dirIn, f3 dirOut) // :88-101
{
    const f3 Ks = matGetFloat3(mat.Ks, hit.uv, mat.map_Ks, s);
    const float Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(Ks);
    const float basePdf = pdfDiffuse(hit, dirOut);
    const float coatingPdf = pdfGGXReflect(hit, mat, dirIn, dirOut);
    const float cosTh = -dot(dirIn, hit.N);
    const float F = fresnelDielectric(cosTh, 1.0f, Ni);
    return (1 - F) * basePdf + F * coatingPdf;
}

// ---- ideal_reflection.cl:9-21, ideal_dielectric.cl:10-44 --------------------------------------
static f3 sampleIdealReflection(const Hit &hit, const Mat &m, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW)
{
    *dirOut = reflect0(dirIn, hit.N);
    *pdfW = 1.0f;
    const f3 ks = matGetFloat3(m.Ks, hit.uv, m.map_Ks, s);
    const float cosO = dot(normalize(*dirOut), hit.N);
    return (cosO != 0.0f) ? ks / cosO : mk3(0.0f);
}
static f3 sampleIdealDielectric(const Hit &hit, const Mat &m, bool backface, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed)
{
    f3 bsdf = mk3(1.0f);
    const float cosI = -dot(dirIn, hit.N);
    float n1 = 1.0f, n2 = m.Ni;
    if (backface) { const float t = n1; n1 = n2; n2 = t; }
    const float eta = n1 / n2;
    float cosThetaT;
    const float fr = fresnelDielectric1(cosI, n1, n2, &cosThetaT);
    if (rnd(seed) < fr)
    {
        *dirOut = reflectDot(dirIn, hit.N, cosI);
    }
    else
    {
        *dirOut = refract1(dirIn, hit.N, eta, cosI, cosThetaT);
        bsdf *= eta * eta;
        bsdf *= matGetFloat3(m.Ks, hit.uv, m.map_Ks, s);
    }
    *pdfW = 1.0f;
    const float cosO = dot(normalize(*dirOut), hit.N);
    return bsdf / cosO;
}

// ---- bxdf_partial.cl dispatch: BXDF_MIXED is remapped to DIFFUSE on entry (:32-34, :140-142,
// :249-251) so the Fresnel-mixture block below it is dead code and is not restated. -------------
static inline int effType(int t) { return t == FLX_BXDF_MIXED ? FLX_BXDF_DIFFUSE : t; }

f3 bxdfSample(const Hit &hit, const flx_material &fm, bool backface, const ShadeScene &s, f3 dirIn, f3 *dirOut, float *pdfW, uint32_t *seed)
{
    const Mat m = loadMat(fm);
    switch (effType(m.type))
    {
    case FLX_BXDF_DIFFUSE: return sampleDiffuse(hit, m, s, dirOut, pdfW, seed);
    case FLX_BXDF_GLOSSY: return sampleGlossy(hit, m, s, dirIn, dirOut, pdfW, seed);
    case FLX_BXDF_GGX_ROUGH_REFLECTION: return sampleGGXReflect(hit, m, s, dirIn, dirOut, pdfW, seed);
    case FLX_BXDF_IDEAL_REFLECTION: return sampleIdealReflection(hit, m, s, dirIn, dirOut, pdfW);
    case FLX_BXDF_GGX_ROUGH_DIELECTRIC: return sampleGGXRefract(hit, m, backface, s, dirIn, dirOut, pdfW, seed);
    case FLX_BXDF_IDEAL_DIELECTRIC: return sampleIdealDielectric(hit, m, backface, s, dirIn, dirOut, pdfW, seed);
    case FLX_BXDF_EMISSIVE: return mk3(1.0f); // dirOut/pdfW untouched (:118-121); caller pins them to 0
    }
    return mk3(0.0f);
}

f3 bxdfEval(const Hit &hit, const flx_material &fm, bool backface, const ShadeScene &s, f3 dirIn, f3 dirOut)
{
    const Mat m = loadMat(fm);
    switch (effType(m.type))
    {
    case FLX_BXDF_DIFFUSE: return evalDiffuse(hit, m, s);
    case FLX_BXDF_GLOSSY: return evalGlossy(hit, m, s, dirIn, dirOut);
    case FLX_BXDF_GGX_ROUGH_REFLECTION: return evalGGXReflect(hit, m, s, dirIn, dirOut);
    case FLX_BXDF_GGX_ROUGH_DIELECTRIC: return evalGGXRefract(hit, m, backface, s, dirIn, dirOut);
    case FLX_BXDF_EMISSIVE: return m.Ke;
    }
    return mk3(0.0f); // ideal reflection / dielectric evaluate to 0
}

float bxdfPdf(const Hit &hit, const flx_material &fm, bool backface, const ShadeScene &s, f3 dirIn, f3 dirOut)
{
    const Mat m = loadMat(fm);
    switch (effType(m.type))
    {
    case FLX_BXDF_DIFFUSE: return pdfDiffuse(hit, dirOut);
    case FLX_BXDF_GLOSSY: return pdfGlossy(hit, m, s, dirIn, dirOut);
    case FLX_BXDF_GGX_ROUGH_REFLECTION: return pdfGGXReflect(hit, m, dirIn, dirOut);
    case FLX_BXDF_GGX_ROUGH_DIELECTRIC: return pdfGGXRefract(hit, m, backface, dirIn, dirOut);
    }
    return 0.0f;
}

} // namespace orc
