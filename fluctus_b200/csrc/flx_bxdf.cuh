// Device BxDF library: Lambert, LuxRender-style glossy coat, GGX rough reflection / refraction
// (Walter 07), mirror, ideal dielectric, emissive.  Behavioural contract: reference
// src/diffuse.cl, glossy.cl, ggx.cl, ideal_reflection.cl, ideal_dielectric.cl and the dispatch in
// src/bxdf_partial.cl.  The reference specialises its material kernels with -DBXDF_USE_* at JIT
// time; here the dispatch is a template on a compile-time type mask so each queue kernel carries
// only its own branch and all variants are built ahead of time.
#pragma once
#include "flx_shade.cuh"

struct Mat // register copy of flx_material (glossy.cl:28 edits its private copy too)
{
    f3 Kd, Ks, Ke; // Kd: already raised to 2.2 (the only consumer is matGetAlbedo; see MaterialsDev::gammaLut)
    float Ns, Ni;
    int map_Kd, map_Ks, type;
};
FLX_DEV Mat loadMat(const flx_material *__restrict__ m)
{
    const float4 *p = (const float4 *)m;
#ifdef FLX_GAMMA_PER_PATH
    const float4 a = __ldg(p);
#else
    const float4 a = __ldg(p + 3); // Kt slot of the device copy = powf(Kd, 2.2f) (k_prepare_materials)
#endif
    const float4 b = __ldg(p + 1), c = __ldg(p + 2), e = __ldg(p + 4), f = __ldg(p + 5);
    Mat r;
    r.Kd = mk3(a); r.Ks = mk3(b); r.Ke = mk3(c);
    r.Ns = e.x; r.Ni = e.y; r.map_Kd = __float_as_int(e.w);
    r.map_Ks = __float_as_int(f.x); r.type = __float_as_int(f.z);
    return r;
}

struct SurfHit { f3 N; float uvx, uvy; };

// ---- diffuse.cl:9-26 ---------------------------------------------------------------------------
FLX_DEV f3 evalDiffuse(const SurfHit &h, const Mat &m, const MaterialsDev &md)
{
    return matGetAlbedo(m.Kd, h.uvx, h.uvy, m.map_Kd, md) * FLX_INV_PI;
}
FLX_DEV f3 sampleDiffuse(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 &dirOut, float &pdfW, uint32_t &seed)
{
    dirOut = cosSampleHemisphere(h.N, seed, pdfW);
    return evalDiffuse(h, m, md);
}
FLX_DEV float pdfDiffuse(const SurfHit &h, f3 dirOut) { return dot(h.N, dirOut) * FLX_INV_PI; }

// ---- ggx.cl ------------------------------------------------------------------------------------
FLX_DEV f3 ggxSampleLobe(float alpha, f3 N, uint32_t &seed) // :20-37
{
    f3 X, Y;
    makeOrthoBasis(N, X, Y);
    const float rx = rnd(seed), ry = rnd(seed);
    const float theta = flx_atan2f(alpha * sqrtf(rx), sqrtf(1.0f - rx));
    const float phi = FLX_2PI * ry;
    float sinTheta, cosTheta, sinPhi, cosPhi;
    flx_sincosf(theta, &sinTheta, &cosTheta);
    flx_sincosf(phi, &sinPhi, &cosPhi);
    return (X * sinTheta * cosPhi + Y * sinTheta * sinPhi + N * cosTheta);
}
FLX_DEV float ggxG1(float alpha, f3 v, f3 n, f3 m) // :41-54
{
    const float mDotV = dot(m, v), nDotV = dot(n, v);
    if (nDotV * mDotV <= 0.0f) return 0.0f;
    const float cosThSq = nDotV * nDotV;
    const float tanSq = (cosThSq > 0.0f) ? ((1.0f - cosThSq) / cosThSq) : 0.0f;
    return 2.0f / (1.0f + sqrtf(1.0f + alpha * alpha * tanSq));
}
FLX_DEV float ggxG(float alpha, f3 dirIn, f3 dirOut, f3 n, f3 m) { return ggxG1(alpha, dirIn, n, m) * ggxG1(alpha, dirOut, n, m); }
FLX_DEV float ggxD(float alpha, f3 n, f3 m) // :65-78 (GTR2)
{
    const float nDotM = dot(n, m);
    if (nDotM <= 0.0f) return 0.0f;
    const float nDotMSq = nDotM * nDotM;
    const float tanSq = ((1.0f - nDotMSq) / nDotMSq);
    const float aSq = alpha * alpha;
    const float denom = FLX_PI * nDotMSq * nDotMSq * (aSq + tanSq) * (aSq + tanSq);
    return (aSq / denom);
}
FLX_DEV float ggxPdfReflect(float alpha, f3 dirOut, f3 N, f3 H) // :81-89
{
    const float nDotH = dot(N, H), oDotH = dot(dirOut, H);
    return oDotH == 0.0f ? 0.0f : ggxD(alpha, N, H) * fabsf(nDotH * 0.25f / oDotH);
}
FLX_DEV f3 ggxReflTerm(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 dirInN, f3 dirOut, f3 H, float alpha)
{
    const float iDotN = dot(dirInN, h.N), oDotN = dot(dirOut, h.N);
    const float F = (m.Ni > 1.0f) ? fresnelDielectric(iDotN, 1.0f, m.Ni) : 1.0f;
    const f3 Ks = matGetFloat3(m.Ks, h.uvx, h.uvy, m.map_Ks, md);
    const float D = ggxD(alpha, h.N, H);
    const float G = ggxG(alpha, dirInN, dirOut, h.N, H);
    const float den = (iDotN * oDotN);
    return (den != 0.0f) ? (Ks * F * G * D * 0.25f / den) : mk3(0.0f);
}
FLX_DEV f3 sampleGGXReflect(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW, uint32_t &seed) // :91-114
{
    const float alpha = m.Ns;
    const f3 H = ggxSampleLobe(alpha, h.N, seed);
    dirOut = reflect0(dirIn, H);
    pdfW = ggxPdfReflect(alpha, dirOut, h.N, H);
    return ggxReflTerm(h, m, md, -dirIn, dirOut, H, alpha);
}
FLX_DEV f3 evalGGXReflect(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 dirIn, f3 dirOut) // :116-136
{
    const f3 H = normalize(dirOut - dirIn);
    return ggxReflTerm(h, m, md, -dirIn, dirOut, H, m.Ns);
}
FLX_DEV float pdfGGXReflect(const SurfHit &h, const Mat &m, f3 dirIn, f3 dirOut) // :138-143
{
    const f3 H = normalize(dirOut - dirIn);
    return ggxPdfReflect(m.Ns, dirOut, h.N, H);
}
FLX_DEV float ggxPdfRefract(float alpha, float etaI, float etaO, f3 dirIn, f3 dirOut, f3 N, f3 H) // :146-154
{
    const float nDotH = fabsf(dot(N, H)), iDotH = fabsf(dot(dirIn, H)), oDotH = fabsf(dot(dirOut, H));
    const float sqrtJInv = etaI * iDotH + etaO * oDotH;
    return sqrtJInv == 0.0f ? 0.0f : ggxD(alpha, N, H) * nDotH * oDotH * etaO * etaO / (sqrtJInv * sqrtJInv);
}
// shared tail of the refraction branch of sample/eval (:186-221, :249-274)
FLX_DEV f3 ggxRefrTerm(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 dirIn, f3 dirOut, f3 H, f3 Nn,
                       float etaI, float etaO, float iDotN, float oDotN, float F)
{
    const float eta = etaI / etaO;
    f3 bsdf = mk3(eta * eta);
    bsdf *= matGetFloat3(m.Ks, h.uvx, h.uvy, m.map_Ks, md);
    const float iDotH = fabsf(dot(dirIn, H)), oDotH = fabsf(dot(dirOut, H));
    const float s = etaI * iDotH + etaO * oDotH;
    const float focusTermDenom = iDotN * oDotN * s * s;
    if (focusTermDenom == 0.0f) return mk3(0.0f);
    const float focusTerm = etaO * etaO * iDotH * oDotH / focusTermDenom;
    const float D = ggxD(m.Ns, Nn, H);
    const float G = ggxG(m.Ns, -dirIn, dirOut, Nn, H);
    return (1.0f - F) * bsdf * D * G * focusTerm;
}
FLX_DEV f3 sampleGGXRefract(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW, uint32_t &seed) // :156-222
{
    const f3 dirInN = -dirIn;
    const float alpha = m.Ns;
    const float etaI = backface ? m.Ni : 1.0f, etaO = backface ? 1.0f : m.Ni;
    const float iDotN = dot(dirInN, h.N);
    f3 H = ggxSampleLobe(alpha, h.N, seed);
    float cosThetaT;
    const float F = fresnelDielectric1(iDotN, etaI, etaO, cosThetaT);
    if (rnd(seed) < F)
    {
        dirOut = reflect0(dirIn, H);
        pdfW = ggxPdfReflect(alpha, dirOut, h.N, H);
        const float oDotN = dot(dirOut, h.N);
        const float D = ggxD(alpha, h.N, H);
        const float G = ggxG(alpha, dirInN, dirOut, h.N, H);
        const float den = (iDotN * oDotN);
        return mk3((den != 0.0f) ? (F * G * D * 0.25f / den) : 0.0f);
    }
    dirOut = refract1(dirIn, h.N, etaI / etaO, iDotN, cosThetaT);
    H = normalize((dirIn * etaI - dirOut * etaO));
    const f3 Nn = backface ? -h.N : h.N;
    pdfW = ggxPdfRefract(alpha, etaI, etaO, dirInN, dirOut, Nn, H);
    return ggxRefrTerm(h, m, md, dirIn, dirOut, H, Nn, etaI, etaO, iDotN, dot(dirOut, h.N), F);
}
FLX_DEV f3 evalGGXRefract(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 dirOut) // :224-275
{
    const f3 dirInN = -dirIn;
    const float etaI = backface ? m.Ni : 1.0f, etaO = backface ? 1.0f : m.Ni;
    const float iDotN = dot(dirInN, h.N), oDotN = dot(dirOut, h.N);
    const float F = fresnelDielectric(iDotN, etaI, etaO);
    if (!backface)
    {
        const f3 H = normalize(dirOut - dirIn);
        const float D = ggxD(m.Ns, h.N, H);
        const float G = ggxG(m.Ns, dirInN, dirOut, h.N, H);
        const float den = (iDotN * oDotN);
        return (den != 0.0f) ? mk3(F * G * D * 0.25f / den) : mk3(0.0f);
    }
    const f3 H = normalize((dirIn * etaI - dirOut * etaO));
    return ggxRefrTerm(h, m, md, dirIn, dirOut, H, -h.N, etaI, etaO, iDotN, oDotN, F);
}
FLX_DEV float pdfGGXRefract(const SurfHit &h, const Mat &m, bool backface, f3 dirIn, f3 dirOut) // :277-296
{
    if (!backface)
    {
        const f3 H = normalize(dirOut - dirIn);
        return ggxPdfReflect(m.Ns, dirOut, h.N, H);
    }
    const float etaI = m.Ni, etaO = 1.0f;
    const f3 H = normalize((dirIn * etaI - dirOut * etaO));
    return ggxPdfRefract(m.Ns, etaI, etaO, -dirIn, dirOut, -h.N, H);
}

// ---- glossy.cl ---------------------------------------------------------------------------------
FLX_DEV f3 etaToKs(float eta) { const float r = (eta > 0.0f) ? ((eta - 1.0f) / (eta + 1.0f)) : 0.0f; return mk3(r * r); }
FLX_DEV float ksToEta(f3 Ks)
{
    const float k = clampf((Ks.x + Ks.y + Ks.z) / 3.0f, 0.0f, 0.99f);
    return (sqrtf(k) + 1.0f) / (1.0f - sqrtf(k));
}
FLX_DEV f3 sampleGlossy(const SurfHit &h, const Mat &mat, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW, uint32_t &seed) // :22-64
{
    Mat m = mat;
    m.Ks = matGetFloat3(mat.Ks, h.uvx, h.uvy, mat.map_Ks, md);
    m.Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(m.Ks);
    if (isZero(m.Ks)) m.Ks = etaToKs(m.Ni);
    const float cosTh = -dot(dirIn, h.N);
    const float F = fresnelDielectric(cosTh, 1.0f, m.Ni);
    float basePdf, coatingPdf;
    f3 baseBrdf, coatingBrdf;
    if (rnd(seed) < F)
    {
        coatingBrdf = sampleGGXReflect(h, m, md, dirIn, dirOut, coatingPdf, seed);
        baseBrdf = evalDiffuse(h, m, md);
        basePdf = pdfDiffuse(h, dirOut);
    }
    else
    {
        baseBrdf = sampleDiffuse(h, m, md, dirOut, basePdf, seed);
        coatingBrdf = evalGGXReflect(h, m, md, dirIn, dirOut);
        coatingPdf = pdfGGXReflect(h, m, dirIn, dirOut);
    }
    if (dot(h.N, dirOut) < 1e-5f) return mk3(0.0f); // pdfW stays at the caller's 0
    pdfW = (1.0f - F) * basePdf + F * coatingPdf;
    return (baseBrdf * (1.0f - F) + coatingBrdf);
}
FLX_DEV f3 evalGlossy(const SurfHit &h, const Mat &mat, const MaterialsDev &md, f3 dirIn, f3 dirOut) // :66-86
{
    Mat m = mat;
    m.Ks = matGetFloat3(mat.Ks, h.uvx, h.uvy, mat.map_Ks, md);
    m.Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(m.Ks);
    if (length(m.Ks) == 0.0f) m.Ks = etaToKs(m.Ni);
    const f3 baseBrdf = evalDiffuse(h, m, md);
    const f3 coatingBrdf = evalGGXReflect(h, m, md, dirIn, dirOut);
    const float cosTh = -dot(dirIn, h.N);
    const float F = fresnelDielectric(cosTh, 1.0f, m.Ni);
    return (baseBrdf * (1.0f - F) + coatingBrdf);
}
FLX_DEV float pdfGlossy(const SurfHit &h, const Mat &mat, const MaterialsDev &md, f3 dirIn, f3 dirOut) // :88-101
{
    const f3 Ks = matGetFloat3(mat.Ks, h.uvx, h.uvy, mat.map_Ks, md);
    const float Ni = (mat.Ni > 0.0f) ? mat.Ni : ksToEta(Ks);
    const float basePdf = pdfDiffuse(h, dirOut);
    const float coatingPdf = pdfGGXReflect(h, mat, dirIn, dirOut);
    const float cosTh = -dot(dirIn, h.N);
    const float F = fresnelDielectric(cosTh, 1.0f, Ni);
    return (1.0f - F) * basePdf + F * coatingPdf;
}

// ---- ideal_reflection.cl:9-21, ideal_dielectric.cl:10-44 ---------------------------------------
FLX_DEV f3 sampleIdealReflection(const SurfHit &h, const Mat &m, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW)
{
    dirOut = reflect0(dirIn, h.N);
    pdfW = 1.0f;
    const f3 ks = matGetFloat3(m.Ks, h.uvx, h.uvy, m.map_Ks, md);
    const float cosO = dot(normalize(dirOut), h.N);
    return (cosO != 0.0f) ? ks / cosO : mk3(0.0f);
}
FLX_DEV f3 sampleIdealDielectric(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW, uint32_t &seed)
{
    f3 bsdf = mk3(1.0f);
    const float cosI = -dot(dirIn, h.N);
    const float n1 = backface ? m.Ni : 1.0f, n2 = backface ? 1.0f : m.Ni;
    const float eta = n1 / n2;
    float cosThetaT;
    const float fr = fresnelDielectric1(cosI, n1, n2, cosThetaT);
    if (rnd(seed) < fr)
    {
        dirOut = reflectDot(dirIn, h.N, cosI);
    }
    else
    {
        dirOut = refract1(dirIn, h.N, eta, cosI, cosThetaT);
        bsdf *= eta * eta;
        bsdf *= matGetFloat3(m.Ks, h.uvx, h.uvy, m.map_Ks, md);
    }
    pdfW = 1.0f;
    const float cosO = dot(normalize(dirOut), h.N);
    return bsdf / cosO;
}

// ---- dispatch (bxdf_partial.cl:19-342).  BXDF_MIXED is shaded as diffuse (:32-34). --------------
#define FLX_HAS(T) ((TYPES & (T)) != 0)
FLX_DEV int effType(int t) { return t == FLX_BXDF_MIXED ? FLX_BXDF_DIFFUSE : t; }

template <int TYPES>
FLX_DEV f3 bxdfSample(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 &dirOut, float &pdfW, uint32_t &seed)
{
    const int t = effType(m.type);
    if (FLX_HAS(FLX_BXDF_DIFFUSE) && t == FLX_BXDF_DIFFUSE) return sampleDiffuse(h, m, md, dirOut, pdfW, seed);
    if (FLX_HAS(FLX_BXDF_GLOSSY) && t == FLX_BXDF_GLOSSY) return sampleGlossy(h, m, md, dirIn, dirOut, pdfW, seed);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_REFLECTION) && t == FLX_BXDF_GGX_ROUGH_REFLECTION) return sampleGGXReflect(h, m, md, dirIn, dirOut, pdfW, seed);
    if (FLX_HAS(FLX_BXDF_IDEAL_REFLECTION) && t == FLX_BXDF_IDEAL_REFLECTION) return sampleIdealReflection(h, m, md, dirIn, dirOut, pdfW);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_DIELECTRIC) && t == FLX_BXDF_GGX_ROUGH_DIELECTRIC) return sampleGGXRefract(h, m, backface, md, dirIn, dirOut, pdfW, seed);
    if (FLX_HAS(FLX_BXDF_IDEAL_DIELECTRIC) && t == FLX_BXDF_IDEAL_DIELECTRIC) return sampleIdealDielectric(h, m, backface, md, dirIn, dirOut, pdfW, seed);
    if (FLX_HAS(FLX_BXDF_EMISSIVE) && t == FLX_BXDF_EMISSIVE) return mk3(1.0f); // dirOut/pdfW untouched (:118-121)
    return mk3(0.0f);
}
template <int TYPES>
FLX_DEV f3 bxdfEval(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 dirOut)
{
    const int t = effType(m.type);
    if (FLX_HAS(FLX_BXDF_DIFFUSE) && t == FLX_BXDF_DIFFUSE) return evalDiffuse(h, m, md);
    if (FLX_HAS(FLX_BXDF_GLOSSY) && t == FLX_BXDF_GLOSSY) return evalGlossy(h, m, md, dirIn, dirOut);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_REFLECTION) && t == FLX_BXDF_GGX_ROUGH_REFLECTION) return evalGGXReflect(h, m, md, dirIn, dirOut);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_DIELECTRIC) && t == FLX_BXDF_GGX_ROUGH_DIELECTRIC) return evalGGXRefract(h, m, backface, md, dirIn, dirOut);
    if (FLX_HAS(FLX_BXDF_EMISSIVE) && t == FLX_BXDF_EMISSIVE) return m.Ke;
    return mk3(0.0f);
}
template <int TYPES>
FLX_DEV float bxdfPdf(const SurfHit &h, const Mat &m, bool backface, const MaterialsDev &md, f3 dirIn, f3 dirOut)
{
    const int t = effType(m.type);
    if (FLX_HAS(FLX_BXDF_DIFFUSE) && t == FLX_BXDF_DIFFUSE) return pdfDiffuse(h, dirOut);
    if (FLX_HAS(FLX_BXDF_GLOSSY) && t == FLX_BXDF_GLOSSY) return pdfGlossy(h, m, md, dirIn, dirOut);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_REFLECTION) && t == FLX_BXDF_GGX_ROUGH_REFLECTION) return pdfGGXReflect(h, m, dirIn, dirOut);
    if (FLX_HAS(FLX_BXDF_GGX_ROUGH_DIELECTRIC) && t == FLX_BXDF_GGX_ROUGH_DIELECTRIC) return pdfGGXRefract(h, m, backface, dirIn, dirOut);
    return 0.0f;
}

// utils.cl:174-207 — only reached for materials with a normal map
FLX_DEV f3 tangentSpaceNormal(f3 hitN, float uvx, float uvy, int hitI, int mapN, const MaterialsDev &md)
{
    f3 texNormal = readTexture(uvx, uvy, md.textures[mapN], md.texels);
    texNormal = texNormal + texNormal - mk3(1.0f, 1.0f, 1.0f);
    const flx_triangle *t = md.tris160 + hitI;
    const f3 e1 = ld3(t->v1.p) - ld3(t->v0.p);
    const f3 e2 = ld3(t->v2.p) - ld3(t->v0.p);
    const f3 t1 = ld3(t->v1.t) - ld3(t->v0.t);
    const f3 t2 = ld3(t->v2.t) - ld3(t->v0.t);
    const float det = (t1.x * t2.y - t1.y * t2.x);
    if (det == 0.0f) return hitN;
    const float invDet = recip(det);
    const f3 T = normalize(invDet * (e1 * t2.y - e2 * t1.y));
    const f3 B = normalize(invDet * (e2 * t1.x - e1 * t2.x));
    f3 N;
    N.x = T.x * texNormal.x + B.x * texNormal.y + hitN.x * texNormal.z;
    N.y = T.y * texNormal.x + B.y * texNormal.y + hitN.y * texNormal.z;
    N.z = T.z * texNormal.x + B.z * texNormal.y + hitN.z * texNormal.z;
    return normalize(N);
}

