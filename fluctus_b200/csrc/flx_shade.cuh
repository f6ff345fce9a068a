// Device shading library: RNG, sampling helpers, textures, environment map, Fresnel.
// Behavioural contract: reference src/random.cl, src/utils.cl, src/env_map.cl, src/fresnel.cl.
// Operation order follows those files so that fixed-seed path state agrees with the CPU oracle to
// 1e-5 relative (the only differences are libm ulps in sinf/cosf/atan2f/acosf/powf).
#pragma once
#include "flx_internal.h"
#include "flx_libm.cuh"

// ---- src/random.cl:7-22 (Wang hash chain) ------------------------------------------------------
FLX_DEV uint32_t wangHash(uint32_t s)
{
    s = (s ^ 61u) ^ (s >> 16);
    s *= 9u;
    s = s ^ (s >> 4);
    s *= 0x27d4eb2du;
    s = s ^ (s >> 15);
    return s;
}
FLX_DEV float rnd(uint32_t &seed)
{
    seed = wangHash(seed);
    return (float)seed * FLX_ONE_2P32; // can be exactly 1.0f, as in the reference
}

FLX_DEV float luminance(f3 v) { return 0.212671f * v.x + 0.715160f * v.y + 0.072169f * v.z; }
FLX_DEV f3 reflect0(f3 dir, f3 n) { return dir - dot(dir, n) * (n + n); }   // utils.cl:30-33
FLX_DEV f3 reflectDot(f3 dir, f3 n, float dt) { return dir + (dt + dt) * n; } // utils.cl:48-51
FLX_DEV f3 refract1(f3 wi, f3 n, float eta, float iDotN, float cosT) { return wi * eta + n * (eta * iDotN - cosT); }
FLX_DEV float pdfAtoW(float pdf, float dist, float cosine) { return pdf * (dist * dist) / fabsf(cosine); }
FLX_DEV f3 ld3(const flx_float3 &v) { return mk3(v.x, v.y, v.z); }

// utils.cl:74-81
FLX_DEV void makeOrthoBasis(f3 N, f3 &a, f3 &b)
{
    const int sign = (N.z > 0.0f) * 2 - 1;
    const float fs = (float)sign;
    const float aa = -recip(fs + N.z);
    const float bb = N.x * N.y * aa;
    a = mk3(1.0f + fs * N.x * N.x * aa, fs * bb, (float)(-sign) * N.x);
    b = mk3(bb, fs + N.y * N.y * aa, -N.y);
}

// utils.cl:96-108
FLX_DEV void uniformSampleDisk(uint32_t &seed, float &dx, float &dy)
{
    const float sqrt_r = sqrtf(rnd(seed));
    const float th = FLX_2PI * rnd(seed);
    float s, c;
    flx_sincosf(th, &s, &c);
    dx = sqrt_r * c; dy = sqrt_r * s;
}

// utils.cl:111-137
FLX_DEV f3 cosSampleHemisphere(f3 n, uint32_t &seed, float &pdf)
{
    const float r1 = FLX_2PI * rnd(seed);
    const float r2 = rnd(seed);
    const float r2s = sqrtf(r2);
    f3 w = n, u, v;
    makeOrthoBasis(w, u, v);
    float sinR1, cosR1;
    flx_sincosf(r1, &sinR1, &cosR1);
    u *= (cosR1 * r2s);
    v *= (sinR1 * r2s);
    w *= sqrtf(1.0f - r2);
    const f3 dir = u + v + w;
    pdf = dot(n, dir) * FLX_INV_PI;
    return dir;
}

// utils.cl:139-158: nearest texel, repeat wrap (unsigned modulo as OpenCL evaluates int % uint)
FLX_DEV uchar4 readTexel(float uvx, float uvy, const flx_tex_desc &tex, const uint8_t *__restrict__ data)
{
    const uint32_t width = tex.width, height = tex.height;
    uvx *= (float)width;
    uvy *= (float)height;
    const float fx = floorf(uvx), fy = floorf(uvy);
    const int tx = (int)(((uint32_t)(int)fx % width + width) % width);
    const int ty = (int)(((uint32_t)(int)fy % height + height) % height);
    int cx = (int)((float)tx + uvx - fx);
    int cy = (int)((float)ty + uvy - fy);
    cx = min(max(cx, 0), (int)width - 1);
    cy = min(max(cy, 0), (int)height - 1);
    return *(const uchar4 *)(data + tex.offset + (uint32_t)cx * 4u + (uint32_t)cy * width * 4u);
}
FLX_DEV f3 readTexture(float uvx, float uvy, const flx_tex_desc &tex, const uint8_t *__restrict__ data)
{
    const uchar4 p = readTexel(uvx, uvy, tex, data);
    f3 c = mk3((float)p.x, (float)p.y, (float)p.z);
    c *= FLX_ONE_255;
    return c;
}

// utils.cl:161-171
FLX_DEV f3 matGetFloat3(f3 fallback, float uvx, float uvy, int idx, const MaterialsDev &m)
{
    return (idx != -1) ? readTexture(uvx, uvy, m.textures[idx], m.texels) : fallback;
}
// pow(matGetFloat3(...), 2.2f) of utils.cl:161-166 without a powf per path: `fallbackGamma` is the material's
// pre-raised constant, texels go through the 256-entry table (MaterialsDev::gammaLut) — same bits either way.
FLX_DEV f3 matGetAlbedo(f3 fallbackGamma, float uvx, float uvy, int idx, const MaterialsDev &m)
{
#ifdef FLX_GAMMA_PER_PATH
    return pow3x(matGetFloat3(fallbackGamma, uvx, uvy, idx, m), 2.2f); // fallbackGamma = raw Kd in this build (loadMat)
#else
    if (idx == -1) return fallbackGamma;
    const uchar4 p = readTexel(uvx, uvy, m.textures[idx], m.texels);
    return mk3(__ldg(m.gammaLut + p.x), __ldg(m.gammaLut + p.y), __ldg(m.gammaLut + p.z));
#endif
}

// utils.cl:251-259
FLX_DEV void sampleAreaLight(const flx_area_light &light, float &pdf, f3 &p, uint32_t &seed)
{
    pdf = recip(4.0f * light.size.x * light.size.y);
    p = ld3(light.pos);
    const float r1 = rnd(seed);
    const float r2 = rnd(seed);
    p += (r1 + r1 - 1.0f) * light.size.x * ld3(light.right);
    p += (r2 + r2 - 1.0f) * light.size.y * ld3(light.up);
}

// ---- src/env_map.cl -----------------------------------------------------------------------------
FLX_DEV void directionToUV(f3 dir, float &u, float &v) // :14-17, atan2pi/acospi as atan2f/acosf * (1/pi)
{
    u = (flx_atan2f(dir.x, -dir.z) * FLX_INV_PI) * 0.5f + 0.5f;
    v = flx_acosf(dir.y) * FLX_INV_PI;
}
FLX_DEV f3 uvToDirection(float u, float v, float &sinPhi) // :21-33
{
    float cosPhi, sinTh, cosTh;
    flx_sincosf(v * FLX_PI, &sinPhi, &cosPhi);
    flx_sincosf(u * FLX_2PI - FLX_PI, &sinTh, &cosTh);
    return mk3(sinPhi * sinTh, cosPhi, -sinPhi * cosTh);
}
// read_imagef(normalised coords, clamp-to-edge, linear) in full fp32: the texture unit's 8-bit
// filter weights would break the 1e-5 state tolerance, so the four texels are blended by hand.
FLX_DEV f3 envLinear(const EnvMapDev &e, float su, float sv)
{
    const int w = e.w, h = e.h;
    const float u = su * (float)w, v = sv * (float)h;
    const float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    const int i0r = (int)fu, j0r = (int)fv;
    const float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    const int i0 = min(max(i0r, 0), w - 1), i1 = min(max(i0r + 1, 0), w - 1);
    const int j0 = min(max(j0r, 0), h - 1), j1 = min(max(j0r + 1, 0), h - 1);
    const float4 T00 = __ldg(e.rgba + (size_t)j0 * w + i0), T10 = __ldg(e.rgba + (size_t)j0 * w + i1);
    const float4 T01 = __ldg(e.rgba + (size_t)j1 * w + i0), T11 = __ldg(e.rgba + (size_t)j1 * w + i1);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return mk3(w00 * T00.x + w10 * T10.x + w01 * T01.x + w11 * T11.x,
               w00 * T00.y + w10 * T10.y + w01 * T01.y + w11 * T11.y,
               w00 * T00.z + w10 * T10.z + w01 * T01.z + w11 * T11.z);
}
FLX_DEV f3 evalEnvMapDir(const EnvMapDev &e, f3 dir) // :35-39
{
    float u, v;
    directionToUV(dir, u, v);
    return envLinear(e, u, v);
}
// :95-109 with the (u, v) of the direction supplied by the caller (same directionToUV, evaluated once when the
// radiance and the pdf of one direction are both needed)
FLX_DEV float envMapPdfUV(const EnvMapDev &e, f3 direction, float u, float v)
{
    if (direction.y > 0.99f) return 0.0f;
    const int iu = min((int)floorf(u * (float)e.w), e.w - 1);
    const int iv = min((int)floorf(v * (float)e.h), e.h - 1);
    return __ldg(e.pdf + iv * e.w + iu) * rsqrt_(1.0f - direction.y * direction.y);
}
// :63-92 alias-method sample; three dependent table reads per sample
FLX_DEV void sampleEnvMapAlias(float r, f3 &L, float &pdfW, const EnvMapDev &e)
{
    const int width = e.w, wh = e.w * e.h;
    const float x = r * (float)wh;
    const int i = min((int)floorf(x), wh - 1);
    const float mProb = __ldg(e.prob + i);
    const int uvInd = (x - (float)i < mProb) ? i : __ldg(e.alias + i);
    const float u = ((float)(uvInd % width) + 0.5f) * recip((float)width);
    const float v = ((float)uvInd + 0.5f) * recip((float)wh);
    float sinTh;
    L = uvToDirection(u, v, sinTh);
    pdfW = (sinTh != 0.0f) ? __ldg(e.pdf + uvInd) * recip(sinTh) : 0.0f;
}
FLX_DEV float envMapPdf(const EnvMapDev &e, f3 direction) // :95-109
{
    if (direction.y > 0.99f) return 0.0f;
    float u, v;
    directionToUV(direction, u, v);
    const int iu = min((int)floorf(u * (float)e.w), e.w - 1);
    const int iv = min((int)floorf(v * (float)e.h), e.h - 1);
    return __ldg(e.pdf + iv * e.w + iu) * rsqrt_(1.0f - direction.y * direction.y);
}

// ---- src/fresnel.cl:5-60 ------------------------------------------------------------------------
FLX_DEV float fresnelDielectric1(float cosThI, float etaI, float etaT, float &cosThetaT)
{
    const float sinThetaI = sqrtf(fmaxf(0.0f, 1.0f - cosThI * cosThI));
    const float sinThetaT = etaI / etaT * sinThetaI;
    cosThetaT = sqrtf(fmaxf(0.0f, 1.0f - sinThetaT * sinThetaT));
    if (sinThetaT >= 1.0f) return 1.0f;
    const float a = etaT * cosThI, b = etaI * cosThetaT;
    const float parl = (a - b) / (a + b);
    const float c = etaI * cosThI, d = etaT * cosThetaT;
    const float perp = (c - d) / (c + d);
    return 0.5f * (parl * parl + perp * perp);
}
FLX_DEV float fresnelDielectric(float cosThI, float etaI, float etaT)
{
    float c;
    return fresnelDielectric1(cosThI, etaI, etaT, c);
}
