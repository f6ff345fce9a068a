// TEST INFRASTRUCTURE — CPU oracle.  Restates reference src/utils.cl, env_map.cl, fresnel.cl,
// tonemap.cl and the per-pixel body of mk_postprocess.cl.
#include "o_shade.h"

namespace orc {

// utils.cl:74-81 (Duff/Frisvad ONB).  `sign + N.z` is int+float -> float.
void makeOrthoBasis(f3 N, f3 *a, f3 *b)
{
    const int sign = (N.z > 0.0f) * 2 - 1;
    const float fs = (float)sign;
    const float aa = -recip(fs + N.z);
    const float bb = N.x * N.y * aa;
    *a = mk3(1.0f + fs * N.x * N.x * aa, fs * bb, (float)(-sign) * N.x);
    *b = mk3(bb, fs + N.y * N.y * aa, -N.y);
}

// utils.cl:96-108
f2 uniformSampleDisk(uint32_t *seed)
{
    const float sqrt_r = sqrtf(rnd(seed));
    const float th = k2PI * rnd(seed);
    f2 r; r.x = sqrt_r * o_cosf(th); r.y = sqrt_r * o_sinf(th);
    return r;
}

// utils.cl:111-137
f3 cosSampleHemisphere(f3 n, uint32_t *seed, float *p)
{
    const float r1 = k2PI * rnd(seed);
    const float r2 = rnd(seed);
    const float r2s = sqrtf(r2);
    f3 w = n, u, v;
    makeOrthoBasis(w, &u, &v);
    const float cosR1 = o_cosf(r1), sinR1 = o_sinf(r1);
    u *= (cosR1 * r2s);
    v *= (sinR1 * r2s);
    w *= sqrtf(1 - r2);
    const f3 dir = u + v + w;
    *p = dot(n, dir) * kInvPI;
    return dir;
}

// utils.cl:139-158.  `(int)floor % width` is int % uint, i.e. evaluated in unsigned arithmetic.
f3 readTexture(f2 uv, const flx_tex_desc &tex, const uint8_t *data)
{
    const uint32_t width = tex.width, height = tex.height;
    uv.x *= (float)width;
    uv.y *= (float)height;
    const int tx = (int)(((uint32_t)(int)floorf(uv.x) % width + width) % width);
    const int ty = (int)(((uint32_t)(int)floorf(uv.y) % height + height) % height);
    int cx = (int)((float)tx + uv.x - floorf(uv.x));
    int cy = (int)((float)ty + uv.y - floorf(uv.y));
    cx = cx < 0 ? 0 : (cx > (int)width - 1 ? (int)width - 1 : cx);
    cy = cy < 0 ? 0 : (cy > (int)height - 1 ? (int)height - 1 : cy);
    const uint8_t *pix = data + tex.offset + (uint32_t)cx * 4u + (uint32_t)cy * width * 4u;
    f3 c = mk3((float)pix[0], (float)pix[1], (float)pix[2]);
    c *= kOne255;
    return c;
}

// utils.cl:161-171
f3 matGetAlbedo(f3 fallback, f2 uv, int idx, const ShadeScene &s)
{
    const f3 val = (idx != -1) ? readTexture(uv, s.textures[idx], s.texData) : fallback;
    return pow3(val, 2.2f);
}
f3 matGetFloat3(f3 fallback, f2 uv, int idx, const ShadeScene &s)
{
    return (idx != -1) ? readTexture(uv, s.textures[idx], s.texData) : fallback;
}

// utils.cl:174-207
f3 tangentSpaceNormal(const Hit &hit, const flx_material &mat, const ShadeScene &s)
{
    if (mat.map_N == -1) return hit.N;
    f3 texNormal = matGetFloat3(mk3(0.5f, 0.5f, 1.0f), hit.uv, mat.map_N, s);
    texNormal = texNormal + texNormal - mk3(1.0f, 1.0f, 1.0f);

    const flx_triangle &t = s.tris[hit.i];
    const f3 e1 = ld3(t.v1.p) - ld3(t.v0.p);
    const f3 e2 = ld3(t.v2.p) - ld3(t.v0.p);
    const f3 t1 = ld3(t.v1.t) - ld3(t.v0.t);
    const f3 t2 = ld3(t.v2.t) - ld3(t.v0.t);
    const float det = (t1.x * t2.y - t1.y * t2.x);
    if (det == 0.0f) return hit.N;

    const float invDet = recip(det);
    const f3 T = normalize(invDet * (e1 * t2.y - e2 * t1.y));
    const f3 B = normalize(invDet * (e2 * t1.x - e1 * t2.x));
    f3 N;
    N.x = T.x * texNormal.x + B.x * texNormal.y + hit.N.x * texNormal.z;
    N.y = T.y * texNormal.x + B.y * texNormal.y + hit.N.y * texNormal.z;
    N.z = T.z * texNormal.x + B.z * texNormal.y + hit.N.z * texNormal.z;
    return normalize(N);
}

// utils.cl:251-259
void sampleAreaLight(const flx_area_light &light, float *pdf, f3 *p, uint32_t *seed)
{
    *pdf = recip(4.0f * light.size.x * light.size.y);
    *p = ld3(light.pos);
    const float r1 = rnd(seed);
    const float r2 = rnd(seed);
    *p += (r1 + r1 - 1.0f) * light.size.x * ld3(light.right);
    *p += (r2 + r2 - 1.0f) * light.size.y * ld3(light.up);
}

// env_map.cl:14-17; atan2pi(y,x) := atan2f(y,x)*(1/pi), acospi(x) := acosf(x)*(1/pi) (pinned)
f2 directionToUV(f3 dir)
{
    f2 uv;
    uv.x = (o_atan2f(dir.x, -dir.z) * kInvPI) * 0.5f + 0.5f;
    uv.y = o_acosf(dir.y) * kInvPI;
    return uv;
}

// env_map.cl:21-33
f3 uvToDirection(float u, float v, float *sinPhi)
{
    const float phi = v * kPI;
    *sinPhi = o_sinf(phi);
    const float cosPhi = o_cosf(phi);
    const float th = u * k2PI - kPI;
    const float sinTh = o_sinf(th), cosTh = o_cosf(th);
    return mk3(*sinPhi * sinTh, cosPhi, -*sinPhi * cosTh);
}

// read_imagef with CLK_NORMALIZED_COORDS_TRUE | CLK_ADDRESS_CLAMP_TO_EDGE | CLK_FILTER_LINEAR,
// OpenCL 1.2 spec §8.2 formula evaluated in full fp32 (hardware samplers use 8-bit weights).
f3 envLinear(const ShadeScene &s, float su, float sv)
{
    const int w = s.envW, h = s.envH;
    const float u = su * (float)w, v = sv * (float)h;
    const float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    int i0 = (int)fu, j0 = (int)fv;
    int i1 = i0 + 1, j1 = j0 + 1;
    const float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    i0 = i0 < 0 ? 0 : (i0 > w - 1 ? w - 1 : i0); i1 = i1 < 0 ? 0 : (i1 > w - 1 ? w - 1 : i1);
    j0 = j0 < 0 ? 0 : (j0 > h - 1 ? h - 1 : j0); j1 = j1 < 0 ? 0 : (j1 > h - 1 ? h - 1 : j1);
    const float *T00 = s.envRGBA + 4 * ((size_t)j0 * w + i0), *T10 = s.envRGBA + 4 * ((size_t)j0 * w + i1);
    const float *T01 = s.envRGBA + 4 * ((size_t)j1 * w + i0), *T11 = s.envRGBA + 4 * ((size_t)j1 * w + i1);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    f3 r;
    r.x = w00 * T00[0] + w10 * T10[0] + w01 * T01[0] + w11 * T11[0];
    r.y = w00 * T00[1] + w10 * T10[1] + w01 * T01[1] + w11 * T11[1];
    r.z = w00 * T00[2] + w10 * T10[2] + w01 * T01[2] + w11 * T11[2];
    return r;
}

f3 evalEnvMapDir(const ShadeScene &s, f3 dir) // env_map.cl:35-39
{
    const f2 uv = directionToUV(dir);
    return envLinear(s, uv.x, uv.y);
}

// env_map.cl:63-92 (alias method; v uses the flat index — SURVEY §8g.6)
void sampleEnvMapAlias(float r, f3 *L, float *pdfW, const ShadeScene &s)
{
    const int width = s.envW, height = s.envH;
    const int wh = width * height;
    const float x = r * (float)wh;
    int i = (int)floorf(x);
    if (i > wh - 1) i = wh - 1;
    const float mProb = s.probTable[i];
    const int uvInd = (x - (float)i < mProb) ? i : s.aliasTable[i];
    const float u = ((float)(uvInd % width) + 0.5f) * recip((float)width);
    const float v = ((float)uvInd + 0.5f) * recip((float)wh);
    float sinTh;
    *L = uvToDirection(u, v, &sinTh);
    if (sinTh != 0.0f) *pdfW = s.pdfTable[uvInd] * recip(sinTh);
    else *pdfW = 0.0f;
}

// env_map.cl:95-109
float envMapPdf(const ShadeScene &s, f3 direction)
{
    if (direction.y > 0.99f) return 0.0f;
    const f2 uv = directionToUV(direction);
    int iu = (int)floorf(uv.x * (float)s.envW); if (iu > s.envW - 1) iu = s.envW - 1;
    int iv = (int)floorf(uv.y * (float)s.envH); if (iv > s.envH - 1) iv = s.envH - 1;
    return s.pdfTable[iv * s.envW + iu] * rsqrt_(1.0f - direction.y * direction.y);
}

// fresnel.cl:5-31 / :34-60
float fresnelDielectric1(float cosThI, float etaI, float etaT, float *cosThetaT)
{
    const float sinThetaI = sqrtf(fmaxf(0.0f, 1.0f - cosThI * cosThI));
    const float sinThetaT = etaI / etaT * sinThetaI;
    *cosThetaT = sqrtf(fmaxf(0.0f, 1.0f - sinThetaT * sinThetaT));
    if (sinThetaT >= 1.0f) return 1.0f;
    const float etaT_cosThI = etaT * cosThI;
    const float etaI_cosThetaT = etaI * *cosThetaT;
    const float parl = (etaT_cosThI - etaI_cosThetaT) / (etaT_cosThI + etaI_cosThetaT);
    const float etaI_cosThI = etaI * cosThI;
    const float etaT_cosThetaT = etaT * *cosThetaT;
    const float perp = (etaI_cosThI - etaT_cosThetaT) / (etaI_cosThI + etaT_cosThetaT);
    return 0.5f * (parl * parl + perp * perp);
}
float fresnelDielectric(float cosThI, float etaI, float etaT)
{
    float c;
    return fresnelDielectric1(cosThI, etaI, etaT, &c);
}

// tonemap.cl:8-33
static inline f3 uc2Func(f3 x)
{
    const float A = 0.22f, B = 0.30f, CB = 0.03f, DE = 0.002f, DF = 0.06f, ExF = 0.033333333333333333333f;
    return ((x * (A * x + mk3(CB)) + mk3(DE)) / (x * (A * x + mk3(B)) + mk3(DF))) - mk3(ExF);
}

// mk_postprocess.cl:27-47
void postprocessPixel(const float in[4], float out[4], float exposure, uint32_t tmOperator)
{
    float c[4] = {in[0], in[1], in[2], in[3]};
    if (c[3] > 0.0f) { const float w = c[3]; for (int k = 0; k < 4; k++) c[k] = c[k] / w; }
    f3 col = mk3(c[0], c[1], c[2]);
    col *= exposure;
    if (tmOperator == 1) col = col / (mk3(1.0f) + col);
    if (tmOperator == 2) col = uc2Func(2.0f * col) / uc2Func(mk3(11.2f));
    if (tmOperator != 3) col = pow3(col, kOne2_2);
    out[0] = col.x; out[1] = col.y; out[2] = col.z; out[3] = c[3];
}

} // namespace orc
