// TEST INFRASTRUCTURE — an OpenCL C dialect on top of C++ so that the reference's UNMODIFIED device
// sources (/root/reference/src/*.cl + geom.h) compile with g++ into oracle/_ref/ and run on the host.
// Nothing of the reference is copied here: this header only supplies what the OpenCL C language and
// its built-in library give a kernel (vector types with swizzles, address-space keywords, work-item
// functions, atomics, images, math built-ins).  The one thing C++ cannot parse — the vector literal
// `(float3)(a, b, c)`, a comma expression in C++ — is rewritten to `float3{a, b, c}` on the fly
// by oracle/ref_cl/clpp.py into oracle/_ref/cl_gen/ (git-ignored).
//
// Built-in semantics (the same pinned reading the oracle states in o_math.h; the reference itself is
// JIT-compiled with -cl-fast-relaxed-math, clcontext.cpp:147, so its own bits are device-defined):
// IEEE binary32, no contraction (the TU is compiled with -ffp-contract=off), native_recip(x) = 1.0f/x,
// rsqrt(x) = 1.0f/sqrtf(x), length = sqrtf(dot), normalize(v) = v * (1.0f/length(v)), dot / cross
// evaluated left to right, fmin/fmax NaN-ignoring, atan2pi = atan2f * (1/pi), acospi = acosf * (1/pi),
// sincos = sinf + cosf, read_imagef(linear) per the OpenCL 1.2 specification formula in fp32.
#pragma once
#include <cmath>
#include <cstdint>
#include <cfloat>
#include <cstdio>
#include <cstring>
#include <cstddef>

typedef unsigned int uint;
typedef unsigned char uchar;
typedef unsigned short ushort;
typedef unsigned long ulong;

// ---- vector types ------------------------------------------------------------------------------------
// OpenCL's sizes and alignments (float3 occupies 16 bytes).  The constructors are what the rewritten
// vector literals `float3{a, b, c}` call; a scalar widens implicitly, as in OpenCL C (ggx.cl:189,251).
// Swizzles: the sources use `lerp(...).xy` (bvh.cl:48,278), `read_imagef(...).xyz` (env_map.cl:40-50),
// `val.xyz = pow(val.xyz, 2.2f)` on a float3 (utils.cl:164) and `color.xyz op= ...` on a float4
// (mk_postprocess.cl:34-44).  C++ has no swizzle members, so clpp.py rewrites `.xy` / `.xyz` to the
// member calls `.xy_()` / `.xyz_()`; the latter returns a proxy that reads and writes lanes 0..2.
struct alignas(8) float2
{
    float x, y;
    float2() = default;
    inline float2(float a, float b) : x(a), y(b) {}
    inline float2(float s) : x(s), y(s) {}
};
struct float3;
struct lanes3
{
    float *p;
    inline operator float3() const;
    inline lanes3 &operator=(const float3 &v);
    inline lanes3 &operator=(const lanes3 &o) { const float a = o.p[0], b = o.p[1], c = o.p[2]; p[0] = a; p[1] = b; p[2] = c; return *this; }
    inline lanes3 &operator*=(float s) { p[0] *= s; p[1] *= s; p[2] *= s; return *this; }
};
struct alignas(16) float3
{
    float x, y, z, w_;
    float3() = default;
    inline float3(float a, float b, float c) : x(a), y(b), z(c), w_(0.0f) {}
    inline float3(float s) : x(s), y(s), z(s), w_(0.0f) {}
    inline float2 xy_() const { return float2(x, y); }
    inline lanes3 xyz_() { lanes3 l; l.p = &x; return l; }
};
struct alignas(16) float4
{
    float x, y, z, w;
    float4() = default;
    inline float4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    inline float4(float s) : x(s), y(s), z(s), w(s) {}
    inline float4(const float3 &v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    inline lanes3 xyz_() { lanes3 l; l.p = &x; return l; }
};
struct alignas(8) int2
{
    int x, y;
    int2() = default;
    inline int2(int a, int b) : x(a), y(b) {}
    inline int2(int s) : x(s), y(s) {}
    inline int2(uint a, uint b) : x((int)a), y((int)b) {}
    inline int2(float a, float b) : x((int)a), y((int)b) {} // scalar float -> int: truncation (utils.cl:146)
};

static_assert(sizeof(float2) == 8 && sizeof(float3) == 16 && sizeof(float4) == 16 && sizeof(int2) == 8, "OpenCL vector sizes");

static inline float2 make_float2(float a, float b) { return float2(a, b); }
static inline float3 make_float3(float a, float b, float c) { return float3(a, b, c); }
static inline float4 make_float4(float a, float b, float c, float d) { return float4(a, b, c, d); }
static inline int2 make_int2(int a, int b) { return int2(a, b); }
inline lanes3::operator float3() const { return float3(p[0], p[1], p[2]); }
inline lanes3 &lanes3::operator=(const float3 &v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; return *this; }

// ---- arithmetic ----------------------------------------------------------------------------------------------
#define CLSHIM_BINOP3(op) \
    static inline float3 operator op(const float3 &a, const float3 &b) { return make_float3(a.x op b.x, a.y op b.y, a.z op b.z); } \
    static inline float3 operator op(const float3 &a, float s) { return make_float3(a.x op s, a.y op s, a.z op s); } \
    static inline float3 operator op(float s, const float3 &a) { return make_float3(s op a.x, s op a.y, s op a.z); } \
    static inline float3 &operator op##=(float3 &a, const float3 &b) { a.x op## = b.x; a.y op## = b.y; a.z op## = b.z; return a; } \
    static inline float3 &operator op##=(float3 &a, float s) { a.x op## = s; a.y op## = s; a.z op## = s; return a; }
CLSHIM_BINOP3(+)
CLSHIM_BINOP3(-)
CLSHIM_BINOP3(*)
CLSHIM_BINOP3(/)
static inline float3 operator-(const float3 &a) { return make_float3(-a.x, -a.y, -a.z); }

#define CLSHIM_BINOP4(op) \
    static inline float4 operator op(const float4 &a, const float4 &b) { return make_float4(a.x op b.x, a.y op b.y, a.z op b.z, a.w op b.w); } \
    static inline float4 operator op(const float4 &a, float s) { return make_float4(a.x op s, a.y op s, a.z op s, a.w op s); } \
    static inline float4 operator op(float s, const float4 &a) { return make_float4(s op a.x, s op a.y, s op a.z, s op a.w); } \
    static inline float4 &operator op##=(float4 &a, const float4 &b) { a.x op## = b.x; a.y op## = b.y; a.z op## = b.z; a.w op## = b.w; return a; } \
    static inline float4 &operator op##=(float4 &a, float s) { a.x op## = s; a.y op## = s; a.z op## = s; a.w op## = s; return a; }
CLSHIM_BINOP4(+)
CLSHIM_BINOP4(-)
CLSHIM_BINOP4(*)
CLSHIM_BINOP4(/)

#define CLSHIM_BINOP2(op) \
    static inline float2 operator op(const float2 &a, const float2 &b) { return make_float2(a.x op b.x, a.y op b.y); } \
    static inline float2 operator op(const float2 &a, float s) { return make_float2(a.x op s, a.y op s); } \
    static inline float2 operator op(float s, const float2 &a) { return make_float2(s op a.x, s op a.y); }
CLSHIM_BINOP2(+)
CLSHIM_BINOP2(-)
CLSHIM_BINOP2(*)
CLSHIM_BINOP2(/)

// ---- math built-ins ------------------------------------------------------------------------------------------
static inline float native_recip(float x) { return 1.0f / x; }
static inline float3 native_recip(const float3 &v) { return make_float3(1.0f / v.x, 1.0f / v.y, 1.0f / v.z); }
static inline float rsqrt(float x) { return 1.0f / sqrtf(x); }
static inline float native_powr(float x, float y) { return powf(x, y); }
static inline float sqrt(float x) { return sqrtf(x); }
static inline float fabs(float x) { return fabsf(x); }
static inline float floor(float x) { return floorf(x); }
static inline float sin(float x) { return sinf(x); }
static inline float cos(float x) { return cosf(x); }
static inline float tan(float x) { return tanf(x); }
static inline float atan2(float y, float x) { return atan2f(y, x); }
static inline float pow(float x, float y) { return powf(x, y); }
static inline float sincos(float x, float *c) { *c = cosf(x); return sinf(x); }
static const float CLSHIM_INV_PI = 0.3183098861837907f; // M_INV_PI of geom.h:30
static inline float atan2pi(float y, float x) { return atan2f(y, x) * CLSHIM_INV_PI; }
static inline float acospi(float x) { return acosf(x) * CLSHIM_INV_PI; }
static inline float fmin(float a, float b) { return fminf(a, b); }
static inline float fmax(float a, float b) { return fmaxf(a, b); }
static inline float3 fmin(const float3 &a, const float3 &b) { return make_float3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline float3 fmax(const float3 &a, const float3 &b) { return make_float3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline float3 pow(const float3 &v, float e) { return make_float3(powf(v.x, e), powf(v.y, e), powf(v.z, e)); }

// OpenCL min/max on gentype (integer and float); float versions are the plain comparisons the spec gives
static inline int min(int a, int b) { return b < a ? b : a; }
static inline int max(int a, int b) { return a < b ? b : a; }
static inline uint min(uint a, uint b) { return b < a ? b : a; }
static inline uint max(uint a, uint b) { return a < b ? b : a; }
static inline float min(float a, float b) { return b < a ? b : a; }
static inline float max(float a, float b) { return a < b ? b : a; }
static inline float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static inline int clamp(int x, int lo, int hi) { return min(max(x, lo), hi); }
static inline int2 clamp(int2 v, int2 lo, int2 hi) { return make_int2(clamp(v.x, lo.x, hi.x), clamp(v.y, lo.y, hi.y)); }
static inline int clz(uint x) { return x ? __builtin_clz(x) : 32; }

static inline float dot(const float3 &a, const float3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float dot(const float4 &a, const float4 &b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
static inline float3 cross(const float3 &a, const float3 &b)
{
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(const float3 &v) { return sqrtf(dot(v, v)); }
static inline float3 normalize(const float3 &v) { return v * (1.0f / length(v)); }

#define M_PI_F 3.14159274101257f

// ---- memory built-ins ----------------------------------------------------------------------------------------
static inline float4 vload4(size_t i, const float *p) { return make_float4(p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]); }
static inline void vstore4(const float4 &v, size_t i, float *p) { p[4 * i] = v.x; p[4 * i + 1] = v.y; p[4 * i + 2] = v.z; p[4 * i + 3] = v.w; }
static inline void vstore(uint v, size_t i, uint *p) { p[i] = v; } // wf_reset.cl:26 (scalar store)

// ---- work-item functions ---------------------------------------------------------------------------------------
static thread_local size_t clshim_gid = 0, clshim_gid1 = 0; // dimension 0 / dimension 1 of the NDRange
static inline size_t get_global_id(int dim) { return dim == 0 ? clshim_gid : (dim == 1 ? clshim_gid1 : 0); }
static inline size_t get_local_id(int) { return 0; }
static inline size_t get_local_size(int) { return 1; }
static inline size_t get_group_id(int) { return clshim_gid; }
#define CLK_LOCAL_MEM_FENCE 1
#define CLK_GLOBAL_MEM_FENCE 2
static inline void barrier(int) {}

// ---- atomics (sequentially consistent; the launchers may run work-items on several host threads) -------------------
static inline uint atomic_inc(volatile uint *p) { return __atomic_fetch_add(p, 1u, __ATOMIC_SEQ_CST); }
static inline uint atomic_add(volatile uint *p, uint v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline int atomic_add(volatile int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline float atomic_xchg(volatile float *p, float v)
{
    uint in, out; memcpy(&in, &v, 4);
    out = __atomic_exchange_n((volatile uint *)p, in, __ATOMIC_SEQ_CST);
    float r; memcpy(&r, &out, 4); return r;
}
static inline uint atomic_cmpxchg(volatile uint *p, uint cmp, uint val)
{
    __atomic_compare_exchange_n(p, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;
}

// ---- images ------------------------------------------------------------------------------------------------------
typedef unsigned int sampler_t;
#define CLK_NORMALIZED_COORDS_FALSE 0u
#define CLK_NORMALIZED_COORDS_TRUE 1u
#define CLK_ADDRESS_CLAMP_TO_EDGE 2u
#define CLK_FILTER_NEAREST 0u
#define CLK_FILTER_LINEAR 16u
struct clshim_image2d { const float *rgba; int width, height; };
typedef const clshim_image2d *image2d_t;
static inline int2 get_image_dim(image2d_t img) { return make_int2(img->width, img->height); }
static inline float4 clshim_texel(image2d_t img, int i, int j)
{
    i = i < 0 ? 0 : (i > img->width - 1 ? img->width - 1 : i);   // CLK_ADDRESS_CLAMP_TO_EDGE
    j = j < 0 ? 0 : (j > img->height - 1 ? img->height - 1 : j);
    const float *p = img->rgba + 4 * ((size_t)j * img->width + i);
    return make_float4(p[0], p[1], p[2], p[3]);
}
// OpenCL 1.2 specification 8.2: u = s*w, i0 = floor(u - 0.5), a = frac(u - 0.5),
// T = (1-a)(1-b) T00 + a(1-b) T10 + (1-a) b T01 + a b T11
static inline float4 read_imagef(image2d_t img, sampler_t smp, float2 uv)
{
    float u = uv.x, v = uv.y;
    if (smp & CLK_NORMALIZED_COORDS_TRUE) { u *= (float)img->width; v *= (float)img->height; }
    if (!(smp & CLK_FILTER_LINEAR)) return clshim_texel(img, (int)floorf(u), (int)floorf(v));
    const float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    const float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    const int i0 = (int)fu, j0 = (int)fv;
    const float4 t00 = clshim_texel(img, i0, j0), t10 = clshim_texel(img, i0 + 1, j0);
    const float4 t01 = clshim_texel(img, i0, j0 + 1), t11 = clshim_texel(img, i0 + 1, j0 + 1);
    return ((1.0f - a) * (1.0f - b)) * t00 + (a * (1.0f - b)) * t10 + ((1.0f - a) * b) * t01 + (a * b) * t11;
}
static inline float4 read_imagef(image2d_t img, sampler_t, int2 ij) { return clshim_texel(img, ij.x, ij.y); }

// ---- address-space and function qualifiers (defined LAST: no system header is included after this point) ---------
#define global
#define __global
#define local
#define __local
#define constant static const
#define __constant static const
#define kernel
#define __kernel
#define read_only
#define write_only
#define __read_only
#define __write_only
