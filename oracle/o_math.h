// TEST INFRASTRUCTURE — CPU oracle for the fluctus wavefront path (never linked by the product).
// Pinned float semantics for the restatement of the reference's OpenCL C:
// IEEE-754 binary32, round-to-nearest-even, no FMA contraction (-ffp-contract=off), denormals
// kept, native_recip(x) = 1.0f/x, rsqrt(x) = 1.0f/sqrtf(x), length = sqrtf(dot),
// normalize(v) = v * (1.0f/length(v)), dot/cross written out left to right, fmin/fmax
// NaN-ignoring (C fminf/fmaxf).  Parity status: PINNED — tests/test_reference_source_cpu.py asserts
// this restatement bit-identical to the reference's own .cl sources compiled as C++ (oracle/refcl.py), and
// tests/test_opencl_reference_gpu.py checks the CUDA path against the reference's untouched kernels on the
// B200's OpenCL platform — see DESIGN.md §4.
#pragma once
#include <cmath>
#include <cstdint>
#include <cfloat>

namespace orc {

struct f3 { float x, y, z; };
struct f2 { float x, y; };

static inline f3 mk3(float x, float y, float z) { f3 r = {x, y, z}; return r; }
static inline f3 mk3(float s) { f3 r = {s, s, s}; return r; }
static inline f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
static inline f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
static inline f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
static inline f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
static inline f3 &operator+=(f3 &a, f3 b) { a = a + b; return a; }
static inline f3 &operator*=(f3 &a, f3 b) { a = a * b; return a; }
static inline f3 &operator*=(f3 &a, float s) { a = a * s; return a; }
static inline f3 &operator/=(f3 &a, float s) { a = a / s; return a; }

static inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline f3 cross(f3 a, f3 b)
{
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(f3 v) { return sqrtf(dot(v, v)); }
static inline f3 normalize(f3 v) { return v * (1.0f / length(v)); }
static inline float recip(float x) { return 1.0f / x; }
static inline float rsqrt_(float x) { return 1.0f / sqrtf(x); }
static inline f3 fmin3(f3 a, f3 b) { return mk3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline f3 fmax3(f3 a, f3 b) { return mk3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static inline bool isZero(f3 v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }
// transcendental functions: see fluctus_b200/csrc/flx_math.cuh (FLX_EXACT_LIBM) — the same switch on this side
#ifdef ORC_EXACT_LIBM
static inline float o_atan2f(float y, float x) { return (float)atan2((double)y, (double)x); }
static inline float o_acosf(float x) { return (float)acos((double)x); }
static inline float o_powf(float x, float y) { return (float)pow((double)x, (double)y); }
static inline float o_sinf(float x) { return (float)sin((double)x); }
static inline float o_cosf(float x) { return (float)cos((double)x); }
#else
static inline float o_atan2f(float y, float x) { return atan2f(y, x); }
static inline float o_acosf(float x) { return acosf(x); }
static inline float o_powf(float x, float y) { return powf(x, y); }
static inline float o_sinf(float x) { return sinf(x); }
static inline float o_cosf(float x) { return cosf(x); }
#endif
static inline f3 pow3(f3 v, float e) { return mk3(o_powf(v.x, e), o_powf(v.y, e), o_powf(v.z, e)); }

// constants of reference src/geom.h:11-35
static const float kPI = 3.14159265358979323846f;
static const float kInvPI = 0.3183098861837907f;
static const float k2PI = 6.2831853071795864f;
static const float kOne255 = 0.003921568627451f;
static const float kOne2_2 = 0.454545454545454545f;
static const float kOne2p32 = 2.3283064365386963e-10f;
static const float kInv2PiPi = 0.05066059182116888572193973160486f;

} // namespace orc
