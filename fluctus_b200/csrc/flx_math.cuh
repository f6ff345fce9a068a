// Device-side float3 arithmetic with a fixed operation order.  The library is compiled with
// -fmad=false -prec-div=true -prec-sqrt=true -ftz=false so that +,-,*,/ and sqrt are the IEEE
// binary32 operations the parity contract is stated in: no FMA
// contraction, no flush-to-zero.  fminf/fmaxf ignore NaNs exactly like OpenCL fmin/fmax.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define FLX_DEV __device__ __forceinline__

struct f3 { float x, y, z; };

FLX_DEV f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
FLX_DEV f3 mk3(float s) { return mk3(s, s, s); }
FLX_DEV f3 mk3(float4 v) { return mk3(v.x, v.y, v.z); }
FLX_DEV f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
FLX_DEV f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
FLX_DEV f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
FLX_DEV f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
FLX_DEV f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
FLX_DEV f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
FLX_DEV f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
FLX_DEV f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
FLX_DEV f3 &operator+=(f3 &a, f3 b) { a = a + b; return a; }
FLX_DEV f3 &operator*=(f3 &a, f3 b) { a = a * b; return a; }
FLX_DEV f3 &operator*=(f3 &a, float s) { a = a * s; return a; }
FLX_DEV f3 &operator/=(f3 &a, float s) { a = a / s; return a; }

FLX_DEV float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
FLX_DEV f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
FLX_DEV float length(f3 v) { return sqrtf(dot(v, v)); }
FLX_DEV f3 normalize(f3 v) { return v * (1.0f / length(v)); }
FLX_DEV float recip(float x) { return 1.0f / x; }
FLX_DEV float rsqrt_(float x) { return 1.0f / sqrtf(x); }
FLX_DEV bool isZero(f3 v) { return v.x == 0.0f && v.y == 0.0f && v.z == 0.0f; }
FLX_DEV float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
FLX_DEV f3 pow3(f3 v, float e) { return mk3(powf(v.x, e), powf(v.y, e), powf(v.z, e)); }
FLX_DEV float4 mk4(f3 v, float w) { return make_float4(v.x, v.y, v.z, w); }
FLX_DEV float4 mk4u(f3 v, uint32_t w) { return make_float4(v.x, v.y, v.z, __uint_as_float(w)); }
FLX_DEV float4 mk4i(f3 v, int32_t w) { return make_float4(v.x, v.y, v.z, __int_as_float(w)); }

// constants of reference src/geom.h:11-35
#define FLX_PI 3.14159265358979323846f
#define FLX_INV_PI 0.3183098861837907f
#define FLX_2PI 6.2831853071795864f
#define FLX_ONE_255 0.003921568627451f
#define FLX_ONE_2_2 0.454545454545454545f
#define FLX_ONE_2P32 2.3283064365386963e-10f
#define FLX_FLT_MAX 3.402823466e+38f
