// The transcendental functions of the path.  They are the ONLY operations whose result is not pinned by IEEE 754
// (CUDA's libdevice and the host's libm round differently in the last place now and then), i.e. the only source of the
// 1e-5 per-stage tolerance.  -DFLX_EXACT_LIBM=1 builds a second library (libfluctus_cuda_exactlibm.so, tests only) that
// evaluates them in double precision and rounds once; the oracle has the same switch (oracle/o_math.h), and with both
// on the stage-by-stage comparison is bit-exact (tests/test_exact_libm_gpu.py) — the proof that nothing else differs.
#pragma once
#include "flx_math.cuh"

#ifndef FLX_EXACT_LIBM
#define FLX_EXACT_LIBM 0
#endif
#if FLX_EXACT_LIBM
FLX_DEV float flx_atan2f(float y, float x) { return (float)atan2((double)y, (double)x); }
FLX_DEV float flx_acosf(float x) { return (float)acos((double)x); }
FLX_DEV float flx_powf(float x, float y) { return (float)pow((double)x, (double)y); }
FLX_DEV void flx_sincosf(float x, float *s, float *c) { double ds, dc; sincos((double)x, &ds, &dc); *s = (float)ds; *c = (float)dc; }
#else
FLX_DEV float flx_atan2f(float y, float x) { return atan2f(y, x); }
FLX_DEV float flx_acosf(float x) { return acosf(x); }
FLX_DEV float flx_powf(float x, float y) { return powf(x, y); }
FLX_DEV void flx_sincosf(float x, float *s, float *c) { sincosf(x, s, c); }
#endif
FLX_DEV f3 pow3x(f3 v, float e) { return mk3(flx_powf(v.x, e), flx_powf(v.y, e), flx_powf(v.z, e)); } // pow3 of flx_math.cuh through the switch
