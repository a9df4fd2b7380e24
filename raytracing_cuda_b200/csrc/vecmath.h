/*
 * vecmath.h -- float3 algebra for the hot path, host and device.
 *
 * Semantics follow the reference's operator set (reference structs.h:54-101)
 * because bit-level parity depends on them:
 *   - dot/cross are written as plain a*b+c*d chains so nvcc contracts them to
 *     FMA the same way it does for the reference (-fmad=true);
 *   - normalize() is "scale by the float-rounded FP64 reciprocal of the norm"
 *     (reference structs.h:82-84).  (float)(1.0/(double)n) equals the
 *     correctly rounded FP32 reciprocal for every finite n (a 24-bit n cannot
 *     put 1/n within 2^-53 of a float midpoint), so the device uses
 *     __frcp_rn and the host uses IEEE 1.0f/n; tests/test_numerics.py checks
 *     the identity exhaustively over sampled floats;
 *   - length() is norm3df on the device and sqrtf(dot) on the host, as in the
 *     reference (structs.h:75-81).
 */
#ifndef RTB_VECMATH_H
#define RTB_VECMATH_H

#include <math.h>
#include "../../include/rtb_structs.h"

#ifdef __CUDACC__
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

typedef rtb_float3 vec3;

RT_HD vec3 v3(float x, float y, float z) { vec3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD vec3 vadd(const vec3& a, const vec3& b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD vec3 vsub(const vec3& a, const vec3& b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD float vdot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RT_HD vec3 vscale(const vec3& a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
RT_HD vec3 vcross(const vec3& a, const vec3& b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
RT_HD vec3 vhadamard(const vec3& a, const vec3& b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }

RT_HD float vlength(const vec3& v) {
#ifdef __CUDA_ARCH__
    return norm3df(v.x, v.y, v.z);
#else
    return sqrtf(vdot(v, v));
#endif
}

/* (float)(1.0 / (double)n) */
RT_HD float recip_f64_rounded(float n) {
#ifdef __CUDA_ARCH__
    return __frcp_rn(n);
#else
    return (float)(1.0 / (double)n);
#endif
}

RT_HD vec3 vnormalize(const vec3& v) { return vscale(v, recip_f64_rounded(vlength(v))); }

RT_HD float clampf(float x, float a, float b) { return fmaxf(a, fminf(b, x)); }

#endif /* RTB_VECMATH_H */
