/*
 * fxaa_core.cuh -- the FXAA arithmetic (reference kernel.cu:286-402) shared by the FXAA kernels (fxaa.cu) and
 * the FXAA phase of the persistent tracer (trace_small.cu).
 *
 * Spelled with explicit rounding intrinsics in the fused shape the rebuilt reference has in SASS
 * (luminance = min(255, fma(b,c3, fma(r,c1, g*c2))) * (1/255)f, blend = fma(1-b, centre, b*neighbour),
 * smoothstep = (3-(f+f))*(f*f), ...), so every kernel and the reference agree bit for bit on identical input.
 */
#ifndef RTB_FXAA_CORE_CUH
#define RTB_FXAA_CORE_CUH

#include <cuda_runtime.h>

/* byte k of c as a float, exactly: the byte is permuted into the mantissa of 2^23 and 2^23 subtracted (PRMT on the
 * ALU pipe + FADD on the FMA pipe, both full rate; I2F.U8 issues at a quarter of that and bounded the luminance
 * phase) */
template <int K>
__device__ __forceinline__ float fx_byte(unsigned int c) {
    return __fadd_rn(__uint_as_float(__byte_perm(c, 0x4B000000u, 0x7650 + K)), -8388608.0f);
}

__device__ __forceinline__ float fx_lum(unsigned int c) {
    const float r = fx_byte<0>(c), g = fx_byte<1>(c), b = fx_byte<2>(c);
    float l = __fmaf_rn(b, 0.0721750f, __fmaf_rn(r, 0.2126729f, __fmul_rn(g, 0.7151522f)));
    return __fmul_rn(fminf(l, 255.0f), 0.0039215688593685626984f);
}

/* rgbToInt (reference kernel.cu:26-32: clamp to [0, 255], truncate, pack) for the blend of fx_blend, whose channels
 * are convex combinations of two bytes: (1 - b) c + b n with b in [0, 1] lies in [0, 255.0001], so the clamps never
 * act and truncation is all that is left.  v + 2^23 rounded TOWARDS ZERO leaves trunc(v) in the low mantissa bits
 * (FADD.RZ, full rate; F2I issues at a quarter of that, like I2F); two PRMTs gather the three low bytes (bytes 1 and
 * 2 of 0x4B0000xx are zero).  tests/test_numerics.py checks the identity for every byte pair and weight class. */
__device__ __forceinline__ unsigned int fx_pack(float r, float g, float b) {
    const unsigned int ri = __float_as_uint(__fadd_rz(r, 8388608.0f));
    const unsigned int gi = __float_as_uint(__fadd_rz(g, 8388608.0f));
    const unsigned int bi = __float_as_uint(__fadd_rz(b, 8388608.0f));
    return __byte_perm(__byte_perm(ri, gi, 0x1140), bi, 0x3410);
}

/* Decide from the nine luminances.  Returns false when the pixel passes through
 * (kernel.cu:349-354); else the blend factor and the neighbour offset (kernel.cu:369-392). */
__device__ __forceinline__ bool fx_decide(float M, float E, float W, float N, float S, float NE, float NW,
                                          float SE, float SW, float& b, int& dx, int& dy) {
    const float high = fmaxf(fmaxf(fmaxf(fmaxf(E, W), N), S), M);
    const float low = fminf(fminf(fminf(fminf(E, W), N), S), M);
    const float contrast = __fadd_rn(high, -low);
    const float threshold = fmaxf(0.0312f, __fmul_rn(0.063f, high));
    if (contrast < threshold) return false;

    float sum = __fadd_rn(__fadd_rn(__fadd_rn(E, W), S), N);
    sum = __fadd_rn(__fadd_rn(__fadd_rn(__fmaf_rn(sum, 2.0f, NE), NW), SE), SW);
    /* (float)((double)sum / 12.0) as the reference computes it (kernel.cu:372) equals the correctly rounded FP32
     * quotient: sum / 12 = m / 3 * 2^k is either exact or has the period-2 binary expansion of a third, which
     * cannot round to a 24-bit midpoint at 53 bits, so rounding twice changes nothing.  The FP32 quotient by FMA
     * correction of sum * RN(1/12) (Markstein): q0 + (sum - q0 * 12) * rc, three FP32 instructions and no
     * conversion to and from FP64 (two quarter-rate F2F).  tests/test_numerics.py checks both steps exactly. */
    const float q0 = __fmul_rn(sum, 0.083333335816860198974609375f);
    float filter = __fmaf_rn(__fmaf_rn(-q0, 12.0f, sum), 0.083333335816860198974609375f, q0);
    float rc;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(contrast));
    filter = fminf(__fmul_rn(fabsf(__fadd_rn(filter, -M)), rc), 1.0f);
    b = __fmul_rn(__fadd_rn(3.0f, -__fadd_rn(filter, filter)), __fmul_rn(filter, filter));

    const float hor = __fadd_rn(__fmaf_rn(fabsf(__fadd_rn(__fadd_rn(N, S), -__fadd_rn(M, M))), 2.0f,
                                          fabsf(__fadd_rn(__fadd_rn(NE, SE), -__fadd_rn(E, E)))),
                                fabsf(__fadd_rn(__fadd_rn(NW, SW), -__fadd_rn(W, W))));
    const float ver = __fadd_rn(__fmaf_rn(fabsf(__fadd_rn(__fadd_rn(E, W), -__fadd_rn(M, M))), 2.0f,
                                          fabsf(__fadd_rn(__fadd_rn(NE, NW), -__fadd_rn(N, N)))),
                                fabsf(__fadd_rn(__fadd_rn(SE, SW), -__fadd_rn(S, S))));
    dx = 0;
    dy = 0;
    if (hor >= ver) dy = fabsf(__fadd_rn(N, -M)) >= fabsf(__fadd_rn(S, -M)) ? -1 : 1;
    else            dx = fabsf(__fadd_rn(E, -M)) >= fabsf(__fadd_rn(W, -M)) ? 1 : -1;
    return true;
}

__device__ __forceinline__ unsigned int fx_blend(unsigned int centre, unsigned int nb, float b) {
    const float omb = __fadd_rn(1.0f, -b);
    const float r = __fmaf_rn(omb, fx_byte<0>(centre), __fmul_rn(b, fx_byte<0>(nb)));
    const float g = __fmaf_rn(omb, fx_byte<1>(centre), __fmul_rn(b, fx_byte<1>(nb)));
    const float bl = __fmaf_rn(omb, fx_byte<2>(centre), __fmul_rn(b, fx_byte<2>(nb)));
    return fx_pack(r, g, bl);
}

#endif
