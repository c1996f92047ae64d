// gmm_math.cuh -- bit-exact device restatement of the reference's SIMD CDF math.
//
// What is matched (reference paths under /root/reference):
//   exp256_ps                      compressai/cpp_exts/rans/avx_mathfun.h:250-304
//   _polya/_as/_logistic_approx_*  compressai/cpp_exts/rans/rans_interface.cpp:135-146, 154-186, 208-214
//   _fast_gmm_cdf<4> (SIMD branch) compressai/cpp_exts/rans/rans_interface.cpp:259-284
//   uint16 quantisation            compressai/cpp_exts/rans/rans_interface.cpp:509-510
// as compiled by GCC -O3 with FMA available (-ffp-contract=fast fuses 9 mul+add
// pairs). Every operation below is an explicit
// round-to-nearest intrinsic, which nvcc never contracts or reassociates, so the
// results do not depend on -fmad / -use_fast_math.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace flashgmm {

enum ApproxMode : int { MODE_POLYA = 0, MODE_AS = 1, MODE_LOGISTIC = 2 };

// Constants are given as the binary32 bit patterns the reference's literals
// round to (double literal -> float, as _PS256_CONST does).
#define FGMM_F(bits) __uint_as_float(bits)
#define FGMM_EXP_HI    FGMM_F(0x42b0c0a5u)  //  88.3762626647949f
#define FGMM_EXP_LO    FGMM_F(0xc2b0c0a5u)  // -88.3762626647949f
#define FGMM_LOG2EF    FGMM_F(0x3fb8aa3bu)  //  1.44269504088896341
#define FGMM_EXP_C1    FGMM_F(0x3f318000u)  //  0.693359375
#define FGMM_EXP_C2    FGMM_F(0xb95e8083u)  // -2.12194440e-4
#define FGMM_EXP_P0    FGMM_F(0x39506967u)  //  1.9875691500E-4
#define FGMM_EXP_P1    FGMM_F(0x3ab743ceu)  //  1.3981999507E-3
#define FGMM_EXP_P2    FGMM_F(0x3c088908u)  //  8.3334519073E-3
#define FGMM_EXP_P3    FGMM_F(0x3d2aa9c1u)  //  4.1665795894E-2
#define FGMM_EXP_P4    FGMM_F(0x3e2aaaaau)  //  1.6666665459E-1
#define FGMM_EXP_P5    FGMM_F(0x3f000000u)  //  5.0000001201E-1 (== 0.5f)
#define FGMM_M2_OVER_PI FGMM_F(0xbf22f983u) // -2.0f / 3.14159274f
#define FGMM_INV_SQRT_2PI FGMM_F(0x3ecc422au) // 0.3989422804014327f
#define FGMM_AS_P      FGMM_F(0x3e6d3389u)  //  0.2316419f
#define FGMM_AS_B1     FGMM_F(0x3ea385fau)  //  0.319381530f
#define FGMM_AS_B2     FGMM_F(0xbeb68f87u)  // -0.356563782f
#define FGMM_AS_B3     FGMM_F(0x3fe40778u)  //  1.781477937f
#define FGMM_AS_B4     FGMM_F(0xbfe91eeau)  // -1.821255978f
#define FGMM_AS_B5     FGMM_F(0x3faa466fu)  //  1.330274429f
#define FGMM_LOGI_K    FGMM_F(0x3fd9db23u)  //  1.702f

// exp256_ps, one lane. NaN handling follows _mm256_min_ps/_mm256_max_ps
// (second operand wins), which is also what fminf/fmaxf give for a NaN first operand.
__device__ __forceinline__ float exp256(float x) {
    x = fminf(x, FGMM_EXP_HI);
    x = fmaxf(x, FGMM_EXP_LO);
    float fx = __fmaf_rn(x, FGMM_LOG2EF, 0.5f);
    // floorf + __float2int_rz are FRND + F2I on the quarter-rate XU pipe. A magic-number form on the FP32 / integer
    // pipes (add 1.5 * 2^23, read the integer from the mantissa, one compare; exact for every finite input,
    // tests/native/check_exact_math.c) was measured in round 2: it relieves the XU pipe (64-71 % busy in the
    // startfreq kernel) but costs three more issue slots per exp, and these kernels are issue-bound (0.81-0.88
    // warp-instructions per cycle and scheduler): startfreq +4 %, rows mode 2 +9 % slower. Not used.
    fx = floorf(fx);                       // floor, then "if greater subtract 1" is a no-op
    x = __fmaf_rn(-fx, FGMM_EXP_C1, x);
    x = __fmaf_rn(-fx, FGMM_EXP_C2, x);
    float z = __fmul_rn(x, x);
    float y = FGMM_EXP_P0;
    y = __fmaf_rn(y, x, FGMM_EXP_P1);
    y = __fmaf_rn(y, x, FGMM_EXP_P2);
    y = __fmaf_rn(y, x, FGMM_EXP_P3);
    y = __fmaf_rn(y, x, FGMM_EXP_P4);
    y = __fmaf_rn(y, x, FGMM_EXP_P5);
    y = __fmaf_rn(y, z, x);
    y = __fadd_rn(y, 1.0f);
    int n = __float2int_rz(fx);            // |fx| <= 128: no saturation difference vs cvttps2dq
    return __fmul_rn(y, __int_as_float((n + 127) << 23));
}

template <int MODE>
__device__ __forceinline__ float phi(float t);

// (Round 2 also tried a branch-free square root for 1 - exp(.) in [0, 1] -- MUFU.RSQ seed, one coupled Newton step, one
// residual correction -- to drop __fsqrt_rn's range check and slow-path call, 10 % of the rows kernel's instructions
// being BSSY / BRA / BSYNC. The exhaustive device comparison against __fsqrt_rn found mismatches, and it was 4 % SLOWER
// in both the startfreq and the rows kernel. Rejected.)
template <>
__device__ __forceinline__ float phi<MODE_POLYA>(float t) {
    float a = __fmul_rn(FGMM_M2_OVER_PI, __fmul_rn(t, t));
    float s = __fsqrt_rn(__fsub_rn(1.0f, exp256(a)));
    s = __uint_as_float((__float_as_uint(t) & 0x80000000u) | (__float_as_uint(s) & 0x7fffffffu));
    return __fmul_rn(0.5f, __fadd_rn(1.0f, s));
}

template <>
__device__ __forceinline__ float phi<MODE_AS>(float t) {
    float ax = fabsf(t);
    float zx = __fmul_rn(FGMM_INV_SQRT_2PI, exp256(__fmul_rn(__fmul_rn(t, t), -0.5f)));
    float u = __fdiv_rn(1.0f, __fmaf_rn(FGMM_AS_P, ax, 1.0f));
    float p = __fmaf_rn(FGMM_AS_B5, u, FGMM_AS_B4);
    p = __fmaf_rn(p, u, FGMM_AS_B3);
    p = __fmaf_rn(p, u, FGMM_AS_B2);
    p = __fmaf_rn(p, u, FGMM_AS_B1);
    p = __fmul_rn(p, u);
    float rp = __fmaf_rn(-zx, p, 1.0f);
    float rn = __fsub_rn(1.0f, rp);
    return (__float_as_uint(t) & 0x80000000u) ? rn : rp;
}

template <>
__device__ __forceinline__ float phi<MODE_LOGISTIC>(float t) {
    float e = exp256(-__fmul_rn(FGMM_LOGI_K, t));   // (-1)*(k*t): the sign flip is exact
    return __fdiv_rn(1.0f, __fadd_rn(1.0f, e));
}

// 4-component mixture CDF at x with the reference's (p0+p1)+(p2+p3) association.
template <int MODE>
__device__ __forceinline__ float gmm_cdf4(float x, const float (&mu)[4], const float (&sg)[4],
                                          const float (&w)[4]) {
    float p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float t = __fdiv_rn(__fsub_rn(x, mu[k]), sg[k]);
        p[k] = __fmul_rn(w[k], phi<MODE>(t));
    }
    return __fadd_rn(__fadd_rn(p[0], p[1]), __fadd_rn(p[2], p[3]));
}

// a / s for a divisor whose correctly rounded reciprocal y = RN(1 / s) is shared by many dividends (the 33 boundaries of
// a symbol's window all divide by the same sigma): multiply, then two residual corrections -- after the first the
// quotient is faithful, and Markstein's theorem makes the second correctly rounded. Five FP32-pipe operations instead
// of MUFU.RCP + Newton steps + range check + slow-path call. Bit-identical to __fdiv_rn for 2^-10 <= s <= 2^16 and the
// dividends this path sees (no overflow / underflow of quotient or residual): tests/native/check_exact_math.c,
// 2 * 10^9 random pairs incl. all-ones and power-of-two significands, 0 mismatches.
__device__ __forceinline__ float div_by_shared_rcp(float a, float s, float y) {
    const float q0 = __fmul_rn(a, y);
    const float q1 = __fmaf_rn(__fmaf_rn(-q0, s, a), y, q0);
    return __fmaf_rn(__fmaf_rn(-q1, s, a), y, q1);
}

// mixture CDF with the reciprocals of the four scales precomputed (rcp[k] = __frcp_rn(sg[k]))
template <int MODE>
__device__ __forceinline__ uint32_t gmm_q16_rcp(float x, const float (&mu)[4], const float (&sg)[4], const float (&rcp)[4],
                                                const float (&w)[4]);

// static_cast<uint16_t>(cdf * 65535): cvttss2si then low 16 bits. x86 returns
// 0x80000000 for NaN / |v| >= 2^31 (low half 0); __float2int_rz saturates, so
// out-of-range is mapped to 0 explicitly.
__device__ __forceinline__ uint32_t q16(float cdf) {
    float v = __fmul_rn(cdf, 65535.0f);
    int i = __float2int_rz(v);
    if (!(fabsf(v) < 2147483648.0f)) i = 0;
    return (uint32_t)i & 0xffffu;
}

template <int MODE>
__device__ __forceinline__ uint32_t gmm_q16(float x, const float (&mu)[4], const float (&sg)[4],
                                            const float (&w)[4]) {
    return q16(gmm_cdf4<MODE>(x, mu, sg, w));
}

template <int MODE>
__device__ __forceinline__ uint32_t gmm_q16_rcp(float x, const float (&mu)[4], const float (&sg)[4], const float (&rcp)[4],
                                                const float (&w)[4]) {
    float p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float t = div_by_shared_rcp(__fsub_rn(x, mu[k]), sg[k], rcp[k]);
        p[k] = __fmul_rn(w[k], phi<MODE>(t));
    }
    return q16(__fadd_rn(__fadd_rn(p[0], p[1]), __fadd_rn(p[2], p[3])));
}

// Packed per-symbol record handed to the host coder: start | freq << 16.
// freq == 0 never occurs for a coded symbol, so it marks the bypass case
// (reference: pmf == 0 -> marker (65535,1) + raw nibbles, rans_interface.cpp:512-552).
#define FGMM_PACK_BYPASS 0x0000ffffu

template <int MODE>
__device__ __forceinline__ uint32_t gmm_startfreq1(int v, const float (&mu)[4], const float (&sg)[4],
                                                   const float (&w)[4]) {
    float x1 = __fsub_rn((float)v, 0.5f);
    float x2 = __fadd_rn(x1, 1.0f);
    uint32_t c1 = gmm_q16<MODE>(x1, mu, sg, w);
    uint32_t c2 = gmm_q16<MODE>(x2, mu, sg, w);
    uint32_t pmf = (c2 - c1) & 0xffffu;
    return pmf == 0 ? FGMM_PACK_BYPASS : (c1 | (pmf << 16));
}

}  // namespace flashgmm
