// gmm_likelihood.cu -- fused forward / backward of GaussianMixtureConditional's
// likelihood (training + eval path), sm_100a.
//
// Reference (paths under /root/reference/compressai):
//   entropy_models/entropy_models.py:774-790   GaussianMixtureConditional._likelihood
//   entropy_models/entropy_models.py:720-737   GaussianConditional._likelihood
//   entropy_models/entropy_models.py:674-678   _standardized_cumulative = 0.5 * erfc(-x / sqrt(2))
//   entropy_models/entropy_models.py:792-808   forward (quantize "noise" | "dequantize", LowerBound 1e-9)
//   ops/bound_ops.py:36-42                     LowerBound forward / custom backward
//
// The reference runs K passes of ~12 elementwise kernels each over (B,M,H,W); this
// is one pass: 4 + 12K bytes in, 8 bytes out per element (60 B at K=4) -> HBM-bound.
// Layout is the reference's own: x is (B, M, H*W) contiguous; component k of
// channel c lives at channel k*M + c of the (B, K*M, H*W) parameter tensors
// (entropy_models.py:784-787), i.e. K planes of M*HW floats per batch item.
// Numerics follow the reference's operation order in fp32 with the exact erfcf
// (contract: 1e-5 relative; not a bit-exact path).
#include <stdlib.h>

#include "gmm_kernels.h"

namespace flashgmm {

namespace {

constexpr int kLkThreads = 256;
constexpr int kMaxK = 8;
#define FGMM_NEG_RSQRT2 (-0.70710678118654752440f)
#define FGMM_RSQRT_2PI 0.39894228040143267794f

struct LkGeom {
    int64_t plane;      // M * HW: elements per batch item of x, and stride between components
    int64_t total;      // B * plane
    int K;
};

__device__ __forceinline__ float std_cum(float t) {           // 0.5 * erfc(-(2^-0.5) * t)
    return 0.5f * erfcf(FGMM_NEG_RSQRT2 * t);
}

// one element: returns the (unbounded) mixture likelihood
__device__ __forceinline__ float lik1(float y, const float* __restrict__ sc, const float* __restrict__ mu,
                                      const float* __restrict__ w, int64_t pidx, int64_t plane, int K,
                                      float scale_bound) {
    float acc = 0.0f;
    for (int k = 0; k < K; ++k) {
        const float s = fmaxf(__ldg(sc + pidx + k * plane), scale_bound);
        const float v = fabsf(y - __ldg(mu + pidx + k * plane));
        const float up = std_cum((0.5f - v) / s);
        const float lo = std_cum((-0.5f - v) / s);
        acc = acc + (up - lo) * __ldg(w + pidx + k * plane);
    }
    return acc;
}

// quant: 0 = y = x (already quantised), 1 = y = x + noise, 2 = y = rint(x)
template <int VEC>
__global__ void __launch_bounds__(kLkThreads)
gmm_likelihood_fwd_kernel(const float* __restrict__ x, const float* __restrict__ noise, int quant,
                          const float* __restrict__ sc, const float* __restrict__ mu,
                          const float* __restrict__ w, LkGeom g, float scale_bound, float lik_bound,
                          float* __restrict__ y_out, float* __restrict__ lik_out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * VEC;
    for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; e < g.total; e += stride) {
        const int64_t b = e / g.plane, i = e - b * g.plane;
        const int64_t pidx = b * g.plane * g.K + i;
        float xv[VEC], yv[VEC], lv[VEC];
        if (VEC == 4) {
            float4 t = __ldcs(reinterpret_cast<const float4*>(x + e));
            xv[0] = t.x; xv[1] = t.y; xv[2] = t.z; xv[3] = t.w;
            if (quant == 1) {
                float4 nz = __ldcs(reinterpret_cast<const float4*>(noise + e));
                xv[0] += nz.x; xv[1] += nz.y; xv[2] += nz.z; xv[3] += nz.w;
            }
        } else {
            xv[0] = x[e];
            if (quant == 1) xv[0] += noise[e];
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) yv[j] = quant == 2 ? rintf(xv[j]) : xv[j];

        if (VEC == 4) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            for (int k = 0; k < g.K; ++k) {
                const float4 s4 = __ldcs(reinterpret_cast<const float4*>(sc + pidx + k * g.plane));
                const float4 m4 = __ldcs(reinterpret_cast<const float4*>(mu + pidx + k * g.plane));
                const float4 w4 = __ldcs(reinterpret_cast<const float4*>(w + pidx + k * g.plane));
                const float ss[4] = {s4.x, s4.y, s4.z, s4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w},
                            ww[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float s = fmaxf(ss[j], scale_bound);
                    const float v = fabsf(yv[j] - mm[j]);
                    const float up = std_cum((0.5f - v) / s);
                    const float lo = std_cum((-0.5f - v) / s);
                    acc[j] = acc[j] + (up - lo) * ww[j];
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) lv[j] = acc[j];
        } else {
            lv[0] = lik1(yv[0], sc, mu, w, pidx, g.plane, g.K, scale_bound);
        }
        if (lik_bound > 0.0f) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) lv[j] = fmaxf(lv[j], lik_bound);
        }
        if (VEC == 4) {
            __stcs(reinterpret_cast<float4*>(y_out + e), make_float4(yv[0], yv[1], yv[2], yv[3]));
            __stcs(reinterpret_cast<float4*>(lik_out + e), make_float4(lv[0], lv[1], lv[2], lv[3]));
        } else {
            y_out[e] = yv[0];
            lik_out[e] = lv[0];
        }
    }
}

// VEC consecutive floats with one streaming load / store (VEC = 1, 2 or 4; 4 * VEC-byte aligned)
template <int VEC>
__device__ __forceinline__ void ldv(const float* p, float (&v)[VEC]) {
    if constexpr (VEC == 4) { const float4 t = __ldcs(reinterpret_cast<const float4*>(p)); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
    else if constexpr (VEC == 2) { const float2 t = __ldcs(reinterpret_cast<const float2*>(p)); v[0] = t.x; v[1] = t.y; }
    else v[0] = *p;
}
template <int VEC>
__device__ __forceinline__ void stv(float* p, const float (&v)[VEC]) {
    if constexpr (VEC == 4) __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
    else if constexpr (VEC == 2) __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1]));
    else *p = v[0];
}

// Backward. Given g = dL/d(likelihood after the bound):
//   g' = g if (lik_raw >= bound) or (g < 0) else 0                         (bound_ops.py:40-42)
//   l_k = Phi(u) - Phi(d), u = (0.5 - v)/s, d = (-0.5 - v)/s, v = |y - mu_k|, s = max(sigma_k, sb)
//   dL/dw_k  = g' l_k
//   dl/dv    = (phi(d) - phi(u)) / s ;  dL/dy += g' w_k dl/dv sign(y - mu_k) ; dL/dmu_k = -that
//   dl/ds    = (phi(d) d - phi(u) u) / s ;  dL/dsigma_k = LowerBound-bwd(sigma_k, sb, g' w_k dl/ds)
// KT = compile-time K (register arrays sized exactly) or 0 = runtime K <= kMaxK.
template <int VEC, int KT>
__global__ void __launch_bounds__(kLkThreads)
gmm_likelihood_bwd_kernel(const float* __restrict__ y, const float* __restrict__ sc,
                          const float* __restrict__ mu, const float* __restrict__ w,
                          const float* __restrict__ grad, LkGeom g, float scale_bound, float lik_bound,
                          float* __restrict__ gy, float* __restrict__ gsc, float* __restrict__ gmu,
                          float* __restrict__ gw) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * VEC;
    for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; e < g.total; e += stride) {
        const int64_t b = e / g.plane, i = e - b * g.plane;
        const int64_t pidx = b * g.plane * g.K + i;
        float yv[VEC], gv[VEC];
        ldv<VEC>(y + e, yv);
        ldv<VEC>(grad + e, gv);
        constexpr int KA = KT ? KT : kMaxK;
        const int K = KT ? KT : g.K;
        float lk[KA][VEC], dv[KA][VEC], ds[KA][VEC], wk[KA][VEC], sraw[KA][VEC];
        float tot[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) tot[j] = 0.0f;
#pragma unroll
        for (int k = 0; k < KA; ++k) {
            if (k < K) {
                float ss[VEC], mm[VEC], ww[VEC];
                ldv<VEC>(sc + pidx + k * g.plane, ss);
                ldv<VEC>(mu + pidx + k * g.plane, mm);
                ldv<VEC>(w + pidx + k * g.plane, ww);
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    const float s = fmaxf(ss[j], scale_bound);
                    const float diff = yv[j] - mm[j];
                    const float v = fabsf(diff);
                    const float u = (0.5f - v) / s, d = (-0.5f - v) / s;
                    const float pu = FGMM_RSQRT_2PI * expf(-0.5f * u * u);
                    const float pd = FGMM_RSQRT_2PI * expf(-0.5f * d * d);
                    const float sgn = diff > 0.0f ? 1.0f : (diff < 0.0f ? -1.0f : 0.0f);
                    lk[k][j] = std_cum(u) - std_cum(d);
                    dv[k][j] = (pd - pu) / s * sgn;           // d l_k / d y
                    ds[k][j] = (pd * d - pu * u) / s;         // d l_k / d s
                    wk[k][j] = ww[j];
                    sraw[k][j] = ss[j];
                    tot[j] = tot[j] + lk[k][j] * ww[j];
                }
            }
        }
        float gyv[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (lik_bound > 0.0f && !(tot[j] >= lik_bound || gv[j] < 0.0f)) gv[j] = 0.0f;
            gyv[j] = 0.0f;
        }
#pragma unroll
        for (int k = 0; k < KA; ++k) {
            if (k < K) {
                float a[VEC], bq[VEC], c[VEC];
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    const float gl = gv[j] * wk[k][j];
                    const float gyk = gl * dv[k][j];
                    gyv[j] += gyk;
                    float gs = gl * ds[k][j];
                    if (!(sraw[k][j] >= scale_bound || gs < 0.0f)) gs = 0.0f;
                    a[j] = gs; bq[j] = -gyk; c[j] = gv[j] * lk[k][j];
                }
                const int64_t o = pidx + k * g.plane;
                stv<VEC>(gsc + o, a);
                stv<VEC>(gmu + o, bq);
                stv<VEC>(gw + o, c);
            }
        }
        stv<VEC>(gy + e, gyv);
    }
}

inline int lk_grid(int64_t items) {
    int64_t b = (items + kLkThreads - 1) / kLkThreads;
    const int64_t cap = 148LL * 32;
    return (int)(b < 1 ? 1 : (b < cap ? b : cap));
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

cudaError_t launch_gmm_likelihood_fwd(const float* x, const float* noise, int quant, const float* sc,
                                      const float* mu, const float* w, int64_t B, int64_t plane, int K,
                                      float scale_bound, float lik_bound, float* y_out, float* lik_out,
                                      cudaStream_t st) {
    LkGeom g{plane, B * plane, K};
    if (g.total <= 0) return cudaSuccess;
    const bool vec = (plane % 4 == 0) && al16(x) && al16(sc) && al16(mu) && al16(w) && al16(y_out) &&
                     al16(lik_out) && (quant != 1 || al16(noise));
    if (vec)
        gmm_likelihood_fwd_kernel<4><<<lk_grid(g.total / 4), kLkThreads, 0, st>>>(x, noise, quant, sc, mu, w, g,
                                                                                scale_bound, lik_bound, y_out, lik_out);
    else
        gmm_likelihood_fwd_kernel<1><<<lk_grid(g.total), kLkThreads, 0, st>>>(x, noise, quant, sc, mu, w, g,
                                                                            scale_bound, lik_bound, y_out, lik_out);
    return cudaGetLastError();
}

cudaError_t launch_gmm_likelihood_bwd(const float* y, const float* sc, const float* mu, const float* w,
                                      const float* grad, int64_t B, int64_t plane, int K, float scale_bound,
                                      float lik_bound, float* gy, float* gsc, float* gmu, float* gw,
                                      cudaStream_t st) {
    LkGeom g{plane, B * plane, K};
    if (g.total <= 0) return cudaSuccess;
    const bool vec = (plane % 4 == 0) && al16(y) && al16(sc) && al16(mu) && al16(w) && al16(grad) && al16(gy) &&
                     al16(gsc) && al16(gmu) && al16(gw);
    // two elements per thread: the per-element state of the second pass (5 x K values) is what limits occupancy --
    // 116 registers and 23 % warps active with four elements (ncu r02), FLASHGMM_LK_BWD_VEC=4 brings that form back
    static const int bwd_vec = [] { const char* v = getenv("FLASHGMM_LK_BWD_VEC"); return v && *v == '4' ? 4 : 2; }();
    if (vec && K == 4 && bwd_vec == 4)
        gmm_likelihood_bwd_kernel<4, 4><<<lk_grid(g.total / 4), kLkThreads, 0, st>>>(y, sc, mu, w, grad, g, scale_bound,
                                                                                   lik_bound, gy, gsc, gmu, gw);
    else if (vec && K == 4)
        gmm_likelihood_bwd_kernel<2, 4><<<lk_grid(g.total / 2), kLkThreads, 0, st>>>(y, sc, mu, w, grad, g, scale_bound,
                                                                                   lik_bound, gy, gsc, gmu, gw);
    else if (K == 4)
        gmm_likelihood_bwd_kernel<1, 4><<<lk_grid(g.total), kLkThreads, 0, st>>>(y, sc, mu, w, grad, g, scale_bound,
                                                                               lik_bound, gy, gsc, gmu, gw);
    else
        gmm_likelihood_bwd_kernel<1, 0><<<lk_grid(g.total), kLkThreads, 0, st>>>(y, sc, mu, w, grad, g, scale_bound,
                                                                               lik_bound, gy, gsc, gmu, gw);
    return cudaGetLastError();
}

}  // namespace flashgmm
