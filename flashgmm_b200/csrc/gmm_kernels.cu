// gmm_kernels.cu -- sm_100a kernels for the reference-compatible (single-stream) path.
//
//  gmm_startfreq_kernel : per symbol, the two quantised mixture-CDF values the
//      encoder needs -> packed (start | freq<<16), bypass marked by freq==0.
//      Replaces the per-symbol body of BufferedRansEncoder::encode_with_indexes_gmm<4>
//      (/root/reference/compressai/cpp_exts/rans/rans_interface.cpp:486-522).
//  gmm_cdf_table_kernel : per symbol, the quantised CDF at every integer boundary
//      v-0.5, v in [-R, R+1] -> uint16 row the host decoder searches. Replaces the
//      CDF evaluations inside RansDecoder::decode_with_indexes_gmm<4>'s binary search
//      (rans_interface.cpp:833-862); the search's control flow is replayed on the row.
//
// Not GEMM-shaped: no tensor cores. HBM-streaming + FP32 pipe. Layout is the
// reference's own K-major [K][N] planes (entropy_models.py:813-827 hands over
// (N,K) views with strides (1,N)), read with 128-bit loads, 4 symbols per thread.
#include "gmm_kernels.h"
#include "gmm_math.cuh"

namespace flashgmm {

namespace {

constexpr int kSfThreads = 256;

// ---- parameter fetch for one symbol under the three supported layouts ----
__device__ __forceinline__ void load_generic(const GmmParamsView& p, int64_t i, float (&mu)[4],
                                             float (&sg)[4], float (&w)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int64_t o = i * p.stride_n + k * p.stride_k;
        sg[k] = __ldg(p.scales + o);
        mu[k] = __ldg(p.means + o);
        w[k] = __ldg(p.weights + o);
    }
}

__device__ __forceinline__ void load_aos(const GmmParamsView& p, int64_t i, float (&mu)[4],
                                         float (&sg)[4], float (&w)[4]) {
    float4 a = __ldcs(reinterpret_cast<const float4*>(p.scales) + i);
    float4 b = __ldcs(reinterpret_cast<const float4*>(p.means) + i);
    float4 c = __ldcs(reinterpret_cast<const float4*>(p.weights) + i);
    sg[0] = a.x; sg[1] = a.y; sg[2] = a.z; sg[3] = a.w;
    mu[0] = b.x; mu[1] = b.y; mu[2] = b.z; mu[3] = b.w;
    w[0] = c.x; w[1] = c.y; w[2] = c.z; w[3] = c.w;
}

// K-major planes, 4 consecutive symbols: 12 x 128-bit loads.
struct Quad {
    float sg[4][4], mu[4][4], w[4][4];   // [symbol][component]
};
__device__ __forceinline__ void load_kmajor4(const GmmParamsView& p, int64_t i4, Quad& q) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float4 a = __ldcs(reinterpret_cast<const float4*>(p.scales + k * p.stride_k) + i4);
        float4 b = __ldcs(reinterpret_cast<const float4*>(p.means + k * p.stride_k) + i4);
        float4 c = __ldcs(reinterpret_cast<const float4*>(p.weights + k * p.stride_k) + i4);
        q.sg[0][k] = a.x; q.sg[1][k] = a.y; q.sg[2][k] = a.z; q.sg[3][k] = a.w;
        q.mu[0][k] = b.x; q.mu[1][k] = b.y; q.mu[2][k] = b.z; q.mu[3][k] = b.w;
        q.w[0][k] = c.x; q.w[1][k] = c.y; q.w[2][k] = c.z; q.w[3][k] = c.w;
    }
}

// Tell the host that this chunk holds at least one bypass symbol (it then needs the
// symbol values on its side). One atomic per launch in the common case: later
// hits see the flag already set and only read it.
__device__ __forceinline__ void note_bypass(uint32_t* flag, bool hit) {
    if (flag == nullptr) return;
    if (hit && *reinterpret_cast<volatile uint32_t*>(flag) == 0u) atomicOr(flag, 1u);
}

// ---- encode side ----
template <int MODE>
__global__ void __launch_bounds__(kSfThreads)
gmm_startfreq_kmajor4_kernel(const int32_t* __restrict__ sym, GmmParamsView p, int64_t n,
                             uint32_t* __restrict__ out, uint32_t* bypass_flag) {
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i4 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i4 < n4; i4 += stride) {
        Quad q;
        load_kmajor4(p, i4, q);
        int4 v = __ldcs(reinterpret_cast<const int4*>(sym) + i4);
        uint4 r;
        r.x = gmm_startfreq1<MODE>(v.x, q.mu[0], q.sg[0], q.w[0]);
        r.y = gmm_startfreq1<MODE>(v.y, q.mu[1], q.sg[1], q.w[1]);
        r.z = gmm_startfreq1<MODE>(v.z, q.mu[2], q.sg[2], q.w[2]);
        r.w = gmm_startfreq1<MODE>(v.w, q.mu[3], q.sg[3], q.w[3]);
        __stcs(reinterpret_cast<uint4*>(out) + i4, r);
        note_bypass(bypass_flag, r.x == FGMM_PACK_BYPASS || r.y == FGMM_PACK_BYPASS ||
                                     r.z == FGMM_PACK_BYPASS || r.w == FGMM_PACK_BYPASS);
    }
    // ragged tail (n % 4 symbols): first threads of block 0
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        int64_t i = (n4 << 2) + threadIdx.x;
        float mu[4], sg[4], w[4];
        load_generic(p, i, mu, sg, w);
        uint32_t r = gmm_startfreq1<MODE>(sym[i], mu, sg, w);
        out[i] = r;
        if (bypass_flag && r == FGMM_PACK_BYPASS) atomicOr(bypass_flag, 1u);
    }
}

template <int MODE, bool AOS>
__global__ void __launch_bounds__(kSfThreads)
gmm_startfreq_scalar_kernel(const int32_t* __restrict__ sym, GmmParamsView p, int64_t n,
                            uint32_t* __restrict__ out, uint32_t* bypass_flag) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float mu[4], sg[4], w[4];
        if (AOS) load_aos(p, i, mu, sg, w); else load_generic(p, i, mu, sg, w);
        uint32_t r = gmm_startfreq1<MODE>(__ldcs(sym + i), mu, sg, w);
        out[i] = r;
        note_bypass(bypass_flag, r == FGMM_PACK_BYPASS);
    }
}

// ---- decode side: quantised CDF rows ----
// Row layout (uint16 elements), R = search radius, W = 2R+2 entries:
//   [0, W)            T[j] = q16(cdf(j - R - 0.5))
//   [W]               1 if T is not non-decreasing (host then replays the reference's search)
//   (W, lut_off)      zero padding, lut_off = round_up(W+1, 8)
//   [lut_off, +32)    only when the host asks for it (long rows): jump table,
//                     lut[b] = max{ j in [0,2R] : T[j] <= b*2048 } (0 if none).
//
// Work decomposition: one thread per PAIR of adjacent entries (one 32-bit word of
// the row), so a stream of n symbols exposes n*(R+1) independent work items -- small
// streams (12 288 symbols in ELIC's first groups) still fill 148 SMs -- and the word
// index maps linearly onto the global address: stores are coalesced by construction.
// A CTA owns S consecutive symbols; their 12 parameters are staged once in shared
// memory (coalesced plane reads), the row image is assembled there so that the
// monotonicity flag and the jump table can be derived, then streamed out.
constexpr int kRowThreads = 128;
constexpr int kLutBuckets = 32;
constexpr int kLutShift = 11;

template <int MODE>
__device__ __forceinline__ void
gmm_cdf_rows_block(const GmmParamsView& p, int64_t n, int R, uint32_t* __restrict__ table32, int row_words,
                   int lut_off, int S, int64_t block, uint32_t* smem_u32, int lane_slots = 0) {
    float* prm = reinterpret_cast<float*>(smem_u32);       // [12][S]: sg[4], mu[4], w[4]
    uint32_t* tile = smem_u32 + 12 * S;                    // [S][row_words]
    uint32_t* bad = tile + (size_t)S * row_words;          // [S]

    const int64_t row0 = block * S;
    const int live = (int)min((int64_t)S, n - row0);
    const int W = 2 * R + 2;

    for (int idx = threadIdx.x; idx < 12 * S; idx += kRowThreads) {
        const int plane = idx / S, s = idx - plane * S;
        float v = plane < 4 ? 1.0f : 0.0f;
        if (s < live) {
            const int k = plane & 3;
            const float* base = plane < 4 ? p.scales : (plane < 8 ? p.means : p.weights);
            v = __ldg(base + (row0 + s) * p.stride_n + k * p.stride_k);
        }
        prm[idx] = v;
    }
    for (int s = threadIdx.x; s < S; s += kRowThreads) bad[s] = 0;
    __syncthreads();

    const int items = live * row_words;
    const int step_s = kRowThreads / row_words, step_w = kRowThreads - step_s * row_words;
    {
        int s = threadIdx.x / row_words, wd = threadIdx.x - s * row_words;
        for (int id = threadIdx.x; id < items; id += kRowThreads) {
            const int j = 2 * wd;
            uint32_t e0 = 0, e1 = 0;
            if (j < W) {                                   // W is even: j < W implies j + 1 < W
                float mu[4], sg[4], w[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    sg[k] = prm[k * S + s];
                    mu[k] = prm[(4 + k) * S + s];
                    w[k] = prm[(8 + k) * S + s];
                }
                // v - 0.5 with v = j - R; identical to (v-1) - 0.5 + 1.0 for |v| < 2^22
                e0 = gmm_q16<MODE>(__fsub_rn((float)(j - R), 0.5f), mu, sg, w);
                e1 = gmm_q16<MODE>(__fsub_rn((float)(j + 1 - R), 0.5f), mu, sg, w);
                if (e1 < e0) atomicOr(&bad[s], 1u);
            }
            tile[id] = e0 | (e1 << 16);
            s += step_s; wd += step_w;
            if (wd >= row_words) { wd -= row_words; ++s; }
        }
    }
    __syncthreads();
    {   // pairs that straddle two words
        int s = threadIdx.x / row_words, wd = threadIdx.x - s * row_words;
        for (int id = threadIdx.x; id < items; id += kRowThreads) {
            if (2 * wd + 2 < W && (tile[id + 1] & 0xffffu) < (tile[id] >> 16)) atomicOr(&bad[s], 1u);
            s += step_s; wd += step_w;
            if (wd >= row_words) { wd -= row_words; ++s; }
        }
    }
    __syncthreads();
    if (lut_off >= 0) {
        for (int idx = threadIdx.x; idx < live * kLutBuckets; idx += kRowThreads) {
            const int s = idx / kLutBuckets, b = idx - s * kLutBuckets;
            uint16_t* row16 = reinterpret_cast<uint16_t*>(tile + (size_t)s * row_words);
            const uint32_t key = (uint32_t)b << kLutShift;
            int lo = 0, hi = 2 * R + 1;                    // count of candidates j in [0,2R] with T[j] <= key
            while (lo < hi) {
                int mid = (lo + hi) >> 1;
                if (row16[mid] <= key) lo = mid + 1; else hi = mid;
            }
            row16[lut_off + b] = (uint16_t)(lo > 0 ? lo - 1 : 0);
        }
        __syncthreads();
    }
    if (lane_slots == kPopWindowed) {
        // Windowed record for the device pop kernel (gmm_pop.cu), any R: the 32 consecutive
        // intervals [j0, j0 + 32) that hold the most probability mass, one per lane. 68 words:
        //   word 0      j0 - R (the symbol value of lane 0's interval), words 1..3 zero
        //   word 4+2l   start T[j0+l]        word 5+2l   width T[j0+l+1] - T[j0+l]
        // Intervals past 2R, and every interval of a row that is not monotone, are (65535, 0): no
        // lane can hit. A symbol outside the window is found by the pop kernel's literal search.
        int* j0s = reinterpret_cast<int*>(bad + S);
        const int span = 2 * R + 1 - 32;                    // last admissible j0 (<= 0: the row fits)
        for (int sidx = threadIdx.x; sidx < live; sidx += kRowThreads) {
            const uint16_t* row16 = reinterpret_cast<const uint16_t*>(tile + (size_t)sidx * row_words);
            int best = 0, mass = -1;
            for (int j0 = 0; j0 <= span; ++j0) {
                const int m = (int)row16[j0 + 32] - (int)row16[j0];
                if (m > mass) { mass = m; best = j0; }
            }
            j0s[sidx] = best;
        }
        __syncthreads();
        uint32_t* dst = table32 + row0 * kPopWindowWords;
        for (int id = threadIdx.x; id < live * kPopWindowWords; id += kRowThreads) {
            const int s = id / kPopWindowWords, t = id - s * kPopWindowWords;
            const int j0 = j0s[s];
            uint32_t v = 0;
            if (t == 0) v = (uint32_t)(j0 - R);
            else if (t >= 4) {
                const int j = j0 + ((t - 4) >> 1);
                const uint16_t* row16 = reinterpret_cast<const uint16_t*>(tile + (size_t)s * row_words);
                const bool ok = j <= 2 * R && bad[s] == 0;
                const uint32_t lo = ok ? row16[j] : 65535u, hi = ok ? row16[j + 1] : 65535u;
                v = (t & 1) ? hi - lo : lo;
            }
            __stcs(dst + id, v);
        }
    } else if (lane_slots > 0) {
        // Lane-row record for the device pop kernel (gmm_pop.cu): per symbol `lane_slots` uint32
        // words, word l*NI + k (NI = lane_slots / 32) = start | width << 16 of interval
        // j = l + 32k, i.e. T[j] | (T[j+1] - T[j]) << 16 -- lane l reads its NI intervals with one
        // vector load. Slots past 2R, and every slot of a row that is not monotone, hold
        // (65535, 0): no lane can hit, the pop kernel then evaluates the reference's search itself.
        uint32_t* dst = table32 + row0 * lane_slots;
        const int ni = lane_slots >> 5;
        for (int id = threadIdx.x; id < live * lane_slots; id += kRowThreads) {
            const int s = id / lane_slots, t = id - s * lane_slots;
            const int l = t / ni, k = t - l * ni;
            const int j = l + 32 * k;
            const uint16_t* row16 = reinterpret_cast<const uint16_t*>(tile + (size_t)s * row_words);
            uint32_t v = 65535u;
            if (j <= 2 * R && bad[s] == 0) {
                const uint32_t lo = row16[j], hi = row16[j + 1];
                v = lo | ((hi - lo) << 16);
            }
            __stcs(dst + id, v);
        }
    } else {
        uint32_t* dst = table32 + row0 * row_words;
        int s = threadIdx.x / row_words, wd = threadIdx.x - s * row_words;
        const int flag_word = W >> 1;
        for (int id = threadIdx.x; id < items; id += kRowThreads) {
            uint32_t v = tile[id];
            if (wd == flag_word) v |= bad[s];
            __stcs(dst + id, v);
            s += step_s; wd += step_w;
            if (wd >= row_words) { wd -= row_words; ++s; }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kRowThreads)
gmm_cdf_rows_kernel(GmmParamsView p, int64_t n, int R, uint32_t* __restrict__ table32, int row_words,
                    int lut_off, int S) {
    extern __shared__ uint32_t smem_u32[];
    gmm_cdf_rows_block<MODE>(p, n, R, table32, row_words, lut_off, S, (int64_t)blockIdx.x, smem_u32);
}

// Every (stream, chunk) entry of a batch in ONE launch: blockIdx.x -> entry by binary
// search over the entries' first_block prefix, then the same block body.
template <int MODE>
__global__ void __launch_bounds__(kRowThreads)
gmm_cdf_rows_batched_kernel(const RowsDesc* __restrict__ descs, int n_desc, int S, int lane_slots) {
    extern __shared__ uint32_t smem_u32[];
    int lo = 0, hi = n_desc - 1;
    const int b = (int)blockIdx.x;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_block <= b) lo = mid; else hi = mid - 1;
    }
    const RowsDesc d = descs[lo];
    gmm_cdf_rows_block<MODE>(d.p, (int64_t)d.n, d.R, d.table32, d.row_words, -1, S,
                             (int64_t)(b - d.first_block), smem_u32, lane_slots);
}

// ---- decode side, device pop path: the WINDOWED record computed directly (no full row) ----
// The record the pop kernel reads (gmm_pop.cu) holds 32 consecutive intervals of the symbol's quantised CDF.
// The first version evaluated the whole row (2R+2 points) and then picked the 32-interval window with the
// most mass: O(R) work per symbol. Here the window is chosen FROM THE PARAMETERS and only its 33 boundaries
// are evaluated, so the cost does not grow with the search radius.
//
// Why any window is exact: for sane parameters (finite, 0 < sigma <= 65536, w >= 0, sum of weights below the
// 16-bit wrap) every row is non-decreasing -- the three Phi approximations are monotone over every finite
// binary32 argument (oracle/check_monotone.c, exhaustive) and subtraction, division by sigma > 0,
// multiplication by w >= 0, the pairwise sums and the truncation are monotone too -- so the reference's binary
// search (rans_interface.cpp:826-862) ends on THE interval that contains cum, and a lane of the window that
// contains cum is that interval. A cum outside the window is "no lane hits": the pop kernel then replays the
// reference's search literally (pop_slow). Symbols whose parameters are not sane get a blank record (every
// symbol of theirs takes the literal path). The window choice is therefore only a matter of speed; it uses
// plain fast arithmetic: centre of [min_k mu_k - 3.9 sigma_k, max_k mu_k + 3.9 sigma_k] over the components
// that carry weight (miss rate 6e-5 on the synthetic mixture of BASELINE config 4).
// |x - mu| <= 2^21 + 1e30 and sigma >= 2^-10 keep every quotient finite; a sane symbol whose quotient would
// overflow does not exist for |mu| <= 1e30 (1e30 * 2^10 < FLT_MAX).
constexpr int kWinThreads = 96;          // 32 symbols x 3 threads; a thread evaluates 11 of its symbol's 33 boundaries
constexpr int kWinSyms = 32;
constexpr int kWinPts = 33;
constexpr int kWinPerThread = 11;

template <int MODE>
__global__ void __launch_bounds__(kWinThreads)
gmm_cdf_window_batched_kernel(const RowsDesc* __restrict__ descs, int n_desc) {
    __shared__ uint32_t T[kWinSyms][kWinPts];                 // stride 33: conflict-free both ways
    __shared__ int j0s[kWinSyms];                             // first boundary index of the window, -1 = blank record
    int lo = 0, hi = n_desc - 1;
    const int b = (int)blockIdx.x;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_block <= b) lo = mid; else hi = mid - 1;
    }
    const RowsDesc d = descs[lo];
    const int64_t row0 = (int64_t)(b - d.first_block) * kWinSyms;
    const int live = (int)min((int64_t)kWinSyms, (int64_t)d.n - row0);
    const int R = d.R;
    const int t = (int)threadIdx.x;
    const int sidx = t / 3, part = t - 3 * sidx;              // the symbol's 3 threads are neighbours: their loads coalesce

    // The symbol's 12 parameters stay in registers for all 11 boundaries, with the reciprocals of the scales: every
    // boundary divides by the same four sigmas (div_by_shared_rcp, bit-identical to __fdiv_rn in the range checked below).
    // Loaded plane by plane, 32 consecutive symbols per request (page-locked host tensors are read in place over
    // PCIe: 128-byte requests, each parameter exactly once), handed to the symbol's threads through shared memory.
    float mu[4], sg[4], w[4], rcp[4];
    bool sane = sidx < live;
    {
        float* prm = reinterpret_cast<float*>(&T[0][0]);      // 12 x 32 floats, before T is written
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = t + kWinThreads * i;              // 384 = 12 planes x 32 symbols
            const int plane = idx >> 5, s2 = idx & 31;
            float v = plane < 4 ? 1.0f : 0.0f;
            if (s2 < live) {
                const float* base = plane < 4 ? d.p.scales : (plane < 8 ? d.p.means : d.p.weights);
                v = __ldg(base + (row0 + s2) * d.p.stride_n + (plane & 3) * d.p.stride_k);
            }
            prm[idx] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 4; ++k) { sg[k] = prm[32 * k + sidx]; mu[k] = prm[32 * (4 + k) + sidx]; w[k] = prm[32 * (8 + k) + sidx]; }
        __syncthreads();                                      // T is written next
    }
    float lo_x = 3.0e38f, hi_x = -3.0e38f, lo_all = 3.0e38f, hi_all = -3.0e38f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        // sane: finite, scale in the range the shared-reciprocal division is checked for, weight a probability,
        // |x - mu| small enough that neither quotient nor residual leaves the normal range
        sane = sane && (sg[k] >= 0.0009765625f) && (sg[k] <= 65536.0f) && (w[k] >= 0.0f) && (w[k] <= 2.0f) &&
               (fabsf(mu[k]) <= 1.0e30f);
        rcp[k] = __frcp_rn(sg[k]);
        const float a = mu[k] - 3.9f * sg[k], c = mu[k] + 3.9f * sg[k];
        lo_all = fminf(lo_all, a); hi_all = fmaxf(hi_all, c);
        if (w[k] > 1.0e-3f) { lo_x = fminf(lo_x, a); hi_x = fmaxf(hi_x, c); }
    }
    // the largest value any boundary can take is the sum of the weights (same association as the CDF): it must
    // stay below the 16-bit wrap of the quantiser
    sane = sane && (__fmul_rn(__fadd_rn(__fadd_rn(w[0], w[1]), __fadd_rn(w[2], w[3])), 65535.0f) < 65536.0f);
    if (hi_x < lo_x) { lo_x = lo_all; hi_x = hi_all; }
    int j0 = -1;
    if (sane) {
        const float centre = fminf(fmaxf(0.5f * (lo_x + hi_x), -3.0e6f), 3.0e6f);
        j0 = min(max(__float2int_rn(centre) - 16 + R, 0), max(0, 2 * R + 1 - 32));
    }
    if (part == 0) j0s[sidx] = j0;
    // Saturated boundaries are not evaluated. Each Phi approximation reaches EXACTLY 1.0f at t >= kSatHi and (modes 0, 1)
    // exactly +0.0f at t <= -kSatLo (thresholds found by bisection over the bit patterns with the CPU restatement; Phi is
    // monotone over every finite argument, oracle/check_monotone.c, so it stays there). Beyond x_hi every component is at
    // 1.0: the boundary is q16((w0 + w1) + (w2 + w3)); below x_lo every component is at +0: the boundary is 0 -- the very
    // values the full evaluation produces. The 0.05 on the thresholds and the slack of half a symbol plus 2^-18 relative
    // cover the rounding of this float arithmetic. Real latents have small sigmas: most of a 33-boundary window is
    // saturated (the window is centred on the active span, so the same iterations are skipped by all lanes of a warp).
    constexpr float kSatHi = (MODE == MODE_POLYA ? 4.9690776f : MODE == MODE_AS ? 5.4203959f : 9.7741079f) + 0.05f;
    constexpr float kSatLo = (MODE == MODE_POLYA ? 5.2172627f : 5.4203959f) + 0.05f;
    float x_lo = 3.0e38f, x_hi = -3.0e38f;
    bool skip_ok = true;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        skip_ok = skip_ok && fabsf(mu[k]) <= 2097152.0f;
        x_lo = fminf(x_lo, mu[k] - kSatLo * sg[k]);
        x_hi = fmaxf(x_hi, mu[k] + kSatHi * sg[k]);
    }
    x_lo = (skip_ok && MODE != MODE_LOGISTIC) ? x_lo - 0.5f - fabsf(x_lo) * 3.8147e-6f : -3.0e38f;
    x_hi = skip_ok ? x_hi + 0.5f + fabsf(x_hi) * 3.8147e-6f : 3.0e38f;
    const uint32_t top = q16(__fadd_rn(__fadd_rn(w[0], w[1]), __fadd_rn(w[2], w[3])));
#pragma unroll 1
    for (int i = 0; i < kWinPerThread; ++i) {
        const int l = part + 3 * i;                           // interleaved: the symbol's active boundaries are shared out evenly
        uint32_t e = 65535u;
        if (j0 >= 0 && j0 + l <= 2 * R + 1) {
            const float x = __fsub_rn((float)(j0 + l - R), 0.5f);
            if (x >= x_hi) e = top;
            else if (x <= x_lo) e = 0u;
            else e = gmm_q16_rcp<MODE>(x, mu, sg, rcp, w);
        }
        T[sidx][l] = e;
    }
    __syncthreads();
    uint32_t* dst = d.table32 + row0 * kPopWindowWords;
    for (int id = t; id < live * kPopWindowWords; id += kWinThreads) {
        const int s2 = id / kPopWindowWords, q = id - s2 * kPopWindowWords;
        const int jj = j0s[s2];
        uint32_t v = 0;
        if (q == 0) v = (uint32_t)((jj < 0 ? 0 : jj) - R);
        else if (q >= 4) {
            const int l = (q - 4) >> 1;
            const bool ok = jj >= 0 && jj + l <= 2 * R;       // interval jj + l exists in the row
            const uint32_t a = ok ? T[s2][l] : 65535u, c = ok ? T[s2][l + 1] : 65535u;
            v = (q & 1) ? c - a : a;
        }
        __stcs(dst + id, v);
    }
}

inline int grid_for(int64_t work, int threads) {
    int64_t b = (work + threads - 1) / threads;
    const int64_t cap = 148LL * 64;   // grid-stride beyond this
    if (b < 1) b = 1;
    return (int)(b < cap ? b : cap);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

int gmm_layout_of(const GmmParamsView& p, const void* sym, const void* out) {
    bool a = aligned16(p.scales) && aligned16(p.means) && aligned16(p.weights);
    if (p.stride_n == 1 && (p.stride_k % 4) == 0 && a && aligned16(sym) && aligned16(out)) return 0;
    if (p.stride_k == 1 && p.stride_n == 4 && a) return 1;
    return 2;
}

template <int MODE>
static cudaError_t launch_startfreq_mode(const int32_t* sym, const GmmParamsView& p, int64_t n,
                                         uint32_t* out, uint32_t* bypass_flag, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    int layout = gmm_layout_of(p, sym, out);
    // Small streams: 4 symbols per thread would leave most SMs idle (12 288 symbols = 12 CTAs);
    // one symbol per thread in 128-thread CTAs exposes 4x the parallelism and the
    // 32-bit plane loads are still coalesced.
    if (layout == 0 && n < 148LL * 2 * kSfThreads * 4) layout = 2;
    switch (layout) {
    case 0:
        gmm_startfreq_kmajor4_kernel<MODE><<<grid_for((n + 3) / 4, kSfThreads), kSfThreads, 0, st>>>(sym, p, n, out, bypass_flag);
        break;
    case 1:
        gmm_startfreq_scalar_kernel<MODE, true><<<grid_for(n, 128), 128, 0, st>>>(sym, p, n, out, bypass_flag);
        break;
    default:
        gmm_startfreq_scalar_kernel<MODE, false><<<grid_for(n, 128), 128, 0, st>>>(sym, p, n, out, bypass_flag);
    }
    return cudaGetLastError();
}

cudaError_t launch_gmm_startfreq(const int32_t* sym, const GmmParamsView& p, int64_t n, int mode,
                                 uint32_t* out, uint32_t* bypass_flag, cudaStream_t st) {
    switch (mode) {
    case MODE_AS: return launch_startfreq_mode<MODE_AS>(sym, p, n, out, bypass_flag, st);
    case MODE_LOGISTIC: return launch_startfreq_mode<MODE_LOGISTIC>(sym, p, n, out, bypass_flag, st);
    default: return launch_startfreq_mode<MODE_POLYA>(sym, p, n, out, bypass_flag, st);
    }
}

// Every stream of a batch in one launch (one symbol per thread, any layout).
template <int MODE>
__global__ void __launch_bounds__(128)
gmm_startfreq_batched_kernel(const SfDesc* __restrict__ descs, int n_desc) {
    int lo = 0, hi = n_desc - 1;
    const int b = (int)blockIdx.x;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_block <= b) lo = mid; else hi = mid - 1;
    }
    const SfDesc d = descs[lo];
    const int i = (b - d.first_block) * 128 + (int)threadIdx.x;
    if (i >= d.n) return;
    float mu[4], sg[4], w[4];
    load_generic(d.p, i, mu, sg, w);
    const int32_t v = __ldg(d.sym + i);
    const uint32_t r = gmm_startfreq1<MODE>(v, mu, sg, w);
    d.packed[i] = r;
    if (r == FGMM_PACK_BYPASS && d.flag != nullptr) {        // rare: hand the pusher the value it will need
        const uint32_t j = atomicAdd(d.flag, 1u);
        if (j < (uint32_t)kSfBypassSlots) { d.flag[1 + 2 * j] = (uint32_t)i; d.flag[2 + 2 * j] = (uint32_t)v; }
    }
}

namespace {
__global__ void gmm_sf_clear_flags_kernel(const SfDesc* __restrict__ descs, int n_desc) {
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i < n_desc && descs[i].flag != nullptr) descs[i].flag[threadIdx.x & 31] = 0u;
}
}  // namespace

// zeroes the 32-word bypass tail of every descriptor that has one (before the launches that fill them)
cudaError_t launch_gmm_sf_clear_flags(const SfDesc* descs_d, int n_desc, cudaStream_t st) {
    if (n_desc <= 0) return cudaSuccess;
    gmm_sf_clear_flags_kernel<<<(n_desc + 7) / 8, 256, 0, st>>>(descs_d, n_desc);
    return cudaGetLastError();
}

cudaError_t launch_gmm_startfreq_batched(const SfDesc* descs_d, int n_desc, int total_blocks, int mode, cudaStream_t st) {
    if (n_desc <= 0 || total_blocks <= 0) return cudaSuccess;
    switch (mode) {
    case MODE_AS: gmm_startfreq_batched_kernel<MODE_AS><<<total_blocks, 128, 0, st>>>(descs_d, n_desc); break;
    case MODE_LOGISTIC: gmm_startfreq_batched_kernel<MODE_LOGISTIC><<<total_blocks, 128, 0, st>>>(descs_d, n_desc); break;
    default: gmm_startfreq_batched_kernel<MODE_POLYA><<<total_blocks, 128, 0, st>>>(descs_d, n_desc);
    }
    return cudaGetLastError();
}

int gmm_table_lut_offset(int R) { return ((2 * R + 3) + 7) & ~7; }

namespace {
constexpr size_t kRowSmemBudget = 96 * 1024;

template <int MODE>
cudaError_t launch_rows_mode(const GmmParamsView& p, int64_t n, int R, uint16_t* table, int row_stride,
                             bool with_lut, cudaStream_t st) {
    const int row_words = row_stride >> 1;
    const size_t per_sym = (size_t)(12 + row_words + 1) * 4;
    int64_t S = n / (148 * 4);                     // enough CTAs to cover the SMs a few times
    S = S < 8 ? 8 : (S > 64 ? 64 : S);
    const int64_t fit = (int64_t)(kRowSmemBudget / per_sym);
    if (S > fit) S = fit;
    if (S < 1) return cudaErrorInvalidValue;
    const int64_t blocks = (n + S - 1) / S;
    const size_t smem = (size_t)S * per_sym;
    if (smem > 48 * 1024) {   // long rows only (R > ~750): opt in to large dynamic shared memory on this device
        cudaError_t e = cudaFuncSetAttribute(gmm_cdf_rows_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kRowSmemBudget);
        if (e != cudaSuccess) return e;
    }
    gmm_cdf_rows_kernel<MODE><<<(unsigned)blocks, kRowThreads, smem, st>>>(
        p, n, R, reinterpret_cast<uint32_t*>(table), row_words, with_lut ? gmm_table_lut_offset(R) : -1, (int)S);
    return cudaGetLastError();
}
}  // namespace

cudaError_t launch_gmm_cdf_rows_batched(const RowsDesc* descs_d, int n_desc, int total_blocks, int S,
                                        int max_row_words, int lane_slots, int mode, cudaStream_t st) {
    if (n_desc <= 0 || total_blocks <= 0) return cudaSuccess;
    if (lane_slots == kPopWindowed) {
        if (S != kRowsWindowSyms) return cudaErrorInvalidValue;      // first_block prefixes are in units of 32 symbols
        switch (mode) {
        case MODE_AS: gmm_cdf_window_batched_kernel<MODE_AS><<<total_blocks, kWinThreads, 0, st>>>(descs_d, n_desc); break;
        case MODE_LOGISTIC: gmm_cdf_window_batched_kernel<MODE_LOGISTIC><<<total_blocks, kWinThreads, 0, st>>>(descs_d, n_desc); break;
        default: gmm_cdf_window_batched_kernel<MODE_POLYA><<<total_blocks, kWinThreads, 0, st>>>(descs_d, n_desc);
        }
        return cudaGetLastError();
    }
    const size_t smem = (size_t)S * (12 + max_row_words + 2) * 4;    // + bad flag + window base per symbol
    if (smem > 48 * 1024) return cudaErrorInvalidValue;      // callers size S for the widest row
    switch (mode) {
    case MODE_AS: gmm_cdf_rows_batched_kernel<MODE_AS><<<total_blocks, kRowThreads, smem, st>>>(descs_d, n_desc, S, lane_slots); break;
    case MODE_LOGISTIC: gmm_cdf_rows_batched_kernel<MODE_LOGISTIC><<<total_blocks, kRowThreads, smem, st>>>(descs_d, n_desc, S, lane_slots); break;
    default: gmm_cdf_rows_batched_kernel<MODE_POLYA><<<total_blocks, kRowThreads, smem, st>>>(descs_d, n_desc, S, lane_slots);
    }
    return cudaGetLastError();
}

cudaError_t launch_gmm_cdf_table(const GmmParamsView& p, int64_t n, int mode, int R, uint16_t* table,
                                 int row_stride, bool with_lut, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    switch (mode) {
    case MODE_AS: return launch_rows_mode<MODE_AS>(p, n, R, table, row_stride, with_lut, st);
    case MODE_LOGISTIC: return launch_rows_mode<MODE_LOGISTIC>(p, n, R, table, row_stride, with_lut, st);
    default: return launch_rows_mode<MODE_POLYA>(p, n, R, table, row_stride, with_lut, st);
    }
}

}  // namespace flashgmm
