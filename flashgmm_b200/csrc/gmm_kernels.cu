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
// Row layout (uint16 elements), R = search radius, W = 2R+2:
//   [0, W)            T[j] = q16(cdf(j - R - 0.5))
//   [W]               1 if T is not non-decreasing (host then replays the reference's search)
//   (W, lut_off)      zero padding, lut_off = round_up(W+1, 8)
//   [lut_off, +32)    jump table: lut[b] = max{ j in [0,2R] : T[j] <= b*2048 } (0 if none), so the
//                     host starts its scan for cum at lut[cum >> 11] instead of at -R.
constexpr int kTabThreads = 128;   // symbols per block
constexpr int kTabTile = 64;       // uint16 entries per tile (32 words)
constexpr int kTabTileWords = kTabTile / 2;
constexpr int kLutBuckets = 32;
constexpr int kLutShift = 11;

template <int MODE>
__global__ void __launch_bounds__(kTabThreads)
gmm_cdf_table_kernel(GmmParamsView p, int64_t n, int R, uint16_t* __restrict__ table, int row_stride) {
    __shared__ uint32_t tile[kTabThreads][kTabTileWords + 1];
    __shared__ uint32_t lut_s[kTabThreads][kLutBuckets / 2 + 1];
    const int64_t row0 = (int64_t)blockIdx.x * kTabThreads;
    const int64_t i = row0 + threadIdx.x;
    const bool live = i < n;
    float mu[4] = {0, 0, 0, 0}, sg[4] = {1, 1, 1, 1}, w[4] = {0, 0, 0, 0};
    if (live) load_generic(p, i, mu, sg, w);

    const int W = 2 * R + 2;
    const int lut_off = (W + 1 + 7) & ~7;
    uint32_t prev = 0, nonmono = 0;
    int cn = 0;                                   // next jump-table bucket to fill
    uint16_t* my_lut = reinterpret_cast<uint16_t*>(lut_s[threadIdx.x]);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* tab32 = reinterpret_cast<uint32_t*>(table);
    const int row_words = row_stride >> 1;

    for (int j0 = 0; j0 < lut_off; j0 += kTabTile) {
        if (live) {
#pragma unroll 2
            for (int jj = 0; jj < kTabTile; jj += 2) {
                uint32_t e[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    int j = j0 + jj + h;
                    uint32_t c = 0;
                    if (j < W) {
                        // v - 0.5 with v = j - R; identical to (v-1) - 0.5 + 1.0 for |v| < 2^22
                        c = gmm_q16<MODE>(__fsub_rn((float)(j - R), 0.5f), mu, sg, w);
                        nonmono |= (c < prev);
                        prev = c;
                        if (j < W - 1) {          // candidates are j in [0, 2R]
                            while (cn < kLutBuckets && (uint32_t)(cn << kLutShift) < c) {
                                my_lut[cn++] = (uint16_t)(j > 0 ? j - 1 : 0);
                            }
                        }
                    } else if (j == W) {
                        c = nonmono;
                    }
                    e[h] = c;
                }
                tile[threadIdx.x][jj >> 1] = e[0] | (e[1] << 16);
            }
        }
        __syncthreads();
        // coalesced write-out: one warp per row segment (<=128 B)
        const int wbase = j0 >> 1;
        for (int r = warp; r < kTabThreads; r += kTabThreads / 32) {
            int64_t row = row0 + r;
            if (row < n && wbase + lane < (lut_off >> 1))
                tab32[row * row_words + wbase + lane] = tile[r][lane];
        }
        __syncthreads();
    }
    if (live) while (cn < kLutBuckets) my_lut[cn++] = (uint16_t)(2 * R);
    __syncthreads();
    // jump tables: 16 words per row, two rows per warp pass
    for (int r = warp * 2 + (lane >> 4); r < kTabThreads; r += (kTabThreads / 32) * 2) {
        int64_t row = row0 + r;
        if (row < n) tab32[row * row_words + (lut_off >> 1) + (lane & 15)] = lut_s[r][lane & 15];
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
    switch (gmm_layout_of(p, sym, out)) {
    case 0:
        gmm_startfreq_kmajor4_kernel<MODE><<<grid_for((n + 3) / 4, kSfThreads), kSfThreads, 0, st>>>(sym, p, n, out, bypass_flag);
        break;
    case 1:
        gmm_startfreq_scalar_kernel<MODE, true><<<grid_for(n, kSfThreads), kSfThreads, 0, st>>>(sym, p, n, out, bypass_flag);
        break;
    default:
        gmm_startfreq_scalar_kernel<MODE, false><<<grid_for(n, kSfThreads), kSfThreads, 0, st>>>(sym, p, n, out, bypass_flag);
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

int gmm_table_lut_offset(int R) { return ((2 * R + 3) + 7) & ~7; }
int gmm_table_row_stride(int R) { return gmm_table_lut_offset(R) + kLutBuckets; }

cudaError_t launch_gmm_cdf_table(const GmmParamsView& p, int64_t n, int mode, int R, uint16_t* table,
                                 int row_stride, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    int64_t blocks = (n + kTabThreads - 1) / kTabThreads;
    switch (mode) {
    case MODE_AS:
        gmm_cdf_table_kernel<MODE_AS><<<(unsigned)blocks, kTabThreads, 0, st>>>(p, n, R, table, row_stride);
        break;
    case MODE_LOGISTIC:
        gmm_cdf_table_kernel<MODE_LOGISTIC><<<(unsigned)blocks, kTabThreads, 0, st>>>(p, n, R, table, row_stride);
        break;
    default:
        gmm_cdf_table_kernel<MODE_POLYA><<<(unsigned)blocks, kTabThreads, 0, st>>>(p, n, R, table, row_stride);
    }
    return cudaGetLastError();
}

}  // namespace flashgmm
