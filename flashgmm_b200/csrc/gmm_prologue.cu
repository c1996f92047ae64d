// gmm_prologue.cu -- the glue of GaussianMixtureConditional.compress / decompress ON the GPU
// (sm_100a): what the reference does with a dozen tensor ops, three host syncs and four
// device->host copies per call (/root/reference/compressai/entropy_models/entropy_models.py):
//   compress  :833-851  abs_max from max / min of y, y_quantized = round(y), per-channel
//                       zero bitmap (sum |y_quantized| over H, W == 0), kept channels compacted
//                       into the symbol vector, parameters reshaped to (Nsym, K) K-major views
//                       with the scale clamp (reshape_entropy_parameters, :810-828)
//   decompress:902-909  decoded symbols scattered back into zeros (1, C, H, W)
// Three kernels, all plain streaming (HBM-bound, coalesced, no reuse -> no shared-memory
// staging, no tensor cores):
//   scan    one CTA per channel: round, "any nonzero", running max / min of the unrounded y
//   gather  one CTA per (plane, channel): kept channels' rows copied to their compacted place
//           (exclusive prefix of the bitmap recomputed per CTA: C is a few hundred)
//   scatter one CTA per channel: symbols back to float rows, zero rows for dropped channels
// Layouts are the reference's: y (1, C, H, W) contiguous; parameters (1, K*C, H, W) contiguous,
// component k of channel c at channel k*C + c; outputs [K][Nkept*HW] planes, i.e. the (Nsym, K)
// views with strides (1, Nsym) the coder reads.
#include "gmm_kernels.h"

#include <mutex>

namespace flashgmm {

namespace {

constexpr int kPrThreads = 256;

// order-preserving map float -> uint32 (so that atomicMax / atomicMin work on floats of either sign)
__device__ __forceinline__ uint32_t ordered_bits(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void __launch_bounds__(kPrThreads)
gmm_prologue_scan_kernel(const float* __restrict__ y, int C, int64_t HW, float* __restrict__ yq,
                         long long* __restrict__ bitmap64, int32_t* __restrict__ bitmap32,
                         uint32_t* __restrict__ extrema) {
    __shared__ float s_mx[kPrThreads / 32], s_mn[kPrThreads / 32];
    __shared__ int s_any[kPrThreads / 32];
    const int64_t base = (int64_t)blockIdx.x * HW;
    float mx = -INFINITY, mn = INFINITY;
    int any = 0;
    for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) {
        const float v = __ldg(y + base + i);
        const float r = rintf(v);                          // torch.round: half to even
        yq[base + i] = r;
        any |= (r != 0.0f) ? 1 : 0;                        // sum |r| == 0  <=>  every r == +-0 (NaN counts as nonzero, as there)
        mx = fmaxf(mx, v);
        mn = fminf(mn, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        any |= __shfl_xor_sync(0xffffffffu, any, o);
    }
    const int wid = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_mx[wid] = mx; s_mn[wid] = mn; s_any[wid] = any; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kPrThreads / 32; ++w) { mx = fmaxf(mx, s_mx[w]); mn = fminf(mn, s_mn[w]); any |= s_any[w]; }
        if (bitmap64) bitmap64[blockIdx.x] = any;
        bitmap32[blockIdx.x] = any;
        if (HW > 0) {
            const int b = (int)blockIdx.x / C;                 // image of this channel: extrema are per image
            atomicMax(extrema + 2 * b, ordered_bits(mx));
            atomicMin(extrema + 2 * b + 1, ordered_bits(mn));
        }
    }
}

// exclusive prefix of the 0/1 bitmap at channel c, by the whole CTA
__device__ __forceinline__ int kept_before(const long long* __restrict__ bitmap, int c, int* s_part) {
    int cnt = 0;
    for (int i = threadIdx.x; i < c; i += kPrThreads) cnt += bitmap[i] != 0 ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = cnt;
    __syncthreads();
    int tot = 0;
    for (int w = 0; w < kPrThreads / 32; ++w) tot += s_part[w];
    return tot;
}

// grid (C, 3K + 1, images): plane p < K scales (clamped), < 2K means, < 3K weights, == 3K symbols.
// Image b's kept symbols start at symbol offset meta.off[b] of the outputs and its K-major block
// [K][meta.n[b]] at K * meta.off[b] of each parameter output.
__global__ void __launch_bounds__(kPrThreads)
gmm_prologue_gather_kernel(const float* __restrict__ yq, const float* __restrict__ scales,
                           const float* __restrict__ means, const float* __restrict__ weights,
                           const long long* __restrict__ bitmap, int C, int64_t HW, int K, PrologueBatch meta,
                           int32_t* __restrict__ sym_out, float* __restrict__ scales_out,
                           float* __restrict__ means_out, float* __restrict__ weights_out) {
    __shared__ int s_part[kPrThreads / 32];
    const int c = blockIdx.x, p = blockIdx.y, b = blockIdx.z;
    const int64_t n_sym = meta.n[b], off = meta.off[b];
    if (n_sym <= 0) return;
    bitmap += (int64_t)b * C;
    if (yq) yq += (int64_t)b * C * HW;
    scales += (int64_t)b * K * C * HW; means += (int64_t)b * K * C * HW; weights += (int64_t)b * K * C * HW;
    if (sym_out) sym_out += off;
    scales_out += (int64_t)K * off; means_out += (int64_t)K * off; weights_out += (int64_t)K * off;
    if (bitmap[c] == 0) return;
    if (p == 3 * K && (yq == nullptr || sym_out == nullptr)) return;
    const int64_t j = kept_before(bitmap, c, s_part);
    if (p == 3 * K) {
        const float* src = yq + (int64_t)c * HW;
        int32_t* dst = sym_out + j * HW;
        for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) dst[i] = (int32_t)__ldg(src + i);   // .int(): truncation
        return;
    }
    const int t = p / K, k = p - t * K;
    const float* src = (t == 0 ? scales : t == 1 ? means : weights) + ((int64_t)k * C + c) * HW;
    float* dst = (t == 0 ? scales_out : t == 1 ? means_out : weights_out) + (int64_t)k * n_sym + j * HW;
    if (t == 0) {
        for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) {
            const float v = __ldg(src + i);
            dst[i] = v < 0.11f ? 0.11f : (v > 256.0f ? 256.0f : v);      // torch.clamp(0.11, 256): NaN stays NaN
        }
    } else {
        for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) dst[i] = __ldg(src + i);
    }
}

__global__ void __launch_bounds__(kPrThreads)
gmm_epilogue_scatter_kernel(const int32_t* __restrict__ symbols, const long long* __restrict__ bitmap,
                            int C, int64_t HW, PrologueBatch meta, float* __restrict__ y_hat) {
    __shared__ int s_part[kPrThreads / 32];
    const int c = blockIdx.x, b = blockIdx.y;
    bitmap += (int64_t)b * C;
    if (symbols) symbols += meta.off[b];
    float* dst = y_hat + ((int64_t)b * C + c) * HW;
    if (bitmap[c] == 0) {
        for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) dst[i] = 0.0f;
        return;
    }
    const int64_t j = kept_before(bitmap, c, s_part);
    const int32_t* src = symbols + j * HW;
    for (int64_t i = threadIdx.x; i < HW; i += kPrThreads) dst[i] = (float)__ldg(src + i);
}

}  // namespace

cudaError_t launch_gmm_prologue_scan(const float* y, int B, int C, int64_t HW, float* yq, long long* bitmap64,
                                     int32_t* bitmap32, uint32_t* extrema, cudaStream_t st) {
    // per image: extrema[2b] collects the max (starts at the smallest code), extrema[2b+1] the min (the largest)
    if (B <= 0) return cudaSuccess;
    cudaError_t e = cudaMemset2DAsync(extrema, 8, 0x00, 4, (size_t)B, st);
    if (e != cudaSuccess) return e;
    e = cudaMemset2DAsync(extrema + 1, 8, 0xff, 4, (size_t)B, st);
    if (e != cudaSuccess) return e;
    if (C <= 0) return cudaSuccess;
    gmm_prologue_scan_kernel<<<(unsigned)(B * C), kPrThreads, 0, st>>>(y, C, HW, yq, bitmap64, bitmap32, extrema);
    return cudaGetLastError();
}

cudaError_t launch_gmm_prologue_gather(const float* yq, const float* scales, const float* means, const float* weights,
                                       const long long* bitmap, int B, int C, int64_t HW, int K, const PrologueBatch& meta,
                                       int32_t* sym_out, float* scales_out, float* means_out, float* weights_out,
                                       cudaStream_t st) {
    if (B <= 0 || C <= 0 || HW <= 0) return cudaSuccess;
    gmm_prologue_gather_kernel<<<dim3((unsigned)C, (unsigned)(3 * K + 1), (unsigned)B), kPrThreads, 0, st>>>(
        yq, scales, means, weights, bitmap, C, HW, K, meta, sym_out, scales_out, means_out, weights_out);
    return cudaGetLastError();
}

cudaError_t launch_gmm_epilogue_scatter(const int32_t* symbols, const long long* bitmap, int B, int C, int64_t HW,
                                        const PrologueBatch& meta, float* y_hat, cudaStream_t st) {
    if (B <= 0 || C <= 0 || HW <= 0) return cudaSuccess;
    gmm_epilogue_scatter_kernel<<<dim3((unsigned)C, (unsigned)B), kPrThreads, 0, st>>>(symbols, bitmap, C, HW, meta, y_hat);
    return cudaGetLastError();
}

// ---- the scan with its scratch in a file-scope array: nothing is allocated, ever ------------------
// (no counterpart in the reference; for callers that cannot call cudaMalloc). Layout of the scratch,
// as in the allocating path: [0, 2B) extrema codes, then B*C bitmap words. Every GPU that loads the
// library holds these kStaticScanMaxWords * 4 bytes until exit; every call shares them, the array keeps
// the previous call's contents (the extrema codes are reset by the kernel below on every call, the
// bitmap words of the call's B*C channels are all overwritten), and two calls that overlap on the GPU
// race on it.
static __device__ __align__(256) uint32_t g_static_scan_scratch[kStaticScanMaxWords];

namespace {

// extrema[2b] collects the max (starts at the smallest code), extrema[2b+1] the min (the largest)
__global__ void gmm_static_scan_reset_kernel(int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) { g_static_scan_scratch[2 * b] = 0u; g_static_scan_scratch[2 * b + 1] = 0xffffffffu; }
}

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
uint32_t* g_static_scan_addr[kMaxDevices] = {};       // DEVICE address of the array, per device

}  // namespace

cudaError_t launch_gmm_prologue_scan_static(const float* y, int B, int C, int64_t HW, float* yq, long long* bitmap64,
                                            const uint32_t** scratch_out, cudaStream_t st) {
    // In host code `g_static_scan_scratch` is a host-side stand-in, not a device address: the address
    // the kernel gets comes from cudaGetSymbolAddress, resolved once per device with that device current.
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    uint32_t* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_scan_addr[dev]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scan_scratch);
            if (e != cudaSuccess) return e;
            g_static_scan_addr[dev] = static_cast<uint32_t*>(p);
        }
        scratch = g_static_scan_addr[dev];
    }
    if (scratch_out) *scratch_out = scratch;
    if (B <= 0) return cudaSuccess;
    if ((int64_t)B * C + 2 * (int64_t)B > kStaticScanMaxWords) return cudaErrorInvalidValue;
    gmm_static_scan_reset_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(B);
    e = cudaGetLastError();
    if (e != cudaSuccess || C <= 0) return e;
    gmm_prologue_scan_kernel<<<(unsigned)(B * C), kPrThreads, 0, st>>>(y, C, HW, yq, bitmap64,
                                                                      reinterpret_cast<int32_t*>(scratch + 2 * B), scratch);
    return cudaGetLastError();
}

float gmm_extremum_from_bits(uint32_t code) {
    const uint32_t b = (code & 0x80000000u) ? (code & 0x7fffffffu) : ~code;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

}  // namespace flashgmm
