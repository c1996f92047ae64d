// gmm_table.cu -- the hyper-latent (z) branch's table coder, CDF half, on the GPU (sm_100a).
//
// The reference marshals five Python lists per image (symbols, indexes, the whole CDF matrix,
// sizes, offsets: /root/reference/compressai/entropy_models/entropy_models.py:237-270, called from
// EntropyBottleneck.compress :605-611) into BufferedRansEncoder::encode_with_indexes
// (compressai/cpp_exts/rans/rans_interface.cpp:334-399), which looks every symbol up in its
// channel's quantised CDF. Here z stays on the device: one thread per element
//   symbol = int(round(z - median_c))                       (EntropyModel.quantize "symbols", :157-182)
//   value  = symbol - offset_c;  out of [0, max) -> sentinel slot max + raw bypass value  (RI:356-363)
//   record = cdf_c[value] | (cdf_c[value+1] - cdf_c[value]) << 16 | raw << 32 | bypass << 63
// and only the 8-byte records go to the host, whose worker threads run the serial rANS push.
// Plain gather: HBM/L2-bound, coalesced over HW, the CDF rows (a few KB) stay in L1/L2.
#include "gmm_kernels.h"

namespace flashgmm {

namespace {

__global__ void __launch_bounds__(256)
eb_records_kernel(const float* __restrict__ z, const float* __restrict__ medians, const int32_t* __restrict__ cdfs,
                  const int32_t* __restrict__ sizes, const int32_t* __restrict__ offsets, int C, int64_t HW,
                  int cdf_stride, unsigned long long* __restrict__ rec, int32_t* __restrict__ symbols_out,
                  uint32_t* __restrict__ err) {
    const int64_t plane = blockIdx.y;                         // (image, channel)
    const int c = (int)(plane % C);
    const float med = __ldg(medians + c);
    const int32_t* cdf = cdfs + (int64_t)c * cdf_stride;
    const int32_t max_value = __ldg(sizes + c) - 2;
    const int32_t off = __ldg(offsets + c);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = rintf(__fsub_rn(__ldg(z + plane * HW + i), med));     // torch.round: half to even
        const int32_t sym = (int32_t)v;
        if (symbols_out) symbols_out[plane * HW + i] = sym;
        int32_t value = sym - off;
        uint32_t raw = 0;
        bool bypass = false;
        if (value < 0) { raw = (uint32_t)(-2 * value - 1); value = max_value; }
        else if (value >= max_value) { raw = (uint32_t)(2 * (value - max_value)); value = max_value; }
        bypass = value == max_value;
        if (raw >> 31) { atomicOr(err, 1u); raw &= 0x7fffffffu; }
        const uint32_t lo = (uint32_t)__ldg(cdf + value), hi = (uint32_t)__ldg(cdf + value + 1);
        const uint32_t packed = (lo & 0xffffu) | (((hi - lo) & 0xffffu) << 16);
        rec[plane * HW + i] = (unsigned long long)packed | ((unsigned long long)raw << 32) |
                              (bypass ? (1ull << 63) : 0ull);
    }
}

}  // namespace

cudaError_t launch_eb_records(const float* z, const float* medians, const int32_t* cdfs, const int32_t* sizes,
                              const int32_t* offsets, int64_t n_img, int C, int64_t HW, int cdf_stride,
                              unsigned long long* rec, int32_t* symbols_out, uint32_t* err, cudaStream_t st) {
    if (n_img <= 0 || C <= 0 || HW <= 0) return cudaSuccess;
    const unsigned gx = (unsigned)((HW + 255) / 256 < 64 ? (HW + 255) / 256 : 64);
    eb_records_kernel<<<dim3(gx, (unsigned)(n_img * C)), 256, 0, st>>>(z, medians, cdfs, sizes, offsets, C, HW, cdf_stride,
                                                                       rec, symbols_out, err);
    return cudaGetLastError();
}

}  // namespace flashgmm
