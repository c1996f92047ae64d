// abi_glue.cu -- C ABI of the GaussianMixtureConditional.compress / decompress glue
// (gmm_prologue.cu): round + zero bitmap + extrema, channel compaction into K-major planes,
// scatter of decoded symbols. Single-image entry points (the reference's call shape,
// /root/reference/compressai/entropy_models/entropy_models.py:833-910) and their batch forms
// (B images per pass of a checkerboard / channel-group codec, one launch each).
#include "abi_common.hpp"

namespace {

int scan_impl(flashgmm_ctx* ctx, const float* y_d, int64_t B, int64_t C, int64_t HW, float* yq_out_d,
              int64_t* bitmap_out_d, int32_t* bitmap_out_h, float* extrema_h, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || C < 0 || HW < 0 || C > (1 << 20) || B > (1 << 16))
        return fail(FLASHGMM_EINVAL, "bad shape B=%lld C=%lld HW=%lld", (long long)B, (long long)C, (long long)HW);
    if (B > 0 && C > 0 && HW > 0 && (!y_d || !yq_out_d)) return fail(FLASHGMM_EINVAL, "null tensor");
    if (B > 0 && (!bitmap_out_h || !extrema_h)) return fail(FLASHGMM_EINVAL, "null host output");
    if (B == 0) return FLASHGMM_OK;
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t words = (size_t)(B * C + 2 * B);
    CU_TRY(ctx->pr_dev.need(words * 4));
    CU_TRY(ctx->pr_host.need(words * 4));
    uint32_t* d = static_cast<uint32_t*>(ctx->pr_dev.p);          // [0, 2B) extrema codes, then bitmap32 (B*C)
    CU_TRY(launch_gmm_prologue_scan(y_d, (int)B, (int)C, HW, yq_out_d, reinterpret_cast<long long*>(bitmap_out_d),
                                    reinterpret_cast<int32_t*>(d + 2 * B), d, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(ctx->pr_host.p, d, words * 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    const uint32_t* h = static_cast<const uint32_t*>(ctx->pr_host.p);
    const bool empty = C == 0 || HW == 0;
    for (int64_t b = 0; b < B; ++b) {
        extrema_h[2 * b] = empty ? 0.0f : gmm_extremum_from_bits(h[2 * b]);
        extrema_h[2 * b + 1] = empty ? 0.0f : gmm_extremum_from_bits(h[2 * b + 1]);
    }
    memcpy(bitmap_out_h, h + 2 * B, (size_t)(B * C) * 4);
    return FLASHGMM_OK;
}

// per-image symbol counts -> kernel argument blocks of <= kPrologueMaxBatch images
template <class F>
int for_image_blocks(int64_t B, const int64_t* n_sym_h, int64_t HW, F&& body) {
    int64_t off = 0;
    for (int64_t b0 = 0; b0 < B; b0 += kPrologueMaxBatch) {
        PrologueBatch m;
        const int nb = (int)std::min<int64_t>(kPrologueMaxBatch, B - b0);
        const int64_t off0 = off;
        for (int i = 0; i < kPrologueMaxBatch; ++i) {
            m.off[i] = 0; m.n[i] = 0;
            if (i < nb) {
                const int64_t n = n_sym_h[b0 + i];
                if (n < 0 || (HW > 0 && n % HW != 0)) return fail(FLASHGMM_EINVAL, "n_sym[%lld]=%lld is not a multiple of HW", (long long)(b0 + i), (long long)n);
                m.off[i] = off - off0;      // relative to the block's first image: the body offsets the outputs by off0
                m.n[i] = n;
                off += n;
            }
        }
        const int rc = body(b0, nb, off0, m);
        if (rc) return rc;
    }
    return FLASHGMM_OK;
}

}  // namespace

extern "C" {

int flashgmm_gmm_prologue_scan(flashgmm_ctx* ctx, const float* y_d, int64_t C, int64_t HW, float* yq_out_d,
                               int64_t* bitmap_out_d, int32_t* bitmap_out_h, float* y_max_h, float* y_min_h,
                               void* stream) {
    if (!y_max_h || !y_min_h) return fail(FLASHGMM_EINVAL, "null host output");
    float ext[2] = {0.0f, 0.0f};
    const int rc = scan_impl(ctx, y_d, 1, C, HW, yq_out_d, bitmap_out_d, bitmap_out_h, ext, stream);
    *y_max_h = ext[0]; *y_min_h = ext[1];
    return rc;
}

int flashgmm_gmm_prologue_scan_batch(flashgmm_ctx* ctx, const float* y_d, int64_t B, int64_t C, int64_t HW,
                                     float* yq_out_d, int64_t* bitmap_out_d, int32_t* bitmap_out_h,
                                     float* extrema_out_h, void* stream) {
    return scan_impl(ctx, y_d, B, C, HW, yq_out_d, bitmap_out_d, bitmap_out_h, extrema_out_h, stream);
}

// No counterpart in the reference: the batch scan for callers that
// cannot allocate. The scratch is the file-scope device array of gmm_prologue.cu; the host results go straight into
// the caller's buffers. No cudaMalloc / cudaFree / cudaHostAlloc, no ctx (a ctx owns growable buffers).
static_assert(kStaticScanMaxWords == FLASHGMM_NOALLOC_SCAN_MAX_WORDS, "the header states the array's size");
int flashgmm_gmm_prologue_scan_batch_noalloc(int device, const float* y_d, int64_t B, int64_t C, int64_t HW,
                                             float* yq_out_d, int64_t* bitmap_out_d, int32_t* bitmap_out_h,
                                             float* extrema_out_h, void* stream) {
    if (B < 0 || C < 0 || HW < 0 || C > (1 << 20) || B > (1 << 16))
        return fail(FLASHGMM_EINVAL, "bad shape B=%lld C=%lld HW=%lld", (long long)B, (long long)C, (long long)HW);
    if (B * C + 2 * B > kStaticScanMaxWords)
        return fail(FLASHGMM_EINVAL, "B*(C+2) = %lld words of scratch exceed the static maximum of %lld (use "
                    "flashgmm_gmm_prologue_scan_batch, which allocates)", (long long)(B * C + 2 * B),
                    (long long)kStaticScanMaxWords);
    if (B > 0 && C > 0 && HW > 0 && (!y_d || !yq_out_d)) return fail(FLASHGMM_EINVAL, "null tensor");
    if (B > 0 && (!bitmap_out_h || !extrema_out_h)) return fail(FLASHGMM_EINVAL, "null host output");
    if (B == 0) return FLASHGMM_OK;
    CU_TRY(cudaSetDevice(device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint32_t* scratch = nullptr;
    CU_TRY(launch_gmm_prologue_scan_static(y_d, (int)B, (int)C, HW, yq_out_d, reinterpret_cast<long long*>(bitmap_out_d),
                                           &scratch, st));
    g_launches.fetch_add(C > 0 ? 2 : 1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(extrema_out_h, scratch, (size_t)(2 * B) * 4, cudaMemcpyDeviceToHost, st));
    if (C > 0) CU_TRY(cudaMemcpyAsync(bitmap_out_h, scratch + 2 * B, (size_t)(B * C) * 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    const bool empty = C == 0 || HW == 0;
    for (int64_t i = 0; i < 2 * B; ++i) {                    // order-preserving codes -> floats, in place
        uint32_t code;
        memcpy(&code, extrema_out_h + i, 4);
        extrema_out_h[i] = empty ? 0.0f : gmm_extremum_from_bits(code);
    }
    return FLASHGMM_OK;
}

int flashgmm_gmm_prologue_gather_batch(flashgmm_ctx* ctx, const float* yq_d, const float* scales_d, const float* means_d,
                                       const float* weights_d, const int64_t* bitmap_d, int64_t B, int64_t C, int64_t HW,
                                       int K, const int64_t* n_sym_h, int32_t* symbols_out_d, float* scales_out_d,
                                       float* means_out_d, float* weights_out_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || C < 0 || HW < 0 || K < 1 || K > 8 || C > 65535 || (B > 0 && !n_sym_h))
        return fail(FLASHGMM_EINVAL, "bad shape B=%lld C=%lld HW=%lld K=%d", (long long)B, (long long)C, (long long)HW, K);
    int64_t total = 0;
    for (int64_t b = 0; b < B; ++b) total += n_sym_h[b] > 0 ? n_sym_h[b] : 0;
    if (total > 0 && (!scales_d || !means_d || !weights_d || !bitmap_d || !scales_out_d || !means_out_d || !weights_out_d))
        return fail(FLASHGMM_EINVAL, "null tensor");
    if (total == 0) return FLASHGMM_OK;
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return for_image_blocks(B, n_sym_h, HW, [&](int64_t b0, int nb, int64_t off0, const PrologueBatch& m) -> int {
        CU_TRY(launch_gmm_prologue_gather(yq_d ? yq_d + b0 * C * HW : nullptr, scales_d + b0 * K * C * HW,
                                          means_d + b0 * K * C * HW, weights_d + b0 * K * C * HW,
                                          reinterpret_cast<const long long*>(bitmap_d) + b0 * C, nb, (int)C, HW, K, m,
                                          symbols_out_d ? symbols_out_d + off0 : nullptr, scales_out_d + K * off0,
                                          means_out_d + K * off0, weights_out_d + K * off0, st));
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return FLASHGMM_OK;
    });
}

int flashgmm_gmm_prologue_gather(flashgmm_ctx* ctx, const float* yq_d, const float* scales_d, const float* means_d,
                                 const float* weights_d, const int64_t* bitmap_d, int64_t C, int64_t HW, int K,
                                 int64_t n_sym, int32_t* symbols_out_d, float* scales_out_d, float* means_out_d,
                                 float* weights_out_d, void* stream) {
    if (n_sym < 0) return fail(FLASHGMM_EINVAL, "bad n_sym=%lld", (long long)n_sym);
    return flashgmm_gmm_prologue_gather_batch(ctx, yq_d, scales_d, means_d, weights_d, bitmap_d, 1, C, HW, K, &n_sym,
                                              symbols_out_d, scales_out_d, means_out_d, weights_out_d, stream);
}

int flashgmm_gmm_epilogue_scatter_batch(flashgmm_ctx* ctx, const int32_t* symbols_d, const int64_t* bitmap_d,
                                        const int64_t* n_sym_h, int64_t B, int64_t C, int64_t HW, float* y_hat_out_d,
                                        void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || C < 0 || HW < 0 || C > 65535 || B > 65535 || (B > 0 && !n_sym_h))
        return fail(FLASHGMM_EINVAL, "bad shape B=%lld C=%lld HW=%lld", (long long)B, (long long)C, (long long)HW);
    if (B > 0 && C > 0 && HW > 0 && (!bitmap_d || !y_hat_out_d)) return fail(FLASHGMM_EINVAL, "null tensor");
    if (B == 0 || C == 0 || HW == 0) return FLASHGMM_OK;
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return for_image_blocks(B, n_sym_h, HW, [&](int64_t b0, int nb, int64_t off0, const PrologueBatch& m) -> int {
        CU_TRY(launch_gmm_epilogue_scatter(symbols_d ? symbols_d + off0 : nullptr,
                                           reinterpret_cast<const long long*>(bitmap_d) + b0 * C, nb, (int)C, HW, m,
                                           y_hat_out_d + b0 * C * HW, st));
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return FLASHGMM_OK;
    });
}

int flashgmm_gmm_epilogue_scatter(flashgmm_ctx* ctx, const int32_t* symbols_d, const int64_t* bitmap_d, int64_t C,
                                  int64_t HW, float* y_hat_out_d, void* stream) {
    // the single-image form does not know the kept count: the kernel only needs the image's base offset (0)
    const int64_t n = 0;
    return flashgmm_gmm_epilogue_scatter_batch(ctx, symbols_d, bitmap_d, &n, 1, C, HW, y_hat_out_d, stream);
}

}  // extern "C"
