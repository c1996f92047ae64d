/*
 * flashgmm.h -- C ABI of libflashgmm_b200.so (sm_100a), the drop-in boundary for
 * FlashGMM's Gaussian-mixture entropy-coding hot path.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the
 * reference repository root, tokkiwa/FlashGMM):
 *   RI  = compressai/cpp_exts/rans/rans_interface.cpp
 *   EM  = compressai/entropy_models/entropy_models.py
 *
 * Conventions
 *  - plain pointers and sizes only; no ownership crosses the ABI (caller allocates).
 *  - every function returns a flashgmm_status; flashgmm_last_error() gives the
 *    message for the calling thread.
 *  - GMM parameters are three float32 (N, 4) views given as base pointer + element
 *    strides (stride_n, stride_k). The reference's callers pass K-major planes,
 *    stride_n = 1, stride_k = N (EM:813-827); (N,4) row-major is stride_n = 4,
 *    stride_k = 1. Any strides are accepted for device memory; host memory must be
 *    one of those two layouts.
 *  - `mem` says where symbols / parameters / decoded output live:
 *    FLASHGMM_MEM_HOST (page-locked buffers are read by the kernels in place over PCIe,
 *    pageable ones are staged in chunks inside the call) or
 *    FLASHGMM_MEM_DEVICE (used in place on the ctx's device). Bitstreams are
 *    always host memory (they are what leaves the machine); the serial rANS state
 *    recurrence runs on the host for single streams and on the GPU, one warp per
 *    stream, for batches (see the batch calls below).
 *  - `mode` is APPROX_MODE with the reference's CODE numbering (RI:98, 223-233):
 *    0 Polya, 1 Abramowitz-Stegun, 2 logistic.
 *  - K is fixed at 4, the only instantiation the reference binds (RI:60, 982).
 *  - a ctx serialises its calls internally; use one ctx per host thread for
 *    concurrency. There is no CPU fallback: without a CUDA device ctx_create fails.
 */
#ifndef FLASHGMM_H_
#define FLASHGMM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct flashgmm_ctx flashgmm_ctx;

typedef enum {
    FLASHGMM_OK = 0,
    FLASHGMM_EINVAL = 1,   /* bad argument (null pointer, unsupported strides, mode, K) */
    FLASHGMM_ECUDA = 2,    /* CUDA runtime error; message in flashgmm_last_error() */
    FLASHGMM_ETRUNC = 3,   /* bitstream shorter than the reads it implies */
    FLASHGMM_ENOSPC = 4    /* caller's output buffer too small; *out_len = needed */
} flashgmm_status;

enum { FLASHGMM_MEM_HOST = 0, FLASHGMM_MEM_DEVICE = 1 };

int flashgmm_version(void);
const char* flashgmm_last_error(void);

/* Optional, before the process creates its CUDA context: ask the driver for n hardware queues
 * (CUDA_DEVICE_MAX_CONNECTIONS, unless the user has set it). The library keeps ~20 CUDA streams busy;
 * with the default 8 queues it is correct but copies can queue behind long kernels. */
int flashgmm_request_hw_queues(int n);

/* Context = device + streams + staging buffers + host worker threads. */
int flashgmm_ctx_create(int device, flashgmm_ctx** out);
void flashgmm_ctx_destroy(flashgmm_ctx* ctx);

/* Upper bound of an encoded stream for n symbols (every symbol bypassed):
 * cf. the reference's n_entries + 16 words buffer, RI:563. */
size_t flashgmm_max_encoded_bytes(int64_t n);

/* ---- kernel-level entry points (device pointers, asynchronous on `stream`) ---- */

/* Per-symbol rANS (start, freq): out[i] = start | freq << 16; freq == 0 marks a
 * symbol the reference sends through bypass (pmf == 0).  Replaces the loop body of
 * BufferedRansEncoder::encode_with_indexes_gmm<4>, RI:486-522 (+ _fast_gmm_cdf<4>
 * RI:259-284 and the approximations RI:135-214). `stream` is a cudaStream_t (NULL =
 * the ctx's own stream, synchronised before return). */
int flashgmm_gmm_startfreq(flashgmm_ctx* ctx, const int32_t* symbols_d, const float* scales_d,
                           const float* means_d, const float* weights_d, int64_t n,
                           int64_t stride_n, int64_t stride_k, int mode, uint32_t* packed_out_d,
                           void* stream);

/* Row stride, in uint16 elements, of the decode table for search radius R. */
int flashgmm_gmm_table_row_stride(int32_t R);

/* Quantised CDF rows for the search decoder: table[i*row_stride + j] =
 * uint16(cdf_i(v - 0.5) * 65535), v = j - R, j in [0, 2R+2); slot 2R+2 is 1 when the
 * row is not non-decreasing. Replaces the CDF evaluations inside the binary search
 * of RansDecoder::decode_with_indexes_gmm<4>, RI:833-862. */
int flashgmm_gmm_cdf_table(flashgmm_ctx* ctx, const float* scales_d, const float* means_d,
                           const float* weights_d, int64_t n, int64_t stride_n, int64_t stride_k,
                           int mode, int32_t R, uint16_t* table_d, int64_t row_stride, void* stream);

/* ---- whole-call entry points (what the reference's pybind methods do) ---- */

/* RansEncoder.encode_with_indexes_gmm(symbols, scales, means, weights, max_value)
 * -> bytes, RI:609-617 (+ RI:458-585). Byte-identical output. max_value is unused by
 * the reference and therefore not a parameter. */
int flashgmm_gmm_encode(flashgmm_ctx* ctx, const int32_t* symbols, const float* scales,
                        const float* means, const float* weights, int64_t n, int64_t stride_n,
                        int64_t stride_k, int mode, int mem, uint8_t* out, size_t out_cap,
                        size_t* out_len);

/* RansDecoder.decode_with_indexes_gmm(encoded, scales, means, weights, max_bs_value)
 * -> int32[N], RI:766-883. symbols_out lives in `mem`. */
int flashgmm_gmm_decode(flashgmm_ctx* ctx, const uint8_t* encoded, size_t encoded_len,
                        const float* scales, const float* means, const float* weights, int64_t n,
                        int64_t stride_n, int64_t stride_k, int mode, int mem, int32_t max_bs_value,
                        int32_t* symbols_out);

/* Many independent streams in one call (one per checkerboard pass / channel group /
 * image). The rANS recurrence is serial within a stream, so the stream is the unit of
 * parallelism:
 *   decode: batches of >= 4 streams (FLASHGMM_POP_MIN_STREAMS) never leave the GPU, whatever the search radius -- a
 *           batched rows kernel writes, per symbol, a 32-interval window of the quantised CDF chosen from the parameters
 *           (33 exact boundaries; saturated ones from the closed form), a pop kernel walks each stream's state with one
 *           warp and replays the reference's search literally for whatever falls outside the window (gmm_pop.cu);
 *           smaller batches run the full-row kernel per stream and pop on a host thread (lowest latency for one stream);
 *   encode: one startfreq pass, then the serial push of the longest streams on the worker
 *           threads and of the rest on the GPU (push kernel, one warp per stream, gmm_push.cu),
 *           both at once, split by a cost model (FLASHGMM_PUSH=host|device|auto).
 * Either way the bytes / symbols are the reference's. Page-locked host buffers are read by
 * the kernels in place; pageable ones are staged. */
typedef struct {
    const int32_t* symbols;   /* encode: in.  decode: unused (may be NULL) */
    const float* scales;
    const float* means;
    const float* weights;
    int64_t n;
    int64_t stride_n;
    int64_t stride_k;
    int32_t max_bs_value;     /* decode only */
    uint8_t* bytes;           /* encode: out buffer.  decode: in stream */
    size_t bytes_cap;         /* encode: capacity.    decode: stream length */
    size_t bytes_len;         /* encode: out, produced length */
    int32_t* symbols_out;     /* decode: out */
    int status;               /* out: per-stream flashgmm_status */
} flashgmm_stream;

int flashgmm_gmm_encode_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode,
                              int mem);
int flashgmm_gmm_decode_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode,
                              int mem);

/* ---- opt-in interleaved multi-lane rANS (own container format, GPU encodes AND decodes) ----
 * No reference counterpart (the reference's single 64-bit state, rans64.h, is serial). Same
 * per-symbol model as the compatible path -- the quantised CDF of RI:498-517 -- coded with
 * 32 independent 32-bit rANS states per segment of 2^segment_log2 symbols (10..16;
 * 0 = default 13). Symbols outside [-R, R] or with pmf == 0 are escaped (raw int32).
 * Format: flashgmm_b200/csrc/gmm_ileave.cu. Bitstreams are host memory. */
size_t flashgmm_gmm_ileave_max_bytes(int64_t n, int segment_log2);
int flashgmm_gmm_ileave_encode(flashgmm_ctx* ctx, const int32_t* symbols, const float* scales,
                               const float* means, const float* weights, int64_t n, int64_t stride_n,
                               int64_t stride_k, int mode, int mem, int32_t R, int segment_log2,
                               uint8_t* out, size_t out_cap, size_t* out_len);
/* Reads n / mode / R / segment size from the container header (NULL outputs are skipped). */
int flashgmm_gmm_ileave_info(const uint8_t* encoded, size_t encoded_len, int64_t* n, int* mode,
                             int32_t* R, int* segment_log2);
/* n must equal the header's; the APPROX_MODE used is the header's. */
int flashgmm_gmm_ileave_decode(flashgmm_ctx* ctx, const uint8_t* encoded, size_t encoded_len,
                               const float* scales, const float* means, const float* weights, int64_t n,
                               int64_t stride_n, int64_t stride_k, int mem, int32_t* symbols_out);

/* ---- training / evaluation likelihood (device pointers, asynchronous on `stream`) ---- */

/* GaussianMixtureConditional.forward, EM:792-808 (+ _likelihood EM:774-790, 720-737,
 * _standardized_cumulative EM:674-678, quantize EM:157-182, LowerBound ops/bound_ops.py:36-42).
 *   x, noise, y_out, lik_out : (B, M*HW) float32 contiguous
 *   scales, means, weights   : (B, K, M*HW) float32 contiguous -- the reference's
 *                              (B, K*M, H, W) tensors, component k of channel c at channel k*M + c
 *   quant : 0 y = x | 1 y = x + noise (training "noise") | 2 y = round-half-even(x) ("dequantize")
 *   scale_bound (0.11) and lik_bound (1e-9; <= 0 disables) are the two LowerBounds.
 * K in [1, 8]. Contract: 1e-5 relative against the reference's fp32 result. */
int flashgmm_gmm_likelihood_fwd(flashgmm_ctx* ctx, const float* x_d, const float* noise_d, int quant,
                                const float* scales_d, const float* means_d, const float* weights_d,
                                int64_t B, int64_t plane, int K, float scale_bound, float lik_bound,
                                float* y_out_d, float* lik_out_d, void* stream);

/* Backward of the above through the likelihood output: grad_lik -> grad_y (add to the
 * incoming grad of y_out for quant 0/1; round has zero gradient), grad_scales,
 * grad_means, grad_weights, with LowerBound's custom gradient on both bounds
 * (ops/bound_ops.py:40-42). y_d is the forward's y_out. */
int flashgmm_gmm_likelihood_bwd(flashgmm_ctx* ctx, const float* y_d, const float* scales_d,
                                const float* means_d, const float* weights_d, const float* grad_lik_d,
                                int64_t B, int64_t plane, int K, float scale_bound, float lik_bound,
                                float* grad_y_d, float* grad_scales_d, float* grad_means_d,
                                float* grad_weights_d, void* stream);

/* ---- GaussianMixtureConditional.compress / decompress glue on the device ----
 * EM = compressai/entropy_models/entropy_models.py. Layouts are the reference's: y (1, C, H, W)
 * contiguous, parameters (1, K*C, H, W) contiguous (component k of channel c at channel k*C + c),
 * HW = H*W. All pointers are device pointers unless named _h. */

/* EM:834-845: y_quantized = round(y); zero bitmap (1 = the channel has a nonzero symbol), as int64
 * on the device (bitmap_out_d, may be NULL) and as int32 on the host; max and min of the UNROUNDED
 * y (y_max_h, y_min_h; the caller forms abs_max as EM:834-837 does). One kernel, one small copy,
 * one synchronisation of `stream`. */
int flashgmm_gmm_prologue_scan(flashgmm_ctx* ctx, const float* y_d, int64_t C, int64_t HW,
                               float* y_quantized_out_d, int64_t* bitmap_out_d, int32_t* bitmap_out_h,
                               float* y_max_h, float* y_min_h, void* stream);

/* EM:846-847 + reshape_entropy_parameters EM:810-828: the kept channels' symbols as int32 and
 * their parameters as K-major planes [K][n_sym] (n_sym = kept channels * HW; scales clamped to
 * [0.11, 256]) -- the (n_sym, K) views with strides (1, n_sym) the coder reads. y_quantized_d and
 * symbols_out_d may both be NULL (decompress needs the parameters only). Asynchronous on `stream`. */
int flashgmm_gmm_prologue_gather(flashgmm_ctx* ctx, const float* y_quantized_d, const float* scales_d,
                                 const float* means_d, const float* weights_d, const int64_t* bitmap_d,
                                 int64_t C, int64_t HW, int K, int64_t n_sym, int32_t* symbols_out_d,
                                 float* scales_out_d, float* means_out_d, float* weights_out_d, void* stream);

/* EM:902-909: y_hat (1, C, H, W) = decoded symbols of the kept channels as float, zeros elsewhere.
 * Asynchronous on `stream`. */
int flashgmm_gmm_epilogue_scatter(flashgmm_ctx* ctx, const int32_t* symbols_d, const int64_t* bitmap_d,
                                  int64_t C, int64_t HW, float* y_hat_out_d, void* stream);

/* Batch forms: B images of one pass in ONE launch each. y / y_quantized / y_hat are (B, C, H, W) contiguous,
 * parameters (B, K*C, H, W), bitmaps (B, C). Per-image outputs are concatenated: image b's kept symbols
 * start at symbol offset off_b = n_sym_h[0] + ... + n_sym_h[b-1] of symbols_out_d, and its K-major block
 * [K][n_sym_h[b]] starts at element K * off_b of each parameter output (so the coder's view of image b is
 * base + K*off_b, stride_n = 1, stride_k = n_sym_h[b]). extrema_out_h receives 2 floats per image
 * (max, min of the unrounded y); bitmap_out_h B*C int32. */
int flashgmm_gmm_prologue_scan_batch(flashgmm_ctx* ctx, const float* y_d, int64_t B, int64_t C, int64_t HW,
                                     float* y_quantized_out_d, int64_t* bitmap_out_d, int32_t* bitmap_out_h,
                                     float* extrema_out_h, void* stream);
/* flashgmm_gmm_prologue_scan_batch for callers that cannot allocate (code under CUDA-graph capture set-up, tight
 * serving loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as flashgmm_gmm_prologue_scan_batch, but no ctx (a ctx owns
 * growable buffers): the kernel's scratch (2 extrema codes per image + one bitmap word per channel) lives in a
 * file-scope __device__ array of FLASHGMM_NOALLOC_SCAN_MAX_WORDS 32-bit words (256 KiB; every GPU that loads the
 * library holds it until exit), and the host results are copied straight into bitmap_out_h / extrema_out_h. The call
 * makes no cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine. B * (C + 2) must not
 * exceed FLASHGMM_NOALLOC_SCAN_MAX_WORDS: FLASHGMM_EINVAL otherwise (24 ELIC images are 24 * 322 = 7728).
 * `device`: CUDA device ordinal of y_d; made current. Two launches, two small copies, one synchronisation of
 * `stream`. All calls on a device share the array (it keeps the previous call's contents; what must start clean
 * is reset on every call): calls that OVERLAP on the GPU -- two host threads, two streams -- race; serialise them.
 * The first call on a device resolves the array's device address (cudaGetSymbolAddress). */
#define FLASHGMM_NOALLOC_SCAN_MAX_WORDS 65536
int flashgmm_gmm_prologue_scan_batch_noalloc(int device, const float* y_d, int64_t B, int64_t C, int64_t HW,
                                             float* y_quantized_out_d, int64_t* bitmap_out_d, int32_t* bitmap_out_h,
                                             float* extrema_out_h, void* stream);
int flashgmm_gmm_prologue_gather_batch(flashgmm_ctx* ctx, const float* y_quantized_d, const float* scales_d,
                                       const float* means_d, const float* weights_d, const int64_t* bitmap_d,
                                       int64_t B, int64_t C, int64_t HW, int K, const int64_t* n_sym_h,
                                       int32_t* symbols_out_d, float* scales_out_d, float* means_out_d,
                                       float* weights_out_d, void* stream);
int flashgmm_gmm_epilogue_scatter_batch(flashgmm_ctx* ctx, const int32_t* symbols_d, const int64_t* bitmap_d,
                                        const int64_t* n_sym_h, int64_t B, int64_t C, int64_t HW,
                                        float* y_hat_out_d, void* stream);

/* ---- hyper-latent z branch: EntropyBottleneck.compress / decompress for B images, z on the device ----
 * EM:605-618 -> EntropyModel.compress / decompress EM:237-327 -> the table coder RI:334-399, 619-688,
 * without the five .tolist() marshals per image: the CDF lookup runs in a kernel (gmm_table.cu), the
 * serial push / pop on the ctx's worker threads (one image each). Element (b, c, i) uses CDF row c
 * (EntropyBottleneck._build_indexes, EM:556-567). Byte-identical streams.
 *   compress:   z_d (B, C, HW) float32, medians_d (C), cdfs_d (C, cdf_stride) int32, cdf_sizes_d / offsets_d (C):
 *               device. streams[b].bytes / bytes_cap: host output buffer; bytes_len / status: out.
 *               symbols_out_d (optional, device int32 (B, C, HW)) receives round(z - median).
 *   decompress: streams[b].bytes / bytes_cap: the stream; cdfs / sizes / offsets: HOST copies;
 *               symbols_out_d (device int32 (B, C, HW)); the caller adds the medians (EM:185-194). */
int flashgmm_eb_compress_batch(flashgmm_ctx* ctx, const float* z_d, const float* medians_d, const int32_t* cdfs_d,
                               const int32_t* cdf_sizes_d, const int32_t* offsets_d, int64_t B, int64_t C, int64_t HW,
                               int64_t cdf_stride, int32_t* symbols_out_d, flashgmm_stream* streams, void* stream);
int flashgmm_eb_decompress_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int64_t B, const int32_t* cdfs_h,
                                 const int32_t* cdf_sizes_h, const int32_t* offsets_h, int64_t C, int64_t HW,
                                 int64_t cdf_stride, int32_t* symbols_out_d, void* stream);

/* EntropyBottleneck._likelihood (EM:434-463 with _logits_cumulative EM:421-440) for filters (3, 3, 3, 3) as one kernel:
 * values_d / out_d (C, N) float32 channel-major (what EntropyBottleneck.forward builds, EM:479-481); params_d (C, 58) =
 * per channel softplus(matrix), bias, tanh(factor) of the five layers in order (layout: csrc/eb_likelihood.cu).
 * out = sigmoid(logits(v + 1/2)) - sigmoid(logits(v - 1/2)); the caller applies the likelihood bound. Asynchronous on
 * `stream`. Contract: 1e-5 relative against the reference's tensor ops. */
int flashgmm_eb_likelihood(flashgmm_ctx* ctx, const float* values_d, const float* params_d, int64_t C, int64_t N,
                           float* out_d, void* stream);

/* ---- host-only halves (no GPU): the serial rANS recurrence ---- */

/* Push packed (start|freq<<16, freq==0 = bypass -> needs symbols[i]) records:
 * BufferedRansEncoder::flush, RI:557-585, incl. the bypass payload RI:524-552. */
int flashgmm_rans_push_packed(const uint32_t* packed, const int32_t* symbols, int64_t n,
                              uint8_t* out, size_t out_cap, size_t* out_len);

/* Pop n symbols using precomputed rows (flashgmm_gmm_cdf_table layout):
 * the state/search/advance loop of RI:793-883. */
int flashgmm_rans_pop_table(const uint8_t* encoded, size_t encoded_len, const uint16_t* table,
                            int64_t row_stride, int64_t n, int32_t R, int32_t* symbols_out);

/* Table-driven coder kept for the hyper-latent z branch:
 * RansEncoder.encode_with_indexes(symbols, indexes, cdfs, cdfs_sizes, offsets), RI:334-399, 587-598
 * RansDecoder.decode_with_indexes(encoded, indexes, cdfs, cdfs_sizes, offsets), RI:619-688.
 * cdfs is a dense [n_cdfs][cdf_stride] int32 matrix (CompressAI's _quantized_cdf). */
int flashgmm_rans_encode_table(const int32_t* symbols, const int32_t* indexes, int64_t n,
                               const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, uint8_t* out,
                               size_t out_cap, size_t* out_len);
int flashgmm_rans_decode_table(const uint8_t* encoded, size_t encoded_len, const int32_t* indexes,
                               int64_t n, const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, int32_t* symbols_out);

/* Scalar single-Gaussian search coder, host only (the reference's is scalar libm code on the CPU and no
 * shipped model calls it): RansEncoder.encode_with_indexes(symbols, scales, max_value) RI:401-456, 600-607 and
 * RansDecoder.decode_with_indexes(encoded, scales, max_bs_value) RI:690-764, zero-mean Gaussian of scale
 * scales[i] per symbol through the SCALAR approximations RI:148-150, 188-204, 216-219 (libm expf / sqrtf).
 * Parity is unpinned across hosts (depends on the host's libm); byte-identical to the reference built on
 * the same machine. */
int flashgmm_scalar_gaussian_encode(const int32_t* symbols, const float* scales, int64_t n, int mode,
                                    uint8_t* out, size_t out_cap, size_t* out_len);
int flashgmm_scalar_gaussian_decode(const uint8_t* encoded, size_t encoded_len, const float* scales, int64_t n,
                                    int mode, int32_t max_bs_value, int32_t* symbols_out);

/* ---- buffered encoder / streaming decoder sessions (host state objects) ----
 * BufferedRansEncoder (RI:334-399, 458-585, 977-984): symbols of several calls accumulate
 * into one stream produced by flush(). */
typedef struct flashgmm_buffered_encoder flashgmm_buffered_encoder;
int flashgmm_benc_create(flashgmm_buffered_encoder** out);
void flashgmm_benc_destroy(flashgmm_buffered_encoder* enc);
/* BufferedRansEncoder::encode_with_indexes(symbols, indexes, cdfs, cdfs_sizes, offsets), RI:334-399 */
int flashgmm_benc_push_table(flashgmm_buffered_encoder* enc, const int32_t* symbols, const int32_t* indexes,
                             int64_t n, const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                             const int32_t* cdf_sizes, const int32_t* offsets);
/* BufferedRansEncoder::encode_with_indexes_gmm, RI:458-554, fed with the packed records of
 * flashgmm_gmm_startfreq (host copies); symbols are needed for bypass records only. */
int flashgmm_benc_push_gmm_packed(flashgmm_buffered_encoder* enc, const uint32_t* packed,
                                  const int32_t* symbols, int64_t n);
/* BufferedRansEncoder::flush, RI:557-585: emits the stream and clears the session. */
int flashgmm_benc_flush(flashgmm_buffered_encoder* enc, uint8_t* out, size_t out_cap, size_t* out_len);
int64_t flashgmm_benc_entries(const flashgmm_buffered_encoder* enc);

/* RansDecoder::set_stream / decode_stream, RI:886-956: the state persists between calls. */
typedef struct flashgmm_stream_decoder flashgmm_stream_decoder;
int flashgmm_sdec_create(const uint8_t* encoded, size_t encoded_len, flashgmm_stream_decoder** out);
void flashgmm_sdec_destroy(flashgmm_stream_decoder* dec);
int flashgmm_sdec_decode_table(flashgmm_stream_decoder* dec, const int32_t* indexes, int64_t n,
                               const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, int32_t* symbols_out);

/* Kernel launches issued by this process through the library (bench.py's gpu_launches). */
uint64_t flashgmm_launch_count(void);

/* Optional live kernel timing for roofline reports: when enabled, every kernel the
 * whole-call entry points launch is bracketed by CUDA events on its own stream.
 * kind 0 = gmm_startfreq (algorithmic bytes 56/symbol), 1 = gmm_cdf_table
 * / the batched rows kernel of the device decoder (48 B/symbol in + the record it writes),
 * 2 = interleaved encode (52 B/symbol + stream), 3 = interleaved decode (48 B/symbol +
 * stream + 4), 4 = gmm_pop_rows, the device decoder's serial half (record + 4 B/symbol),
 * 5 = gmm_push, the device encoder's serial half (4 B/symbol + the words it writes twice;
 * timed by the kernel's own %globaltimer stamps). */
void flashgmm_profile_enable(int on);
void flashgmm_profile_reset(void);
int flashgmm_profile_read(int kind, double* total_ms, uint64_t* launches, double* algorithmic_bytes);

#ifdef __cplusplus
}
#endif
#endif /* FLASHGMM_H_ */
