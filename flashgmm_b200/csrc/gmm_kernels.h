// gmm_kernels.h -- internal launcher interface between the C-ABI layer and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <stdint.h>

namespace flashgmm {

// (N, 4) parameter views in element strides. The reference's callers hand over
// K-major planes: stride_n == 1, stride_k == N (entropy_models.py:813-827).
struct GmmParamsView {
    const float* scales;
    const float* means;
    const float* weights;
    int64_t stride_n;
    int64_t stride_k;
};

// 0 = K-major 128-bit path, 1 = AoS (N,4) contiguous, 2 = generic strided.
int gmm_layout_of(const GmmParamsView& p, const void* sym, const void* out);

// bypass_flag (nullable, device, pre-zeroed): set to 1 if any symbol needs bypass.
cudaError_t launch_gmm_startfreq(const int32_t* sym, const GmmParamsView& p, int64_t n, int mode,
                                 uint32_t* out, uint32_t* bypass_flag, cudaStream_t st);

// Decode-table row for search radius R (uint16 elements): 2R+2 CDF entries, 1 flag
// slot, zero padding up to lut_offset (multiple of 8), then -- if with_lut -- a
// 32-entry jump table. The row stride is decided by the host search code
// (rans_host.hpp: table_row_stride / table_has_lut).
int gmm_table_lut_offset(int R);

cudaError_t launch_gmm_cdf_table(const GmmParamsView& p, int64_t n, int mode, int R, uint16_t* table,
                                 int row_stride, bool with_lut, cudaStream_t st);

// ---- device-resident decode: batched rows + one-warp-per-stream pop (gmm_pop.cu) ----
// One entry per (stream, chunk): the chunk's rows are produced by the batched rows kernel
// and consumed by the pop kernel. Plain device pointers; arrays live in device memory.
struct RowsDesc {
    GmmParamsView p;          // already offset to the chunk's first symbol
    uint32_t* table32;        // rows out
    int32_t n;                // symbols in this chunk
    int32_t R;
    int32_t row_words;        // compact row (internal tile) width in uint32
    int32_t first_block;      // exclusive prefix of ceil(n / S) over the launch's entries
};
struct PopDesc {
    const uint16_t* rows;     // lane-row records (gmm_kernels.cu), 16-byte aligned
    const uint32_t* words;    // the stream's 32-bit words (device copy)
    int32_t* out;             // decoded symbols of this chunk
    uint64_t* state;          // [0] = x, [1] = word cursor | overrun << 32 (carried between chunks)
    GmmParamsView p;          // the chunk's parameters (symbols the lanes cannot place are searched exactly)
    int32_t n;                // symbols in this chunk (<= 0: nothing to do)
    int32_t R;
    uint32_t n_words;
    int32_t first;            // 1 = initialise the state from the stream's first two words
};
constexpr int kPopMaxR = 63;                       // lane-row records: 2R + 1 <= 128 intervals = 4 per lane
constexpr int kPopWindowed = 1;                    // lane_slots value selecting the windowed record (any R)
constexpr int kPopWindowWords = 68;                // 272 bytes per symbol
constexpr int kPopWindowMaxR = 1 << 20;           // the windowed record costs the same for any search radius
constexpr int kRowsWindowSyms = 32;                // symbols per CTA of the windowed rows kernel
inline int gmm_pop_lane_slots(int max_R) {             // 1, 2 or 4 intervals per lane
    const int ni = (2 * max_R + 1 + 31) / 32;
    return 32 * (ni == 3 ? 4 : ni);
}
// descs_d[0 .. n_desc) with first_block prefix; total_blocks = sum ceil(n / S).
// lane_slots == 0: compact rows (row_words uint32 per symbol); kPopWindowed: windowed records;
// 32 / 64 / 128: lane-row records.
cudaError_t launch_gmm_cdf_rows_batched(const RowsDesc* descs_d, int n_desc, int total_blocks, int S,
                                        int max_row_words, int lane_slots, int mode, cudaStream_t st);
// claim: 1025 zeroed ints per launch (next stream, streams taken per SM) or NULL (CTA i = stream i);
// quota: streams one SM may take
constexpr int kPopClaimInts = 1025;
cudaError_t launch_gmm_pop_rows(const PopDesc* descs_d, int n_desc, int lane_slots, int mode, int* claim, int quota,
                                cudaStream_t st);

// ---- device-resident encode: batched startfreq + one-warp-per-stream push (gmm_push.cu) ----
constexpr int kSfBypassSlots = 15;
struct SfDesc {
    GmmParamsView p;
    const int32_t* sym;
    uint32_t* packed;         // out: start | freq << 16 per symbol, 0x0000ffff = bypass
    int32_t n;
    int32_t first_block;      // exclusive prefix of ceil(n / 128)
    uint32_t* flag;           // optional, 32 words, zeroed by the caller: [0] counts the bypass records of the range,
                              // [1 + 2j], [2 + 2j] = index and symbol value of the j-th (first kSfBypassSlots only)
};
cudaError_t launch_gmm_startfreq_batched(const SfDesc* descs_d, int n_desc, int total_blocks, int mode, cudaStream_t st);
cudaError_t launch_gmm_sf_clear_flags(const SfDesc* descs_d, int n_desc, cudaStream_t st);

// Exact reciprocal of every 16-bit frequency: q = (umul64hi(x, rcp) >> shift) == x / freq for
// x < 2^63 (rans_host.hpp::RcpEntry, built on the host, uploaded once per ctx).
struct PushRcp { unsigned long long rcp; uint32_t cmpl; uint16_t shift; uint16_t bias; };   // == rans_host.hpp::RcpEntry
struct PushDesc {
    const uint32_t* packed;   // n records (device)
    const int32_t* sym;       // symbol values (bypass records only)
    uint32_t* buf;            // scratch for the stream's words, filled from the end
    uint32_t cap_words;       // scratch capacity == the caller's output capacity in words
    int32_t n;
    uint32_t* result;         // 8 words: [0] words produced (needed, if it overflowed), [1] offset in `out`, [2] overflow, [4..7] %globaltimer at entry / exit
};
// All streams' words end up back to back in `out` (claimed with one atomicAdd per stream on
// *out_used, in words); result[1] says where.
cudaError_t launch_gmm_push(const PushDesc* descs_d, int n_desc, const PushRcp* rcp_d, uint32_t* out,
                            unsigned long long* out_used, unsigned int* started, cudaStream_t st);
// blocks `st` on the device until *started >= target (every push warp resident) or 3 ms have passed
cudaError_t launch_gmm_wait_started(const unsigned int* started, unsigned int target, cudaStream_t st);

// Interleaved multi-lane rANS (gmm_ileave.cu). scratch: n_seg * gmm_ileave_scratch_words()
// uint16; seg_words: n_seg uint32; out: device buffer for the whole container.
int64_t gmm_ileave_scratch_words(int seg_log2);
cudaError_t launch_gmm_ileave_encode(const int32_t* sym, const GmmParamsView& p, int64_t n, int mode, int R,
                                     int seg_log2, uint16_t* scratch, uint32_t* seg_words, uint8_t* out,
                                     unsigned long long* total_bytes, cudaStream_t st);
cudaError_t launch_gmm_ileave_decode(const uint8_t* enc, const GmmParamsView& p, int64_t n, int mode, int R,
                                     int seg_log2, uint32_t n_seg, int32_t* out, uint32_t* err, cudaStream_t st);

// GaussianMixtureConditional likelihood (gmm_likelihood.cu). x / y / lik / grad are
// (B, plane) contiguous; parameters are (B, K, plane) contiguous (plane = M*H*W).
// quant: 0 = y = x, 1 = y = x + noise, 2 = y = rint(x).
cudaError_t launch_gmm_likelihood_fwd(const float* x, const float* noise, int quant, const float* sc,
                                      const float* mu, const float* w, int64_t B, int64_t plane, int K,
                                      float scale_bound, float lik_bound, float* y_out, float* lik_out,
                                      cudaStream_t st);
cudaError_t launch_gmm_likelihood_bwd(const float* y, const float* sc, const float* mu, const float* w,
                                      const float* grad, int64_t B, int64_t plane, int K, float scale_bound,
                                      float lik_bound, float* gy, float* gsc, float* gmu, float* gw,
                                      cudaStream_t st);

// ---- compress / decompress glue (gmm_prologue.cu), batches of images ----
// One launch covers up to kPrologueMaxBatch images (the per-image output offsets travel as a kernel
// argument). Image b: y / y_quantized / y_hat at b*C*HW, parameters at b*K*C*HW, bitmap at b*C;
// its kept symbols at symbol offset off[b] of the outputs, its K-major block [K][n[b]] at K*off[b].
constexpr int kPrologueMaxBatch = 64;
struct PrologueBatch { int64_t off[kPrologueMaxBatch]; int64_t n[kPrologueMaxBatch]; };
// extrema: 2 words per image (order-preserving codes of max and min of the unrounded y)
cudaError_t launch_gmm_prologue_scan(const float* y, int B, int C, int64_t HW, float* yq, long long* bitmap64,
                                     int32_t* bitmap32, uint32_t* extrema, cudaStream_t st);
// The same scan with its scratch (extrema codes, bitmap words: 2B + B*C uint32) in a file-scope __device__ array of
// kStaticScanMaxWords words instead of a caller / ctx buffer: allocates nothing. *scratch_out = the array's DEVICE
// address on the current device ([0, 2B) extrema codes, then the B*C bitmap words). cudaErrorInvalidValue beyond the
// maximum. Calls share the array: calls that overlap on the GPU race.
constexpr int64_t kStaticScanMaxWords = 65536;          // 256 KiB per GPU, held until exit
cudaError_t launch_gmm_prologue_scan_static(const float* y, int B, int C, int64_t HW, float* yq, long long* bitmap64,
                                            const uint32_t** scratch_out, cudaStream_t st);
cudaError_t launch_gmm_prologue_gather(const float* yq, const float* scales, const float* means, const float* weights,
                                       const long long* bitmap, int B, int C, int64_t HW, int K, const PrologueBatch& meta,
                                       int32_t* sym_out, float* scales_out, float* means_out, float* weights_out,
                                       cudaStream_t st);
cudaError_t launch_gmm_epilogue_scatter(const int32_t* symbols, const long long* bitmap, int B, int C, int64_t HW,
                                        const PrologueBatch& meta, float* y_hat, cudaStream_t st);
float gmm_extremum_from_bits(uint32_t code);      // inverse of the order-preserving code the scan kernel reduces with


// ---- hyper-latent (z) branch: factorised-prior table coder records (gmm_table.cu) ----
// rec[i] (i over (B, C, HW)): bits 0-15 start, 16-31 freq, 32-62 raw bypass value, bit 63 = the symbol took the
// table's sentinel slot and carries a bypass payload. symbols_out (optional) receives round(z - median).
// *err is set to 1 if a bypass value does not fit 31 bits.
cudaError_t launch_eb_records(const float* z, const float* medians, const int32_t* cdfs, const int32_t* sizes,
                              const int32_t* offsets, int64_t n_img, int C, int64_t HW, int cdf_stride,
                              unsigned long long* rec, int32_t* symbols_out, uint32_t* err, cudaStream_t st);

// EntropyBottleneck._likelihood for filters (3, 3, 3, 3) (eb_likelihood.cu): values / out (C, N), params (C, 58).
cudaError_t launch_eb_likelihood(const float* values, const float* params, int64_t C, int64_t N, float* out, cudaStream_t st);

}  // namespace flashgmm
