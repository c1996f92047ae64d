// gmm_ileave.cu -- opt-in interleaved multi-lane rANS coder (north_star item 4), sm_100a.
//
// No reference counterpart: the reference's stream is one serial 64-bit rANS state
// (third_party/ryg_rans/rans64.h) and therefore cannot be produced or consumed in
// parallel. This format keeps the reference's per-symbol model bit for bit -- the same
// quantised mixture CDF as gmm_math.cuh, the same 65 535 total with slot 65 535 as the
// escape marker (rans_interface.cpp:56, 509-517) -- but codes it with many independent
// rANS states so that both directions run entirely on the GPU:
//
//   * the stream is cut into segments of S = 2^segment_log2 symbols; one warp owns one
//     segment, lane l codes symbols l, l+32, l+64, ... of it with its own 32-bit state
//     (L = 2^16, 16-bit renormalisation words: at most one word per symbol per lane);
//   * per step the lanes that renormalise emit/read their words in lane order
//     (ballot + popc prefix), so a segment is ONE word stream read strictly forward;
//   * a symbol whose pmf is 0 or that lies outside [-R, R] is escaped: marker
//     (65535, 1) + its raw int32 as two words placed after the step's renorm words;
//   * the decoder finds each symbol by binary search over [-R, R] on the monotone
//     quantised CDF (monotonicity of all three Phi approximations over every binary32
//     input is checked exhaustively by oracle/check_monotone.c).
//
// Container (little endian):
//   header 32 B : "FGIL", u8 version=1, u8 mode, u8 segment_log2, u8 0, i64 n, i32 R,
//                 u32 n_seg, u64 payload_bytes
//   u32 offset[n_seg + 1]   byte offsets of the segments inside the payload
//   payload                 per segment: u32 state[32], u16 pad (0|1), u16 0, `pad` zero words, then the
//                           16-bit words in reading order; every segment is 4-byte aligned
#include "gmm_kernels.h"
#include "gmm_math.cuh"

namespace flashgmm {

namespace {

constexpr uint32_t kIlL = 1u << 16;        // lower bound of the normalised state interval
constexpr int kIlWarpsPerCta = 4;

__device__ __forceinline__ void il_load(const GmmParamsView& p, int64_t i, float (&mu)[4], float (&sg)[4],
                                        float (&w)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int64_t o = i * p.stride_n + k * p.stride_k;
        sg[k] = __ldg(p.scales + o);
        mu[k] = __ldg(p.means + o);
        w[k] = __ldg(p.weights + o);
    }
}

// scratch layout: segment s owns words [s*cap, (s+1)*cap); it is filled from the END.
template <int MODE>
__global__ void __launch_bounds__(32 * kIlWarpsPerCta)
gmm_ileave_encode_kernel(const int32_t* __restrict__ sym, GmmParamsView p, int64_t n, int R, int seg_log2,
                         uint16_t* __restrict__ scratch, int64_t cap, uint32_t* __restrict__ seg_words) {
    const int lane = threadIdx.x & 31;
    const int64_t seg = (int64_t)blockIdx.x * kIlWarpsPerCta + (threadIdx.x >> 5);
    const int64_t seg0 = seg << seg_log2;
    if (seg0 >= n) return;
    const int64_t seg_n = min((int64_t)1 << seg_log2, n - seg0);
    const int steps = (int)((seg_n + 31) >> 5);
    uint16_t* buf = scratch + seg * cap;
    int64_t pos = cap;                               // next free word is buf[pos - 1]
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t x = kIlL;

    for (int t = steps - 1; t >= 0; --t) {
        const int64_t i = seg0 + ((int64_t)t << 5) + lane;
        const bool valid = i < seg0 + seg_n;
        uint32_t start = 0, freq = 1;
        bool esc = false;
        int32_t v = 0;
        if (valid) {
            float mu[4], sg[4], w[4];
            il_load(p, i, mu, sg, w);
            v = __ldg(sym + i);
            const uint32_t pk = gmm_startfreq1<MODE>(v, mu, sg, w);
            // escape also when start + freq passes 65535: only parameters that are not a probability mixture (weights
            // summing above 1, NaN) produce that -- the 16-bit CDF wraps there, the interval would overlap the escape
            // slot and carry into the state. For sane parameters every row is non-decreasing (oracle/check_monotone.c),
            // which is what the decoder's search relies on; everything else travels raw.
            // ... and when the parameters are not a sane mixture at all (a scale that is not positive and finite, a negative
            // or NaN weight): only sane parameters guarantee the non-decreasing row the decoder's search assumes
            bool sane = true;
#pragma unroll
            for (int k = 0; k < 4; ++k) sane = sane && (sg[k] > 0.0f) && (sg[k] <= 3.0e38f) && (w[k] >= 0.0f) && (fabsf(mu[k]) <= 3.0e38f);
            sane = sane && (__fmul_rn(__fadd_rn(__fadd_rn(w[0], w[1]), __fadd_rn(w[2], w[3])), 65535.0f) < 65536.0f);   // no 16-bit wrap anywhere in the row
            esc = (pk == FGMM_PACK_BYPASS) || v < -R || v > R || ((pk & 0xffffu) + (pk >> 16) > 65535u) || !sane;
            start = esc ? 65535u : (pk & 0xffffu);
            freq = esc ? 1u : (pk >> 16);
        }
        // escape payloads sit after this step's renorm words in reading order -> written first
        const uint32_t me = __ballot_sync(0xffffffffu, esc);
        if (me) {
            pos -= 2 * __popc(me);
            if (esc) {
                const int r = __popc(me & lt);
                buf[pos + 2 * r] = (uint16_t)((uint32_t)v & 0xffffu);
                buf[pos + 2 * r + 1] = (uint16_t)((uint32_t)v >> 16);
            }
        }
        const bool emit = valid && x >= (freq << 16);
        const uint32_t mr = __ballot_sync(0xffffffffu, emit);
        pos -= __popc(mr);
        if (emit) {
            buf[pos + __popc(mr & lt)] = (uint16_t)(x & 0xffffu);
            x >>= 16;
        }
        if (valid) x = ((x / freq) << 16) + (x % freq) + start;
    }
    // 4-byte alignment: an odd word count gets one zero word in FRONT of the stream; a two-word
    // mini header after the states tells the decoder how many words to skip (0 or 1).
    const int pad = (int)(pos & 1);
    if (pad) { --pos; if (lane == 0) buf[pos] = 0; }
    pos -= 2;
    if (lane == 0) { buf[pos] = (uint16_t)pad; buf[pos + 1] = 0; }
    pos -= 64;
    buf[pos + 2 * lane] = (uint16_t)(x & 0xffffu);
    buf[pos + 2 * lane + 1] = (uint16_t)(x >> 16);
    if (lane == 0) seg_words[seg] = (uint32_t)(cap - pos);
}

// One CTA: exclusive scan of the segment sizes -> header + offset table at the front of `out`.
__global__ void __launch_bounds__(1024)
gmm_ileave_layout_kernel(const uint32_t* __restrict__ seg_words, uint32_t n_seg, int mode, int seg_log2, int64_t n,
                         int R, uint8_t* __restrict__ out, unsigned long long* __restrict__ total_bytes) {
    __shared__ uint32_t part[1024];
    __shared__ uint32_t carry;
    uint32_t* offs = reinterpret_cast<uint32_t*>(out + 32);
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_seg; base += 1024) {
        const uint32_t idx = base + threadIdx.x;
        const uint32_t mine = idx < n_seg ? seg_words[idx] * 2u : 0u;
        part[threadIdx.x] = mine;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) {               // Hillis-Steele inclusive scan
            uint32_t add = threadIdx.x >= d ? part[threadIdx.x - d] : 0u;
            __syncthreads();
            part[threadIdx.x] += add;
            __syncthreads();
        }
        if (idx < n_seg) offs[idx] = carry + part[threadIdx.x] - mine;
        __syncthreads();
        if (threadIdx.x == 1023) carry += part[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        offs[n_seg] = carry;
        out[0] = 'F'; out[1] = 'G'; out[2] = 'I'; out[3] = 'L';
        out[4] = 1; out[5] = (uint8_t)mode; out[6] = (uint8_t)seg_log2; out[7] = 0;
        *reinterpret_cast<int64_t*>(out + 8) = n;
        *reinterpret_cast<int32_t*>(out + 16) = R;
        *reinterpret_cast<uint32_t*>(out + 20) = n_seg;
        *reinterpret_cast<unsigned long long*>(out + 24) = carry;
        *total_bytes = 32ull + 4ull * (n_seg + 1) + carry;
    }
}

// One CTA per segment: scratch tail -> final position (32-bit copies; both sides 4-byte aligned).
__global__ void __launch_bounds__(256)
gmm_ileave_gather_kernel(const uint16_t* __restrict__ scratch, int64_t cap, const uint32_t* __restrict__ seg_words,
                         uint32_t n_seg, uint8_t* __restrict__ out) {
    const uint32_t seg = blockIdx.x;
    if (seg >= n_seg) return;
    const uint32_t words = seg_words[seg];
    const uint32_t* offs = reinterpret_cast<const uint32_t*>(out + 32);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(scratch + (int64_t)seg * cap + (cap - words));
    uint32_t* dst = reinterpret_cast<uint32_t*>(out + 32 + 4ull * (n_seg + 1) + offs[seg]);
    for (uint32_t k = threadIdx.x; k < words / 2; k += blockDim.x) dst[k] = src[k];
}

// max{ j in [0, 2R] : T[j] <= cum } by bisection on the monotone quantised CDF; returns the
// symbol v = j - R with its (start, freq). T[j] = q16(cdf(j - R - 0.5)).
template <int MODE>
__device__ __forceinline__ int32_t il_search(uint32_t cum, int R, const float (&mu)[4], const float (&sg)[4],
                                             const float (&w)[4], uint32_t* start, uint32_t* freq) {
    int lo = 0, hi = 2 * R;
    uint32_t c_lo = 0xffffffffu, c_hi = 0xffffffffu;          // T[lo], T[hi + 1] when known
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        const uint32_t c = gmm_q16<MODE>(__fsub_rn((float)(mid - R), 0.5f), mu, sg, w);
        if (c <= cum) { lo = mid; c_lo = c; } else { hi = mid - 1; c_hi = c; }
    }
    if (c_lo == 0xffffffffu) c_lo = gmm_q16<MODE>(__fsub_rn((float)(lo - R), 0.5f), mu, sg, w);
    if (c_hi == 0xffffffffu) c_hi = gmm_q16<MODE>(__fsub_rn((float)(lo + 1 - R), 0.5f), mu, sg, w);
    uint32_t f = (c_hi - c_lo) & 0xffffu;
    *start = c_lo;
    *freq = f ? f : 1u;                                        // unreachable for streams this encoder wrote
    return lo - R;
}

template <int MODE>
__global__ void __launch_bounds__(32 * kIlWarpsPerCta)
gmm_ileave_decode_kernel(const uint8_t* __restrict__ enc, GmmParamsView p, int64_t n, int R, int seg_log2,
                         uint32_t n_seg, int32_t* __restrict__ out, uint32_t* __restrict__ err) {
    const int lane = threadIdx.x & 31;
    const int64_t seg = (int64_t)blockIdx.x * kIlWarpsPerCta + (threadIdx.x >> 5);
    if (seg >= n_seg) return;
    const int64_t seg0 = seg << seg_log2;
    const int64_t seg_n = min((int64_t)1 << seg_log2, n - seg0);
    const int steps = (int)((seg_n + 31) >> 5);
    const uint32_t* offs = reinterpret_cast<const uint32_t*>(enc + 32);
    const uint32_t o0 = offs[seg], o1 = offs[seg + 1];
    const uint16_t* buf = reinterpret_cast<const uint16_t*>(enc + 32 + 4ull * (n_seg + 1) + o0);
    const int64_t end = ((int64_t)o1 - (int64_t)o0) >> 1;      // words in this segment
    if (end < 66 || buf[64] > 1) { if (lane == 0) atomicOr(err, 1u); return; }
    uint32_t x = (uint32_t)buf[2 * lane] | ((uint32_t)buf[2 * lane + 1] << 16);
    int64_t pos = 66 + buf[64];
    const uint32_t lt = (1u << lane) - 1u;
    bool bad = false;

    for (int t = 0; t < steps; ++t) {
        const int64_t i = seg0 + ((int64_t)t << 5) + lane;
        const bool valid = i < seg0 + seg_n;
        bool esc = false;
        int32_t v = 0;
        if (valid) {
            const uint32_t cum = x & 0xffffu;
            uint32_t start = 65535u, freq = 1u;
            if (cum == 65535u) {
                esc = true;
            } else {
                float mu[4], sg[4], w[4];
                il_load(p, i, mu, sg, w);
                v = il_search<MODE>(cum, R, mu, sg, w, &start, &freq);
            }
            x = freq * (x >> 16) + cum - start;
        }
        const bool need = valid && x < kIlL;
        const uint32_t mr = __ballot_sync(0xffffffffu, need);
        if (need) {
            const int64_t q = pos + __popc(mr & lt);
            if (q < end) x = (x << 16) | buf[q]; else bad = true;
        }
        pos += __popc(mr);
        const uint32_t me = __ballot_sync(0xffffffffu, esc);
        if (me) {
            if (esc) {
                const int64_t q = pos + 2 * __popc(me & lt);
                if (q + 1 < end) v = (int32_t)((uint32_t)buf[q] | ((uint32_t)buf[q + 1] << 16)); else bad = true;
            }
            pos += 2 * __popc(me);
        }
        if (valid) out[i] = v;
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(err, 2u);
}

}  // namespace

int64_t gmm_ileave_scratch_words(int seg_log2) { return 3 * ((int64_t)1 << seg_log2) + 64 + 4; }

cudaError_t launch_gmm_ileave_encode(const int32_t* sym, const GmmParamsView& p, int64_t n, int mode, int R,
                                     int seg_log2, uint16_t* scratch, uint32_t* seg_words, uint8_t* out,
                                     unsigned long long* total_bytes, cudaStream_t st) {
    const int64_t n_seg = (n + ((int64_t)1 << seg_log2) - 1) >> seg_log2;
    const int64_t cap = gmm_ileave_scratch_words(seg_log2);
    if (n_seg > 0) {
        const unsigned blocks = (unsigned)((n_seg + kIlWarpsPerCta - 1) / kIlWarpsPerCta);
        const int threads = 32 * kIlWarpsPerCta;
        switch (mode) {
        case MODE_AS: gmm_ileave_encode_kernel<MODE_AS><<<blocks, threads, 0, st>>>(sym, p, n, R, seg_log2, scratch, cap, seg_words); break;
        case MODE_LOGISTIC: gmm_ileave_encode_kernel<MODE_LOGISTIC><<<blocks, threads, 0, st>>>(sym, p, n, R, seg_log2, scratch, cap, seg_words); break;
        default: gmm_ileave_encode_kernel<MODE_POLYA><<<blocks, threads, 0, st>>>(sym, p, n, R, seg_log2, scratch, cap, seg_words);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    gmm_ileave_layout_kernel<<<1, 1024, 0, st>>>(seg_words, (uint32_t)n_seg, mode, seg_log2, n, R, out, total_bytes);
    if (n_seg > 0) gmm_ileave_gather_kernel<<<(unsigned)n_seg, 256, 0, st>>>(scratch, cap, seg_words, (uint32_t)n_seg, out);
    return cudaGetLastError();
}

cudaError_t launch_gmm_ileave_decode(const uint8_t* enc, const GmmParamsView& p, int64_t n, int mode, int R,
                                     int seg_log2, uint32_t n_seg, int32_t* out, uint32_t* err, cudaStream_t st) {
    if (n_seg == 0) return cudaSuccess;
    const unsigned blocks = (n_seg + kIlWarpsPerCta - 1) / kIlWarpsPerCta;
    const int threads = 32 * kIlWarpsPerCta;
    switch (mode) {
    case MODE_AS: gmm_ileave_decode_kernel<MODE_AS><<<blocks, threads, 0, st>>>(enc, p, n, R, seg_log2, n_seg, out, err); break;
    case MODE_LOGISTIC: gmm_ileave_decode_kernel<MODE_LOGISTIC><<<blocks, threads, 0, st>>>(enc, p, n, R, seg_log2, n_seg, out, err); break;
    default: gmm_ileave_decode_kernel<MODE_POLYA><<<blocks, threads, 0, st>>>(enc, p, n, R, seg_log2, n_seg, out, err);
    }
    return cudaGetLastError();
}

}  // namespace flashgmm
