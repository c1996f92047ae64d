// gmm_pop.cu -- the search decoder's serial half ON the GPU (sm_100a): one warp per
// stream walks the reference's 64-bit rANS state through the stream's quantised CDF
// rows (gmm_kernels.cu) and finds each symbol's interval with all 32 lanes at once.
//
// Replaces the loop of RansDecoder::decode_with_indexes_gmm<4>
// (/root/reference/compressai/cpp_exts/rans/rans_interface.cpp:793-883) together with
// Rans64DecGet / Rans64DecAdvance (third_party/ryg_rans/rans64.h:118-142) and
// Rans64DecGetBits (rans_interface.cpp:318-331), symbol for symbol:
//   cum = x & 0xffff
//   cum == 65535            -> bypass payload (count nibbles, value nibbles)
//   else search mid in [-R, R] with q(mid-.5) <= cum < q(mid+.5)      (:826-862)
//   pmf == 0 -> 1                                                     (:865-875)
//   x = pmf * (x >> 16) + cum - start; if x < 2^31: x = x << 32 | next word
//
// The recurrence is serial per stream (symbol i+1's cum needs the state after symbol
// i), so the unit of parallelism is the STREAM: a batch of images brings tens to
// hundreds of streams, one warp each, spread over the SMs. A lone warp pays the full
// latency of every dependent instruction, ~12 cycles from a predicate to the branch that
// reads it and ~25 for a taken branch, so the symbol loop is built around its chain
//   LOP (cum) -> IADD/ISETP (does cum fall in this lane's interval?) ->
//   STS (the winner posts its new state) -> LDS (every lane reads it) -> refill select
// and keeps everything else off it:
//   * the rows kernel hands each lane its interval ready-made (start and width as two 32-bit
//     words, one LDS.64, no unpacking) and rows reach shared memory through
//     a ring of 1-D TMA bulk copies (cp.async.bulk + mbarrier) issued by lane 0;
//   * every lane forms its candidate state beside the compare;
//   * symbols run in groups of 32 with NO branch inside a group: after a symbol no lane can
//     place the group runs on with garbage; the lane that places a symbol posts it into a
//     shared-memory buffer, and a slot left untouched rolls the group back to be redone symbol
//     by symbol; a placed group is written out with one coalesced store;
//   * stream words come from a 128-word shared-memory window refilled between groups.
//
// Lane l owns the intervals j = l + 32k (k < NI): hit iff T[j] <= cum < T[j+1]. On a
// monotone row at most one lane hits and that is the `mid` the reference's binary
// search ends on. Whatever the lanes cannot place -- cum outside every interval
// (malformed streams), rows that are not monotone (the rows kernel blanks them), the
// bypass marker -- takes the literal path: the reference's binary search with the CDF
// evaluated on the spot (gmm_math.cuh), uniformly in all lanes.
#include <stdlib.h>

#include "gmm_kernels.h"
#include "gmm_math.cuh"

namespace flashgmm {

namespace {

constexpr int kPopStages = 4;
constexpr int kPopStageBytes = 8704;      // 32 windowed records, 32 / 16 lane-row records
constexpr uint32_t kFull = 0xffffffffu;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// 1-D TMA: global -> shared, completion counted in bytes on `bar`.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// zero-extending 16-bit shared load straight into a 32-bit register (no mask instruction)
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}

// rANS state + word cursor of the careful path (uniform: all lanes hold the same values).
struct PopState {
    uint32_t xlo, xhi;       // x
    uint32_t ptr;            // index of the next stream word
    const uint32_t* words;
    uint32_t n_words;

    __device__ __forceinline__ uint32_t fetch(uint32_t i) const { return i < n_words ? __ldg(words + i) : 0u; }
    __device__ __forceinline__ bool overrun() const { return ptr > n_words; }   // a word past the end was taken
    // x = xn, renormalised (Rans64DecAdvance's tail, rans64.h:136-141)
    __device__ __forceinline__ void settle(uint32_t nlo, uint32_t nhi) {
        if (nhi == 0u && nlo < 0x80000000u) { xhi = nlo; xlo = fetch(ptr); ++ptr; }
        else { xhi = nhi; xlo = nlo; }
    }
    __device__ __forceinline__ void advance(uint32_t start_, uint32_t freq) {   // Rans64DecAdvance, scale_bits 16
        const uint64_t x = ((uint64_t)xhi << 32) | xlo;
        const uint64_t xn = (uint64_t)freq * (x >> 16) + (uint64_t)(xlo & 0xffffu) - (uint64_t)start_;
        settle((uint32_t)xn, (uint32_t)(xn >> 32));
    }
    __device__ __forceinline__ uint32_t get_bits4() {                           // Rans64DecGetBits, 4 bits
        const uint32_t val = xlo & 15u;
        const uint64_t x = (((uint64_t)xhi << 32) | xlo) >> 4;
        settle((uint32_t)x, (uint32_t)(x >> 32));
        return val;
    }
    __device__ __forceinline__ uint32_t bypass_raw() {                          // rans_interface.cpp:808-823
        int32_t vb = (int32_t)get_bits4(), nb = vb;
        while (vb == 15 && !overrun()) { vb = (int32_t)get_bits4(); nb += vb; }
        uint32_t raw = 0;
        for (int32_t j = 0; j < nb && !overrun(); ++j) raw |= get_bits4() << ((j * 4) & 31);
        return raw;
    }
};

__device__ __noinline__ uint32_t q16_at(int mode, int mid, const float (&mu)[4], const float (&sg)[4],
                                        const float (&w)[4]) {
    const float x = __fsub_rn((float)mid, 0.5f);     // x2 of mid == x1 of mid+1 bit for bit, |mid| < 2^22
    switch (mode) {
    case MODE_AS: return gmm_q16<MODE_AS>(x, mu, sg, w);
    case MODE_LOGISTIC: return gmm_q16<MODE_LOGISTIC>(x, mu, sg, w);
    default: return gmm_q16<MODE_POLYA>(x, mu, sg, w);
    }
}

// One symbol through the literal reference path (rans_interface.cpp:805-877), uniform in all
// lanes. State in and out BY VALUE: a reference would pin the caller's state to local memory.
struct SlowResult { PopState S; int32_t v; };
__device__ __noinline__ SlowResult pop_slow(PopState S, GmmParamsView p, int64_t i, int32_t R, int mode) {
    SlowResult o;
    const uint32_t cum = S.xlo & 0xffffu;
    if (cum == 65535u) {
        S.advance(65535u, 1u);
        o.v = (int32_t)S.bypass_raw();
    } else {
        float mu[4], sg[4], w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int64_t off = i * p.stride_n + k * p.stride_k;
            sg[k] = __ldg(p.scales + off); mu[k] = __ldg(p.means + off); w[k] = __ldg(p.weights + off);
        }
        int32_t s = -R, e = R, mid = 0;
        uint32_t c1 = 0, c2 = 0;
        while (s <= e) {
            mid = s + (e - s) / 2;
            c1 = q16_at(mode, mid, mu, sg, w);
            c2 = q16_at(mode, mid + 1, mu, sg, w);
            if (c1 <= cum && c2 > cum) break;
            else if (c1 > cum) e = mid - 1;
            else s = mid + 1;
        }
        uint32_t pmf = (c2 - c1) & 0xffffu;            // the re-evaluation at :856-862 gives c1, c2 again
        if (pmf == 0) pmf = 1;
        S.advance(c1, pmf);
        o.v = mid;
    }
    o.S = S;
    return o;
}

constexpr int kWinWords = 128;       // shared-memory window over the stream's words (4 blocks of 32)

// this lane's NI interval words (start | width << 16) of the row at shared address `a`
template <int NI>
__device__ __forceinline__ void lds_row(uint32_t a, uint32_t (&w)[NI]) {
    if constexpr (NI == 1) {
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[0]) : "r"(a));
    } else if constexpr (NI == 2) {
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "r"(a));
    } else {
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(a));
    }
}

// One interval of a lane, ready for the hit test.
struct Ival { uint32_t lo, wd; };

// Record of one symbol
//   WIN == false: 32 lanes x NI uint32 words, word l*NI + k = start | width << 16 of interval
//                 l + 32k (128*NI bytes): every interval of the row, R <= 63;
//   WIN == true : windowed record (NI == 1, 272 bytes, any R): word 0 = symbol value of lane 0's
//                 interval, then {start, width} as two uint32 per lane for the 32 consecutive
//                 intervals that hold the most mass -- one LDS.64 and no unpacking per symbol.
// SAFE == true puts a __syncwarp() between the winner's shared-memory post and the read-back (and before the
// group's output buffer is read): the ordering the fast variant gets from the LSU handling one warp's shared
// accesses in program order is then guaranteed by the memory model. The fast variant is what ships; the safe
// one (FLASHGMM_POP_SAFE=1) exists so that every toolkit bump can cross-check the two on the golden set
// (tests/test_pop_gpu.py::test_safe_and_fast_pop_variants_agree; the SASS the ordering was verified on is
// recorded in profiles/r02_pop_sass_fingerprint.txt).
template <int NI, bool WIN, bool SAFE>
__global__ void __launch_bounds__(32)
gmm_pop_rows_kernel(const PopDesc* __restrict__ descs, int mode, int n_desc, int quota, int* __restrict__ claim) {
    extern __shared__ __align__(128) uint8_t ring[];
    __shared__ uint64_t bars[kPopStages];
    __shared__ __align__(16) uint32_t bcast[4];
    __shared__ __align__(16) uint32_t win[2 * kWinWords];    // every block twice, kWinWords apart: a group reads on past the wrap
    __shared__ uint32_t obuf[32];                            // the group's symbols (lane-local part), posted by the lanes that placed them
    static_assert(!WIN || NI == 1, "the windowed record has one interval per lane");
    constexpr uint32_t kRowBytes = WIN ? 4u * kPopWindowWords : 128u * NI;
    // symbols between two looks at the "somebody could not be placed" count = rows per stage: the
    // per-group bookkeeping (~100 instructions) is paid once per 32 symbols
    constexpr int kGroup = (kPopStageBytes / kRowBytes) >= 32 ? 32 : 16;
    constexpr int kRps = kGroup;

    const int lane = threadIdx.x;
    // Which stream is this CTA's? One-warp CTAs must be SPREAD over the SMs (a warp of this kernel
    // uses about a third of a sub-partition's issue slots; six on one SM run far slower each), but
    // while the rows kernels keep every SM nearly full the block scheduler drops a launch's CTAs
    // onto the few SMs that have room at that instant (measured: 48 warps on 8 SMs). So the grid is
    // several times the number of streams and a CTA only takes a stream if its SM is below its
    // quota of them; otherwise it lingers a few microseconds (keeping the slot busy, so that the
    // CTAs behind it go elsewhere) and leaves. The grid's last n_desc CTAs take whatever is left.
    int idx = (int)blockIdx.x;
    if (claim != nullptr) {
        int got = -1;
        bool linger = false;
        if (lane == 0 && *reinterpret_cast<volatile int*>(claim) < n_desc) {
            uint32_t smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            int* cnt = claim + 1 + (smid & 1023u);
            const bool last_resort = (int)blockIdx.x >= (int)gridDim.x - n_desc;
            if (atomicAdd(cnt, 1) < quota || last_resort) {
                got = atomicAdd(claim, 1);
                if (got >= n_desc) got = -1;
            } else {
                atomicSub(cnt, 1);
                linger = true;
            }
        }
        if (linger) __nanosleep(4000);
        idx = __shfl_sync(kFull, got, 0);
        if (idx < 0) return;
    }
    const PopDesc d = descs[idx];
    if (d.n <= 0) return;
    const int R = d.R;
    const int n_stages = (d.n + kRps - 1) / kRps;
    const uint8_t* src = reinterpret_cast<const uint8_t*>(d.rows);

    auto issue = [&](int st) {                                       // lane 0 only
        const int slot = st % kPopStages;
        const int cnt = min(kRps, d.n - st * kRps);
        const uint32_t bytes = (uint32_t)cnt * kRowBytes;
        mbar_expect_tx(&bars[slot], bytes);
        bulk_g2s(ring + slot * kPopStageBytes, src + (size_t)st * kRps * kRowBytes, bytes, &bars[slot]);
    };
    if (lane == 0) {
        for (int s = 0; s < kPopStages; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int s = 0; s < min(kPopStages, n_stages); ++s) issue(s);
        bcast[0] = bcast[1] = bcast[2] = bcast[3] = 0u;
    }
    __syncwarp();

    PopState S;
    S.words = d.words; S.n_words = d.n_words;
    if (d.first) {                                                   // Rans64DecInit, rans64.h:118-124
        S.xlo = S.fetch(0); S.xhi = S.fetch(1);
        S.ptr = 2;
    } else {
        S.xlo = (uint32_t)d.state[0]; S.xhi = (uint32_t)(d.state[0] >> 32);
        S.ptr = (uint32_t)d.state[1];
    }
    const uint32_t sticky_overrun = d.first ? (d.n_words < 2 ? 1u : 0u) : (uint32_t)(d.state[1] >> 32) & 1u;

    // word window: blocks [next_blk - 4, next_blk) of 32 words are in win[]; a group needs the
    // words [ptr, ptr + kGroup] and blocks are fetched two ahead of that
    const uint32_t win_base = smem_addr(win);
    uint32_t next_blk = S.ptr >> 5;
    auto feed_window = [&]() {
        const uint32_t want = ((S.ptr + kGroup + 1) >> 5) + 2u;      // first block NOT needed yet
        if (next_blk + 4u < want) next_blk = want - 4u;              // the careful path jumped far ahead
        if (__builtin_expect(next_blk < want, 0)) {
            while (next_blk < want) {
                const uint32_t wv = S.fetch(next_blk * 32u + lane);
                win[(next_blk & 3u) * 32u + lane] = wv;
                win[(next_blk & 3u) * 32u + lane + kWinWords] = wv;
                ++next_blk;
            }
            __syncwarp();
        }
    };
    feed_window();

    const uint32_t bc = smem_addr(bcast);
    const uint32_t ob = smem_addr(obuf);
    constexpr uint32_t kUnplaced = 0x80000000u;                       // never a symbol's lane-local part (|.| < 2^11)
    int32_t* out = d.out;
    int sym0 = 0;                                                    // chunk-local index of the stage's first symbol
    int vbase[NI];
#pragma unroll
    for (int k = 0; k < NI; ++k) vbase[k] = WIN ? lane : lane + 32 * k - R;
    // this lane's intervals of the row at shared address `a` (a = row base + lane offset)
    auto load_ivals = [&](uint32_t a, Ival (&iv)[NI]) {
        if constexpr (WIN) {
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(iv[0].lo), "=r"(iv[0].wd) : "r"(a));
        } else {
            uint32_t w[NI];
            lds_row<NI>(a, w);
#pragma unroll
            for (int k = 0; k < NI; ++k) { iv[k].lo = w[k] & 0xffffu; iv[k].wd = w[k] >> 16; }
        }
    };
    constexpr uint32_t kLaneOff0 = WIN ? 16u : 0u, kLaneStride = WIN ? 8u : 4u * NI;

    for (int st = 0; st < n_stages; ++st) {
        const int slot = st % kPopStages;
        while (!mbar_try_wait(&bars[slot], (uint32_t)(st / kPopStages) & 1u)) {}
        const int cnt = min(kRps, d.n - st * kRps);
        const uint32_t stage_base = smem_addr(ring + slot * kPopStageBytes) + kLaneOff0 + kLaneStride * lane;   // this lane's words

        for (int r = 0; r < cnt; r += kGroup) {
            const int g = min(kGroup, cnt - r);
            uint32_t rp = stage_base + (uint32_t)r * kRowBytes;
            bool redo = g < kGroup;
            if (!redo) {
                // ---- kGroup symbols, no branch ----
                uint32_t xlo = S.xlo, xhi = S.xhi;
                Ival cur[NI], nxt[NI];
                load_ivals(rp, cur);
                // next stream word: the window holds every block twice, so the cursor just walks on
                const uint32_t w0 = win_base + ((S.ptr << 2) & (4u * kWinWords - 4u));
                uint32_t waddr = w0;
                uint32_t wnext = lds_u32(waddr);
                obuf[lane] = kUnplaced;
                __syncwarp();
#pragma unroll
                for (int i = 0; i < kGroup; ++i) {
                    rp += kRowBytes;
                    if (i + 1 < kGroup) load_ivals(rp, nxt);         // next row: independent of the chain
                    const uint32_t cum = xlo & 0xffffu;
                    const uint32_t xs_lo = __funnelshift_r(xlo, xhi, 16), xs_hi = xhi >> 16;
                    uint32_t clo = 0, chi = 0;
                    uint64_t c64 = 0;                                // NI == 1: the candidate as one register pair
                    int v = lane;                                    // windowed: the record's word 0 is added when the group is written out
                    bool hit = false;
#pragma unroll
                    for (int k = 0; k < NI; ++k) {
                        const uint32_t lo = cur[k].lo, wd = cur[k].wd;
                        const uint32_t dk = cum - lo;
                        const bool hk = dk < wd;         // never true for cum == 65535: start + width <= 65535
                        if constexpr (NI == 1) {
                            // the candidate is formed BESIDE the compare (the empty volatile asm pins it
                            // there: the compiler would otherwise sink both multiplies under the hit
                            // predicate, behind the compare)
                            c64 = (uint64_t)wd * xs_lo + dk + ((uint64_t)(wd * xs_hi) << 32);
                            asm volatile("" : "+l"(c64));
                            if (!WIN) v = vbase[0];
                        } else {
                            const uint64_t ck = (uint64_t)wd * xs_lo + dk + ((uint64_t)(wd * xs_hi) << 32);
                            if (hk) { clo = (uint32_t)ck; chi = (uint32_t)(ck >> 32); v = vbase[k]; }
                        }
                        hit |= hk;
                    }
                    // the winner posts its state, everyone reads it back (the LSU keeps one warp's shared
                    // accesses in program order). If no lane can place cum the slot keeps its old content
                    // and the group runs on with garbage: the symbol's obuf slot stays kUnplaced and the
                    // group is redone below. (Measured against REDUX.OR x2, a 128-bit slot that also
                    // carries the symbol, and posting the row word with the multiply after the read-back:
                    // this is the fastest of the four by 5-12 %, profiles/r01_s12_pop_variants.txt.)
                    uint32_t rlo, rhi;
                    if constexpr (NI == 1) {
                        if (hit) asm volatile("st.shared.u64 [%0], %1;" ::"r"(bc), "l"(c64) : "memory");
                    } else {
                        if (hit) asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(bc), "r"(clo), "r"(chi) : "memory");
                    }
                    if constexpr (SAFE) __syncwarp();
                    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(rlo), "=r"(rhi) : "r"(bc) : "memory");
                    if constexpr (SAFE) __syncwarp();                // nobody posts the next state before everyone has read this one
                    if (hit) asm volatile("st.shared.u32 [%0], %1;" ::"r"(ob + 4u * (uint32_t)i), "r"(v) : "memory");
                    const bool need = rhi == 0u && rlo < 0x80000000u;                // rans64.h:136-141
                    xhi = need ? rlo : rhi;
                    xlo = need ? wnext : rlo;
                    waddr += need ? 4u : 0u;                         // unconditional reload: two instructions instead of
                    wnext = lds_u32(waddr);                          // five predicated ones, and it is needed a symbol later at the earliest
#pragma unroll
                    for (int k = 0; k < NI; ++k) cur[k] = nxt[k];
                }
                __syncwarp();
                const uint32_t mine = obuf[lane & (kGroup - 1)];     // lane i answers for symbol i of the group
                if (__builtin_expect(__all_sync(kFull, mine != kUnplaced), 1)) {     // every symbol placed
                    S.xlo = xlo; S.xhi = xhi; S.ptr += (waddr - w0) >> 2;
                    if (lane < kGroup) {
                        const uint32_t row0 = stage_base - kLaneOff0 - kLaneStride * lane + (uint32_t)(r + lane) * kRowBytes;
                        const int v0 = WIN ? (int)lds_u32(row0) : 0;
                        out[r + lane] = (int)mine + v0;              // one coalesced store per group
                    }
                } else {
                    redo = true;
                    rp = stage_base + (uint32_t)r * kRowBytes;
                }
                __syncwarp();                                        // obuf is rewritten by the next group
            }
            if (redo) {
                // ---- symbol by symbol from the group's first state: literal path where needed ----
                for (int i = 0; i < g; ++i, rp += kRowBytes) {
                    const uint32_t cum = S.xlo & 0xffffu;
                    const uint32_t xs_lo = __funnelshift_r(S.xlo, S.xhi, 16), xs_hi = S.xhi >> 16;
                    uint32_t clo = 0, chi = 0;
                    int v = 0;
                    bool hit = false;
                    Ival cur[NI];
                    load_ivals(rp, cur);
                    const int v0 = WIN ? (int)lds_u32(rp - kLaneOff0 - kLaneStride * lane) : 0;
#pragma unroll
                    for (int k = 0; k < NI; ++k) {
                        const uint32_t lo = cur[k].lo, wd = cur[k].wd;
                        const uint32_t dk = cum - lo;
                        const uint64_t ck = (uint64_t)wd * xs_lo + dk + ((uint64_t)(wd * xs_hi) << 32);
                        if (dk < wd) { clo = (uint32_t)ck; chi = (uint32_t)(ck >> 32); v = vbase[k] + v0; hit = true; }
                    }
                    const uint32_t rlo = __reduce_or_sync(kFull, clo);
                    const uint32_t rhi = __reduce_or_sync(kFull, chi);
                    if (!__any_sync(kFull, hit)) {
                        const SlowResult o = pop_slow(S, d.p, sym0 + r + i, R, mode);
                        S = o.S;
                        if (lane == 0) out[r + i] = o.v;
                    } else {
                        if (hit) out[r + i] = v;
                        S.settle(rlo, rhi);
                    }
                }
            }
            feed_window();
        }
        out += cnt;
        sym0 += cnt;
        __syncwarp();
        if (lane == 0 && st + kPopStages < n_stages) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(st + kPopStages);
        }
    }
    if (lane == 0) {
        d.state[0] = ((uint64_t)S.xhi << 32) | S.xlo;
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));           // bits 40+: where the stream's last launch ran (trace only)
        d.state[1] = ((uint64_t)smid << 40) | ((uint64_t)(sticky_overrun | (S.overrun() ? 1u : 0u)) << 32) | S.ptr;
    }
}

}  // namespace

cudaError_t launch_gmm_pop_rows(const PopDesc* descs_d, int n_desc, int lane_slots, int mode, int* claim, int quota,
                                cudaStream_t st) {
    if (n_desc <= 0) return cudaSuccess;
    const size_t smem = (size_t)kPopStages * kPopStageBytes;
    const int grid = claim ? 4 * n_desc : n_desc;             // see the kernel's first lines
    static const bool safe = [] { const char* v = getenv("FLASHGMM_POP_SAFE"); return v && *v && *v != '0'; }();
#define FGMM_POP_LAUNCH(NI, WIN)                                                                                     \
    do {                                                                                                             \
        if (safe) gmm_pop_rows_kernel<NI, WIN, true><<<grid, 32, smem, st>>>(descs_d, mode, n_desc, quota, claim);   \
        else gmm_pop_rows_kernel<NI, WIN, false><<<grid, 32, smem, st>>>(descs_d, mode, n_desc, quota, claim);       \
    } while (0)
    switch (lane_slots) {
    case kPopWindowed: FGMM_POP_LAUNCH(1, true); break;
    case 32: FGMM_POP_LAUNCH(1, false); break;
    case 64: FGMM_POP_LAUNCH(2, false); break;
    case 128: FGMM_POP_LAUNCH(4, false); break;
    default: return cudaErrorInvalidValue;
    }
#undef FGMM_POP_LAUNCH
    return cudaGetLastError();
}

}  // namespace flashgmm
