// gmm_push.cu -- the encoder's serial half ON the GPU (sm_100a): one warp per stream
// walks the packed (start, freq) records of the startfreq kernel backwards through the
// reference's 64-bit rANS state and writes the stream's 32-bit words.
//
// Replaces BufferedRansEncoder::flush (/root/reference/compressai/cpp_exts/rans/
// rans_interface.cpp:557-585) with Rans64EncPut / Rans64EncFlush
// (third_party/ryg_rans/rans64.h:59-105), Rans64EncPutBits (rans_interface.cpp:295-316) and
// the bypass payload of encode_with_indexes_gmm (rans_interface.cpp:512-552):
//   symbols last to first:
//     x >= freq << 47      -> emit the low 32 bits, x >>= 32
//     x = ((x / freq) << 16) + x % freq + start
//   pmf == 0 (record 0x0000ffff): value nibbles high to low, the nibble count, then the
//     marker put(65535, 1) -- the reverse of the order the decoder reads them in
//   flush: the two halves of x
// x / freq is an exact reciprocal multiply (umul64hi + shift; table built on the host by
// rans_host.hpp::rcp_table, the same one the host encoder uses), so the words are the
// reference's bit for bit.
//
// Like the decoder's pop kernel this is one serial chain per stream, all lanes carrying
// the same state; the lanes earn their keep by fetching: 32 records and their 32
// reciprocal entries are loaded in one coalesced read and one 32-wide gather (L2-resident
// 1 MB table) a group ahead, parked in shared memory, and read back one 16-byte broadcast
// per symbol. Used by the batch encoder when the host has fewer cores than the batch
// needs (several ranks on one box); on a free host the worker-thread encoder is faster
// per stream.
#include "gmm_kernels.h"

namespace flashgmm {

namespace {

constexpr uint32_t kFull = 0xffffffffu;

// rANS state + output cursor. Passed and returned BY VALUE everywhere: taking its address (a
// member function called out of line) would pin it to local memory, and the symbol chain would
// pay an LDL for the buffer pointer at every renormalisation.
struct PushState {
    uint32_t xlo, xhi;
    uint32_t pos;            // words are written at buf[--pos]; counts down from cap
    uint32_t need;           // words the stream needs in total (keeps counting after an overflow)
};

__device__ __forceinline__ void emit(PushState& S, uint32_t* buf, bool lane0, uint32_t w) {
    ++S.need;
    if (S.pos > 0u) { --S.pos; if (lane0) buf[S.pos] = w; }
}
__device__ __forceinline__ void put_bits4(PushState& S, uint32_t* buf, bool lane0, uint32_t val) {   // Rans64EncPutBits, 4 bits
    if (S.xhi >= (1u << 27)) { emit(S, buf, lane0, S.xlo); S.xlo = S.xhi; S.xhi = 0u; }            // x >= 2^59
    S.xhi = (S.xhi << 4) | (S.xlo >> 28);
    S.xlo = (S.xlo << 4) | val;
}
// rans_interface.cpp:524-552, reversed: value nibbles high to low, the count, then put(65535, 1)
__device__ __noinline__ PushState put_bypass(PushState S, uint32_t* buf, bool lane0, int32_t v) {
    const uint32_t raw = (uint32_t)v;
    int nb = 0;
    for (uint32_t t = raw; t != 0u; t >>= 4) ++nb;
    for (int j = nb - 1; j >= 0; --j) put_bits4(S, buf, lane0, (raw >> (4 * j)) & 15u);
    put_bits4(S, buf, lane0, (uint32_t)nb);
    if (S.xhi >= (1u << 15)) { emit(S, buf, lane0, S.xlo); S.xlo = S.xhi; S.xhi = 0u; }            // x >= 1 << 47
    S.xhi = (S.xhi << 16) | (S.xlo >> 16);
    S.xlo = (S.xlo << 16) | 65535u;
    return S;
}

// One symbol of Rans64EncPut (scale_bits 16) on registers, no branch: renormalise when
// x >= freq << 47, then x = x + start + bias + (x / freq) * (65536 - freq).
__device__ __forceinline__ void put_symbol(PushState& S, uint32_t* buf, bool lane0, const uint4 e) {
    const uint32_t freq = e.z >> 16, start = e.z & 0xffffu;
    const bool spill = S.xhi >= (freq << 15);
    const bool wr = spill && S.pos > 0u;
    S.need += spill ? 1u : 0u;
    S.pos -= wr ? 1u : 0u;
    if (wr && lane0) buf[S.pos] = S.xlo;
    const uint32_t lo = spill ? S.xhi : S.xlo, hi = spill ? 0u : S.xhi;
    const unsigned long long x = ((unsigned long long)hi << 32) | lo;
    const unsigned long long rcp = ((unsigned long long)e.y << 32) | e.x;
    const unsigned long long q = __umul64hi(x, rcp) >> (e.w & 0xffffu);
    const unsigned long long xn = x + start + (e.w >> 16) + q * (unsigned long long)(65536u - freq);
    S.xlo = (uint32_t)xn; S.xhi = (uint32_t)(xn >> 32);
}

__global__ void __launch_bounds__(32)
gmm_push_kernel(const PushDesc* __restrict__ descs, const PushRcp* __restrict__ rcp_tab,
                uint32_t* __restrict__ out, unsigned long long* __restrict__ out_used, unsigned int* __restrict__ started) {
    __shared__ __align__(16) uint4 park[2][32];      // {rcp.lo, rcp.hi, packed, shift | bias << 16}

    const PushDesc d = descs[blockIdx.x];
    const int lane = threadIdx.x;
    if (lane == 0 && started != nullptr) atomicAdd(started, 1u);      // "this warp is resident": see gmm_wait_started_kernel
    unsigned long long t_begin;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
    PushState S;
    S.xlo = 0x80000000u; S.xhi = 0u;                      // RANS64_L = 2^31
    S.pos = d.cap_words; S.need = 0u;
    uint32_t* const buf = d.buf;
    const bool lane0 = lane == 0;

    const int n = d.n;
    const int n_groups = (n + 31) >> 5;
    // group g covers symbols [n - 32(g+1), n - 32g) (clipped at 0); lane l holds symbol n - 1 - 32g - l.
    // Two loads feed a group: the record, then the reciprocal entry it indexes. Both are issued
    // well ahead and land in registers while the warp is busy with the chain: records two groups
    // ahead, reciprocal entries one group ahead; they are parked in shared memory at the END of the
    // group in between, so the chain never waits for global memory.
    auto load_pk = [&](int g) -> uint32_t {
        const int i = n - 1 - 32 * g - lane;
        return (g < n_groups && i >= 0) ? __ldg(d.packed + i) : 0x00010000u;       // inert filler, never pushed
    };
    auto load_rcp = [&](uint32_t pk) -> uint4 {
        const PushRcp r = rcp_tab[pk >> 16];
        return make_uint4((uint32_t)r.rcp, (uint32_t)(r.rcp >> 32), pk, (uint32_t)r.shift | ((uint32_t)r.bias << 16));
    };
    uint32_t pk_next = load_pk(1);
    park[0][lane] = load_rcp(load_pk(0));
    __syncwarp();

    for (int g = 0; g < n_groups; ++g) {
        const uint4 e_next = load_rcp(pk_next);              // group g+1's entries: consumed at the end of this group
        pk_next = load_pk(g + 2);                            // group g+2's records: consumed one group from now
        const int cnt = min(32, n - 32 * g);
        const uint4* slot = park[g & 1];
        // does the group hold a bypass record? (known from the records, not from the state: the
        // common group then runs without a single branch)
        const bool plain = __all_sync(kFull, lane >= cnt || (slot[lane].z >> 16) != 0u);
        if (plain && cnt == 32) {
            uint4 e = slot[0];
#pragma unroll
            for (int l = 0; l < 32; ++l) {
                const uint4 e_cur = e;                       // one 16-byte broadcast, fetched a symbol ahead
                if (l + 1 < 32) e = slot[l + 1];
                put_symbol(S, buf, lane0, e_cur);
            }
        } else {
            for (int l = 0; l < cnt; ++l) {
                const uint4 e = slot[l];
                if ((e.z >> 16) == 0u) S = put_bypass(S, buf, lane0, __ldg(d.sym + (n - 1 - 32 * g - l)));
                else put_symbol(S, buf, lane0, e);
            }
        }
        park[(g + 1) & 1][lane] = e_next;
        __syncwarp();
    }
    // Rans64EncFlush: x.hi then x.lo walking backwards -> [x.lo, x.hi] lead the stream
    emit(S, buf, lane0, S.xhi);
    emit(S, buf, lane0, S.xlo);

    // hand the words over: claim a slice of `out`, copy the tail of the scratch buffer into it
    const uint32_t used = d.cap_words - S.pos;
    const bool overflow = S.need > d.cap_words;
    unsigned long long off = 0;
    if (lane == 0) off = atomicAdd(out_used, (unsigned long long)used);
    off = __shfl_sync(kFull, off, 0);
    __syncwarp();
    for (uint32_t k = lane; k < used; k += 32) out[off + k] = buf[S.pos + k];
    if (lane == 0) {
        d.result[0] = S.need;
        d.result[1] = (uint32_t)off;
        d.result[2] = overflow ? 1u : 0u;
        // the kernel times itself (ns): an event recorded behind it would be one more dependent
        // entry in the hardware queue
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        d.result[4] = (uint32_t)t_begin; d.result[5] = (uint32_t)(t_begin >> 32);
        d.result[6] = (uint32_t)t_end; d.result[7] = (uint32_t)(t_end >> 32);
    }
}

// One-warp CTAs must be SPREAD over the SMs: a warp of this kernel uses about a third of a
// sub-partition's issue slots, and eight of them on one sub-partition run 2-5x slower each. The
// block scheduler spreads them when it places them on an idle GPU, but packs them onto whichever
// SMs free up first when a wide kernel that fills every SM (the startfreq kernels: 16 CTAs x 128
// threads) is draining at that moment -- measured: the push kernel then takes 10.2 ms instead of
// 2.2 ms. So the stream that issues such wide kernels beside the push kernel first waits, ON THE
// DEVICE, until every push warp has reported in. Bounded: gives up after `timeout_ns`.
// (Round 1 and most of round 2. Superseded by the shared-memory reservation below, which needs no inter-kernel
// progress assumption; kept behind FLASHGMM_PUSH_SPREAD_KB=0.)
__global__ void gmm_wait_started_kernel(const unsigned int* started, unsigned int target, unsigned long long timeout_ns) {
    if (threadIdx.x != 0) return;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*reinterpret_cast<const volatile unsigned int*>(started) < target) {
        __nanosleep(500);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) break;
    }
}

}  // namespace

// Spread BY CONSTRUCTION (default): every push CTA reserves FLASHGMM_PUSH_SPREAD_KB (default 40) KB of dynamic shared
// memory it never touches, so that at most floor(227 / 40) = 5 of them fit on one SM wherever the block scheduler puts
// them (740 resident on the GPU), and nothing has to wait for anything. Measured with 6 batches in flight and the device
// push: 2524 Msym/s with the wait kernel below, 3100-3130 with 28 / 40 / 56 KB, 2970-3000 with 75 / 100 KB.
// FLASHGMM_PUSH_SPREAD_KB=0 restores the wait kernel.
static int push_spread_bytes() {
    static const int v = [] {
        const char* e = getenv("FLASHGMM_PUSH_SPREAD_KB");
        long kb = e && *e ? strtol(e, nullptr, 10) : 40;
        if (kb < 0) kb = 0;
        if (kb > 200) kb = 200;
        return (int)(kb * 1024);
    }();
    return v;
}

cudaError_t launch_gmm_push(const PushDesc* descs_d, int n_desc, const PushRcp* rcp_d, uint32_t* out,
                            unsigned long long* out_used, unsigned int* started, cudaStream_t st) {
    if (n_desc <= 0) return cudaSuccess;
    const int spread = push_spread_bytes();
    if (spread > 0) {
        cudaError_t e = cudaFuncSetAttribute(gmm_push_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, spread);
        if (e != cudaSuccess) return e;
    }
    gmm_push_kernel<<<n_desc, 32, (size_t)spread, st>>>(descs_d, rcp_d, out, out_used, started);
    return cudaGetLastError();
}

cudaError_t launch_gmm_wait_started(const unsigned int* started, unsigned int target, cudaStream_t st) {
    if (push_spread_bytes() > 0) return cudaSuccess;          // the push CTAs cannot be packed: nothing to wait for
    gmm_wait_started_kernel<<<1, 32, 0, st>>>(started, target, 3000000ull);      // 3 ms at most
    return cudaGetLastError();
}

}  // namespace flashgmm
