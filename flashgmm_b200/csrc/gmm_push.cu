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

struct PushState {
    uint32_t xlo, xhi;
    uint32_t* buf;           // words are written at buf[--pos]
    uint32_t pos;            // next free slot is pos - 1; counts down from cap
    uint32_t need;           // words the stream needs in total (keeps counting after an overflow)
    bool lane0;

    __device__ __forceinline__ void emit(uint32_t w) {
        ++need;
        if (pos > 0u) { --pos; if (lane0) buf[pos] = w; }
    }
    __device__ __forceinline__ void put_bits4(uint32_t val) {              // Rans64EncPutBits, 4 bits
        if (xhi >= (1u << 27)) { emit(xlo); xlo = xhi; xhi = 0u; }        // x >= 2^59
        xhi = (xhi << 4) | (xlo >> 28);
        xlo = (xlo << 4) | val;
    }
    __device__ __forceinline__ void put_marker() {                         // put(65535, 1)
        if (xhi >= (1u << 15)) { emit(xlo); xlo = xhi; xhi = 0u; }        // x >= 1 << 47
        xhi = (xhi << 16) | (xlo >> 16);
        xlo = (xlo << 16) | 65535u;
    }
    __device__ __noinline__ void put_bypass(int32_t v) {                // rans_interface.cpp:524-552, reversed
        const uint32_t raw = (uint32_t)v;
        int nb = 0;
        for (uint32_t t = raw; t != 0u; t >>= 4) ++nb;
        for (int j = nb - 1; j >= 0; --j) put_bits4((raw >> (4 * j)) & 15u);
        put_bits4((uint32_t)nb);
        put_marker();
    }
};

__global__ void __launch_bounds__(32)
gmm_push_kernel(const PushDesc* __restrict__ descs, const PushRcp* __restrict__ rcp_tab,
                uint32_t* __restrict__ out, unsigned long long* __restrict__ out_used) {
    __shared__ __align__(16) uint4 park[2][32];      // {rcp.lo, rcp.hi, packed, shift | bias << 16}

    const PushDesc d = descs[blockIdx.x];
    const int lane = threadIdx.x;
    unsigned long long t_begin;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
    PushState S;
    S.xlo = 0x80000000u; S.xhi = 0u;                      // RANS64_L = 2^31
    S.buf = d.buf; S.pos = d.cap_words; S.need = 0u; S.lane0 = lane == 0;

    const int n = d.n;
    const int n_groups = (n + 31) >> 5;
    // group g covers symbols [n - 32(g+1), n - 32g) (clipped at 0); lane l holds symbol hi - 1 - l
    auto fetch = [&](int g, int slot) {
        const int i = n - 1 - 32 * g - lane;
        uint4 e = make_uint4(0u, 0u, 0x00010000u, 0u);     // inert: never read (i < 0 is past the loop bound)
        if (i >= 0) {
            const uint32_t pk = __ldg(d.packed + i);
            const PushRcp r = rcp_tab[pk >> 16];
            e = make_uint4((uint32_t)r.rcp, (uint32_t)(r.rcp >> 32), pk, (uint32_t)r.shift | ((uint32_t)r.bias << 16));
        }
        park[slot][lane] = e;
    };
    if (n_groups > 0) fetch(0, 0);
    __syncwarp();

    for (int g = 0; g < n_groups; ++g) {
        if (g + 1 < n_groups) fetch(g + 1, (g + 1) & 1);
        const int cnt = min(32, n - 32 * g);
        const uint4* slot = park[g & 1];
        // does the group hold a bypass record? (known from the records, not from the state: the
        // common group then runs without a single branch)
        const bool plain = __all_sync(kFull, lane >= cnt || (slot[lane].z >> 16) != 0u);
        if (plain && cnt == 32) {
            uint32_t xlo = S.xlo, xhi = S.xhi, pos = S.pos, need = S.need;
#pragma unroll
            for (int l = 0; l < 32; ++l) {
                const uint4 e = slot[l];                    // one 16-byte broadcast
                const uint32_t freq = e.z >> 16, start = e.z & 0xffffu;
                // Rans64EncPut, scale_bits 16: renormalise when x >= freq << 47
                const bool spill = xhi >= (freq << 15);
                const bool wr = spill && pos > 0u;
                need += spill ? 1u : 0u;
                pos -= wr ? 1u : 0u;
                if (wr && S.lane0) S.buf[pos] = xlo;
                const uint32_t lo = spill ? xhi : xlo, hi = spill ? 0u : xhi;
                const unsigned long long x = ((unsigned long long)hi << 32) | lo;
                const unsigned long long rcp = ((unsigned long long)e.y << 32) | e.x;
                const unsigned long long q = __umul64hi(x, rcp) >> (e.w & 0xffffu);
                const unsigned long long xn = x + start + (e.w >> 16) + q * (unsigned long long)(65536u - freq);
                xlo = (uint32_t)xn; xhi = (uint32_t)(xn >> 32);
            }
            S.xlo = xlo; S.xhi = xhi; S.pos = pos; S.need = need;
        } else {
            for (int l = 0; l < cnt; ++l) {
                const uint4 e = slot[l];
                const uint32_t freq = e.z >> 16, start = e.z & 0xffffu;
                if (freq == 0u) {
                    S.put_bypass(__ldg(d.sym + (n - 1 - 32 * g - l)));
                    continue;
                }
                if (S.xhi >= (freq << 15)) { S.emit(S.xlo); S.xlo = S.xhi; S.xhi = 0u; }
                const unsigned long long x = ((unsigned long long)S.xhi << 32) | S.xlo;
                const unsigned long long rcp = ((unsigned long long)e.y << 32) | e.x;
                const unsigned long long q = __umul64hi(x, rcp) >> (e.w & 0xffffu);
                const unsigned long long xn = x + start + (e.w >> 16) + q * (unsigned long long)(65536u - freq);
                S.xlo = (uint32_t)xn; S.xhi = (uint32_t)(xn >> 32);
            }
        }
        __syncwarp();
    }
    // Rans64EncFlush: x.hi then x.lo walking backwards -> [x.lo, x.hi] lead the stream
    S.emit(S.xhi);
    S.emit(S.xlo);

    // hand the words over: claim a slice of `out`, copy the tail of the scratch buffer into it
    const uint32_t used = d.cap_words - S.pos;
    const bool overflow = S.need > d.cap_words;
    unsigned long long off = 0;
    if (lane == 0) off = atomicAdd(out_used, (unsigned long long)used);
    off = __shfl_sync(kFull, off, 0);
    __syncwarp();
    for (uint32_t k = lane; k < used; k += 32) out[off + k] = d.buf[S.pos + k];
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

}  // namespace

cudaError_t launch_gmm_push(const PushDesc* descs_d, int n_desc, const PushRcp* rcp_d, uint32_t* out,
                            unsigned long long* out_used, cudaStream_t st) {
    if (n_desc <= 0) return cudaSuccess;
    gmm_push_kernel<<<n_desc, 32, 0, st>>>(descs_d, rcp_d, out, out_used);
    return cudaGetLastError();
}

}  // namespace flashgmm
