// rans_host.hpp -- the serial 64-bit rANS state push/pop, host side.
//
// Bit-identical restatement of (paths under /root/reference):
//   third_party/ryg_rans/rans64.h:59-142           Rans64Enc{Init,Put,Flush}, Rans64Dec{Init,Get,Advance}
//   compressai/cpp_exts/rans/rans_interface.cpp:295-331   Rans64EncPutBits / Rans64DecGetBits
//   compressai/cpp_exts/rans/rans_interface.cpp:524-585   GMM bypass payload + BufferedRansEncoder::flush
// The encoder replaces the reference's `x / freq`, `x % freq` by an exact
// 64x64->128 reciprocal multiply (same idea as rans64.h:152-278, Rans64EncSymbol);
// results are identical for every x < 2^63, which tests/test_host_rans.py checks
// against plain division.
//
// The GPU produces one packed word per symbol (start | freq<<16, freq==0 = bypass);
// only this serial recurrence stays on the CPU, as BASELINE.json's north_star (4) says.
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>
#if defined(__AVX2__)
#include <immintrin.h>
#endif
#include <vector>

namespace flashgmm {

constexpr uint32_t kPrecision = 16;
constexpr uint32_t kMaxCdf = 65535;
constexpr uint32_t kBypassBits = 4;
constexpr uint32_t kBypassMax = 15;
constexpr uint64_t kRansL = 1ull << 31;
constexpr uint32_t kPackBypass = 0x0000ffffu;

struct RcpEntry {
    uint64_t rcp;
    uint32_t cmpl;       // 65536 - freq
    uint16_t shift;
    uint16_t bias;       // 65535 for freq==1 else 0
};

inline const RcpEntry* rcp_table() {
    static RcpEntry* tab = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        RcpEntry* t = static_cast<RcpEntry*>(aligned_alloc(64, sizeof(RcpEntry) * 65536));
        t[0] = RcpEntry{0, 65536, 0, 0};
        t[1] = RcpEntry{~0ull, 65535, 0, 65535};
        for (uint32_t f = 2; f < 65536; ++f) {
            uint32_t shift = 0;
            while (f > (1u << shift)) ++shift;
            unsigned __int128 num = ((unsigned __int128)1 << (shift + 63)) + (f - 1);
            t[f].rcp = (uint64_t)(num / f);
            t[f].cmpl = 65536 - f;
            t[f].shift = (uint16_t)(shift - 1);
            t[f].bias = 0;
        }
        tab = t;
    });
    return tab;
}

// Backward-growing word buffer: rANS writes from the end towards the start.
class BackBuffer {
public:
    explicit BackBuffer(size_t words_hint = 1024) { reset(words_hint); }
    ~BackBuffer() { free(buf_); }
    BackBuffer(const BackBuffer&) = delete;
    BackBuffer& operator=(const BackBuffer&) = delete;
    void reset(size_t words_hint) {
        if (words_hint < 64) words_hint = 64;
        if (words_hint > cap_) { free(buf_); buf_ = static_cast<uint32_t*>(malloc(words_hint * 4)); cap_ = words_hint; }
        ptr_ = buf_ + cap_;
    }
    // make sure `need` more words can be written before ptr_
    inline void reserve(size_t need) {
        if ((size_t)(ptr_ - buf_) >= need) return;
        size_t used = (buf_ + cap_) - ptr_;
        size_t ncap = cap_ * 2 + need;
        uint32_t* nb = static_cast<uint32_t*>(malloc(ncap * 4));
        memcpy(nb + ncap - used, ptr_, used * 4);
        free(buf_);
        buf_ = nb; cap_ = ncap; ptr_ = nb + ncap - used;
    }
    uint32_t*& ptr() { return ptr_; }
    size_t bytes() const { return ((buf_ + cap_) - ptr_) * 4; }
    const uint8_t* data() const { return reinterpret_cast<const uint8_t*>(ptr_); }
private:
    uint32_t* buf_ = nullptr;
    uint32_t* ptr_ = nullptr;
    size_t cap_ = 0;
};

struct RansEncoder64 {
    uint64_t x = kRansL;
    BackBuffer* out = nullptr;
    const RcpEntry* rcp = nullptr;

    void init(BackBuffer* b) { x = kRansL; out = b; rcp = rcp_table(); }

    inline void put(uint32_t start, uint32_t freq) {          // Rans64EncPut, scale_bits=16
        uint64_t xx = x;
        if (xx >= ((uint64_t)freq << 47)) { *--out->ptr() = (uint32_t)xx; xx >>= 32; }
        const RcpEntry& e = rcp[freq];
        uint64_t q = (uint64_t)(((unsigned __int128)xx * e.rcp) >> 64) >> e.shift;
        x = xx + start + e.bias + q * e.cmpl;
    }
    inline void put_bits4(uint32_t val) {                     // Rans64EncPutBits, nbits=4
        uint64_t xx = x;
        if (xx >= (1ull << 59)) { *--out->ptr() = (uint32_t)xx; xx >>= 32; }   // ((L>>16)<<32) * (1<<12)
        x = (xx << kBypassBits) | val;
    }
    // the reference's bypass payload for int32 value v, walked in reverse emission order
    inline void put_bypass_reverse(int32_t v) {
        uint32_t raw = (uint32_t)v;
        int nb = 0;
        for (uint32_t t = raw; t != 0; t >>= 4) ++nb;         // 0..8 nibbles
        for (int j = nb - 1; j >= 0; --j) put_bits4((raw >> (4 * j)) & kBypassMax);
        put_bits4((uint32_t)nb);                              // nb < 15: a single count nibble
        put(kMaxCdf, 1);
    }
    // symbols [0, cnt) of one chunk, last first. `room` words must be reserved by the caller
    // (<= 11 words per symbol). sym_at(i) is asked for the value of a bypassed symbol only.
    template <class SymAt>
    inline void push_chunk_reverse_with(const uint32_t* packed, int64_t cnt, SymAt sym_at) {
        constexpr int64_t kBlk = 4096;
        for (int64_t hi = cnt; hi > 0; hi -= kBlk) {
            int64_t lo = hi > kBlk ? hi - kBlk : 0;
            out->reserve((size_t)(hi - lo) * 11 + 4);
            for (int64_t i = hi - 1; i >= lo; --i) {
                // the records sit in page-locked memory the GPU has just written and the walk is
                // backwards: ask for the lines well ahead
                if ((i & 15) == 15) __builtin_prefetch(packed + i - 1024, 0, 0);
                uint32_t pk = packed[i];
                uint32_t freq = pk >> 16;
                if (__builtin_expect(freq != 0, 1)) put(pk & 0xffffu, freq);
                else put_bypass_reverse(sym_at(i));
            }
        }
    }
    inline void push_chunk_reverse(const uint32_t* packed, const int32_t* sym, int64_t cnt) {
        push_chunk_reverse_with(packed, cnt, [sym](int64_t i) { return sym[i]; });
    }
    inline void flush() {                                      // Rans64EncFlush
        out->reserve(2);
        uint32_t*& p = out->ptr();
        p -= 2;
        p[0] = (uint32_t)x;
        p[1] = (uint32_t)(x >> 32);
    }
};

struct RansDecoder64 {
    uint64_t x = 0;
    const uint32_t* p = nullptr;
    const uint32_t* end = nullptr;
    bool overrun = false;

    inline uint32_t word() {
        if (__builtin_expect(p >= end, 0)) { overrun = true; return 0; }
        return *p++;
    }
    void init(const uint8_t* enc, size_t len) {               // Rans64DecInit
        p = reinterpret_cast<const uint32_t*>(enc);
        end = p + len / 4;
        overrun = false;
        uint64_t lo = word(), hi = word();
        x = lo | (hi << 32);
    }
    inline uint32_t get() const { return (uint32_t)(x & 0xffffu); }            // Rans64DecGet
    inline void advance(uint32_t start, uint32_t freq) {                        // Rans64DecAdvance
        uint64_t xx = (uint64_t)freq * (x >> kPrecision) + (x & 0xffffu) - start;
        if (xx < kRansL) xx = (xx << 32) | word();
        x = xx;
    }
    inline uint32_t get_bits4() {                                               // Rans64DecGetBits
        uint32_t val = (uint32_t)(x & kBypassMax);
        uint64_t xx = x >> kBypassBits;
        if (xx < kRansL) xx = (xx << 32) | word();
        x = xx;
        return val;
    }
    inline uint32_t bypass_raw() {                             // rans_interface.cpp:808-823 payload
        int32_t vb = (int32_t)get_bits4();
        int32_t nb = vb;
        while (vb == (int32_t)kBypassMax && !overrun) { vb = (int32_t)get_bits4(); nb += vb; }
        uint32_t raw = 0;
        for (int32_t j = 0; j < nb && !overrun; ++j) raw |= get_bits4() << ((j * 4) & 31);
        return raw;
    }
};

// Replays RansDecoder::decode_with_indexes_gmm<4> (rans_interface.cpp:766-883) on
// the GPU-computed rows (layout: gmm_kernels.cu). row[j] = q16(cdf(v - 0.5)),
// v = j - R, j in [0, 2R+2); row[2R+2] != 0 flags a row that is not non-decreasing:
// then the reference's binary search is replayed probe by probe. Otherwise the
// search result is max{ j in [0,2R] : row[j] <= cum } (0 if none) -- the same `mid`
// the reference reaches, including its fall-off-the-end cases -- found by a short
// forward scan from the row's jump table entry for cum's bucket.
constexpr int kLutShift = 11;
constexpr int kLutBuckets = 32;
inline int table_lut_offset(int32_t R) { return ((2 * R + 3) + 7) & ~7; }
// Short rows (<= 64 candidates) are searched whole with AVX2 and carry no jump table.
#if defined(__AVX2__)
inline bool table_has_lut(int32_t R) { return 2 * R + 1 > 64; }
#else
inline bool table_has_lut(int32_t) { return true; }
#endif
inline int table_row_stride(int32_t R) { return table_lut_offset(R) + (table_has_lut(R) ? kLutBuckets : 0); }

// Per-call search plan: short rows are counted whole with AVX2 (no dependent jump-table
// load, no data-dependent branch); long rows start from the jump table.
struct SearchPlan {
    int32_t R = 0;
    int lut_off = 0;
    int nvec = 0;                 // 16-lane vectors covering the candidates j in [0, 2R]; 0 = use the jump table
    uint32_t lane_mask[4] = {0, 0, 0, 0};   // movemask bits (2 per lane) of valid candidates per vector
    explicit SearchPlan(int32_t r) : R(r), lut_off(table_lut_offset(r)) {
        const int cand = 2 * r + 1;
#if defined(__AVX2__)
        if (!table_has_lut(r)) {
            nvec = (cand + 15) / 16;
            for (int v = 0; v < nvec; ++v) {
                int lanes = cand - 16 * v;
                if (lanes > 16) lanes = 16;
                lane_mask[v] = lanes == 16 ? 0xffffffffu : ((1u << (2 * lanes)) - 1u);
            }
        }
#endif
    }
};

inline int32_t search_row(const uint16_t* row, const SearchPlan& P, uint32_t cum, uint32_t* start,
                          uint32_t* freq) {
    const int32_t R = P.R;
    int32_t j;
    if (__builtin_expect(row[2 * R + 2] == 0, 1)) {
#if defined(__AVX2__)
        if (P.nvec) {
            const __m256i c = _mm256_set1_epi16((short)cum);
            uint32_t bits = 0;
            for (int v = 0; v < P.nvec; ++v) {
                __m256i t = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(row + 16 * v));
                __m256i le = _mm256_cmpeq_epi16(_mm256_min_epu16(t, c), t);      // t <= cum (unsigned)
                bits += (uint32_t)__builtin_popcount((uint32_t)_mm256_movemask_epi8(le) & P.lane_mask[v]);
            }
            const int32_t cnt = (int32_t)(bits >> 1);
            j = cnt > 0 ? cnt - 1 : 0;
        } else
#endif
        {
            j = row[P.lut_off + (cum >> kLutShift)];
            const int32_t jmax = 2 * R;
            while (j < jmax && row[j + 1] <= cum) ++j;
        }
    } else {
        int32_t s = -R, e = R, mid = 0;
        while (s <= e) {
            mid = s + (e - s) / 2;
            uint32_t c1 = row[mid + R], c2 = row[mid + R + 1];
            if (c1 <= cum && c2 > cum) break;
            else if (c1 > cum) e = mid - 1;
            else s = mid + 1;
        }
        j = mid + R;
    }
    uint32_t c1 = row[j], c2 = row[j + 1];
    uint32_t pmf = (c2 - c1) & 0xffffu;
    if (pmf == 0) pmf = 1;               // rans_interface.cpp:866-875 (the start clamp there can never trigger)
    *start = c1; *freq = pmf;
    return j - R;
}

// Pop `cnt` symbols whose rows start at `row` (the loop of rans_interface.cpp:798-881).
inline void pop_rows(RansDecoder64& dec, const uint16_t* row, int64_t rs, int32_t R, int64_t cnt, int32_t* out) {
    const SearchPlan plan(R);
    for (int64_t i = 0; i < cnt; ++i, row += rs) {
        __builtin_prefetch(row + 8 * rs);
        __builtin_prefetch(row + 8 * rs + 32);
        const uint32_t cum = dec.get();
        int32_t v;
        if (__builtin_expect(cum == kMaxCdf, 0)) {
            dec.advance(kMaxCdf, 1);
            v = (int32_t)dec.bypass_raw();
        } else {
            uint32_t st, fr;
            if (__builtin_expect(i + 1 == cnt && 16 * plan.nvec > rs, 0)) {
                // the vector search reads whole 16-entry vectors: the last row of a table may end before the
                // last vector does (rows are padded to 8 entries), so it is searched in a padded copy
                uint16_t last[16 * 4 + 16] = {0};
                memcpy(last, row, (size_t)rs * sizeof(uint16_t));
                v = search_row(last, plan, cum, &st, &fr);
            } else {
                v = search_row(row, plan, cum, &st, &fr);
            }
            dec.advance(st, fr);
        }
        out[i] = v;
    }
}

}  // namespace flashgmm
