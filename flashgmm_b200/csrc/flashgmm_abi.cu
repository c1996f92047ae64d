// flashgmm_abi.cu -- the C ABI declared in include/flashgmm.h: contexts, chunked
// host<->device pipelines, and the hand-off between the sm_100a kernels
// (gmm_kernels.cu) and the serial host rANS recurrence (rans_host.hpp).
//
// Pipeline shape (one "lane" = 3 slots, one CUDA stream + event per slot):
//   encode: chunks are issued LAST FIRST (rANS encodes in reverse symbol order):
//           [H2D params] -> startfreq kernel -> D2H packed words; the host pushes
//           chunk c while chunks c-1, c-2 are in flight on the other slots.
//   decode: chunks in forward order: [H2D params] -> cdf-table kernel -> D2H rows;
//           the host pops chunk c while c+1, c+2 are in flight.
// Batch calls spread independent streams over the ctx's lanes (one worker thread each).
#include "abi_common.hpp"

namespace flashgmm_abi {
thread_local std::string g_err;
std::atomic<uint64_t> g_launches{0};
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
double g_prof_ms[PROF_KINDS] = {};
double g_prof_bytes[PROF_KINDS] = {};
uint64_t g_prof_launches[PROF_KINDS] = {};
}  // namespace flashgmm_abi

namespace {

inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

// Chunk plan of a stream: (offset, count) in PROCESSING order. Chunks grow
// geometrically from `first` to `cap` symbols so that the host starts its serial
// pass after a short first kernel instead of after the whole stream; every offset
// is a multiple of `first` (itself a multiple of 128), which keeps 128-bit loads
// aligned. reverse = the encoder's order (last symbols first, small chunks at the tail).
struct Chunk { int64_t c0, cnt; };
std::vector<Chunk> plan_chunks(int64_t n, int64_t first, int64_t cap, bool reverse) {
    std::vector<Chunk> out;
    if (n <= 0) return out;
    const int64_t m = n / first, r = n % first;           // whole blocks of `first` symbols + ragged rest
    if (m == 0) { out.push_back({0, n}); return out; }
    std::vector<int64_t> blocks;                          // chunk sizes in blocks: 1, 2, 4, ... <= cap/first
    const int64_t cap_b = std::max<int64_t>(1, cap / first);
    for (int64_t left = m, b = 1; left > 0; b = std::min(cap_b, b * 2)) {
        int64_t take = std::min(left, b);
        blocks.push_back(take);
        left -= take;
    }
    if (!reverse) {
        int64_t c0 = 0;
        for (size_t i = 0; i < blocks.size(); ++i) {
            int64_t cnt = blocks[i] * first + (i + 1 == blocks.size() ? r : 0);
            out.push_back({c0, cnt});
            c0 += cnt;
        }
    } else {
        int64_t end = n;
        for (size_t i = 0; i < blocks.size(); ++i) {
            int64_t cnt = blocks[i] * first + (i == 0 ? r : 0);
            out.push_back({end - cnt, cnt});
            end -= cnt;
        }
    }
    return out;
}

inline void prof_begin(Slot& S, int kind, double bytes) {
    S.prof_kind = -1;
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    if (cudaEventRecord(S.k0, S.st) == cudaSuccess) { S.prof_kind = kind; S.prof_bytes = bytes; }
}
inline void prof_end(Slot& S) {
    if (S.prof_kind >= 0) cudaEventRecord(S.k1, S.st);
}
inline void prof_collect(Slot& S) {   // call after the slot's work is known to be complete
    if (S.prof_kind < 0) return;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, S.k0, S.k1) == cudaSuccess) {
        std::lock_guard<std::mutex> g(g_prof_mu);
        g_prof_ms[S.prof_kind] += ms;
        g_prof_bytes[S.prof_kind] += S.prof_bytes;
        g_prof_launches[S.prof_kind] += 1;
    }
    S.prof_kind = -1;
}

int check_gmm_args(const float* sc, const float* mu, const float* w, int64_t n, int64_t sn, int64_t sk,
                   int mode, int mem) {
    if (n < 0) return fail(FLASHGMM_EINVAL, "n must be >= 0");
    if (mode < 0 || mode > 2) return fail(FLASHGMM_EINVAL, "mode must be 0 (Polya), 1 (A&S) or 2 (logistic)");
    if (mem != FLASHGMM_MEM_HOST && mem != FLASHGMM_MEM_DEVICE) return fail(FLASHGMM_EINVAL, "bad mem kind");
    if (n > 0 && (!sc || !mu || !w)) return fail(FLASHGMM_EINVAL, "null parameter pointer");
    if (mem == FLASHGMM_MEM_HOST && n > 0) {
        bool kmajor = (sn == 1 && sk >= n);
        bool aos = (sk == 1 && sn == 4);
        if (!kmajor && !aos)
            return fail(FLASHGMM_EINVAL,
                        "host parameters must be K-major planes (stride_n=1, stride_k>=n) or (N,4) "
                        "row-major (stride_n=4, stride_k=1); got stride_n=%lld stride_k=%lld",
                        (long long)sn, (long long)sk);
    }
    return FLASHGMM_OK;
}

// Stage parameters (and optionally symbols) of chunk [c0, c0+cnt) from host memory
// into S.d_in; returns the device view.
int stage_chunk(Slot& S, const flashgmm_stream& s, int64_t c0, int64_t cnt, bool with_symbols,
                GmmParamsView* view, const int32_t** sym_d) {
    const int64_t cntp = round_up(cnt, 4);
    const size_t plane = (size_t)cntp * 4 * sizeof(float);
    const size_t sym_bytes = with_symbols ? (size_t)cntp * 4 : 0;
    CU_TRY(S.d_in.need(sym_bytes + 3 * plane));
    char* base = static_cast<char*>(S.d_in.p);
    float* d_sc = reinterpret_cast<float*>(base + sym_bytes);
    float* d_mu = reinterpret_cast<float*>(base + sym_bytes + plane);
    float* d_w = reinterpret_cast<float*>(base + sym_bytes + 2 * plane);
    if (with_symbols) {
        CU_TRY(cudaMemcpyAsync(base, s.symbols + c0, (size_t)cnt * 4, cudaMemcpyHostToDevice, S.st));
        *sym_d = reinterpret_cast<const int32_t*>(base);
    }
    if (s.stride_n == 1) {   // K-major planes -> [4][cntp]
        const size_t spitch = (size_t)s.stride_k * 4, dpitch = (size_t)cntp * 4, width = (size_t)cnt * 4;
        CU_TRY(cudaMemcpy2DAsync(d_sc, dpitch, s.scales + c0, spitch, width, 4, cudaMemcpyHostToDevice, S.st));
        CU_TRY(cudaMemcpy2DAsync(d_mu, dpitch, s.means + c0, spitch, width, 4, cudaMemcpyHostToDevice, S.st));
        CU_TRY(cudaMemcpy2DAsync(d_w, dpitch, s.weights + c0, spitch, width, 4, cudaMemcpyHostToDevice, S.st));
        *view = GmmParamsView{d_sc, d_mu, d_w, 1, cntp};
    } else {                 // (N,4) row-major
        const size_t bytes = (size_t)cnt * 16;
        CU_TRY(cudaMemcpyAsync(d_sc, s.scales + c0 * 4, bytes, cudaMemcpyHostToDevice, S.st));
        CU_TRY(cudaMemcpyAsync(d_mu, s.means + c0 * 4, bytes, cudaMemcpyHostToDevice, S.st));
        CU_TRY(cudaMemcpyAsync(d_w, s.weights + c0 * 4, bytes, cudaMemcpyHostToDevice, S.st));
        *view = GmmParamsView{d_sc, d_mu, d_w, 4, 1};
    }
    return FLASHGMM_OK;
}

// Host buffers that are page-locked (cudaHostAlloc / cudaHostRegister, e.g. torch pin_memory)
// are read by the kernels in place over PCIe: every parameter is read exactly once, in
// coalesced 128-byte requests, so a staging copy would only add API calls and a second pass.
// Returns the device alias of `p`, or nullptr when the memory is pageable.
inline const void* pinned_alias(const void* p) {
    if (!p) return nullptr;
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return a.devicePointer;
}
// All host tensors of a stream pinned? Then fill `v` (and `sym_d`) with their device aliases.
inline bool zero_copy_view(const flashgmm_stream& s, bool with_symbols, GmmParamsView* v, const int32_t** sym_d) {
    static const bool off = env_long("FLASHGMM_ZERO_COPY", 1) == 0;
    if (off || s.n <= 0) return false;
    const float* sc = static_cast<const float*>(pinned_alias(s.scales));
    const float* mu = static_cast<const float*>(pinned_alias(s.means));
    const float* w = static_cast<const float*>(pinned_alias(s.weights));
    if (!sc || !mu || !w) return false;
    if (with_symbols) {
        const int32_t* sy = static_cast<const int32_t*>(pinned_alias(s.symbols));
        if (!sy) return false;
        *sym_d = sy;
    }
    *v = GmmParamsView{sc, mu, w, s.stride_n, s.stride_k};
    return true;
}

inline GmmParamsView device_view(const flashgmm_stream& s, int64_t c0) {
    return GmmParamsView{s.scales + c0 * s.stride_n, s.means + c0 * s.stride_n,
                         s.weights + c0 * s.stride_n, s.stride_n, s.stride_k};
}

// ---------------- encode one stream on one lane ----------------
int encode_stream(flashgmm_ctx* ctx, Lane& L, flashgmm_stream& s, int mode, int mem) {
    int rc = check_gmm_args(s.scales, s.means, s.weights, s.n, s.stride_n, s.stride_k, mode, mem);
    if (rc) return rc;
    if (s.n > 0 && !s.symbols) return fail(FLASHGMM_EINVAL, "null symbols");
    if (!s.bytes && s.bytes_cap) return fail(FLASHGMM_EINVAL, "null output buffer");
    if ((rc = L.init())) return rc;

    const int64_t n = s.n;
    const std::vector<Chunk> plan = plan_chunks(n, ctx->enc_first, ctx->enc_chunk, true);
    const int64_t nch = (int64_t)plan.size();
    GmmParamsView zc_view{};
    const int32_t* zc_sym = nullptr;
    const bool zc = mem == FLASHGMM_MEM_HOST && zero_copy_view(s, true, &zc_view, &zc_sym);
    L.bb.reset((size_t)(n * 9 / 16) + 64);
    RansEncoder64 enc;
    enc.init(&L.bb);

    auto issue = [&](int64_t q) -> int {
        const int64_t c0 = plan[q].c0, cnt = plan[q].cnt;
        Slot& S = L.slot[q % kSlots];
        const int64_t cntp = round_up(cnt, 4);
        CU_TRY(S.d_out.need((size_t)(cntp + 4) * 4));
        CU_TRY(S.h_out.need((size_t)(cntp + 4) * 4));
        uint32_t* d_packed = static_cast<uint32_t*>(S.d_out.p);
        uint32_t* d_flag = d_packed + cntp;
        GmmParamsView view;
        const int32_t* sym_d;
        if (zc) {
            view = GmmParamsView{zc_view.scales + c0 * s.stride_n, zc_view.means + c0 * s.stride_n,
                                 zc_view.weights + c0 * s.stride_n, s.stride_n, s.stride_k};
            sym_d = zc_sym + c0;
        } else if (mem == FLASHGMM_MEM_HOST) {
            int r = stage_chunk(S, s, c0, cnt, true, &view, &sym_d);
            if (r) return r;
        } else {
            view = device_view(s, c0);
            sym_d = s.symbols + c0;
        }
        CU_TRY(cudaMemsetAsync(d_flag, 0, 4, S.st));
        prof_begin(S, PROF_STARTFREQ, 56.0 * (double)cnt);
        CU_TRY(launch_gmm_startfreq(sym_d, view, cnt, mode, d_packed, d_flag, S.st));
        prof_end(S);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CU_TRY(cudaMemcpyAsync(S.h_out.p, d_packed, (size_t)(cntp + 1) * 4, cudaMemcpyDeviceToHost, S.st));
        CU_TRY(cudaEventRecord(S.ev, S.st));
        return FLASHGMM_OK;
    };

    for (int64_t q = 0; q < std::min<int64_t>(kSlots, nch); ++q)
        if ((rc = issue(q))) return rc;
    for (int64_t q = 0; q < nch; ++q) {
        const int64_t c0 = plan[q].c0, cnt = plan[q].cnt;
        Slot& S = L.slot[q % kSlots];
        CU_TRY(cudaEventSynchronize(S.ev));
        prof_collect(S);
        const uint32_t* packed = static_cast<const uint32_t*>(S.h_out.p);
        const int32_t* sym_h = nullptr;
        if (mem == FLASHGMM_MEM_HOST) {
            sym_h = s.symbols + c0;
        } else if (packed[round_up(cnt, 4)] != 0) {   // chunk has bypass symbols: fetch their values
            CU_TRY(S.h_sym.need((size_t)cnt * 4));
            CU_TRY(cudaMemcpyAsync(S.h_sym.p, s.symbols + c0, (size_t)cnt * 4, cudaMemcpyDeviceToHost, S.st));
            CU_TRY(cudaStreamSynchronize(S.st));
            sym_h = static_cast<const int32_t*>(S.h_sym.p);
        }
        enc.push_chunk_reverse(packed, sym_h, cnt);
        if (q + kSlots < nch && (rc = issue(q + kSlots))) return rc;
    }
    enc.flush();
    s.bytes_len = L.bb.bytes();
    if (s.bytes_len > s.bytes_cap)
        return fail(FLASHGMM_ENOSPC, "output buffer too small: need %zu bytes, have %zu", s.bytes_len, s.bytes_cap);
    memcpy(s.bytes, L.bb.data(), s.bytes_len);
    return FLASHGMM_OK;
}

// ---------------- decode one stream on one lane ----------------
constexpr int32_t kMaxTableR = 16383;

int decode_stream(flashgmm_ctx* ctx, Lane& L, flashgmm_stream& s, int mode, int mem) {
    int rc = check_gmm_args(s.scales, s.means, s.weights, s.n, s.stride_n, s.stride_k, mode, mem);
    if (rc) return rc;
    if (!s.bytes) return fail(FLASHGMM_EINVAL, "null bitstream");
    if (s.n > 0 && !s.symbols_out) return fail(FLASHGMM_EINVAL, "null symbols_out");
    const int32_t R = s.max_bs_value;
    if (R < 0 || R > kMaxTableR)
        return fail(FLASHGMM_EINVAL, "max_bs_value %d outside supported search radius [0, %d]", R, kMaxTableR);
    if ((rc = L.init())) return rc;

    const int64_t n = s.n;
    const int rs = table_row_stride(R);
    const int64_t cap = std::max<int64_t>(128, std::min<int64_t>(ctx->dec_chunk_bytes / (rs * 2), 1 << 20)) / 128 * 128;
    const std::vector<Chunk> plan = plan_chunks(n, std::min<int64_t>(cap, ctx->dec_first), cap, false);
    const int64_t nch = (int64_t)plan.size();

    GmmParamsView zc_view{};
    const int32_t* zc_unused = nullptr;
    const bool zc = mem == FLASHGMM_MEM_HOST && zero_copy_view(s, false, &zc_view, &zc_unused);
    int32_t* out_h = s.symbols_out;
    if (mem == FLASHGMM_MEM_DEVICE && n > 0) {
        CU_TRY(L.h_dec.need((size_t)n * 4));
        out_h = static_cast<int32_t*>(L.h_dec.p);
    }

    auto issue = [&](int64_t q) -> int {
        const int64_t c0 = plan[q].c0, cnt = plan[q].cnt;
        Slot& S = L.slot[q % kSlots];
        const size_t tab_bytes = (size_t)cnt * rs * 2;
        CU_TRY(S.d_out.need(tab_bytes));
        CU_TRY(S.h_out.need(tab_bytes));
        GmmParamsView view;
        if (zc) {
            view = GmmParamsView{zc_view.scales + c0 * s.stride_n, zc_view.means + c0 * s.stride_n,
                                 zc_view.weights + c0 * s.stride_n, s.stride_n, s.stride_k};
        } else if (mem == FLASHGMM_MEM_HOST) {
            const int32_t* unused;
            int r = stage_chunk(S, s, c0, cnt, false, &view, &unused);
            if (r) return r;
        } else {
            view = device_view(s, c0);
        }
        prof_begin(S, PROF_TABLE, (48.0 + 2.0 * rs) * (double)cnt);
        CU_TRY(launch_gmm_cdf_table(view, cnt, mode, R, static_cast<uint16_t*>(S.d_out.p), rs, table_has_lut(R), S.st));
        prof_end(S);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CU_TRY(cudaMemcpyAsync(S.h_out.p, S.d_out.p, tab_bytes, cudaMemcpyDeviceToHost, S.st));
        CU_TRY(cudaEventRecord(S.ev, S.st));
        return FLASHGMM_OK;
    };

    RansDecoder64 dec;
    dec.init(s.bytes, s.bytes_cap);
    for (int64_t q = 0; q < std::min<int64_t>(kSlots, nch); ++q)
        if ((rc = issue(q))) return rc;
    for (int64_t q = 0; q < nch; ++q) {
        const int64_t c0 = plan[q].c0, cnt = plan[q].cnt;
        Slot& S = L.slot[q % kSlots];
        CU_TRY(cudaEventSynchronize(S.ev));
        prof_collect(S);
        const uint16_t* row = static_cast<const uint16_t*>(S.h_out.p);
        int32_t* o = out_h + c0;
        pop_rows(dec, row, rs, R, cnt, o);
        if (mem == FLASHGMM_MEM_DEVICE)
            CU_TRY(cudaMemcpyAsync(s.symbols_out + c0, o, (size_t)cnt * 4, cudaMemcpyHostToDevice, S.st));
        if (q + kSlots < nch && (rc = issue(q + kSlots))) return rc;
    }
    if (mem == FLASHGMM_MEM_DEVICE)
        for (auto& S : L.slot) CU_TRY(cudaStreamSynchronize(S.st));
    if (dec.overrun) return fail(FLASHGMM_ETRUNC, "bitstream truncated: decoder read past %zu bytes", s.bytes_cap);
    return FLASHGMM_OK;
}

Lane& lane_of(flashgmm_ctx* ctx, int idx) {
    while ((int)ctx->lanes.size() <= idx) ctx->lanes.emplace_back(new Lane());
    return *ctx->lanes[idx];
}


// ---------------- decode a batch of streams entirely on the device ----------------
// rows kernel (all streams of a chunk in one launch) -> pop kernel (one warp per stream);
// chunk c+1's rows are computed while chunk c is popped. Only bitstreams go up and decoded
// symbols come down (mem == HOST additionally stages the parameters chunk by chunk).
constexpr int32_t kMaxDevicePopR = kPopWindowMaxR;   // wider searches stay on the host path

int decode_batch_device(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem) {
    std::lock_guard<std::mutex> lock(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    if (!ctx->dd) ctx->dd.reset(new DevDecode());
    DevDecode& D = *ctx->dd;
    int rc = D.init();
    if (rc) return rc;

    // ---- validate, size ----
    std::vector<int> act;                       // streams that take part
    int worst = FLASHGMM_OK;
    int64_t max_n = 0, total_n = 0;
    size_t total_words = 0;
    int32_t max_R = 0;
    for (int i = 0; i < n_streams; ++i) {
        flashgmm_stream& s = streams[i];
        s.status = check_gmm_args(s.scales, s.means, s.weights, s.n, s.stride_n, s.stride_k, mode, mem);
        if (!s.status && !s.bytes) s.status = fail(FLASHGMM_EINVAL, "null bitstream");
        if (!s.status && s.n > 0 && !s.symbols_out) s.status = fail(FLASHGMM_EINVAL, "null symbols_out");
        if (!s.status && (s.max_bs_value < 0 || s.max_bs_value > kMaxDevicePopR))
            s.status = fail(FLASHGMM_EINVAL, "max_bs_value %d outside [0, %d]", s.max_bs_value, kMaxDevicePopR);
        if (!s.status && s.n > INT32_MAX / 2) s.status = fail(FLASHGMM_EINVAL, "stream too long for one call");
        if (s.status) { worst = s.status; continue; }
        if (s.n == 0) continue;
        act.push_back(i);
        max_n = std::max(max_n, s.n);
        total_n += s.n;
        total_words += (s.bytes_cap / 4 + 3) / 4 * 4;       // 16-byte aligned slots
        max_R = std::max(max_R, s.max_bs_value);
    }
    const int na = (int)act.size();
    if (na == 0) return worst;
    // Streams by length, longest first, cut into groups where the length halves. The longest streams'
    // pop is the critical path of the call (serial, ~58 ns per symbol): a group's rows are never
    // made to wait for the rows of shorter streams, and every group pops on its own CUDA stream.
    std::stable_sort(act.begin(), act.end(), [&](int a, int b) { return streams[a].n > streams[b].n; });
    std::vector<int> grp_begin{0};
    for (int a = 1; a < na; ++a)
        if ((int)grp_begin.size() < kPopGroups && streams[act[a]].n * 2 <= streams[act[grp_begin.back()]].n)
            grp_begin.push_back(a);
    const int ng = (int)grp_begin.size();
    grp_begin.push_back(na);

    std::vector<int32_t> rs(na);
    std::vector<GmmParamsView> zc(na);                       // host callers: device aliases of page-locked tensors
    std::vector<char> is_zc(na, 0);
    if (mem == FLASHGMM_MEM_HOST)
        for (int a = 0; a < na; ++a) {
            const int32_t* unused = nullptr;
            is_zc[a] = zero_copy_view(streams[act[a]], false, &zc[a], &unused) ? 1 : 0;
        }
    std::vector<size_t> word_off(na), out_off(na), par_off(na), npad(na);
    {
        size_t wo = 0, oo = 0, po = 0;
        for (int a = 0; a < na; ++a) {
            const flashgmm_stream& s = streams[act[a]];
            rs[a] = gmm_table_lut_offset(s.max_bs_value);        // compact tile width (uint16) inside the rows kernel
            word_off[a] = wo; wo += (s.bytes_cap / 4 + 3) / 4 * 4;
            out_off[a] = oo; oo += (size_t)round_up(s.n, 4);
            npad[a] = (size_t)round_up(s.n, 4);
            par_off[a] = po; po += is_zc[a] ? 0 : 12 * npad[a];
        }
    }
    const int max_row_words = gmm_table_lut_offset(max_R) / 2;
    // windowed records (one interval per lane, any R) unless FLASHGMM_POP_WINDOW=0 asks for the
    // every-interval lane rows (R <= 63)
    static const bool windowed_dflt = env_long("FLASHGMM_POP_WINDOW", 1) != 0;
    const bool windowed = windowed_dflt || max_R > kPopMaxR;
    const int lane_slots = windowed ? kPopWindowed : gmm_pop_lane_slots(max_R);
    const int64_t rec_bytes = windowed ? 4 * (int64_t)kPopWindowWords : 4 * (int64_t)lane_slots;   // record per symbol
    int S = (int)std::min<int64_t>(32, (48 * 1024) / ((12 + max_row_words + 2) * 4));
    if (total_n < 148 * 8 * 32) S = std::min(S, 8);
    if (windowed) S = kRowsWindowSyms;                       // the windowed kernel's CTA shape is fixed

    // ---- work items: (group, symbol range common to the group's streams) ----
    struct Item { int g; int64_t lo, hi, left; int k, ring_prev; };   // k: index within the group; ring_prev: the item whose row buffer it reuses
    auto item_bytes = [&](int g, int64_t lo, int64_t hi) {
        int64_t b = 0;
        for (int a = grp_begin[g]; a < grp_begin[g + 1]; ++a) {
            const int64_t n = streams[act[a]].n;
            if (n > lo) b += (std::min(n, hi) - lo) * rec_bytes;
        }
        return b;
    };
    std::vector<Item> items;
    for (int g = 0; g < ng; ++g) {
        const int64_t gmax = streams[act[grp_begin[g]]].n;
        int64_t step = 4096, lo = 0;
        while (lo < gmax) {
            int64_t st = step;
            while (st > 256 && item_bytes(g, lo, lo + st) > ctx->pop_rows_budget) st /= 2;
            const int64_t hi = std::min(gmax, lo + st);
            items.push_back(Item{g, lo, hi, gmax - lo, 0, -1});
            lo = hi;
            step = std::min<int64_t>(32768, step * 3);
        }
    }
    // earliest deadline first: the item with the most symbols still to pop behind it is the most
    // urgent (within a group that is the natural order)
    std::stable_sort(items.begin(), items.end(), [](const Item& x, const Item& y) { return x.left > y.left; });
    const int nch = (int)items.size();
    {
        std::vector<std::vector<int>> of_group(ng);
        for (int c = 0; c < nch; ++c) {
            std::vector<int>& v = of_group[items[c].g];
            items[c].k = (int)v.size();
            items[c].ring_prev = items[c].k >= kRowBufs ? v[items[c].k - kRowBufs] : -1;
            v.push_back(c);
        }
    }
    if ((rc = D.need_events(nch))) return rc;
    const bool prof = g_prof_on.load(std::memory_order_relaxed) != 0;
    static const bool trace = env_long("FLASHGMM_TRACE", 0) != 0;   // per-chunk timeline on stderr (with profiling on)
    if (prof && (rc = D.need_timers(3 * (size_t)nch + 1))) return rc;

    // ---- device buffers ----
    for (int g = 0; g < ng; ++g) {
        int64_t biggest = 0;
        int n_items = 0;
        for (int c = 0; c < nch; ++c)
            if (items[c].g == g) { biggest = std::max(biggest, item_bytes(g, items[c].lo, items[c].hi)); ++n_items; }
        for (int b = 0; b < std::min(n_items, kRowBufs); ++b) CU_TRY(D.rows[g][b].need((size_t)biggest));
    }
    static const bool spread = env_long("FLASHGMM_POP_SPREAD", 0) != 0;     // SM-aware placement of the pop warps (opt-in: spreads them, measured no gain)
    if (D.n_sm == 0) { CU_TRY(cudaDeviceGetAttribute(&D.n_sm, cudaDevAttrMultiProcessorCount, ctx->device)); D.n_sm = std::max(1, D.n_sm); }
    if (spread) CU_TRY(D.claim.need((size_t)nch * kPopClaimInts * sizeof(int)));
    CU_TRY(D.words.need(total_words * 4 + 16));
    CU_TRY(D.h_words.need(total_words * 4 + 16));
    CU_TRY(D.state.need((size_t)na * 16));
    CU_TRY(D.h_state.need((size_t)na * 16));
    int32_t* d_out = nullptr;
    float* d_par = nullptr;
    if (mem == FLASHGMM_MEM_HOST) {
        CU_TRY(D.out.need((out_off[na - 1] + npad[na - 1]) * 4));
        CU_TRY(D.params.need((par_off[na - 1] + 12 * npad[na - 1]) * 4 + 16));
        d_out = static_cast<int32_t*>(D.out.p);
        d_par = static_cast<float*>(D.params.p);
    }

    // ---- descriptors for every chunk, one upload ----
    struct ChunkInfo { int first, count, blocks; };
    std::vector<ChunkInfo> ci(nch);
    size_t n_desc = 0;
    for (int c = 0; c < nch; ++c)
        for (int a = grp_begin[items[c].g]; a < grp_begin[items[c].g + 1]; ++a) n_desc += streams[act[a]].n > items[c].lo ? 1 : 0;
    const size_t rows_desc_bytes = (n_desc * sizeof(RowsDesc) + 15) / 16 * 16;
    CU_TRY(D.h_desc.need(rows_desc_bytes + n_desc * sizeof(PopDesc)));
    CU_TRY(D.desc.need(rows_desc_bytes + n_desc * sizeof(PopDesc)));
    RowsDesc* h_rd = static_cast<RowsDesc*>(D.h_desc.p);
    PopDesc* h_pd = reinterpret_cast<PopDesc*>(static_cast<char*>(D.h_desc.p) + rows_desc_bytes);
    RowsDesc* d_rd = static_cast<RowsDesc*>(D.desc.p);
    PopDesc* d_pd = reinterpret_cast<PopDesc*>(static_cast<char*>(D.desc.p) + rows_desc_bytes);
    auto stream_view = [&](int a, int64_t c0) -> GmmParamsView {
        const flashgmm_stream& s = streams[act[a]];
        if (mem == FLASHGMM_MEM_DEVICE) return device_view(s, c0);
        if (is_zc[a])
            return GmmParamsView{zc[a].scales + c0 * s.stride_n, zc[a].means + c0 * s.stride_n,
                                 zc[a].weights + c0 * s.stride_n, s.stride_n, s.stride_k};
        float* b = d_par + par_off[a];
        if (s.stride_n == 1)
            return GmmParamsView{b + c0, b + 4 * npad[a] + c0, b + 8 * npad[a] + c0, 1, (int64_t)npad[a]};
        return GmmParamsView{b + 4 * c0, b + 4 * npad[a] + 4 * c0, b + 8 * npad[a] + 4 * c0, 4, 1};
    };
    {
        size_t k = 0;
        for (int c = 0; c < nch; ++c) {
            ci[c].first = (int)k;
            int blocks = 0;
            size_t roff = 0;
            char* rbase = static_cast<char*>(D.rows[items[c].g][items[c].k % kRowBufs].p);
            for (int a = grp_begin[items[c].g]; a < grp_begin[items[c].g + 1]; ++a) {
                const flashgmm_stream& s = streams[act[a]];
                if (s.n <= items[c].lo) continue;
                const int64_t c0 = items[c].lo, cnt = std::min(s.n, items[c].hi) - c0;
                RowsDesc& r = h_rd[k];
                r.p = stream_view(a, c0);
                r.table32 = reinterpret_cast<uint32_t*>(rbase + roff);
                r.n = (int32_t)cnt; r.R = s.max_bs_value; r.row_words = rs[a] / 2; r.first_block = blocks;
                PopDesc& q = h_pd[k];
                q.rows = reinterpret_cast<const uint16_t*>(rbase + roff);
                q.words = static_cast<const uint32_t*>(D.words.p) + word_off[a];
                q.out = (mem == FLASHGMM_MEM_DEVICE ? s.symbols_out : d_out + out_off[a]) + c0;
                q.state = static_cast<uint64_t*>(D.state.p) + 2 * (size_t)a;
                q.p = r.p;
                q.n = (int32_t)cnt; q.R = s.max_bs_value;
                q.n_words = (uint32_t)(s.bytes_cap / 4); q.first = c0 == 0 ? 1 : 0;
                blocks += (int)((cnt + S - 1) / S);
                roff += (size_t)cnt * rec_bytes;
                ++k;
            }
            ci[c].count = (int)k - ci[c].first;
            ci[c].blocks = blocks;
        }
    }
    CU_TRY(cudaMemcpyAsync(D.desc.p, D.h_desc.p, rows_desc_bytes + n_desc * sizeof(PopDesc), cudaMemcpyHostToDevice,
                           D.st_copy));

    // ---- parameter staging per chunk (host callers with pageable tensors) ----
    // page-locked tensors are not staged at all: the rows kernel reads them in place over PCIe
    // (each parameter exactly once, 128-byte coalesced requests, thousands of warps in flight).
    auto upload_chunk = [&](int c) -> int {
        if (mem == FLASHGMM_MEM_HOST) {
            for (int a = grp_begin[items[c].g]; a < grp_begin[items[c].g + 1]; ++a) {
                const flashgmm_stream& s = streams[act[a]];
                if (s.n <= items[c].lo || is_zc[a]) continue;
                const int64_t c0 = items[c].lo, cnt = std::min(s.n, items[c].hi) - c0;
                float* b = d_par + par_off[a];
                const float* src[3] = {s.scales, s.means, s.weights};
                for (int t = 0; t < 3; ++t) {
                    float* dst = b + 4 * npad[a] * t;
                    if (s.stride_n == 1)
                        CU_TRY(cudaMemcpy2DAsync(dst + c0, npad[a] * 4, src[t] + c0, (size_t)s.stride_k * 4,
                                                 (size_t)cnt * 4, 4, cudaMemcpyHostToDevice, D.st_copy));
                    else
                        CU_TRY(cudaMemcpyAsync(dst + 4 * c0, src[t] + 4 * c0, (size_t)cnt * 16, cudaMemcpyHostToDevice,
                                               D.st_copy));
                }
            }
        }
        CU_TRY(cudaEventRecord(D.ev_up[c], D.st_copy));
        if (prof) CU_TRY(cudaEventRecord(D.t0[2 * nch + c], D.st_copy));
        return FLASHGMM_OK;
    };

    auto launch_rows = [&](int c) -> int {
        CU_TRY(cudaStreamWaitEvent(D.st_rows, D.ev_up[c], 0));
        if (items[c].ring_prev >= 0) CU_TRY(cudaStreamWaitEvent(D.st_rows, D.ev_pop[items[c].ring_prev], 0));
        if (prof) CU_TRY(cudaEventRecord(D.t0[2 * c], D.st_rows));
        CU_TRY(launch_gmm_cdf_rows_batched(d_rd + ci[c].first, ci[c].count, ci[c].blocks, S, max_row_words, lane_slots,
                                           mode, D.st_rows));
        if (prof) CU_TRY(cudaEventRecord(D.t1[2 * c], D.st_rows));
        CU_TRY(cudaEventRecord(D.ev_rows[c], D.st_rows));
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return FLASHGMM_OK;
    };
    auto launch_pop = [&](int c) -> int {
        cudaStream_t sp = D.st_pop[items[c].g];
        CU_TRY(cudaStreamWaitEvent(sp, D.ev_rows[c], 0));
        if (items[c].lo == 0) CU_TRY(cudaStreamWaitEvent(sp, D.ev_words[items[c].g], 0));
        if (prof) CU_TRY(cudaEventRecord(D.t0[2 * c + 1], sp));
        CU_TRY(launch_gmm_pop_rows(d_pd + ci[c].first, ci[c].count, lane_slots, mode,
                                   spread ? static_cast<int*>(D.claim.p) + (size_t)c * kPopClaimInts : nullptr,
                                   (ci[c].count + D.n_sm - 1) / D.n_sm, sp));
        if (prof) CU_TRY(cudaEventRecord(D.t1[2 * c + 1], sp));
        CU_TRY(cudaEventRecord(D.ev_pop[c], sp));
        if (items[c].hi >= streams[act[grp_begin[items[c].g]]].n) {
            // the group's last item: its final states and (host callers) its symbols go home behind it
            const int a0 = grp_begin[items[c].g], a1 = grp_begin[items[c].g + 1];
            CU_TRY(cudaMemcpyAsync(static_cast<char*>(D.h_state.p) + 16 * (size_t)a0, static_cast<char*>(D.state.p) + 16 * (size_t)a0,
                                   16 * (size_t)(a1 - a0), cudaMemcpyDeviceToHost, sp));
            if (mem == FLASHGMM_MEM_HOST)
                for (int a = a0; a < a1; ++a) {
                    const flashgmm_stream& s = streams[act[a]];
                    CU_TRY(cudaMemcpyAsync(s.symbols_out, d_out + out_off[a], (size_t)s.n * 4, cudaMemcpyDeviceToHost, sp));
                }
        }
        g_launches.fetch_add(1, std::memory_order_relaxed);
        return FLASHGMM_OK;
    };

    // The first chunk's rows launch goes out first, then the bitstreams (their own stream: pop(0)
    // must not wait behind parameter traffic), then the staging copies of pageable tensors; rows
    // kernels run up to kRowBufs - 1 items ahead of their group's pop kernel.
    if (prof) CU_TRY(cudaEventRecord(D.t0[3 * nch], D.st_copy));
    if ((rc = upload_chunk(0))) return rc;
    if ((rc = launch_rows(0))) return rc;
    {
        if (spread) CU_TRY(cudaMemsetAsync(D.claim.p, 0, (size_t)nch * kPopClaimInts * sizeof(int), D.st_words));
        // group by group, the longest streams' words first: their pop is the critical path
        uint32_t* hw = static_cast<uint32_t*>(D.h_words.p);
        for (int g = 0; g < ng; ++g) {
            const int a0 = grp_begin[g], a1 = grp_begin[g + 1];
            for (int a = a0; a < a1; ++a) {
                const flashgmm_stream& s = streams[act[a]];
                memcpy(hw + word_off[a], s.bytes, s.bytes_cap / 4 * 4);
            }
            const size_t w0 = word_off[a0], w1 = a1 < na ? word_off[a1] : total_words;
            CU_TRY(cudaMemcpyAsync(static_cast<uint32_t*>(D.words.p) + w0, hw + w0, (w1 - w0) * 4, cudaMemcpyHostToDevice,
                                   D.st_words));
            CU_TRY(cudaEventRecord(D.ev_words[g], D.st_words));
        }
    }
    for (int c = 1; c < nch; ++c)
        if ((rc = upload_chunk(c))) return rc;
    {
        // enqueue order: a rows launch that reuses a buffer goes out after the pop launch that frees it
        // (a wait on an event not yet recorded would be a no-op); otherwise pops trail rows by two items
        int next_pop = 0;
        for (int c = 1; c < nch; ++c) {
            while (next_pop <= items[c].ring_prev)
                if ((rc = launch_pop(next_pop++))) return rc;
            if ((rc = launch_rows(c))) return rc;
            while (next_pop + 2 <= c)
                if ((rc = launch_pop(next_pop++))) return rc;
        }
        while (next_pop < nch)
            if ((rc = launch_pop(next_pop++))) return rc;
    }
    for (int g = 0; g < ng; ++g) CU_TRY(cudaStreamSynchronize(D.st_pop[g]));
    CU_TRY(cudaStreamSynchronize(D.st_rows));
    CU_TRY(cudaStreamSynchronize(D.st_copy));
    CU_TRY(cudaStreamSynchronize(D.st_words));

    if (prof && trace) {
        for (int c = 0; c < nch; ++c) {
            float up = 0, r0 = 0, r1 = 0, p0 = 0, p1 = 0;
            cudaEventElapsedTime(&up, D.t0[3 * nch], D.t0[2 * nch + c]);
            cudaEventElapsedTime(&r0, D.t0[3 * nch], D.t0[2 * c]);
            cudaEventElapsedTime(&r1, D.t0[3 * nch], D.t1[2 * c]);
            cudaEventElapsedTime(&p0, D.t0[3 * nch], D.t0[2 * c + 1]);
            cudaEventElapsedTime(&p1, D.t0[3 * nch], D.t1[2 * c + 1]);
            fprintf(stderr, "[flashgmm trace] item %d group %d [%lld,%lld) streams %d: staged %.2f | rows %.2f-%.2f | pop %.2f-%.2f ms\n",
                    c, items[c].g, (long long)items[c].lo, (long long)items[c].hi, ci[c].count, up, r0, r1, p0, p1);
        }
    }
    if (prof) {
        std::lock_guard<std::mutex> pg(g_prof_mu);
        for (int c = 0; c < nch; ++c) {
            float ms = 0;
            double rows_b = 0, pop_b = 0;
            for (int k = ci[c].first; k < ci[c].first + ci[c].count; ++k) {
                rows_b += (48.0 + (double)rec_bytes) * h_rd[k].n;
                pop_b += ((double)rec_bytes + 4.0) * h_rd[k].n;
            }
            if (cudaEventElapsedTime(&ms, D.t0[2 * c], D.t1[2 * c]) == cudaSuccess) {
                g_prof_ms[PROF_TABLE] += ms; g_prof_bytes[PROF_TABLE] += rows_b; g_prof_launches[PROF_TABLE] += 1;
            }
            if (cudaEventElapsedTime(&ms, D.t0[2 * c + 1], D.t1[2 * c + 1]) == cudaSuccess) {
                g_prof_ms[PROF_POP] += ms; g_prof_bytes[PROF_POP] += pop_b; g_prof_launches[PROF_POP] += 1;
            }
        }
    }
    const uint64_t* hs = static_cast<const uint64_t*>(D.h_state.p);
    if (prof && trace) {                                     // how the groups' last pop launches were spread over the SMs
        for (int g2 = 0; g2 < ng; ++g2) {
            int per_sm[1024] = {0}, used = 0, worst_sm = 0;
            for (int a = grp_begin[g2]; a < grp_begin[g2 + 1]; ++a) ++per_sm[(hs[2 * a + 1] >> 40) & 1023];
            for (int v : per_sm) { used += v ? 1 : 0; worst_sm = std::max(worst_sm, v); }
            fprintf(stderr, "[flashgmm trace] group %d: %d pop warps on %d SMs, at most %d on one SM\n", g2,
                    grp_begin[g2 + 1] - grp_begin[g2], used, worst_sm);
        }
    }
    for (int a = 0; a < na; ++a) {
        flashgmm_stream& s = streams[act[a]];
        if (((hs[2 * a + 1] >> 32) & 1u) != 0) {
            s.status = fail(FLASHGMM_ETRUNC, "bitstream truncated: decoder read past %zu bytes", s.bytes_cap);
            worst = s.status;
        }
    }
    return worst;
}

// Device pop (one warp per stream, ~53 ns per symbol whatever the number of streams) or host pop (full rows come down
// over PCIe, a worker thread pops a stream at a few ns per symbol)? The device path's time is its longest stream; the
// host path's is what the rows cost on the link (measured ~30 GB/s with several streams in flight) or the workers'
// rounds. Small batches -- a few images of one pass, a single image's streams -- are faster on the host (measured for
// 8 / 12 ELIC images per pass: 15.0 / 19.1 ms against 25.1 / 25.5 ms per batch); from ~16 images on, and whenever the
// host has few cores, the device wins. FLASHGMM_POP_POLICY=device|host overrides, FLASHGMM_POP_MIN_STREAMS stays a floor.
bool want_device_pop(const flashgmm_ctx* ctx, const flashgmm_stream* streams, int n_streams) {
    if (n_streams < ctx->pop_min_streams) return false;
    for (int i = 0; i < n_streams; ++i)
        if (streams[i].max_bs_value > kMaxDevicePopR) return false;
    if (ctx->pop_policy == 1) return true;
    if (ctx->pop_policy == 2) return false;
    double row_bytes = 0, total_n = 0, max_n = 0;
    for (int i = 0; i < n_streams; ++i) {
        const double n = (double)streams[i].n;
        if (streams[i].max_bs_value > kMaxTableR) return true;          // the host table is limited in radius
        row_bytes += n * 2.0 * table_row_stride(streams[i].max_bs_value);
        total_n += n;
        max_n = std::max(max_n, n);
    }
    const double workers = (double)std::max(1, ctx->n_workers);
    const double rounds = std::ceil((double)n_streams / workers);
    const double host_s = std::max(row_bytes / 30e9, rounds * max_n * 6e-9) + 150e-6;
    const double dev_s = max_n * 53e-9 + 200e-6;
    return !(host_s < 0.8 * dev_s);
}


// The library keeps up to ~20 CUDA streams busy at once (one per worker thread, the device
// encoder's, the device decoder's). With the driver's default of 8 hardware queues streams share
// queues, and a copy or event queued behind a long kernel stalls every stream that shares its
// queue (measured: batch encode 6.1 ms -> 5.0 ms with 32 queues). The setting is read when the
// CUDA context is created, so the host must ask for it BEFORE that (flashgmm_request_hw_queues);
// the library no longer touches the process environment on load, and works with the default.

// ---------------- encode a batch of streams entirely on the device ----------------
// batched startfreq kernel -> push kernel (one warp per stream) -> one D2H of the words.
// Returns -1 when the batch is not eligible (pageable host tensors): the caller then takes
// the worker-thread path.
// Two phases so that the caller's thread can queue the device part, push its share of the host
// part, and only then wait: _begin queues everything (asynchronous), _finish waits and hands out
// the streams. The caller holds ctx->de_mu across both.
int encode_batch_device_begin(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem) {
    CU_TRY(cudaSetDevice(ctx->device));
    if (!ctx->de) ctx->de.reset(new DevEncode());
    DevEncode& E = *ctx->de;
    int rc = E.init();
    if (rc) return rc;

    std::vector<int>& act = E.act;
    act.clear();
    E.worst = FLASHGMM_OK;
    std::vector<GmmParamsView> view(n_streams);
    std::vector<const int32_t*> sym(n_streams, nullptr);
    int& worst = E.worst;
    for (int i = 0; i < n_streams; ++i) {
        flashgmm_stream& s = streams[i];
        s.bytes_len = 0;
        s.status = check_gmm_args(s.scales, s.means, s.weights, s.n, s.stride_n, s.stride_k, mode, mem);
        if (!s.status && s.n > 0 && !s.symbols) s.status = fail(FLASHGMM_EINVAL, "null symbols");
        if (!s.status && !s.bytes && s.bytes_cap) s.status = fail(FLASHGMM_EINVAL, "null output buffer");
        if (!s.status && s.n > INT32_MAX / 2) s.status = fail(FLASHGMM_EINVAL, "stream too long for one call");
        if (s.status) { worst = s.status; continue; }
        if (mem == FLASHGMM_MEM_DEVICE) {
            view[i] = device_view(s, 0);
            sym[i] = s.symbols;
        } else if (s.n > 0 && !zero_copy_view(s, true, &view[i], &sym[i])) {
            return -1;
        }
        act.push_back(i);
    }
    const int na = (int)act.size();
    if (na == 0) return FLASHGMM_OK;

    std::vector<size_t> pk_off(na), sc_off(na);
    size_t pk_total = 0, sc_total = 0;
    int blocks = 0;
    for (int a = 0; a < na; ++a) {
        const flashgmm_stream& s = streams[act[a]];
        pk_off[a] = pk_total; pk_total += (size_t)round_up(std::max<int64_t>(s.n, 1), 4);
        sc_off[a] = sc_total; sc_total += s.bytes_cap / 4;
    }
    CU_TRY(E.packed.need(pk_total * 4));
    CU_TRY(E.scratch.need(sc_total * 4 + 16));
    // words and per-stream results are written by the kernel straight into page-locked host
    // memory: no copy is queued behind the long push kernel (a dependent entry would stall every
    // other stream that shares its hardware queue, the worker threads' among them)
    CU_TRY(E.h_out.need(sc_total * 4 + 16));
    CU_TRY(E.result.need(16));
    CU_TRY(E.h_result.need((size_t)na * 32 + 16));
    const size_t sf_bytes = ((size_t)na * sizeof(SfDesc) + 15) / 16 * 16;
    CU_TRY(E.desc.need(sf_bytes + (size_t)na * sizeof(PushDesc)));
    CU_TRY(E.h_desc.need(sf_bytes + (size_t)na * sizeof(PushDesc)));
    if (!E.rcp_ready) {
        static_assert(sizeof(PushRcp) == sizeof(RcpEntry), "device table mirrors the host one");
        CU_TRY(E.rcp.need(65536 * sizeof(PushRcp)));
        CU_TRY(cudaMemcpyAsync(E.rcp.p, rcp_table(), 65536 * sizeof(PushRcp), cudaMemcpyHostToDevice, E.st));
        E.rcp_ready = true;
    }
    SfDesc* h_sf = static_cast<SfDesc*>(E.h_desc.p);
    PushDesc* h_pd = reinterpret_cast<PushDesc*>(static_cast<char*>(E.h_desc.p) + sf_bytes);
    uint32_t* d_packed = static_cast<uint32_t*>(E.packed.p);
    uint32_t* d_result = static_cast<uint32_t*>(E.h_result.p);
    unsigned long long* d_used = static_cast<unsigned long long*>(E.result.p);
    double total_n = 0;
    for (int a = 0; a < na; ++a) {
        const flashgmm_stream& s = streams[act[a]];
        h_sf[a] = SfDesc{view[act[a]], sym[act[a]], d_packed + pk_off[a], (int32_t)s.n, blocks, nullptr};
        blocks += (int)((s.n + 127) / 128);
        h_pd[a] = PushDesc{d_packed + pk_off[a], sym[act[a]], static_cast<uint32_t*>(E.scratch.p) + sc_off[a],
                           (uint32_t)(s.bytes_cap / 4), (int32_t)s.n, d_result + 8 * (size_t)a};
        total_n += (double)s.n;
    }
    CU_TRY(cudaMemcpyAsync(E.desc.p, E.h_desc.p, sf_bytes + (size_t)na * sizeof(PushDesc), cudaMemcpyHostToDevice, E.st));
    CU_TRY(cudaMemsetAsync(d_used, 0, 16, E.st));                  // words claimed (u64) + push warps resident (u32)
    const bool prof = E.prof = g_prof_on.load(std::memory_order_relaxed) != 0;
    E.total_n = total_n;
    if (prof) CU_TRY(cudaEventRecord(E.t[0], E.st));
    CU_TRY(launch_gmm_startfreq_batched(static_cast<const SfDesc*>(E.desc.p), na, blocks, mode, E.st));
    if (prof) CU_TRY(cudaEventRecord(E.t[1], E.st));
    CU_TRY(cudaEventRecord(E.ev_sf, E.st));
    CU_TRY(launch_gmm_push(reinterpret_cast<const PushDesc*>(static_cast<char*>(E.desc.p) + sf_bytes), na,
                           static_cast<const PushRcp*>(E.rcp.p), static_cast<uint32_t*>(E.h_out.p), d_used,
                           reinterpret_cast<unsigned int*>(d_used + 1), E.st));
    g_launches.fetch_add(2, std::memory_order_relaxed);
    return FLASHGMM_OK;
}

int encode_batch_device_finish(flashgmm_ctx* ctx, flashgmm_stream* streams) {
    DevEncode& E = *ctx->de;
    const std::vector<int>& act = E.act;
    const int na = (int)act.size();
    int worst = E.worst;
    if (na == 0) return worst;
    const bool prof = E.prof;
    const double total_n = E.total_n;
    CU_TRY(cudaStreamSynchronize(E.st));
    const uint32_t* hr = static_cast<const uint32_t*>(E.h_result.p);
    unsigned long long used_words = 0;
    unsigned long long t_first = ~0ull, t_last = 0;
    for (int a = 0; a < na; ++a) {
        used_words += std::min<unsigned long long>(hr[8 * a], streams[act[a]].bytes_cap / 4);
        t_first = std::min(t_first, (unsigned long long)hr[8 * a + 4] | ((unsigned long long)hr[8 * a + 5] << 32));
        t_last = std::max(t_last, (unsigned long long)hr[8 * a + 6] | ((unsigned long long)hr[8 * a + 7] << 32));
    }
    {
        // the cost model's device figure follows what the push warps actually did (their own %globaltimer stamps):
        // exponential average of ns per symbol over the streams long enough to be past start-up effects
        double ns_sum = 0;
        int cnt = 0;
        for (int a = 0; a < na; ++a) {
            const unsigned long long b = (unsigned long long)hr[8 * a + 4] | ((unsigned long long)hr[8 * a + 5] << 32);
            const unsigned long long e = (unsigned long long)hr[8 * a + 6] | ((unsigned long long)hr[8 * a + 7] << 32);
            const int64_t n = streams[act[a]].n;
            if (n >= 8192 && e > b) { ns_sum += (double)(e - b) / (double)n; ++cnt; }
        }
        if (cnt > 0) {
            const double seen = std::min(200.0, std::max(10.0, ns_sum / cnt));
            ctx->push_dev_ns = 0.75 * ctx->push_dev_ns + 0.25 * seen;
        }
    }
    static const bool dtrace = env_long("FLASHGMM_TRACE", 0) != 0;
    if (dtrace) {
        unsigned long long latest_start = 0, longest = 0;
        for (int a = 0; a < na; ++a) {
            const unsigned long long b = (unsigned long long)hr[8 * a + 4] | ((unsigned long long)hr[8 * a + 5] << 32);
            const unsigned long long e = (unsigned long long)hr[8 * a + 6] | ((unsigned long long)hr[8 * a + 7] << 32);
            latest_start = std::max(latest_start, b - t_first);
            longest = std::max(longest, e - b);
        }
        fprintf(stderr, "[flashgmm trace] push kernel ran %.3f ms (its own %%globaltimer stamps): last warp started at %.3f ms, longest warp %.3f ms;",
                (double)(t_last - t_first) * 1e-6, (double)latest_start * 1e-6, (double)longest * 1e-6);
        for (int a : {0, na / 4, na / 2, 3 * na / 4, na - 1}) {
            const unsigned long long b = (unsigned long long)hr[8 * a + 4] | ((unsigned long long)hr[8 * a + 5] << 32);
            const unsigned long long e = (unsigned long long)hr[8 * a + 6] | ((unsigned long long)hr[8 * a + 7] << 32);
            fprintf(stderr, " n=%lld: %.3f-%.3f", (long long)streams[act[a]].n, (double)(b - t_first) * 1e-6, (double)(e - t_first) * 1e-6);
        }
        fprintf(stderr, "\n");
    }
    if (prof) {
        float a = 0;
        cudaEventElapsedTime(&a, E.t[0], E.t[1]);
        const double b = (double)(t_last - t_first) * 1e-6;          // the push kernel's own %globaltimer stamps
        std::lock_guard<std::mutex> pg(g_prof_mu);
        g_prof_ms[PROF_STARTFREQ] += a; g_prof_bytes[PROF_STARTFREQ] += 56.0 * total_n; g_prof_launches[PROF_STARTFREQ] += 1;
        g_prof_ms[PROF_PUSH] += b; g_prof_bytes[PROF_PUSH] += 4.0 * total_n + 8.0 * (double)used_words; g_prof_launches[PROF_PUSH] += 1;
    }
    const uint32_t* words = static_cast<const uint32_t*>(E.h_out.p);
    for (int a = 0; a < na; ++a) {
        flashgmm_stream& s = streams[act[a]];
        const uint32_t need = hr[8 * a], off = hr[8 * a + 1], over = hr[8 * a + 2];
        s.bytes_len = (size_t)need * 4;
        if (over) {
            s.status = fail(FLASHGMM_ENOSPC, "output buffer too small: need %zu bytes, have %zu", s.bytes_len, s.bytes_cap);
            worst = s.status;
        } else {
            memcpy(s.bytes, words + off, s.bytes_len);
        }
    }
    return worst;
}

// Host workers or device warps for the serial push? A host worker does ~4.5 ns per symbol (waves path,
// tail effects included), a device warp ~46 ns but every stream has its own. Streams are sorted by length; the `k` longest go
// to the worker threads, the rest to the device, both at once: k minimises the later finish.
// (One free box: the long streams stay on the host, the many short ones go to the GPU. Several
// ranks sharing few cores: most or all of the batch goes to the GPU.)
int plan_push_split(const flashgmm_ctx* ctx, const flashgmm_stream* streams, int n_streams, int mem, std::vector<int>* order) {
    order->resize(n_streams);
    for (int i = 0; i < n_streams; ++i) (*order)[i] = i;
    std::sort(order->begin(), order->end(), [&](int a, int b) { return streams[a].n > streams[b].n; });
    if (ctx->push_policy == 1 || n_streams < 4) return n_streams;
    if (ctx->push_policy == 2) return 0;
    // page-locked host inputs: the call is bound by PCIe (52 B per symbol), and >= 6 workers push
    // faster than the link delivers; a device part only adds its push as a tail (measured 14.4 ms
    // split against 12.7 ms all-host for 24 images)
    if (mem == FLASHGMM_MEM_HOST && ctx->n_workers >= 6) return n_streams;
    // Not lane-aware on purpose. With several ctxs coding at once (one per host thread) the whole batch is better off on
    // the push kernel -- its latency hides behind the other lanes and the host cores stay free: 6 batches in flight, 16
    // cores, 2500-2800 Msym/s with this split against 2830-3130 with FLASHGMM_PUSH=device -- but deciding that per call
    // from the number of other batch calls in flight made lanes flip between the two paths and halved the rate
    // (1270-1480 Msym/s, measured); the application states it instead (FLASHGMM_PUSH=device when it creates the ctxs of
    // its lanes: bench.py does).
    const double workers = (double)std::max(1, ctx->n_workers);
    double best = 1e300, host_sum = 0;
    int best_k = n_streams;
    for (int k = 0; k <= n_streams; ++k) {
        // host: k longest streams over the workers (never faster than its longest stream)
        const double host = k == 0 ? 0.0
            : std::max(host_sum / std::min(workers, (double)k), (double)streams[(*order)[0]].n) * ctx->push_host_ns;
        const double dev = k == n_streams ? 0.0 : (double)streams[(*order)[k]].n * ctx->push_dev_ns + 2.5e5;
        const double t = std::max(host, dev);
        if (t < best) { best = t; best_k = k; }
        if (k < n_streams) host_sum += (double)streams[(*order)[k]].n;
    }
    return best_k;
}

typedef int (*stream_fn)(flashgmm_ctx*, Lane&, flashgmm_stream&, int, int);

int run_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem, stream_fn fn) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (n_streams < 0 || (n_streams > 0 && !streams)) return fail(FLASHGMM_EINVAL, "bad stream array");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    if (n_streams <= 1 || ctx->n_workers <= 1) {
        int worst = FLASHGMM_OK;
        for (int i = 0; i < n_streams; ++i) {
            streams[i].status = fn(ctx, lane_of(ctx, 0), streams[i], mode, mem);
            if (streams[i].status) worst = streams[i].status;
        }
        return worst;
    }
    // longest stream first, dealt to whichever worker is free
    std::vector<int> order(n_streams);
    for (int i = 0; i < n_streams; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return streams[a].n > streams[b].n; });
    if (!ctx->pool) ctx->pool.reset(new Pool(ctx->n_workers));
    const int nw = std::min(ctx->n_workers, n_streams);
    for (int i = 0; i < nw; ++i) lane_of(ctx, i + 1);
    std::atomic<int> next{0};
    std::vector<std::string> errs(nw);
    std::function<void(int)> work = [&](int wi) {
        cudaSetDevice(ctx->device);
        Lane& L = *ctx->lanes[wi + 1];
        for (;;) {
            int k = next.fetch_add(1);
            if (k >= n_streams) break;
            flashgmm_stream& s = streams[order[k]];
            s.status = fn(ctx, L, s, mode, mem);
            if (s.status) errs[wi] = g_err;
        }
    };
    ctx->pool->run(work, nw);
    int worst = FLASHGMM_OK;
    for (int i = 0; i < n_streams; ++i) if (streams[i].status) worst = streams[i].status;
    if (worst) for (auto& e : errs) if (!e.empty()) { g_err = e; break; }
    return worst;
}

// ---------------- encode a batch: batched CDF waves on the device, push on the workers ----------------
// The worker-thread path of run_batch(encode_stream) launches one small kernel and one copy per
// stream chunk and waits for each: with page-locked host inputs (kernels read them over PCIe) the
// link idles whenever the workers are busy pushing. Here ONE stream of batched startfreq launches
// covers every stream of the batch wave by wave -- a wave is a symbol range counted from the END of
// each stream, because the encoder walks backwards -- and writes the packed records straight into
// page-locked host memory; the workers take (wave, stream) tasks in wave-major order and push each
// as soon as its wave's event has fired and the stream's previous wave is done. PCIe (or the FP32
// pipe) never waits for a push, and a worker makes no CUDA call per task beyond the event wait.
// `gate` (optional): an event the first wave waits for (the device part of a split batch goes first);
// `wait_ctr` (optional): a device counter the first wave waits for on the device (>= wait_target).
// Returns -1 when the batch is not eligible (pageable host tensors), before touching anything.
int encode_batch_waves(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem, cudaEvent_t gate,
                       const unsigned int* wait_ctr, unsigned int wait_target) {
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    if (!ctx->hw) ctx->hw.reset(new HostWaves());
    HostWaves& H = *ctx->hw;
    int rc = H.init();
    if (rc) return rc;

    std::vector<int> act;
    std::vector<GmmParamsView> view(n_streams);
    std::vector<const int32_t*> sym(n_streams, nullptr);
    int worst = FLASHGMM_OK;
    if (mem == FLASHGMM_MEM_HOST)                          // eligibility first: nothing is modified on -1
        for (int i = 0; i < n_streams; ++i) {
            const flashgmm_stream& s = streams[i];
            GmmParamsView v{};
            const int32_t* sp = nullptr;
            if (s.n > 0 && s.scales && s.means && s.weights && s.symbols && !zero_copy_view(s, true, &v, &sp)) return -1;
        }
    int64_t max_n = 0;
    for (int i = 0; i < n_streams; ++i) {
        flashgmm_stream& s = streams[i];
        s.bytes_len = 0;
        s.status = check_gmm_args(s.scales, s.means, s.weights, s.n, s.stride_n, s.stride_k, mode, mem);
        if (!s.status && s.n > 0 && !s.symbols) s.status = fail(FLASHGMM_EINVAL, "null symbols");
        if (!s.status && !s.bytes && s.bytes_cap) s.status = fail(FLASHGMM_EINVAL, "null output buffer");
        if (!s.status && s.n > INT32_MAX / 2) s.status = fail(FLASHGMM_EINVAL, "stream too long for one call");
        if (s.status) { worst = s.status; continue; }
        if (mem == FLASHGMM_MEM_DEVICE) { view[i] = device_view(s, 0); sym[i] = s.symbols; }
        else if (s.n > 0) zero_copy_view(s, true, &view[i], &sym[i]);
        act.push_back(i);
        max_n = std::max(max_n, s.n);
    }
    const int na = (int)act.size();
    if (na == 0) return worst;

    // ---- waves (distances from the end of each stream) and tasks ----
    std::vector<int64_t> wb{0};
    for (int64_t step = 8192; wb.back() < max_n; step = 16384) wb.push_back(wb.back() + step);
    const int nw_waves = (int)wb.size() - 1;
    struct Task { int w, a; int64_t lo, hi; };
    std::vector<Task> tasks;
    std::vector<size_t> t_off;                               // wave-major record layout: one copy per wave
    size_t pk_total = 0;
    std::vector<int> wave_first(nw_waves + 1, 0), wave_blocks(nw_waves, 0), last_wave(na, -1);
    for (int w = 0; w < nw_waves; ++w) {
        wave_first[w] = (int)tasks.size();
        for (int a = 0; a < na; ++a) {
            const int64_t n = streams[act[a]].n;
            if (n <= wb[w]) continue;
            tasks.push_back(Task{w, a, n - std::min(n, wb[w + 1]), n - wb[w]});
            t_off.push_back(pk_total);
            pk_total += (size_t)round_up(tasks.back().hi - tasks.back().lo, 32) + 32;   // 128-byte aligned slots; the word after the records is the bypass flag
            last_wave[a] = w;
        }
    }
    wave_first[nw_waves] = (int)tasks.size();
    const int nt = (int)tasks.size();

    const bool prof = g_prof_on.load(std::memory_order_relaxed) != 0;
    if ((rc = H.need_events((size_t)std::max(1, nw_waves), prof))) return rc;
    t_off.push_back(pk_total);
    CU_TRY(H.h_packed.need(pk_total * 4 + 16));
    CU_TRY(H.d_packed.need(pk_total * 4 + 16));
    CU_TRY(H.h_desc.need((size_t)std::max(1, nt) * sizeof(SfDesc)));
    CU_TRY(H.desc.need((size_t)std::max(1, nt) * sizeof(SfDesc)));
    // records go to device memory and come down with one copy per wave (a kernel writing 4-byte
    // records straight into page-locked host memory managed ~4 GB/s)
    uint32_t* packed = static_cast<uint32_t*>(H.h_packed.p);
    uint32_t* d_packed = static_cast<uint32_t*>(H.d_packed.p);
    SfDesc* h_sf = static_cast<SfDesc*>(H.h_desc.p);
    for (int w = 0; w < nw_waves; ++w) {
        int blocks = 0;
        for (int t = wave_first[w]; t < wave_first[w + 1]; ++t) {
            const Task& k = tasks[t];
            const flashgmm_stream& s = streams[act[k.a]];
            const GmmParamsView& v = view[act[k.a]];
            h_sf[t] = SfDesc{GmmParamsView{v.scales + k.lo * s.stride_n, v.means + k.lo * s.stride_n,
                                           v.weights + k.lo * s.stride_n, s.stride_n, s.stride_k},
                             sym[act[k.a]] + k.lo, d_packed + t_off[t], (int32_t)(k.hi - k.lo), blocks,
                             d_packed + t_off[t] + round_up(k.hi - k.lo, 32)};
            blocks += (int)((k.hi - k.lo + 127) / 128);
        }
        wave_blocks[w] = blocks;
    }
    CU_TRY(cudaMemcpyAsync(H.desc.p, H.h_desc.p, (size_t)nt * sizeof(SfDesc), cudaMemcpyHostToDevice, H.st));
    CU_TRY(launch_gmm_sf_clear_flags(static_cast<const SfDesc*>(H.desc.p), nt, H.st));   // the tasks' bypass tails start empty
    if (gate) CU_TRY(cudaStreamWaitEvent(H.st, gate, 0));
    // the wide kernels below must not start while the device part's one-warp push CTAs are being
    // placed (gmm_push.cu::gmm_wait_started_kernel)
    if (wait_ctr) CU_TRY(launch_gmm_wait_started(wait_ctr, wait_target, H.st));
    for (int w = 0; w < nw_waves; ++w) {
        if (prof) CU_TRY(cudaEventRecord(H.t0[w], H.st));
        CU_TRY(launch_gmm_startfreq_batched(static_cast<const SfDesc*>(H.desc.p) + wave_first[w],
                                            wave_first[w + 1] - wave_first[w], wave_blocks[w], mode, H.st));
        if (prof) CU_TRY(cudaEventRecord(H.t1[w], H.st));
        CU_TRY(cudaEventRecord(H.evk[w], H.st));
        CU_TRY(cudaStreamWaitEvent(H.st_down, H.evk[w], 0));
        const size_t w0 = t_off[wave_first[w]], w1 = t_off[wave_first[w + 1]];
        CU_TRY(cudaMemcpyAsync(packed + w0, d_packed + w0, (w1 - w0) * 4, cudaMemcpyDeviceToHost, H.st_down));
        CU_TRY(cudaEventRecord(H.ev[w], H.st_down));
        g_launches.fetch_add(1, std::memory_order_relaxed);
    }

    static const bool wtrace = env_long("FLASHGMM_TRACE", 0) != 0;
    const auto tr0 = std::chrono::steady_clock::now();
    auto ms_since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tr0).count(); };
    std::vector<double> t_landed(std::max(1, nw_waves), 0.0);
    std::vector<double> acc(4 * 64, 0.0);                     // trace: per worker ms in event wait / order wait / scan / push
    // ---- workers ----
    while ((int)H.bb.size() < na) H.bb.emplace_back(new HostWaves::PaddedBuffer());
    std::vector<RansEncoder64> enc(na);
    for (int a = 0; a < na; ++a) {
        H.bb[a]->b.reset((size_t)(streams[act[a]].n * 9 / 16) + 64);
        enc[a].init(&H.bb[a]->b);
    }
    std::unique_ptr<std::atomic<int>[]> done(new std::atomic<int>[na]);
    for (int a = 0; a < na; ++a) done[a].store(0, std::memory_order_relaxed);
    std::unique_ptr<std::atomic<int>[]> landed(new std::atomic<int>[std::max(1, nw_waves)]);
    for (int w = 0; w < nw_waves; ++w) landed[w].store(0, std::memory_order_relaxed);
    auto finish_stream = [&](int a) {
        flashgmm_stream& s = streams[act[a]];
        enc[a].flush();
        s.bytes_len = H.bb[a]->b.bytes();
        if (s.bytes_len > s.bytes_cap)
            s.status = fail(FLASHGMM_ENOSPC, "output buffer too small: need %zu bytes, have %zu", s.bytes_len, s.bytes_cap);
        else
            memcpy(s.bytes, H.bb[a]->b.data(), s.bytes_len);
    };
    for (int a = 0; a < na; ++a) if (last_wave[a] < 0) finish_stream(a);     // empty streams: flush only
    if (!ctx->pool) ctx->pool.reset(new Pool(ctx->n_workers));
    const int nw = std::max(1, std::min(ctx->n_workers, nt));
    std::atomic<int> next{0};
    std::vector<std::string> errs(nw);
    std::vector<int> cuda_fail(nw, 0);
    std::function<void(int)> work = [&](int wi) {
        cudaSetDevice(ctx->device);
        std::vector<int32_t> tmp;
        for (;;) {
            const int t = next.fetch_add(1);
            if (t >= nt) break;
            const Task& k = tasks[t];
            flashgmm_stream& s = streams[act[k.a]];
            const double c0 = wtrace ? ms_since() : 0.0;
            if (!landed[k.w].load(std::memory_order_acquire)) {
                if (cudaEventSynchronize(H.ev[k.w]) != cudaSuccess) cuda_fail[wi] = 1;
                if (wtrace && t_landed[k.w] == 0.0) t_landed[k.w] = ms_since();
                landed[k.w].store(1, std::memory_order_release);
            }
            const double cA = wtrace ? ms_since() : 0.0;
            // the stream's earlier waves (taken by other workers before this task was) must be pushed first
            for (int spin = 0; done[k.a].load(std::memory_order_acquire) != k.w; ++spin)
                if (spin > 64) std::this_thread::yield();
            const double c1 = wtrace ? ms_since() : 0.0;
            const int64_t cnt = k.hi - k.lo;
            const uint32_t* pk = packed + t_off[t];
            const int32_t* sym_h = nullptr;
            if (mem == FLASHGMM_MEM_HOST) {
                sym_h = s.symbols + k.lo;
            } else {
                // bypass records (rare) need the symbol's value: the kernel left the first few
                // (index, value) pairs behind the records; more than that, fetch the range
                const uint32_t* tail = pk + round_up(cnt, 32);
                if (tail[0] > (uint32_t)kSfBypassSlots) {
                    tmp.resize((size_t)cnt);
                    if (cudaMemcpy(tmp.data(), s.symbols + k.lo, (size_t)cnt * 4, cudaMemcpyDeviceToHost) != cudaSuccess)
                        cuda_fail[wi] = 1;
                    sym_h = tmp.data();
                }
            }
            const double c2 = wtrace ? ms_since() : 0.0;
            if (!cuda_fail[wi]) {
                RansEncoder64 e = enc[k.a];                  // state on this worker's stack while it runs:
                if (sym_h != nullptr) {                      // neighbouring streams' states share cache lines
                    e.push_chunk_reverse(pk, sym_h, cnt);
                } else {
                    const uint32_t* tail = pk + round_up(cnt, 32);
                    e.push_chunk_reverse_with(pk, cnt, [tail](int64_t i) {
                        for (uint32_t j = 0; j < tail[0]; ++j)
                            if (tail[1 + 2 * j] == (uint32_t)i) return (int32_t)tail[2 + 2 * j];
                        return 0;
                    });
                }
                enc[k.a] = e;
            }
            if (wtrace) {
                const double c3 = ms_since();
                acc[4 * wi] += cA - c0; acc[4 * wi + 1] += c1 - cA; acc[4 * wi + 2] += c2 - c1; acc[4 * wi + 3] += c3 - c2;
            }
            if (k.w == last_wave[k.a]) {
                finish_stream(k.a);
                if (s.status) errs[wi] = g_err;
            }
            done[k.a].store(k.w + 1, std::memory_order_release);
        }
    };
    const double t_setup = ms_since();
    ctx->pool->run(work, nw);
    if (wtrace) {
        fprintf(stderr, "[flashgmm trace] encode waves: %d streams, %d tasks, %d workers; workers start %.2f ms, done %.2f ms; waves landed at",
                na, nt, nw, t_setup, ms_since());
        for (int w = 0; w < nw_waves; ++w) fprintf(stderr, " %.2f", t_landed[w]);
        fprintf(stderr, "\n[flashgmm trace]   per worker ms (event wait / order wait / scan / push):");
        for (int i = 0; i < std::min(nw, 64); ++i) fprintf(stderr, " %.2f/%.2f/%.2f/%.2f", acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
        fprintf(stderr, "\n");
    }
    CU_TRY(cudaStreamSynchronize(H.st));
    CU_TRY(cudaStreamSynchronize(H.st_down));
    for (int f : cuda_fail) if (f) return fail(FLASHGMM_ECUDA, "CUDA error in the batched encode path: %s", cudaGetErrorString(cudaGetLastError()));
    if (prof) {
        std::lock_guard<std::mutex> pg(g_prof_mu);
        for (int w = 0; w < nw_waves; ++w) {
            float ms = 0;
            double n_w = 0;
            for (int t = wave_first[w]; t < wave_first[w + 1]; ++t) n_w += (double)(tasks[t].hi - tasks[t].lo);
            if (cudaEventElapsedTime(&ms, H.t0[w], H.t1[w]) == cudaSuccess) {
                g_prof_ms[PROF_STARTFREQ] += ms; g_prof_bytes[PROF_STARTFREQ] += 56.0 * n_w; g_prof_launches[PROF_STARTFREQ] += 1;
            }
        }
    }
    for (int a = 0; a < na; ++a) if (streams[act[a]].status) worst = streams[act[a]].status;
    if (worst) for (auto& e : errs) if (!e.empty()) { g_err = e; break; }
    return worst;
}

}  // namespace

// ============================ extern "C" surface ============================
extern "C" {

int flashgmm_version(void) { return 200; }

int flashgmm_request_hw_queues(int n) {
    if (n < 1 || n > 32) return fail(FLASHGMM_EINVAL, "n must be in [1, 32]");
    char buf[16];
    snprintf(buf, sizeof buf, "%d", n);
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", buf, 0);           // a value set by the user wins
    return FLASHGMM_OK;
}

const char* flashgmm_last_error(void) { return g_err.c_str(); }

uint64_t flashgmm_launch_count(void) { return g_launches.load(); }

void flashgmm_profile_enable(int on) { g_prof_on.store(on ? 1 : 0); }

void flashgmm_profile_reset(void) {
    std::lock_guard<std::mutex> g(g_prof_mu);
    for (int k = 0; k < PROF_KINDS; ++k) { g_prof_ms[k] = 0; g_prof_bytes[k] = 0; g_prof_launches[k] = 0; }
}

int flashgmm_profile_read(int kind, double* total_ms, uint64_t* launches, double* algorithmic_bytes) {
    if (kind < 0 || kind >= PROF_KINDS) return fail(FLASHGMM_EINVAL, "kind must be 0 (startfreq), 1 (cdf rows), 2 (ileave encode), 3 (ileave decode), 4 (pop) or 5 (push)");
    std::lock_guard<std::mutex> g(g_prof_mu);
    if (total_ms) *total_ms = g_prof_ms[kind];
    if (launches) *launches = g_prof_launches[kind];
    if (algorithmic_bytes) *algorithmic_bytes = g_prof_bytes[kind];
    return FLASHGMM_OK;
}

size_t flashgmm_max_encoded_bytes(int64_t n) {
    if (n < 0) n = 0;
    return (size_t)4 * ((size_t)10 * (size_t)n + 16);
}

int flashgmm_ctx_create(int device, flashgmm_ctx** out) {
    if (!out) return fail(FLASHGMM_EINVAL, "null out pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(FLASHGMM_ECUDA, "no CUDA device available (%s); flashgmm_b200 has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count) return fail(FLASHGMM_EINVAL, "device %d out of range [0,%d)", device, count);
    CU_TRY(cudaSetDevice(device));
    CU_TRY(cudaFree(0));
    flashgmm_ctx* c = new flashgmm_ctx();
    c->device = device;
    c->enc_first = std::max<long>(128, env_long("FLASHGMM_ENC_FIRST", 16384)) / 128 * 128;
    c->enc_chunk = std::max<long>(c->enc_first, env_long("FLASHGMM_ENC_CHUNK", 1 << 20)) / 128 * 128;
    c->dec_first = std::max<long>(128, env_long("FLASHGMM_DEC_FIRST", 8192)) / 128 * 128;
    c->dec_chunk_bytes = std::max<long>(1 << 16, env_long("FLASHGMM_DEC_CHUNK_BYTES", 16 << 20));
    long hw = (long)std::thread::hardware_concurrency();
    c->n_workers = (int)std::max<long>(1, env_long("FLASHGMM_WORKERS", std::min<long>(16, hw > 1 ? hw - 1 : 1)));
    c->pop_min_streams = (int)std::max<long>(1, env_long("FLASHGMM_POP_MIN_STREAMS", 4));
    c->pop_rows_budget = std::max<long>(1 << 20, env_long("FLASHGMM_POP_ROWS_BYTES", 256l << 20));
    {
        const char* pp = getenv("FLASHGMM_POP_POLICY");      // device | host | (unset) cost model
        c->pop_policy = (pp && pp[0] == 'd') ? 1 : (pp && pp[0] == 'h') ? 2 : 0;
    }
    {
        const char* pp = getenv("FLASHGMM_PUSH");            // host | device | (unset) auto
        c->push_policy = (pp && pp[0] == 'h') ? 1 : (pp && pp[0] == 'd') ? 2 : 0;
    }
    rcp_table();
    if (!getenv("FLASHGMM_PUSH_HOST_NS")) {
        // calibrate the cost model's host figure on THIS machine's cores: push 2 x 64 Ki synthetic records
        // (~0.5 ms), keep the faster pass; x1.25 for the waves path's tail effects (measured 3.7 vs 4.5 ns)
        std::vector<uint32_t> rec(1 << 16);
        uint32_t lcg = 12345u;
        for (auto& r : rec) {
            lcg = lcg * 1664525u + 1013904223u;
            const uint32_t freq = 1u + ((lcg >> 8) % 6000u), start = (lcg >> 20) % (65535u - freq);
            r = start | (freq << 16);
        }
        double best = 1e30;
        for (int pass = 0; pass < 2; ++pass) {
            BackBuffer bb(rec.size());
            RansEncoder64 enc;
            enc.init(&bb);
            const auto t0 = std::chrono::steady_clock::now();
            enc.push_chunk_reverse(rec.data(), nullptr, (int64_t)rec.size());
            enc.flush();
            best = std::min(best, std::chrono::duration<double, std::nano>(std::chrono::steady_clock::now() - t0).count());
        }
        c->push_host_ns = std::min(40.0, std::max(1.5, 1.25 * best / (double)rec.size()));
    } else {
        c->push_host_ns = std::max(0.5, atof(getenv("FLASHGMM_PUSH_HOST_NS")));
    }
    if (getenv("FLASHGMM_PUSH_DEV_NS")) c->push_dev_ns = std::max(1.0, atof(getenv("FLASHGMM_PUSH_DEV_NS")));
    *out = c;
    return FLASHGMM_OK;
}

void flashgmm_ctx_destroy(flashgmm_ctx* ctx) {
    if (!ctx) return;
    ctx->pool.reset();
    cudaSetDevice(ctx->device);
    for (auto& l : ctx->lanes) l->destroy();
    if (ctx->dd) ctx->dd->destroy();
    if (ctx->de) ctx->de->destroy();
    if (ctx->hw) ctx->hw->destroy();
    ctx->pr_dev.release(); ctx->eb_dev.release(); ctx->eb_host.release(); ctx->pr_host.release();
    ctx->il_scratch.release(); ctx->il_segw.release(); ctx->il_out.release(); ctx->il_err.release();
    ctx->il_host.release();
    delete ctx;
}

int flashgmm_gmm_startfreq(flashgmm_ctx* ctx, const int32_t* symbols_d, const float* scales_d,
                           const float* means_d, const float* weights_d, int64_t n, int64_t stride_n,
                           int64_t stride_k, int mode, uint32_t* packed_out_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    int rc = check_gmm_args(scales_d, means_d, weights_d, n, stride_n, stride_k, mode, FLASHGMM_MEM_DEVICE);
    if (rc) return rc;
    if (n > 0 && (!symbols_d || !packed_out_d)) return fail(FLASHGMM_EINVAL, "null symbols / output");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Lane& L = lane_of(ctx, 0);
    if (!st) { if ((rc = L.init())) return rc; st = L.slot[0].st; }
    GmmParamsView v{scales_d, means_d, weights_d, stride_n, stride_k};
    CU_TRY(launch_gmm_startfreq(symbols_d, v, n, mode, packed_out_d, nullptr, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!stream) CU_TRY(cudaStreamSynchronize(st));
    return FLASHGMM_OK;
}

int flashgmm_gmm_table_row_stride(int32_t R) { return R < 0 ? -1 : table_row_stride(R); }

int flashgmm_gmm_cdf_table(flashgmm_ctx* ctx, const float* scales_d, const float* means_d,
                           const float* weights_d, int64_t n, int64_t stride_n, int64_t stride_k,
                           int mode, int32_t R, uint16_t* table_d, int64_t row_stride, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    int rc = check_gmm_args(scales_d, means_d, weights_d, n, stride_n, stride_k, mode, FLASHGMM_MEM_DEVICE);
    if (rc) return rc;
    if (R < 0 || R > kMaxTableR) return fail(FLASHGMM_EINVAL, "R outside [0, %d]", kMaxTableR);
    if (row_stride != table_row_stride(R))
        return fail(FLASHGMM_EINVAL, "row_stride must be flashgmm_table_row_stride(R) = %d", table_row_stride(R));
    if (n > 0 && !table_d) return fail(FLASHGMM_EINVAL, "null table");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Lane& L = lane_of(ctx, 0);
    if (!st) { if ((rc = L.init())) return rc; st = L.slot[0].st; }
    GmmParamsView v{scales_d, means_d, weights_d, stride_n, stride_k};
    CU_TRY(launch_gmm_cdf_table(v, n, mode, R, table_d, (int)row_stride, table_has_lut(R), st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!stream) CU_TRY(cudaStreamSynchronize(st));
    return FLASHGMM_OK;
}

int flashgmm_gmm_encode(flashgmm_ctx* ctx, const int32_t* symbols, const float* scales,
                        const float* means, const float* weights, int64_t n, int64_t stride_n,
                        int64_t stride_k, int mode, int mem, uint8_t* out, size_t out_cap,
                        size_t* out_len) {
    if (!out_len) return fail(FLASHGMM_EINVAL, "null out_len");
    flashgmm_stream s{};
    s.symbols = symbols; s.scales = scales; s.means = means; s.weights = weights;
    s.n = n; s.stride_n = stride_n; s.stride_k = stride_k;
    s.bytes = out; s.bytes_cap = out_cap;
    int rc = run_batch(ctx, &s, 1, mode, mem, encode_stream);
    *out_len = s.bytes_len;
    return rc;
}

int flashgmm_gmm_decode(flashgmm_ctx* ctx, const uint8_t* encoded, size_t encoded_len,
                        const float* scales, const float* means, const float* weights, int64_t n,
                        int64_t stride_n, int64_t stride_k, int mode, int mem, int32_t max_bs_value,
                        int32_t* symbols_out) {
    flashgmm_stream s{};
    s.scales = scales; s.means = means; s.weights = weights;
    s.n = n; s.stride_n = stride_n; s.stride_k = stride_k;
    s.max_bs_value = max_bs_value;
    s.bytes = const_cast<uint8_t*>(encoded); s.bytes_cap = encoded_len;
    s.symbols_out = symbols_out;
    if (ctx && want_device_pop(ctx, &s, 1)) return decode_batch_device(ctx, &s, 1, mode, mem);
    return run_batch(ctx, &s, 1, mode, mem, decode_stream);
}

int flashgmm_gmm_encode_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (n_streams < 0 || (n_streams > 0 && !streams)) return fail(FLASHGMM_EINVAL, "bad stream array");
    std::vector<int> order;
    const int k_host = plan_push_split(ctx, streams, n_streams, mem, &order);
    static const bool waves = env_long("FLASHGMM_ENC_WAVES", 1) != 0;
    // host part: batched CDF waves + worker pushes; falls back to one kernel per stream chunk for
    // pageable host tensors (-1) or when switched off
    auto host_encode = [&](flashgmm_stream* part, int n_part, cudaEvent_t gate, const unsigned int* wait_ctr = nullptr,
                           unsigned int wait_target = 0) -> int {
        if (waves && n_part >= 2) {
            const int rc = encode_batch_waves(ctx, part, n_part, mode, mem, gate, wait_ctr, wait_target);
            if (rc != -1) return rc;
        }
        return run_batch(ctx, part, n_part, mode, mem, encode_stream);
    };
    if (k_host >= n_streams) return host_encode(streams, n_streams, nullptr);
    // the n_streams - k_host shortest streams on the device (own CUDA stream) while the
    // k_host longest are pushed by the worker threads
    std::vector<flashgmm_stream> dev_part, host_part;
    for (int j = k_host; j < n_streams; ++j) dev_part.push_back(streams[order[j]]);
    for (int j = 0; j < k_host; ++j) host_part.push_back(streams[order[j]]);
    // the calling thread queues the device part (asynchronous), takes its share of the host part,
    // then collects the device part
    std::lock_guard<std::mutex> dg(ctx->de_mu);
    static const bool btrace = env_long("FLASHGMM_TRACE", 0) != 0;
    const auto bt0 = std::chrono::steady_clock::now();
    auto bms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - bt0).count(); };
    int dev_rc = encode_batch_device_begin(ctx, dev_part.data(), (int)dev_part.size(), mode, mem);
    const double bt_begin = bms();
    const bool dev_queued = dev_rc == FLASHGMM_OK;
    // host inputs: the device part's parameters cross PCIe first (its push then runs while the host
    // part's cross), instead of both halving the link
    cudaEvent_t gate = (dev_queued && mem == FLASHGMM_MEM_HOST && !ctx->de->act.empty()) ? ctx->de->ev_sf : nullptr;
    // ... and its startfreq kernel always does (event), then its push warps are placed (device counter)
    const bool dev_live = dev_queued && !ctx->de->act.empty();
    if (dev_live) gate = ctx->de->ev_sf;
    const unsigned int* wait_ctr = dev_live ? reinterpret_cast<const unsigned int*>(static_cast<const unsigned long long*>(ctx->de->result.p) + 1) : nullptr;
    int host_rc = host_part.empty() ? FLASHGMM_OK
                                    : host_encode(host_part.data(), (int)host_part.size(), gate, wait_ctr,
                                                  dev_live ? (unsigned int)ctx->de->act.size() : 0u);
    const double bt_host = bms();
    if (dev_queued) {
        const std::string host_err = host_rc ? g_err : std::string();
        dev_rc = encode_batch_device_finish(ctx, dev_part.data());
        if (btrace) fprintf(stderr, "[flashgmm trace] encode batch: device part queued %.2f ms, host part done %.2f ms, device part collected %.2f ms\n",
                            bt_begin, bt_host, bms());
        if (host_rc) g_err = host_err;
    } else if (dev_rc < 0) {                                 // pageable host tensors: the workers take them too
        dev_rc = host_encode(dev_part.data(), (int)dev_part.size(), nullptr);
    } else if (ctx->de && ctx->de->st) {
        cudaStreamSynchronize(ctx->de->st);                  // a failed begin may have queued part of its work
    }
    for (int j = 0; j < k_host; ++j) streams[order[j]] = host_part[j];
    for (int j = k_host; j < n_streams; ++j) streams[order[j]] = dev_part[j - k_host];
    return host_rc ? host_rc : dev_rc;
}

int flashgmm_gmm_decode_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int n_streams, int mode, int mem) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (n_streams < 0 || (n_streams > 0 && !streams)) return fail(FLASHGMM_EINVAL, "bad stream array");
    if (want_device_pop(ctx, streams, n_streams)) return decode_batch_device(ctx, streams, n_streams, mode, mem);
    return run_batch(ctx, streams, n_streams, mode, mem, decode_stream);
}

// ---- interleaved multi-lane rANS ----
static int il_seg_log2(int v) { return v == 0 ? 13 : v; }

size_t flashgmm_gmm_ileave_max_bytes(int64_t n, int segment_log2) {
    const int sl = il_seg_log2(segment_log2);
    if (n < 0 || sl < 10 || sl > 16) return 0;
    const size_t n_seg = (size_t)((n + ((int64_t)1 << sl) - 1) >> sl);
    return 32 + 4 * (n_seg + 1) + n_seg * 2 * (size_t)gmm_ileave_scratch_words(sl);
}

int flashgmm_gmm_ileave_info(const uint8_t* enc, size_t len, int64_t* n, int* mode, int32_t* R, int* segment_log2) {
    if (!enc || len < 36) return fail(FLASHGMM_ETRUNC, "interleaved container shorter than its header");
    if (memcmp(enc, "FGIL", 4) != 0 || enc[4] != 1) return fail(FLASHGMM_EINVAL, "not a FGIL v1 container");
    int64_t hn; int32_t hR; uint32_t nseg; uint64_t pay;
    memcpy(&hn, enc + 8, 8); memcpy(&hR, enc + 16, 4); memcpy(&nseg, enc + 20, 4); memcpy(&pay, enc + 24, 8);
    const int sl = enc[6];
    if (enc[5] > 2 || sl < 10 || sl > 16 || hn < 0 || hR < 0 || (uint64_t)((hn + ((int64_t)1 << sl) - 1) >> sl) != nseg)
        return fail(FLASHGMM_EINVAL, "corrupt FGIL header");
    if (len < 32 + 4 * ((size_t)nseg + 1) + pay) return fail(FLASHGMM_ETRUNC, "interleaved container truncated");
    if (n) *n = hn;
    if (mode) *mode = enc[5];
    if (R) *R = hR;
    if (segment_log2) *segment_log2 = sl;
    return FLASHGMM_OK;
}

int flashgmm_gmm_ileave_encode(flashgmm_ctx* ctx, const int32_t* symbols, const float* scales,
                               const float* means, const float* weights, int64_t n, int64_t stride_n,
                               int64_t stride_k, int mode, int mem, int32_t R, int segment_log2,
                               uint8_t* out, size_t out_cap, size_t* out_len) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (!out_len) return fail(FLASHGMM_EINVAL, "null out_len");
    int rc = check_gmm_args(scales, means, weights, n, stride_n, stride_k, mode, mem);
    if (rc) return rc;
    if (n > 0 && !symbols) return fail(FLASHGMM_EINVAL, "null symbols");
    const int sl = il_seg_log2(segment_log2);
    if (sl < 10 || sl > 16) return fail(FLASHGMM_EINVAL, "segment_log2 must be 0 (default) or in [10, 16]");
    if (R < 0 || R > (1 << 22)) return fail(FLASHGMM_EINVAL, "R outside [0, 2^22]");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    Lane& L = lane_of(ctx, 0);
    if ((rc = L.init())) return rc;
    Slot& S = L.slot[0];
    const int64_t n_seg = (n + ((int64_t)1 << sl) - 1) >> sl;
    const int64_t cap = gmm_ileave_scratch_words(sl);
    const size_t max_bytes = flashgmm_gmm_ileave_max_bytes(n, sl);
    CU_TRY(ctx->il_scratch.need((size_t)std::max<int64_t>(1, n_seg) * cap * 2));
    CU_TRY(ctx->il_segw.need((size_t)(n_seg + 1) * 4));
    CU_TRY(ctx->il_out.need(max_bytes + 64));
    CU_TRY(ctx->il_err.need(64));
    CU_TRY(ctx->il_host.need(64));

    flashgmm_stream s{};
    s.symbols = symbols; s.scales = scales; s.means = means; s.weights = weights;
    s.n = n; s.stride_n = stride_n; s.stride_k = stride_k;
    GmmParamsView view{scales, means, weights, stride_n, stride_k};
    const int32_t* sym_d = symbols;
    if (mem == FLASHGMM_MEM_HOST && n > 0) {
        if ((rc = stage_chunk(S, s, 0, n, true, &view, &sym_d))) return rc;
    }
    unsigned long long* d_total = static_cast<unsigned long long*>(ctx->il_err.p);
    prof_begin(S, PROF_IL_ENC, 52.0 * (double)n);
    CU_TRY(launch_gmm_ileave_encode(sym_d, view, n, mode, R, sl, static_cast<uint16_t*>(ctx->il_scratch.p),
                                    static_cast<uint32_t*>(ctx->il_segw.p), static_cast<uint8_t*>(ctx->il_out.p),
                                    d_total, S.st));
    prof_end(S);
    g_launches.fetch_add(n_seg > 0 ? 3 : 1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(ctx->il_host.p, d_total, 8, cudaMemcpyDeviceToHost, S.st));
    CU_TRY(cudaStreamSynchronize(S.st));
    prof_collect(S);
    const size_t total = (size_t)*static_cast<unsigned long long*>(ctx->il_host.p);
    *out_len = total;
    if (total > out_cap) return fail(FLASHGMM_ENOSPC, "output buffer too small: need %zu bytes, have %zu", total, out_cap);
    if (!out) return fail(FLASHGMM_EINVAL, "null output buffer");
    CU_TRY(cudaMemcpyAsync(out, ctx->il_out.p, total, cudaMemcpyDeviceToHost, S.st));
    CU_TRY(cudaStreamSynchronize(S.st));
    return FLASHGMM_OK;
}

int flashgmm_gmm_ileave_decode(flashgmm_ctx* ctx, const uint8_t* encoded, size_t encoded_len,
                               const float* scales, const float* means, const float* weights, int64_t n,
                               int64_t stride_n, int64_t stride_k, int mem, int32_t* symbols_out) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    int64_t hn; int mode, sl; int32_t R;
    int rc = flashgmm_gmm_ileave_info(encoded, encoded_len, &hn, &mode, &R, &sl);
    if (rc) return rc;
    if (hn != n) return fail(FLASHGMM_EINVAL, "container holds %lld symbols, caller expects %lld", (long long)hn, (long long)n);
    if ((rc = check_gmm_args(scales, means, weights, n, stride_n, stride_k, mode, mem))) return rc;
    if (n > 0 && !symbols_out) return fail(FLASHGMM_EINVAL, "null symbols_out");
    if (n == 0) return FLASHGMM_OK;
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    Lane& L = lane_of(ctx, 0);
    if ((rc = L.init())) return rc;
    Slot& S = L.slot[0];
    const uint32_t n_seg = (uint32_t)((n + ((int64_t)1 << sl) - 1) >> sl);
    uint64_t pay; memcpy(&pay, encoded + 24, 8);
    const size_t used = 32 + 4 * ((size_t)n_seg + 1) + (size_t)pay;
    // host-side validation of the offset table (the kernel bounds-checks word reads inside a segment)
    {
        const uint8_t* ot = encoded + 32;
        uint32_t prev = 0;
        for (uint32_t k = 0; k <= n_seg; ++k) {
            uint32_t o; memcpy(&o, ot + 4 * (size_t)k, 4);
            if ((o & 3u) || o < prev || o > pay || (k == 0 && o != 0))
                return fail(FLASHGMM_EINVAL, "corrupt FGIL offset table at segment %u", k);
            prev = o;
        }
        if (prev != pay) return fail(FLASHGMM_EINVAL, "corrupt FGIL offset table (total)");
    }
    CU_TRY(ctx->il_out.need(used + 64));
    CU_TRY(ctx->il_err.need(64));
    CU_TRY(ctx->il_host.need(64));
    CU_TRY(cudaMemcpyAsync(ctx->il_out.p, encoded, used, cudaMemcpyHostToDevice, S.st));
    CU_TRY(cudaMemsetAsync(ctx->il_err.p, 0, 16, S.st));
    flashgmm_stream s{};
    s.scales = scales; s.means = means; s.weights = weights; s.n = n; s.stride_n = stride_n; s.stride_k = stride_k;
    GmmParamsView view{scales, means, weights, stride_n, stride_k};
    int32_t* out_d = symbols_out;
    if (mem == FLASHGMM_MEM_HOST) {
        const int32_t* unused;
        if ((rc = stage_chunk(S, s, 0, n, false, &view, &unused))) return rc;
        CU_TRY(S.d_out.need((size_t)n * 4));
        out_d = static_cast<int32_t*>(S.d_out.p);
    }
    uint32_t* d_err = static_cast<uint32_t*>(ctx->il_err.p) + 2;
    prof_begin(S, PROF_IL_DEC, 52.0 * (double)n);
    CU_TRY(launch_gmm_ileave_decode(static_cast<const uint8_t*>(ctx->il_out.p), view, n, mode, R, sl, n_seg, out_d,
                                    d_err, S.st));
    prof_end(S);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(ctx->il_host.p, d_err, 4, cudaMemcpyDeviceToHost, S.st));
    if (mem == FLASHGMM_MEM_HOST)
        CU_TRY(cudaMemcpyAsync(symbols_out, out_d, (size_t)n * 4, cudaMemcpyDeviceToHost, S.st));
    CU_TRY(cudaStreamSynchronize(S.st));
    prof_collect(S);
    if (*static_cast<uint32_t*>(ctx->il_host.p) != 0)
        return fail(FLASHGMM_ETRUNC, "interleaved stream is truncated or corrupt (segment reads ran past their end)");
    return FLASHGMM_OK;
}

// ---- likelihood ----
static int lk_stream(flashgmm_ctx* ctx, void* stream, cudaStream_t* st) {
    *st = static_cast<cudaStream_t>(stream);
    if (*st) return FLASHGMM_OK;
    Lane& L = lane_of(ctx, 0);
    int rc = L.init();
    if (rc) return rc;
    *st = L.slot[0].st;
    return FLASHGMM_OK;
}

int flashgmm_gmm_likelihood_fwd(flashgmm_ctx* ctx, const float* x_d, const float* noise_d, int quant,
                                const float* scales_d, const float* means_d, const float* weights_d,
                                int64_t B, int64_t plane, int K, float scale_bound, float lik_bound,
                                float* y_out_d, float* lik_out_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || plane < 0) return fail(FLASHGMM_EINVAL, "negative size");
    if (K < 1 || K > 8) return fail(FLASHGMM_EINVAL, "K must be in [1, 8], got %d", K);
    if (quant < 0 || quant > 2) return fail(FLASHGMM_EINVAL, "quant must be 0, 1 or 2");
    if (B * plane > 0 && (!x_d || !scales_d || !means_d || !weights_d || !y_out_d || !lik_out_d || (quant == 1 && !noise_d)))
        return fail(FLASHGMM_EINVAL, "null pointer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st;
    int rc = lk_stream(ctx, stream, &st);
    if (rc) return rc;
    CU_TRY(launch_gmm_likelihood_fwd(x_d, noise_d, quant, scales_d, means_d, weights_d, B, plane, K, scale_bound,
                                     lik_bound, y_out_d, lik_out_d, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!stream) CU_TRY(cudaStreamSynchronize(st));
    return FLASHGMM_OK;
}

int flashgmm_gmm_likelihood_bwd(flashgmm_ctx* ctx, const float* y_d, const float* scales_d,
                                const float* means_d, const float* weights_d, const float* grad_lik_d,
                                int64_t B, int64_t plane, int K, float scale_bound, float lik_bound,
                                float* grad_y_d, float* grad_scales_d, float* grad_means_d,
                                float* grad_weights_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || plane < 0) return fail(FLASHGMM_EINVAL, "negative size");
    if (K < 1 || K > 8) return fail(FLASHGMM_EINVAL, "K must be in [1, 8], got %d", K);
    if (B * plane > 0 && (!y_d || !scales_d || !means_d || !weights_d || !grad_lik_d || !grad_y_d || !grad_scales_d ||
                          !grad_means_d || !grad_weights_d))
        return fail(FLASHGMM_EINVAL, "null pointer");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st;
    int rc = lk_stream(ctx, stream, &st);
    if (rc) return rc;
    CU_TRY(launch_gmm_likelihood_bwd(y_d, scales_d, means_d, weights_d, grad_lik_d, B, plane, K, scale_bound, lik_bound,
                                     grad_y_d, grad_scales_d, grad_means_d, grad_weights_d, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!stream) CU_TRY(cudaStreamSynchronize(st));
    return FLASHGMM_OK;
}

}  // extern "C"
