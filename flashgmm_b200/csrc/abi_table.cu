// abi_table.cu -- host halves of the C ABI: the serial rANS push / pop over packed records and
// rows, the table-driven coder kept for the hyper-latent z branch (single calls, buffered encoder
// and streaming decoder sessions, and the batched EntropyBottleneck compress / decompress whose
// CDF lookup runs on the GPU, gmm_table.cu), and the scalar single-Gaussian search coder.
#include "abi_common.hpp"

#include <cmath>

extern "C" {

// ---- host-only halves ----
static int emit(BackBuffer& bb, uint8_t* out, size_t cap, size_t* out_len) {
    *out_len = bb.bytes();
    if (*out_len > cap) return fail(FLASHGMM_ENOSPC, "output buffer too small: need %zu bytes, have %zu", *out_len, cap);
    memcpy(out, bb.data(), *out_len);
    return FLASHGMM_OK;
}

int flashgmm_rans_push_packed(const uint32_t* packed, const int32_t* symbols, int64_t n, uint8_t* out,
                              size_t out_cap, size_t* out_len) {
    if (!out_len || (n > 0 && !packed) || n < 0 || (!out && out_cap)) return fail(FLASHGMM_EINVAL, "bad argument");
    if (!symbols)
        for (int64_t i = 0; i < n; ++i)
            if ((packed[i] >> 16) == 0) return fail(FLASHGMM_EINVAL, "bypass record at %lld needs symbols", (long long)i);
    BackBuffer bb((size_t)(n * 9 / 16) + 64);
    RansEncoder64 enc;
    enc.init(&bb);
    enc.push_chunk_reverse(packed, symbols, n);
    enc.flush();
    return emit(bb, out, out_cap, out_len);
}

int flashgmm_rans_pop_table(const uint8_t* encoded, size_t encoded_len, const uint16_t* table,
                            int64_t row_stride, int64_t n, int32_t R, int32_t* symbols_out) {
    if (!encoded || n < 0 || (n > 0 && (!table || !symbols_out)) || R < 0 || row_stride != table_row_stride(R))
        return fail(FLASHGMM_EINVAL, "bad argument");
    RansDecoder64 dec;
    dec.init(encoded, encoded_len);
    pop_rows(dec, table, row_stride, R, n, symbols_out);
    if (dec.overrun) return fail(FLASHGMM_ETRUNC, "bitstream truncated");
    return FLASHGMM_OK;
}

static int check_table_args(const int32_t* indexes, int64_t n, const int32_t* cdfs, int64_t n_cdfs,
                            int64_t cdf_stride, const int32_t* sizes, const int32_t* offsets) {
    if (n < 0 || n_cdfs < 0 || (n > 0 && (!indexes || !cdfs || !sizes || !offsets)))
        return fail(FLASHGMM_EINVAL, "bad argument");
    for (int64_t c = 0; c < n_cdfs; ++c)
        if (sizes[c] < 2 || sizes[c] > cdf_stride) return fail(FLASHGMM_EINVAL, "cdf_sizes[%lld]=%d invalid", (long long)c, sizes[c]);
    for (int64_t i = 0; i < n; ++i)
        if (indexes[i] < 0 || indexes[i] >= n_cdfs) return fail(FLASHGMM_EINVAL, "indexes[%lld]=%d out of range", (long long)i, indexes[i]);
    return FLASHGMM_OK;
}

int flashgmm_rans_encode_table(const int32_t* symbols, const int32_t* indexes, int64_t n,
                               const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, uint8_t* out,
                               size_t out_cap, size_t* out_len) {
    if (!out_len || (n > 0 && !symbols)) return fail(FLASHGMM_EINVAL, "bad argument");
    int rc = check_table_args(indexes, n, cdfs, n_cdfs, cdf_stride, cdf_sizes, offsets);
    if (rc) return rc;
    BackBuffer bb((size_t)(n * 9 / 16) + 64);
    RansEncoder64 enc;
    enc.init(&bb);
    for (int64_t i = n - 1; i >= 0; --i) {             // reverse walk of rans_interface.cpp:343-398
        bb.reserve(24);
        const int32_t ci = indexes[i];
        const int32_t* cdf = cdfs + (int64_t)ci * cdf_stride;
        const int32_t max_value = cdf_sizes[ci] - 2;
        int32_t value = symbols[i] - offsets[ci];
        uint32_t raw = 0;
        if (value < 0) { raw = (uint32_t)(-2 * value - 1); value = max_value; }
        else if (value >= max_value) { raw = (uint32_t)(2 * (value - max_value)); value = max_value; }
        if (value == max_value) {
            int nb = 0;
            while (nb < 8 && (raw >> (nb * 4)) != 0) ++nb;
            for (int j = nb - 1; j >= 0; --j) enc.put_bits4((raw >> (4 * j)) & kBypassMax);
            enc.put_bits4((uint32_t)nb);
        }
        const uint32_t start = (uint32_t)cdf[value] & 0xffffu;
        const uint32_t freq = (uint32_t)(cdf[value + 1] - cdf[value]) & 0xffffu;
        if (freq == 0) return fail(FLASHGMM_EINVAL, "zero-frequency symbol at %lld (invalid CDF)", (long long)i);
        enc.put(start, freq);
    }
    enc.flush();
    return emit(bb, out, out_cap, out_len);
}

static void decode_table_run(RansDecoder64& dec, const int32_t* indexes, int64_t n, const int32_t* cdfs,
                             int64_t cdf_stride, const int32_t* cdf_sizes, const int32_t* offsets,
                             int32_t* symbols_out) {
    for (int64_t i = 0; i < n; ++i) {
        const int32_t ci = indexes[i];
        const int32_t* cdf = cdfs + (int64_t)ci * cdf_stride;
        const int32_t sz = cdf_sizes[ci], max_value = sz - 2;
        const uint32_t cum = dec.get();
        const int32_t* it = std::lower_bound(cdf, cdf + sz, (int32_t)(cum + 1));
        int32_t sidx = (int32_t)(it - cdf) - 1;
        if (sidx < 0) sidx = 0;
        if (sidx > max_value) sidx = max_value;
        dec.advance((uint32_t)cdf[sidx], (uint32_t)(cdf[sidx + 1] - cdf[sidx]));
        int32_t value = sidx;
        if (value == max_value) {
            int32_t vb = (int32_t)dec.get_bits4(), nb = vb;
            while (vb == (int32_t)kBypassMax && !dec.overrun) { vb = (int32_t)dec.get_bits4(); nb += vb; }
            int32_t raw = 0;
            for (int32_t j = 0; j < nb && !dec.overrun; ++j) raw |= (int32_t)(dec.get_bits4() << ((j * 4) & 31));
            value = raw >> 1;
            if (raw & 1) value = -value - 1; else value += max_value;
        }
        symbols_out[i] = value + offsets[ci];
    }
}

int flashgmm_rans_decode_table(const uint8_t* encoded, size_t encoded_len, const int32_t* indexes,
                               int64_t n, const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, int32_t* symbols_out) {
    if (!encoded || (n > 0 && !symbols_out)) return fail(FLASHGMM_EINVAL, "bad argument");
    int rc = check_table_args(indexes, n, cdfs, n_cdfs, cdf_stride, cdf_sizes, offsets);
    if (rc) return rc;
    RansDecoder64 dec;
    dec.init(encoded, encoded_len);
    decode_table_run(dec, indexes, n, cdfs, cdf_stride, cdf_sizes, offsets, symbols_out);
    if (dec.overrun) return fail(FLASHGMM_ETRUNC, "bitstream truncated");
    return FLASHGMM_OK;
}

// ---- sessions ----
struct flashgmm_buffered_encoder {
    struct Sym { uint16_t start, range; uint8_t bypass; };     // rans_interface.hpp:48-52
    std::vector<Sym> syms;
    inline void raw_nibbles(uint32_t raw) {                     // count prefix + nibbles, RI:372-397 / 524-552
        int32_t nb = 0;
        while (nb < 8 && (raw >> (nb * 4)) != 0) ++nb;
        int32_t v = nb;
        while (v >= (int32_t)kBypassMax) { syms.push_back({(uint16_t)kBypassMax, (uint16_t)(kBypassMax + 1), 1}); v -= kBypassMax; }
        syms.push_back({(uint16_t)v, (uint16_t)(v + 1), 1});
        for (int32_t j = 0; j < nb; ++j) {
            uint16_t q = (uint16_t)((raw >> (j * 4)) & kBypassMax);
            syms.push_back({q, (uint16_t)(q + 1), 1});
        }
    }
};

int flashgmm_benc_create(flashgmm_buffered_encoder** out) {
    if (!out) return fail(FLASHGMM_EINVAL, "null out pointer");
    *out = new flashgmm_buffered_encoder();
    rcp_table();
    return FLASHGMM_OK;
}
void flashgmm_benc_destroy(flashgmm_buffered_encoder* enc) { delete enc; }
int64_t flashgmm_benc_entries(const flashgmm_buffered_encoder* enc) { return enc ? (int64_t)enc->syms.size() : -1; }

int flashgmm_benc_push_table(flashgmm_buffered_encoder* enc, const int32_t* symbols, const int32_t* indexes,
                             int64_t n, const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                             const int32_t* cdf_sizes, const int32_t* offsets) {
    if (!enc || (n > 0 && !symbols)) return fail(FLASHGMM_EINVAL, "bad argument");
    int rc = check_table_args(indexes, n, cdfs, n_cdfs, cdf_stride, cdf_sizes, offsets);
    if (rc) return rc;
    for (int64_t i = 0; i < n; ++i) {
        const int32_t ci = indexes[i];
        const int32_t* cdf = cdfs + (int64_t)ci * cdf_stride;
        const int32_t max_value = cdf_sizes[ci] - 2;
        int32_t value = symbols[i] - offsets[ci];
        uint32_t raw = 0;
        if (value < 0) { raw = (uint32_t)(-2 * value - 1); value = max_value; }
        else if (value >= max_value) { raw = (uint32_t)(2 * (value - max_value)); value = max_value; }
        const uint16_t range = (uint16_t)(cdf[value + 1] - cdf[value]);
        if (range == 0) return fail(FLASHGMM_EINVAL, "zero-frequency symbol at %lld (invalid CDF)", (long long)i);
        enc->syms.push_back({(uint16_t)cdf[value], range, 0});
        if (value == max_value) enc->raw_nibbles(raw);
    }
    return FLASHGMM_OK;
}

int flashgmm_benc_push_gmm_packed(flashgmm_buffered_encoder* enc, const uint32_t* packed,
                                  const int32_t* symbols, int64_t n) {
    if (!enc || n < 0 || (n > 0 && !packed)) return fail(FLASHGMM_EINVAL, "bad argument");
    for (int64_t i = 0; i < n; ++i) {
        const uint32_t pk = packed[i], freq = pk >> 16;
        if (freq) { enc->syms.push_back({(uint16_t)(pk & 0xffffu), (uint16_t)freq, 0}); continue; }
        if (!symbols) return fail(FLASHGMM_EINVAL, "bypass record at %lld needs symbols", (long long)i);
        enc->syms.push_back({(uint16_t)kMaxCdf, 1, 0});        // RI:513-521
        enc->raw_nibbles((uint32_t)symbols[i]);
    }
    return FLASHGMM_OK;
}

int flashgmm_benc_flush(flashgmm_buffered_encoder* enc, uint8_t* out, size_t out_cap, size_t* out_len) {
    if (!enc || !out_len || (!out && out_cap)) return fail(FLASHGMM_EINVAL, "bad argument");
    BackBuffer bb(enc->syms.size() / 2 + 64);
    RansEncoder64 r;
    r.init(&bb);
    for (size_t k = enc->syms.size(); k-- > 0;) {              // reverse walk, RI:569-577
        if ((k & 1023) == 1023 || k + 1 == enc->syms.size()) bb.reserve(1024 + 8);
        const auto& s = enc->syms[k];
        if (s.bypass) r.put_bits4(s.start); else r.put(s.start, s.range);
    }
    r.flush();
    int rc = emit(bb, out, out_cap, out_len);
    if (rc == FLASHGMM_OK) enc->syms.clear();                  // on ENOSPC the caller retries with a larger buffer
    return rc;
}

struct flashgmm_stream_decoder {
    std::vector<uint8_t> stream;
    RansDecoder64 dec;
};

int flashgmm_sdec_create(const uint8_t* encoded, size_t encoded_len, flashgmm_stream_decoder** out) {
    if (!out || !encoded) return fail(FLASHGMM_EINVAL, "bad argument");
    auto* d = new flashgmm_stream_decoder();
    d->stream.assign(encoded, encoded + encoded_len);          // set_stream copies, RI:887
    d->dec.init(d->stream.data(), d->stream.size());
    *out = d;
    return FLASHGMM_OK;
}
void flashgmm_sdec_destroy(flashgmm_stream_decoder* dec) { delete dec; }

int flashgmm_sdec_decode_table(flashgmm_stream_decoder* dec, const int32_t* indexes, int64_t n,
                               const int32_t* cdfs, int64_t n_cdfs, int64_t cdf_stride,
                               const int32_t* cdf_sizes, const int32_t* offsets, int32_t* symbols_out) {
    if (!dec || (n > 0 && !symbols_out)) return fail(FLASHGMM_EINVAL, "bad argument");
    int rc = check_table_args(indexes, n, cdfs, n_cdfs, cdf_stride, cdf_sizes, offsets);
    if (rc) return rc;
    decode_table_run(dec->dec, indexes, n, cdfs, cdf_stride, cdf_sizes, offsets, symbols_out);
    if (dec->dec.overrun) return fail(FLASHGMM_ETRUNC, "bitstream truncated");
    return FLASHGMM_OK;
}


// ---- EntropyBottleneck.compress / decompress for a batch of images, z on the device ----
// compress  (entropy_models.py:605-611 -> :237-270 -> RI:334-399): gmm_table.cu turns z into 8-byte
//   records on the device, one copy brings them to page-locked memory, the ctx's worker threads run
//   the serial push (one image each). No Python lists, no per-symbol marshalling.
// decompress (entropy_models.py:613-618 -> :272-327 -> RI:619-688): the workers pop the symbols with
//   the table search (std::lower_bound like RI:651), one copy takes them to the device; the caller
//   adds the medians (EntropyModel.dequantize, :185-194).
// Indexes are implicit: element (b, c, i) uses CDF row c (EntropyBottleneck._build_indexes, :556-567).
int flashgmm_eb_compress_batch(flashgmm_ctx* ctx, const float* z_d, const float* medians_d, const int32_t* cdfs_d,
                               const int32_t* cdf_sizes_d, const int32_t* offsets_d, int64_t B, int64_t C, int64_t HW,
                               int64_t cdf_stride, int32_t* symbols_out_d, flashgmm_stream* streams, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || C <= 0 || HW < 0 || cdf_stride < 2 || B * C > 65535 || (B > 0 && !streams))
        return fail(FLASHGMM_EINVAL, "bad shape B=%lld C=%lld HW=%lld", (long long)B, (long long)C, (long long)HW);
    if (B > 0 && HW > 0 && (!z_d || !medians_d || !cdfs_d || !cdf_sizes_d || !offsets_d)) return fail(FLASHGMM_EINVAL, "null tensor");
    if (B == 0) return FLASHGMM_OK;
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t per = C * HW, total = B * per;
    CU_TRY(ctx->eb_dev.need((size_t)total * 8 + 16));
    CU_TRY(ctx->eb_host.need((size_t)total * 8 + 16));
    unsigned long long* d_rec = static_cast<unsigned long long*>(ctx->eb_dev.p);
    uint32_t* d_err = reinterpret_cast<uint32_t*>(d_rec + total);
    CU_TRY(cudaMemsetAsync(d_err, 0, 4, st));
    CU_TRY(launch_eb_records(z_d, medians_d, cdfs_d, cdf_sizes_d, offsets_d, B, (int)C, HW, (int)cdf_stride, d_rec,
                             symbols_out_d, d_err, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(ctx->eb_host.p, d_rec, (size_t)total * 8 + 4, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    const unsigned long long* rec = static_cast<const unsigned long long*>(ctx->eb_host.p);
    if (*reinterpret_cast<const uint32_t*>(rec + total) != 0)
        return fail(FLASHGMM_EINVAL, "a hyper-latent symbol is outside the coder's range (|z| > 2^29)");
    auto push_one = [&](int64_t b, std::string* err) {
        flashgmm_stream& s = streams[b];
        BackBuffer bb((size_t)(per * 9 / 16) + 64);
        RansEncoder64 enc;
        enc.init(&bb);
        const unsigned long long* r = rec + b * per;
        for (int64_t i = per - 1; i >= 0; --i) {              // reverse walk of RI:343-398
            bb.reserve(24);
            const unsigned long long v = r[i];
            if (v >> 63) {
                const uint32_t raw = (uint32_t)(v >> 32) & 0x7fffffffu;
                int nb = 0;
                while (nb < 8 && (raw >> (nb * 4)) != 0) ++nb;
                for (int j = nb - 1; j >= 0; --j) enc.put_bits4((raw >> (4 * j)) & kBypassMax);
                enc.put_bits4((uint32_t)nb);
            }
            const uint32_t freq = ((uint32_t)v >> 16) & 0xffffu;
            if (freq == 0) { s.status = FLASHGMM_EINVAL; *err = "zero-frequency symbol (invalid CDF)"; return; }
            enc.put((uint32_t)v & 0xffffu, freq);
        }
        enc.flush();
        s.bytes_len = bb.bytes();
        if (s.bytes_len > s.bytes_cap || !s.bytes) { s.status = FLASHGMM_ENOSPC; *err = "output buffer too small"; return; }
        memcpy(s.bytes, bb.data(), s.bytes_len);
        s.status = FLASHGMM_OK;
    };
    std::vector<std::string> errs((size_t)B);
    if (B == 1 || ctx->n_workers <= 1) {
        for (int64_t b = 0; b < B; ++b) push_one(b, &errs[b]);
    } else {
        if (!ctx->pool) ctx->pool.reset(new Pool(ctx->n_workers));
        std::atomic<int64_t> next{0};
        std::function<void(int)> work = [&](int) {
            for (;;) { const int64_t b = next.fetch_add(1); if (b >= B) break; push_one(b, &errs[b]); }
        };
        ctx->pool->run(work, (int)std::min<int64_t>(ctx->n_workers, B));
    }
    int worst = FLASHGMM_OK;
    for (int64_t b = 0; b < B; ++b) if (streams[b].status) { worst = streams[b].status; g_err = errs[b]; }
    return worst;
}

int flashgmm_eb_decompress_batch(flashgmm_ctx* ctx, flashgmm_stream* streams, int64_t B, const int32_t* cdfs_h,
                                 const int32_t* cdf_sizes_h, const int32_t* offsets_h, int64_t C, int64_t HW,
                                 int64_t cdf_stride, int32_t* symbols_out_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (B < 0 || C <= 0 || HW < 0 || cdf_stride < 2 || (B > 0 && (!streams || !cdfs_h || !cdf_sizes_h || !offsets_h)))
        return fail(FLASHGMM_EINVAL, "bad argument");
    for (int64_t c = 0; c < C; ++c)
        if (cdf_sizes_h[c] < 2 || cdf_sizes_h[c] > cdf_stride) return fail(FLASHGMM_EINVAL, "cdf_sizes[%lld]=%d invalid", (long long)c, cdf_sizes_h[c]);
    if (B == 0 || HW == 0) return FLASHGMM_OK;
    if (!symbols_out_d) return fail(FLASHGMM_EINVAL, "null symbols_out");
    std::lock_guard<std::mutex> g(ctx->mu);
    CU_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t per = C * HW, total = B * per;
    CU_TRY(ctx->eb_host.need((size_t)total * 4));
    int32_t* h_sym = static_cast<int32_t*>(ctx->eb_host.p);
    auto pop_one = [&](int64_t b) {
        flashgmm_stream& s = streams[b];
        if (!s.bytes) { s.status = FLASHGMM_EINVAL; return; }
        RansDecoder64 dec;
        dec.init(s.bytes, s.bytes_cap);
        int32_t* out = h_sym + b * per;
        for (int64_t c = 0; c < C; ++c) {
            const int32_t* cdf = cdfs_h + c * cdf_stride;
            const int32_t sz = cdf_sizes_h[c], max_value = sz - 2, off = offsets_h[c];
            for (int64_t i = 0; i < HW; ++i) {
                const uint32_t cum = dec.get();
                const int32_t* it = std::lower_bound(cdf, cdf + sz, (int32_t)(cum + 1));
                int32_t sidx = (int32_t)(it - cdf) - 1;
                if (sidx < 0) sidx = 0;
                if (sidx > max_value) sidx = max_value;
                dec.advance((uint32_t)cdf[sidx], (uint32_t)(cdf[sidx + 1] - cdf[sidx]));
                int32_t value = sidx;
                if (value == max_value) {                     // RI:662-681
                    int32_t vb = (int32_t)dec.get_bits4(), nb = vb;
                    while (vb == (int32_t)kBypassMax && !dec.overrun) { vb = (int32_t)dec.get_bits4(); nb += vb; }
                    int32_t raw = 0;
                    for (int32_t j = 0; j < nb && !dec.overrun; ++j) raw |= (int32_t)(dec.get_bits4() << ((j * 4) & 31));
                    value = raw >> 1;
                    if (raw & 1) value = -value - 1; else value += max_value;
                }
                out[c * HW + i] = value + off;
            }
        }
        s.status = dec.overrun ? FLASHGMM_ETRUNC : FLASHGMM_OK;
    };
    if (B == 1 || ctx->n_workers <= 1) {
        for (int64_t b = 0; b < B; ++b) pop_one(b);
    } else {
        if (!ctx->pool) ctx->pool.reset(new Pool(ctx->n_workers));
        std::atomic<int64_t> next{0};
        std::function<void(int)> work = [&](int) {
            for (;;) { const int64_t b = next.fetch_add(1); if (b >= B) break; pop_one(b); }
        };
        ctx->pool->run(work, (int)std::min<int64_t>(ctx->n_workers, B));
    }
    CU_TRY(cudaMemcpyAsync(symbols_out_d, h_sym, (size_t)total * 4, cudaMemcpyHostToDevice, st));
    CU_TRY(cudaStreamSynchronize(st));                        // eb_host is reused by the next call
    int worst = FLASHGMM_OK;
    for (int64_t b = 0; b < B; ++b)
        if (streams[b].status) {
            worst = streams[b].status;
            fail(worst, worst == FLASHGMM_ETRUNC ? "hyper bitstream %lld truncated" : "hyper bitstream %lld: bad argument", (long long)b);
        }
    return worst;
}

// EntropyBottleneck._likelihood, EM:434-463 (+ _logits_cumulative EM:421-440), filters (3, 3, 3, 3): device pointers,
// asynchronous on `stream`.
int flashgmm_eb_likelihood(flashgmm_ctx* ctx, const float* values_d, const float* params_d, int64_t C, int64_t N,
                           float* out_d, void* stream) {
    if (!ctx) return fail(FLASHGMM_EINVAL, "null ctx");
    if (C < 0 || N < 0 || C > 65535) return fail(FLASHGMM_EINVAL, "bad shape C=%lld N=%lld", (long long)C, (long long)N);
    if (C > 0 && N > 0 && (!values_d || !params_d || !out_d)) return fail(FLASHGMM_EINVAL, "null tensor");
    CU_TRY(cudaSetDevice(ctx->device));
    CU_TRY(launch_eb_likelihood(values_d, params_d, C, N, out_d, static_cast<cudaStream_t>(stream)));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return FLASHGMM_OK;
}

// ---- scalar single-Gaussian search coder (a14) ----
// RansEncoder.encode_with_indexes(symbols, scales, max_value) / RansDecoder.decode_with_indexes(encoded,
// scales, max_bs_value): RI:401-456, 600-607, 690-764 with the SCALAR approximations RI:148-150, 188-204,
// 216-219. The reference evaluates them with libm on the CPU (std::exp, std::sqrt), one symbol at a time; no
// shipped model calls this overload. It stays a host routine here too, in scalar_gaussian.cpp (built by g++ with
// the reference's own flags so that the compiler contracts the same expressions): its results depend
// on the host's libm, so parity is "unpinned" across machines and checked only against the reference compiled
// on the same box.
extern "C" float flashgmm_scalar_gaussian_cdf(float t, int mode);

static inline uint32_t scalar_q16(float t, int mode) {
    const float v = flashgmm_scalar_gaussian_cdf(t, mode) * 65535.0f;
    // static_cast<uint16_t>(float): cvttss2si then the low 16 bits (0x80000000 for NaN / out of range)
    const int32_t i = (v >= -2147483648.0f && v < 2147483648.0f) ? (int32_t)v : INT32_MIN;
    return (uint32_t)i & 0xffffu;
}

int flashgmm_scalar_gaussian_encode(const int32_t* symbols, const float* scales, int64_t n, int mode, uint8_t* out,
                                    size_t out_cap, size_t* out_len) {
    if (!out_len || n < 0 || (n > 0 && (!symbols || !scales)) || mode < 0 || mode > 2) return fail(FLASHGMM_EINVAL, "bad argument");
    BackBuffer bb((size_t)(n * 9 / 16) + 64);
    RansEncoder64 enc;
    enc.init(&bb);
    for (int64_t i = n - 1; i >= 0; --i) {
        bb.reserve(24);
        const int32_t v = symbols[i];
        const uint32_t c1 = scalar_q16(((float)v - 0.5f) / scales[i], mode);
        const uint32_t c2 = scalar_q16(((float)v - 0.5f + 1.0f) / scales[i], mode);
        const uint32_t pmf = (c2 - c1) & 0xffffu;
        if (pmf == 0) enc.put_bypass_reverse(v);              // marker (65535, 1) + nibbles, RI:417-453
        else enc.put(c1, pmf);
    }
    enc.flush();
    return emit(bb, out, out_cap, out_len);
}

int flashgmm_scalar_gaussian_decode(const uint8_t* encoded, size_t encoded_len, const float* scales, int64_t n,
                                    int mode, int32_t max_bs_value, int32_t* symbols_out) {
    if (!encoded || n < 0 || (n > 0 && (!scales || !symbols_out)) || mode < 0 || mode > 2 || max_bs_value < 0)
        return fail(FLASHGMM_EINVAL, "bad argument");
    RansDecoder64 dec;
    dec.init(encoded, encoded_len);
    for (int64_t i = 0; i < n; ++i) {
        const uint32_t cum = dec.get();
        int32_t value;
        if (cum == kMaxCdf) {
            dec.advance(kMaxCdf, 1);
            value = (int32_t)dec.bypass_raw();
        } else {
            int32_t s = -max_bs_value, e = max_bs_value, mid = 0;
            uint32_t c1 = 0, c2 = 0;
            while (s <= e) {
                mid = s + (e - s) / 2;
                c1 = scalar_q16(((float)mid - 0.5f) / scales[i], mode);
                c2 = scalar_q16(((float)mid - 0.5f + 1.0f) / scales[i], mode);
                if (c1 <= cum && c2 > cum) break;
                else if (c1 > cum) e = mid - 1;
                else s = mid + 1;
            }
            uint32_t pmf = (c2 - c1) & 0xffffu;
            if (pmf == 0 && c1 <= cum) {                      // RI:746-749
                pmf = 1;
                if (c1 + pmf > kMaxCdf) c1 = kMaxCdf - 1;
            }
            // pmf == 0 with c1 > cum: the reference advances with frequency 0 (state collapses); reproduce it
            dec.advance(c1, pmf);
            value = mid;
        }
        symbols_out[i] = value;
        if (dec.overrun) return fail(FLASHGMM_ETRUNC, "bitstream truncated");
    }
    return FLASHGMM_OK;
}

}  // extern "C"
