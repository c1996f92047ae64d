// abi_common.hpp -- what the translation units of the C ABI share: error reporting, grow-only
// buffers, the worker pool, the per-pipeline state objects and struct flashgmm_ctx.
#pragma once
#include <atomic>
#include <chrono>
#include <algorithm>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/flashgmm.h"
#include "gmm_kernels.h"
#include "rans_host.hpp"

using namespace flashgmm;


namespace flashgmm_abi {

extern thread_local std::string g_err;
extern std::atomic<uint64_t> g_launches;

// Optional per-kernel timing (bench.py's live roofline): CUDA events around each
// launch on the launching stream, read back when the chunk is consumed.
enum { PROF_STARTFREQ = 0, PROF_TABLE = 1, PROF_IL_ENC = 2, PROF_IL_DEC = 3, PROF_POP = 4, PROF_PUSH = 5, PROF_KINDS = 6 };
extern std::atomic<int> g_prof_on;
extern std::mutex g_prof_mu;
extern double g_prof_ms[PROF_KINDS];
extern double g_prof_bytes[PROF_KINDS];
extern uint64_t g_prof_launches[PROF_KINDS];

inline int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU_TRY(expr)                                                                          \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return fail(FLASHGMM_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                        __FILE__, __LINE__);                                                  \
    } while (0)

constexpr int kSlots = 3;

inline long env_long(const char* name, long dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atol(v) : dflt;
}

// Grow-only device / pinned buffers.
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t need(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t need(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4;
        cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct Slot {
    cudaStream_t st = nullptr;
    cudaEvent_t ev = nullptr;
    cudaEvent_t k0 = nullptr, k1 = nullptr;   // kernel timing (profiling only)
    int prof_kind = -1;
    double prof_bytes = 0;
    DevBuf d_in;      // staged symbols + parameters (host-memory callers)
    DevBuf d_out;     // packed words / table rows
    PinBuf h_out;     // D2H landing zone
    PinBuf h_sym;     // lazily fetched symbols (device-memory callers, bypass only)
};

struct Lane {
    Slot slot[kSlots];
    // ONE CUDA stream per lane, shared by its slots (the slots only rotate buffers and events):
    // the whole library then stays below the 32 hardware queues it asks for, and no worker's
    // copy ever queues behind another stream's long kernel (see flashgmm_more_hw_queues).
    cudaStream_t st = nullptr;
    BackBuffer bb;
    PinBuf h_dec;     // decoded symbols before H2D (device-memory callers)
    bool ready = false;
    int init() {
        if (ready) return FLASHGMM_OK;
        CU_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (auto& s : slot) {
            s.st = st;
            CU_TRY(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming));
            CU_TRY(cudaEventCreate(&s.k0));
            CU_TRY(cudaEventCreate(&s.k1));
        }
        ready = true;
        return FLASHGMM_OK;
    }
    void destroy() {
        if (st) cudaStreamDestroy(st);
        st = nullptr;
        for (auto& s : slot) {
            if (s.ev) cudaEventDestroy(s.ev);
            if (s.k0) cudaEventDestroy(s.k0);
            if (s.k1) cudaEventDestroy(s.k1);
            s.d_in.release(); s.d_out.release(); s.h_out.release(); s.h_sym.release();
            s.st = nullptr; s.ev = nullptr;
        }
        h_dec.release();
        ready = false;
    }
};

// Minimal persistent worker pool: run(fn, count) executes fn(worker_index) on
// `count` workers and waits.
class Pool {
public:
    explicit Pool(int n) : n_(n) {
        for (int i = 0; i < n; ++i) th_.emplace_back([this, i] { loop(i); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> g(m_); stop_ = true; ++gen_; }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    int size() const { return n_; }
    void run(const std::function<void(int)>& fn, int count) {
        std::unique_lock<std::mutex> g(m_);
        fn_ = &fn; want_ = std::min(count, n_); done_ = 0; ++gen_;
        cv_.notify_all();
        cv_done_.wait(g, [this] { return done_ == want_; });
        fn_ = nullptr;
    }
private:
    void loop(int i) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* fn;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                if (i >= want_) continue;
                fn = fn_;
            }
            (*fn)(i);
            { std::lock_guard<std::mutex> g(m_); ++done_; }
            cv_done_.notify_one();
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, cv_done_;
    const std::function<void(int)>* fn_ = nullptr;
    int want_ = 0, done_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

// Buffers, streams and events of the device-resident decode path.
constexpr int kRowBufs = 3;        // per group: rows run up to 2 work items ahead of the group's pop kernel
constexpr int kPopGroups = 6;      // streams are grouped by length; every group has its own pop stream
struct DevDecode {
    cudaStream_t st_rows = nullptr, st_copy = nullptr, st_words = nullptr;
    cudaStream_t st_pop[kPopGroups] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::vector<cudaEvent_t> ev_rows, ev_pop, ev_up;      // one per chunk (grown on demand)
    std::vector<cudaEvent_t> t0, t1;                      // kernel timing pairs (profiling only)
    cudaEvent_t ev_words[kPopGroups] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // a group's bitstreams are up
    DevBuf rows[kPopGroups][kRowBufs], words, desc, state, params, out, claim;
    int n_sm = 0;
    PinBuf h_desc, h_words, h_state;
    int init() {
        if (st_rows) return FLASHGMM_OK;
        int lo_prio = 0, hi_prio = 0;                     // the serial pop kernel is the critical path
        CU_TRY(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
        for (auto& sp : st_pop) CU_TRY(cudaStreamCreateWithPriority(&sp, cudaStreamNonBlocking, hi_prio));
        CU_TRY(cudaStreamCreateWithPriority(&st_rows, cudaStreamNonBlocking, hi_prio < lo_prio ? hi_prio + 1 : lo_prio));
        CU_TRY(cudaStreamCreateWithPriority(&st_copy, cudaStreamNonBlocking, lo_prio));
        CU_TRY(cudaStreamCreateWithPriority(&st_words, cudaStreamNonBlocking, hi_prio));
        for (auto& e : ev_words) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        return FLASHGMM_OK;
    }
    int need_events(size_t n) {
        while (ev_rows.size() < n) {
            cudaEvent_t a, b, c;
            CU_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
            CU_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
            CU_TRY(cudaEventCreateWithFlags(&c, cudaEventDisableTiming));
            ev_rows.push_back(a); ev_pop.push_back(b); ev_up.push_back(c);
        }
        return FLASHGMM_OK;
    }
    int need_timers(size_t n) {
        while (t0.size() < n) {
            cudaEvent_t a, b;
            CU_TRY(cudaEventCreate(&a));
            CU_TRY(cudaEventCreate(&b));
            t0.push_back(a); t1.push_back(b);
        }
        return FLASHGMM_OK;
    }
    void destroy() {
        for (auto v : {&ev_rows, &ev_pop, &ev_up, &t0, &t1}) { for (auto e : *v) cudaEventDestroy(e); v->clear(); }
        for (auto& e : ev_words) { if (e) cudaEventDestroy(e); e = nullptr; }
        if (st_rows) cudaStreamDestroy(st_rows);
        for (auto& sp : st_pop) { if (sp) cudaStreamDestroy(sp); sp = nullptr; }
        if (st_copy) cudaStreamDestroy(st_copy);
        if (st_words) cudaStreamDestroy(st_words);
        st_rows = st_copy = st_words = nullptr;
        for (auto& gr : rows) for (auto& r : gr) r.release();
        words.release(); desc.release(); state.release(); claim.release();
        params.release(); out.release(); h_desc.release(); h_words.release(); h_state.release();
    }
};

// Buffers of the device-resident encode path (batched startfreq kernel -> push kernel).
struct DevEncode {
    cudaStream_t st = nullptr;
    cudaEvent_t t[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_sf = nullptr;          // the device part's CDF pass is done (its inputs have crossed PCIe)
    DevBuf packed, scratch, desc, rcp, result;
    PinBuf h_desc, h_result, h_out;
    bool rcp_ready = false;
    // a batch in flight between encode_batch_device_begin and _finish
    std::vector<int> act;
    int worst = FLASHGMM_OK;
    bool prof = false;
    double total_n = 0;
    int init() {
        if (st) return FLASHGMM_OK;
        CU_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (auto& e : t) CU_TRY(cudaEventCreate(&e));
        CU_TRY(cudaEventCreateWithFlags(&ev_sf, cudaEventDisableTiming));
        return FLASHGMM_OK;
    }
    void destroy() {
        if (st) cudaStreamDestroy(st);
        for (auto& e : t) if (e) { cudaEventDestroy(e); e = nullptr; }
        if (ev_sf) { cudaEventDestroy(ev_sf); ev_sf = nullptr; }
        st = nullptr; rcp_ready = false;
        packed.release(); scratch.release(); desc.release(); rcp.release(); result.release();
        h_desc.release(); h_result.release(); h_out.release();
    }
};

// Buffers of the batched host-push path (one startfreq launch per wave of all streams, records
// written straight into page-locked host memory, worker threads push them as the waves land).
struct HostWaves {
    cudaStream_t st = nullptr, st_down = nullptr;    // CDF kernels | copies of their records to the host
    std::vector<cudaEvent_t> ev, evk, t0, t1;        // per wave: records are on the host | kernel done
    DevBuf desc, d_packed;
    PinBuf h_desc, h_packed;
    // one word buffer per stream, reused across calls; each on its own cache lines (workers on
    // neighbouring streams would otherwise fight over the line holding the write cursors)
    struct alignas(128) PaddedBuffer { BackBuffer b; };
    std::vector<std::unique_ptr<PaddedBuffer>> bb;
    int init() {
        if (st) return FLASHGMM_OK;
        CU_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        CU_TRY(cudaStreamCreateWithFlags(&st_down, cudaStreamNonBlocking));
        return FLASHGMM_OK;
    }
    int need_events(size_t n, bool timers) {
        while (ev.size() < n) {
            cudaEvent_t e, k;
            CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            CU_TRY(cudaEventCreateWithFlags(&k, cudaEventDisableTiming));
            ev.push_back(e); evk.push_back(k);
        }
        while (timers && t0.size() < n) {
            cudaEvent_t a, b;
            CU_TRY(cudaEventCreate(&a));
            CU_TRY(cudaEventCreate(&b));
            t0.push_back(a); t1.push_back(b);
        }
        return FLASHGMM_OK;
    }
    void destroy() {
        for (auto v : {&ev, &evk, &t0, &t1}) { for (auto e : *v) cudaEventDestroy(e); v->clear(); }
        if (st) cudaStreamDestroy(st);
        if (st_down) cudaStreamDestroy(st_down);
        st = st_down = nullptr;
        desc.release(); d_packed.release(); h_desc.release(); h_packed.release();
        bb.clear();
    }
};

}  // namespace flashgmm_abi

using namespace flashgmm_abi;

struct flashgmm_ctx {
    int device = 0;
    std::mutex mu;
    std::vector<std::unique_ptr<Lane>> lanes;   // lane 0 = caller's thread; 1.. = workers
    std::unique_ptr<Pool> pool;
    int64_t enc_chunk = 1 << 20, enc_first = 16384;
    int64_t dec_chunk_bytes = 16 << 20, dec_first = 8192;
    int n_workers = 8;
    // interleaved coder (whole-stream buffers, lane 0's first stream)
    DevBuf il_scratch, il_segw, il_out, il_err;
    PinBuf il_host;
    // device-resident decode (rows kernel -> pop kernel, nothing but symbols leaves the GPU)
    std::unique_ptr<DevDecode> dd;
    int pop_min_streams = 4;          // batches at least this large may decode on the device
    int pop_policy = 0;               // 0 cost model (want_device_pop), 1 always device, 2 always host
    int64_t pop_rows_budget = 256ll << 20;
    // device-resident encode (startfreq kernel -> push kernel): chosen per call by a cost model
    std::unique_ptr<DevEncode> de;
    std::unique_ptr<HostWaves> hw;
    DevBuf pr_dev;                    // compress prologue: extrema codes + bitmap
    PinBuf pr_host;
    DevBuf eb_dev;                    // hyper-latent table coder: records on the device
    PinBuf eb_host;                   //   ... and their page-locked landing zone / decoded symbols
    std::mutex de_mu;                 // the device encode path owns its buffers and stream: it runs beside the workers
    int push_policy = 0;              // 0 auto, 1 always host workers, 2 always device
    double push_host_ns = 4.5, push_dev_ns = 46.0;   // per symbol: one host worker (calibrated at ctx creation) / one device warp (follows the push kernel's own stamps)
};
