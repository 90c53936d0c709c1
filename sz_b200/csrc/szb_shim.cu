// szb_shim.cu -- the C-ABI shim (include/szb200.h): per-device context (stream + grow-only workspace), kernel
// launch sequences for one blocked-regression job, stream-section assembly.  No CPU fallback anywhere: every stage of
// the hot path is a kernel from szb_stage1.cuh / szb_wavefront.cuh / szb_huffman.cuh.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -lineinfo -Xcompiler -ffp-contract=off ...
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <chrono>
#include <vector>

#include "../../include/szb200.h"
#include "szb_common.cuh"
#include "szb_huffman.cuh"
#include "szb_stage1.cuh"
#include "szb_chain.cuh"
#include "szb_wavefront.cuh"

using namespace szb;

#define SZ_SCES 0
#define SZ_NSCS -1
#define SZ_DERR -4
#define SZ_MERR -5

static thread_local char g_err[512] = "";
static void set_err(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
static void set_err(const char *fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
#define CU(call)                                                                                              \
    do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { set_err("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); return SZ_NSCS; } } while (0)
#define KCHECK()                                                                                              \
    do { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { set_err("%s:%d kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return SZ_NSCS; } c.launches++; } while (0)

enum Stage { ST_OPT = 0, ST_MEAN, ST_FIT, ST_CHAIN, ST_WAVE, ST_HIST, ST_TREE, ST_ENC, ST_EMIT, ST_PARSE, ST_DECODE, ST_DCHAIN, ST_RECON, ST_JOIN, ST_COUNT };
static const char *k_stage_names[16] = {"opt_intervals", "use_mean_mean", "fit_select", "coeff_chain", "wavefront_quant", "histogram", "huffman_tree",
                                        "encode", "emit", "d_parse", "d_huffman_decode", "d_coeff_chain", "d_wavefront_recon", "wait_coeff_trees_side_stream",
                                        "coeff_trees_kernel_side_stream", "coeff_hist_and_encode_side_stream"};   // [14], [15]: the coefficient huff_build_kernel itself, and the histograms before / the encode launch behind it (they overlap the main stream's stages and are not part of their sum)
extern "C" const char *szb200_stage_name(int i) { return (i >= 0 && i < 16) ? k_stage_names[i] : ""; }

// ---------------------------------------------------------------------------------------------------------------------
struct Buf { void *p = nullptr; size_t cap = 0; };

struct Ctx {
    int dev = -1;
    cudaStream_t stream = nullptr, stream2 = nullptr;   // stream2: side branch (coefficient histograms / trees / bit lengths)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_tree0 = nullptr, ev_tree1 = nullptr;  // around the coefficient-tree kernel on the side stream
    cudaEvent_t ev_side0 = nullptr, ev_side1 = nullptr;  // first and last thing on the side stream
    std::map<std::string, Buf> bufs;
    uint32_t wf_epoch = 0;                               // tag of the most recent wavefront launch (see szb_wavefront.cuh)
    void *pinned = nullptr; size_t pinned_cap = 0;
    unsigned long long *pinned_words = nullptr;          // 32 words of pinned memory for small read-backs that must not block the host
    unsigned char *bounce = nullptr;                     // 2 x BOUNCE_CHUNK pinned bytes: staging of device-to-host copies into pageable memory
    cudaEvent_t ev_bounce[2] = {nullptr, nullptr};
    cudaEvent_t ev[24];
    bool ev_ok = false;
    unsigned int launches = 0;
    int sm_count = 148;
    size_t smem_optin = 0;
    szb200_stats last; bool have_last = false;
    // debug bookkeeping of the last compress job
    Geom dbg_g; int dbg_dtype = -1; unsigned int dbg_intervals = 0; size_t dbg_reg_count = 0;
    // cached sample-position table
    std::vector<uint64_t> pos_cache; size_t pos_key[4] = {0, 0, 0, 0};

    int init(int d) {
        dev = d;
        CU(cudaSetDevice(d));
        CU(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        // the side stream carries the few-CTA serial kernels (coefficient trees and chains): highest priority, so that their CTAs are
        // placed as soon as an SM has room instead of queueing behind the big grids of the main stream
        { int lo_p = 0, hi_p = 0; CU(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p)); CU(cudaStreamCreateWithPriority(&stream2, cudaStreamNonBlocking, hi_p)); }
        CU(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
        for (auto &e : ev) CU(cudaEventCreate(&e));
        CU(cudaEventCreate(&ev_tree0)); CU(cudaEventCreate(&ev_tree1)); CU(cudaEventCreate(&ev_side0)); CU(cudaEventCreate(&ev_side1));
        {   // Kernels launched while huff_build_kernel already sits on its SMs waiting for them (compress_impl) are loaded here: with lazy
            // module loading a kernel's first launch synchronises with the kernels that are running, i.e. with the one that waits for it.
            cudaFuncAttributes fa;
            CU(cudaFuncGetAttributes(&fa, coeff_lattice_kernel<float>)); CU(cudaFuncGetAttributes(&fa, coeff_lattice_kernel<double>));
            CU(cudaFuncGetAttributes(&fa, reg_gather_kernel<float>)); CU(cudaFuncGetAttributes(&fa, reg_gather_kernel<double>));
        }
        ev_ok = true;
        cudaDeviceProp pr; CU(cudaGetDeviceProperties(&pr, d));
        sm_count = pr.multiProcessorCount; smem_optin = pr.sharedMemPerBlockOptin;
        return 0;
    }
    void *get(const char *name, size_t bytes) {
        Buf &b = bufs[name];
        if (b.cap < bytes) {
            if (b.p) cudaFree(b.p);
            size_t want = bytes + bytes / 8 + 256;
            if (cudaMalloc(&b.p, want) != cudaSuccess) { b.p = nullptr; b.cap = 0; set_err("cudaMalloc(%zu) for %s failed", want, name); return nullptr; }
            b.cap = want;
        }
        return b.p;
    }
    void *pin(size_t bytes) {
        if (pinned_cap < bytes) {
            if (pinned) cudaFreeHost(pinned);
            if (cudaMallocHost(&pinned, bytes + 4096) != cudaSuccess) { pinned = nullptr; pinned_cap = 0; return nullptr; }
            pinned_cap = bytes + 4096;
        }
        return pinned;
    }
    int need_pinned_words() {
        if (!pinned_words && cudaMallocHost((void **)&pinned_words, 256) != cudaSuccess) { pinned_words = nullptr; cudaGetLastError(); set_err("pinned alloc failed"); return SZ_NSCS; }
        return 0;
    }
    void release() {
        for (auto &kv : bufs) if (kv.second.p) cudaFree(kv.second.p);
        bufs.clear();
        if (pinned) { cudaFreeHost(pinned); pinned = nullptr; pinned_cap = 0; }
        if (pinned_words) { cudaFreeHost(pinned_words); pinned_words = nullptr; }
        if (bounce) { cudaFreeHost(bounce); bounce = nullptr; }
    }
};

static std::mutex g_mu;
static std::map<int, Ctx *> g_ctx;

// A context (streams + grow-only workspace) exists per (device, lane).  Lane 0 is what every single call uses; the batch entry
// points give each of their worker threads its own lane, so that the jobs of a batch overlap on the GPU.
static thread_local int g_lane = 0;
constexpr int SZB_MAX_LANES = 16;
static Ctx *ctx() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess) { set_err("no CUDA device: %s", cudaGetErrorString(cudaGetLastError())); return nullptr; }
    std::lock_guard<std::mutex> lk(g_mu);
    const int key = d * SZB_MAX_LANES + g_lane;
    auto it = g_ctx.find(key);
    if (it != g_ctx.end()) return it->second;
    Ctx *c = new Ctx();
    if (c->init(d) != 0) { delete c; return nullptr; }
    g_ctx[key] = c;
    return c;
}
extern "C" int szb200_set_lane(int lane) { if (lane < 0 || lane >= SZB_MAX_LANES) return SZ_NSCS; g_lane = lane; return 0; }
extern "C" int szb200_current_device(void) { int d = -1; if (cudaGetDevice(&d) != cudaSuccess) { cudaGetLastError(); return -1; } return d; }

extern "C" int szb200_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }
extern "C" int szb200_set_device(int device) { CU(cudaSetDevice(device)); return ctx() ? 0 : SZ_NSCS; }
extern "C" const char *szb200_last_error(void) { return g_err; }
extern "C" void szb200_release(void) { Ctx *c = ctx(); if (c) { cudaStreamSynchronize(c->stream); c->release(); } }
extern "C" void *szb200_stream(void) { Ctx *c = ctx(); return c ? (void *)c->stream : nullptr; }
// Ordering against the caller's own streams: the library's work runs on its private non-blocking stream, which nothing orders
// against the stream a caller's producer kernel (or an NCCL collective) was launched on.
extern "C" int szb200_wait_event(void *cuda_event) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (!cuda_event) return 0;
    CU(cudaStreamWaitEvent(c->stream, (cudaEvent_t)cuda_event, 0));
    return 0;
}
extern "C" int szb200_record_event(void *cuda_event) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (!cuda_event) return SZ_NSCS;
    CU(cudaEventRecord((cudaEvent_t)cuda_event, c->stream));
    return 0;
}
extern "C" int szb200_sync(void) { Ctx *c = ctx(); if (!c) return SZ_NSCS; CU(cudaStreamSynchronize(c->stream)); return 0; }
extern "C" int szb200_last_stats(szb200_stats *out) { Ctx *c = ctx(); if (!c || !c->have_last) return SZ_NSCS; *out = c->last; return 0; }

static bool is_pinned_host(const void *p) {
    cudaPointerAttributes at;
    const bool pinned = cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
}
static int ensure_bounce(Ctx &c);
// Host-to-device copy from PAGEABLE memory through the pinned bounce buffers: two host threads fill a chunk while the DMA of the
// previous one runs (the driver's own staging of pageable memory is a single-threaded copy).
static int h2d_bounced(Ctx &c, void *dst, const void *src, size_t bytes);

extern "C" void *szb200_stage_input(const void *host_ptr, size_t bytes) {
    Ctx *c = ctx(); if (!c) return nullptr;
    void *d = c->get("in_stage", bytes); if (!d) return nullptr;
    if (bytes >= ((size_t)4 << 20) && !is_pinned_host(host_ptr)) { if (h2d_bounced(*c, d, host_ptr, bytes)) return nullptr; return d; }
    if (cudaMemcpyAsync(d, host_ptr, bytes, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { set_err("H2D failed"); return nullptr; }
    return d;
}

extern "C" void *szb200_output_buffer(size_t bytes) { Ctx *c = ctx(); return c ? c->get("out_stage", bytes) : nullptr; }

// Device-to-host copy into PAGEABLE memory through two pinned bounce buffers: the DMA of chunk i+1 runs while the CPU moves
// chunk i to its destination.  (A direct cudaMemcpy into pageable memory reaches ~4.7 GB/s here; the DMA alone does > 50 GB/s,
// so the copy becomes as fast as the host can write its own pages.)
constexpr size_t BOUNCE_CHUNK = (size_t)8 << 20;
static int ensure_bounce(Ctx &c) {
    if (!c.bounce) {
        if (cudaMallocHost((void **)&c.bounce, 2 * BOUNCE_CHUNK) != cudaSuccess) { c.bounce = nullptr; cudaGetLastError(); }
        else { cudaEventCreateWithFlags(&c.ev_bounce[0], cudaEventDisableTiming); cudaEventCreateWithFlags(&c.ev_bounce[1], cudaEventDisableTiming); }
    }
    return c.bounce ? 0 : SZ_NSCS;
}
static int copy_threads() {                                     // host threads that share one bounce chunk (SZB200_COPY_THREADS, default 4, at most 8)
    static int n = 0;
    if (n == 0) { const char *e = getenv("SZB200_COPY_THREADS"); n = e ? atoi(e) : 4; if (n < 1) n = 1; if (n > 8) n = 8; }
    return n;
}
// The host side of a bounced copy is a memcpy between pinned and pageable memory that one core cannot do at the DMA's pace: several
// threads share every chunk.  They are kept (creating them per chunk cost more than the copy: 7 x ~20 us for 8 MB at 20 GB/s), asleep
// between copies and spinning for the next chunk while one is under way.  One bounced copy uses them at a time; a second one, from
// another host thread, copies its chunks alone.
struct CopyPool {
    std::mutex owner;                                  // held by the bounced copy that uses the pool
    std::mutex m; std::condition_variable cv;
    std::vector<std::thread> th;
    std::atomic<int> active{0};
    std::atomic<unsigned> gen{0};
    std::atomic<int> pending{0};
    unsigned char *d = nullptr; const unsigned char *s = nullptr; size_t part = 0, n = 0; int parts = 0;
    void worker(int k) {
        unsigned seen = 0;
        for (;;) {
            { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return active.load(std::memory_order_acquire) != 0; }); }
            while (active.load(std::memory_order_acquire)) {
                const unsigned g = gen.load(std::memory_order_acquire);
                if (g == seen) { std::this_thread::yield(); continue; }
                seen = g;
                if (k < parts) { const size_t off = (size_t)k * part, len = (k == parts - 1) ? n - off : part; memcpy(d + off, s + off, len); }
                pending.fetch_sub(1, std::memory_order_release);
            }
        }
    }
    void begin() {
        if (th.empty()) for (int k = 1; k < copy_threads(); k++) { th.emplace_back([this, k] { worker(k); }); th.back().detach(); }
        { std::lock_guard<std::mutex> l(m); active.store(1, std::memory_order_release); }
        cv.notify_all();
    }
    void end() { active.store(0, std::memory_order_release); }
    void copy(void *dst, const void *src, size_t bytes) {
        const int nt = copy_threads();
        if (bytes < ((size_t)1 << 20) || nt == 1) { memcpy(dst, src, bytes); return; }
        d = (unsigned char *)dst; s = (const unsigned char *)src; n = bytes; parts = nt; part = (bytes / (size_t)nt) & ~(size_t)4095;
        pending.store(nt - 1, std::memory_order_relaxed);
        gen.fetch_add(1, std::memory_order_release);
        memcpy(d, s, part);                                // part 0 is this thread's
        while (pending.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }
};
static CopyPool &g_pool = *new CopyPool;             // never destroyed: its threads outlive static destruction
struct PoolUse {                                       // scope of one bounced copy
    bool mine;
    PoolUse() : mine(copy_threads() > 1 && g_pool.owner.try_lock()) { if (mine) g_pool.begin(); }
    ~PoolUse() { if (mine) { g_pool.end(); g_pool.owner.unlock(); } }
    void copy(void *d, const void *s, size_t n) { if (mine) g_pool.copy(d, s, n); else memcpy(d, s, n); }
};
static int h2d_bounced(Ctx &c, void *dst, const void *src, size_t bytes) {
    if (ensure_bounce(c)) { CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.stream)); return 0; }
    const size_t nchunks = (bytes + BOUNCE_CHUNK - 1) / BOUNCE_CHUNK;
    PoolUse pool;
    for (size_t i = 0; i < nchunks; i++) {
        const size_t off = i * BOUNCE_CHUNK, n = std::min(BOUNCE_CHUNK, bytes - off);
        if (i >= 2) CU(cudaEventSynchronize(c.ev_bounce[i & 1]));       // the DMA that last read this buffer has finished
        pool.copy(c.bounce + (i & 1) * BOUNCE_CHUNK, (const unsigned char *)src + off, n);
        CU(cudaMemcpyAsync((unsigned char *)dst + off, c.bounce + (i & 1) * BOUNCE_CHUNK, n, cudaMemcpyHostToDevice, c.stream));
        CU(cudaEventRecord(c.ev_bounce[i & 1], c.stream));
    }
    // the bounce buffers are reused by the next copy: let the last two DMAs finish before returning
    CU(cudaEventSynchronize(c.ev_bounce[0])); if (nchunks > 1) CU(cudaEventSynchronize(c.ev_bounce[1]));
    return 0;
}
static int d2h_bounced(Ctx &c, void *dst, const void *src, size_t bytes) {
    (void)ensure_bounce(c);
    if (!c.bounce) { CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c.stream)); CU(cudaStreamSynchronize(c.stream)); return 0; }
    const size_t nchunks = (bytes + BOUNCE_CHUNK - 1) / BOUNCE_CHUNK;
    auto issue = [&](size_t i) -> int {
        const size_t off = i * BOUNCE_CHUNK, n = std::min(BOUNCE_CHUNK, bytes - off);
        CU(cudaMemcpyAsync(c.bounce + (i & 1) * BOUNCE_CHUNK, (const unsigned char *)src + off, n, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaEventRecord(c.ev_bounce[i & 1], c.stream));
        return 0;
    };
    const bool hprof = getenv("SZB200_HOST_PROFILE") != nullptr;
    double t_wait = 0, t_copy = 0;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    PoolUse pool;
    if (issue(0)) return SZ_NSCS;
    for (size_t i = 0; i < nchunks; i++) {
        if (i + 1 < nchunks && issue(i + 1)) return SZ_NSCS;      // buffer (i+1)&1 was drained at step i-1
        const double t0 = hprof ? now() : 0;
        CU(cudaEventSynchronize(c.ev_bounce[i & 1]));
        const double t1 = hprof ? now() : 0;
        const size_t off = i * BOUNCE_CHUNK, n = std::min(BOUNCE_CHUNK, bytes - off);
        // the host side of the copy is what bounds it (one core writes ~10 GB/s): several threads share every chunk (measured on the
        // 512^3 stream, 38.5 MB: 2.4 ms with two threads, 1.6 ms with four, 1.65 ms with eight)
        pool.copy((unsigned char *)dst + off, c.bounce + (i & 1) * BOUNCE_CHUNK, n);
        if (hprof) { t_wait += t1 - t0; t_copy += now() - t1; }
    }
    if (hprof) fprintf(stderr, "[host-prof] D2H of %zu bytes through the bounce buffers: waited for DMA %.2f ms, host copies %.2f ms\n", bytes, t_wait, t_copy);
    return 0;
}

extern "C" int szb200_copy(void *dst, const void *src, size_t bytes, int kind) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (kind == 1 && bytes >= ((size_t)1 << 20)) {
        cudaPointerAttributes at;
        const bool pinned_dst = cudaPointerGetAttributes(&at, dst) == cudaSuccess && at.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (!pinned_dst) return d2h_bounced(*c, dst, src, bytes);
    }
    cudaMemcpyKind k = kind == 0 ? cudaMemcpyHostToDevice : kind == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    CU(cudaMemcpyAsync(dst, src, bytes, k, c->stream));
    if (kind != 2) CU(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// K1
template <typename T>
static int range_impl(Ctx &c, const T *d, size_t n, double *vmin, double *vmax) {
    int grid = c.sm_count * 4;
    size_t need = ((n * sizeof(T)) / 16 + 511) / 512; if ((size_t)grid > need) grid = (int)(need ? need : 1);
    MinMax<T> *part = (MinMax<T> *)c.get("mm_part", sizeof(MinMax<T>) * (size_t)(c.sm_count * 4 + 1) + 64);
    if (!part) return SZ_NSCS;
    MinMax<T> *res = part + c.sm_count * 4;
    unsigned int *done = (unsigned int *)c.get("mm_done", 64); if (!done) return SZ_NSCS;
    CU(cudaMemsetAsync(done, 0, 64, c.stream));
    range_kernel<T><<<grid, 512, 0, c.stream>>>(d, (uint64_t)n, part, done, res); KCHECK();
    MinMax<T> h;
    CU(cudaMemcpyAsync(&h, res, sizeof(h), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    // sign of zero: the reference keeps the FIRST occurrence among equal values (ref: dataCompression.c:108-113)
    T ext[2] = {h.mn, h.mx};
    for (int q = 0; q < 2; q++) {
        if (ext[q] == (T)0) {
            unsigned long long *idx = (unsigned long long *)done;
            CU(cudaMemsetAsync(idx, 0xff, 8, c.stream));
            first_equal_kernel<T><<<c.sm_count * 2, 256, 0, c.stream>>>(d, (uint64_t)n, (T)0, idx); KCHECK();
            unsigned long long hi = 0;
            CU(cudaMemcpyAsync(&hi, idx, 8, cudaMemcpyDeviceToHost, c.stream));
            CU(cudaStreamSynchronize(c.stream));
            if (hi != ~0ULL) { CU(cudaMemcpyAsync(&ext[q], d + hi, sizeof(T), cudaMemcpyDeviceToHost, c.stream)); CU(cudaStreamSynchronize(c.stream)); }
            CU(cudaMemsetAsync(done, 0, 64, c.stream));
        }
    }
    *vmin = (double)ext[0]; *vmax = (double)ext[1];
    return 0;
}

extern "C" int szb200_range(int dtype, const void *d_data, size_t n, double *vmin, double *vmax) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (n == 0) { set_err("empty input"); return SZ_NSCS; }
    return dtype == 0 ? range_impl<float>(*c, (const float *)d_data, n, vmin, vmax) : range_impl<double>(*c, (const double *)d_data, n, vmin, vmax);
}

// ---------------------------------------------------------------------------------------------------------------------
// host-side helpers (scalar arithmetic of the reference that is not data-parallel)
static unsigned int round_up_pow2(unsigned int base) {   // ref: conf.c:35-46
    base -= 1; base |= base >> 1; base |= base >> 2; base |= base >> 4; base |= base >> 8; base |= base >> 16; return base + 1;
}

// positions of the sampling-mean walk (ref: sz_float.c:6400-6420 3D; :5409-5418 2D) -- shape-only, cached
static void sample_positions(Ctx &c, const Geom &g, std::vector<uint64_t> &out) {
    size_t key[4] = {(size_t)g.dim, g.ax[0].n, g.ax[1].n, g.ax[2].n};
    if (memcmp(key, c.pos_key, sizeof(key)) == 0 && !c.pos_cache.empty()) { out = c.pos_cache; return; }
    out.clear();
    const uint64_t len = g.N;
    const uint64_t md = (uint64_t)(int)sqrt((double)len);
    if (g.dim == 2) { for (uint64_t p = 0; p < len; p += md) out.push_back(p); }
    else {
        const uint64_t r3 = g.ax[2].n, r23 = (uint64_t)g.ax[1].n * g.ax[2].n;
        uint64_t pos = 0, oc = 0, oc2 = 0;
        while (pos < len) {
            out.push_back(pos);
            pos += md; oc += md; oc2 += md;
            if (oc >= r3) { oc = 0; pos -= 1; }
            if (oc2 >= r23) { oc2 = 0; pos -= 1; }
        }
    }
    c.pos_cache = out; memcpy(c.pos_key, key, sizeof(key));
}

// number of sample ROWS the interval-optimiser walk visits before its first address >= N
static uint64_t opt_row_limit(const Geom &g, uint32_t sd) {
    const uint64_t N = g.N, n1 = g.ax[1].n, n2 = g.ax[2].n, s0 = n1 * n2;
    if (g.dim == 3) {
        const uint64_t n0 = g.ax[0].n;
        if (n0 < 2 || n1 < 2) return 0;
        uint64_t total = (n0 - 1) * (n1 - 1);
        // rows whose first sample could spill past N: only those within (sd + s0) elements of the end
        uint64_t safe_i = 1;
        if (N > (uint64_t)sd + 2 * s0) safe_i = (N - sd - s0) / s0;        // rows with i < safe_i cannot reach N
        if (safe_i < 1) safe_i = 1;
        uint64_t r = (safe_i - 1) * (n1 - 1);
        for (; r < total; r++) {
            uint64_t i = 1 + r / (n1 - 1), j = 1 + r % (n1 - 1);
            int64_t c0 = (r == 0) ? (int64_t)sd - 2 : (int64_t)(sd - (uint32_t)((i + j) % sd));
            if (i * s0 + j * n2 + (uint64_t)c0 >= N) return r;
        }
        return total;
    } else {
        if (n1 < 2) return 0;
        uint64_t total = n1 - 1;
        uint64_t r = 0;
        if (N > (uint64_t)sd + 2 * n2) { uint64_t safe_i = (N - sd - n2) / n2; if (safe_i > 1) r = safe_i - 1; }
        for (; r < total; r++) {
            uint64_t i = 1 + r;
            int64_t c0 = (r == 0) ? (int64_t)sd - 1 : (int64_t)(sd - (uint32_t)(i % sd));
            if (i * n2 + (uint64_t)c0 >= N) return r;
        }
        return total;
    }
}

// anti-diagonal pencil order for the wavefront ticket dispenser
static void pencil_order(uint32_t A, uint32_t B, std::vector<uint32_t> &order) {
    order.clear(); order.reserve((size_t)A * B);
    for (uint32_t d = 0; d + 1 < A + B; d++) {
        uint32_t a_lo = d >= B ? d - B + 1 : 0, a_hi = d < A ? d : A - 1;
        for (uint32_t a = a_lo; a <= a_hi; a++) order.push_back(a * B + (d - a));
    }
}

static void put_be32(unsigned char *b, uint32_t v) { b[0] = v >> 24; b[1] = v >> 16; b[2] = v >> 8; b[3] = v; }
static void put_be64(unsigned char *b, uint64_t v) { for (int i = 0; i < 8; i++) b[i] = (unsigned char)(v >> (56 - 8 * i)); }
static uint32_t get_be32(const unsigned char *b) { return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3]; }
static uint64_t get_be64(const unsigned char *b) { uint64_t v = 0; for (int i = 0; i < 8; i++) v = (v << 8) | b[i]; return v; }
template <typename T> static void put_beT(unsigned char *b, T v) {
    if (sizeof(T) == 4) { uint32_t u; memcpy(&u, &v, 4); put_be32(b, u); } else { uint64_t u; memcpy(&u, &v, 8); put_be64(b, u); }
}
template <typename T> static T get_beT(const unsigned char *b) {
    T v; if (sizeof(T) == 4) { uint32_t u = get_be32(b); memcpy(&v, &u, 4); } else { uint64_t u = get_be64(b); memcpy(&v, &u, 8); } return v;
}

// stream assembly: a list of pieces, each either host bytes (copied into a pinned arena first) or a device section
struct Piece { bool dev; size_t off_or_ptr; const void *dptr; size_t len; };
struct Emitter {
    std::vector<unsigned char> host; std::vector<Piece> pieces; size_t total = 0;
    unsigned char *h(size_t n) { size_t o = host.size(); host.resize(o + n); pieces.push_back({false, o, nullptr, n}); total += n; return host.data() + o; }
    void d(const void *p, size_t n) { if (n) { pieces.push_back({true, 0, p, n}); total += n; } }
};

// One launch assembles the body: grid.y = piece, grid.x strides over the piece.  Host-made pieces (a few header bytes each) travel in
// the kernel's arguments; device sections are copied 16 destination-aligned bytes per thread from whatever alignment the source has.
constexpr int EMIT_MAX_PIECES = 40, EMIT_INLINE = 1024;
struct EmitPiece { const unsigned char *src; unsigned long long dst_off, len; };       // src == nullptr: `len` bytes of inl[] from inl_off
struct EmitArgs { EmitPiece p[EMIT_MAX_PIECES]; unsigned int inl_off[EMIT_MAX_PIECES]; unsigned char inl[EMIT_INLINE]; unsigned char *out; };
__global__ void __launch_bounds__(256) emit_kernel(const __grid_constant__ EmitArgs A) {
    const EmitPiece P = A.p[blockIdx.y];
    unsigned char *d = A.out + P.dst_off;
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (unsigned long long)gridDim.x * blockDim.x;
    if (P.src == nullptr) { for (unsigned long long i = tid; i < P.len; i += nthr) d[i] = A.inl[A.inl_off[blockIdx.y] + i]; return; }
    const unsigned char *s = P.src;
    unsigned long long head = (16u - (unsigned)((uintptr_t)d & 15u)) & 15u; if (head > P.len) head = P.len;
    const unsigned long long nvec = (P.len - head) / 16, tail0 = head + nvec * 16;
    for (unsigned long long i = tid; i < head; i += nthr) d[i] = s[i];
    for (unsigned long long i = tail0 + tid; i < P.len; i += nthr) d[i] = s[i];
    const unsigned char *sv = s + head;
    const unsigned int shift = (unsigned)((uintptr_t)sv & 3u), sel = 0x3210u + 0x1111u * shift;
    const uint32_t *sw = reinterpret_cast<const uint32_t *>(sv - shift);              // aligned; reads at most 3 bytes before / after the section
    uint4 *dv = reinterpret_cast<uint4 *>(d + head);
    for (unsigned long long j = tid; j < nvec; j += nthr) {
        const uint32_t *w = sw + 4 * j;
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2], w3 = w[3], w4 = shift ? w[4] : 0u;
        dv[j] = make_uint4(__byte_perm(w0, w1, sel), __byte_perm(w1, w2, sel), __byte_perm(w2, w3, sel), __byte_perm(w3, w4, sel));
    }
}

static int emit_all(Ctx &c, Emitter &E, void *dst, int dst_on_device) {
    // The body is always assembled in device memory; a host destination then receives it with ONE device-to-host copy.  (Copying ~20
    // sections separately into pageable host memory cost 17 ms for a 38 MB body; ~20 device-to-device copies 0.2 ms of launches.)
    unsigned char *out = dst_on_device ? (unsigned char *)dst : (unsigned char *)c.get("emit_stage", E.total + 64);
    if (!out) return SZ_NSCS;
    if (E.pieces.size() > (size_t)EMIT_MAX_PIECES || E.host.size() > (size_t)EMIT_INLINE) { set_err("emit: too many pieces"); return SZ_NSCS; }
    EmitArgs A; memset(&A, 0, sizeof(A));
    A.out = out;
    memcpy(A.inl, E.host.data(), E.host.size());
    size_t o = 0, longest = 0; int n = 0;
    for (auto &p : E.pieces) {
        A.p[n].src = p.dev ? (const unsigned char *)p.dptr : nullptr; A.p[n].dst_off = o; A.p[n].len = p.len; A.inl_off[n] = p.dev ? 0u : (unsigned int)p.off_or_ptr;
        o += p.len; longest = std::max(longest, p.len); n++;
    }
    if (n) {
        const unsigned gx = (unsigned)std::min<size_t>(std::max<size_t>((longest / 16 + 255) / 256, 1), (size_t)c.sm_count * 8);
        emit_kernel<<<dim3(gx, n), 256, 0, c.stream>>>(A); KCHECK();
    }
    if (!dst_on_device) CU(cudaMemcpyAsync(dst, out, E.total, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
struct TreeWs {              // device workspace + descriptor for one Huffman tree
    HuffTreeDesc d; uint32_t *hist;
};

static int tree_ws(Ctx &c, int idx, uint32_t nbins, TreeWs &w) {
    char name[32]; snprintf(name, sizeof(name), "tree%d", idx);
    const uint32_t cap = 2 * nbins + 2;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    size_t o_hist = take(4ull * nbins), o_left = take(4ull * cap), o_right = take(4ull * cap), o_sizeL = take(4ull * cap), o_sym = take(4ull * cap),
           o_pre = take(4ull * cap), o_code_n = take(8ull * cap), o_len_n = take(cap), o_front = take(8ull * cap), o_heap = take(8ull * cap + 64),
           o_code = take(8ull * nbins), o_clen = take(nbins), o_tree = take(16 + 13ull * cap), o_out = take(64);
    unsigned char *base = (unsigned char *)c.get(name, off); if (!base) return SZ_NSCS;
    w.hist = (uint32_t *)(base + o_hist);
    w.d.hist = w.hist; w.d.nbins = nbins; w.d.capacity = cap; w.d.wait_flag = nullptr;
    w.d.n_left = (uint32_t *)(base + o_left); w.d.n_right = (uint32_t *)(base + o_right); w.d.n_sizeL = (uint32_t *)(base + o_sizeL); w.d.n_sym = (uint32_t *)(base + o_sym);
    w.d.n_pre = (uint32_t *)(base + o_pre); w.d.n_code = (uint64_t *)(base + o_code_n); w.d.n_len = (uint8_t *)(base + o_len_n);
    w.d.frontier = (uint32_t *)(base + o_front); w.d.heap_g = (uint32_t *)(base + o_heap);
    w.d.code = (uint64_t *)(base + o_code); w.d.clen = (uint8_t *)(base + o_clen); w.d.tree_bytes = base + o_tree; w.d.out = (uint32_t *)(base + o_out);
    return 0;
}

// exclusive scan helper (tile sums -> offsets -> apply); `slot` selects a private tile-sum buffer so that scans on different
// streams never share scratch
static int launch_scan(Ctx &c, cudaStream_t st, int slot, const uint32_t *in, uint64_t n, unsigned long long *out, unsigned long long *total) {
    if (n == 0) { if (total) CU(cudaMemsetAsync(total, 0, 8, st)); return 0; }
    const uint32_t ntiles = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    char nm[32]; snprintf(nm, sizeof(nm), "scan_tiles%d", slot);
    unsigned long long *tiles = (unsigned long long *)c.get(nm, 8ull * ntiles + 64); if (!tiles) return SZ_NSCS;
    scan_tile_sum_kernel<uint32_t><<<ntiles, SCAN_TB, 0, st>>>(in, n, tiles); KCHECK();
    scan_tile_offsets_kernel<<<1, SCAN_TB, 0, st>>>(tiles, ntiles, total); KCHECK();
    scan_apply_kernel<uint32_t><<<ntiles, SCAN_TB, 0, st>>>(in, n, tiles, out); KCHECK();
    return 0;
}

static int launch_hist(Ctx &c, cudaStream_t st, const uint16_t *s, uint64_t n, uint32_t nbins, uint32_t *hist) {
    CU(cudaMemsetAsync(hist, 0, 4ull * nbins, st));
    if (n == 0) return 0;
    int use_smem = nbins <= 16384;
    size_t sm = use_smem ? 4ull * nbins : 0;
    if (sm > 48 * 1024) CU(cudaFuncSetAttribute(hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    uint64_t want = (n / 8 + 511) / 512; int grid = (int)std::min<uint64_t>(std::max<uint64_t>(want, 1), (uint64_t)c.sm_count * 4);
    hist_kernel<<<grid, 512, sm, st>>>(s, n, nbins, hist, use_smem); KCHECK();
    return 0;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn)p;
        else cudaGetLastError();
    });
    return fn;
}
// the merged wavefront input `xin` ([rows = n0 * n1][pitch]) as a tensor {pitch, n1, n0} (2D: {pitch, n1}); box = what one warp's
// pencil needs of it for two chunks: {2 CH, PJ, PI} (2D: {2 CH, PJ}).  This GPU's memory only; no swizzle, out-of-bounds = zero.
template <typename T, int DIM>
static int make_xin_map(CUtensorMap *tm, const T *d_xin, const Geom &g, uint32_t pitch) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) { set_err("cuTensorMapEncodeTiled is not available from this driver"); return SZ_NSCS; }
    const cuuint64_t dims[3] = {pitch, g.ax[1].n, g.ax[0].n};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch * sizeof(T), (cuuint64_t)pitch * sizeof(T) * g.ax[1].n};
    const cuuint32_t box[3] = {(cuuint32_t)WfSmem<T, DIM>::HALF, (cuuint32_t)Pencil<DIM>::PJ, (cuuint32_t)Pencil<DIM>::PI};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(tm, sizeof(T) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, DIM == 3 ? 3 : 2, (void *)d_xin, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_err("cuTensorMapEncodeTiled failed (%d) for xin %u x %u x pitch %u", (int)r, g.ax[0].n, g.ax[1].n, pitch); return SZ_NSCS; }
    return 0;
}

template <typename T, int DIM>
static int launch_wavefront(Ctx &c, const Geom &g, bool decomp, int use_mean, int cap, const unsigned char *indicator, const T *x, uint16_t *codes,
                            T *recon, const T *coeff_last, T eb, T recip, T mean, int write_all) {
    constexpr int PI = Pencil<DIM>::PI, PJ = Pencil<DIM>::PJ;
    WfParams P; memset(&P, 0, sizeof(P));
    P.g = g; P.cap = cap; P.radius = cap / 2; P.cap_l = cap - 2; P.use_mean = use_mean;
    P.pencils_a = (g.ax[0].n + PI - 1) / PI; P.pencils_b = (g.ax[1].n + PJ - 1) / PJ; P.num_pencils = P.pencils_a * P.pencils_b;
    std::vector<uint32_t> order; pencil_order(P.pencils_a, P.pencils_b, order);
    uint32_t *d_order = (uint32_t *)c.get("wf_order", 4ull * P.num_pencils); if (!d_order) return SZ_NSCS;
    // tagged boundary words: zeroed when (re)allocated, afterwards every launch uses a fresh epoch as its tag
    const size_t bnd_bytes = wf_bnd_bytes<T, DIM>(P.num_pencils, g.ax[2].n);
    unsigned long long *d_bnd;
    {
        Buf &bb = c.bufs["wf_bnd"];
        void *before = bb.p; const size_t cap_before = bb.cap;
        d_bnd = (unsigned long long *)c.get("wf_bnd", bnd_bytes); if (!d_bnd) return SZ_NSCS;
        bool fresh = (bb.p != before) || (bb.cap != cap_before);
        if (++c.wf_epoch == 0) { c.wf_epoch = 1; fresh = true; }
        if (fresh) CU(cudaMemsetAsync(d_bnd, 0, bb.cap, c.stream));
    }
    unsigned int *d_misc = (unsigned int *)c.get("wf_misc", 64); if (!d_misc) return SZ_NSCS;
    // pageable source: the runtime stages it before returning, so `order` may die at scope exit
    CU(cudaMemcpyAsync(d_order, order.data(), 4ull * P.num_pencils, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemsetAsync(d_misc, 0, 64, c.stream));
    P.prof = nullptr;
#ifdef SZB_WF_PROFILE
    { unsigned long long *pr = (unsigned long long *)c.get("wf_prof", 64 + 32ull * P.num_pencils + 2048); if (!pr) return SZ_NSCS; CU(cudaMemsetAsync(pr, 0, 64 + 32ull * P.num_pencils + 2048, c.stream)); P.prof = pr; }
#endif
    P.order = d_order; P.bnd = d_bnd; P.epoch = c.wf_epoch; P.ticket = d_misc; P.err = (int *)(d_misc + 4); P.write_all_recon = write_all;
    // ---- pre-pass: neighbour-independent points -> merged input `xin` + one "given" bit per point ----
    const uint32_t n2 = g.ax[2].n, rows = g.ax[0].n * g.ax[1].n, wpr = (n2 + 31) / 32;
    std::vector<uint32_t> ktab(n2);
    for (uint32_t k = 0; k < n2; k++) { uint32_t kk; const uint32_t bk = g.ax[2].block_of(k, kk); ktab[k] = (bk << 8) | kk; }
    uint32_t *d_ktab = (uint32_t *)c.get("wf_ktab", 4ull * n2);
    const uint32_t pitch = (n2 + WfCfg<DIM>::CH - 1) / WfCfg<DIM>::CH * WfCfg<DIM>::CH;      // rows of xin are whole chunks: 16-byte aligned bulk copies
    T *d_xin = (T *)c.get("wf_xin", sizeof(T) * (size_t)rows * pitch + 256);
    uint32_t *d_given = (uint32_t *)c.get("wf_given", 4ull * rows * wpr);
    if (!d_ktab || !d_xin || !d_given) return SZ_NSCS;
    CU(cudaMemcpyAsync(d_ktab, ktab.data(), 4ull * n2, cudaMemcpyHostToDevice, c.stream));
    {
        const int pgrid = (int)std::min<uint32_t>((rows + 7) / 8, (uint32_t)c.sm_count * 16u);
        if (!decomp) {
            if (use_mean) prepass_compress_kernel<T, DIM, true><<<pgrid, 256, 0, c.stream>>>(g, x, indicator, coeff_last, d_ktab, codes, d_xin, pitch, d_given, wpr, eb, recip, mean, cap, cap / 2);
            else prepass_compress_kernel<T, DIM, false><<<pgrid, 256, 0, c.stream>>>(g, x, indicator, coeff_last, d_ktab, codes, d_xin, pitch, d_given, wpr, eb, recip, mean, cap, cap / 2);
        } else {
            if (use_mean) prepass_decompress_kernel<T, DIM, true><<<pgrid, 256, 0, c.stream>>>(g, codes, recon, indicator, coeff_last, d_ktab, d_xin, pitch, d_given, wpr, eb, mean, cap / 2);
            else prepass_decompress_kernel<T, DIM, false><<<pgrid, 256, 0, c.stream>>>(g, codes, recon, indicator, coeff_last, d_ktab, d_xin, pitch, d_given, wpr, eb, mean, cap / 2);
        }
        KCHECK();
    }
    P.given = d_given; P.given_wpr = wpr; P.pitch = pitch;
    CUtensorMap xin_map;
    if (make_xin_map<T, DIM>(&xin_map, d_xin, g, pitch)) return SZ_NSCS;
    const size_t per_warp = wf_smem_per_warp<T>(DIM);
    int warps = sizeof(T) == 4 ? 4 : 2;
    while (warps > 1 && per_warp * warps > c.smem_optin) warps--;
    const size_t smem = per_warp * warps;
    int ctas_per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (c.smem_optin + 1024) / (smem + 1024)));
    uint32_t need_ctas = (P.num_pencils + warps - 1) / warps;
    int grid = (int)std::min<uint32_t>(need_ctas, (uint32_t)(c.sm_count * ctas_per_sm));
    auto run = [&](auto kern) -> int {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, warps * 32, smem, c.stream>>>(P, xin_map, codes, recon, eb, recip); KCHECK();
        return 0;
    };
    int rc;
    if (decomp) rc = run(wavefront_kernel<T, DIM, false, true>);
    else rc = use_mean ? run(wavefront_kernel<T, DIM, true, false>) : run(wavefront_kernel<T, DIM, false, false>);
    if (rc) return rc;
#ifdef SZB_WF_PROFILE
    { unsigned long long h[8]; CU(cudaMemcpyAsync(h, P.prof, 64, cudaMemcpyDeviceToHost, c.stream)); CU(cudaStreamSynchronize(c.stream));
      fprintf(stderr, "[wf-prof] pencils=%u warps/cta=%d grid=%d | per pencil: total=%.0f cyc, neighbour wait + halo fetch=%.0f | steps/pencil=%.0f -> %.0f cyc/step outside the fetch\n",
              P.num_pencils, warps, grid, (double)h[3] / P.num_pencils, (double)h[0] / P.num_pencils,
              (double)h[4] / P.num_pencils, ((double)h[3] - (double)h[0]) / (double)h[4]);
      if (const char *tl = getenv("SZB200_WF_TIMELINE")) {       // per-pencil timeline: ticket time, first-chunk start, end (ns), SM id
          std::vector<unsigned long long> t(4ull * P.num_pencils);
          CU(cudaMemcpy(t.data(), P.prof + 8, 32ull * P.num_pencils, cudaMemcpyDeviceToHost));
          { unsigned long long tr[256]; CU(cudaMemcpy(tr, P.prof + 8 + 4ull * P.num_pencils, 2048, cudaMemcpyDeviceToHost));
            fprintf(stderr, "[wf-trace] chunk: enter(+cycles since previous exit) fetch-cycles spins stage-cycles\n");
            for (int q = 0; q < 40 && tr[4*q+1]; q++) fprintf(stderr, "  c=%d +%lld fetch=%lld spins=%llu stage=%lld\n", q, q ? (long long)(tr[4*q] - tr[4*q-1]) : 0LL, (long long)(tr[4*q+1] - tr[4*q]), tr[4*q+2], (long long)(tr[4*q+3] - tr[4*q+1])); }
          if (FILE *f = fopen(tl, "w")) { fprintf(f, "%u %u\n", P.pencils_a, P.pencils_b); for (uint32_t q = 0; q < P.num_pencils; q++) fprintf(f, "%llu %llu %llu %llu\n", t[4ull*q], t[4ull*q+1], t[4ull*q+2], t[4ull*q+3]); fclose(f); } } }
#endif
    return 0;
}

static int check_wf_err(Ctx &c) {
    unsigned int *d_misc = (unsigned int *)c.get("wf_misc", 64);
    int e = 0;
    CU(cudaMemcpyAsync(&e, d_misc + 4, 4, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (e) { set_err("wavefront kernel: dependency wait timed out (internal error)"); return SZ_NSCS; }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
extern "C" size_t szb200_body_capacity(const szb200_job *job) {
    Geom g = make_geom(job->dim, job->n0, job->n1, job->n2);
    size_t st = job->dtype == 0 ? 4 : 8;
    // mirrors the reference's calloc (ref: sz_float.c:7394) plus slack for the coefficient trees
    return 8 + 4 + st + 12 + 5ull * (16 + 13ull * 2 * 65536) + 1 + st + g.num_blocks * (4 * 4 + 2 + 2 + st) + 8 + g.N * st + g.N * 4 + 4096;
}

// one cluster of two CTAs per tree (the second one only works for trees too large for one SM's shared memory, see huff_build_kernel)
// max_leaves: upper bound of the leaves of any tree of the launch.  Below 32768 every tree fits one SM's shared memory and the grid is
// one plain CTA per tree (a cluster needs two free SMs of one GPC at the same moment: under several worker lanes that wait costs).
static cudaError_t launch_huff_build(cudaStream_t st, const HuffTreeDesc *d_desc, int ntrees, size_t heap_sm, uint32_t heap_cap, uint64_t max_leaves) {
    const unsigned csize = max_leaves >= 32768u ? 2u : 1u;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(csize * (unsigned)ntrees); cfg.blockDim = dim3(HUFF_TB); cfg.dynamicSmemBytes = heap_sm; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    static int form = -1;                                        // SZB200_HEAP=twowarp: the older two-warp heap for every tree (A/B runs)
    if (form < 0) { const char *e = getenv("SZB200_HEAP"); form = (e && strcmp(e, "twowarp") == 0) ? 1 : 0; }
    return cudaLaunchKernelEx(&cfg, huff_build_kernel, d_desc, heap_cap, (uint32_t)form);
}

template <typename T>
static int compress_impl(Ctx &c, const szb200_job &job, const T *d_x, void *dst, int dst_on_device, size_t dst_cap, size_t *body_size, szb200_stats *stats) {
    const Geom g = make_geom(job.dim, job.n0, job.n1, job.n2);
    const int ncoef = g.ncoef;
    const uint32_t nb = g.num_blocks;
    const uint64_t N = g.N;
    if (N >= (1ull << 32)) { set_err("arrays of 2^32 or more elements must be cut into slabs (N=%llu)", (unsigned long long)N); return SZ_DERR; }
    if (job.max_quant_intervals > 65536 || job.quantization_intervals > 65536) { set_err("more than 65536 quantization intervals is not supported"); return SZ_NSCS; }
    if (job.sampleDistance < 2) { set_err("sampleDistance must be >= 2"); return SZ_NSCS; }
    c.launches = 0;
    CU(cudaStreamSynchronize(c.stream2));       // a call that failed half-way may have left work on the side stream
    int evi = 0;
    CU(cudaEventRecord(c.ev[evi++], c.stream));
    float ms_stage[16]; memset(ms_stage, 0, sizeof(ms_stage));
    int stage_of_ev[24];
    auto mark = [&](int st) -> int { stage_of_ev[evi] = st; CU(cudaEventRecord(c.ev[evi++], c.stream)); return 0; };

    const T eb = (T)job.eb;                      // float path: the kernels see the narrowed value (ref: sz_float.c:6527 signature)
    const double ebw = (double)eb;
    const T recip_eb = (T)1 / eb;
    unsigned int intervals;
    int use_mean = 0;
    T mean = (T)0;
    const uint32_t maxRangeRadius = job.max_quant_intervals / 2;

    // ---- precisions (ref: sz_float.c:6639-6645, 2D :5608-5612) and noise (:6697 / :5674) ----
    T prec[4] = {0, 0, 0, 0};
    {
        const T rel = (job.dim == 3) ? (T)0.025 : (T)(0.15 / 3);
        const T t = rel * eb;
        if (job.dim == 3) { prec[0] = t / (T)g.ax[0].late; prec[1] = t / (T)g.ax[1].late; prec[2] = t / (T)g.ax[2].late; prec[3] = t; }
        else { prec[0] = t / (T)g.ax[1].late; prec[1] = t / (T)g.ax[2].late; prec[2] = t; }
    }
    const T noise = (job.dim == 3) ? (T)((double)eb * 1.22) : (T)((double)eb * 0.81);

    // ---- fit + selection, started beside the interval optimiser (which goes to the side stream): it needs nothing of the optimiser but the answer
    //      to "predict blocks from the mean too?", which is almost always no; when it is yes, the stage is run again below ----
    T *d_reg = (T *)c.get("reg", sizeof(T) * 4ull * nb), *d_regc = (T *)c.get("reg_c", sizeof(T) * 4ull * nb), *d_clast = (T *)c.get("coeff_last", sizeof(T) * 4ull * nb),
      *d_cunp = (T *)c.get("coeff_unpred", sizeof(T) * 4ull * nb);
    uint16_t *d_ctype = (uint16_t *)c.get("coeff_type", 2ull * 4 * nb + 64);
    unsigned char *d_ind = (unsigned char *)c.get("indicator", nb + 64);
    uint32_t *d_flags = (uint32_t *)c.get("reg_flags", 4ull * nb + 64), *d_list = (uint32_t *)c.get("reg_list", 4ull * nb + 64);
    unsigned long long *d_rank = (unsigned long long *)c.get("reg_rank", 8ull * nb + 64);
    unsigned long long *d_tot = (unsigned long long *)c.get("totals", 256);
    if (!d_reg || !d_regc || !d_clast || !d_cunp || !d_ctype || !d_ind || !d_flags || !d_list || !d_rank || !d_tot) return SZ_NSCS;
    if (c.need_pinned_words()) return SZ_NSCS;
    auto launch_fit = [&](cudaStream_t st, int scan_slot, int um, T mn) -> int {
        int grid = (nb + 127) / 128;
        if (job.dim == 3) fit_select_kernel<T, 3><<<grid, 128, 0, st>>>(d_x, g, noise, um, mn, d_reg, d_ind);
        else fit_select_kernel<T, 2><<<grid, 128, 0, st>>>(d_x, g, noise, um, mn, d_reg, d_ind);
        KCHECK();
        reg_flag_kernel<<<(nb + 255) / 256, 256, 0, st>>>(d_ind, nb, d_flags); KCHECK();
        if (launch_scan(c, st, scan_slot, d_flags, nb, d_rank, d_tot)) return SZ_NSCS;
        CU(cudaMemcpyAsync(c.pinned_words, d_tot, 8, cudaMemcpyDeviceToHost, st));      // pinned: the call returns at once
        return 0;
    };
    const bool fit_early = job.quantization_intervals == 0;
    if (fit_early) {
        CU(cudaEventRecord(c.ev_fork, c.stream));
        CU(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));        // the optimiser below runs on the side stream
        if (launch_fit(c.stream, 0, 0, (T)0)) return SZ_NSCS;
    }

    // ---- interval optimiser (ref: sz_float.c:6647-6655) ----
    T dense_pos = (T)0;
    if (job.quantization_intervals == 0) {
        const cudaStream_t so = c.stream2;          // the optimiser's small kernels: high-priority side stream, beside the fit on the main one
        std::vector<uint64_t> pos; sample_positions(c, g, pos);
        const uint32_t cnt = (uint32_t)pos.size();
        uint64_t *d_pos = (uint64_t *)c.get("opt_pos", 8ull * cnt + 8); T *d_vals = (T *)c.get("opt_vals", sizeof(T) * (size_t)cnt + 64);
        unsigned char *ob = (unsigned char *)c.get("opt_misc", 4ull * 32768 + 4ull * 8192 + 256);
        if (!d_pos || !d_vals || !ob) return SZ_NSCS;
        uint32_t *d_int = (uint32_t *)ob, *d_freq = d_int + 32768; OptAccum *d_acc = (OptAccum *)(d_freq + 8192);
        OptOut *d_out = (OptOut *)((unsigned char *)d_acc + 64); T *d_mean = (T *)((unsigned char *)d_acc + 128);
        void *pin = c.pin(8ull * cnt + 64); if (!pin) return SZ_NSCS;
        memcpy(pin, pos.data(), 8ull * cnt);
        CU(cudaMemcpyAsync(d_pos, pin, 8ull * cnt, cudaMemcpyHostToDevice, so));
        CU(cudaMemsetAsync(ob, 0, 4ull * 32768 + 4ull * 8192 + 256, so));
        {
            const size_t sm = sizeof(T) * (size_t)cnt + 16;
            const int in_smem = sm <= c.smem_optin - 1024 ? 1 : 0;
            if (in_smem) CU(cudaFuncSetAttribute(sample_mean_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
            sample_mean_kernel<T><<<1, 1024, in_smem ? sm : 0, so>>>(d_x, d_pos, cnt, d_vals, d_mean, in_smem); KCHECK();
        }
        const uint64_t rows = opt_row_limit(g, (uint32_t)job.sampleDistance);
        if (rows > 0) {
            int grid = (int)std::min<uint64_t>((rows + 255) / 256, (uint64_t)c.sm_count * 8);
            if (job.dim == 3) opt_sample_kernel<T, 3><<<grid, 256, 0, so>>>(d_x, g, (uint32_t)job.sampleDistance, rows, d_mean, ebw, maxRangeRadius, d_int, d_freq, d_acc);
            else opt_sample_kernel<T, 2><<<grid, 256, 0, so>>>(d_x, g, (uint32_t)job.sampleDistance, rows, d_mean, ebw, maxRangeRadius, d_int, d_freq, d_acc);
            KCHECK();
        }
        opt_finish_kernel<T><<<1, 1024, 0, so>>>(d_int, d_freq, d_acc, d_mean, maxRangeRadius, job.predThreshold, d_out); KCHECK();
        OptOut ho;
        CU(cudaMemcpyAsync(&ho, d_out, sizeof(ho), cudaMemcpyDeviceToHost, so));
        CU(cudaStreamSynchronize(so));
        // scalar tail of optimize_intervals_* (ref: sz_float.c:6486-6520)
        unsigned int acc = 2 * (ho.first_index + 1);
        unsigned int p2 = round_up_pow2(acc); if (p2 < 32) p2 = 32;
        intervals = p2;
        const T smean = (T)ho.mean;
        const T max_freq = (T)((double)ho.hits * 1.0 / (double)ho.samples);
        dense_pos = (T)((double)smean + ebw * (double)(long long)((long long)ho.max_index + 1 - 4096));
        const T mean_freq = (T)((double)ho.max_sum * 1.0 / (double)ho.samples);
        if ((double)mean_freq > 0.5 || mean_freq > max_freq) use_mean = 1;
        if (job.dim == 2) use_mean = 0;             // ref: sz_float.c:5615
    } else intervals = job.quantization_intervals;
    if (mark(ST_OPT)) return SZ_NSCS;

    if (use_mean) {                                 // ref: sz_float.c:6657-6669
        T *d_m = (T *)c.get("mean_out", 64); if (!d_m) return SZ_NSCS;
        mean_kernel<T><<<1, 1024, 0, c.stream>>>(d_x, N, dense_pos, eb, d_m); KCHECK();
        CU(cudaMemcpyAsync(&mean, d_m, sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
    }
    if (mark(ST_MEAN)) return SZ_NSCS;

    // ---- fit + selection: the early run stands unless blocks are predicted from the mean as well ----
    if (fit_early) CU(cudaStreamSynchronize(c.stream));          // (the optimiser's stream was synchronised for its result)
    if (!fit_early || use_mean) {
        if (launch_fit(c.stream, 0, use_mean, mean)) return SZ_NSCS;
        CU(cudaStreamSynchronize(c.stream));
    }
    const uint32_t reg_count = (uint32_t)c.pinned_words[0];
    if (mark(ST_FIT)) return SZ_NSCS;

    // ---- coefficient chain (its proof stage also counts the codes: the histograms of the four coefficient trees) ----
    TreeWs tw[5];
    const int ntrees = 1 + (reg_count > 0 ? ncoef : 0);
    if (tree_ws(c, 0, intervals, tw[0])) return SZ_NSCS;
    for (int e = 0; e < ntrees - 1; e++) if (tree_ws(c, 1 + e, 65536, tw[1 + e])) return SZ_NSCS;
    // everything the side branch needs is allocated, cleared and uploaded here, before the chain is launched: the host then queues the tree
    // kernel while the chain still runs and the GPU starts it the moment the chain ends
    HuffTreeDesc *d_desc = (HuffTreeDesc *)c.get("tree_desc", sizeof(HuffTreeDesc) * 8); if (!d_desc) return SZ_NSCS;
    const uint32_t heap_cap = (uint32_t)((c.smem_optin - 8192) / 8);
    const size_t heap_sm = 8ull * heap_cap;
    CU(cudaFuncSetAttribute(huff_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)heap_sm));
    const uint64_t nchunk_main = (N + ENC_CHUNK - 1) / ENC_CHUNK, nchunk_c = reg_count ? ((uint64_t)reg_count + ENC_CHUNK - 1) / ENC_CHUNK : 0;
    const uint64_t nch_all = nchunk_main + 4 * nchunk_c + 8;
    uint32_t *d_cb = (uint32_t *)c.get("enc_cb", 4 * nch_all), *d_cz = (uint32_t *)c.get("enc_cz", 4 * nch_all);
    unsigned long long *d_cbo = (unsigned long long *)c.get("enc_cbo", 8 * nch_all), *d_czo = (unsigned long long *)c.get("enc_czo", 8 * nch_all);
    uint16_t *d_codes = (uint16_t *)c.get("codes", 2ull * N + 64);
    const int write_all = getenv("SZB200_DEBUG_RECON") ? 1 : 0;      // debug: keep the full reconstructed field (szb200_debug_fetch(4))
    T *d_recon = write_all ? (T *)c.get("recon", sizeof(T) * N + 64) : nullptr;
    if (!d_cb || !d_cz || !d_cbo || !d_czo || !d_codes || (write_all && !d_recon)) return SZ_NSCS;
    // The coefficient streams: ONE launch behind the tree kernel on the side stream (coef_encode_kernel: lengths, look-back prefix,
    // bit scatter).  A Huffman code averages less than entropy + 1 <= 17 bits per symbol: that bounds the buffers, zeroed here.
    uint32_t *d_cbits[4] = {nullptr, nullptr, nullptr, nullptr};
    unsigned long long *d_cstate = nullptr;
    if (ntrees > 1) {
        const size_t bound = (17ull * reg_count + 7) / 8 + 64;
        for (int e = 0; e < ntrees - 1; e++) {
            char name[32]; snprintf(name, sizeof(name), "enc_cbits%d", e);
            d_cbits[e] = (uint32_t *)c.get(name, bound); if (!d_cbits[e]) return SZ_NSCS;
            CU(cudaMemsetAsync(d_cbits[e], 0, bound, c.stream));
        }
        d_cstate = (unsigned long long *)c.get("enc_cstate", 8 * (4 * nchunk_c + 8)); if (!d_cstate) return SZ_NSCS;
        CU(cudaMemsetAsync(d_cstate, 0, 8 * (4 * nchunk_c + 8), c.stream));
    }
    unsigned int *d_done = (unsigned int *)c.get("chain_done", 64); if (!d_done) return SZ_NSCS;
    CU(cudaMemsetAsync(d_done, 0, 64, c.stream));
    ChainOut *d_chain = (ChainOut *)c.get("chain_out", 128); if (!d_chain) return SZ_NSCS;
    CU(cudaMemsetAsync(d_chain, 0, 128, c.stream));
    for (int e = 0; e < ntrees - 1; e++) CU(cudaMemsetAsync(tw[1 + e].hist, 0, 4ull * 65536, c.stream));      // counted by the chain's proof stage
    for (int e = 0; e < ntrees - 1; e++) tw[1 + e].d.wait_flag = d_done + e;      // tree e starts when chain e has set its flag
    {
        HuffTreeDesc hd[8];
        for (int i = 0; i < ntrees; i++) hd[i] = tw[i].d;
        CU(cudaMemcpyAsync(d_desc, hd, sizeof(HuffTreeDesc) * ntrees, cudaMemcpyHostToDevice, c.stream));   // pageable: staged before return
    }
    if (ntrees > 1) {                                  // the side stream starts from here: everything above is ready, the chain is not awaited
        CU(cudaEventRecord(c.ev_fork, c.stream));
        CU(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));
    }
    if (reg_count > 0) {
        LatHist hists; memset(&hists, 0, sizeof(hists));
        for (int e = 0; e < ncoef; e++) hists.h[e] = tw[1 + e].hist;
        reg_gather_kernel<T><<<(nb + 255) / 256, 256, 0, c.stream>>>(d_ind, d_rank, nb, ncoef, d_reg, d_regc, d_list); KCHECK();
        const int use_div = (sizeof(T) == 4 && !use_mean) ? 1 : 0;     // ref: sz_float.c:7133 vs :6795; double always multiplies
        // one CTA per coefficient: lattice guess, add chain, exact proof (szb_chain.cuh); SZB200_CHAIN=exact runs every step with the
        // reference's complete formula (the form a failed window falls back to), for tests and timing
        static const int chain_mode = [] { const char *e = getenv("SZB200_CHAIN"); return e && !strcmp(e, "exact") ? 2 : 0; }();
        static const float chain_band = [] { const char *e = getenv("SZB200_LAT_BAND"); return e ? (float)atof(e) : LAT_BAND; }();
        coeff_lattice_kernel<T><<<ncoef, LAT_THREADS, 0, c.stream>>>(d_regc, d_list, nb, reg_count, prec[0], prec[1], prec[2], prec[3], use_div, d_ctype, d_clast, d_cunp, hists, d_done, d_chain, chain_mode, chain_band); KCHECK();
    }
    // ---- side branch on stream2: the coefficient Huffman trees (exact heap emulation, the longest serial stage) and the coefficient bit
    //      streams, concurrent with the wavefront / main-tree / main-encode work.  The tree kernel does not wait for the chain's end in stream order (it was
    //      forked before the chain): its CTAs sit on their SMs and wait for the chain's per-coefficient flags ----
    if (ntrees > 1) {
        CU(cudaEventRecord(c.ev_side0, c.stream2));
        CU(cudaEventRecord(c.ev_tree0, c.stream2));
        CU(launch_huff_build(c.stream2, d_desc + 1, ntrees - 1, heap_sm, heap_cap, std::min<uint64_t>(65536u, reg_count))); KCHECK();
        CU(cudaEventRecord(c.ev_tree1, c.stream2));
        {
            CoefEncArgs A; memset(&A, 0, sizeof(A));
            for (int e = 0; e < ntrees - 1; e++) { A.c[e].syms = d_ctype + (size_t)e * nb; A.c[e].code = tw[1 + e].d.code; A.c[e].clen = tw[1 + e].d.clen; A.c[e].out_words = d_cbits[e]; A.c[e].total_bits = d_tot + 3 + e; }
            A.n = reg_count; A.nchunks = (uint32_t)nchunk_c; A.state = d_cstate; A.ticket = (unsigned int *)(d_cstate + 4 * nchunk_c);
            coef_encode_kernel<<<dim3((unsigned)nchunk_c, ntrees - 1), ENC_THREADS, 0, c.stream2>>>(A); KCHECK();
        }
        CU(cudaEventRecord(c.ev_side1, c.stream2));
        CU(cudaEventRecord(c.ev_join, c.stream2));
    }

    if (mark(ST_CHAIN)) return SZ_NSCS;

    // ---- wavefront predict + quantise ----
    {
        int rc = (job.dim == 3) ? launch_wavefront<T, 3>(c, g, false, use_mean, (int)intervals, d_ind, d_x, d_codes, d_recon, d_clast, eb, recip_eb, mean, write_all)
                                : launch_wavefront<T, 2>(c, g, false, use_mean, (int)intervals, d_ind, d_x, d_codes, d_recon, d_clast, eb, recip_eb, mean, write_all);
        if (rc) return rc;
    }
    if (mark(ST_WAVE)) return SZ_NSCS;

    // ---- main histogram, main tree, main bit lengths ----
    if (launch_hist(c, c.stream, d_codes, N, intervals, tw[0].hist)) return SZ_NSCS;
    if (mark(ST_HIST)) return SZ_NSCS;
    CU(launch_huff_build(c.stream, d_desc, 1, heap_sm, heap_cap, std::min<uint64_t>(intervals, N))); KCHECK();
    if (mark(ST_TREE)) return SZ_NSCS;
    enc_len_kernel<true><<<(unsigned)nchunk_main, ENC_THREADS, 0, c.stream>>>(d_codes, N, g, tw[0].d.clen, d_cb, d_cz); KCHECK();
    if (launch_scan(c, c.stream, 0, d_cb, nchunk_main, d_cbo, d_tot + 1)) return SZ_NSCS;
    if (launch_scan(c, c.stream, 0, d_cz, nchunk_main, d_czo, d_tot + 2)) return SZ_NSCS;
    // ---- sizes of the main stream (the side branch with the coefficient trees is still running; nothing here waits for it) ----
    unsigned long long h_tot[8]; uint32_t h_tree[5][16]; ChainOut h_chain;
    CU(cudaMemcpyAsync(h_tot, d_tot, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaMemcpyAsync(h_tree[0], tw[0].d.out, 64, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (check_wf_err(c)) return SZ_NSCS;
    const uint64_t main_bits = h_tot[1], total_unpred = h_tot[2];
    const uint64_t main_bytes = (main_bits + 7) / 8;

    // ---- encode pass 2, main codes + ordered unpredictable values + mode bits ----
    uint32_t *d_bits = (uint32_t *)c.get("enc_bits", main_bytes + 64);
    T *d_unp = (T *)c.get("unpred", sizeof(T) * (total_unpred + 8));
    if (!d_bits || !d_unp) return SZ_NSCS;
    CU(cudaMemsetAsync(d_bits, 0, main_bytes + 64, c.stream));
    enc_write_kernel<T, true><<<(unsigned)nchunk_main, ENC_THREADS, 0, c.stream>>>(d_codes, N, g, tw[0].d.code, tw[0].d.clen, d_cbo, d_czo, d_bits, d_x, d_unp); KCHECK();
    unsigned char *d_indbits = (unsigned char *)c.get("ind_bits", (nb + 7) / 8 + 64); if (!d_indbits) return SZ_NSCS;
    indicator_pack_kernel<<<((nb + 7) / 8 + 255) / 256, 256, 0, c.stream>>>(d_ind, nb, d_indbits); KCHECK();
    if (mark(ST_ENC)) return SZ_NSCS;

    // ---- join the side branch (coefficient trees), then the coefficient bit streams ----
    if (ntrees > 1) { CU(cudaStreamWaitEvent(c.stream, c.ev_join, 0)); if (mark(ST_JOIN)) return SZ_NSCS; }
    CU(cudaMemcpyAsync(h_tot + 3, d_tot + 3, 5 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    for (int i = 1; i < ntrees; i++) CU(cudaMemcpyAsync(h_tree[i], tw[i].d.out, 64, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaMemcpyAsync(&h_chain, d_chain, sizeof(h_chain), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (getenv("SZB200_TREE_PROFILE")) fprintf(stderr, "[chain-prof] steps=%u passes=%u (reruns: %u) kcycles: serial thread busy=%u, helper busy=%u, total=%u; last coefficient: pair-form steps=%u, full steps=%u\n", reg_count, h_chain.passes, h_chain.slow_chunks, h_chain.kcyc_serial, h_chain.kcyc_helper, h_chain.kcyc_total, h_chain.n_pair, h_chain.n_full);
    if (getenv("SZB200_TREE_PROFILE") && h_chain.prof[0] + h_chain.prof[2]) fprintf(stderr, "[chain-prof] serial thread {count, cycles}: clean words %u %u, clean groups %u %u, pair groups %u %u, step-by-step groups %u %u\n", h_chain.prof[0], h_chain.prof[1], h_chain.prof[2], h_chain.prof[3], h_chain.prof[4], h_chain.prof[5], h_chain.prof[6], h_chain.prof[7]);
    if (getenv("SZB200_TREE_PROFILE")) for (int i = 0; i < ntrees; i++)
        fprintf(stderr, "[tree-prof] tree %d: leaves=%u kcycles: compaction=%u leaf-inserts=%u inserts+merges=%u level-passes=%u; whole CTA %u kcycles in %u ns, after waiting %u us for its input (the chain ended %d us before that wait did)\n", i, h_tree[i][2], h_tree[i][4], h_tree[i][5], h_tree[i][6], h_tree[i][7], h_tree[i][13], h_tree[i][12], h_tree[i][15] - h_tree[i][14], i ? (int)(h_tree[i][15] - h_chain.prof[7]) : 0);
    if (h_chain.err) { set_err("coefficient chain: a step left its tabulated interval (internal error)"); return SZ_NSCS; }
    for (int i = 0; i < ntrees; i++) if (h_tree[i][3]) { set_err(h_tree[i][3] == 2 ? "Huffman tree build gave up (tree %d)" : "Huffman code longer than 64 bits (tree %d)", i); return SZ_NSCS; }
    uint64_t cbytes[4] = {0, 0, 0, 0};
    for (int e = 0; e < ntrees - 1; e++) cbytes[e] = (h_tot[3 + e] + 7) / 8;
    if (mark(ST_ENC)) return SZ_NSCS;

    // ---- stream body (SURVEY.md Appendix A from offset 4+MD; ref: sz_float.c:7392-7489) ----
    Emitter E;
    { unsigned char *p = E.h(8 + 4 + sizeof(T) + 12);
      put_be64(p, N); p += 8; put_be32(p, job.dim == 3 ? 6 : 16); p += 4; put_beT<T>(p, eb); p += sizeof(T);
      put_be32(p, intervals); p += 4; put_be32(p, h_tree[0][0]); p += 4; put_be32(p, h_tree[0][1]); }
    E.d(tw[0].d.tree_bytes, h_tree[0][0]);
    { unsigned char *p = E.h(1 + sizeof(T)); p[0] = (unsigned char)use_mean; memcpy(p + 1, &mean, sizeof(T)); }
    E.d(d_indbits, (nb + 7) / 8);
    for (int e = 0; e < ntrees - 1; e++) {
        { unsigned char *p = E.h(sizeof(T) + 12); put_beT<T>(p, prec[e]); p += sizeof(T); put_be32(p, 32768); p += 4; put_be32(p, h_tree[1 + e][0]); p += 4; put_be32(p, h_tree[1 + e][1]); }
        E.d(tw[1 + e].d.tree_bytes, h_tree[1 + e][0]);
        { unsigned char *p = E.h(8); put_be64(p, cbytes[e]); }
        E.d(d_cbits[e], cbytes[e]);
        { unsigned char *p = E.h(4); put_be32(p, h_chain.unpred_count[e]); }
        E.d(d_cunp + (size_t)e * nb, sizeof(T) * (size_t)h_chain.unpred_count[e]);
    }
    { unsigned char *p = E.h(8); uint64_t tu = total_unpred; memcpy(p, &tu, 8); }
    E.d(d_unp, sizeof(T) * total_unpred);
    E.d(d_bits, main_bytes);
    if (E.total > dst_cap) { set_err("stream body (%zu bytes) exceeds the destination capacity (%zu)", E.total, dst_cap); *body_size = E.total; return SZB200_BODY_TOO_LARGE; }
    if (emit_all(c, E, dst, dst_on_device)) return SZ_NSCS;
    *body_size = E.total;
    if (mark(ST_EMIT)) return SZ_NSCS;
    CU(cudaStreamSynchronize(c.stream));

    szb200_stats st; memset(&st, 0, sizeof(st));
    st.intervals = intervals; st.use_mean = use_mean; st.mean = (double)mean; st.num_blocks = nb; st.reg_count = reg_count; st.total_unpred = total_unpred;
    st.tree_bytes = h_tree[0][0]; st.node_count = h_tree[0][1]; st.type_bits_bytes = main_bytes; st.body_bytes = E.total; st.launches = c.launches;
    for (int i = 1; i < evi; i++) { float ms = 0; cudaEventElapsedTime(&ms, c.ev[i - 1], c.ev[i]); st.ms_stage[stage_of_ev[i]] += ms; }
    cudaEventElapsedTime(&st.ms_total, c.ev[0], c.ev[evi - 1]);
    if (ntrees > 1) {
        for (int i = 1; i < ntrees; i++) st.ms_stage[14] = std::max(st.ms_stage[14], (float)h_tree[i][12] * 1e-6f);      // measured inside the kernel, behind its wait for the chain
        float b = 0; cudaEventElapsedTime(&b, c.ev_tree1, c.ev_side1); st.ms_stage[15] = b;
    }
    c.last = st; c.have_last = true;
    c.dbg_g = g; c.dbg_dtype = job.dtype; c.dbg_intervals = intervals; c.dbg_reg_count = reg_count;
    if (stats) *stats = st;
    return 0;
}

extern "C" int szb200_compress_blocked(const szb200_job *job, const void *d_data, void *dst, int dst_on_device, size_t dst_capacity, size_t *body_size, szb200_stats *stats) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (job->dim != 2 && job->dim != 3) { set_err("dim must be 2 or 3"); return SZ_DERR; }
    if (job->dtype == 0) return compress_impl<float>(*c, *job, (const float *)d_data, dst, dst_on_device, dst_capacity, body_size, stats);
    if (job->dtype == 1) return compress_impl<double>(*c, *job, (const double *)d_data, dst, dst_on_device, dst_capacity, body_size, stats);
    set_err("dtype must be SZ_FLOAT or SZ_DOUBLE"); return SZ_NSCS;
}

// ---------------------------------------------------------------------------------------------------------------------
// Huffman decode of several symbol arrays driven together (main type array + the coefficient arrays): the Jacobi
// synchronisation rounds of all of them are launched back to back and their "changed" flags are read with ONE host sync per
// pair of rounds.  tree bytes + payloads are device pointers into the (aligned copy of the) stream.
struct DecJob {
    int slot; const unsigned char *d_tree; uint32_t nodeCount; const unsigned char *d_payload; uint64_t payload_bytes, nsym;
    bool blockmajor; uint16_t *d_out;
    // internal
    uint32_t *lut, *L, *R, *C; uint8_t *leaf; uint32_t *words; uint64_t total_bits, nsub;
    unsigned long long *ep, *ec, *start_used, *sym_off, *zero_off; uint32_t *count, *zcount;
    bool single_leaf, converged; uint32_t root_sym;
};

template <typename T>
static int huff_decode_all(Ctx &c, DecJob *jobs, int njobs, const Geom &g, const T *d_unpred, uint64_t unpred_count, T *d_field,
                           const std::function<int(cudaStream_t)> &side_tail, unsigned long long **d_totals_out, uint64_t *expect /*[5], 0 = not checked*/) {
    char nm[48];
    struct RootInfo { uint8_t leaf; uint32_t sym; } roots[8];
    const bool prof = getenv("SZB200_DEC_PROFILE") != nullptr;
    cudaEvent_t pe[6]; int npe = 0;
    auto pmark = [&]() { if (prof && npe < 6) { cudaEventCreate(&pe[npe]); cudaEventRecord(pe[npe], c.stream); npe++; } };
    pmark();
    for (int j = 0; j < njobs; j++) {
        DecJob &J = jobs[j];
        snprintf(nm, sizeof(nm), "dec_tree%d", J.slot);
        unsigned char *tb = (unsigned char *)c.get(nm, 16ull * J.nodeCount + 1024 + 4ull * DEC_LUT_WORDS); if (!tb) return SZ_NSCS;
        // the three look-up tables (dec_lut_kernel), then the unpacked tree
        J.lut = (uint32_t *)tb; J.L = J.lut + DEC_LUT_WORDS; J.R = J.L + J.nodeCount; J.C = J.R + J.nodeCount; J.leaf = (uint8_t *)(J.C + J.nodeCount);
        dec_tree_unpack_kernel<<<(J.nodeCount + 255) / 256, 256, 0, c.stream>>>(J.d_tree, J.nodeCount, J.L, J.R, J.C, J.leaf); KCHECK();
        // a tree whose root is a leaf has exactly one node (ref: Huffman.c:316-321); only then is the root's symbol needed on the host
        roots[j].leaf = J.nodeCount == 1 ? 1 : 0; roots[j].sym = 0;
        if (roots[j].leaf) { CU(cudaMemcpyAsync(&roots[j].sym, J.C, 4, cudaMemcpyDeviceToHost, c.stream)); CU(cudaStreamSynchronize(c.stream)); }
    }
    pmark();
    unsigned int *d_changed = (unsigned int *)c.get("dec_changed", 256); if (!d_changed) return SZ_NSCS;
    for (int j = 0; j < njobs; j++) {
        DecJob &J = jobs[j];
        J.single_leaf = roots[j].leaf != 0; J.root_sym = roots[j].sym; J.converged = false;
        if (J.single_leaf) {            // constant symbol array, empty payload (ref: Huffman.c:316-321)
            fill_u16_kernel<<<c.sm_count * 4, 256, 0, c.stream>>>(J.d_out, J.nsym, (uint16_t)J.root_sym); KCHECK();
            if (J.blockmajor && J.root_sym == 0 && d_field) {       // every point unpredictable: the raw values are the field
                unpred_all_kernel<T><<<(unsigned)((J.nsym + 4095) / 4096), 256, 0, c.stream>>>(J.nsym, g, d_unpred, unpred_count, d_field); KCHECK();
            }
            J.converged = true;
            continue;
        }
        dec_lut_kernel<<<((1u << DEC_K) + 255) / 256, 256, 0, c.stream>>>(J.L, J.R, J.C, J.leaf, J.lut); KCHECK();
        snprintf(nm, sizeof(nm), "dec_words%d", J.slot);
        J.total_bits = J.payload_bytes * 8;
        const size_t wbytes = ((J.payload_bytes + 3) & ~(size_t)3) + 256;
        J.words = (uint32_t *)c.get(nm, wbytes); if (!J.words) return SZ_NSCS;
        CU(cudaMemsetAsync((unsigned char *)J.words + (J.payload_bytes & ~(size_t)3), 0, wbytes - (J.payload_bytes & ~(size_t)3), c.stream));
        CU(cudaMemcpyAsync(J.words, J.d_payload, J.payload_bytes, cudaMemcpyDeviceToDevice, c.stream));
        J.nsub = (J.total_bits + DEC_SUB_BITS - 1) / DEC_SUB_BITS;
        if (J.nsub == 0) { set_err("empty Huffman payload for %llu symbols", (unsigned long long)J.nsym); return SZ_NSCS; }
        snprintf(nm, sizeof(nm), "dec_sub%d", J.slot);
        unsigned char *sb = (unsigned char *)c.get(nm, (8 * 5 + 4 * 2) * (J.nsub + 8) + 256); if (!sb) return SZ_NSCS;
        J.ep = (unsigned long long *)sb; J.ec = J.ep + J.nsub + 1; J.start_used = J.ec + J.nsub + 1; J.sym_off = J.start_used + J.nsub + 1; J.zero_off = J.sym_off + J.nsub + 1;
        J.count = (uint32_t *)(J.zero_off + J.nsub + 1); J.zcount = J.count + J.nsub + 1;
    }
    pmark();
    // the prefix sums also give the number of symbols every stream really holds; the caller compares them with the counts the
    // header promises once the job is done (a truncated stream otherwise decodes "successfully" into too few symbols)
    unsigned long long *d_dtot = (unsigned long long *)c.get("dec_totals", 64); if (!d_dtot) return SZ_NSCS;
    CU(cudaMemsetAsync(d_dtot, 0, 64, c.stream));
    *d_totals_out = d_dtot;
    for (int j = 0; j < 5; j++) expect[j] = (j < njobs && !jobs[j].single_leaf) ? jobs[j].nsym : 0;
    // The coefficient streams' prefix sums and write passes (one launch each for all of them) and the work that only needs the
    // coefficient codes (the coefficient chain) go to the second stream, forked from the main one where this is called.
    bool side_used = false;
    auto launch_side = [&](bool fork) -> int {
        if (fork) { CU(cudaEventRecord(c.ev_fork, c.stream)); CU(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0)); }
        DecJobList L; memset(&L, 0, sizeof(L));
        uint32_t ctas = 0;
        for (int j = 0; j < njobs; j++) {
            DecJob &J = jobs[j];
            if (J.single_leaf || J.blockmajor) continue;
            DecJobDev &D = L.j[L.n++];
            D.words = J.words; D.total_bits = J.total_bits; D.nsub = J.nsub; D.tr = DecTree{J.L, J.R, J.C, J.leaf, J.nodeCount, J.lut};
            D.start_used = J.start_used; D.count = J.count; D.sym_off = J.sym_off; D.nsym = J.nsym; D.out = J.d_out; D.total = d_dtot + j;
            D.cta_begin = ctas; D.active = 1;
            ctas += (uint32_t)((J.nsub + 255) / 256);
        }
        if (L.n > 0) {
            DecJobList *d_clist = (DecJobList *)c.get("dec_joblist_coef", sizeof(DecJobList) + 64); if (!d_clist) return SZ_NSCS;
            CU(cudaMemcpyAsync(d_clist, &L, sizeof(L), cudaMemcpyHostToDevice, c.stream2));
            dec_coef_scan_kernel<<<L.n, 1024, 0, c.stream2>>>(d_clist); KCHECK();
            dec_coef_write_kernel<<<ctas, 256, 0, c.stream2>>>(d_clist); KCHECK();
            side_used = true;
        }
        if (side_tail) { if (side_tail(c.stream2)) return SZ_NSCS; side_used = true; }
        return 0;
    };
    // Jacobi rounds: round 0 is the speculative decode; a job has converged when a whole round moved no start position.
    // One launch per round covers every job that has not converged yet.  Round 0 starts every CTA's first sub-sequence at a guess, round 1
    // hands the true starts across the CTA boundaries and is nearly always the last one: the coefficient streams' side branch is queued
    // right behind it, before its verdict is read back, and queued again only if the verdict asks for another round on one of them.
    if (c.need_pinned_words()) return SZ_NSCS;
    DecJobList *d_list = (DecJobList *)c.get("dec_joblist", sizeof(DecJobList) + 64), *d_list2 = (DecJobList *)c.get("dec_joblist2", sizeof(DecJobList) + 64);
    if (!d_list || !d_list2) return SZ_NSCS;
    unsigned int *d_changed2 = d_changed + 16;
    // one round for the jobs `take` selects (not yet converged ones), on stream `st`
    auto launch_round = [&](cudaStream_t st, const std::function<bool(const DecJob &)> &take, bool first, DecJobList *d_l, unsigned int *d_ch, bool &any) -> int {
        DecJobList L; memset(&L, 0, sizeof(L));
        uint32_t ctas = 0; any = false;
        for (int j = 0; j < njobs; j++) {
            DecJob &J = jobs[j];
            DecJobDev &D = L.j[j];
            D.cta_begin = ctas; D.active = 0;
            if (J.single_leaf || !take(J)) continue;
            D.words = J.words; D.total_bits = J.total_bits; D.nsub = J.nsub; D.tr = DecTree{J.L, J.R, J.C, J.leaf, J.nodeCount, J.lut};
            D.end_prev = J.ep; D.end_cur = J.ec; D.start_used = J.start_used; D.count = J.count; D.zcount = J.zcount; D.changed = d_ch + j;
            ctas += (uint32_t)((J.nsub + 255) / 256);
            if (J.converged) continue;
            D.active = 1; any = true;
            std::swap(J.ep, J.ec);
        }
        L.n = njobs;
        if (!any) return 0;
        CU(cudaMemsetAsync(d_ch, 0, 4 * 8, st));
        CU(cudaMemcpyAsync(d_l, &L, sizeof(L), cudaMemcpyHostToDevice, st));       // pageable source: staged before the call returns
        dec_sync_kernel<<<ctas, 256, 0, st>>>(d_l, first ? 1 : 0); KCHECK();
        return 0;
    };
    auto is_coef = [](const DecJob &J) { return !J.blockmajor; };
    auto is_main = [](const DecJob &J) { return J.blockmajor; };
    auto is_any = [](const DecJob &) { return true; };
    // The coefficient streams (a few CTAs each) go through their two rounds, their prefix sums, write passes and the coefficient chain on
    // the second stream, from the start, beside the main stream's rounds.  Round 1's verdict on them is read back when the main
    // stream is done; should it ask for another round (not seen), the rounds go on below and the side branch is queued once more.
    bool side_early = false, coef_moved = false;
    {
        bool have_coef = false;
        for (int j = 0; j < njobs; j++) if (!jobs[j].single_leaf && !jobs[j].blockmajor) have_coef = true;
        if (have_coef) {
            bool any = false;
            CU(cudaEventRecord(c.ev_fork, c.stream));
            CU(cudaStreamWaitEvent(c.stream2, c.ev_fork, 0));
            if (launch_round(c.stream2, is_coef, true, d_list2, d_changed2, any)) return SZ_NSCS;
            if (launch_round(c.stream2, is_coef, false, d_list2, d_changed2, any)) return SZ_NSCS;
            CU(cudaMemcpyAsync(c.pinned_words + 8, d_changed2, 4 * 8, cudaMemcpyDeviceToHost, c.stream2));
            CU(cudaEventRecord(c.ev_side0, c.stream2));
            if (launch_side(false)) return SZ_NSCS;
            side_early = true;
        }
    }
    for (int round = 0;; round++) {
        bool any = false;
        if (launch_round(c.stream, side_early ? std::function<bool(const DecJob &)>(is_main) : std::function<bool(const DecJob &)>(is_any), round == 0, d_list, d_changed, any)) return SZ_NSCS;
        if (!any) break;
        if (round == 0) continue;                                // the first round always changes something
        unsigned int ch[8];
        CU(cudaMemcpyAsync(ch, d_changed, 4 * njobs, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        for (int j = 0; j < njobs; j++) if ((!side_early || jobs[j].blockmajor) && !jobs[j].converged && !ch[j]) jobs[j].converged = true;
        if (round > 100000) { set_err("Huffman decode failed to synchronise"); return SZ_NSCS; }
    }
    if (side_early) {
        CU(cudaEventSynchronize(c.ev_side0));
        const unsigned int *ch2 = reinterpret_cast<const unsigned int *>(c.pinned_words + 8);
        for (int j = 0; j < njobs; j++) if (!jobs[j].single_leaf && !jobs[j].blockmajor) { if (ch2[j]) coef_moved = true; else jobs[j].converged = true; }
        for (int round = 2; coef_moved; round++) {               // more rounds for the coefficient streams (on the main stream, in order behind the side branch)
            bool any = false;
            CU(cudaEventRecord(c.ev_join, c.stream2)); CU(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
            if (launch_round(c.stream, is_coef, false, d_list, d_changed, any)) return SZ_NSCS;
            if (!any) break;
            unsigned int ch[8];
            CU(cudaMemcpyAsync(ch, d_changed, 4 * njobs, cudaMemcpyDeviceToHost, c.stream));
            CU(cudaStreamSynchronize(c.stream));
            for (int j = 0; j < njobs; j++) if (!jobs[j].blockmajor && !jobs[j].converged && !ch[j]) jobs[j].converged = true;
            if (round > 100000) { set_err("Huffman decode failed to synchronise"); return SZ_NSCS; }
        }
    }
    pmark();
    if (!side_early || coef_moved) { if (launch_side(true)) return SZ_NSCS; }     // (again: the coefficient streams needed more rounds)
    for (int j = 0; j < njobs; j++) {
        DecJob &J = jobs[j];
        if (J.single_leaf || !J.blockmajor) continue;
        DecTree tr{J.L, J.R, J.C, J.leaf, J.nodeCount, J.lut};
        if (launch_scan(c, c.stream, 0, J.count, J.nsub, J.sym_off, d_dtot + j)) return SZ_NSCS;
        if (launch_scan(c, c.stream, 0, J.zcount, J.nsub, J.zero_off, nullptr)) return SZ_NSCS;
        const unsigned grid = (unsigned)((J.nsub + DECW_TB - 1) / DECW_TB);
        CU(cudaFuncSetAttribute(dec_write_bm_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DECW_SMEM));
        dec_write_bm_kernel<T><<<grid, DECW_TB, DECW_SMEM, c.stream>>>(J.words, J.total_bits, J.nsub, tr, J.start_used, J.sym_off, J.zero_off, J.nsym, g, J.d_out, d_unpred, unpred_count, d_field);
        KCHECK();
    }
    pmark();
    if (side_used) { CU(cudaEventRecord(c.ev_join, c.stream2)); CU(cudaStreamWaitEvent(c.stream, c.ev_join, 0)); }
    pmark();
    if (prof) {
        cudaStreamSynchronize(c.stream);
        float ms[6] = {0, 0, 0, 0, 0, 0};
        for (int q = 1; q < npe; q++) cudaEventElapsedTime(&ms[q], pe[q - 1], pe[q]);
        fprintf(stderr, "[dec-prof] tree unpack + root read-back %.3f | luts, payload copies %.3f | sync rounds %.3f | main prefix sums + write %.3f | wait for the side stream %.3f ms\n", ms[1], ms[2], ms[3], ms[4], ms[5]);
        for (int q = 0; q < npe; q++) cudaEventDestroy(pe[q]);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// The SZ 1.4 Huffman container, stand-alone (ref: Huffman.c:790-816 encode_withTree, :865-889 decode_withTree; used by the
// reference's 1D / withLinearRegression = NO storage, TightDataPointStorageF.c:296-301):
//     [nodeCount : 4 BE][stateNum / 2 : 4 BE][tree bytes][Huffman bits of the n symbols]
// Same kernels as the blocked path: histogram, exact heap, code table + serialised tree, bit lengths, prefix sum, bit scatter; and the
// marker-less parallel decode on the way back.  Symbols are 16-bit (stateNum <= 65536) in device memory, in their own (linear) order.
extern "C" int szb200_encode_with_tree(const void *d_symbols_u16, size_t n, unsigned int stateNum, void *dst, int dst_on_device,
                                       size_t dst_capacity, size_t *out_size) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS;
    Ctx &c = *cp;
    if (!d_symbols_u16 || !dst || !out_size || n == 0 || stateNum == 0 || stateNum > 65536 || n >= (1ull << 32)) { set_err("encode_with_tree: bad arguments"); return SZ_NSCS; }
    const uint16_t *d_sym = (const uint16_t *)d_symbols_u16;
    TreeWs tw;
    if (tree_ws(c, 0, stateNum, tw)) return SZ_NSCS;
    HuffTreeDesc *d_desc = (HuffTreeDesc *)c.get("tree_desc", sizeof(HuffTreeDesc) * 8); if (!d_desc) return SZ_NSCS;
    const uint32_t heap_cap = (uint32_t)((c.smem_optin - 8192) / 8);
    const size_t heap_sm = 8ull * heap_cap;
    CU(cudaFuncSetAttribute(huff_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)heap_sm));
    CU(cudaMemcpyAsync(d_desc, &tw.d, sizeof(HuffTreeDesc), cudaMemcpyHostToDevice, c.stream));
    if (launch_hist(c, c.stream, d_sym, n, stateNum, tw.hist)) return SZ_NSCS;
    CU(launch_huff_build(c.stream, d_desc, 1, heap_sm, heap_cap, std::min<uint64_t>(stateNum, n))); KCHECK();
    const uint64_t nchunk = (n + ENC_CHUNK - 1) / ENC_CHUNK;
    uint32_t *d_cb = (uint32_t *)c.get("enc_cb", 4 * (nchunk + 8)), *d_cz = (uint32_t *)c.get("enc_cz", 4 * (nchunk + 8));
    unsigned long long *d_cbo = (unsigned long long *)c.get("enc_cbo", 8 * (nchunk + 8)), *d_czo = (unsigned long long *)c.get("enc_czo", 8 * (nchunk + 8));
    unsigned long long *d_tot = (unsigned long long *)c.get("totals", 256);
    if (!d_cb || !d_cz || !d_cbo || !d_czo || !d_tot) return SZ_NSCS;
    Geom g = make_geom(2, 1, 1, 1);
    enc_len_kernel<false><<<(unsigned)nchunk, ENC_THREADS, 0, c.stream>>>(d_sym, n, g, tw.d.clen, d_cb, d_cz); KCHECK();
    if (launch_scan(c, c.stream, 0, d_cb, nchunk, d_cbo, d_tot + 1)) return SZ_NSCS;
    unsigned long long h_bits = 0; uint32_t h_tree[8];
    CU(cudaMemcpyAsync(&h_bits, d_tot + 1, 8, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaMemcpyAsync(h_tree, tw.d.out, 32, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (h_tree[3]) { set_err(h_tree[3] == 2 ? "Huffman tree build gave up" : "Huffman code longer than 64 bits"); return SZ_NSCS; }
    const uint64_t nbytes = (h_bits + 7) / 8;
    uint32_t *d_bits = (uint32_t *)c.get("enc_bits", nbytes + 64); if (!d_bits) return SZ_NSCS;
    CU(cudaMemsetAsync(d_bits, 0, nbytes + 64, c.stream));
    enc_write_kernel<float, false><<<(unsigned)nchunk, ENC_THREADS, 0, c.stream>>>(d_sym, n, g, tw.d.code, tw.d.clen, d_cbo, d_czo, d_bits, (const float *)nullptr, (float *)nullptr); KCHECK();
    Emitter E;
    { unsigned char *p = E.h(8); put_be32(p, h_tree[1]); put_be32(p + 4, stateNum / 2); }
    E.d(tw.d.tree_bytes, h_tree[0]);
    E.d(d_bits, nbytes);
    if (E.total > dst_capacity) { set_err("encode_with_tree: %zu bytes do not fit the destination (%zu)", E.total, dst_capacity); *out_size = E.total; return SZB200_BODY_TOO_LARGE; }
    if (emit_all(c, E, dst, dst_on_device)) return SZ_NSCS;
    *out_size = E.total;
    return 0;
}

extern "C" int szb200_decode_with_tree(const void *container, size_t len, int on_device, size_t n, void *d_symbols_out_u16) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS;
    Ctx &c = *cp;
    if (!container || !d_symbols_out_u16 || len < 8 || n == 0 || n >= (1ull << 32)) { set_err("decode_with_tree: bad arguments"); return SZ_NSCS; }
    unsigned char *d_body = (unsigned char *)c.get("dec_body", len + 256); if (!d_body) return SZ_NSCS;
    CU(cudaMemcpyAsync(d_body, container, len, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c.stream));
    unsigned char head[8];
    CU(cudaMemcpyAsync(head, d_body, 8, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    const uint32_t nodeCount = ((uint32_t)head[0] << 24) | ((uint32_t)head[1] << 16) | ((uint32_t)head[2] << 8) | head[3];
    const uint64_t w = nodeCount <= 256 ? 1 : (nodeCount <= 65536 ? 2 : 4), tree_bytes = 1 + (uint64_t)nodeCount * (2 * w + 5);   // ref: Huffman.c:881-886
    if (nodeCount == 0 || nodeCount > 2 * 65536 || 8 + tree_bytes > len) { set_err("decode_with_tree: not a Huffman container (nodeCount %u)", nodeCount); return SZ_NSCS; }
    DecJob job; memset(&job, 0, sizeof(job));
    job.slot = 1; job.d_tree = d_body + 8; job.nodeCount = nodeCount; job.d_payload = d_body + 8 + tree_bytes; job.payload_bytes = len - 8 - tree_bytes;
    job.nsym = n; job.blockmajor = false; job.d_out = (uint16_t *)d_symbols_out_u16;
    unsigned long long *d_totals = nullptr; uint64_t expect[5] = {0, 0, 0, 0, 0};
    const Geom g = make_geom(2, 1, 1, 1);
    c.launches = 0;
    if (huff_decode_all<float>(c, &job, 1, g, (const float *)nullptr, 0, (float *)nullptr, nullptr, &d_totals, expect)) return SZ_NSCS;
    unsigned long long h_tot[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (d_totals) CU(cudaMemcpyAsync(h_tot, d_totals, 64, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (expect[0] && h_tot[0] < expect[0]) { set_err("decode_with_tree: the container holds %llu symbols, %zu expected", h_tot[0], n); return SZ_NSCS; }
    return 0;
}

template <typename T>
static int decompress_impl(Ctx &c, const szb200_job &job, const void *body, size_t body_len, int body_on_device, T *d_out, szb200_stats *stats) {
    const Geom g = make_geom(job.dim, job.n0, job.n1, job.n2);
    const int ncoef = g.ncoef; const uint32_t nb = g.num_blocks; const uint64_t N = g.N;
    if (N >= (1ull << 32)) { set_err("arrays of 2^32 or more elements must be cut into slabs"); return SZ_DERR; }
    c.launches = 0;
    CU(cudaStreamSynchronize(c.stream2));       // a call that failed half-way may have left work on the side stream
    int evi = 0; int stage_of_ev[24];
    CU(cudaEventRecord(c.ev[evi++], c.stream));
    auto mark = [&](int st) -> int { stage_of_ev[evi] = st; CU(cudaEventRecord(c.ev[evi++], c.stream)); return 0; };
    // ---- device copy of the body (aligned), host copy of its parseable head ----
    unsigned char *d_body = (unsigned char *)c.get("dec_body", body_len + 256); if (!d_body) return SZ_NSCS;
    if (!body_on_device && body_len >= ((size_t)4 << 20) && !is_pinned_host(body)) { if (h2d_bounced(c, d_body, body, body_len)) return SZ_NSCS; }
    else CU(cudaMemcpyAsync(d_body, body, body_len, body_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c.stream));
    // the section headers are walked on the device (dec_parse_kernel); the host reads the result back once
    DecHead *d_head = (DecHead *)c.get("dec_head", sizeof(DecHead) + 64); if (!d_head) return SZ_NSCS;
    dec_parse_kernel<<<1, 256, 0, c.stream>>>(d_body, (unsigned long long)body_len, (int)sizeof(T), ncoef, nb, d_head); KCHECK();
    DecHead H;
    CU(cudaMemcpyAsync(&H, d_head, sizeof(H), cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    if (H.err) { set_err(H.err == 1 ? "truncated stream" : "unsupported stream"); return SZ_NSCS; }
    if (H.block_size != (uint32_t)(job.dim == 3 ? 6 : 16)) { set_err("unexpected block size %u in stream", H.block_size); return SZ_NSCS; }
    auto bitsT = [](unsigned long long b) -> T { T v; if (sizeof(T) == 4) { uint32_t u = (uint32_t)b; memcpy(&v, &u, 4); } else memcpy(&v, &b, 8); return v; };
    const T eb = bitsT(H.eb_bits);
    const uint32_t intervals = H.intervals, tree_size = H.tree_size, nodeCount = H.node_count;
    if (intervals > 65536 || nodeCount == 0 || nodeCount > 2 * 65536) { set_err("unsupported stream: intervals=%u nodeCount=%u", intervals, nodeCount); return SZ_NSCS; }
    const size_t o_tree = (size_t)H.o_tree, o_ind = (size_t)H.o_ind;
    const int use_mean = (int)H.use_mean;
    const T mean = bitsT(H.mean_bits);
    const uint32_t reg_count = H.reg_count;
    T prec[4] = {0, 0, 0, 0}; int crad[4] = {0, 0, 0, 0}; size_t o_ctree[4] = {0, 0, 0, 0}, o_cbits[4] = {0, 0, 0, 0}, o_cunp[4] = {0, 0, 0, 0};
    uint32_t cnode[4] = {0, 0, 0, 0}, cunp_cnt[4] = {0, 0, 0, 0}; uint64_t cbits_bytes[4] = {0, 0, 0, 0};
    if (reg_count > 0) {
        for (int e = 0; e < ncoef; e++) {
            prec[e] = bitsT(H.prec_bits[e]); crad[e] = (int)H.crad[e]; cnode[e] = H.cnode[e]; cunp_cnt[e] = H.cunp_cnt[e]; cbits_bytes[e] = H.cbits_bytes[e];
            o_ctree[e] = (size_t)H.o_ctree[e]; o_cbits[e] = (size_t)H.o_cbits[e]; o_cunp[e] = (size_t)H.o_cunp[e];
            if (crad[e] != 32768 || cnode[e] == 0 || cnode[e] > 2 * 65536) { set_err("unsupported coefficient section"); return SZ_NSCS; }
        }
    }
    const uint64_t total_unpred = H.total_unpred;
    const size_t o_unp = (size_t)H.o_unp, o_bits = (size_t)H.o_bits;
    const uint64_t main_bytes = body_len - o_bits;
    // Nothing read from the stream sizes a copy or a loop before it has been checked against what the job and the buffer allow:
    // a cut or corrupt stream is refused here instead of writing past a workspace (ref: the reference trusts these fields).
    {
        auto tree_bytes_of = [](uint32_t n) -> uint64_t { const uint64_t w = n <= 256 ? 1 : (n <= 65536 ? 2 : 4); return 1 + (uint64_t)n * (2 * w + 5); };
        bool ok = reg_count <= nb && total_unpred <= N && o_bits <= body_len && o_unp <= body_len && o_unp + (uint64_t)sizeof(T) * total_unpred <= body_len
                  && tree_bytes_of(nodeCount) <= tree_size && o_tree + (uint64_t)tree_size <= body_len && o_ind + (nb + 7) / 8 <= body_len;
        for (int e = 0; ok && reg_count > 0 && e < ncoef; e++)
            ok = cunp_cnt[e] <= reg_count && o_ctree[e] + tree_bytes_of(cnode[e]) <= body_len && o_cbits[e] + cbits_bytes[e] <= body_len
                 && o_cunp[e] + (uint64_t)sizeof(T) * cunp_cnt[e] <= body_len;
        if (!ok) { set_err("corrupt stream: a section header does not fit the stream or the array"); return SZ_NSCS; }
    }
    if (mark(ST_PARSE)) return SZ_NSCS;

    // ---- aligned copies of the T-typed sections ----
    T *d_unp = (T *)c.get("dec_unpred", sizeof(T) * (total_unpred + 8)); if (!d_unp) return SZ_NSCS;
    if (total_unpred) CU(cudaMemcpyAsync(d_unp, d_body + o_unp, sizeof(T) * total_unpred, cudaMemcpyDeviceToDevice, c.stream));
    unsigned char *d_ind = (unsigned char *)c.get("indicator", nb + 64); if (!d_ind) return SZ_NSCS;
    indicator_unpack_kernel<<<(nb + 255) / 256, 256, 0, c.stream>>>(d_body + o_ind, nb, d_ind); KCHECK();

    // ---- Huffman decode: main codes -> raster code array (+ unpredictable values pre-placed in the output) ----
    uint16_t *d_codes = (uint16_t *)c.get("codes", 2ull * N + 64); if (!d_codes) return SZ_NSCS;
    T *d_clast = (T *)c.get("coeff_last", sizeof(T) * 4ull * nb); if (!d_clast) return SZ_NSCS;
    uint16_t *d_ctype = (uint16_t *)c.get("coeff_type", 2ull * 4 * nb + 64); if (!d_ctype) return SZ_NSCS;
    // ---- coefficient chain: needs the coefficient codes only, so it runs on the second stream beside the main write pass ----
    // the list of regression blocks and the raw coefficient values do not depend on any decoded code: prepared up front
    uint32_t *d_list = nullptr; T *d_cu = nullptr;
    if (reg_count > 0) {
        uint32_t *d_flags = (uint32_t *)c.get("reg_flags", 4ull * nb + 64);
        d_list = (uint32_t *)c.get("reg_list", 4ull * nb + 64);
        unsigned long long *d_rank = (unsigned long long *)c.get("reg_rank", 8ull * nb + 64);
        d_cu = (T *)c.get("coeff_unpred", sizeof(T) * 4ull * nb);
        T *d_dummy = (T *)c.get("reg", sizeof(T) * 4ull * nb), *d_dummy2 = (T *)c.get("reg_c", sizeof(T) * 4ull * nb);
        if (!d_flags || !d_list || !d_rank || !d_cu || !d_dummy || !d_dummy2) return SZ_NSCS;
        reg_flag_kernel<<<(nb + 255) / 256, 256, 0, c.stream>>>(d_ind, nb, d_flags); KCHECK();
        if (launch_scan(c, c.stream, 0, d_flags, nb, d_rank, nullptr)) return SZ_NSCS;
        reg_gather_kernel<T><<<(nb + 255) / 256, 256, 0, c.stream>>>(d_ind, d_rank, nb, 0, d_dummy, d_dummy2, d_list); KCHECK();
        for (int e = 0; e < ncoef; e++) if (cunp_cnt[e]) CU(cudaMemcpyAsync(d_cu + (size_t)e * nb, d_body + o_cunp[e], sizeof(T) * cunp_cnt[e], cudaMemcpyDeviceToDevice, c.stream));
    }
    unsigned long long *d_dec_totals = nullptr; uint64_t dec_expect[5] = {0, 0, 0, 0, 0};
    auto chain_tail = [&](cudaStream_t st) -> int {
        if (reg_count == 0) return 0;
        // The serial chains run beside the main write pass.  They ask for shared memory they do not use, so that at most one CTA of that
        // pass fits on an SM next to a chain: the chain's speed is the issue rate its one thread gets.
        const int chain_pad = (int)std::min<size_t>(c.smem_optin > 80 * 1024 ? c.smem_optin - 80 * 1024 : 0, 150 * 1024);
        CU(cudaFuncSetAttribute(coeff_chain_decode_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, chain_pad));
        coeff_chain_decode_kernel<T><<<ncoef, LAT_THREADS, chain_pad, st>>>(d_ctype, d_cu, nb, reg_count, prec[0], prec[1], prec[2], prec[3], crad[0], crad[1], crad[2], crad[3], d_list, d_clast); KCHECK();
        return 0;
    };
    {
        DecJob jobs[5]; memset(jobs, 0, sizeof(jobs));
        int nj = 0;
        jobs[nj].slot = 0; jobs[nj].d_tree = d_body + o_tree; jobs[nj].nodeCount = nodeCount; jobs[nj].d_payload = d_body + o_bits; jobs[nj].payload_bytes = main_bytes;
        jobs[nj].nsym = N; jobs[nj].blockmajor = true; jobs[nj].d_out = d_codes; nj++;
        if (reg_count > 0) for (int e = 0; e < ncoef; e++) {
            jobs[nj].slot = 1 + e; jobs[nj].d_tree = d_body + o_ctree[e]; jobs[nj].nodeCount = cnode[e]; jobs[nj].d_payload = d_body + o_cbits[e]; jobs[nj].payload_bytes = cbits_bytes[e];
            jobs[nj].nsym = reg_count; jobs[nj].blockmajor = false; jobs[nj].d_out = d_ctype + (size_t)e * nb; nj++;
        }
        if (huff_decode_all<T>(c, jobs, nj, g, d_unp, total_unpred, d_out, chain_tail, &d_dec_totals, dec_expect)) return SZ_NSCS;
    }
    if (mark(ST_DECODE)) return SZ_NSCS;

    if (mark(ST_DCHAIN)) return SZ_NSCS;

    // ---- wavefront reconstruction ----
    {
        const T recip = (T)1 / eb;
        int rc = (job.dim == 3) ? launch_wavefront<T, 3>(c, g, true, use_mean, (int)intervals, d_ind, (const T *)nullptr, d_codes, d_out, d_clast, eb, recip, mean, 0)
                                : launch_wavefront<T, 2>(c, g, true, use_mean, (int)intervals, d_ind, (const T *)nullptr, d_codes, d_out, d_clast, eb, recip, mean, 0);
        if (rc) return rc;
    }
    if (mark(ST_RECON)) return SZ_NSCS;
    unsigned long long h_dec_totals[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (d_dec_totals) CU(cudaMemcpyAsync(h_dec_totals, d_dec_totals, sizeof(h_dec_totals), cudaMemcpyDeviceToHost, c.stream));
    if (check_wf_err(c)) return SZ_NSCS;          // (synchronises the stream)
    for (int j = 0; j < 5; j++)
        if (dec_expect[j] && h_dec_totals[j] < dec_expect[j]) {
            set_err("truncated stream: Huffman section %d holds %llu symbols, %llu expected", j, h_dec_totals[j], (unsigned long long)dec_expect[j]);
            return SZ_NSCS;
        }
    szb200_stats st; memset(&st, 0, sizeof(st));
    st.intervals = intervals; st.use_mean = use_mean; st.mean = (double)mean; st.num_blocks = nb; st.reg_count = reg_count; st.total_unpred = total_unpred;
    st.tree_bytes = tree_size; st.node_count = nodeCount; st.type_bits_bytes = main_bytes; st.body_bytes = body_len; st.launches = c.launches;
    for (int i = 1; i < evi; i++) { float ms = 0; cudaEventElapsedTime(&ms, c.ev[i - 1], c.ev[i]); st.ms_stage[stage_of_ev[i]] += ms; }
    cudaEventElapsedTime(&st.ms_total, c.ev[0], c.ev[evi - 1]);
    c.last = st; c.have_last = true;
    if (stats) *stats = st;
    return 0;
}

extern "C" int szb200_decompress_blocked(const szb200_job *job, const void *body, size_t body_len, int body_on_device, void *d_out, szb200_stats *stats) {
    Ctx *c = ctx(); if (!c) return SZ_NSCS;
    if (job->dim != 2 && job->dim != 3) { set_err("dim must be 2 or 3"); return SZ_DERR; }
    if (job->dtype == 0) return decompress_impl<float>(*c, *job, body, body_len, body_on_device, (float *)d_out, stats);
    if (job->dtype == 1) return decompress_impl<double>(*c, *job, body, body_len, body_on_device, (double *)d_out, stats);
    set_err("dtype must be SZ_FLOAT or SZ_DOUBLE"); return SZ_NSCS;
}

// ---------------------------------------------------------------------------------------------------------------------
// small element-wise kernels for the constant / lossless containers and protectValueRange
template <typename U> __global__ void swap_copy_kernel(const unsigned char *__restrict__ src, unsigned char *__restrict__ dst, uint64_t n) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        unsigned char b[sizeof(U)];
#pragma unroll
        for (int q = 0; q < (int)sizeof(U); q++) b[q] = src[i * sizeof(U) + q];
#pragma unroll
        for (int q = 0; q < (int)sizeof(U); q++) dst[i * sizeof(U) + q] = b[sizeof(U) - 1 - q];
    }
}
template <typename T> __global__ void fill_kernel(T *out, uint64_t n, T v) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = v;
}
template <typename T> __global__ void clamp_kernel(T *out, uint64_t n, T mn, T mx) {     // ref: szd_float.c:161-176
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        T v = out[i];
        if (v <= mx && v >= mn) continue;
        if (v < mn) out[i] = mn; else if (v > mx) out[i] = mx;
    }
}

extern "C" int szb200_swap_copy(int dtype, const void *src, int src_on_device, void *dst, int dst_on_device, size_t n) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS; Ctx &c = *cp;
    const size_t es = dtype == 0 ? 4 : 8, bytes = n * es;
    const unsigned char *s = (const unsigned char *)src; unsigned char *d = (unsigned char *)dst;
    if (!src_on_device) { unsigned char *t = (unsigned char *)c.get("swap_in", bytes + 64); if (!t) return SZ_NSCS; CU(cudaMemcpyAsync(t, src, bytes, cudaMemcpyHostToDevice, c.stream)); s = t; }
    if (!dst_on_device) { d = (unsigned char *)c.get("swap_out", bytes + 64); if (!d) return SZ_NSCS; }
    if (dtype == 0) swap_copy_kernel<uint32_t><<<c.sm_count * 8, 256, 0, c.stream>>>(s, d, n); else swap_copy_kernel<uint64_t><<<c.sm_count * 8, 256, 0, c.stream>>>(s, d, n);
    KCHECK();
    if (!dst_on_device) CU(cudaMemcpyAsync(dst, d, bytes, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    return 0;
}
extern "C" int szb200_fill(int dtype, double value, void *d_out, size_t n) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS; Ctx &c = *cp;
    if (dtype == 0) fill_kernel<float><<<c.sm_count * 8, 256, 0, c.stream>>>((float *)d_out, n, (float)value); else fill_kernel<double><<<c.sm_count * 8, 256, 0, c.stream>>>((double *)d_out, n, value);
    KCHECK(); CU(cudaStreamSynchronize(c.stream)); return 0;
}
extern "C" int szb200_clamp(int dtype, void *d_out, size_t n, double vmin, double vmax) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS; Ctx &c = *cp;
    if (dtype == 0) clamp_kernel<float><<<c.sm_count * 8, 256, 0, c.stream>>>((float *)d_out, n, (float)vmin, (float)vmax);
    else clamp_kernel<double><<<c.sm_count * 8, 256, 0, c.stream>>>((double *)d_out, n, vmin, vmax);
    KCHECK(); CU(cudaStreamSynchronize(c.stream)); return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
extern "C" long long szb200_debug_fetch(int which, void *host_dst, size_t cap) {
    Ctx *cp = ctx(); if (!cp) return SZ_NSCS; Ctx &c = *cp;
    if (c.dbg_dtype < 0) return SZ_NSCS;
    const size_t ts = c.dbg_dtype == 0 ? 4 : 8; const Geom &g = c.dbg_g; const size_t nb = g.num_blocks;
    const char *name = nullptr; size_t bytes = 0;
    switch (which) {
        case 0: name = "reg"; bytes = ts * g.ncoef * nb; break;
        case 1: name = "indicator"; bytes = nb; break;
        case 2: name = "coeff_last"; bytes = ts * g.ncoef * nb; break;
        case 3: name = "codes"; bytes = 2 * g.N; break;
        case 4: name = "recon"; bytes = ts * g.N; break;
        case 5: name = "coeff_type"; bytes = 2 * g.ncoef * nb; break;
        case 6: name = "tree0"; bytes = 4ull * c.dbg_intervals; break;
        default: return SZ_NSCS;
    }
    auto it = c.bufs.find(name); if (it == c.bufs.end() || !it->second.p) return SZ_NSCS;
    if (bytes > cap) bytes = cap;
    CU(cudaMemcpyAsync(host_dst, it->second.p, bytes, cudaMemcpyDeviceToHost, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    return (long long)bytes;
}
