// szb_stage1.cuh -- kernels that run before the wavefront: range scan (K1), sampling mean, interval-optimiser
// sampling + finish (K3), use_mean mean (K4), regression fit + predictor selection (K2), regression-block
// compaction and the coefficient DPCM chain (K5).  See DESIGN.md for the roofline of each.
#pragma once
#include <string.h>
#include "szb_common.cuh"

namespace szb {

// =====================================================================================================================
// K1: computeRangeSize_float/_double (ref: sz/src/dataCompression.c:102-119, :149-166)
// value-only min/max: 128-bit loads, warp-shuffle reduction, per-CTA partials, last-CTA final reduction.
// (+-0 ties: the reference keeps the FIRST occurrence; resolved by first_zero_kernel only when min or max == 0.)
// =====================================================================================================================
template <typename T> struct MinMax { T mn, mx; };

template <typename T>
__device__ __forceinline__ void mm_update(T v, T &mn, T &mx) { mn = v < mn ? v : mn; mx = v > mx ? v : mx; }

template <typename T>
__global__ void __launch_bounds__(512) range_kernel(const T *__restrict__ x, uint64_t n, MinMax<T> *partials,
                                                    unsigned int *done, MinMax<T> *result) {
    constexpr int VEC = 16 / sizeof(T);
    T mn = x[0], mx = x[0];
    // scalar head up to 16-byte alignment
    uint64_t head = ((16 - ((uintptr_t)x & 15)) & 15) / sizeof(T);
    if (head > n) head = n;
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    if (tid < head) mm_update(x[tid], mn, mx);
    const T *xa = x + head;
    uint64_t nvec = (n - head) / VEC;
    if (sizeof(T) == 4) {
        const float4 *xv = reinterpret_cast<const float4 *>(xa);
        for (uint64_t i = tid; i < nvec; i += nthreads) {
            float4 v = __ldg(xv + i);
            mm_update((T)v.x, mn, mx); mm_update((T)v.y, mn, mx); mm_update((T)v.z, mn, mx); mm_update((T)v.w, mn, mx);
        }
    } else {
        const double2 *xv = reinterpret_cast<const double2 *>(xa);
        for (uint64_t i = tid; i < nvec; i += nthreads) {
            double2 v = __ldg(xv + i);
            mm_update((T)v.x, mn, mx); mm_update((T)v.y, mn, mx);
        }
    }
    uint64_t tail0 = head + nvec * VEC;
    if (tail0 + tid < n) mm_update(x[tail0 + tid], mn, mx);   // tail < VEC elements

    for (int o = 16; o > 0; o >>= 1) {
        T a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
        mn = a < mn ? a : mn; mx = b > mx ? b : mx;
    }
    __shared__ T s_mn[16], s_mx[16];
    __shared__ bool s_last;
    int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    if (l == 0) { s_mn[w] = mn; s_mx[w] = mx; }
    __syncthreads();
    if (w == 0) {
        mn = l < nw ? s_mn[l] : s_mn[0]; mx = l < nw ? s_mx[l] : s_mx[0];
        for (int o = 8; o > 0; o >>= 1) {
            T a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
            mn = a < mn ? a : mn; mx = b > mx ? b : mx;
        }
        if (l == 0) {
            partials[blockIdx.x].mn = mn; partials[blockIdx.x].mx = mx;
            __threadfence();
            unsigned int t = atomicAdd(done, 1u);
            s_last = (t == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (s_last && w == 0) {
        __threadfence();
        mn = partials[0].mn; mx = partials[0].mx;
        for (unsigned int i = l; i < gridDim.x; i += 32) {
            T a = __ldcg(&partials[i].mn), b = __ldcg(&partials[i].mx);
            mn = a < mn ? a : mn; mx = b > mx ? b : mx;
        }
        for (int o = 16; o > 0; o >>= 1) {
            T a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
            mn = a < mn ? a : mn; mx = b > mx ? b : mx;
        }
        if (l == 0) { result->mn = mn; result->mx = mx; *done = 0; }
    }
}

// first index whose value compares equal to `v` (only launched when min or max is a zero: sign-of-zero tie-break)
template <typename T>
__global__ void first_equal_kernel(const T *__restrict__ x, uint64_t n, T v, unsigned long long *idx) {
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long best = ~0ULL;
    for (uint64_t i = tid; i < n; i += nthreads) if (x[i] == v) { best = i; break; }
    if (best != ~0ULL) atomicMin(idx, best);
}

// =====================================================================================================================
// sampling mean of the interval optimiser (S3; ref: sz_float.c:6400-6421 / 2D :5409-5419): sequential T sum over
// ~sqrt(N) strided samples.  Positions depend on the shape only and are precomputed on the host (pos[]).
// Gather in parallel, then ONE thread adds in the reference's order (float addition is not associative).
// =====================================================================================================================
template <typename T>
__global__ void __launch_bounds__(1024) sample_mean_kernel(const T *__restrict__ x, const uint64_t *__restrict__ pos, uint32_t count,
                                                           T *vals, T *mean_out, int in_smem) {
    // the gathered samples are staged in shared memory when they fit (the serial adder then reads at shared-memory latency,
    // eight values per batch), else in the global scratch array
    extern __shared__ __align__(16) unsigned char s_raw[];
    T *buf = in_smem ? reinterpret_cast<T *>(s_raw) : vals;
    for (uint32_t i = threadIdx.x; i < count; i += blockDim.x) buf[i] = x[pos[i]];
    __syncthreads();
    if (threadIdx.x == 0) {
        T mean = (T)0;
        uint32_t i = 0;
        for (; i + 8 <= count; i += 8) {
            T v0 = buf[i], v1 = buf[i + 1], v2 = buf[i + 2], v3 = buf[i + 3], v4 = buf[i + 4], v5 = buf[i + 5], v6 = buf[i + 6], v7 = buf[i + 7];
            mean += v0; mean += v1; mean += v2; mean += v3; mean += v4; mean += v5; mean += v6; mean += v7;
        }
        for (; i < count; i++) mean += buf[i];
        if (count > 0) mean = mean / (T)(unsigned long long)count;
        *mean_out = mean;
    }
}

// =====================================================================================================================
// K3: optimize_intervals_*_with_freq_and_dense_pos (ref: sz_float.c:6396-6523, 2D :5405-5512; double sz_double.c:5773, :4790)
// One thread per sample ROW (closed form of the pointer walk, SURVEY.md Appendix C, validated in tests):
//   3D: rows (i,j), i>=1, j=1..n1-1; first column c0 = sd - ((i+j) mod sd)   (very first row: sd-2)
//   2D: rows i>=1;                    first column c0 = sd - (i mod sd)       (very first row: sd-1)
//   samples c0, c0+sd, ... < n2; the first sample of a row is always taken (it may spill into following rows).
// Histograms live in shared memory (low bins) and are flushed with one global atomic per non-zero bin.
// =====================================================================================================================
struct OptAccum {          // device-resident accumulator, zeroed before the launch
    unsigned long long hits, samples;
};
constexpr int OPT_SMALL_BINS = 2048;

template <typename T, int DIM>
__global__ void __launch_bounds__(256) opt_sample_kernel(const T *__restrict__ x, Geom g, uint32_t sd, uint64_t row_limit,
                                                         const T *__restrict__ mean_ptr, double eb, uint32_t maxRangeRadius,
                                                         uint32_t *g_intervals, uint32_t *g_freq, OptAccum *acc) {
    __shared__ uint32_t s_int[OPT_SMALL_BINS];
    __shared__ uint32_t s_freq[8192];
    __shared__ unsigned int s_hits, s_samples;
    for (int i = threadIdx.x; i < OPT_SMALL_BINS; i += blockDim.x) s_int[i] = 0;
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) s_freq[i] = 0;
    if (threadIdx.x == 0) { s_hits = 0; s_samples = 0; }
    __syncthreads();
    const T mean = *mean_ptr;
    const uint64_t n1 = g.ax[1].n, n2 = g.ax[2].n;
    const int64_t s1 = (int64_t)n2, s0 = (int64_t)(n1 * n2);
    unsigned int hits = 0, samples = 0;
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < row_limit; r += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t base; int64_t c0;
        if (DIM == 3) {
            uint64_t i = 1 + r / (n1 - 1), j = 1 + r % (n1 - 1);
            c0 = (r == 0) ? (int64_t)sd - 2 : (int64_t)(sd - (uint32_t)((i + j) % sd));
            base = i * (uint64_t)s0 + j * (uint64_t)s1;
        } else {
            uint64_t i = 1 + r;
            c0 = (r == 0) ? (int64_t)sd - 1 : (int64_t)(sd - (uint32_t)(i % sd));
            base = i * (uint64_t)s1;
        }
        for (int64_t c = c0, m = 0; m == 0 || c < (int64_t)n2; c += sd, m++) {
            const T *d = x + base + c;
            T cur = d[0];
            T pv;
            if (DIM == 3) {   // term order of the OPTIMISER (ref: sz_float.c:6448), differs from the main loop
                pv = d[-1] + d[-s1];
                pv = pv + d[-s0];
                pv = pv - d[-1 - s0];
                pv = pv - d[-s1 - 1];
                pv = pv - d[-s1 - s0];
                pv = pv + d[-s1 - s0 - 1];
            } else {          // ref: sz_float.c:5444
                pv = d[-1] + d[-s1];
                pv = pv - d[-s1 - 1];
            }
            T dif = pv - cur;
            T pred_err = TT<T>::abs(dif);
            if ((double)pred_err < eb) hits++;
            unsigned long long ri = x86_d2u64(((double)pred_err / eb + 1.0) / 2.0);
            if (ri >= maxRangeRadius) ri = maxRangeRadius - 1;
            if (ri < OPT_SMALL_BINS) atomicAdd(&s_int[ri], 1u); else atomicAdd(&g_intervals[ri], 1u);
            T mean_diff = cur - mean;
            long long q = x86_cvttsd2si((double)mean_diff / eb);
            unsigned long long fu = (mean_diff > (T)0) ? (unsigned long long)q + 4096ULL : (unsigned long long)q - 1ULL + 4096ULL;
            long long fi = (long long)fu;
            int bin = (fi <= 0) ? 0 : ((unsigned long long)fi >= 8192ULL ? 8191 : (int)fi);
            atomicAdd(&s_freq[bin], 1u);
            samples++;
        }
    }
    atomicAdd(&s_hits, hits); atomicAdd(&s_samples, samples);
    __syncthreads();
    for (int i = threadIdx.x; i < OPT_SMALL_BINS; i += blockDim.x) { uint32_t v = s_int[i]; if (v && (uint32_t)i < maxRangeRadius) atomicAdd(&g_intervals[i], v); }
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) { uint32_t v = s_freq[i]; if (v) atomicAdd(&g_freq[i], v); }
    if (threadIdx.x == 0) { atomicAdd(&acc->hits, (unsigned long long)s_hits); atomicAdd(&acc->samples, (unsigned long long)s_samples); }
}

struct OptOut {            // read back by the host, which finishes the (scalar, double) arithmetic of ref :6486-6520
    uint32_t first_index;  // smallest i with cumulative count > targetCount (already clamped to maxRangeRadius-1)
    uint32_t max_index;
    unsigned long long max_sum, hits, samples;
    double mean;           // sampling mean widened to double (exact)
};

template <typename T>
__global__ void __launch_bounds__(1024) opt_finish_kernel(const uint32_t *__restrict__ g_intervals, const uint32_t *__restrict__ g_freq,
                                                          const OptAccum *acc, const T *mean_ptr, uint32_t maxRangeRadius,
                                                          float predThreshold, OptOut *out) {
    __shared__ unsigned long long s_part[1024];
    __shared__ unsigned int s_first;
    __shared__ unsigned long long s_best;       // packed (sum << 13 | (8191 - index)) so that max picks larger sum, then SMALLER index
    const unsigned long long samples = acc->samples;
    // size_t targetCount = sample_count * predThreshold  -> float product (ref: sz_float.c:6489)
    float tcf = (float)samples * predThreshold;
    unsigned long long target = x86_d2u64((double)tcf);
    int t = threadIdx.x;
    uint32_t per = (maxRangeRadius + 1023) / 1024;
    uint32_t lo = t * per, hi = min(lo + per, maxRangeRadius);
    unsigned long long local = 0;
    for (uint32_t i = lo; i < hi; i++) local += g_intervals[i];
    s_part[t] = local;
    if (t == 0) { s_first = 0xffffffffu; s_best = 0; }
    __syncthreads();
    // exclusive prefix of the per-thread sums (serial by one warp-free thread: 1024 adds, negligible)
    if (t == 0) { unsigned long long run = 0; for (int i = 0; i < 1024; i++) { unsigned long long v = s_part[i]; s_part[i] = run; run += v; } }
    __syncthreads();
    unsigned long long run = s_part[t];
    for (uint32_t i = lo; i < hi; i++) { run += g_intervals[i]; if (run > target) { atomicMin(&s_first, i); break; } }
    // densest adjacent pair, i = 1 .. 8189, FIRST maximum (ref: sz_float.c:6509-6516)
    unsigned long long best = 0;
    for (uint32_t i = 1 + t; i < 8192 - 2; i += 1024) {
        unsigned long long s = (unsigned long long)g_freq[i] + g_freq[i + 1];
        unsigned long long key = (s << 13) | (unsigned long long)(8191 - i);
        if (s > 0 && key > best) best = key;
    }
    if (best) atomicMax(&s_best, best);
    __syncthreads();
    if (t == 0) {
        uint32_t fi = s_first;
        if (fi == 0xffffffffu || fi >= maxRangeRadius) fi = maxRangeRadius - 1;
        out->first_index = fi;
        out->max_sum = s_best >> 13;
        out->max_index = s_best ? (uint32_t)(8191 - (s_best & 8191ULL)) : 0u;
        out->hits = acc->hits; out->samples = samples;
        out->mean = (double)(*mean_ptr);
    }
}

// =====================================================================================================================
// K4: the use_mean mean (S2; ref: sz_float.c:6657-6669): sequential T sum of all x with |x - dense_pos| < eb.
// Order matters (float addition), so ONE thread accumulates; the CTA stages tiles in shared memory so that the
// loads are coalesced and off the dependent chain.  Rare path (background-dominated 3D fields only).
// =====================================================================================================================
template <typename T>
__global__ void __launch_bounds__(1024) mean_kernel(const T *__restrict__ x, uint64_t n, T dense_pos, T eb, T *mean_out) {
    constexpr int TILE = 16384 / (int)sizeof(T);   // 2 x 16 KB of values + flags stays under the 48 KB static limit
    __shared__ T s_val[2][TILE];
    __shared__ unsigned char s_flag[2][TILE];
    T sum = (T)0;
    unsigned long long cnt = 0;
    int buf = 0;
    auto load = [&](uint64_t base, int b) {
        for (int i = threadIdx.x; i < TILE; i += blockDim.x) {
            uint64_t idx = base + i;
            T v = idx < n ? x[idx] : (T)0;
            T df = v - dense_pos;
            s_val[b][i] = v;
            s_flag[b][i] = (idx < n && TT<T>::abs(df) < eb) ? 1 : 0;
        }
    };
    load(0, 0);
    __syncthreads();
    for (uint64_t base = 0; base < n; base += TILE) {
        if (threadIdx.x >= 32) {                       // the other warps prefetch the next tile
            if (base + TILE < n) {
                for (int i = threadIdx.x - 32; i < TILE; i += blockDim.x - 32) {
                    uint64_t idx = base + TILE + i;
                    T v = idx < n ? x[idx] : (T)0;
                    T df = v - dense_pos;
                    s_val[buf ^ 1][i] = v;
                    s_flag[buf ^ 1][i] = (idx < n && TT<T>::abs(df) < eb) ? 1 : 0;
                }
            }
        } else if (threadIdx.x == 0) {
            for (int i = 0; i < TILE; i++) if (s_flag[buf][i]) { sum += s_val[buf][i]; cnt++; }
        }
        __syncthreads();
        buf ^= 1;
    }
    if (threadIdx.x == 0) *mean_out = cnt > 0 ? sum / (T)cnt : (T)0;
}

// =====================================================================================================================
// K2: regression fit (ref: sz_float.c:6586-6637, 2D :5569-5605) fused with the predictor selection by diagonal
// sampling (ref: sz_float.c:7083-7123 no-mean, :6747-6786 mean, 2D :6000-6028).  One thread per block; the nested
// accumulation order of the reference is kept exactly.  Adjacent threads own adjacent blocks along the fast axis,
// so a warp's loads of one (ii,jj) row touch one contiguous 32*edge*sizeof(T) span.
// =====================================================================================================================
template <typename T, int DIM>
__global__ void __launch_bounds__(128) fit_select_kernel(const T *__restrict__ x, Geom g, T noise, int use_mean, T mean,
                                                         T *__restrict__ reg /*[ncoef][num_blocks]*/, unsigned char *__restrict__ indicator) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= g.num_blocks) return;
    uint32_t b2 = b % g.ax[2].nb, t = b / g.ax[2].nb;
    uint32_t b1 = t % g.ax[1].nb, b0 = t / g.ax[1].nb;
    const uint32_t bx = g.ax[0].edge(b0), by = g.ax[1].edge(b1), bz = g.ax[2].edge(b2);
    const int64_t s0 = (int64_t)g.s0, s1 = (int64_t)g.s1;
    const T *blk = x + (uint64_t)g.ax[0].off(b0) * g.s0 + (uint64_t)g.ax[1].off(b1) * g.s1 + g.ax[2].off(b2);
    const uint32_t nb = g.num_blocks;
    T pa, pb, pc, pd = (T)0;
    if (DIM == 3) {
        T fx = (T)0, fy = (T)0, fz = (T)0, f = (T)0;
        for (uint32_t i = 0; i < bx; i++) {
            T sum_x = (T)0;
            for (uint32_t j = 0; j < by; j++) {
                T sum_y = (T)0;
                const T *row = blk + i * s0 + j * s1;
                for (uint32_t k = 0; k < bz; k++) {
                    T cur = __ldg(row + k);
                    sum_y = sum_y + cur;
                    T pr = cur * (T)k;
                    fz = fz + pr;
                }
                T pj = sum_y * (T)j;
                fy = fy + pj;
                sum_x = sum_x + sum_y;
            }
            T pi = sum_x * (T)i;
            fx = fx + pi;
            f = f + sum_x;
        }
        T coeff = (T)(1.0 / (double)((unsigned long long)bx * by * bz));
        T u;
        u = (T)2 * fx; u = u / (T)(bx - 1); u = u - f; u = u * (T)6; u = u * coeff; pa = u / (T)(bx + 1);
        u = (T)2 * fy; u = u / (T)(by - 1); u = u - f; u = u * (T)6; u = u * coeff; pb = u / (T)(by + 1);
        u = (T)2 * fz; u = u / (T)(bz - 1); u = u - f; u = u * (T)6; u = u * coeff; pc = u / (T)(bz + 1);
        T ta = (T)(bx - 1) * pa; ta = ta / (T)2;
        T tb = (T)(by - 1) * pb; tb = tb / (T)2;
        T tc = (T)(bz - 1) * pc; tc = tc / (T)2;
        T s = ta + tb; s = s + tc;
        T fc = f * coeff;
        pd = fc - s;
        reg[b] = pa; reg[nb + b] = pb; reg[2 * nb + b] = pc; reg[3 * nb + b] = pd;
    } else {
        // 2D: axis 1 is the slow one ("x" of the reference), axis 2 the fast one
        T fx = (T)0, fy = (T)0, f = (T)0;
        for (uint32_t i = 0; i < by; i++) {
            T sum_x = (T)0;
            const T *row = blk + i * s1;
            for (uint32_t j = 0; j < bz; j++) {
                T cur = __ldg(row + j);
                sum_x = sum_x + cur;
                T pr = cur * (T)j;
                fy = fy + pr;
            }
            T pi = sum_x * (T)i;
            fx = fx + pi;
            f = f + sum_x;
        }
        T coeff = (T)(1.0 / (double)((unsigned long long)by * bz));
        T u;
        u = (T)2 * fx; u = u / (T)(by - 1); u = u - f; u = u * (T)6; u = u * coeff; pa = u / (T)(by + 1);
        u = (T)2 * fy; u = u / (T)(bz - 1); u = u - f; u = u * (T)6; u = u * coeff; pb = u / (T)(bz + 1);
        T ta = (T)(by - 1) * pa; ta = ta / (T)2;
        T tb = (T)(bz - 1) * pb; tb = tb / (T)2;
        T s = ta + tb;
        T fc = f * coeff;
        pc = fc - s;
        reg[b] = pa; reg[nb + b] = pb; reg[2 * nb + b] = pc;
    }
    // ---- predictor selection on the ORIGINAL data ----
    T err_sz = (T)0, err_reg = (T)0;
    if (DIM == 3) {
        int bs = (int)min(bx, min(by, bz));
        for (int d = 1; d < bs; d++) {
            int bmi = bs - d;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                int pj = (q & 2) ? bmi : d, pk = (q & 1) ? bmi : d;
                const T *c = blk + (int64_t)d * s0 + (int64_t)pj * s1 + pk;
                T cur = __ldg(c);
                T ps = __ldg(c - 1) + __ldg(c - s1);
                ps = ps + __ldg(c - s0);
                ps = ps - __ldg(c - s1 - 1);
                ps = ps - __ldg(c - s0 - 1);
                ps = ps - __ldg(c - s0 - s1);
                ps = ps + __ldg(c - s0 - s1 - 1);
                T r0 = pa * (T)d, r1 = pb * (T)pj, r2 = pc * (T)pk;
                T pr = r0 + r1; pr = pr + r2; pr = pr + pd;
                T e1 = ps - cur; e1 = TT<T>::abs(e1); e1 = e1 + noise;
                if (use_mean) { T e2 = mean - cur; e2 = TT<T>::abs(e2); if (!(e1 < e2)) e1 = e2; }
                err_sz = err_sz + e1;
                T er = pr - cur; er = TT<T>::abs(er);
                err_reg = err_reg + er;
            }
        }
    } else {
        int bs = (int)min(by, bz);
        for (int d = 1; d < bs; d++) {
            int bmi = bs - d;
            const T *c = blk + (int64_t)d * s1 + d;
            T cur = __ldg(c);
            T ps = __ldg(c - 1) + __ldg(c - s1); ps = ps - __ldg(c - s1 - 1);
            T r0 = pa * (T)d, r1 = pb * (T)d;
            T pr = r0 + r1; pr = pr + pc;
            T e1 = ps - cur; e1 = TT<T>::abs(e1); e1 = e1 + noise; err_sz = err_sz + e1;
            T er = pr - cur; er = TT<T>::abs(er); err_reg = err_reg + er;
            c = blk + (int64_t)d * s1 + bmi;
            cur = __ldg(c);
            ps = __ldg(c - 1) + __ldg(c - s1); ps = ps - __ldg(c - s1 - 1);
            r0 = pa * (T)(d - 1); r1 = pb * (T)bmi;                 // quirk of the reference (sz_float.c:6023)
            pr = r0 + r1; pr = pr + pc;
            e1 = ps - cur; e1 = TT<T>::abs(e1); e1 = e1 + noise; err_sz = err_sz + e1;
            er = pr - cur; er = TT<T>::abs(er); err_reg = err_reg + er;
        }
    }
    indicator[b] = (err_reg < err_sz) ? 0 : 1;      // 1 = Lorenzo (ref: sz_float.c:7352)
}

// =====================================================================================================================
// exclusive scan (u32 in -> u64 out) over tiles of 8192 elements: tile sums, scan of the tile sums, per-tile scan.
// Used for block compaction, unpredictable ranks and bit offsets.
// =====================================================================================================================
constexpr int SCAN_TB = 1024, SCAN_PER = 8, SCAN_TILE = SCAN_TB * SCAN_PER;

__device__ __forceinline__ unsigned long long block_exclusive_scan_1024(unsigned long long v, unsigned long long *s_warp /*[32]*/, unsigned long long &block_total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = s_warp[lane], winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned long long u = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += u; }
        s_warp[lane] = winc - w;                    // exclusive prefix of the warp totals
        if (lane == 31) s_warp[32] = winc;          // block total
    }
    __syncthreads();
    block_total = s_warp[32];
    return s_warp[wid] + inc - v;
}

template <typename InT>
__global__ void __launch_bounds__(SCAN_TB) scan_tile_sum_kernel(const InT *__restrict__ in, uint64_t n, unsigned long long *__restrict__ tile_sums) {
    __shared__ unsigned long long s_warp[33];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_PER;
    unsigned long long v = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++) if (base + k < n) v += in[base + k];
    unsigned long long tot;
    (void)block_exclusive_scan_1024(v, s_warp, tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(SCAN_TB) scan_tile_offsets_kernel(unsigned long long *tile_sums, uint32_t ntiles, unsigned long long *total) {
    __shared__ unsigned long long s_warp[33];
    const uint32_t per = (ntiles + SCAN_TB - 1) / SCAN_TB;
    const uint32_t lo = min(ntiles, threadIdx.x * per), hi = min(ntiles, lo + per);
    unsigned long long v = 0;
    for (uint32_t i = lo; i < hi; i++) v += tile_sums[i];
    unsigned long long tot;
    unsigned long long run = block_exclusive_scan_1024(v, s_warp, tot);
    for (uint32_t i = lo; i < hi; i++) { unsigned long long t = tile_sums[i]; tile_sums[i] = run; run += t; }
    if (threadIdx.x == 0 && total) *total = tot;
}

template <typename InT>
__global__ void __launch_bounds__(SCAN_TB) scan_apply_kernel(const InT *__restrict__ in, uint64_t n, const unsigned long long *__restrict__ tile_offsets,
                                                              unsigned long long *__restrict__ out) {
    __shared__ unsigned long long s_warp[33];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_PER;
    InT e[SCAN_PER];
    unsigned long long v = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++) { e[k] = (base + k < n) ? in[base + k] : (InT)0; v += e[k]; }
    unsigned long long tot;
    unsigned long long run = block_exclusive_scan_1024(v, s_warp, tot) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_PER; k++) if (base + k < n) { out[base + k] = run; run += e[k]; }
}

// indicator -> (is regression) flags, and scatter of the compacted regression-block list + coefficients
__global__ void reg_flag_kernel(const unsigned char *__restrict__ indicator, uint32_t nb, uint32_t *flags) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb) flags[b] = indicator[b] ? 0u : 1u;
}
template <typename T>
__global__ void reg_gather_kernel(const unsigned char *__restrict__ indicator, const unsigned long long *__restrict__ rank, uint32_t nb, int ncoef,
                                  const T *__restrict__ reg, T *__restrict__ reg_c /*[ncoef][nb]*/, uint32_t *__restrict__ reg_list) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb || indicator[b]) return;
    uint32_t r = (uint32_t)rank[b];
    reg_list[r] = b;
    for (int e = 0; e < ncoef; e++) reg_c[(size_t)e * nb + r] = reg[(size_t)e * nb + b];
}

// =====================================================================================================================
// K5: coefficient DPCM chain (S1; ref: sz_float.c:7126-7152 no-mean "/precision", :6792-6812 mean "*recip", 2D :6035-6055;
// double always "*recip").  last[e] threads through every regression block in raster order: a nonlinear float
// recurrence.  The kernel is in szb_chain.cuh; here its result block and the reference's step.
// =====================================================================================================================
struct ChainOut { unsigned int unpred_count[4]; unsigned int err; unsigned int passes, kcyc_serial, kcyc_total, kcyc_helper, slow_chunks, n_full, n_pair, prof[8]; };   // from `passes` on: profiling aid

// one step of the reference's coefficient quantiser as a function of `last` (exact formula, ref: sz_float.c:7130-7150)
template <typename T>
__device__ __forceinline__ int chain_step_exact(T cur, T last, T prec, T rprec, int use_div, T &nl) {
    T diff = cur - last;
    T itv = use_div ? TT<T>::abs(diff) / prec : TT<T>::abs(diff) * rprec;
    itv = itv + (T)1;
    if (itv < (T)65536) {
        if (diff < (T)0) itv = -itv;
        int ct = (int)(itv / (T)2) + 32768;
        T dq = (T)(2 * (ct - 32768));
        T v = dq * prec;
        v = last + v;
        T chk = cur - v;
        if (TT<T>::abs(chk) > prec) { nl = cur; return 0; }
        nl = v;
        return ct;
    }
    nl = cur;
    return 0;
}

}  // namespace szb
