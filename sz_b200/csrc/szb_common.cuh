// szb_common.cuh -- geometry, IEEE/x86 conversion helpers and small device utilities shared by all kernels.
// Compiled with --fmad=false: every T operation below is a single IEEE-754 op, written one per statement in the
// order of the reference, so results match the x86 (SSE2, no FMA) reference bit for bit.
#pragma once
#include <type_traits>
#include <cuda_runtime.h>
#include <stdint.h>

namespace szb {

// ---- block geometry per axis (ref: sz/include/sz.h:93-123, SZ_COMPUTE_*_NUMBER_OF_BLOCKS / SZ_COMPUTE_BLOCKCOUNT) ----
struct Axis {
    uint32_t n, nb, early, late, split;
    __host__ __device__ __forceinline__ uint32_t edge(uint32_t b) const { return b < split ? early : late; }
    __host__ __device__ __forceinline__ uint32_t off(uint32_t b) const { return b < split ? b * early : b * late + split; }
    // inverse: coordinate -> (block, index inside block)
    __host__ __device__ __forceinline__ uint32_t block_of(uint32_t x, uint32_t &inner) const {
        uint32_t cut = split * early;
        if (x < cut) { uint32_t b = x / early; inner = x - b * early; return b; }
        uint32_t y = x - cut; uint32_t b = y / late; inner = y - b * late; return split + b;
    }
};

inline Axis make_axis(size_t n, size_t B) {
    Axis a; a.n = (uint32_t)n; a.nb = (n <= B) ? 1u : (uint32_t)(n / B);
    a.early = a.late = (uint32_t)(n / a.nb); a.split = (uint32_t)(n % a.nb);
    if (a.split) a.early++;
    return a;
}

// Unified 3-axis view: 3D uses (n0,n1,n2); 2D uses a dummy axis 0 of extent 1 and (n1,n2) = (r2,r1).
struct Geom {
    int dim;                 // 2 or 3
    int ncoef;               // 3 or 4
    Axis ax[3];
    uint64_t N, s0, s1;      // element strides of axis 0 and 1 (axis 2 is contiguous)
    uint32_t num_blocks;
    // block-major offset of block (b0,b1,b2) in the type[] array (ref: sz_float.c:7064 type_offset + running pointer)
    __host__ __device__ __forceinline__ uint64_t block_type_off(uint32_t b0, uint32_t b1, uint32_t b2) const {
        uint64_t sx = ax[0].edge(b0), sy = ax[1].edge(b1);
        return (uint64_t)ax[0].off(b0) * s0 + sx * ((uint64_t)ax[1].off(b1) * ax[2].n + sy * (uint64_t)ax[2].off(b2));
    }
    __host__ __device__ __forceinline__ uint32_t block_id(uint32_t b0, uint32_t b1, uint32_t b2) const {
        return (b0 * ax[1].nb + b1) * ax[2].nb + b2;
    }
};

inline Geom make_geom(int dim, size_t n0, size_t n1, size_t n2) {
    Geom g; g.dim = dim; g.ncoef = (dim == 3) ? 4 : 3;
    size_t B = (dim == 3) ? 6 : 16;      // ref: sz_float.c:6546, :5537
    if (dim == 2) n0 = 1;
    g.ax[0] = (dim == 3) ? make_axis(n0, B) : make_axis(1, 1);
    g.ax[1] = make_axis(n1, B);
    g.ax[2] = make_axis(n2, B);
    g.N = (uint64_t)n0 * n1 * n2; g.s0 = (uint64_t)n1 * n2; g.s1 = n2;
    g.num_blocks = g.ax[0].nb * g.ax[1].nb * g.ax[2].nb;
    return g;
}

// ---- x86-64 (gcc, SSE2) float->integer conversion semantics ------------------------------------
// cvttsd2si: truncation; out-of-range or NaN gives 0x8000000000000000
__device__ __forceinline__ long long x86_cvttsd2si(double d) {
    if (d >= -9223372036854775808.0 && d < 9223372036854775808.0) return __double2ll_rz(d);
    return (long long)0x8000000000000000ULL;
}
// (size_t)d as gcc emits it: below 2^63 a plain cvttsd2si, else cvttsd2si(d - 2^63) ^ 2^63
__device__ __forceinline__ unsigned long long x86_d2u64(double d) {
    if (d < 9223372036854775808.0) return (unsigned long long)x86_cvttsd2si(d);
    return (unsigned long long)x86_cvttsd2si(d - 9223372036854775808.0) ^ 0x8000000000000000ULL;
}

// ---- type traits -------------------------------------------------------------------------------------------------
template <typename T> struct TT;
template <> struct TT<float> {
    typedef uint32_t bits_t;
    static __device__ __forceinline__ float abs(float v) { return fabsf(v); }
    static __device__ __forceinline__ float trunc(float v) { return truncf(v); }
};
template <> struct TT<double> {
    typedef uint64_t bits_t;
    static __device__ __forceinline__ double abs(double v) { return fabs(v); }
    static __device__ __forceinline__ double trunc(double v) { return ::trunc(v); }
};

// ---- linear-scaling quantiser (ref: sz_float.c:7160-7180, :7272-7290) ---------------------------------
// returns the code (0 = unpredictable) and the reconstructed value
template <typename T>
__device__ __forceinline__ int quantize(T x, T pred, T recip, T step, int cap, int radius, T &rec) {
    T diff = x - pred;
    T itv = TT<T>::abs(diff) * recip;
    itv = itv + (T)1;
    if (itv < (T)cap) {
        if (diff < (T)0) itv = -itv;
        int q = (int)(itv / (T)2) + radius;
        T dq = (T)(2 * (q - radius));
        T r = dq * step;
        r = pred + r;
        T chk = x - r;
        if (TT<T>::abs(chk) > step) { rec = x; return 0; }
        rec = r;
        return q;
    }
    rec = x;
    return 0;
}

__device__ __forceinline__ uint32_t bswap32(uint32_t v) { return __byte_perm(v, 0, 0x0123); }
__device__ __forceinline__ uint64_t bswap64(uint64_t v) {
    return ((uint64_t)bswap32((uint32_t)v) << 32) | bswap32((uint32_t)(v >> 32));
}

__device__ __forceinline__ int ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

}  // namespace szb
