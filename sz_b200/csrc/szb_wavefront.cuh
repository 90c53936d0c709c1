// szb_wavefront.cuh -- K6 / D4: prediction + linear-scaling quantisation (compress) and reconstruction
// (decompress) on RECONSTRUCTED neighbours, scheduled as a wavefront over hyperplanes i+j+k = const.
//
// ref: sz/src/sz_float.c:7054-7373 (no-mean), :6721-7050 (mean), 2D :5990-6290; szd_float.c:3597-3748 (readable form)
//
// Mapping (DESIGN.md "wavefront"): one warp owns a PENCIL of PI x PJ rows (3D: 8 x 4 rows of axis-2 lines; 2D: 1 x 32)
// and sweeps it along the fast axis.  Lane (li,lj) handles point k = t - li - lj at step t, so every lane's seven
// Lorenzo neighbours were produced at steps <= t-1 and are exchanged through a small shared-memory ring (plus four
// carried registers).
//
// Neighbouring pencils exchange their boundary rows (li = PI-1 or lj = PJ-1) through a global scratch of TAGGED 64-bit
// words: {tag = launch epoch, 32 payload bits}, one word per fp32 value, two per fp64 value.  A 64-bit relaxed store / load
// at gpu scope is single-copy atomic, so a consumer that sees the current epoch in a word also sees its payload: the data
// carries its own "ready" bit.  There are no flags, no fences and no separate publish step -- the producer just stores, the
// consumer loads the CH points of the next chunk from its up / left / up-left neighbours (one L2 round trip) and retries
// until every tag matches.  Stale words of earlier launches carry older epochs and never match; the scratch is zeroed
// when it is (re)allocated and epochs start at 1.
// Pencils are handed out by a global ticket counter in anti-diagonal order, so a waiting warp only ever waits for a
// ticket that has already been taken: no deadlock regardless of how many CTAs are resident.  Every spin is bounded
// and raises an error flag instead of hanging the GPU.
//
// The run time of this kernel is its dependency chain, (hops x (CH + DMAX + hand-off latency) + n2) steps, times the
// latency of one step -- not bandwidth.  Hence everything that does NOT depend on reconstructed neighbours is taken out
// of the step: a fully parallel pre-pass (prepass_*_kernel below) resolves regression blocks, mean hits and (decompress)
// unpredictable values and code de-quantisation into a merged input array `xin` plus one "given" bit per point.  The
// wavefront step is then: one input (staged by bulk-asynchronous copies half a ring ahead), three ring loads, the Lorenzo sum, the
// branch-free quantiser (compress) or one add (decompress), one select on the given bit.
#pragma once
#include <cuda.h>          // CUtensorMap (the tensor map itself is encoded on the host, szb_shim.cu)
#include "szb_common.cuh"

namespace szb {

template <int DIM> struct Pencil;
template <> struct Pencil<3> { static constexpr int PI = 8, PJ = 4, CH = 16; };
template <> struct Pencil<2> { static constexpr int PI = 1, PJ = 32, CH = 32; };   // CH must be >= DMAX = PI + PJ - 2

template <int DIM> struct WfCfg {
    static constexpr int CH = Pencil<DIM>::CH;     // chunk length along the fast axis (publish / stage granularity)
    static constexpr int W = 2 * CH;               // reconstruction ring length (chunks c-1, c)
    static constexpr int S = W + 1;                // padded stride
    static constexpr int IW = 4 * CH;              // input ring length (chunks c-1 .. c+2: prefetch runs two chunks ahead)
    static constexpr int NSLOT = 2;                // the input ring is filled half by half (two chunks per copy), one mbarrier per half
    static constexpr int ROWS = (Pencil<DIM>::PI + 1) * (Pencil<DIM>::PJ + 1);   // incl. halo rows li = -1 / lj = -1
};

struct WfParams {
    Geom g;
    int cap, radius, cap_l;          // intvCapacity, intvRadius, intvCapacity-2
    int use_mean;
    uint32_t pencils_a, pencils_b;   // pencil grid
    uint32_t num_pencils;
    const uint32_t *order;           // ticket -> pencil id (anti-diagonal order)
    unsigned long long *bnd;         // tagged boundary words [num_pencils][NR][n2][sizeof(T)/4]
    uint32_t epoch;                  // tag of this launch (> 0)
    unsigned int *ticket;            // global ticket counter
    int *err;                        // error flag (1 = spin timeout)
    const uint32_t *given;           // [rows][given_wpr] bit k&31 of word k>>5: point resolved by the pre-pass
    uint32_t given_wpr;              // words per row
    uint32_t pitch;                  // elements per row of xin (n2 rounded up to a whole chunk: every chunk copy is in bounds and aligned)
    int write_all_recon;             // debug: store every reconstructed point in the raster array (compress)
    unsigned long long *prof;        // optional (only with -DSZB_WF_PROFILE): [0] wait+halo cycles [3] pencil total [4] steps; [8+4*pid..] timeline
};

// boundary rows a pencil exports: 3D bottom rows (li = PI-1; slot lj) + right rows (lj = PJ-1, li < PI-1; slot PJ+li); 2D the last row
template <int DIM> struct WfBnd { static constexpr int NR = (DIM == 3) ? Pencil<DIM>::PI + Pencil<DIM>::PJ - 1 : 1; };
template <typename T, int DIM>
inline size_t wf_bnd_bytes(uint32_t num_pencils, uint32_t n2) { return (size_t)num_pencils * WfBnd<DIM>::NR * n2 * (sizeof(T) / 4) * 8; }

// Per-warp shared memory: [reconstruction ring incl. halo rows | input ring | one mbarrier per half of the input ring].
// The input ring holds two HALVES of two chunks each; a half is the destination of ONE tensor-map tile load (TMA): the box
// {2 CH points, PJ rows, PI planes} of the merged input `xin` lands as [lane = li * PJ + lj][2 CH], rows beyond the array are
// zero-filled by the TMA unit.  Lane l reads its point k at half (k / 2CH) & 1, offset l * 2CH + k mod 2CH: the lanes of a warp
// are at different k (their delays differ), which spreads them over the banks without padding.  128-byte aligned (TMA destination).
template <typename T, int DIM> struct WfSmem {
    static constexpr int HALF = 2 * WfCfg<DIM>::CH;                                                   // points per lane and half
    static constexpr size_t RING = (((size_t)WfCfg<DIM>::ROWS * WfCfg<DIM>::S * sizeof(T)) + 127) & ~(size_t)127;
    static constexpr size_t HALF_BYTES = (size_t)32 * HALF * sizeof(T);
    static constexpr size_t INR = 2 * HALF_BYTES;
    static constexpr size_t BARS = 8 * WfCfg<DIM>::NSLOT;
    static constexpr size_t TOTAL = (RING + INR + BARS + 127) & ~(size_t)127;
};
template <typename T>
__host__ __device__ constexpr size_t wf_smem_per_warp(int dim) { return dim == 3 ? WfSmem<T, 3>::TOTAL : WfSmem<T, 2>::TOTAL; }

// ---- bulk-asynchronous copies (TMA, 1D) and their transaction barriers --------------------------------------------------
__device__ __forceinline__ void mbar_init(unsigned int bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned int bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned int bar, unsigned int parity) {
    unsigned int done;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    return done != 0;
}
// one contiguous run of `bytes` (multiple of 16, both addresses 16-byte aligned) global -> shared; completion is counted on `bar`
__device__ __forceinline__ void bulk_g2s(unsigned int smem_dst, const void *gmem_src, unsigned int bytes, unsigned int bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst), "l"(gmem_src), "r"(bytes), "r"(bar) : "memory");
}
// one tile of a tensor map (this GPU's memory, no multicast) global -> shared; the box's bytes are counted on `bar`
__device__ __forceinline__ void tma_tile_3d(unsigned int smem_dst, const CUtensorMap *tm, int c0, int c1, int c2, unsigned int bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_tile_2d(unsigned int smem_dst, const CUtensorMap *tm, int c0, int c1, unsigned int bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ int ld_relaxed(const int *p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// 64-bit relaxed accesses at gpu scope (single-copy atomic; served by L2)
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// tagged boundary words of one value
template <typename T> struct Tagged;
template <> struct Tagged<float> {
    static constexpr int WPE = 1;
    static __device__ __forceinline__ void store(unsigned long long *p, float v, unsigned long long tag_hi) { st_relaxed_u64(p, tag_hi | (unsigned long long)__float_as_uint(v)); }
    // words already loaded -> value; true when the tag matches
    static __device__ __forceinline__ bool decode(const unsigned long long *w, uint32_t epoch, float &v) {
        v = __uint_as_float((uint32_t)w[0]);
        return (uint32_t)(w[0] >> 32) == epoch;
    }
};
template <> struct Tagged<double> {
    static constexpr int WPE = 2;
    static __device__ __forceinline__ void store(unsigned long long *p, double v, unsigned long long tag_hi) {
        const unsigned long long b = (unsigned long long)__double_as_longlong(v);
        st_relaxed_u64(p, tag_hi | (b & 0xffffffffull));
        st_relaxed_u64(p + 1, tag_hi | (b >> 32));
    }
    static __device__ __forceinline__ bool decode(const unsigned long long *w, uint32_t epoch, double &v) {
        v = __longlong_as_double((long long)((w[0] & 0xffffffffull) | (w[1] << 32)));
        return (uint32_t)(w[0] >> 32) == epoch && (uint32_t)(w[1] >> 32) == epoch;
    }
};

// ---- pre-pass: everything that is independent of reconstructed neighbours ------------------------------------------
// One warp per row of the fast axis, 32 consecutive points per iteration (coalesced), one ballot per 32 points builds the
// "given" word.  ktab[k] = (block index along axis 2) << 8 | (offset inside that block), built on the host.
//
// compress (ref: regression blocks sz_float.c:7098-7152 / 2D :6036-6078; mean hit :6929-6933):
//   regression block : code = quantize(x, a*ii + b*jj + c*kk + d) -> codes[], xin = reconstructed value, given
//   Lorenzo block    : mean hit (use_mean) -> code = radius, xin = mean, given;  else xin = x, not given
template <typename T, int DIM, bool MEAN>
__global__ void __launch_bounds__(256) prepass_compress_kernel(Geom g, const T *__restrict__ x, const unsigned char *__restrict__ indicator,
                                                               const T *__restrict__ coeff_last, const uint32_t *__restrict__ ktab,
                                                               uint16_t *__restrict__ codes, T *__restrict__ xin, uint32_t pitch, uint32_t *__restrict__ given,
                                                               uint32_t wpr, T eb, T recip_eb, T mean, int cap, int radius) {
    const int lane = threadIdx.x & 31;
    const uint32_t n1 = g.ax[1].n, n2 = g.ax[2].n, nblk = g.num_blocks;
    const uint32_t rows = g.ax[0].n * n1;
    for (uint32_t row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += gridDim.x * (blockDim.x >> 5)) {
        const uint32_t i = row / n1, j = row - i * n1;
        uint32_t ii, jj;
        const uint32_t b0 = g.ax[0].block_of(i, ii), b1 = g.ax[1].block_of(j, jj);
        const uint32_t bid0 = g.block_id(b0, b1, 0);
        const T iif = (T)ii, jjf = (T)jj;
        const uint64_t base = (uint64_t)row * n2;
        for (uint32_t k0 = 0; k0 < n2; k0 += 32) {
            const uint32_t k = k0 + lane;
            bool giv = false;
            if (k < n2) {
                const uint32_t kt = ktab[k];
                const uint32_t bid = bid0 + (kt >> 8);
                const T cur = x[base + k];
                T v = cur;
                if (indicator[bid] == 0) {
                    T pred;
                    if (DIM == 3) {
                        T m0 = coeff_last[bid] * iif, m1 = coeff_last[nblk + bid] * jjf, m2 = coeff_last[2 * nblk + bid] * (T)(kt & 255u);
                        pred = m0 + m1; pred = pred + m2; pred = pred + coeff_last[3 * nblk + bid];
                    } else {
                        T m1 = coeff_last[bid] * jjf, m2 = coeff_last[nblk + bid] * (T)(kt & 255u);
                        pred = m1 + m2; pred = pred + coeff_last[2 * nblk + bid];
                    }
                    const int q = quantize<T>(cur, pred, recip_eb, eb, cap, radius, v);
                    codes[base + k] = (uint16_t)q;
                    giv = true;
                } else if (MEAN) {
                    const T dm = cur - mean;
                    if (TT<T>::abs(dm) <= eb) { codes[base + k] = (uint16_t)radius; v = mean; giv = true; }
                }
                xin[(uint64_t)row * pitch + k] = v;
            }
            const uint32_t w = __ballot_sync(0xffffffffu, giv);
            if (lane == 0) given[(uint64_t)row * wpr + (k0 >> 5)] = w;
        }
    }
}

// decompress (ref: szd_float.c:3597-3748): out[] holds the unpredictable values at code-0 points (prefilled)
//   regression block : code != 0 -> xin = pred + 2*(code-radius)*eb, given;  code == 0 -> xin = out (prefilled), given
//   Lorenzo block    : code == 0 -> xin = out, given;  mean code -> xin = mean, given;
//                      else xin = 2*(code'-radius)*eb (the term the wavefront adds to its prediction), not given
template <typename T, int DIM, bool MEAN>
__global__ void __launch_bounds__(256) prepass_decompress_kernel(Geom g, const uint16_t *__restrict__ codes_bm /*block-major, as decoded*/, const T *__restrict__ out,
                                                                 const unsigned char *__restrict__ indicator, const T *__restrict__ coeff_last,
                                                                 const uint32_t *__restrict__ ktab, T *__restrict__ xin, uint32_t pitch, uint32_t *__restrict__ given,
                                                                 uint32_t wpr, T eb, T mean, int radius) {
    const int lane = threadIdx.x & 31;
    const uint32_t n1 = g.ax[1].n, n2 = g.ax[2].n, nblk = g.num_blocks;
    const uint32_t rows = g.ax[0].n * n1;
    for (uint32_t row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += gridDim.x * (blockDim.x >> 5)) {
        const uint32_t i = row / n1, j = row - i * n1;
        uint32_t ii, jj;
        const uint32_t b0 = g.ax[0].block_of(i, ii), b1 = g.ax[1].block_of(j, jj);
        const uint32_t bid0 = g.block_id(b0, b1, 0);
        const T iif = (T)ii, jjf = (T)jj;
        const uint64_t base = (uint64_t)row * n2;
        // block-major position of point (row, k): offset of the block (ref: sz_float.c:7064 type_offset) + (ii*sy + jj)*sz + kk
        const uint32_t sx = g.ax[0].edge(b0), sy = g.ax[1].edge(b1);
        const uint64_t bm_row = (uint64_t)g.ax[0].off(b0) * g.s0 + (uint64_t)sx * g.ax[1].off(b1) * n2;
        const uint32_t sxy = sx * sy, rin = ii * sy + jj;
        for (uint32_t k0 = 0; k0 < n2; k0 += 32) {
            const uint32_t k = k0 + lane;
            bool giv = false;
            if (k < n2) {
                const uint32_t kt = ktab[k];
                const uint32_t bid = bid0 + (kt >> 8);
                const uint32_t bk = kt >> 8;
                const uint32_t bm = sxy * g.ax[2].off(bk) + rin * g.ax[2].edge(bk) + (kt & 255u);
                int tcode = (int)codes_bm[bm_row + bm];
                T v;
                if (tcode == 0) { v = out[base + k]; giv = true; }
                else if (indicator[bid] == 0) {
                    T pred;
                    if (DIM == 3) {
                        T m0 = coeff_last[bid] * iif, m1 = coeff_last[nblk + bid] * jjf, m2 = coeff_last[2 * nblk + bid] * (T)(kt & 255u);
                        pred = m0 + m1; pred = pred + m2; pred = pred + coeff_last[3 * nblk + bid];
                    } else {
                        T m1 = coeff_last[bid] * jjf, m2 = coeff_last[nblk + bid] * (T)(kt & 255u);
                        pred = m1 + m2; pred = pred + coeff_last[2 * nblk + bid];
                    }
                    T dq = (T)(2 * (tcode - radius)); T w = dq * eb; v = pred + w;
                    giv = true;
                } else if (MEAN && tcode == radius) { v = mean; giv = true; }
                else {
                    if (MEAN && tcode < radius) tcode += 1;
                    T dq = (T)(2 * (tcode - radius)); v = dq * eb;
                }
                xin[(uint64_t)row * pitch + k] = v;
            }
            const uint32_t w = __ballot_sync(0xffffffffu, giv);
            if (lane == 0) given[(uint64_t)row * wpr + (k0 >> 5)] = w;
        }
    }
}

// ---- the wavefront ----------------------------------------------------------------------------------------------
// Both rings are indexed by TIME, not by position: element k of a row whose lane runs `d` steps behind lane 0 lives in slot
// (k + d) mod ring length, i.e. in the slot of the step at which it is produced / consumed.  At step t every lane therefore
// writes slot t, reads its neighbours at slots t-1 / t-2 and its input at slot t: all warp-uniform.  With the step loop unrolled
// over one ring period (U = 2 chunks) every shared-memory offset and the chunk boundaries are compile-time constants and the
// step body is straight-line code: 4 LDS, the 6-add Lorenzo chain, the quantiser, 1 STS, 1 warp barrier.
template <typename T, int DIM, bool MEAN, bool DECOMP>
__global__ void __launch_bounds__(128) wavefront_kernel(WfParams P, const __grid_constant__ CUtensorMap xin_map /*the merged input of the pre-pass, `xin`,
                                                        as a tensor {pitch, n1, n0} (2D: {pitch, n1}) with box {2 CH, PJ, PI}*/,
                                                        uint16_t *codes /*raster; compress: codes of the not-given points are written here*/,
                                                        T *recon /*raster; decompress: output; compress: only with write_all_recon*/,
                                                        T eb, T recip_eb) {
    constexpr int PI = Pencil<DIM>::PI, PJ = Pencil<DIM>::PJ;
    constexpr int CH = WfCfg<DIM>::CH, W = WfCfg<DIM>::W, S = WfCfg<DIM>::S, IW = WfCfg<DIM>::IW, HALF = WfSmem<T, DIM>::HALF, NSLOT = WfCfg<DIM>::NSLOT;
    constexpr unsigned int HALF_BYTES = (unsigned int)WfSmem<T, DIM>::HALF_BYTES;
    static_assert(IW == 2 * HALF && (HALF & (HALF - 1)) == 0, "the input ring is two halves of two chunks");
    constexpr int U = W;                               // steps per unrolled period (two chunks)
    constexpr int RJ = PJ + 1;                         // ring rows per li (incl. halo lj = -1)
    constexpr int DMAX = PI - 1 + PJ - 1;
    constexpr int NR = WfBnd<DIM>::NR, WPE = Tagged<T>::WPE;
    static_assert(CH >= DMAX && DMAX >= 1 && DMAX < 32, "a chunk must cover the lane skew");
    static_assert(U % 32 == 0, "given-bit windows are whole words");
    static_assert(U - 6 >= DMAX, "the next half of the input ring is requested at step U-6, when every lane has left the previous one");
    constexpr int NW = U / 32;                         // given-bit words per period
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T *ring = reinterpret_cast<T *>(smem_raw + (size_t)warp * wf_smem_per_warp<T>(DIM));
    T *inr = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(ring) + WfSmem<T, DIM>::RING);       // input ring [2 halves][32 lanes][2 CH]
    const unsigned int inr_s32 = (unsigned int)__cvta_generic_to_shared(inr);
    const unsigned int bars_s32 = inr_s32 + (unsigned int)WfSmem<T, DIM>::INR;                          // NSLOT mbarriers
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < NSLOT; q++) mbar_init(bars_s32 + 8u * q, 1u);
        fence_mbar_init();
    }
    __syncwarp();
    unsigned int phases = 0;                           // bit q: parity the next wait on barrier q expects
    const Geom &g = P.g;
    const uint32_t n0 = g.ax[0].n, n1 = g.ax[1].n, n2 = g.ax[2].n;
    const uint32_t pitch = P.pitch;
    const int li = (DIM == 3) ? lane / PJ : 0, lj = (DIM == 3) ? lane % PJ : lane;
    const int delay = li + lj;
    // boundary slot this lane exports (-1: none)
    const int bslot = (DIM == 3) ? (li == PI - 1 ? lj : (lj == PJ - 1 ? PJ + li : -1)) : (lj == PJ - 1 ? 0 : -1);
    T *my_row = ring + ((size_t)(li + 1) * RJ + (lj + 1)) * S;
    const T *row_b = ring + ((size_t)(li + 1) * RJ + lj) * S;        // (li, lj-1)
    const T *row_c = ring + ((size_t)li * RJ + (lj + 1)) * S;        // (li-1, lj)
    const T *row_f = ring + ((size_t)li * RJ + lj) * S;              // (li-1, lj-1)
    const T *my_in = inr + (size_t)lane * HALF;                      // this lane's row of half 0; half 1 is 32 * HALF elements on
    const int radius = P.radius;
    const T caplf = (T)P.cap_l;
    const uint32_t wpr = P.given_wpr;
    const uint32_t epoch = P.epoch;
    const unsigned long long tag_hi = (unsigned long long)epoch << 32;
    unsigned long long *const bnd = P.bnd;
    const uint64_t prow = (uint64_t)n2 * WPE;          // words per boundary row

    for (;;) {
        unsigned int tk = 0;
        if (lane == 0) tk = atomicAdd(P.ticket, 1u);
        tk = __shfl_sync(0xffffffffu, tk, 0);
        if (tk >= P.num_pencils) return;
        const uint32_t pid = P.order[tk];
        const uint32_t pa = pid / P.pencils_b, pb = pid % P.pencils_b;
        const uint32_t i0 = pa * PI, j0 = pb * PJ;
        const uint32_t vi = min((uint32_t)PI, n0 - i0), vj = min((uint32_t)PJ, n1 - j0);     // rows of the pencil inside the domain
        const bool active = ((uint32_t)li < vi) && ((uint32_t)lj < vj);
        const bool store_rec = active && (DECOMP || P.write_all_recon);
        const bool exports = active && bslot >= 0;
        const uint64_t rowid = (uint64_t)(i0 + li) * n1 + (j0 + lj);
        const int64_t rowbase = (int64_t)(rowid * n2) - delay;      // + t = index of this lane's point at step t
        const int64_t bbase = (int64_t)(((uint64_t)pid * NR + (uint32_t)max(bslot, 0)) * n2) - delay;
        const uint32_t *grow = P.given + rowid * wpr;
        // zero the whole reconstruction ring (halo rows of domain-boundary pencils must read as +0.0)
        for (int q = lane; q < WfCfg<DIM>::ROWS * S; q += 32) ring[q] = (T)0;
        // "given" bits: gprev = word before the current period, gnext[] = words of the next period (requested one period ahead).
        // Inactive lanes see all-ones, so that they never store a code.
        uint32_t gprev = 0, gnext[NW];
#pragma unroll
        for (int h = 0; h < NW; h++) gnext[h] = active ? ((uint32_t)h < wpr ? grow[h] : 0u) : 0xffffffffu;
        T ra = (T)0, rd = (T)0, re = (T)0, rg = (T)0;    // carried neighbours: a=(.,.,k-1) own, d=(li,lj-1,k-1), e=(li-1,lj,k-1), g=(li-1,lj-1,k-1)
        bool ok = true;
        __syncwarp();

        // staging of the input: ONE tensor-map tile load per PAIR of chunks (pair h = chunks 2h, 2h+1 = one half of the input ring)
        // brings the 2 CH points of all PI x PJ rows of the pencil; lane 0 tells the half's barrier how many bytes to expect and issues
        // the request.  It is issued late in a period (step U-6), when the slowest lane has left the half that is overwritten.
        // Rows of xin are padded to whole chunks; what lies beyond the array reads as zero.
        auto stage_pair = [&](uint32_t h) {
            const uint32_t k0 = 2 * h * CH;
            if (k0 >= n2) return;
            const unsigned int half = h & 1u, bar = bars_s32 + 8u * half;
            if (lane == 0) {
                mbar_expect_tx(bar, HALF_BYTES);
                if (DIM == 3) tma_tile_3d(inr_s32 + half * HALF_BYTES, &xin_map, (int)k0, (int)j0, (int)i0, bar);
                else tma_tile_2d(inr_s32 + half * HALF_BYTES, &xin_map, (int)k0, (int)j0, bar);
            }
        };
        auto wait_pair = [&](uint32_t h) {
            const unsigned int half = h & 1u, bar = bars_s32 + 8u * half, parity = (phases >> half) & 1u;
            if (!mbar_try_wait(bar, parity)) {
                const long long deadline = clock64() + 2000000000LL;         // ~1 s: a copy that has not landed by then never will
                while (!mbar_try_wait(bar, parity)) {
                    if (clock64() > deadline) { if (lane == 0) atomicExch(P.err, 1); ok = false; break; }
                }
            }
            phases ^= 1u << half;
        };

        // chunk boundary: fetch the CH boundary values of chunk c from the up / up-left / left pencils (tagged words, L2) and
        // retry until every tag carries this launch's epoch; then put them into the halo rows of the ring and stage input
        // chunk c+2.  Halo row (li,lj) with li = -1 or lj = -1 has the virtual delay li + lj: element k -> slot (k + li + lj) mod W.
        const bool has_up = (DIM == 3) && pa > 0, has_left = pb > 0;
        const unsigned long long *b_up = bnd + (uint64_t)(pid - (has_up ? P.pencils_b : 0u)) * NR * prow;
        const unsigned long long *b_ul = bnd + ((uint64_t)(pid - (has_up && has_left ? P.pencils_b + 1u : 0u)) * NR + (PJ - 1)) * prow;
        const unsigned long long *b_left = bnd + (uint64_t)(pid - (has_left ? 1u : 0u)) * NR * prow;
        // tagged words of the halo of one chunk.  Every load is issued before any result is looked at: one L2 round trip for
        // the whole halo.  Rows outside the domain read as {tag ok, +0.0}.
        unsigned long long wu[(PJ + 1) * WPE], wl[PI * WPE];
        bool pre_valid = false;                              // wu / wl hold a speculative fetch of the next chunk
        auto load_halo = [&](uint32_t c) {
            const uint32_t k0 = c * CH;
            if (k0 + lane >= n2 || lane >= CH) return;
            const uint64_t koff = (uint64_t)(k0 + lane) * WPE;
            const unsigned long long fill = tag_hi;
            if (has_up) {
#pragma unroll
                for (int w = 0; w < WPE; w++) wu[w] = has_left ? ld_relaxed_u64(b_ul + koff + w) : fill;                        // row (i0-1, j0-1)
#pragma unroll
                for (int q = 0; q < PJ; q++) {                                                                                    // rows (i0-1, j0+q)
#pragma unroll
                    for (int w = 0; w < WPE; w++) wu[(q + 1) * WPE + w] = ((uint32_t)q < vj) ? ld_relaxed_u64(b_up + (uint64_t)q * prow + koff + w) : fill;
                }
            }
            if (has_left) {
#pragma unroll
                for (int q = 0; q < PI; q++) {                                                                                    // rows (i0+q, j0-1)
                    const int slot = (DIM == 3) ? (q == PI - 1 ? PJ - 1 : PJ + q) : 0;
#pragma unroll
                    for (int w = 0; w < WPE; w++) wl[q * WPE + w] = ((uint32_t)q < vi) ? ld_relaxed_u64(b_left + (uint64_t)slot * prow + koff + w) : fill;
                }
            }
        };
        // speculative fetch half a chunk ahead.  MEASURED SLOWER at 512^3 (the steps after the loads stretch from ~250 to ~800
        // cycles; wavefront 3.2 -> 4.0 ms), so it is compiled out; kept because the trade-off may differ for other pencil shapes.
        constexpr bool PREFETCH = false;
        auto prefetch_halo = [&](uint32_t c) {
            if (PREFETCH && (has_up || has_left) && c * CH < n2) { load_halo(c); pre_valid = true; }
        };
        auto boundary = [&](uint32_t c) {
#ifdef SZB_WF_PROFILE
            long long p0 = clock64();
#endif
            const uint32_t k0 = c * CH;
            const uint32_t kcnt = min((uint32_t)CH, n2 - k0);
            const bool mine = (uint32_t)lane < kcnt;
            if (c == 0 && (has_up || has_left)) {
                // a pencil may wait a long time for its first chunk: one lane polls one word per neighbour (the last point of the
                // chunk in the neighbour's slowest exporting row) before the whole warp starts loading
                if (lane == 0) {
                    const uint64_t klast = (uint64_t)(kcnt - 1) * WPE;
                    const unsigned long long *cu = has_up ? b_up + (uint64_t)(vj - 1) * prow + klast : nullptr;
                    const unsigned long long *cl = nullptr;
                    if (has_left) cl = b_left + (uint64_t)((DIM == 3) ? (vi == (uint32_t)PI ? PJ - 1 : PJ + vi - 1) : 0) * prow + klast;
                    unsigned int spins = 0;
                    for (;;) {
                        const bool gu = !cu || (uint32_t)(ld_relaxed_u64(cu) >> 32) == epoch;
                        const bool gl = !cl || (uint32_t)(ld_relaxed_u64(cl) >> 32) == epoch;
                        if (gu && gl) break;
                        __nanosleep(64);
                        if ((++spins & 0xff) == 0) {
                            if (ld_relaxed(P.err) != 0) { ok = false; break; }
                            if (spins > (1u << 22)) { atomicExch(P.err, 1); ok = false; break; }
                        }
                    }
                }
                ok = __shfl_sync(0xffffffffu, (int)ok, 0) != 0;
            }
            T hu[PJ + 1], hl[PI];
            unsigned int spins = 0;
            bool use_pre = pre_valid;
            pre_valid = false;
            while (ok) {
                bool good = true;
                if (!use_pre) load_halo(c);
                use_pre = false;
                if (mine) {
                    if (has_up) {
#pragma unroll
                        for (int q = 0; q <= PJ; q++) good &= Tagged<T>::decode(wu + q * WPE, epoch, hu[q]);
                    }
                    if (has_left) {
#pragma unroll
                        for (int q = 0; q < PI; q++) good &= Tagged<T>::decode(wl + q * WPE, epoch, hl[q]);
                    }
                }
                if (__all_sync(0xffffffffu, good)) break;
                __nanosleep(40);
                if ((++spins & 0xff) == 0) {
                    if (ld_relaxed(P.err) != 0) ok = false;
                    else if (spins > (1u << 22)) { atomicExch(P.err, 1); ok = false; }
                    ok = __shfl_sync(0xffffffffu, (int)ok, 0) != 0;
                }
            }
#ifdef SZB_WF_PROFILE
            if (lane == 0 && pid == P.num_pencils / 2 + P.pencils_b / 2 && c < 64) {        // per-chunk trace of one mid-field pencil
                unsigned long long *tr = P.prof + 8 + 4 * (size_t)P.num_pencils + 4 * (size_t)c;
                tr[0] = (unsigned long long)p0; tr[1] = (unsigned long long)clock64(); tr[2] = spins;
            }
            if (lane == 0) { atomicAdd(P.prof + 0, (unsigned long long)(clock64() - p0));
                             if (c == 0) { unsigned long long gt; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt)); P.prof[8 + 4 * (size_t)pid + 1] = gt; } }
#endif
            if (ok) {
                if (mine) {
                    if (has_up) {
#pragma unroll
                        for (int q = -1; q < PJ; q++) ring[((size_t)0 * RJ + (q + 1)) * S + ((k0 + lane + (q - 1)) & (W - 1))] = hu[q + 1];
                    }
                    if (has_left) {
#pragma unroll
                        for (int q = 0; q < PI; q++) ring[((size_t)(q + 1) * RJ + 0) * S + ((k0 + lane + (q - 1)) & (W - 1))] = hl[q];
                    }
                }
                if ((c & 1u) == 0) wait_pair(c >> 1);       // the half holding chunks c, c+1 has landed
            }
            __syncwarp();
#ifdef SZB_WF_PROFILE
            if (lane == 0 && pid == P.num_pencils / 2 + P.pencils_b / 2 && c < 64) P.prof[8 + 4 * (size_t)P.num_pencils + 4 * (size_t)c + 3] = (unsigned long long)clock64();
#endif
        };

#ifdef SZB_WF_PROFILE
        const long long pen0 = clock64();
        if (lane == 0) { unsigned long long gt0; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt0)); unsigned int smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
                         P.prof[8 + 4 * (size_t)pid + 0] = gt0; P.prof[8 + 4 * (size_t)pid + 3] = smid; }
#endif
        stage_pair(0);
        const uint32_t tsteps = n2 + DMAX;
        const uint32_t nper = (tsteps + U - 1) / U;

        // one period of U steps.  CHK = false: every lane's point exists (DMAX <= t, t < n2 for all steps of the period)
        auto period = [&](uint32_t m, auto chk_tag) {
            constexpr bool CHK = decltype(chk_tag)::value;
            const uint32_t t0 = m * U;
            // given-bit windows of this period: bit s%32 of win[s/32] <-> this lane's point at step t0 + s
            uint32_t win[NW];
#pragma unroll
            for (int h = 0; h < NW; h++) {
                const uint32_t cur = gnext[h];
                win[h] = __funnelshift_rc(gprev, cur, 32 - delay);
                gprev = cur;
            }
#pragma unroll
            for (int h = 0; h < NW; h++) {
                const uint32_t wi = (m + 1) * NW + h;
                gnext[h] = active ? (wi < wpr ? grow[wi] : 0u) : 0xffffffffu;
            }
            const uint32_t kin = t0 - (uint32_t)delay;                     // this lane's position at step t0 (wraps while t0 < delay: masked below)
            uint16_t *cp = codes + (rowbase + (int64_t)t0);
            T *rp = recon + (rowbase + (int64_t)t0);
            unsigned long long *bp = bnd + (bbase + (int64_t)t0) * WPE;
#pragma unroll
            for (int s = 0; s < U; s++) {
                if (s == 0 || s == CH) {
                    const uint32_t c = 2 * m + (s == CH ? 1u : 0u);
                    if (c * CH < n2 && ok) {
                        if (has_up || has_left) boundary(c);
                        else { if ((c & 1u) == 0) wait_pair(c >> 1); __syncwarp(); }
                    }
                }
                if (s == U - 6) stage_pair(m + 1);
                if (s == CH / 2) prefetch_halo(2 * m + 1);
                if (s == CH + CH / 2) prefetch_halo(2 * m + 2);
                const bool giv = (win[s >> 5] >> (s & 31)) & 1u;
                bool valid = true;
                if (CHK) valid = active && (uint32_t)(t0 + s - (uint32_t)delay) < n2;
                const T rb = row_b[(s + W - 1) & (W - 1)];
                const T rc = (DIM == 3) ? row_c[(s + W - 1) & (W - 1)] : (T)0;
                const T rf = (DIM == 3) ? row_f[(s + W - 2) & (W - 1)] : (T)0;
                const uint32_t ipos = (kin + s) & (IW - 1);
                const T in = my_in[(ipos & (HALF - 1)) + (ipos & HALF) * 32u];         // half (ipos / HALF) starts 32 * HALF elements on
                T pred;
                if (DIM == 3) {   // main-loop term order (ref: sz_float.c:7268-7269)
                    pred = ra + rb; pred = pred + rc; pred = pred - rd; pred = pred - re; pred = pred - rf; pred = pred + rg;
                } else {          // ref: sz_float.c:6204
                    pred = ra + rb; pred = pred - rd;
                }
                T rec;
                if (!DECOMP) {
                    // branch-free form of the quantiser; given / out-of-range lanes compute garbage that is discarded.
                    // 2*trunc(v/2) in floating point equals (T)(2*(int)(v/2)) for every in-range v (up to the sign of a zero, which
                    // no comparison or code can see) and keeps the two conversions off the dependency chain
                    const T cur = in;
                    const T diff = cur - pred;
                    T itv = TT<T>::abs(diff) * recip_eb; itv = itv + (T)1;
                    const bool inrange = itv < caplf;
                    const T sitv = diff < (T)0 ? -itv : itv;
                    const T half = TT<T>::trunc(sitv / (T)2);
                    const T dq = half + half;
                    T r0 = dq * eb; r0 = pred + r0;
                    const T chk = cur - r0;
                    const bool good = inrange && !(TT<T>::abs(chk) > eb);
                    rec = (good && !giv) ? r0 : cur;
                    int q0 = (int)half + radius;
                    if (MEAN) { if (q0 <= radius) q0 -= 1; }      // ref: sz_float.c:6944 (the shift precedes the verification)
                    if (!giv && valid) cp[s] = (uint16_t)(good ? q0 : 0);
                } else {
                    const T r0 = pred + in;                       // in = 2*(code-radius)*eb for the not-given points
                    rec = giv ? in : r0;
                }
                if (CHK) { if (!valid) rec = (T)0; }
                if (exports && valid) Tagged<T>::store(bp + s * WPE, rec, tag_hi);
                if (store_rec && valid) rp[s] = rec;
                my_row[s] = rec;
                ra = rec;
                if (CHK) { rd = valid ? rb : (T)0; re = valid ? rc : (T)0; rg = valid ? rf : (T)0; }
                else { rd = rb; re = rc; rg = rf; }
                __syncwarp();
            }
        };
        for (uint32_t m = 0; m < nper && ok; m++) {
            const bool full = (m > 0) && ((m + 1) * U <= n2);      // t - delay >= 0 and t < n2 for every step and lane
            if (full) period(m, std::false_type{}); else period(m, std::true_type{});
        }
#ifdef SZB_WF_PROFILE
        if (lane == 0) { atomicAdd(P.prof + 3, (unsigned long long)(clock64() - pen0)); atomicAdd(P.prof + 4, (unsigned long long)tsteps);
                         unsigned long long gt; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt)); P.prof[8 + 4 * (size_t)pid + 2] = gt; }
#endif
        if (!ok) return;
        __syncwarp();
    }
}

}  // namespace szb
