// szb_wavefront.cuh -- K6 / D4: prediction + linear-scaling quantisation (compress) and reconstruction
// (decompress) on RECONSTRUCTED neighbours, scheduled as a wavefront over hyperplanes i+j+k = const.
//
// ref: sz/src/sz_float.c:7054-7373 (no-mean), :6721-7050 (mean), 2D :5990-6290; szd_float.c:3597-3748 (readable form)
//
// Mapping (DESIGN.md "wavefront"): one warp owns a PENCIL of PI x PJ rows (3D: 8 x 4 rows of axis-2 lines; 2D: 1 x 32)
// and sweeps it along the fast axis.  Lane (li,lj) handles point k = t - li - lj at step t, so every lane's seven
// Lorenzo neighbours were produced at steps <= t-1 and are exchanged through a small shared-memory ring (plus four
// carried registers).  Neighbouring pencils exchange their boundary rows through the global reconstruction array,
// CH points at a time, guarded by per-pencil progress flags (release/acquire at gpu scope).  Pencils are handed out
// by a global ticket counter in anti-diagonal order, so a waiting warp only ever waits for a ticket that has already
// been taken: no deadlock regardless of how many CTAs are resident.  Every spin is bounded and raises an error flag
// instead of hanging the GPU.
//
// The run time of this kernel is its dependency chain, (hops x (CH + DMAX + flag latency) + n2) steps, times the latency
// of one step -- not bandwidth.  Hence: input rows arrive by LDGSTS two chunks ahead, block modes are fetched one block
// ahead, everything loop-invariant lives in registers and the Lorenzo quantiser is branch-free.
#pragma once
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
    static constexpr int IS = IW + 1;
    static constexpr int ROWS = (Pencil<DIM>::PI + 1) * (Pencil<DIM>::PJ + 1);   // incl. halo rows li = -1 / lj = -1
};

struct WfParams {
    Geom g;
    int cap, radius, cap_l;          // intvCapacity, intvRadius, intvCapacity-2
    int use_mean;
    uint32_t pencils_a, pencils_b;   // pencil grid
    uint32_t num_pencils;
    const uint32_t *order;           // ticket -> pencil id (anti-diagonal order)
    int *flags;                      // [num_pencils] chunks completed
    unsigned int *ticket;            // global ticket counter
    int *err;                        // error flag (1 = spin timeout)
    const unsigned char *indicator;  // [num_blocks], 1 = Lorenzo
    int write_all_recon;             // debug: store every reconstructed point (compress)
    unsigned long long *prof;        // optional cycle counters (only with -DSZB_WF_PROFILE): [0] wait [1] halo [2] publish [3] pencil total [4] steps [5] boundary other
};

template <typename T>
__host__ __device__ constexpr size_t wf_smem_per_warp(int dim) {
    // reconstruction ring incl. halo rows + input ring (x, or codes widened to T-sized slots)
    return dim == 3 ? ((size_t)WfCfg<3>::ROWS * WfCfg<3>::S + (size_t)32 * WfCfg<3>::IS) * sizeof(T)
                    : ((size_t)WfCfg<2>::ROWS * WfCfg<2>::S + (size_t)32 * WfCfg<2>::IS) * sizeof(T);
}

// asynchronous global -> shared copy of one element (LDGSTS): the data never passes through registers, so the warp does not
// stall on the load; completion is observed with cp.async.wait_group
template <typename T>
__device__ __forceinline__ void cp_async_elem(T *smem_dst, const T *gmem_src) {
    const unsigned int d = (unsigned int)__cvta_generic_to_shared(smem_dst);
    if (sizeof(T) == 4) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gmem_src) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ int ld_relaxed(const int *p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// wait until *flag >= need: polled by one lane with RELAXED loads (an acquire load would invalidate the SM's L1 on every
// poll); the caller issues the acquire fence once after the wait.  false on timeout / error.
__device__ __forceinline__ bool wf_wait(const int *flag, int need, int *err) {
    unsigned int spins = 0;
    while (ld_relaxed(flag) < need) {
        __nanosleep(100);
        if ((++spins & 0xff) == 0) {
            if (ld_relaxed(err) != 0) return false;
            if (spins > (1u << 23)) { atomicExch(err, 1); return false; }
        }
    }
    return true;
}

template <typename T, int DIM, bool MEAN, bool DECOMP>
__global__ void __launch_bounds__(128) wavefront_kernel(WfParams P, const T *__restrict__ x /*compress: original field*/,
                                                        uint16_t *codes /*raster; compress: out, decompress: in*/,
                                                        T *recon /*raster; compress: boundary-row scratch, decompress: output (unpredictables prefilled)*/,
                                                        const T *__restrict__ coeff_last /*[ncoef][num_blocks]*/, T eb, T recip_eb, T mean) {
    constexpr int PI = Pencil<DIM>::PI, PJ = Pencil<DIM>::PJ;
    constexpr int CH = WfCfg<DIM>::CH, W = WfCfg<DIM>::W, S = WfCfg<DIM>::S, IW = WfCfg<DIM>::IW, IS = WfCfg<DIM>::IS;
    constexpr int RJ = PJ + 1;                         // ring rows per li (incl. halo lj = -1)
    constexpr int DMAX = PI - 1 + PJ - 1;
    static_assert(CH >= DMAX, "a chunk must cover the lane skew");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T *ring = reinterpret_cast<T *>(smem_raw + (size_t)warp * wf_smem_per_warp<T>(DIM));
    T *inr = ring + (size_t)WfCfg<DIM>::ROWS * S;       // input ring [32][IS]
    const Geom &g = P.g;
    const uint32_t n0 = g.ax[0].n, n1 = g.ax[1].n, n2 = g.ax[2].n;
    const int li = (DIM == 3) ? lane / PJ : 0, lj = (DIM == 3) ? lane % PJ : lane;
    const int delay = li + lj;
    T *my_row = ring + ((size_t)(li + 1) * RJ + (lj + 1)) * S;
    const T *row_b = ring + ((size_t)(li + 1) * RJ + lj) * S;        // (li, lj-1)
    const T *row_c = ring + ((size_t)li * RJ + (lj + 1)) * S;        // (li-1, lj)
    const T *row_f = ring + ((size_t)li * RJ + lj) * S;              // (li-1, lj-1)
    const T *my_in = inr + (size_t)lane * IS;
    const uint32_t nchunks = (n2 + CH - 1) / CH;
    // loop-invariant scalars in registers
    const int radius = P.radius, cap = P.cap;
    const T caplf = (T)P.cap_l;
    const uint32_t nb2 = g.ax[2].nb, nblk = g.num_blocks;
    const uint32_t ax2_split = g.ax[2].split, ax2_early = g.ax[2].early, ax2_late = g.ax[2].late;
    const unsigned char *__restrict__ indicator = P.indicator;
    const bool boundary_lane = (li == PI - 1) || (lj == PJ - 1) || P.write_all_recon;

    for (;;) {
        unsigned int tk = 0;
        if (lane == 0) tk = atomicAdd(P.ticket, 1u);
        tk = __shfl_sync(0xffffffffu, tk, 0);
        if (tk >= P.num_pencils) return;
        const uint32_t pid = P.order[tk];
        const uint32_t pa = pid / P.pencils_b, pb = pid % P.pencils_b;
        const uint32_t i0 = pa * PI, j0 = pb * PJ;
        const uint32_t i = i0 + li, j = j0 + lj;
        const bool active = (i < n0) && (j < n1);
        const uint64_t rowbase = ((uint64_t)i * n1 + j) * n2;
        uint16_t *crow = codes + rowbase;
        T *rrow = recon + rowbase;
        // zero the whole reconstruction ring (halo rows of domain-boundary pencils must read as +0.0)
        for (int q = lane; q < WfCfg<DIM>::ROWS * S; q += 32) ring[q] = (T)0;
        // per-lane block bookkeeping along axes 0 and 1 (fixed for the whole pencil)
        uint32_t ii = 0, jj = 0, b0 = 0, b1 = 0;
        if (active) { b0 = g.ax[0].block_of(i, ii); b1 = g.ax[1].block_of(j, jj); }
        const uint32_t bid0 = g.block_id(b0, b1, 0);     // block ids along axis 2 are consecutive
        const T iif = (T)ii, jjf = (T)jj;
        uint32_t bk = 0, kk = 0, rem = 0;               // rem = points left in the current block
        bool lorenzo = true;
        unsigned char next_ind = 1;                     // raw indicator byte of the next block (compared only when consumed, so the load is not waited for early)
        T nc0 = (T)0, nc1 = (T)0, nc2 = (T)0, nc3 = (T)0;    // quantised coefficients of the next block (prefetched)
        if (active) {
            next_ind = indicator[bid0];
            nc0 = coeff_last[bid0]; nc1 = coeff_last[nblk + bid0]; nc2 = coeff_last[2 * nblk + bid0];
            if (DIM == 3) nc3 = coeff_last[3 * nblk + bid0];
        }
        T c_base = (T)0, c_k = (T)0, c_d = (T)0;        // regression: pred = (c_base + c_k*kk) + c_d
        T ra = (T)0, rd = (T)0, re = (T)0, rg = (T)0;    // carried neighbours: a=(.,.,k-1) own, d=(li,lj-1,k-1), e=(li-1,lj,k-1), g=(li-1,lj-1,k-1)
        bool ok = true;
        __syncwarp();

        // decompress: the u16 codes of a chunk are loaded into registers one chunk before they are needed and widened into
        // the input ring at the next chunk boundary, so the load latency is hidden behind a whole chunk of work
        uint16_t cv[32];
        auto load_codes = [&](uint32_t c) {
            const uint32_t k0 = c * CH;
            const uint32_t kcnt = k0 < n2 ? min((uint32_t)CH, n2 - k0) : 0u;
#pragma unroll
            for (int r = 0; r < 32; r++) {
                const uint32_t ri = i0 + ((DIM == 3) ? r / PJ : 0), rj = j0 + ((DIM == 3) ? r % PJ : r);
                cv[r] = ((uint32_t)lane < kcnt && ri < n0 && rj < n1) ? __ldg(codes + ((uint64_t)ri * n1 + rj) * n2 + k0 + lane) : (uint16_t)0;
            }
        };
        auto store_codes = [&](uint32_t c) {
            const uint32_t slot0 = (c & 3) * CH;
            if (lane < CH) {
#pragma unroll
                for (int r = 0; r < 32; r++) inr[(size_t)r * IS + slot0 + lane] = (T)cv[r];      // codes < 65536: exact in T
            }
        };
        // compress: cooperative LDGSTS staging of the input rows of chunk c (row r of the pencil <-> row r of the ring)
        auto stage_input = [&](uint32_t c) {
            const uint32_t k0 = c * CH;
            if (k0 < n2) {
                const uint32_t kcnt = min((uint32_t)CH, n2 - k0);
                const uint32_t slot0 = (c & 3) * CH;
                if ((uint32_t)lane < kcnt) {
#pragma unroll 8
                    for (int r = 0; r < 32; r++) {
                        const uint32_t ri = i0 + ((DIM == 3) ? r / PJ : 0), rj = j0 + ((DIM == 3) ? r % PJ : r);
                        if (ri < n0 && rj < n1) cp_async_elem<T>(inr + (size_t)r * IS + slot0 + lane, x + ((uint64_t)ri * n1 + rj) * n2 + k0 + lane);
                    }
                }
            }
            cp_async_commit();                      // one group per chunk, even when empty, keeps the wait count uniform
        };
        // halo rows of chunk c from the neighbouring pencils' published reconstruction (L2 loads, never L1): all loads are
        // issued before the first shared-memory store so that they overlap
        auto stage_halo = [&](uint32_t c) {
            const uint32_t k0 = c * CH;
            const uint32_t kcnt = min((uint32_t)CH, n2 - k0);
            const uint32_t slot0 = (c & 1) * CH;
            if ((uint32_t)lane >= kcnt) return;
            T hu[PJ + 1], hl[PI];
            if (DIM == 3 && pa > 0) {
#pragma unroll
                for (int q = -1; q < PJ; q++) {               // rows (i0-1, j0+q)
                    const int64_t rj = (int64_t)j0 + q;
                    hu[q + 1] = (rj >= 0 && rj < (int64_t)n1) ? __ldcg(recon + ((uint64_t)(i0 - 1) * n1 + (uint64_t)rj) * n2 + k0 + lane) : (T)0;
                }
            }
            if (pb > 0) {
#pragma unroll
                for (int q = 0; q < PI; q++) {                // rows (i0+q, j0-1)
                    const uint32_t ri = i0 + q;
                    hl[q] = (ri < n0) ? __ldcg(recon + ((uint64_t)ri * n1 + (j0 - 1)) * n2 + k0 + lane) : (T)0;
                }
            }
            if (DIM == 3 && pa > 0) {
#pragma unroll
                for (int q = -1; q < PJ; q++) ring[((size_t)0 * RJ + (q + 1)) * S + slot0 + lane] = hu[q + 1];
            }
            if (pb > 0) {
#pragma unroll
                for (int q = 0; q < PI; q++) ring[((size_t)(q + 1) * RJ + 0) * S + slot0 + lane] = hl[q];
            }
        };

        const int *flag_up = (DIM == 3 && pa > 0) ? P.flags + (pid - P.pencils_b) : nullptr;
        const int *flag_left = (pb > 0) ? P.flags + (pid - 1) : nullptr;
        const uint32_t tsteps = n2 + DMAX;
#ifdef SZB_WF_PROFILE
        const long long pen0 = clock64();
#endif
        if (!DECOMP) { stage_input(0); stage_input(1); }
        else { load_codes(0); store_codes(0); load_codes(1); }
        uint32_t next_boundary = 0, next_publish = CH - 1 + DMAX, cidx = 0;
        for (uint32_t t = 0; t < tsteps; t++) {
            if (t == next_boundary && t < n2) {
                const uint32_t c = cidx;
#ifdef SZB_WF_PROFILE
                long long p0 = clock64();
#endif
                if (lane == 0) {
                    if (flag_up) ok = ok && wf_wait(flag_up, (int)c + 1, P.err);
                    if (flag_left) ok = ok && wf_wait(flag_left, (int)c + 1, P.err);
                    // acquire side of the hand-off: ONE acquire load per flag after the relaxed polling (no full fence)
                    if (flag_up) (void)ld_acquire(flag_up);
                    if (flag_left) (void)ld_acquire(flag_left);
                }
                ok = __shfl_sync(0xffffffffu, (int)ok, 0) != 0;
                if (!ok) break;
#ifdef SZB_WF_PROFILE
                long long p1 = clock64();
#endif
                stage_halo(c);
#ifdef SZB_WF_PROFILE
                __syncwarp();
                long long p2 = clock64();
#endif
                if (!DECOMP) { stage_input(c + 2); cp_async_wait<2>(); }   // chunk c has landed (groups c+1, c+2 may still be in flight)
                else { if (c > 0) store_codes(c); load_codes(c + 1); }
                __syncwarp();
#ifdef SZB_WF_PROFILE
                if (lane == 0) { long long p3 = clock64(); atomicAdd(P.prof + 0, (unsigned long long)(p1 - p0)); atomicAdd(P.prof + 1, (unsigned long long)(p2 - p1)); atomicAdd(P.prof + 5, (unsigned long long)(p3 - p2)); }
#endif
                next_boundary += CH; cidx++;
            }
            const uint32_t k = t - (uint32_t)delay;           // wraps to a huge value while t < delay
            if (active && k < n2) {
                const uint32_t slot = k & (W - 1), islot = k & (IW - 1);
                if (rem == 0) {                               // enter the next block along the fast axis
                    // mode and quantised coefficients of THIS block were requested one block ago (no stall on the loads here);
                    // request the next block's now
                    lorenzo = next_ind != 0;
                    if (!lorenzo) {
                        if (DIM == 3) { T m0 = nc0 * iif, m1 = nc1 * jjf; c_base = m0 + m1; c_k = nc2; c_d = nc3; }
                        else { c_base = nc0 * jjf; c_k = nc1; c_d = nc2; }
                    }
                    rem = (bk < ax2_split) ? ax2_early : ax2_late;
                    kk = 0;
                    bk++;
                    if (bk < nb2) {
                        const uint32_t bid = bid0 + bk;
                        next_ind = indicator[bid];
                        nc0 = coeff_last[bid]; nc1 = coeff_last[nblk + bid]; nc2 = coeff_last[2 * nblk + bid];
                        if (DIM == 3) nc3 = coeff_last[3 * nblk + bid];
                    }
                }
                const T rb = row_b[slot], rc = (DIM == 3) ? row_c[slot] : (T)0, rf = (DIM == 3) ? row_f[slot] : (T)0;
                const T in = my_in[islot];
                T rec;
                if (!DECOMP) {
                    const T cur = in;
                    int q;
                    if (!lorenzo) {
                        T pred = c_k * (T)kk; pred = c_base + pred; pred = pred + c_d;
                        q = quantize<T>(cur, pred, recip_eb, eb, cap, radius, rec);
                    } else {
                        T pred;
                        if (DIM == 3) {   // main-loop term order (ref: sz_float.c:7268-7269)
                            pred = ra + rb; pred = pred + rc; pred = pred - rd; pred = pred - re; pred = pred - rf; pred = pred + rg;
                        } else {          // ref: sz_float.c:6204
                            pred = ra + rb; pred = pred - rd;
                        }
                        // branch-free form of the quantiser; out-of-range lanes compute garbage that is discarded
                        const T diff = cur - pred;
                        T itv = TT<T>::abs(diff) * recip_eb; itv = itv + (T)1;
                        const bool inrange = itv < caplf;
                        const T sitv = diff < (T)0 ? -itv : itv;
                        int q0 = (int)(sitv / (T)2) + radius;
                        const T dq = (T)(2 * (q0 - radius));
                        T r0 = dq * eb; r0 = pred + r0;
                        if (MEAN) { if (q0 <= radius) q0 -= 1; }      // ref: sz_float.c:6944 (the shift precedes the verification)
                        const T chk = cur - r0;
                        const bool good = inrange && !(TT<T>::abs(chk) > eb);
                        q = good ? q0 : 0;
                        rec = good ? r0 : cur;
                        if (MEAN) { const T dm = cur - mean; if (TT<T>::abs(dm) <= eb) { q = radius; rec = mean; } }   // ref: :6929-6933
                    }
                    crow[k] = (uint16_t)q;
                    if (boundary_lane) rrow[k] = rec;
                } else {
                    int tcode = (int)in;
                    if (!lorenzo) {
                        if (tcode != 0) {
                            T pred = c_k * (T)kk; pred = c_base + pred; pred = pred + c_d;
                            T dq = (T)(2 * (tcode - radius)); T w = dq * eb; rec = pred + w;
                        } else rec = rrow[k];                             // unpredictable value, prefilled
                    } else {
                        if (MEAN && tcode == radius) rec = mean;
                        else if (tcode == 0) rec = rrow[k];
                        else {
                            if (MEAN && tcode < radius) tcode += 1;
                            T pred;
                            if (DIM == 3) { pred = ra + rb; pred = pred + rc; pred = pred - rd; pred = pred - re; pred = pred - rf; pred = pred + rg; }
                            else { pred = ra + rb; pred = pred - rd; }
                            T dq = (T)(2 * (tcode - radius)); T w = dq * eb; rec = pred + w;
                        }
                    }
                    rrow[k] = rec;
                }
                my_row[slot] = rec;
                ra = rec; rd = rb; re = rc; rg = rf;
                kk++; rem--;
            }
            __syncwarp();
            // publish chunk c once the slowest lane has finished it
            if (t == next_publish) {
#ifdef SZB_WF_PROFILE
                long long q0 = clock64();
#endif
                // release side: the warp barrier orders every lane's stores before lane 0's release store (cumulativity), so one
                // release by one lane publishes the whole chunk -- no per-lane full fence
                __syncwarp();
                if (lane == 0) st_release(P.flags + pid, (int)((t - DMAX) / CH) + 1);
                next_publish += CH;
#ifdef SZB_WF_PROFILE
                if (lane == 0) atomicAdd(P.prof + 2, (unsigned long long)(clock64() - q0));
#endif
            }
        }
        if (!DECOMP) cp_async_wait<0>();
#ifdef SZB_WF_PROFILE
        if (lane == 0) { atomicAdd(P.prof + 3, (unsigned long long)(clock64() - pen0)); atomicAdd(P.prof + 4, (unsigned long long)tsteps); }
#endif
        if (!ok) return;
        __syncwarp();
        if (lane == 0) st_release(P.flags + pid, (int)nchunks);
        __syncwarp();
    }
}

}  // namespace szb
