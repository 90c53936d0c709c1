// szb_wavefront.cuh -- K6 / D4: prediction + linear-scaling quantisation (compress) and reconstruction
// (decompress) on RECONSTRUCTED neighbours, scheduled as a wavefront over hyperplanes i+j+k = const.
//
// ref: sz/src/sz_float.c:7054-7373 (no-mean), :6721-7050 (mean), 2D :5990-6290; szd_float.c:3597-3748 (readable form)
//
// Mapping (DESIGN.md "wavefront"): one warp owns a PENCIL of PI x PJ rows (3D: 8 x 4 rows of axis-2 lines; 2D: 1 x 32)
// and sweeps it along the fast axis.  Lane (li,lj) handles point k = t - li - lj at step t, so every lane's seven
// Lorenzo neighbours were produced at steps <= t-1 and are exchanged through a small shared-memory ring (plus three
// carried registers).  Neighbouring pencils exchange their boundary rows through the global reconstruction array,
// CH points at a time, guarded by per-pencil progress flags (release/acquire at gpu scope).  Pencils are handed out
// by a global ticket counter in anti-diagonal order, so a waiting warp only ever waits for a ticket that has already
// been taken: no deadlock regardless of how many CTAs are resident.  Every spin is bounded and raises an error flag
// instead of hanging the GPU.
#pragma once
#include "szb_common.cuh"

namespace szb {

constexpr int WF_CH = 32;            // chunk length along the fast axis
constexpr int WF_W = 2 * WF_CH;      // ring length
constexpr int WF_S = WF_W + 1;       // padded ring stride (elements)
constexpr int WF_IW = 4 * WF_CH;     // input ring length: chunks c-1 .. c+2 (prefetch runs two chunks ahead)
constexpr int WF_IS = WF_IW + 1;     // padded input ring stride

template <int DIM> struct Pencil;
template <> struct Pencil<3> { static constexpr int PI = 8, PJ = 4; };
template <> struct Pencil<2> { static constexpr int PI = 1, PJ = 32; };

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
};

template <typename T>
__host__ __device__ constexpr size_t wf_smem_per_warp(int dim) {
    return (size_t)(dim == 3 ? (9 * 5) : (2 * 33)) * WF_S * sizeof(T)   // reconstruction ring incl. halo rows
           + (size_t)32 * WF_IS * sizeof(T);                              // input ring (x, or codes widened to T-sized slots)
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

// wait until flags[p] >= need; returns false on timeout / error
__device__ __forceinline__ bool wf_wait(const int *flag, int need, int *err) {
    unsigned int spins = 0;
    while (ld_acquire(flag) < need) {
        __nanosleep(64);
        if ((++spins & 0x3ff) == 0) {
            if (ld_acquire(err) != 0) return false;
            if (spins > (1u << 24)) { atomicExch(err, 1); return false; }
        }
    }
    return true;
}

template <typename T, int DIM, bool MEAN, bool DECOMP>
__global__ void __launch_bounds__(256) wavefront_kernel(WfParams P, const T *__restrict__ x /*compress: original field*/,
                                                        uint16_t *codes /*raster; compress: out, decompress: in*/,
                                                        T *recon /*raster; compress: boundary-row scratch, decompress: output (unpredictables prefilled)*/,
                                                        const T *__restrict__ coeff_last /*[ncoef][num_blocks]*/, T eb, T recip_eb, T mean) {
    constexpr int PI = Pencil<DIM>::PI, PJ = Pencil<DIM>::PJ;
    constexpr int RJ = PJ + 1;                         // ring rows per li (incl. halo lj = -1)
    constexpr int DMAX = PI - 1 + PJ - 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    T *ring = reinterpret_cast<T *>(smem_raw + (size_t)warp * wf_smem_per_warp<T>(DIM));
    T *inr = ring + (size_t)(PI + 1) * RJ * WF_S;      // input ring [32][WF_S]
    const Geom &g = P.g;
    const uint32_t n0 = g.ax[0].n, n1 = g.ax[1].n, n2 = g.ax[2].n;
    const int li = (DIM == 3) ? lane / PJ : 0, lj = (DIM == 3) ? lane % PJ : lane;
    const int delay = li + lj;
    T *my_row = ring + ((size_t)(li + 1) * RJ + (lj + 1)) * WF_S;
    const T *row_b = ring + ((size_t)(li + 1) * RJ + lj) * WF_S;        // (li, lj-1)
    const T *row_c = ring + ((size_t)li * RJ + (lj + 1)) * WF_S;        // (li-1, lj)
    const T *row_f = ring + ((size_t)li * RJ + lj) * WF_S;              // (li-1, lj-1)
    const T *my_in = inr + (size_t)lane * WF_IS;
    const uint32_t nchunks = (n2 + WF_CH - 1) / WF_CH;

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
        // zero the whole reconstruction ring (halo rows of domain-boundary pencils must read as +0.0)
        for (int q = lane; q < (PI + 1) * RJ * WF_S; q += 32) ring[q] = (T)0;
        // per-lane block bookkeeping along axes 0 and 1 (fixed for the whole pencil)
        uint32_t ii = 0, jj = 0, b0 = 0, b1 = 0, sy = 1;
        if (active) {
            b0 = g.ax[0].block_of(i, ii);
            b1 = g.ax[1].block_of(j, jj);
            sy = g.ax[1].edge(b1);
        }
        (void)sy;
        uint32_t bk = 0, kk = 0, sz = g.ax[2].edge(0);
        bool lorenzo = true;
        T c_base = (T)0, c_k = (T)0, c_d = (T)0;        // regression: pred = (c_base + c_k*kk) + c_d
        T ra = (T)0, rd = (T)0, re = (T)0, rg = (T)0;    // carried neighbours: a=(.,.,k-1) own, d=(li,lj-1,k-1), e=(li-1,lj,k-1), g=(li-1,lj-1,k-1)
        bool ok = true;
        __syncwarp();

        // cooperative staging of the input rows of chunk c into the input ring (row r of the pencil <-> lane r).
        // compress: LDGSTS (asynchronous); decompress: the u16 codes are loaded in one batch and widened to T.
        // decompress: the u16 codes of a chunk are loaded into registers one chunk before they are needed and widened into
        // the input ring at the next chunk boundary, so the load latency is hidden behind a whole chunk of work
        uint16_t cv[32];
        auto load_codes = [&](uint32_t c) {
            const uint32_t k0 = c * WF_CH;
            const uint32_t kcnt = k0 < n2 ? min((uint32_t)WF_CH, n2 - k0) : 0u;
#pragma unroll
            for (int r = 0; r < 32; r++) {
                const uint32_t ri = i0 + ((DIM == 3) ? r / PJ : 0), rj = j0 + ((DIM == 3) ? r % PJ : r);
                cv[r] = ((uint32_t)lane < kcnt && ri < n0 && rj < n1) ? __ldg(codes + ((uint64_t)ri * n1 + rj) * n2 + k0 + lane) : (uint16_t)0;
            }
        };
        auto store_codes = [&](uint32_t c) {
            const uint32_t slot0 = (c & 3) * WF_CH;
#pragma unroll
            for (int r = 0; r < 32; r++) inr[(size_t)r * WF_IS + slot0 + lane] = (T)cv[r];      // codes < 65536: exact in T
        };
        auto stage_input = [&](uint32_t c) {
            const uint32_t k0 = c * WF_CH;
            if (!DECOMP && k0 < n2) {
                const uint32_t kcnt = min((uint32_t)WF_CH, n2 - k0);
                const uint32_t slot0 = (c & 3) * WF_CH;
                if (!DECOMP) {
                    if ((uint32_t)lane < kcnt) {
#pragma unroll 8
                        for (int r = 0; r < 32; r++) {
                            const uint32_t ri = i0 + ((DIM == 3) ? r / PJ : 0), rj = j0 + ((DIM == 3) ? r % PJ : r);
                            if (ri < n0 && rj < n1) cp_async_elem<T>(inr + (size_t)r * WF_IS + slot0 + lane, x + ((uint64_t)ri * n1 + rj) * n2 + k0 + lane);
                        }
                    }
                }
            }
            if (!DECOMP) cp_async_commit();          // one group per chunk, even when empty, keeps the wait count uniform
        };
        // halo rows of chunk c from the neighbouring pencils' published reconstruction (L2 loads, never L1): all loads are
        // issued before the first shared-memory store so that they overlap
        auto stage_halo = [&](uint32_t c) {
            const uint32_t k0 = c * WF_CH;
            const uint32_t kcnt = min((uint32_t)WF_CH, n2 - k0);
            const uint32_t slot0 = (c & 1) * WF_CH;
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
                for (int q = -1; q < PJ; q++) ring[((size_t)0 * RJ + (q + 1)) * WF_S + slot0 + lane] = hu[q + 1];
            }
            if (pb > 0) {
#pragma unroll
                for (int q = 0; q < PI; q++) ring[((size_t)(q + 1) * RJ + 0) * WF_S + slot0 + lane] = hl[q];
            }
        };

        const int *flag_up = (DIM == 3 && pa > 0) ? P.flags + (pid - P.pencils_b) : nullptr;
        const int *flag_left = (pb > 0) ? P.flags + (pid - 1) : nullptr;
        const uint32_t tsteps = n2 + DMAX;
        // input prefetch runs two chunks ahead of the chunk lane 0 is in
        if (!DECOMP) { stage_input(0); stage_input(1); }
        else { load_codes(0); store_codes(0); load_codes(1); }
        for (uint32_t t = 0; t < tsteps; t++) {
            if ((t % WF_CH) == 0 && t < n2) {
                const uint32_t c = t / WF_CH;
                if (lane == 0) {
                    if (flag_up) ok = ok && wf_wait(flag_up, (int)c + 1, P.err);
                    if (flag_left) ok = ok && wf_wait(flag_left, (int)c + 1, P.err);
                }
                ok = __shfl_sync(0xffffffffu, (int)ok, 0) != 0;
                if (!ok) break;
                stage_halo(c);
                if (!DECOMP) { stage_input(c + 2); cp_async_wait<2>(); }   // chunk c has landed (groups c+1, c+2 may still be in flight)
                else { if (c > 0) store_codes(c); load_codes(c + 1); }
                __syncwarp();
            }
            const int64_t ks = (int64_t)t - delay;
            if (active && ks >= 0 && ks < (int64_t)n2) {
                const uint32_t k = (uint32_t)ks;
                const uint32_t slot = k % WF_W, islot = k % WF_IW;
                if (kk == sz) { bk++; kk = 0; sz = g.ax[2].edge(bk); }
                if (kk == 0) {
                    const uint32_t bid = g.block_id(b0, b1, bk);
                    lorenzo = P.indicator[bid] != 0;
                    if (!lorenzo) {
                        const uint32_t nb = g.num_blocks;
                        if (DIM == 3) {
                            T l0 = coeff_last[bid], l1 = coeff_last[nb + bid];
                            T m0 = l0 * (T)ii, m1 = l1 * (T)jj;
                            c_base = m0 + m1; c_k = coeff_last[2 * nb + bid]; c_d = coeff_last[3 * nb + bid];
                        } else {
                            T l0 = coeff_last[bid];
                            c_base = l0 * (T)jj; c_k = coeff_last[nb + bid]; c_d = coeff_last[2 * nb + bid];
                        }
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
                        q = quantize<T>(cur, pred, recip_eb, eb, P.cap, P.radius, rec);
                    } else {
                        bool done = false;
                        if (MEAN) { T dm = cur - mean; if (TT<T>::abs(dm) <= eb) { q = P.radius; rec = mean; done = true; } }
                        if (!done) {
                            T pred;
                            if (DIM == 3) {   // main-loop term order (ref: sz_float.c:7268-7269)
                                pred = ra + rb; pred = pred + rc; pred = pred - rd; pred = pred - re; pred = pred - rf; pred = pred + rg;
                            } else {          // ref: sz_float.c:6204
                                pred = ra + rb; pred = pred - rd;
                            }
                            if (!MEAN) q = quantize<T>(cur, pred, recip_eb, eb, P.cap_l, P.radius, rec);
                            else {            // ref: sz_float.c:6936-6956 (code shift happens BEFORE the verification)
                                T diff = cur - pred;
                                T itv = TT<T>::abs(diff) * recip_eb; itv = itv + (T)1;
                                if (itv < (T)P.cap_l) {
                                    if (diff < (T)0) itv = -itv;
                                    q = (int)(itv / (T)2) + P.radius;
                                    T dq = (T)(2 * (q - P.radius));
                                    rec = dq * eb; rec = pred + rec;
                                    if (q <= P.radius) q -= 1;
                                    T chk = cur - rec;
                                    if (TT<T>::abs(chk) > eb) { q = 0; rec = cur; }
                                } else { q = 0; rec = cur; }
                            }
                        }
                    }
                    codes[rowbase + k] = (uint16_t)q;
                    if (P.write_all_recon || li == PI - 1 || lj == PJ - 1) recon[rowbase + k] = rec;
                } else {
                    int tcode = (int)in;
                    if (!lorenzo) {
                        if (tcode != 0) {
                            T pred = c_k * (T)kk; pred = c_base + pred; pred = pred + c_d;
                            T dq = (T)(2 * (tcode - P.radius)); T w = dq * eb; rec = pred + w;
                        } else rec = recon[rowbase + k];                  // unpredictable value, prefilled
                    } else {
                        if (MEAN && tcode == P.radius) rec = mean;
                        else if (tcode == 0) rec = recon[rowbase + k];
                        else {
                            if (MEAN && tcode < P.radius) tcode += 1;
                            T pred;
                            if (DIM == 3) { pred = ra + rb; pred = pred + rc; pred = pred - rd; pred = pred - re; pred = pred - rf; pred = pred + rg; }
                            else { pred = ra + rb; pred = pred - rd; }
                            T dq = (T)(2 * (tcode - P.radius)); T w = dq * eb; rec = pred + w;
                        }
                    }
                    recon[rowbase + k] = rec;
                }
                my_row[slot] = rec;
                ra = rec; rd = rb; re = rc; rg = rf;
                kk++;
            }
            __syncwarp();
            // publish chunk c once the slowest lane has finished it
            if (t >= (uint32_t)DMAX && ((t - DMAX) % WF_CH) == WF_CH - 1) {
                __threadfence();
                __syncwarp();
                if (lane == 0) st_release(P.flags + pid, (int)((t - DMAX) / WF_CH) + 1);
            }
        }
        if (!DECOMP) cp_async_wait<0>();
        if (!ok) return;
        __threadfence();
        __syncwarp();
        if (lane == 0) st_release(P.flags + pid, (int)nchunks);
        __syncwarp();
    }
}

}  // namespace szb
