// szb_chain.cuh -- K5, the coefficient DPCM chain (ref: sz/src/sz_float.c:7126-7152 no-mean "/precision", :6792-6812 mean
// "*recip", 2D :6035-6055; sz_double.c the same with "*recip"), as "lattice guess, add chain, exact proof".
//
// The reference threads last[e] through every regression block in raster order:
//     code_i = F(cur_i, last_{i-1}),   last_i = fl(last_{i-1} + fl(2 (code_i - 32768) * prec))      (or cur_i when unpredictable)
// a float recurrence whose step is a dependent chain of a subtraction, a division, two additions, two conversions, a product, a
// sum and a check.  In exact arithmetic every `last` lies on the lattice base + 2 prec Z (base = the value after the latest
// unpredictable coefficient) and last_i is the lattice point nearest to cur_i whatever the history: the lattice index
// k_i = round((cur_i - base) / (2 prec)) has a closed form and code_i = k_i - k_{i-1} + 32768.  Floating point only moves `last`
// off the lattice by a few ulps; that matters for the code when cur_i lies within this drift of the midpoint of two lattice
// points, and it always matters for the bits of last_i.  One CTA per coefficient, windows of LAT_W steps, three stages that run
// side by side on consecutive windows (two __syncthreads per window):
//   (G1) helper thread h guesses the code of step h of the NEXT window from the closed form (double arithmetic, relative to the exact
//       `last` that entered the window (S) is running) and prepares the step in up to three forms: SURE (one increment
//       fl(2 dk * prec)); PAIR when cur_h may lie within the drift band of a midpoint -- two candidate codes of magnitudes n, n+1,
//       told apart by an exact threshold on |cur - last| that the thread finds by evaluating the reference's own quotient at the
//       floats around (2n+1) prec -- or when it follows such a step (two codes again, any threshold between them);
//   (G2) at the top of the next tick, when the exact `last` entering the window is known, the same thread subtracts the measured
//       drift, settles the class with the band of its own window only and writes the step's table entry; FULL when no prepared
//       form is safe (two midpoint steps in a row, near the code range, not finite);
//   (S) ONE thread runs the recurrence: 32 or 8 SURE steps are that many adds; a group of 8 with PAIR steps runs branch-free as
//       subtraction, compare, select, add per step plus the reference's check; a FULL step is the reference's complete formula.  An
//       unpredictable coefficient (new base) ends the window.
//   (V) helper thread h proves step h with the complete formula on the exact `last` that entered it: code and bits of the new
//       `last`.  A window that fails is rerun from its exact entry value with a wider band, then with every step FULL, i.e. as the
//       reference's own loop.
// The output is the reference's sequence by induction over the proved steps, not by the quality of the guess, which only decides
// the speed.
#pragma once

#include "szb_stage1.cuh"

namespace szb {

constexpr int LAT_W = 256;                    // steps per window = helper threads
constexpr int LAT_THREADS = 11 * 32;           // warp 0: the serial thread; warps 1-3, 5-7, 9, 10: helpers; warps 4 and 8 share warp 0's scheduler and stay idle
constexpr float LAT_BAND = 0.35f;             // drift band in units of mag eps sqrt(steps), see (G1)
template <typename T> struct LatEps;
template <> struct LatEps<float> { static constexpr double eps = 1.0 / 8388608.0; };              // 2^-23: ulp(x) <= eps |x|
template <> struct LatEps<double> { static constexpr double eps = 1.0 / 4503599627370496.0; };    // 2^-52

template <typename T> struct __align__(16) LatSlot {        // one window in shared memory
    T cur[LAT_W], inc[LAT_W], incB[LAT_W], thr[LAT_W], out[LAT_W];     // out[i]: last after step i
    int q[LAT_W], qB[LAT_W];
    unsigned int hm[LAT_W / 32], xm[LAT_W / 32];                        // one bit per step: PAIR form / FULL
    unsigned char bigm[LAT_W / 8];                                      // written by (S): PAIR steps that took their second candidate
    unsigned int wend; T exit;                                          // written by (S): steps done, last after them
    float maxmag;                                                       // written by (G1): largest magnitude the window's values can have
};

__device__ __forceinline__ void lat_helper_bar() { asm volatile("bar.sync 1, %0;" ::"n"(LAT_W) : "memory"); }

// magnitude of the code the reference's formula gives for |cur - last| = d (monotone in d; ref: sz_float.c:7133-7139)
template <typename T>
__device__ __forceinline__ int lat_mag(T d, T prec, T rprec, int use_div) {
    T itv = use_div ? d / prec : d * rprec;
    itv = itv + (T)1;
    return itv < (T)65536 ? (int)(itv / (T)2) : 65536;
}

__device__ __forceinline__ void lat_ld8(const float *p, float v[8]) {
    const float4 a = *reinterpret_cast<const float4 *>(p), b = *reinterpret_cast<const float4 *>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void lat_ld8(const double *p, double v[8]) {
#pragma unroll
    for (int t = 0; t < 4; t++) { const double2 a = *reinterpret_cast<const double2 *>(p + 2 * t); v[2 * t] = a.x; v[2 * t + 1] = a.y; }
}
__device__ __forceinline__ void lat_st8(float *p, const float v[8]) {
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]); *reinterpret_cast<float4 *>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void lat_st8(double *p, const double v[8]) {
#pragma unroll
    for (int t = 0; t < 4; t++) *reinterpret_cast<double2 *>(p + 2 * t) = make_double2(v[2 * t], v[2 * t + 1]);
}

// The running `last` goes in and out of these by value (a reference parameter would put it into local memory).
template <typename T> struct LatStep { T s; int ended; };

// Eight steps of which some are PAIR and none FULL, without a branch: every step in the PAIR form (a SURE step has an infinite
// threshold).  ended = 1: some step failed the reference's check (unpredictable coefficient); the caller reruns the group step by step.
template <typename T>
__device__ __noinline__ LatStep<T> lat_pair_group(LatSlot<T> &L, int i0, T prec, T s) {
    T cur[8], thr[8], iA[8], iB[8], o[8];
    lat_ld8(&L.cur[i0], cur); lat_ld8(&L.thr[i0], thr); lat_ld8(&L.inc[i0], iA); lat_ld8(&L.incB[i0], iB);
    unsigned int bigm = 0u; bool unp = false;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const T diff = cur[t] - s;
        const bool big = !(TT<T>::abs(diff) < thr[t]);
        s = s + (big ? iB[t] : iA[t]);
        o[t] = s;
        const T chk = cur[t] - s;
        unp = unp || TT<T>::abs(chk) > prec;
        bigm |= big ? (1u << t) : 0u;
    }
    lat_st8(&L.out[i0], o);
    L.bigm[i0 >> 3] = (unsigned char)bigm;
    LatStep<T> r; r.s = s; r.ended = unp ? 1 : 0;
    return r;
}

// Eight steps one by one: SURE an add, PAIR with the reference's check, FULL the reference's complete formula.  ended: index of the
// first unpredictable coefficient (the window ends with it; the steps behind it are not run), or -1.
template <typename T>
__device__ __noinline__ LatStep<T> lat_slow_group(LatSlot<T> &L, int i0, unsigned int hm8, unsigned int xm8, T prec, T rprec, int use_div, T s) {
    unsigned int bigm = 0u; int ended = -1;
    for (int t = 0; t < 8 && ended < 0; t++) {
        const int i = i0 + t;
        const T cur = L.cur[i];
        if ((xm8 >> t) & 1u) {
            T nl;
            const int ct = chain_step_exact<T>(cur, s, prec, rprec, use_div, nl);
            L.q[i] = ct; s = nl;
            if (ct == 0) ended = t;
        } else if ((hm8 >> t) & 1u) {
            const T diff = cur - s;
            const bool big = !(TT<T>::abs(diff) < L.thr[i]);
            const T v = s + (big ? L.incB[i] : L.inc[i]);
            const T chk = cur - v;
            if (TT<T>::abs(chk) > prec) { L.q[i] = 0; L.qB[i] = 0; s = cur; ended = t; }
            else { s = v; bigm |= big ? (1u << t) : 0u; }
        } else s = s + L.inc[i];
        L.out[i] = s;
    }
    L.bigm[i0 >> 3] = (unsigned char)bigm;
    LatStep<T> r; r.s = s; r.ended = ended;
    return r;
}

// 32 steps with their increments in registers; returns true when the window ends in this word (wend set)
template <typename T>
__device__ __forceinline__ bool lat_word(LatSlot<T> &L, int w, const T (&v)[32], unsigned int hmw, unsigned int xmw, T prec, T rprec, int use_div, T &s, int &wend, unsigned int *prof) {
#ifdef LAT_PROF
    const long long pc0 = clock64();
#define LAT_PROF_ADD(k) { prof[2 * (k)] += 1u; prof[2 * (k) + 1] += (unsigned int)(clock64() - pc1); }
#else
#define LAT_PROF_ADD(k)
#endif
    if ((hmw | xmw) == 0u) {                                  // 32 SURE steps (steps past the window's end add zero)
        T o[32];
#pragma unroll
        for (int t = 0; t < 32; t++) { s = s + v[t]; o[t] = s; }
#pragma unroll
        for (int g = 0; g < 4; g++) lat_st8(&L.out[32 * w + 8 * g], &o[8 * g]);
#ifdef LAT_PROF
        { const long long pc1 = pc0; LAT_PROF_ADD(0) }
#endif
        return false;
    }
#pragma unroll
    for (int g = 0; g < 4; g++) {
        const int i0 = 32 * w + 8 * g;
        const unsigned int h8 = (hmw >> (8 * g)) & 0xffu, x8 = (xmw >> (8 * g)) & 0xffu;
#ifdef LAT_PROF
        const long long pc1 = clock64();
#endif
        if ((h8 | x8) == 0u) {
            T o[8];
#pragma unroll
            for (int t = 0; t < 8; t++) { s = s + v[8 * g + t]; o[t] = s; }
            lat_st8(&L.out[i0], o);
            LAT_PROF_ADD(1)
        } else {
            LatStep<T> r; r.ended = 1;
            if (x8 == 0u) r = lat_pair_group<T>(L, i0, prec, s);
            if (r.ended) {
                r = lat_slow_group<T>(L, i0, h8, x8, prec, rprec, use_div, s);
                if (r.ended >= 0) { wend = i0 + r.ended + 1; return true; }
            }
            s = r.s;
            LAT_PROF_ADD(x8 ? 3 : 2)
        }
    }
    return false;
}
__device__ __forceinline__ void lat_ld32(const float *p, float (&v)[32]) {
#pragma unroll
    for (int g = 0; g < 4; g++) lat_ld8(p + 8 * g, &v[8 * g]);
}
__device__ __forceinline__ void lat_ld32(const double *p, double (&v)[32]) {
#pragma unroll
    for (int g = 0; g < 4; g++) lat_ld8(p + 8 * g, &v[8 * g]);
}

struct LatHist { uint32_t *h[4]; };          // per coefficient: the histogram of its codes (65536 bins, zeroed), counted by the proof stage

template <typename T> struct LatWin { uint32_t p; int cnt, slot, mode; T entry; bool valid, fresh; };
// what (G1) leaves in the registers of helper h for (G2), one tick later
template <typename T> struct LatPrep {
    double base, frac, adk, band;          // the exact `last` the guess is relative to; cur's offset from its lattice point; |dk|; the band of (G2)
    int kprev0;                            // lattice index of the step before the window under `base`
    bool in, valid, pairok, followok;      // step inside the window; numbers finite; PAIR data of a midpoint step / of a step behind one prepared
    int q0, qA, qB, fqA, fqB; T inc0, incA, incB, thr, fincA, fincB, fthr;
};

template <typename T>
__global__ void __launch_bounds__(LAT_THREADS) coeff_lattice_kernel(const T *__restrict__ reg_c /*[ncoef][nb], compacted*/, const uint32_t *__restrict__ reg_list,
                                                                    uint32_t nb, uint32_t reg_count, T p0, T p1, T p2, T p3, int use_div,
                                                                    uint16_t *__restrict__ coeff_type /*[ncoef][nb]*/, T *__restrict__ coeff_last /*[ncoef][nb] by block id*/,
                                                                    T *__restrict__ coeff_unpred /*[ncoef][nb]*/, LatHist hists, unsigned int *done_flags /*[ncoef]: set when the coefficient's codes and histogram are final*/, ChainOut *out, int first_mode, float band_scale) {
    typedef typename TT<T>::bits_t bits_t;
    constexpr int W = LAT_W;
    __shared__ LatSlot<T> s_slot[2];
    __shared__ int s_k[W];
    __shared__ unsigned char s_tie[W];            // (G1) bit 0: may be at a midpoint, bit 1: cur above its lattice point, bit 2: PAIR data ready; (G2) bit 3: at a midpoint
    __shared__ float s_wmag[W / 32];
    __shared__ unsigned int s_wcnt[2][W / 32];    // by tick parity: read after the tick's barrier, written again one tick later
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool helper = warp != 0 && warp != 4 && warp != 8;          // the serial thread keeps its scheduler to itself
    const int hw = warp - 1 - (warp > 4 ? 1 : 0) - (warp > 8 ? 1 : 0), h = hw * 32 + lane;
    const int e = blockIdx.x;
    const T prec = e == 0 ? p0 : e == 1 ? p1 : e == 2 ? p2 : p3;
    const T rprec = (T)1 / prec;
    const double inv2p = 1.0 / (2.0 * (double)prec);
    const T *cur_e = reg_c + (size_t)e * nb;
    uint32_t *hist_e = hists.h[e];
    T t_inf; { const bits_t b = sizeof(T) == 4 ? (bits_t)0x7f800000u : (bits_t)0x7ff0000000000000ull; memcpy(&t_inf, &b, sizeof(T)); }
    const double sqrt_h1 = (double)sqrtf((float)(h > 0 ? h + 1 : 1)) * 1.000001;
    uint32_t ucount = 0;
    unsigned int n_ticks = 0, n_redo = 0, n_pair = 0, n_full = 0;
    long long cyc_s = 0, cyc_h = 0; const long long cyc_0 = clock64();
    unsigned long long dbg_g0; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(dbg_g0));
    unsigned int prof[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // LAT_PROF builds: {count, cycles} of clean words, clean groups, PAIR groups, groups run step by step
    LatWin<T> vw, sw, gw;
    LatPrep<T> P; P.in = false;
    vw.valid = false; sw.valid = false;
    gw.p = 0; gw.cnt = (int)min((uint32_t)W, reg_count); gw.slot = 0; gw.mode = first_mode; gw.entry = (T)0; gw.valid = reg_count > 0; gw.fresh = true;
    while (vw.valid || sw.valid || gw.valid) {
        // ======================== (G2): the window (S) runs in this tick gets its final table ========================
        if (sw.valid) {
            if (helper) {
                LatSlot<T> &L = s_slot[sw.slot];
                // measured drift: where the exact `last` entering the window lies relative to the lattice the guess used
                const double y = ((double)sw.entry - P.base) * inv2p, yk = rint(y), rho = y - yk;
                const bool mismatch = !(fabs(y) < 1073741824.0) || (int)yk != P.kprev0;      // the step before went to another lattice point
                const bool tie = P.in && !(0.5 - fabs(P.frac - rho) > P.band);
                s_tie[h] = (unsigned char)((s_tie[h] & 7) | (tie ? 8 : 0));
                lat_helper_bar();
                const bool tie1 = h >= 1 && (s_tie[h - 1] & 8) != 0, tie2 = h >= 2 && (s_tie[h - 2] & 8) != 0;
                const bool pred_pairok = h >= 1 && (s_tie[h - 1] & 4) != 0;
                bool full = P.in && (sw.mode >= 2 || !P.valid || P.adk >= 32760.0 || (h == 0 && mismatch) ||
                                     (tie && (tie1 || !P.pairok)) || (!tie && tie1 && (tie2 || !pred_pairok || !P.followok)));
                const bool pairform = P.in && !full && (tie || tie1);
                T inc = P.inc0, incB = (T)0, thr = t_inf; int q = P.q0, qB = 0;
                if (pairform && tie) { inc = P.incA; incB = P.incB; thr = P.thr; q = P.qA; qB = P.qB; }
                else if (pairform) { inc = P.fincA; incB = P.fincB; thr = P.fthr; q = P.fqA; qB = P.fqB; }
                if (full || !P.in) { inc = (T)0; q = 0; }
                const unsigned int hm = __ballot_sync(0xffffffffu, pairform), xm = __ballot_sync(0xffffffffu, full);
                if (lane == 0) { L.hm[hw] = hm; L.xm[hw] = xm; }
                L.inc[h] = inc; L.incB[h] = incB; L.thr[h] = thr; L.q[h] = q; L.qB[h] = qB;
                n_pair += __popc(hm); n_full += __popc(xm);
            }
            __syncthreads();
        }
        bool my_fail = false;
        int v_ct = -1;                                     // helpers: the code of the step proved in this tick (counted once the window has passed)
        if (warp == 0) {
            // ================================================ (S) ================================================
            if (lane == 0 && sw.valid) {
                const long long c0 = clock64();
                LatSlot<T> &L = s_slot[sw.slot];
                T s = sw.entry;
                int wend = sw.cnt;
                const int nwords = (sw.cnt + 31) >> 5;
                T a[32], b[32];                                      // increments of two words: one in use, one being fetched
                lat_ld32(&L.inc[0], a);
                unsigned int hm_a = L.hm[0], xm_a = L.xm[0], hm_b = 0u, xm_b = 0u;
                for (int w = 0; w < nwords; w += 2) {
                    if (w + 1 < nwords) { lat_ld32(&L.inc[32 * (w + 1)], b); hm_b = L.hm[w + 1]; xm_b = L.xm[w + 1]; }
                    if (lat_word<T>(L, w, a, hm_a, xm_a, prec, rprec, use_div, s, wend, prof)) break;
                    if (w + 1 >= nwords) break;
                    if (w + 2 < nwords) { lat_ld32(&L.inc[32 * (w + 2)], a); hm_a = L.hm[w + 2]; xm_a = L.xm[w + 2]; }
                    if (lat_word<T>(L, w + 1, b, hm_b, xm_b, prec, rprec, use_div, s, wend, prof)) break;
                }
                L.wend = (unsigned int)wend;
                L.exit = L.out[wend - 1];
                cyc_s += clock64() - c0;
            }
        } else if (helper) {
            const long long c0 = clock64();
            // global loads of this tick first: their latency hides behind (V)
            T cur_g = (T)0; uint32_t blk_v = 0;
            if (gw.valid && h < gw.cnt) cur_g = cur_e[gw.p + h];
            if (vw.valid && h < vw.cnt) blk_v = reg_list[vw.p + h];
            // ================================================ (V) ================================================
            int ct = 1; T lout = (T)0, vcur = (T)0; unsigned int zmask = 0;
            if (vw.valid) {
                LatSlot<T> &L = s_slot[vw.slot];
                if (h < vw.cnt) {
                    vcur = L.cur[h];
                    const T lin = h ? L.out[h - 1] : vw.entry;
                    lout = L.out[h];
                    T nl;
                    ct = chain_step_exact<T>(vcur, lin, prec, rprec, use_div, nl);
                    bits_t bn, bo; memcpy(&bn, &nl, sizeof(T)); memcpy(&bo, &lout, sizeof(T));
                    const bool pairform = (L.hm[h >> 5] >> (h & 31)) & 1u, big = pairform && ((L.bigm[h >> 3] >> (h & 7)) & 1u);
                    my_fail = ct != (big ? L.qB[h] : L.q[h]) || bn != bo;
                    v_ct = ct;
                }
                zmask = __ballot_sync(0xffffffffu, h < vw.cnt && ct == 0);
                if (lane == 0) s_wcnt[n_ticks & 1u][hw] = __popc(zmask);
            }
            lat_helper_bar();
            if (vw.valid && h < vw.cnt) {            // written before the verdict: a failed window is rerun and writes the same places again
                coeff_type[(size_t)e * nb + vw.p + h] = (uint16_t)ct;
                coeff_last[(size_t)e * nb + blk_v] = lout;
                if (ct == 0) {
                    unsigned int before = 0;
                    for (int w = 0; w < hw; w++) before += s_wcnt[n_ticks & 1u][w];
                    coeff_unpred[(size_t)e * nb + ucount + before + __popc(zmask & ((1u << lane) - 1u))] = vcur;
                }
            }
            // ================================================ (G1) ================================================
            if (gw.valid) {
                LatSlot<T> &L = s_slot[gw.slot];
                // the guess is relative to an exact `last`: the window's own entry value (fresh), or the entry value of the window
                // that (S) is running now, one window back
                const double base = (double)(gw.fresh ? gw.entry : sw.entry);
                const int m0 = gw.fresh ? 0 : sw.cnt;
                const bool in = h < gw.cnt;
                const double x = ((double)cur_g - base) * inv2p;
                const double kr = rint(x);
                const bool valid = in && fabs(x) < 1073741824.0;            // false for NaN/inf too
                const int k = valid ? (int)kr : 0;
                const bool up = x > kr;
                L.cur[h] = cur_g; s_k[h] = k;
                {
                    float m = in ? fabsf((float)cur_g) : 0.0f;
                    if (!(m <= 3.0e38f)) m = 3.0e38f;
#pragma unroll
                    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                    if (lane == 0) s_wmag[hw] = m;
                }
                // the step before the window (pipelined windows only): its lattice index under the same base
                int kprev0 = 0; bool valid_prev0 = true;
                double magp = fabs(base);
                if (!gw.fresh) {
                    magp = fmax(magp, (double)s_slot[sw.slot].maxmag);      // the drift is run up in that window too
                    const T pc = s_slot[sw.slot].cur[sw.cnt - 1];
                    const double xp = ((double)pc - base) * inv2p;
                    valid_prev0 = fabs(xp) < 1073741824.0;
                    kprev0 = valid_prev0 ? (int)rint(xp) : 0;
                }
                lat_helper_bar();
                const int kprev = h ? s_k[h - 1] : kprev0;
                const long long dk = (long long)k - (long long)kprev;
                const double adk = fabs((double)dk);
                float wm = s_wmag[0];
#pragma unroll
                for (int w = 1; w < W / 32; w++) wm = fmaxf(wm, s_wmag[w]);
                const double mag = fmax((double)wm * 1.0000002, magp);
                if (h == 0) L.maxmag = (float)fmin(mag * 1.0000002, 3.0e38);
                // may this step be at a midpoint when (G2) looks?  The band of both windows, with a margin.
                const double pband = 1.5 * (gw.mode == 0 ? (double)band_scale : 16.0 * (double)band_scale) * (mag * LatEps<T>::eps) * (double)sqrtf((float)(m0 + h + 1)) * inv2p
                                     + (adk + 2.0) * (2.5 * LatEps<T>::eps) + (fabs(x) + 1.0) * 1e-15;
                const bool ptie = valid && !(0.5 - fabs(x - kr) > pband);
                const bool valid_prev = h ? (fabs(((double)L.cur[h - 1] - base) * inv2p) < 1073741824.0) : valid_prev0;
                // Drift of `last` after m steps from an exact value: two roundings per step (product, sum), each uniform within half an
                // ulp of a value no larger than mag -- a random walk with sigma <= 0.25 mag eps per step, in practice less (values
                // are rarely at the window's maximum).  A window that leaves the band is caught by (V) and rerun with a wider one, so
                // the band is set for speed: LAT_BAND = 0.35 (the 512^3 field: no rerun at 0.35, one in 409 windows at 0.25).
                // Plus the formula's own roundings (difference, quotient, +1 and the reciprocal: at most four half ulps of the
                // quotient 2|dk|, i.e. 2 eps |dk| lattice units) and the error of this double-precision estimate.  (G2) measures the
                // drift up to the window's entry: m = h + 1.
                const double bscale = (gw.mode == 0 ? (double)band_scale : 16.0 * (double)band_scale) * (mag * LatEps<T>::eps) * inv2p;
                const double bfix = (adk + 2.0) * (2.5 * LatEps<T>::eps) + (fabs(x) + 1.0) * 1e-15;
                P.base = base; P.frac = x - kr; P.adk = adk; P.band = bscale * sqrt_h1 + bfix;
                P.kprev0 = kprev0; P.in = in; P.valid = valid && valid_prev;
                P.q0 = (int)dk + 32768; { const T dq = (T)(2 * (int)dk); P.inc0 = dq * prec; }
                P.pairok = false; P.followok = false;
                if (ptie && P.valid && adk < 32760.0) {
                    // the candidates: magnitudes n0 (A) and n0 + 1 (B) with one sign
                    long long dA = dk, dB = dk + (up ? 1 : -1);       // the lattice point on the other side of the midpoint
                    long long aA = dA < 0 ? -dA : dA, aB = dB < 0 ? -dB : dB;
                    if (aA > aB) { long long t = dA; dA = dB; dB = t; t = aA; aA = aB; aB = t; }
                    // exact threshold: the smallest float d with lat_mag(d) >= aB, looked for among the floats around (2 aB - 1) prec
                    const T d0 = (T)((double)(2 * aB - 1) * (double)prec);
                    bits_t b0; memcpy(&b0, &d0, sizeof(T));
                    // (2 aB - 1) prec itself passes unless the rounding of d0 went below it; the threshold is at most a few floats
                    // further down (the quotient and the +1 round to nearest): walk from d0 to the first float that fails, or up to the
                    // first that passes -- two or three evaluations of the monotone function instead of a table of nine
                    bool ok = d0 > (T)0 && b0 >= (bits_t)16;
                    int first = 0;
                    auto passes = [&](int o) { const bits_t b = b0 + (bits_t)(long long)o; T d; memcpy(&d, &b, sizeof(T)); return lat_mag<T>(d, prec, rprec, use_div) >= (int)aB; };
                    if (ok) {
                        if (passes(0)) { while (first > -7 && passes(first - 1)) first--; ok = first > -7; }
                        else { first = 1; while (first < 7 && !passes(first)) first++; ok = first < 7; }
                    }
                    if (ok) {
                        const bits_t b = b0 + (bits_t)(long long)first; memcpy(&P.thr, &b, sizeof(T));
                        P.qA = (int)dA + 32768; P.qB = (int)dB + 32768;
                        const T dqA = (T)(2 * (int)dA), dqB = (T)(2 * (int)dB);
                        P.incA = dqA * prec; P.incB = dqB * prec;
                        P.pairok = true;
                    }
                }
                s_tie[h] = (unsigned char)((ptie ? 1 : 0) | (up ? 2 : 0) | (P.pairok ? 4 : 0));
                lat_helper_bar();
                if (h >= 1 && (s_tie[h - 1] & 1) && P.valid) {
                    // behind a step that may be at a midpoint: that step's candidates in ITS order (A = smaller magnitude of its own
                    // code), this step's code from either, and any threshold between the two magnitudes (they differ by 2 prec in
                    // |cur - last|, the midpoint is a whole prec from both)
                    const unsigned char tp = s_tie[h - 1];
                    const long long kpp = h >= 2 ? (long long)s_k[h - 2] : (long long)kprev0;
                    long long kA = kprev, kB = (long long)kprev + ((tp & 2) ? 1 : -1);
                    const long long pA = kA - kpp, pB = kB - kpp;
                    if ((pA < 0 ? -pA : pA) > (pB < 0 ? -pB : pB)) { const long long t = kA; kA = kB; kB = t; }
                    long long fA = (long long)k - kA, fB = (long long)k - kB;
                    long long aA = fA < 0 ? -fA : fA, aB = fB < 0 ? -fB : fB;
                    if (aA > aB) { long long t = fA; fA = fB; fB = t; t = aA; aA = aB; aB = t; }
                    if (aB < 32760) {
                        P.fqA = (int)fA + 32768; P.fqB = (int)fB + 32768;
                        const T dqA = (T)(2 * (int)fA), dqB = (T)(2 * (int)fB);
                        P.fincA = dqA * prec; P.fincB = dqB * prec;
                        P.fthr = (T)((double)(2 * aB - 1) * (double)prec);
                        P.followok = true;
                    }
                }
            }
            cyc_h += clock64() - c0;
        }
        const int fail = __syncthreads_or(my_fail ? 1 : 0);
        const unsigned int par = n_ticks & 1u;
        n_ticks++;
        // ---- what the next tick does (the same decision in every thread) ----
        if (vw.valid) {
            if (fail) {                                    // rerun the failed window from its exact entry value; everything after it is void
                if (vw.mode >= 2) { if (tid == 0) { out->err = 1u; if (done_flags) atomicExch(done_flags + e, 1u); } return; }     // the complete formula disagrees with itself: cannot happen
                gw.p = vw.p; gw.cnt = (int)min((uint32_t)W, reg_count - vw.p); gw.slot = vw.slot; gw.mode = vw.mode + 1; gw.entry = vw.entry; gw.valid = true; gw.fresh = true;
                vw.valid = false; sw.valid = false; n_redo++;
                continue;
            }
            if (v_ct >= 0) atomicAdd(hist_e + v_ct, 1u);   // the window has passed: its codes are final
            if (helper) {                                  // (the count of unpredictable coefficients is the helpers' business)
                unsigned int tot = 0;
#pragma unroll
                for (int w = 0; w < W / 32; w++) tot += s_wcnt[par][w];
                ucount += tot;
            }
        }
        if (sw.valid) {
            const int wend = (int)s_slot[sw.slot].wend;
            const T exit_v = s_slot[sw.slot].exit;
            const bool early = wend < sw.cnt;
            vw = sw; vw.cnt = wend;
            const uint32_t np = sw.p + (uint32_t)wend;
            if (early) {                                   // a new base: the window prepared behind it is void
                gw.p = np; gw.cnt = (int)min((uint32_t)W, reg_count - np); gw.slot = sw.slot ^ 1; gw.mode = first_mode; gw.entry = exit_v; gw.valid = np < reg_count; gw.fresh = true;
                sw.valid = false;
                continue;
            }
            if (gw.valid) { sw = gw; if (!gw.fresh) sw.entry = exit_v; }
            else sw.valid = false;
        } else {
            vw.valid = false;
            if (gw.valid) sw = gw; else sw.valid = false;      // a fresh window carries its own entry value
        }
        if (sw.valid && sw.p + (uint32_t)sw.cnt < reg_count) {
            gw.p = sw.p + (uint32_t)sw.cnt; gw.cnt = (int)min((uint32_t)W, reg_count - gw.p); gw.slot = sw.slot ^ 1; gw.mode = first_mode; gw.valid = true; gw.fresh = false;
        } else gw.valid = false;
    }
    __threadfence();                                 // every thread's histogram counts and stores, then the flag
    __syncthreads();
    if (tid == 32) out->unpred_count[e] = ucount;          // (helper 0: the serial thread's warp does not keep the count)
    if (tid == 0) {
        if (done_flags) atomicExch(done_flags + e, 1u);
        if (e == (int)gridDim.x - 1) {
            out->passes = n_ticks; out->slow_chunks = n_redo; out->kcyc_serial = (unsigned int)(cyc_s >> 10);
            out->kcyc_total = (unsigned int)((clock64() - cyc_0) >> 10);
            for (int i = 0; i < 8; i++) out->prof[i] = prof[i];
            { unsigned long long g1; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1)); out->prof[6] = (unsigned int)(dbg_g0 / 1000ull); out->prof[7] = (unsigned int)(g1 / 1000ull); }
        }
    }
    if (e == (int)gridDim.x - 1 && helper && lane == 0) { atomicAdd(&out->n_pair, n_pair); atomicAdd(&out->n_full, n_full); if (hw == 0) out->kcyc_helper = (unsigned int)(cyc_h >> 10); }
}

// ---------------------------------------------------------------------------------------------------------------------
// Decoder side (ref: szd_float.c:5809-5819): last = last + 2 (t - R) * prec, or the next stored raw value when t == 0.  Run like (S) above:
// one CTA per coefficient, windows of LAT_W steps; the helpers turn the next window's codes into increments (or raw values, found by
// the ordered rank of the zero code among the zeros before it) and scatter the previous window's results to their block ids; one
// thread adds 32 increments from registers per word (a word with a raw value in it goes step by step).
template <typename T>
__global__ void __launch_bounds__(LAT_THREADS) coeff_chain_decode_kernel(const uint16_t *__restrict__ coeff_type /*[ncoef][nb]*/, const T *__restrict__ cu /*[ncoef][nb] raw values in order*/,
                                                                         uint32_t nb, uint32_t reg_count, T p0, T p1, T p2, T p3, int r0_, int r1_, int r2_, int r3_,
                                                                         const uint32_t *__restrict__ reg_list, T *__restrict__ coeff_last /*[ncoef][nb] by block id*/) {
    constexpr int W = LAT_W;
    __shared__ __align__(16) T s_val[2][W];
    __shared__ __align__(16) T s_out[2][W];
    __shared__ unsigned int s_rm[2][W / 32];              // raw-value flags of a window, one bit per step (also: their count per warp)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool helper = warp != 0 && warp != 4 && warp != 8;          // the serial thread keeps its scheduler to itself
    const int hw = warp - 1 - (warp > 4 ? 1 : 0) - (warp > 8 ? 1 : 0), h = hw * 32 + lane;
    const int e = blockIdx.x;
    const T prec = e == 0 ? p0 : e == 1 ? p1 : e == 2 ? p2 : p3;
    const int rad = e == 0 ? r0_ : e == 1 ? r1_ : e == 2 ? r2_ : r3_;
    const uint16_t *ct_e = coeff_type + (size_t)e * nb;
    const T *cu_e = cu + (size_t)e * nb;
    const uint32_t nwin = (reg_count + W - 1) / W;
    T last = (T)0;
    unsigned int zrun = 0;                                 // helpers: raw values consumed by the windows prepared so far
    // helpers: codes of window `win` -> s_val / s_rm of slot win & 1 (tc: this thread's code, loaded by the caller ahead of time)
    auto prepare = [&](uint32_t win, int tc, bool in) {
        const int sl = (int)(win & 1u);
        const bool z = in && tc == 0;
        const unsigned int zm = __ballot_sync(0xffffffffu, z);
        if (lane == 0) s_rm[sl][hw] = zm;
        lat_helper_bar();
        unsigned int before = 0, tot = 0;
#pragma unroll
        for (int q = 0; q < W / 32; q++) { const unsigned int c = __popc(s_rm[sl][q]); before += q < hw ? c : 0u; tot += c; }
        T v = (T)0;
        if (z) v = cu_e[zrun + before + __popc(zm & ((1u << lane) - 1u))];
        else if (in) { const T dq = (T)(2 * (tc - rad)); v = dq * prec; }
        s_val[sl][h] = v;
        zrun += tot;
    };
    int t_ahead = 1;                                       // helpers: this thread's code of the window after the next one, loaded a tick early
    if (helper) {
        const bool in = (uint32_t)h < reg_count;
        prepare(0u, in ? (int)ct_e[h] : 1, in);
        const uint32_t p1 = (uint32_t)W + (uint32_t)h;
        if (p1 < reg_count) t_ahead = (int)ct_e[p1];
    }
    __syncthreads();
    for (uint32_t w = 0; w <= nwin; w++) {
        if (warp == 0) {
            if (lane == 0 && w < nwin) {
                const int sl = (int)(w & 1u);
                const int cnt = (int)min((uint32_t)W, reg_count - w * W);
                const int nwords = (cnt + 31) >> 5;
                T a[32], b[32];
                lat_ld32(&s_val[sl][0], a);
                auto word = [&](int j, const T (&v)[32]) {
                    const unsigned int rm = s_rm[sl][j];
                    T o[32];
                    if (rm == 0u) {
#pragma unroll
                        for (int t = 0; t < 32; t++) { last = last + v[t]; o[t] = last; }
                    } else {
#pragma unroll
                        for (int t = 0; t < 32; t++) { const T sum = last + v[t]; last = ((rm >> t) & 1u) ? v[t] : sum; o[t] = last; }
                    }
#pragma unroll
                    for (int g = 0; g < 4; g++) lat_st8(&s_out[sl][32 * j + 8 * g], &o[8 * g]);
                };
                for (int j = 0; j < nwords; j += 2) {
                    if (j + 1 < nwords) lat_ld32(&s_val[sl][32 * (j + 1)], b);
                    word(j, a);
                    if (j + 1 >= nwords) break;
                    if (j + 2 < nwords) lat_ld32(&s_val[sl][32 * (j + 2)], a);
                    word(j + 1, b);
                }
            }
        } else if (helper) {
            // the next window's codes were loaded a tick ago; the ones after it are requested now
            const uint32_t pn = (w + 1) * W + (uint32_t)h;
            const bool in_n = w + 1 < nwin && pn < reg_count;
            const int tn = t_ahead;
            { const uint32_t p2 = (w + 2) * W + (uint32_t)h; t_ahead = p2 < reg_count ? (int)ct_e[p2] : 1; }
            if (w >= 1) {                                  // the previous window's results to their blocks
                const uint32_t pp = (w - 1) * W + (uint32_t)h;
                if (pp < reg_count) coeff_last[(size_t)e * nb + reg_list[pp]] = s_out[(w - 1) & 1u][h];
            }
            if (w + 1 < nwin) prepare(w + 1, tn, in_n);
        }
        __syncthreads();
    }
}

}  // namespace szb
