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
// side by side on consecutive windows (one __syncthreads per window):
//   (G) helper thread h guesses the code of step h from the closed form (double arithmetic, relative to the exact `last` that
//       entered the window before) and classes the step: SURE; PAIR when the guess is within a drift band of a midpoint or follows
//       such a step -- two candidate codes, told apart by an exact threshold D on |cur - last| that the thread finds by evaluating
//       the reference's own quotient at the floats around (2n+1) prec; FULL when neither form is safe (two pairs in a row, near the
//       code range, not finite).  Increments fl(2 dk * prec) are formed here.
//   (S) ONE thread runs the recurrence: a SURE step is one add, a PAIR step a subtraction, a compare, a select, an add and the
//       reference's check, a FULL step the reference's complete formula.  An unpredictable coefficient (new base) ends the window.
//   (V) helper thread h proves step h with the complete formula on the exact `last` that entered it: code and bits of the new
//       `last`.  A window that fails is rerun from its exact entry value with a wider band, then with every step FULL, i.e. as the
//       reference's own loop.
// The output is the reference's sequence by induction over the proved steps, not by the quality of the guess, which only decides
// the speed.
#pragma once
#include "szb_stage1.cuh"

namespace szb {

constexpr int LAT_W = 256;                    // steps per window = helper threads
constexpr int LAT_THREADS = LAT_W + 32;       // warp 0: the serial thread; warps 1..8: helpers
constexpr float LAT_BAND = 0.75f;             // drift band in units of mag eps sqrt(steps), see (G)
template <typename T> struct LatEps;
template <> struct LatEps<float> { static constexpr double eps = 1.0 / 8388608.0; };              // 2^-23: ulp(x) <= eps |x|
template <> struct LatEps<double> { static constexpr double eps = 1.0 / 4503599627370496.0; };    // 2^-52

template <typename T> struct __align__(16) LatSlot {        // one window in shared memory
    T cur[LAT_W], inc[LAT_W], incB[LAT_W], thr[LAT_W], out[LAT_W];     // out[i]: last after step i
    int q[LAT_W], qB[LAT_W];
    unsigned int hm[LAT_W / 32], xm[LAT_W / 32];                        // one bit per step: PAIR or FULL / FULL
    unsigned int wend; T exit;                                          // written by (S): steps done, last after them
    float maxmag;                                                       // written by (G): largest magnitude the window's values can have
};

__device__ __forceinline__ void lat_helper_bar() { asm volatile("bar.sync 1, %0;" ::"n"(LAT_W) : "memory"); }

// magnitude of the code the reference's formula gives for |cur - last| = d (monotone in d; ref: sz_float.c:7133-7139)
template <typename T>
__device__ __forceinline__ int lat_mag(T d, T prec, T rprec, int use_div) {
    T itv = use_div ? d / prec : d * rprec;
    itv = itv + (T)1;
    return itv < (T)65536 ? (int)(itv / (T)2) : 65536;
}

// A FULL step: the reference's complete formula; `ended`: the coefficient is unpredictable (the window ends with it).  The running
// `last` goes in and out by value (a reference parameter would put it into local memory).
template <typename T> struct LatStep { T s; int ended; };
template <typename T>
__device__ __noinline__ LatStep<T> lat_full_step(LatSlot<T> &L, int i, T prec, T rprec, int use_div, T s) {
    T nl;
    const int ct = chain_step_exact<T>(L.cur[i], s, prec, rprec, use_div, nl);
    L.q[i] = ct; L.out[i] = nl;
    LatStep<T> r; r.s = nl; r.ended = ct == 0;
    return r;
}
// A PAIR step: the code has magnitude n0 or n0 + 1, told apart by |cur - last| against the exact threshold; then the reference's check.
// When it takes its second candidate, the step behind it starts from the other lattice point: its increment and code are replaced.
template <typename T>
__device__ __forceinline__ LatStep<T> lat_pair_step(LatSlot<T> &L, int i, T prec, T s) {
    const T cur = L.cur[i], thr = L.thr[i], iA = L.inc[i], iB = L.incB[i];
    const T diff = cur - s;
    const bool big = !(TT<T>::abs(diff) < thr);
    const T v = s + (big ? iB : iA);
    const T chk = cur - v;
    const bool unp = TT<T>::abs(chk) > prec;
    if (big) {
        L.q[i] = L.qB[i];
        if (i + 1 < LAT_W) { L.inc[i + 1] = L.incB[i + 1]; L.q[i + 1] = L.qB[i + 1]; }
    }
    LatStep<T> r;
    if (unp) { L.q[i] = 0; r.s = cur; r.ended = 1; } else { r.s = v; r.ended = 0; }
    L.out[i] = r.s;
    return r;
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

// Eight steps of which at least one is PAIR or FULL (ONE copy of this code: the serial thread runs alone on its scheduler and pays for
// every instruction-cache miss).  ended: index in the group of the first unpredictable coefficient, or -1; the steps behind such a
// coefficient are computed for nothing (no exit from the unrolled steps).
template <typename T> struct LatGroup { T s; int ended; };
template <typename T>
__device__ __noinline__ LatGroup<T> lat_mixed_group(LatSlot<T> &L, int i0, unsigned int hm8, unsigned int xm8, T prec, T rprec, int use_div, T s) {
    T v[8];
    lat_ld8(&L.inc[i0], v);
    int ended = -1;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        if ((hm8 >> t) & 1u) {
            LatStep<T> r;
            if ((xm8 >> t) & 1u) r = lat_full_step<T>(L, i0 + t, prec, rprec, use_div, s);
            else {
                r = lat_pair_step<T>(L, i0 + t, prec, s);
                if (t + 1 < 8) v[t + 1] = L.inc[i0 + t + 1];       // it may have replaced the increment of the step behind it
            }
            s = r.s;
            if (r.ended && ended < 0) ended = t;
        } else { s = s + v[t]; L.out[i0 + t] = s; }
    }
    LatGroup<T> g; g.s = s; g.ended = ended;
    return g;
}

template <typename T> struct LatWin { uint32_t p; int cnt, slot, mode; T entry; bool valid, fresh; };

template <typename T>
__global__ void __launch_bounds__(LAT_THREADS) coeff_lattice_kernel(const T *__restrict__ reg_c /*[ncoef][nb], compacted*/, const uint32_t *__restrict__ reg_list,
                                                                    uint32_t nb, uint32_t reg_count, T p0, T p1, T p2, T p3, int use_div,
                                                                    uint16_t *__restrict__ coeff_type /*[ncoef][nb]*/, T *__restrict__ coeff_last /*[ncoef][nb] by block id*/,
                                                                    T *__restrict__ coeff_unpred /*[ncoef][nb]*/, ChainOut *out, int first_mode, float band_scale) {
    typedef typename TT<T>::bits_t bits_t;
    constexpr int W = LAT_W;
    __shared__ __align__(16) LatSlot<T> s_slot[2];
    __shared__ int s_k[W];
    __shared__ unsigned char s_tie[W];            // bit 0: within the band of a midpoint; bit 1: cur above the guessed lattice point
    __shared__ float s_wmag[W / 32];
    __shared__ unsigned int s_wcnt[2][W / 32];      // by tick parity: read after the tick's barrier, written again one tick later
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, h = tid - 32;
    const int e = blockIdx.x;
    const T prec = e == 0 ? p0 : e == 1 ? p1 : e == 2 ? p2 : p3;
    const T rprec = (T)1 / prec;
    const double inv2p = 1.0 / (2.0 * (double)prec);
    const T *cur_e = reg_c + (size_t)e * nb;
    uint32_t ucount = 0;
    unsigned int n_ticks = 0, n_redo = 0, n_pair = 0, n_full = 0;
    long long cyc_s = 0; const long long cyc_0 = clock64();
    LatWin<T> vw, sw, gw;
    vw.valid = false; sw.valid = false;
    gw.p = 0; gw.cnt = (int)min((uint32_t)W, reg_count); gw.slot = 0; gw.mode = first_mode; gw.entry = (T)0; gw.valid = reg_count > 0; gw.fresh = true;
    while (vw.valid || sw.valid || gw.valid) {
        bool my_fail = false;
        if (warp == 0) {
            // ---------------------------------------- (S) ----------------------------------------
            if (lane == 0 && sw.valid) {
                const long long c0 = clock64();
                LatSlot<T> &L = s_slot[sw.slot];
                T s = sw.entry;
                int wend = sw.cnt;
                const int nwords = (sw.cnt + 31) >> 5;
                T nv[8];                                             // increments of the next group, fetched one group ahead
                lat_ld8(&L.inc[0], nv);
                unsigned int hm_n = L.hm[0], xm_n = L.xm[0];
                for (int w = 0; w < nwords; w++) {
                    const unsigned int hmw = hm_n, xmw = xm_n;
                    if (w + 1 < nwords) { hm_n = L.hm[w + 1]; xm_n = L.xm[w + 1]; }
                    int ended_at = -1;
#pragma unroll
                    for (int gg = 0; gg < 4; gg++) {
                        const int i0 = 32 * w + 8 * gg;
                        T v[8];
#pragma unroll
                        for (int t = 0; t < 8; t++) v[t] = nv[t];
                        if (i0 + 8 < LAT_W) lat_ld8(&L.inc[i0 + 8], nv);
                        const unsigned int hm8 = (hmw >> (8 * gg)) & 0xffu;
                        if (hm8 == 0u) {                             // eight SURE steps (steps past the window's end add zero)
#pragma unroll
                            for (int t = 0; t < 8; t++) { s = s + v[t]; v[t] = s; }
                            lat_st8(&L.out[i0], v);
                        } else {
                            const LatGroup<T> r = lat_mixed_group<T>(L, i0, hm8, (xmw >> (8 * gg)) & 0xffu, prec, rprec, use_div, s);
                            s = r.s;
                            if (r.ended >= 0 && ended_at < 0) ended_at = 8 * gg + r.ended;
                            if (i0 + 8 < LAT_W) lat_ld8(&L.inc[i0 + 8], nv);      // a PAIR step at the group's end replaces the next increment
                        }
                    }
                    if (ended_at >= 0) { wend = 32 * w + ended_at + 1; break; }
                }
                L.wend = (unsigned int)wend;
                L.exit = L.out[wend - 1];
                cyc_s += clock64() - c0;
            }
        } else {
            // global loads of this tick first: their latency hides behind (V)
            T cur_g = (T)0; uint32_t blk_v = 0;
            if (gw.valid && h < gw.cnt) cur_g = cur_e[gw.p + h];
            if (vw.valid && h < vw.cnt) blk_v = reg_list[vw.p + h];
            // ---------------------------------------- (V) ----------------------------------------
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
                    my_fail = ct != L.q[h] || bn != bo;
                }
                zmask = __ballot_sync(0xffffffffu, h < vw.cnt && ct == 0);
                if (lane == 0) s_wcnt[n_ticks & 1u][warp - 1] = __popc(zmask);
            }
            lat_helper_bar();
            if (vw.valid && h < vw.cnt) {            // written before the verdict: a failed window is rerun and writes the same places again
                coeff_type[(size_t)e * nb + vw.p + h] = (uint16_t)ct;
                coeff_last[(size_t)e * nb + blk_v] = lout;
                if (ct == 0) {
                    unsigned int before = 0;
                    for (int w = 0; w < warp - 1; w++) before += s_wcnt[n_ticks & 1u][w];
                    coeff_unpred[(size_t)e * nb + ucount + before + __popc(zmask & ((1u << lane) - 1u))] = vcur;
                }
            }
            // ---------------------------------------- (G) ----------------------------------------
            if (gw.valid) {
                LatSlot<T> &L = s_slot[gw.slot];
                // the guess is relative to an exact `last`: the window's own entry value (fresh), or the entry value of the window
                // that (S) is running now, one window back
                const double base = (double)(gw.fresh ? gw.entry : sw.entry);
                const int m0 = gw.fresh ? 0 : sw.cnt;
                const double bandscale = gw.mode == 0 ? (double)band_scale : 16.0 * (double)band_scale;
                const double x = ((double)cur_g - base) * inv2p;
                const double kr = rint(x);
                const bool valid = h < gw.cnt && fabs(x) < 1073741824.0;            // false for NaN/inf too
                const double tiedist = 0.5 - fabs(x - kr);
                const int k = valid ? (int)kr : 0;
                L.cur[h] = cur_g; s_k[h] = k;
                {
                    float m = h < gw.cnt ? fabsf((float)cur_g) : 0.0f;
                    if (!(m <= 3.0e38f)) m = 3.0e38f;
#pragma unroll
                    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                    if (lane == 0) s_wmag[warp - 1] = m;
                }
                // the step before the window (pipelined windows only): its lattice index and midpoint status under the same base
                int kprev0 = 0; bool tie_prev0 = false, up_prev0 = false, valid_prev0 = true;
                double magp = fabs(base);
                if (!gw.fresh) {
                    magp = fmax(magp, (double)s_slot[sw.slot].maxmag);      // the drift is run up in that window too
                    const T pc = s_slot[sw.slot].cur[sw.cnt - 1];
                    const double xp = ((double)pc - base) * inv2p, kp = rint(xp);
                    valid_prev0 = fabs(xp) < 1073741824.0;
                    kprev0 = valid_prev0 ? (int)kp : 0;
                    up_prev0 = xp > kp;
                    magp = fmax(magp, fabs((double)pc));
                    // its own band needs the step before it; the widest quotient stands in
                    tie_prev0 = !(0.5 - fabs(xp - kp) > bandscale * (magp * LatEps<T>::eps) * sqrt((double)m0) * inv2p + 32770.0 * (8.0 * LatEps<T>::eps) + (fabs(xp) + 1.0) * 1e-15);
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
                // Drift of `last` after m steps from the exact base: two roundings per step (product, sum), each uniform within half an ulp
                // of a value no larger than ~mag: a random walk with sigma ~ 0.25 mag eps per step.  LAT_BAND = 0.75 is three sigma; a
                // window that leaves the band is caught by (V) and rerun with a wider one, so the band is set for speed, in lattice
                // units.  Plus the formula's own roundings (difference, quotient, +1; the reciprocal), relative to the quotient |dk|, and
                // the error of this double-precision estimate.
                const double band = bandscale * (mag * LatEps<T>::eps) * sqrt((double)(m0 + h + 1)) * inv2p + (adk + 2.0) * (8.0 * LatEps<T>::eps) + (fabs(x) + 1.0) * 1e-15;
                const bool tie = !(tiedist > band);
                const bool up = x > kr;
                s_tie[h] = (unsigned char)((tie ? 1 : 0) | (up ? 2 : 0));
                lat_helper_bar();
                // midpoint status of the two steps before this one (before a pipelined window: the step before it stands for both)
                const unsigned char tp0 = (unsigned char)((tie_prev0 ? 1 : 0) | (up_prev0 ? 2 : 0));
                const unsigned char tp = h ? s_tie[h - 1] : tp0;
                const bool tie1 = (tp & 1) != 0;
                const bool tie2 = h >= 2 ? (s_tie[h - 2] & 1) != 0 : tie_prev0;
                const bool valid_prev = h ? (fabs(((double)L.cur[h - 1] - base) * inv2p) < 1073741824.0) : valid_prev0;
                const bool in = h < gw.cnt;
                // FULL: nothing safe to prepare.  A step at a midpoint whose predecessor is at one too; the step behind such a step.
                bool full = in && (gw.mode >= 2 || !valid || !valid_prev || adk >= 32760.0 || (tie1 && (tie || tie2 || h == 0)));
                bool pair = in && !full && tie;                      // at a midpoint, predecessor sure: two candidate codes, one threshold
                const bool follow = in && !full && !tie && tie1;     // behind a PAIR step: two candidate increments, chosen by that step
                int qA = (int)dk + 32768, qB = 0; T incA = (T)0, incB = (T)0, thr = (T)0;
                if (pair) {
                    // the candidates: magnitudes n0 (A) and n0 + 1 (B) with one sign
                    long long dA = dk, dB = dk + (up ? 1 : -1);       // the lattice point on the other side of the midpoint
                    long long aA = dA < 0 ? -dA : dA, aB = dB < 0 ? -dB : dB;
                    if (aA > aB) { long long t = dA; dA = dB; dB = t; t = aA; aA = aB; aB = t; }
                    if (aB >= 32760) { full = true; pair = false; }
                    else {
                        // exact threshold: the smallest float d with lat_mag(d) >= aB, looked for among the floats around (2 aB - 1) prec
                        const T d0 = (T)((double)(2 * aB - 1) * (double)prec);
                        bits_t b0; memcpy(&b0, &d0, sizeof(T));
                        int first = 99;
#pragma unroll
                        for (int o = 4; o >= -4; o--) {
                            const bits_t b = b0 + (bits_t)(long long)o; T d; memcpy(&d, &b, sizeof(T));
                            if (lat_mag<T>(d, prec, rprec, use_div) >= (int)aB) first = o;
                        }
                        bool mono_ok = first != 99;
                        {   // below the found float the magnitude must be smaller (monotone function: one more evaluation settles it)
                            const bits_t b = b0 + (bits_t)(long long)(first - 1); T d; memcpy(&d, &b, sizeof(T));
                            if (mono_ok && lat_mag<T>(d, prec, rprec, use_div) >= (int)aB) mono_ok = false;
                        }
                        if (!mono_ok || !(d0 > (T)0) || b0 < (bits_t)16) { full = true; pair = false; }
                        else {
                            const bits_t b = b0 + (bits_t)(long long)first; memcpy(&thr, &b, sizeof(T));
                            qA = (int)dA + 32768; qB = (int)dB + 32768;
                            const T dqA = (T)(2 * (int)dA), dqB = (T)(2 * (int)dB);
                            incA = dqA * prec; incB = dqB * prec;
                        }
                    }
                }
                if (follow) {
                    // the predecessor's candidates in ITS order (A = smaller magnitude of its own code): this step's increment for either
                    const long long kpp = h >= 2 ? (long long)s_k[h - 2] : (long long)kprev0;
                    long long kA = kprev, kB = (long long)kprev + ((tp & 2) ? 1 : -1);
                    const long long pA = kA - kpp, pB = kB - kpp;
                    if ((pA < 0 ? -pA : pA) > (pB < 0 ? -pB : pB)) { const long long t = kA; kA = kB; kB = t; }
                    const long long fA = (long long)k - kA, fB = (long long)k - kB;
                    qA = (int)fA + 32768; qB = (int)fB + 32768;
                    const T dqA = (T)(2 * (int)fA), dqB = (T)(2 * (int)fB);
                    incA = dqA * prec; incB = dqB * prec;
                }
                if (in && !pair && !full && !follow) { const T dq = (T)(2 * (int)dk); incA = dq * prec; }
                // a midpoint step that could not be prepared as a PAIR is FULL, and then so is the step behind it
                if (in && tie && !pair) s_tie[h] = (unsigned char)(s_tie[h] | 4);
                lat_helper_bar();
                if (follow && h && (s_tie[h - 1] & 4)) full = true;
                if (full) { qA = 0; incA = (T)0; }
                const bool handler = pair || full;
                const unsigned int hm = __ballot_sync(0xffffffffu, handler), xm = __ballot_sync(0xffffffffu, full);
                if (lane == 0) { L.hm[warp - 1] = hm; L.xm[warp - 1] = xm; }
                L.inc[h] = in ? incA : (T)0; L.incB[h] = incB; L.thr[h] = thr; L.q[h] = qA; L.qB[h] = qB;
                n_pair += __popc(hm & ~xm); n_full += __popc(xm);
            }
        }
        const int fail = __syncthreads_or(my_fail ? 1 : 0);
        const unsigned int par = n_ticks & 1u;
        n_ticks++;
        // ---- what the next tick does (the same decision in every thread) ----
        if (vw.valid) {
            if (fail) {                                    // rerun the failed window from its exact entry value; everything after it is void
                if (vw.mode >= 2) { if (tid == 0) out->err = 1u; return; }     // the complete formula disagrees with itself: cannot happen
                gw.p = vw.p; gw.cnt = (int)min((uint32_t)W, reg_count - vw.p); gw.slot = vw.slot; gw.mode = vw.mode + 1; gw.entry = vw.entry; gw.valid = true; gw.fresh = true;
                vw.valid = false; sw.valid = false; n_redo++;
                continue;
            }
            unsigned int tot = 0;
#pragma unroll
            for (int w = 0; w < W / 32; w++) tot += s_wcnt[par][w];
            ucount += tot;
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
    if (tid == 0) {
        out->unpred_count[e] = ucount;
        if (e == (int)gridDim.x - 1) {
            out->passes = n_ticks; out->slow_chunks = n_redo; out->kcyc_serial = (unsigned int)(cyc_s >> 10);
            out->kcyc_total = (unsigned int)((clock64() - cyc_0) >> 10);
        }
    }
    if (e == (int)gridDim.x - 1 && warp >= 1 && lane == 0) { atomicAdd(&out->kcyc_helper, n_pair); atomicAdd(&out->n_full, n_full); }
}

}  // namespace szb
