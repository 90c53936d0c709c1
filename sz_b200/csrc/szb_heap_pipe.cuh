// szb_heap_pipe.cuh -- the reference's Huffman heap (ref: sz/src/Huffman.c:76-114 qinsert / qremove, :165-185 init) replayed by ONE
// warp with several heap operations in flight, one per lane.
//
// The tree the reference builds depends on the exact array layout of its binary heap (ties), so every qremove / qinsert has to be
// replayed with its exact effect.  A removal is a chain of dependent steps, one per heap level (which child is smaller decides where
// the next step looks), and a lone warp pays the full shared-memory and ALU latency for each of them.  But the steps of DIFFERENT
// operations are independent as long as a younger operation only reads heap levels the older ones are done with:
//
//   * a removal ("sift": the element taken from the last slot sinks from the root) reads level l+1 and writes level l in the step
//     that has its hole at level l; it moves down one level per step.  A removal launched two steps after the previous one therefore
//     only ever sees final values: the older hole is always two levels further down.
//   * the insertion of a merged node is bottom-up in the reference: the ancestors of the new last slot that are STRICTLY larger move
//     down one level (a bottom prefix of the ancestor chain, the heap being ordered).  It is done in one go by the whole warp (lane j
//     looks at the ancestor j levels up, one vote counts the movers), as soon as no removal in flight has its hole on that chain:
//     then the chain is final, and no older removal will read it either (the children of an off-chain position are off-chain).
//   * the element the insertion would put into the last slot (the parent's element if anything moves, else the new node) is exactly
//     what the next merge's first removal takes out of that slot again, so it is handed over in registers, together with the new root.
//
// One "tick" = every lane that holds a removal advances it by one level (one 128-bit load of the two children, compare, select, one
// 64-bit store).  Per merge: first removal, two ticks, second removal (waits while an older hole is an ancestor of the slot its
// element comes from: that removal could still put its own element there), two ticks, insertion (waits while the chain is busy).
// Measured on recorded coefficient histograms (tools/heap_pipe_sim.cpp models exactly this schedule and checks it against the
// sequential heap): 5.3 ticks per merge, against two full-depth sifts (about 2 x 12 levels) one after the other.
//
// Heap entries are {frequency, node id} pairs, 8 bytes, slot i at base + 8 i (slot 0 unused), children of i = the aligned 16 bytes at
// base + 16 i.  All addresses are 32-bit shared-window addresses.
#pragma once
#include <stdint.h>

namespace szb {

constexpr uint32_t HP_TAIL = 16;          // the last merges (heap smaller than this) are done one after the other

static __device__ __forceinline__ uint4 hp_ld4(unsigned a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
static __device__ __forceinline__ uint2 hp_ld2(unsigned a) { uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v; }
static __device__ __forceinline__ void hp_st2(unsigned a, uint32_t x, uint32_t y) { asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory"); }

// h is x or one of its ancestors (1-based heap positions)
static __device__ __forceinline__ bool hp_anc(uint32_t h, uint32_t x) { const int d = __clz(h) - __clz(x); return d >= 0 && (x >> (d & 31)) == h; }

struct HpLane {
    unsigned ha, pa, ba;    // address of the hole, of its two children, of the last slot this removal may use
    uint32_t ek, en;        // the element that sinks
    bool act;               // the removal is still sinking
    bool in;                // the children's slot pair exists (address not beyond the last slot); kept one step ahead, see hp_tick
};
static __device__ __forceinline__ void hp_launch(HpLane &S, const unsigned base, const uint32_t last, const uint2 e) {
    S.act = true; S.ha = base + 8u; S.pa = base + 16u; S.ba = base + 8u * last; S.in = 2u <= last; S.ek = e.x; S.en = e.y;
}

// One level for every removal in flight.  The dependent chain from one load to the next is: which child is smaller -> select the
// next pair address.  Everything else is kept off it: the next pair address of BOTH children and whether those pairs exist are
// computed while the load is in flight, and the load is issued for every lane that was sinking before this step (the comparison
// with the sinking element, which decides whether it goes on, is not waited for: a lane that stops has loaded a pair it ignores).
static __device__ __forceinline__ void hp_tick(const unsigned base, HpLane &S) {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (S.in) v = hp_ld4(S.pa);
    const unsigned pl = 2u * S.pa - base;                       // children pair of the left child; the right child's is 16 bytes on
    const bool in_l = pl <= S.ba, in_r = pl + 16u <= S.ba;
    const bool right = S.pa + 8u <= S.ba && v.z < v.x;          // right child only if STRICTLY smaller (Huffman.c:98)
    const uint32_t kc = right ? v.z : v.x, nc = right ? v.w : v.y;
    const bool move = S.act && S.in && S.ek > kc;               // sinks only while STRICTLY larger (Huffman.c:101)
    if (S.act) hp_st2(S.ha, move ? kc : S.ek, move ? nc : S.en);
    if (move) S.ha = S.pa + (right ? 8u : 0u);
    S.in = S.act && (right ? in_r : in_l);
    S.pa = right ? pl + 16u : pl;
    S.act = move;
    __syncwarp();
}

// leaf phase: bottom-up insertion by the lanes (lane k looks at the ancestor k levels above the new slot q0)
static __device__ __forceinline__ void hp_insert_warp(const unsigned base, const int q0, const uint32_t freq, const uint32_t node, const int lane) {
    const int p = q0 >> lane;
    const bool anc = lane >= 1 && p >= 1;
    uint2 v = make_uint2(0u, 0u);
    if (anc) v = hp_ld2(base + 8u * (unsigned)p);
    const unsigned mv = __ballot_sync(0xffffffffu, anc && v.x > freq);
    const int fin = q0 >> __popc(mv);
    if ((mv >> lane) & 1u) hp_st2(base + 8u * (unsigned)(q0 >> (lane - 1)), v.x, v.y);
    if (lane == 0) hp_st2(base + 8u * (unsigned)fin, freq, node);
    __syncwarp();
}

// Leaf inserts, merges, and the root: the whole heap part of one tree, run by ONE warp (all 32 lanes call it).
//   base: shared address of the heap, 8 (leaves + 2) bytes; lfreq: leaf frequencies in increasing symbol order (global memory)
//   n_left / n_right: children of the internal nodes (global memory), node ids: leaves 0 .. leaves-1, internal nodes in merge order
static __device__ __noinline__ uint32_t hp_build(unsigned base, const uint32_t leaves, const uint32_t *__restrict__ lfreq, uint32_t *n_left, uint32_t *n_right,
                                                 uint32_t *prof_out, const long long tk1) {
    int lane = threadIdx.x & 31;
    asm volatile("mov.u32 %0, %0;" : "+r"(lane));
    asm volatile("mov.u32 %0, %0;" : "+r"(base));
    // ---- leaves, in increasing symbol order.  32 at a time: a leaf that is not smaller than its parent stays where it is appended,
    //      and the sifts of the other leaves of its group only LOWER the values on its ancestors, so one test against the parents as
    //      they are before the group settles it for the whole group; the rest go one by one, in order ----
    {
        int qend = 1;
        uint32_t cur = (uint32_t)lane < leaves ? lfreq[lane] : 0u;
        for (uint32_t k0 = 0; k0 < leaves; k0 += 32) {
            const uint32_t nxt = k0 + 32u + (uint32_t)lane < leaves ? lfreq[k0 + 32u + lane] : 0u;
            const int cnt = (int)min(32u, leaves - k0);
            const int q0 = qend;
            unsigned sift = cnt == 32 ? 0xffffffffu : (1u << cnt) - 1u;
            if (q0 >= 32) {                                       // the parents of the group's slots all lie below the group
                const bool mine = lane < cnt;
                const uint32_t fp = mine ? hp_ld2(base + 8u * (unsigned)((q0 + lane) >> 1)).x : 0u;
                sift = __ballot_sync(0xffffffffu, mine && fp > cur);
                if (mine && ((sift >> lane) & 1u) == 0u) hp_st2(base + 8u * (unsigned)(q0 + lane), cur, k0 + (uint32_t)lane);
                __syncwarp();
            }
            while (sift) {
                const int j = __ffs((int)sift) - 1; sift &= sift - 1u;
                hp_insert_warp(base, q0 + j, __shfl_sync(0xffffffffu, cur, j), k0 + (uint32_t)j, lane);
            }
            qend = q0 + cnt;
            cur = nxt;
        }
    }
    if (lane == 0) prof_out[5] = (uint32_t)((clock64() - tk1) >> 10);
    // ---- merges ----
    HpLane S; S.act = false; S.in = false; S.ha = base + 8u; S.pa = base + 16u; S.ba = base; S.ek = 0u; S.en = 0u;
    uint32_t m = leaves, k = 0;                                   // m: elements in the heap; the last one travels in e1, not in its slot
    uint2 e1 = make_uint2(0u, 0u), a = make_uint2(0u, 0u);
    const bool piped = m >= HP_TAIL;
    if (piped) { e1 = hp_ld2(base + 8u * m); a = hp_ld2(base + 8u); }
    for (; m >= HP_TAIL; m--, k++) {
        const int la = (int)((2u * k) & 31u);
        // first removal: takes the root `a`, the last element e1 sinks from the root among slots 1 .. m-1
        if (lane == la) hp_launch(S, base, m - 1u, e1);
        hp_tick(base, S);
        hp_tick(base, S);
        // second removal: its element is the one in slot m-1, once no older removal can still end there
        const uint32_t x = m - 1u;
        while (__ballot_sync(0xffffffffu, S.act && hp_anc((S.ha - base) >> 3, x))) hp_tick(base, S);
        const uint2 b = hp_ld2(base + 8u), e2 = hp_ld2(base + 8u * x);
        if (lane == la + 1) hp_launch(S, base, m - 2u, e2);
        hp_tick(base, S);
        hp_tick(base, S);
        // the merged node: first removed -> right, second removed -> left (gcc evaluates the arguments of Huffman.c:181 right to left)
        const uint32_t nn = leaves + k, kn = a.x + b.x;
        if (lane == 0) { n_left[nn] = b.y; n_right[nn] = a.y; }
        // its insertion at slot x (Huffman.c:76-85), by the lanes, once the ancestor chain of x is nobody's
        while (__ballot_sync(0xffffffffu, S.act && hp_anc((S.ha - base) >> 3, x))) hp_tick(base, S);
        const int d = 31 - __clz(x);                              // ancestors: x >> 1 .. x >> d (= the root)
        const int j = lane + 1;
        const bool valid = j <= d;
        uint2 v = make_uint2(0u, 0u);
        if (valid) v = hp_ld2(base + 8u * (x >> j));
        const uint2 root = hp_ld2(base + 8u), par = hp_ld2(base + 8u * (x >> 1));
        const int cnt = __popc(__ballot_sync(0xffffffffu, valid && v.x > kn));     // the ancestors 1 .. cnt levels up move down one level
        if (j >= 2 && j <= cnt) hp_st2(base + 8u * (x >> (j - 1)), v.x, v.y);
        if (lane == 0 && cnt >= 1) hp_st2(base + 8u * (x >> cnt), kn, nn);
        // what would now sit in slot x is what the next first removal takes from there: handed over, never stored
        e1 = cnt >= 1 ? par : make_uint2(kn, nn);
        a = cnt == d ? make_uint2(kn, nn) : root;
        __syncwarp();
    }
    while (__ballot_sync(0xffffffffu, S.act)) hp_tick(base, S);
    // ---- the last few merges, one operation after the other (lane 0) ----
    uint32_t rootn = 0u;
    if (lane == 0) {
        if (piped) hp_st2(base + 8u * m, e1.x, e1.y);
        uint32_t qend = m + 1u, nn = leaves + k;
        auto ld = [&](uint32_t i) { return hp_ld2(base + 8u * i); };
        auto st = [&](uint32_t i, uint2 e) { hp_st2(base + 8u * i, e.x, e.y); };
        auto rem = [&]() {
            const uint2 top = ld(1u); qend--;
            const uint2 mm = ld(qend);
            uint32_t i = 1u, l;
            while ((l = 2u * i) < qend) {
                uint2 c = ld(l);
                if (l + 1u < qend) { const uint2 cr = ld(l + 1u); if (cr.x < c.x) { c = cr; l++; } }
                if (mm.x > c.x) { st(i, c); i = l; } else break;
            }
            st(i, mm);
            return top;
        };
        while (qend > 2u) {
            const uint2 rb = rem(), ra = rem();
            n_left[nn] = ra.y; n_right[nn] = rb.y;
            const uint2 e = make_uint2(ra.x + rb.x, nn);
            uint32_t i = qend++, jj;
            while ((jj = i >> 1)) { const uint2 pj = ld(jj); if (pj.x <= e.x) break; st(i, pj); i = jj; }
            st(i, e);
            nn++;
        }
        rootn = ld(1u).y;
    }
    rootn = __shfl_sync(0xffffffffu, rootn, 0);
    if (lane == 0) prof_out[6] = (uint32_t)((clock64() - tk1) >> 10);
    return rootn;
}

}  // namespace szb
