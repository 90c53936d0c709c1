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

// A lane that holds no removal has ek = 0 (never larger than a child: frequencies are >= 1) and its hole at the unused slot 0: it
// runs the same instructions as the others and changes nothing.
struct HpLane {
    unsigned ha, pa;        // address of the hole, of its two children
    uint32_t ek, en;        // the element that sinks
    uint32_t t0;            // tick count at launch: a sinking element goes down one level per tick, so its hole's level is ticks - t0
};
constexpr uint32_t HP_INF = 0xffffffffu;

// One level for every removal in flight, without a branch or a predicate: slots beyond the heap's current end hold HP_INF (every
// removal marks the slot it empties), and a children pair beyond the array is read from `capa`, a pair of HP_INF entries behind it,
// so "no such child" is an ordinary comparison.  The dependent chain from one load to the next is: which child is smaller -> its
// address -> its children's address -> clamp; whether the element goes on sinking is not waited for (a lane that stops has ek = 0
// from then on and whatever it loads is ignored).
// CHK: the tick also answers "does some removal have its hole on the ancestor chain of slot x (x included) after this tick?"
// (cx = level of x minus the tick count after this tick: a sinking element goes down one level per tick, so its hole's level is
// ticks - t0 and the chain's node at that level is x >> (cx + t0); a lane without a removal has its hole at slot 0, which the
// max() keeps out of the comparison).  Asked inside the tick, the vote's latency is hidden behind the tick's own tail.
template <bool CHK>
static __device__ __forceinline__ bool hp_tick_finish(const unsigned base, const unsigned capa, HpLane &S, const uint4 v, const uint32_t x = 0u, const int cx = 0) {
    const bool right = v.z < v.x;                               // right child only if STRICTLY smaller (Huffman.c:98)
    const uint32_t kc = right ? v.z : v.x, nc = right ? v.w : v.y;
    const unsigned hc = S.pa + (right ? 8u : 0u);
    const bool move = S.ek > kc;                                // sinks only while STRICTLY larger (Huffman.c:101)
    hp_st2(S.ha, move ? kc : S.ek, move ? nc : S.en);
    S.ha = move ? hc : base;
    S.pa = min(2u * hc - base, capa);
    S.ek = move ? S.ek : 0u;
    bool busy = false;
    if (CHK) busy = __any_sync(0xffffffffu, base + 8u * max(x >> ((cx + (int)S.t0) & 31), 1u) == S.ha);
    return busy;
}
template <bool CHK>
static __device__ __forceinline__ bool hp_tick(const unsigned base, const unsigned capa, HpLane &S, const uint32_t x = 0u, const int cx = 0) {
    const uint4 v = hp_ld4(S.pa);
    return hp_tick_finish<CHK>(base, capa, S, v, x, cx);
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
//   base: shared address of the heap, 16-byte aligned, 8 (leaves + 4) bytes; lfreq: leaf frequencies in increasing symbol order (global memory)
//   up: what the passes after the heap need, per node (global memory, written as the merges happen, never read here):
//       {parent:16 | 1 + nodes of the left sibling's subtree if this node is a right child, else 1 : 16}; the root keeps HP_ROOT.
//   Node ids: leaves 0 .. leaves-1 in symbol order, internal nodes in merge order; fewer than 32768 leaves (ids and subtree sizes fit
//   16 bits: the node word of a heap entry is {id:16 | nodes in its subtree:16}, so a merge knows its subtree size without a lookup).
constexpr uint32_t HP_ROOT = 0xffffffffu;
static __device__ __noinline__ uint32_t hp_build(unsigned base, const uint32_t leaves, const uint32_t *__restrict__ lfreq, uint32_t *up,
                                                 uint32_t *loga, uint32_t *logb, uint32_t *prof_out, const long long tk1) {
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
                if (mine && ((sift >> lane) & 1u) == 0u) hp_st2(base + 8u * (unsigned)(q0 + lane), cur, (k0 + (uint32_t)lane) | 0x10000u);
                __syncwarp();
            }
            while (sift) {
                const int j = __ffs((int)sift) - 1; sift &= sift - 1u;
                hp_insert_warp(base, q0 + j, __shfl_sync(0xffffffffu, cur, j), (k0 + (uint32_t)j) | 0x10000u, lane);
            }
            qend = q0 + cnt;
            cur = nxt;
        }
    }
    if (lane == 0) prof_out[5] = (uint32_t)((clock64() - tk1) >> 10);
    // ---- merges ----
    const uint32_t cap = (leaves + 2u) & ~1u;                     // an even slot behind the heap: slots cap, cap + 1 = the HP_INF pair
    const unsigned capa = base + 8u * cap;
    if (lane < 3) hp_st2(base + 8u * (leaves + 1u + (uint32_t)lane), HP_INF, 0u);      // leaves + 1 .. cap + 1 (at most three slots)
    __syncwarp();
    HpLane S; S.ha = base; S.pa = capa; S.ek = 0u; S.en = 0u; S.t0 = 0u;
    uint32_t m = leaves, k = 0, ticks = 0;                        // m: elements in the heap (the last one travels in registers, not in its slot)
#ifdef HP_COUNT
    uint32_t n_wait_e = 0, n_wait_c = 0;
#define HP_CNT(x) (x)++
#else
#define HP_CNT(x)
#endif
    const bool piped = m >= HP_TAIL;
    uint2 pend = make_uint2(0u, 0u);                              // the merged node whose insertion is still to be done when the loop ends
    uint32_t nid1 = leaves + 0x10000u, ra = 0u, rb = 0u;        // next node id (+ the 1 of its subtree size); lane's pending records
    uint32_t nlog = 0u;                                           // merges logged (the last HP_TAIL ones write their `up` records directly)
    if (piped) {
        // first removal of the first merge: takes the root `a`, the last element sinks from the root
        uint2 a = hp_ld2(base + 8u);
        { const uint2 e1 = hp_ld2(base + 8u * m); __syncwarp(); hp_st2(base + 8u * m, HP_INF, 0u);
          if (lane == 0) { S.ha = base + 8u; S.pa = base + 16u; S.ek = e1.x; S.en = e1.y; S.t0 = ticks; } }
        __syncwarp();
        hp_tick<false>(base, capa, S); ticks++;
        for (;; m--, k++) {
            const int la = (int)((2u * k) & 31u);
            // second removal: its element is the one in slot x = m - 1, once no older removal can still end there (hole on x's chain).
            const uint32_t x = m - 1u;
            const int lvx = 31 - __clz(x);
            ticks++;
            bool busy = hp_tick<true>(base, capa, S, x, lvx - (int)ticks);             // the first removal's second level: level 1 is final again
            while (busy) { ticks++; busy = hp_tick<true>(base, capa, S, x, lvx - (int)ticks); HP_CNT(n_wait_e); }
            // Launch and first level in one round trip: the root, slot x and the lanes' children pairs are loaded together.
            if (lane == la + 1) { S.ha = base + 8u; S.pa = base + 16u; S.t0 = ticks; }
            uint4 v = hp_ld4(S.pa);
            const uint2 b = hp_ld2(base + 8u), e2 = hp_ld2(base + 8u * x);
            hp_st2(base + 8u * x, HP_INF, 0u);                    // the heap ends one slot earlier (every lane stores the same)
            if (lane == la + 1) { S.ek = e2.x; S.en = e2.y; }
            hp_tick_finish<false>(base, capa, S, v); ticks++;
            ticks++;
            busy = hp_tick<true>(base, capa, S, x, lvx - (int)ticks);
            // the merged node: first removed -> right, second removed -> left (gcc evaluates the arguments of Huffman.c:181 right to left)
            const uint32_t kn = a.x + b.x;
            const uint32_t nn = (a.y & 0xffff0000u) + (b.y & 0xffff0000u) + nid1;      // {id, nodes of both subtrees + 1}
            nid1++;
            // the children of this merge are kept by lane k mod 32 and logged 32 merges at a time, two coalesced stores (scattered
            // `up` records from here cost the lone warp several hundred cycles per group when other kernels load the memory system);
            // hp_apply_log turns the log into `up` records afterwards, with all threads
            if (lane == (int)(k & 31u)) { ra = a.y; rb = b.y; }
            if ((k & 31u) == 31u) { loga[k - 31u + (uint32_t)lane] = ra; logb[k - 31u + (uint32_t)lane] = rb; }
            if (m - 1u < HP_TAIL) { pend = make_uint2(kn, nn); break; }
            // its insertion at slot x (Huffman.c:76-85), once the ancestor chain of x is nobody's, in one round trip with the launch and
            // first level of the next merge's first removal, which takes what the insertion would leave in slot x out of it again:
            // mostly the new node itself (its parent is not larger), which therefore never touches the heap.
            while (busy) { ticks++; busy = hp_tick<true>(base, capa, S, x, lvx - (int)ticks); HP_CNT(n_wait_c); }
            const int ln = (la + 2) & 31;
            if (lane == ln) { S.ha = base + 8u; S.pa = base + 16u; S.t0 = ticks; }
            v = hp_ld4(S.pa);
            const uint2 par = hp_ld2(base + 8u * (x >> 1));
            a = hp_ld2(base + 8u);
            uint2 e1 = make_uint2(kn, nn);
            if (par.x > kn) {
                // the lanes move the chain: lane j-1 looks at the ancestor j levels up, the larger ones (a bottom prefix) go down a level,
                // the parent's element is the one that leaves
                const int j = lane + 1;
                const bool valid = j <= lvx;                      // ancestors: x >> 1 .. x >> lvx (= the root)
                uint2 w = make_uint2(0u, 0u);
                if (valid) w = hp_ld2(base + 8u * (x >> j));
                const int cnt = __popc(__ballot_sync(0xffffffffu, valid && w.x > kn));
                __syncwarp();
                if (j >= 2 && j <= cnt) hp_st2(base + 8u * (x >> (j - 1)), w.x, w.y);
                if (lane == 0) hp_st2(base + 8u * (x >> cnt), kn, nn);
                e1 = par;
                __syncwarp();
                v = hp_ld4(S.pa);                                 // the top of the heap may have changed
                a = hp_ld2(base + 8u);
            }
            if (lane == ln) { S.ek = e1.x; S.en = e1.y; }
            hp_tick_finish<false>(base, capa, S, v); ticks++;
        }
        m--; k++;                                                 // the merge whose node is still in `pend`
        if ((k & 31u) != 0u && (uint32_t)lane < (k & 31u)) { loga[(k & ~31u) + (uint32_t)lane] = ra; logb[(k & ~31u) + (uint32_t)lane] = rb; }   // the last, incomplete group
        nlog = k;
    }
    while (__any_sync(0xffffffffu, S.ek != 0u)) hp_tick<false>(base, capa, S);
    // ---- the last few merges, one operation after the other (lane 0) ----
    uint32_t rootn = 0u;
    if (lane == 0) {
        // m elements; if the loop above ran, its last merged node `pend` still has to be inserted (then the heap holds m - 1 in slots 1 .. m-1)
        uint32_t qend = piped ? m : m + 1u, nid = leaves + k;
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
        auto ins = [&](const uint2 e) {
            uint32_t i = qend++, jj;
            while ((jj = i >> 1)) { const uint2 pj = ld(jj); if (pj.x <= e.x) break; st(i, pj); i = jj; }
            st(i, e);
        };
        if (piped) ins(pend);
        while (qend > 2u) {
            const uint2 rb = rem(), ra = rem();                  // first removed -> right, second -> left
            const uint32_t sl = ra.y >> 16;
            up[ra.y & 0xffffu] = nid | 0x10000u; up[rb.y & 0xffffu] = nid | ((sl + 1u) << 16);
            ins(make_uint2(ra.x + rb.x, nid | ((sl + (rb.y >> 16) + 1u) << 16)));
            nid++;
        }
        rootn = ld(1u).y & 0xffffu;
        up[rootn] = HP_ROOT;
    }
    rootn = __shfl_sync(0xffffffffu, rootn, 0);
    if (lane == 0) prof_out[6] = (uint32_t)((clock64() - tk1) >> 10);
#ifdef HP_COUNT
    if (lane == 0) { prof_out[8] = ticks; prof_out[9] = n_wait_e; prof_out[10] = n_wait_c; }
#endif
    return rootn | (nlog << 16);                                  // both below 65536 (fewer than 32768 leaves)
}

// merge k of the log made node leaves + k out of a (left; second removed) and b (right): the children's {parent, pre-order step} records
__device__ __forceinline__ void hp_apply_log(uint32_t *up, const uint32_t *loga, const uint32_t *logb, uint32_t leaves, uint32_t k) {
    const uint32_t ra = loga[k], rb = logb[k], nid = leaves + k;
    up[rb & 0xffffu] = nid | 0x10000u; up[ra & 0xffffu] = nid | ((rb & 0xffff0000u) + 0x10000u);
}

}  // namespace szb
