// szb_huffman.cuh -- quant-code histogram, exact SZ Huffman tree (heap emulation), code table + tree serialisation,
// encode (per-chunk bit lengths + prefix sum + bit scatter) and the marker-less parallel decode.
// ref: sz/src/Huffman.c:17-185 (tree), :122-157 (codes), :205-308 (encode), :310-343 (decode), :443-585 (serialise),
//      :656-788 (reconstruct); call sites sz/src/sz_float.c:7379-7471.
#pragma once
#include "szb_common.cuh"
#include "szb_heap_pipe.cuh"

namespace szb {

// =====================================================================================================================
// Block-major <-> raster index iterator.  The reference emits type[] / unpredictable data "block-major": blocks in
// (b0,b1,b2) raster order, points inside a block in (ii,jj,kk) raster order (ref: sz_float.c:7064, :7359).  Our code
// array lives in RASTER order (the wavefront kernels want coalesced rows), so the encoder/decoder walk block-major
// order and translate to raster addresses incrementally.
// =====================================================================================================================
struct BMIter {
    uint32_t b0, b1, b2, ii, jj, kk, sx, sy, sz;
    uint64_t raster;
    __device__ __forceinline__ static void split_div(uint64_t &rem, uint64_t unit, const Axis &a, uint32_t &b, uint32_t &edge) {
        uint64_t cut = (uint64_t)a.split * a.early * unit;
        if (rem < cut) { uint64_t u = (uint64_t)a.early * unit; b = (uint32_t)(rem / u); rem -= (uint64_t)b * u; edge = a.early; }
        else { uint64_t r2 = rem - cut; uint64_t u = (uint64_t)a.late * unit; uint32_t bb = (uint32_t)(r2 / u); rem = r2 - (uint64_t)bb * u; b = a.split + bb; edge = a.late; }
    }
    __device__ __forceinline__ void set_raster(const Geom &g) {
        raster = ((uint64_t)(g.ax[0].off(b0) + ii) * g.ax[1].n + (g.ax[1].off(b1) + jj)) * g.ax[2].n + g.ax[2].off(b2) + kk;
    }
    __device__ __forceinline__ void init(const Geom &g, uint64_t p) {
        uint64_t rem = p;
        split_div(rem, g.s0, g.ax[0], b0, sx);
        split_div(rem, (uint64_t)sx * g.ax[2].n, g.ax[1], b1, sy);
        split_div(rem, (uint64_t)sx * sy, g.ax[2], b2, sz);
        uint64_t plane = (uint64_t)sy * sz;
        ii = (uint32_t)(rem / plane); rem -= (uint64_t)ii * plane;
        jj = (uint32_t)(rem / sz); kk = (uint32_t)(rem - (uint64_t)jj * sz);
        set_raster(g);
    }
    __device__ __forceinline__ void next(const Geom &g) {
        kk++; raster++;
        if (kk == sz) {
            kk = 0; jj++; raster += g.ax[2].n - sz;
            if (jj == sy) {
                jj = 0; ii++; raster += g.s0 - (uint64_t)sy * g.ax[2].n;
                if (ii == sx) {
                    ii = 0; b2++;
                    if (b2 == g.ax[2].nb) { b2 = 0; b1++; if (b1 == g.ax[1].nb) { b1 = 0; b0++; } }
                    if (b0 < g.ax[0].nb) { sx = g.ax[0].edge(b0); sy = g.ax[1].edge(b1); sz = g.ax[2].edge(b2); set_raster(g); }
                }
            }
        }
    }
};

// =====================================================================================================================
// histogram of u16 symbols (order-independent): shared-memory privatisation when the bin count fits, 128-bit loads
// =====================================================================================================================
__global__ void __launch_bounds__(512) hist_kernel(const uint16_t *__restrict__ s, uint64_t n, uint32_t nbins, uint32_t *g_hist, int use_smem) {
    extern __shared__ uint32_t s_hist[];
    if (use_smem) { for (uint32_t i = threadIdx.x; i < nbins; i += blockDim.x) s_hist[i] = 0; __syncthreads(); }
    uint32_t *h = use_smem ? s_hist : g_hist;
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (uint64_t)gridDim.x * blockDim.x;
    uint64_t head = ((16 - ((uintptr_t)s & 15)) & 15) / 2; if (head > n) head = n;
    if (tid < head) atomicAdd(&h[s[tid]], 1u);
    const uint4 *sv = reinterpret_cast<const uint4 *>(s + head);
    uint64_t nvec = (n - head) / 8;
    for (uint64_t i = tid; i < nvec; i += nthreads) {
        uint4 v = __ldg(sv + i);
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; q++) { atomicAdd(&h[w[q] & 0xffffu], 1u); atomicAdd(&h[w[q] >> 16], 1u); }
    }
    uint64_t tail0 = head + nvec * 8;
    if (tail0 + tid < n) atomicAdd(&h[s[tail0 + tid]], 1u);
    if (use_smem) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < nbins; i += blockDim.x) { uint32_t v = s_hist[i]; if (v) atomicAdd(&g_hist[i], v); }
    }
}

// =====================================================================================================================
// Huffman tree build: EXACT emulation of the reference's binary-heap construction, one CTA per tree.
//   leaves inserted in increasing symbol order; merge step as gcc compiles Huffman.c:181: b = qremove(), a = qremove(),
//   parent.left = a, parent.right = b.  The heap holds (freq, node, subtree size) entries in shared memory when it fits
//   (global-memory fallback otherwise); the serial part is done by thread 0, the code assignment and the pre-order
//   serialisation (pad_tree_*, Huffman.c:443-501) are breadth-first over the finished tree with the whole CTA.
// =====================================================================================================================
struct HuffTreeDesc {           // per-tree I/O, all pointers to global memory
    const uint32_t *hist;       // [nbins]
    uint32_t nbins;             // stateNum (symbols are < nbins)
    // workspace, capacity 2*max_leaves nodes
    uint32_t *n_left, *n_right, *n_sizeL, *n_sym;   // internal: left/right/size of left subtree; leaf: symbol in n_sym, left=right=0xffffffff
    uint32_t *n_pre;            // pre-order index
    uint64_t *n_code;           // right-aligned code bits
    uint8_t *n_len;
    uint32_t *frontier;         // [2][capacity]
    uint32_t *heap_g;           // global fallback heap: [3][capacity]
    // outputs
    uint64_t *code;             // [nbins] right-aligned bits
    uint8_t *clen;              // [nbins]
    unsigned char *tree_bytes;  // serialised tree
    uint32_t *out;              // [0]=treeByteSize [1]=nodeCount [2]=leaves [3]=error(1 = code longer than 64 bits)
    uint32_t capacity;          // node capacity
    const unsigned int *wait_flag;  // not nullptr: the kernel was launched ahead of its input; every CTA waits until this word is non-zero
                                    // (set, after a __threadfence, by the kernel that fills `hist`)
};

// The heap is two parallel arrays, frequencies f[] and node ids n[] (1-based, children of i at 2i and 2i+1).  Only the
// frequencies decide the path of a sift; the node ids just follow it.  Slots at or beyond qend always hold HEAP_FINF, so
// "child does not exist" falls out of the ordinary comparisons.
//
// The final array layout of every operation must equal the reference's, ties included, so the operations run one after the other
// and the cost is the instruction count and the dependent chain per heap level.  The single-thread form below (heap_insert /
// heap_remove / heap_build_tree) serves the trees that do not fit into shared memory: its sift-down reads the four GRANDCHILDREN
// f[4i..4i+3] one level ahead and moves the node ids one level late.  Trees whose frequencies fit into shared memory use the
// two-warp form further down (vote-based insertion, three levels per removal trip, node ids replayed by a second warp).
// (Measured slower, and dropped: a lane-per-node removal over 5-level subtrees -- votes on the dependent chain cost more than the
//  levels they save; a multi-warp pipelined heap; fully unrolled predicated sifts; grandchild prefetch on packed 64-bit entries.)
constexpr uint32_t HEAP_FINF = 0xffffffffu;
struct HeapTop { uint32_t f, n; };

// Accessors of the two arrays.  The shared-memory flavour addresses them through 32-bit shared-window addresses kept in
// registers and explicit ld.shared / st.shared: with ordinary pointers the compiler re-derives the window base inside the
// sift loop (a dozen uniform-datapath instructions per level).
struct HeapSmem {
    unsigned int fa, na;                 // shared addresses of f[0] and n[0]
    bool n16 = false;                    // node ids stored as 16-bit values (single-CTA layouts of the two-warp form, trees with fewer than 32768 leaves)
    __device__ __forceinline__ uint32_t ldf(int i) const { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(fa + 4u * (unsigned)i)); return v; }
    __device__ __forceinline__ uint32_t ldn(int i) const {
        uint32_t v;
        if (n16) { unsigned short h; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h) : "r"(na + 2u * (unsigned)i)); v = h; }
        else asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(na + 4u * (unsigned)i));
        return v;
    }
    __device__ __forceinline__ void stf(int i, uint32_t v) const { asm volatile("st.shared.u32 [%0], %1;" ::"r"(fa + 4u * (unsigned)i), "r"(v) : "memory"); }
    __device__ __forceinline__ void stn(int i, uint32_t v) const {
        if (n16) asm volatile("st.shared.u16 [%0], %1;" ::"r"(na + 2u * (unsigned)i), "h"((unsigned short)v) : "memory");
        else asm volatile("st.shared.u32 [%0], %1;" ::"r"(na + 4u * (unsigned)i), "r"(v) : "memory");
    }
    __device__ __forceinline__ uint2 ldf2(int i) const { uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(fa + 4u * (unsigned)i)); return v; }
    __device__ __forceinline__ uint4 ldf4(int i) const { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(fa + 4u * (unsigned)i)); return v; }
    // the same on byte addresses (the removal keeps the hole's ADDRESS, not its index: no index-to-address step per access)
    static __device__ __forceinline__ uint2 ld2a(unsigned a) { uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v; }
    static __device__ __forceinline__ uint4 ld4a(unsigned a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
    static __device__ __forceinline__ void sta(unsigned a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
};
struct HeapGmem {
    uint32_t *f, *n;
    __device__ __forceinline__ uint32_t ldf(int i) const { return f[i]; }
    __device__ __forceinline__ uint32_t ldn(int i) const { return n[i]; }
    __device__ __forceinline__ void stf(int i, uint32_t v) const { f[i] = v; }
    __device__ __forceinline__ void stn(int i, uint32_t v) const { n[i] = v; }
    __device__ __forceinline__ uint2 ldf2(int i) const { return *reinterpret_cast<const uint2 *>(f + i); }
    __device__ __forceinline__ uint4 ldf4(int i) const { return *reinterpret_cast<const uint4 *>(f + i); }
};

template <typename H>
__device__ __forceinline__ void heap_insert(const H &h, int &qend, uint32_t freq, uint32_t node) {
    // ref: Huffman.c:76-85 (bottom-up sift: parents with a STRICTLY larger frequency move down)
    int i = qend++, j;
    int pend_i = 0; uint32_t pend_v = 0;                        // node id whose store is one level late (no store waits for its load)
    while ((j = i >> 1)) {
        const uint32_t fj = h.ldf(j);
        if (fj <= freq) break;
        if (pend_i) h.stn(pend_i, pend_v);
        pend_v = h.ldn(j); pend_i = i;
        h.stf(i, fj); i = j;
    }
    if (pend_i) h.stn(pend_i, pend_v);
    h.stf(i, freq); h.stn(i, node);
}

template <typename H>
__device__ __forceinline__ HeapTop heap_remove(const H &h, int &qend, int cap) {
    // ref: Huffman.c:87-114.  The moved element m is kept in registers ("hole" form of the reference's swaps: identical
    // final layout).  Right child only if STRICTLY smaller; descend only while m is STRICTLY larger than that child.
    HeapTop top; top.f = h.ldf(1); top.n = h.ldn(1);
    qend--;
    const uint32_t mf = h.ldf(qend), mn = h.ldn(qend);
    h.stf(qend, HEAP_FINF);
    if (qend == 1) return top;
    const int lim = cap >> 1;                                   // positions >= lim have no child slots
    const int qmax = cap - 4;                                   // last aligned quad
    int i = 1;
    uint2 pr = h.ldf2(2);                                       // (f[2i], f[2i+1])
    uint4 qa = h.ldf4(4), qb;                                   // f[4i .. 4i+3]; beyond the array a clamped (unused) quad is read
    int pend_i = 0; uint32_t pend_v = 0;                        // node id whose store is one level late
    // two levels per trip, the prefetched quad alternating between qa and qb: a single quad variable makes the compiler
    // copy the freshly loaded registers at the loop edge, which waits for the load and puts its latency back on the chain
#define SZB_SIFT_LEVEL(QCUR, QNEXT)                                                        \
    {                                                                                      \
        const bool r = pr.y < pr.x;                                                        \
        const uint32_t fc = min(pr.x, pr.y);                                               \
        if (!(mf > fc)) break;                                                             \
        const int c = 2 * i + (r ? 1 : 0);                                                 \
        if (pend_i) h.stn(pend_i, pend_v);                                                 \
        pend_v = h.ldn(c); pend_i = i;                                                     \
        h.stf(i, fc);                                                                      \
        i = c;                                                                             \
        pr = r ? make_uint2(QCUR.z, QCUR.w) : make_uint2(QCUR.x, QCUR.y);                  \
        if (i >= lim) pr = make_uint2(HEAP_FINF, HEAP_FINF); /* no child slots: stop at the next compare (select, not branch) */ \
        QNEXT = h.ldf4(min(4 * i, qmax)); /* 4i > qmax only when the next level has no slots: never used then */ \
    }
    for (;;) {
        SZB_SIFT_LEVEL(qa, qb)
        SZB_SIFT_LEVEL(qb, qa)
    }
#undef SZB_SIFT_LEVEL
    if (pend_i) h.stn(pend_i, pend_v);
    h.stf(i, mf); h.stn(i, mn);
    return top;
}

// leaf inserts + merge loop of one tree (one thread); returns the root node
template <typename H>
__device__ __forceinline__ uint32_t heap_build_tree(const H &h, int hcap, uint32_t leaves, const uint32_t *lfreq, const HuffTreeDesc &D, long long tk1) {
    int qend = 1;
    uint32_t k = 0;
    for (; k + 4 <= leaves; k += 4) {                    // leaf frequencies are fetched four at a time, off the sift chain
        uint32_t f0 = lfreq[k], f1 = lfreq[k + 1], f2 = lfreq[k + 2], f3 = lfreq[k + 3];
        heap_insert(h, qend, f0, k); heap_insert(h, qend, f1, k + 1); heap_insert(h, qend, f2, k + 2); heap_insert(h, qend, f3, k + 3);
    }
    for (; k < leaves; k++) heap_insert(h, qend, lfreq[k], k);
    D.out[5] = (uint32_t)((clock64() - tk1) >> 10);
    uint32_t nn = leaves;
    while (qend > 2) {
        const HeapTop b = heap_remove(h, qend, hcap);    // first removed  -> right (gcc evaluates Huffman.c:181's arguments right to left)
        const HeapTop a = heap_remove(h, qend, hcap);    // second removed -> left
        D.n_left[nn] = a.n; D.n_right[nn] = b.n;
        heap_insert(h, qend, a.f + b.f, nn);
        nn++;
    }
    D.out[6] = (uint32_t)((clock64() - tk1) >> 10);
    return h.ldn(1);
}

// ---- two-warp form of the same heap (frequencies in shared memory) ------------------------------------------------------
// Only the frequencies decide what a sift does; the node ids merely follow.  Warp 0 (the producer) therefore runs the sifts on
// f[] alone and describes them to warp 1 (the consumer, another scheduler) with ONE queue entry per merge -- the final positions
// of the two removals and of the insertion; everything else (where an element is taken from or appended, the node number) the
// consumer counts along -- and the consumer replays the node-id moves on n[] and writes the children of every merged node.  The
// producer never reads n[], so the queue is one-way: single producer, single consumer, entries tagged with their sequence number.
constexpr int HQ_CAP = 512;                                   // queue entries (one 64-bit word each)
constexpr unsigned HQ_SPIN_LIMIT = 1u << 24;                  // polls before a waiting side gives up (a merge takes ~40 of them)
enum { HQ_MERGE = 0, HQ_LEAF = 1, HQ_FIN = 2 };
// entry = {low word: p1:16 | p2:16, high word: p3:16 | kind:2 | -:2 | tag:12}, written and read with ONE 64-bit shared-memory
// access (single-copy atomic), so the consumer sees either the old or the new entry, never a mixture, and no fence is needed:
// nothing else is shared between the two warps.  Positions are below 65536 in every shared-memory layout.  tag = (sequence number
// + 1) mod 4096; the ring holds 512 entries, so the entry that occupied a slot one lap earlier differs in the tag, and so does the
// all-zero initial state for the first lap.
struct HeapQueue {
    unsigned int qa;                                          // shared address of the ring
    unsigned int *done;                                       // consumer progress (generic pointer into shared memory)
    uint32_t t;                                               // producer: next sequence number
    bool lane0;                                               // every lane of the producer runs push(), lane 0 stores
    bool remote = false;                                      // the ring lives in the partner CTA of the cluster (qa is a shared::cluster address)
    bool dead = false;                                        // the consumer stopped answering (never seen; keeps a fault from hanging the GPU)
    __device__ __forceinline__ void push(uint32_t kind, uint32_t p1, uint32_t p2, uint32_t p3) {
        if ((t & 127u) == 0 && t >= HQ_CAP - 128 && !dead) {  // back-pressure, checked every 128 entries
            unsigned spins = 0;
            while (t - *(volatile unsigned int *)done > HQ_CAP - 128) { if (++spins > HQ_SPIN_LIMIT) { dead = true; break; } }
        }
        const uint32_t lo = p1 | (p2 << 16), hi = p3 | (kind << 16) | (((t + 1) & 4095u) << 20);
        unsigned long long e;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(e) : "r"(lo), "r"(hi));
        const unsigned slot = qa + 8u * (t & (HQ_CAP - 1));
        if (lane0) {
            if (remote) asm volatile("st.relaxed.cluster.shared::cluster.u64 [%0], %1;" ::"r"(slot), "l"(e) : "memory");
            else asm volatile("st.volatile.shared.u64 [%0], %1;" ::"r"(slot), "l"(e) : "memory");
        }
        t++;
    }
};
// thread-block cluster helpers (the two CTAs of a tree: rank 0 holds the frequencies, rank 1 the node ids when both do not fit one SM)
__device__ __forceinline__ unsigned cluster_size() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned cluster_addr(unsigned saddr, unsigned rank) { unsigned r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// The producer is a WARP whose 32 lanes run the same code on the same values (the removals are simply executed redundantly: same
// addresses, same data).  The lanes serve the insertion: the positions an inserted element can pass are the ancestors of the slot
// it is appended at, a chain known up front, so lane k fetches the ancestor k levels up, one vote finds the first ancestor that
// stays (frequency <= the new one, Huffman.c:76-85), and the lanes below it store their value one level down.  Returns the final
// position.
// WITH_N: the node ids travel in the same pass (leaf phase, n[] in shared memory: the consumer then has nothing to do for it).
// The heap is ordered, so up the chain the frequencies only fall: the ancestors that move are a prefix of the chain and their
// number is a population count.  The removals before an insertion need no barrier (every lane ran all of their stores itself); the
// insertion's stores come from different lanes, hence the barrier after it.
template <bool WITH_N>
__device__ __forceinline__ uint32_t heap2_insert_warp(const HeapSmem &h, const int q0, uint32_t freq, uint32_t node, int lane) {
    const int p = q0 >> lane;                                   // lane 0: the new slot; lane k: its ancestor k levels up (0: above the root)
    uint32_t fk = 0u, nk = 0u;
    const bool anc = lane >= 1 && p >= 1;
    if (anc) { fk = h.ldf(p); if (WITH_N) nk = h.ldn(p); }
    const unsigned mv = __ballot_sync(0xffffffffu, anc && fk > freq);
    const int s = __popc(mv) + 1;                               // ancestors 1 .. s-1 move down, the new element lands on ancestor s-1
    const int fin = q0 >> (s - 1);
    if ((mv >> lane) & 1u) { h.stf(q0 >> (lane - 1), fk); if (WITH_N) h.stn(q0 >> (lane - 1), nk); }
    if (lane == 0) { h.stf(fin, freq); if (WITH_N) h.stn(fin, node); }
    __syncwarp();
    return (uint32_t)fin;
}
// Removal (ref: Huffman.c:87-114), "hole" form: the element taken from the end sinks from the root while it is STRICTLY larger than
// the smaller child (right child only if STRICTLY smaller).  THREE levels per trip: the children f[2i..2i+1], the grandchildren
// f[4i..4i+3] and the great-grandchildren f[8i..8i+7] are fetched together (all addresses follow from i alone), so the serial chain
// holds one shared-memory round trip per three levels; the three decisions are then taken in registers (compare, select, compare,
// select, compare) and the trip ends in ONE branch: the stores of the trip are predicated on how far the element sinks.
// LIM = false: the frequency array has room for the children of every position a sift can reach (cap >= 2 * (leaves + 2)), so the
// "no child slots" tests drop out.  A quad that would start beyond the array is fetched from a clamped address and never used:
// whenever a position is occupied its children's slots exist, and with them the whole aligned quad around them (cap is a
// multiple of 4).  Returns the removed frequency; fin = final position of the moved element (0: heap now empty).
template <bool LIM>
__device__ __forceinline__ uint32_t heap2_remove(const HeapSmem &h, int &qend, int cap, uint32_t &fin) {
    const uint32_t topf = h.ldf(1);
    qend--;
    const uint32_t mf = h.ldf(qend);
    h.stf(qend, HEAP_FINF);
    if (qend == 1) { fin = 0u; return topf; }
    // byte addresses throughout: a = address of f[i]; the children of f[i] start at 2a - base, and so on
    const unsigned fa = h.fa;
    const unsigned alim = fa + 4u * (unsigned)(cap >> 1);       // positions from here on have no child slots
    const unsigned aqmax = fa + 4u * (unsigned)(cap - 4), apmax = fa + 4u * (unsigned)(cap - 2);
    unsigned a = fa + 4u;
    for (;;) {
        const unsigned ap = 2u * a - fa, aq = 2u * ap - fa, ao = 2u * aq - fa;
        uint2 P = HeapSmem::ld2a(LIM ? min(ap, apmax) : ap);
        const uint4 Q = HeapSmem::ld4a(min(aq, aqmax));
        const uint4 O0 = HeapSmem::ld4a(min(ao, aqmax)), O1 = HeapSmem::ld4a(min(ao + 16u, aqmax));
        if (LIM && a >= alim) P = make_uint2(HEAP_FINF, HEAP_FINF);
        const bool r1 = P.y < P.x;
        const uint32_t c1 = min(P.x, P.y);
        unsigned a1 = ap; if (r1) a1 += 4u;
        uint2 P2 = r1 ? make_uint2(Q.z, Q.w) : make_uint2(Q.x, Q.y);
        const uint4 H = r1 ? O1 : O0;
        if (LIM && a1 >= alim) P2 = make_uint2(HEAP_FINF, HEAP_FINF);
        const bool r2 = P2.y < P2.x;
        const uint32_t c2 = min(P2.x, P2.y);
        unsigned a2 = 2u * a1 - fa; if (r2) a2 += 4u;
        uint2 P3 = r2 ? make_uint2(H.z, H.w) : make_uint2(H.x, H.y);
        if (LIM && a2 >= alim) P3 = make_uint2(HEAP_FINF, HEAP_FINF);
        const bool r3 = P3.y < P3.x;
        const uint32_t c3 = min(P3.x, P3.y);
        unsigned a3 = 2u * a2 - fa; if (r3) a3 += 4u;
        const bool g1 = mf > c1, g2 = g1 && mf > c2, g3 = g2 && mf > c3;      // sinks below level 1 / 2 / 3 of this trip
        if (g1) HeapSmem::sta(a, c1);
        if (g2) HeapSmem::sta(a1, c2);
        if (g3) HeapSmem::sta(a2, c3);
        if (!g3) { a = g2 ? a2 : (g1 ? a1 : a); break; }
        a = a3;
    }
    HeapSmem::sta(a, mf);
    const int i = (int)((a - fa) >> 2);
    fin = (uint32_t)i;
    return topf;
}
// The merge loop is a function of its own, not inlined: the code the compiler makes of it then does not depend on what surrounds
// it (the same loop measured between 21.7 and 25.9 M cycles for the 19 k-leaf tree as part of differently shaped callers).
// Everything travels by value in registers.  Returns the queue's sequence number.
template <bool LIM>
__device__ __noinline__ uint32_t heap2_merge_phase(unsigned fa, unsigned qa, bool remote, unsigned int *done, uint32_t qt, int hcap, int qend, int lane) {
    HeapSmem h; h.fa = fa; h.na = 0u;
    HeapQueue Q; Q.qa = qa; Q.remote = remote; Q.done = done; Q.t = qt; Q.lane0 = lane == 0;
    while (qend > 2) {
        uint32_t p1, p2;
        const uint32_t fb = heap2_remove<LIM>(h, qend, hcap, p1);    // first removed  -> right (gcc evaluates Huffman.c:181's arguments right to left)
        const uint32_t fa2 = heap2_remove<LIM>(h, qend, hcap, p2);   // second removed -> left
        const uint32_t p3 = heap2_insert_warp<false>(h, qend, fa2 + fb, 0u, lane); qend++;
        Q.push(HQ_MERGE, p1, p2, p3);
    }
    Q.push(HQ_FIN, 0u, 0u, 0u);
    return Q.t | (Q.dead ? 0x80000000u : 0u);
}
// LEAFN: n[] is in shared memory and the producer fills it during the leaf phase itself (no queue traffic until the first merge)
template <bool LIM, bool LEAFN>
__device__ __forceinline__ void heap2_warp_a(const HeapSmem &h, HeapQueue &Q, int hcap, uint32_t leaves, const uint32_t *lfreq, const HuffTreeDesc &D, long long tk1) {
    int lane = threadIdx.x & 31;
    asm volatile("mov.u32 %0, %0;" : "+r"(lane));                // kept in a register (the special register is not re-read per use)
    Q.lane0 = lane == 0;
    int qend = 1;
    // leaf frequencies: one coalesced load per 32 leaves, the next group in flight while this one is inserted
    uint32_t cur = (uint32_t)lane < leaves ? lfreq[lane] : 0u;
    for (uint32_t k0 = 0; k0 < leaves; k0 += 32) {
        const uint32_t nxt = k0 + 32u + (uint32_t)lane < leaves ? lfreq[k0 + 32u + lane] : 0u;
        const int cnt = (int)min(32u, leaves - k0);
        // The 32 leaves of a group go to the slots q0 .. q0+cnt-1, whose parents all lie below q0 (q0 >= 32).  A leaf that is not
        // smaller than its parent stays where it is appended; a sift of another leaf of the group only LOWERS the values on its
        // ancestors, so a test against the parent as it is before the group holds for the leaf's turn as well.  Those leaves are
        // therefore stored all at once, and only the others (bit set in `sift`) take the serial route, in order.
        const int q0 = qend;
        unsigned sift = cnt == 32 ? 0xffffffffu : (1u << cnt) - 1u;
        if (q0 >= 32) {
            const bool mine = lane < cnt;
            const uint32_t fp = mine ? h.ldf((q0 + lane) >> 1) : 0u;
            sift = __ballot_sync(0xffffffffu, mine && fp > cur);
            if (mine && ((sift >> lane) & 1u) == 0u) { h.stf(q0 + lane, cur); if (LEAFN) h.stn(q0 + lane, k0 + (uint32_t)lane); }
            __syncwarp();
        }
        if (LEAFN) {
            while (sift) {
                const int j = __ffs((int)sift) - 1; sift &= sift - 1u;
                heap2_insert_warp<true>(h, q0 + j, __shfl_sync(0xffffffffu, cur, j), k0 + (uint32_t)j, lane);
            }
        } else {
            for (int j = 0; j < cnt; j++) {                     // the consumer replays every leaf: one entry each, in order
                uint32_t fin = (uint32_t)(q0 + j);
                const uint32_t f = __shfl_sync(0xffffffffu, cur, j);
                if ((sift >> j) & 1u) fin = heap2_insert_warp<false>(h, q0 + j, f, 0u, lane);
                Q.push(HQ_LEAF, 0u, 0u, fin);
            }
        }
        qend = q0 + cnt;
        cur = nxt;
    }
    if (LEAFN) __threadfence_block();                           // n[] as left by the leaf phase, before the first queue entry
    if (lane == 0) D.out[5] = (uint32_t)((clock64() - tk1) >> 10);
    const uint32_t qt = heap2_merge_phase<LIM>(h.fa, Q.qa, Q.remote, Q.done, Q.t, hcap, qend, lane);
    Q.t = qt & 0x7fffffffu; if (qt >> 31) Q.dead = true;
    if (lane == 0) D.out[6] = (uint32_t)((clock64() - tk1) >> 10);
}

// The consumer (warp 1): node ids.  The warp replays each operation with one lane per node of the ancestor chain: all the loads of
// an operation are in flight together, the warp synchronises, every lane stores its value one level up (removal) or down
// (insertion).  n[] is in the shared memory of the CTA the consumer runs in (16- or 32-bit ids).  Returns the root node, or
// 0xffffffff if the producer stopped feeding the queue.
struct NodeIdsSmem {
    HeapSmem h;
    __device__ __forceinline__ uint32_t ld(uint32_t i) const { return h.ldn((int)i); }
    __device__ __forceinline__ void st(uint32_t i, uint32_t v) const { h.stn((int)i, v); }
};
// removal: a = final position of the moved element (0: the heap became empty), b = position it was taken from; the path is the
// chain of ancestors of a, every node of it moves up one level; the old root goes to *out
template <typename NA>
__device__ __forceinline__ void heap2_replay_remove(const NA &na, uint32_t a, uint32_t b, uint32_t *out, int lane) {
    const int depth = a ? 31 - __clz(a) : 0;
    const bool on_path = a != 0u && lane >= 1 && lane <= depth;
    const bool last = a != 0u && lane == 31;
    const uint32_t src = last ? b : (on_path ? a >> (depth - lane) : 1u);                // lane 0: the root
    uint32_t v = 0u;
    if (lane == 0 || on_path || last) v = na.ld(src);
    __syncwarp();
    if (on_path || last) na.st(last ? a : src >> 1, v);
    if (lane == 0) *out = v;
    __syncwarp();
}
// insertion: a = position appended at, b = final position, c = node id; the ancestors in between move down one level
template <typename NA>
__device__ __forceinline__ void heap2_replay_insert(const NA &na, uint32_t a, uint32_t b, uint32_t c, int lane) {
    const int k = __clz(b) - __clz(a);                                                   // depth(a) - depth(b)
    uint32_t v = c;
    if (lane < k) v = na.ld(a >> (lane + 1));
    __syncwarp();
    if (lane < k) na.st(a >> lane, v);
    if (lane == 31) na.st(b, c);
    __syncwarp();
}
template <typename NA>
__device__ __forceinline__ uint32_t heap2_warp_b(const NA &na, unsigned int qaddr, unsigned int done_addr, bool done_remote, uint32_t leaves, bool leaves_done, const HuffTreeDesc &D) {
    const int lane = threadIdx.x & 31;
    uint32_t qend = leaves_done ? leaves + 1u : 1u, nn = leaves, kleaf = 0u;              // counted along with the producer
    unsigned long long ahead;                                                            // the next entry, fetched while the current one is replayed
    asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(ahead) : "r"(qaddr) : "memory");
    for (uint32_t t = 0;; t++) {
        const unsigned int slot = qaddr + 8u * (t & (HQ_CAP - 1));
        unsigned long long e = ahead;                                                     // every lane polls the same word: one broadcast load, uniform exit
        unsigned spins = 0;
        while ((uint32_t)(e >> 52) != ((t + 1) & 4095u)) {
            asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(e) : "r"(slot) : "memory");
            if (++spins > HQ_SPIN_LIMIT) return 0xffffffffu;                             // the producer stopped: report, do not hang
        }
        asm volatile("ld.volatile.shared.u64 %0, [%1];" : "=l"(ahead) : "r"(qaddr + 8u * ((t + 1) & (HQ_CAP - 1))) : "memory");
        const uint32_t kind = (uint32_t)(e >> 48) & 3u;
        const uint32_t p1 = (uint32_t)e & 0xffffu, p2 = ((uint32_t)e >> 16) & 0xffffu, p3 = (uint32_t)(e >> 32) & 0xffffu;
        if (kind == HQ_FIN) break;
        if (kind == HQ_MERGE) {
            qend--; heap2_replay_remove(na, p1, qend, D.n_right + nn, lane);
            qend--; heap2_replay_remove(na, p2, qend, D.n_left + nn, lane);
            heap2_replay_insert(na, qend, p3, nn, lane); qend++; nn++;
        } else {
            heap2_replay_insert(na, qend, p3, kleaf, lane); qend++; kleaf++;
        }
        if (lane == 0 && (t & 63u) == 63u) {
            if (done_remote) asm volatile("st.relaxed.cluster.shared::cluster.u32 [%0], %1;" ::"r"(done_addr), "r"(t + 1) : "memory");
            else asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(done_addr), "r"(t + 1) : "memory");
        }
    }
    __syncwarp();
    return na.ld(1u);
}

constexpr int HUFF_TB = 1024;                                 // the breadth-first passes after the heap use the threads; the heap itself two warps
// heap_form: 0 = the pipelined one-warp heap (szb_heap_pipe.cuh) whenever the tree fits one SM's shared memory as 8-byte entries, 1 = never (the
// two-warp forms below, which also serve the larger trees)
__global__ void __launch_bounds__(HUFF_TB) huff_build_kernel(const HuffTreeDesc *descs, uint32_t smem_heap_capacity, uint32_t heap_form) {
    extern __shared__ __align__(16) uint32_t s_heap[];          // f[cap] then n[cap]
    // launched as clusters of two CTAs per tree: rank 0 builds the tree; rank 1 only works for a tree whose frequencies and node
    // ids do not both fit one SM's shared memory -- it then keeps the node ids and runs the consumer (see "pair layout" below)
    // (the host launches plain one-CTA-per-tree grids when no tree of the launch can have 32768 leaves or more)
    const unsigned csize = cluster_size();
    const unsigned rank = csize > 1 ? cluster_rank() : 0u;
    const HuffTreeDesc D = descs[blockIdx.x / csize];
    __shared__ uint32_t s_scan[HUFF_TB];
    __shared__ uint32_t s_root, s_leaves, s_cnt, s_err, s_lvl[72];
    const int t = threadIdx.x;
    if (t == 0) s_cnt = 0;
    if (D.wait_flag) {
        // launched beside the producer of the histogram and already resident on its SM: the heap starts the moment the counts are final,
        // with no launch in between (a launch behind the producer lost ~0.3 ms to other kernels taking the SMs first)
        if (t == 0) {
            unsigned long long g0, g1; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g0)); g1 = g0;
            if (rank == 0) D.out[14] = (uint32_t)(g0 / 1000ull);
            while (*(volatile const unsigned int *)D.wait_flag == 0u) {
                __nanosleep(200);
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
                if (g1 - g0 > 4000000000ull) { s_cnt = 0xffffffffu; break; }          // 4 s: the producer never came
            }
            if (rank == 0) D.out[15] = (uint32_t)(g1 / 1000ull);
            __threadfence();
        }
        __syncthreads();
        if (s_cnt == 0xffffffffu) { if (t == 0 && rank == 0) { D.out[0] = 0; D.out[1] = 0; D.out[2] = 0; D.out[3] = 3; } return; }
    }
    const long long tk0 = clock64();
    unsigned long long gt0; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt0));
    // ---- phase A (parallel): ordered compaction of the non-empty bins -> leaf nodes 0..leaves-1 in increasing symbol order.
    //      Every warp takes a contiguous range of bins, 32 at a time (coalesced), one vote per row counts and places them ----
    uint32_t *lfreq = D.n_pre;                                  // scratch until phase E
    {
        const int warp = t >> 5, lane = t & 31;
        const uint32_t rows = (D.nbins + HUFF_TB - 1) / HUFF_TB;                   // rows of 32 bins per warp
        const uint32_t bin0 = (uint32_t)warp * rows * 32u;
        uint32_t cnt = 0;
#pragma unroll 8
        for (uint32_t r = 0; r < rows; r++) {
            const uint32_t bin = bin0 + 32u * r + (uint32_t)lane;
            const uint32_t f = bin < D.nbins ? D.hist[bin] : 0u;
            cnt += __popc(__ballot_sync(0xffffffffu, f != 0u));
        }
        if (lane == 0) s_scan[warp] = cnt;
        __syncthreads();
        if (warp == 0) {
            const uint32_t c = s_scan[lane];
            uint32_t inc = c;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
            s_scan[32 + lane] = inc - c;
            if (lane == 31) { s_leaves = inc; s_err = 0; }
        }
        __syncthreads();
        uint32_t k = s_scan[32 + warp];
        if (rank == 0) {
#pragma unroll 8
            for (uint32_t r = 0; r < rows; r++) {                                  // the counts again (cache hits)
                const uint32_t bin = bin0 + 32u * r + (uint32_t)lane;
                const uint32_t f = bin < D.nbins ? D.hist[bin] : 0u;
                const unsigned mk = __ballot_sync(0xffffffffu, f != 0u);
                if (f != 0u) { const uint32_t idx = k + __popc(mk & ((1u << lane) - 1u)); D.n_sym[idx] = bin; lfreq[idx] = f; }
                k += __popc(mk);
            }
        }
        __syncthreads();
    }
    const uint32_t leaves = s_leaves;
    const uint32_t nodeCount = leaves ? 2 * leaves - 1 : 0;
    if (leaves == 0) { if (t == 0 && rank == 0) { D.out[0] = 0; D.out[1] = 0; D.out[2] = 0; D.out[3] = 0; D.out[12] = 0; D.out[13] = 0; } return; }
    // the code table of the absent symbols (never looked up by the encoder; kept defined) -- by the threads that would otherwise
    // only wait for the heap
    if (rank == 0 && t >= 64) for (uint32_t i = t - 64; i < D.nbins; i += blockDim.x - 64) { D.code[i] = 0; D.clen[i] = 0; }
    // ---- phase B (serial): exact emulation of the reference's heap ----
    const long long tk1 = clock64();
    bool piped_tree = false;
    {
        // separate inlined copies of the heap loops per layout, so that the shared-memory ones compile to LDS/STS with 32-bit
        // addresses instead of generic loads (a pointer that may be shared OR global forces the slower generic path)
        __shared__ unsigned int s_qdone;
        // compact layout (trees with fewer than 32768 leaves: node ids fit 16 bits): f gets 2*(leaves+2) entries -- child slots exist for
        // every position a sift can reach, so the level loop needs no "no child slots" test; up to ~22.8 k leaves.
        // tight layout: f gets leaves + 8 entries ("no child slots" tests in the level loop); up to 32767 leaves.
        const uint32_t fcap_c = (2u * (leaves + 2u) + 3u) & ~3u;
        const uint32_t n16_bytes = ((leaves + 8u) * 2u + 7u) & ~7u;
        const bool compact = 2u * leaves < 65536u && (unsigned long long)fcap_c * 4u + n16_bytes + 8u * HQ_CAP <= (unsigned long long)smem_heap_capacity * 8u;
        const uint32_t fcap_t16 = (leaves + 8u + 3u) & ~3u;
        const bool tight16 = 2u * leaves < 65536u && (unsigned long long)fcap_t16 * 4u + n16_bytes + 8u * HQ_CAP <= (unsigned long long)smem_heap_capacity * 8u;
        // pair layout (neither of the above fits, up to ~54 k leaves): rank 0 keeps the queue-sized gap and f[], rank 1 the queue and
        // n[] at the same offsets; the producer writes its queue entries into rank 1's shared memory (one 64-bit remote store per
        // merge, nothing waits for it), the consumer works on its own SM's memory and sends its progress counter and the root back
        const uint32_t fcap_p = (uint32_t)(((smem_heap_capacity * 8u - 8u * HQ_CAP) / 4u) & ~3u);
        const bool piped = heap_form == 0u && leaves + 4u <= smem_heap_capacity && leaves < 32768u;
        piped_tree = piped;
        const bool pair = csize > 1 && !piped && !compact && !tight16 && leaves + 8u <= fcap_p;
        if (rank != 0 && !pair) return;
        if (!piped && rank == 0) { for (uint32_t q = t; q < leaves; q += blockDim.x) { D.n_left[q] = 0xffffffffu; D.n_right[q] = 0xffffffffu; } __syncthreads(); }
        if (piped) {
            // several heap operations in flight in one warp, {frequency, node} entries of 8 bytes (szb_heap_pipe.cuh)
            if (t < 32) {
                const uint32_t r = hp_build((unsigned int)__cvta_generic_to_shared(s_heap), leaves, lfreq, D.n_left, D.n_right, D.n_sizeL, D.out, tk1);
                if (t == 0) { s_root = r & 0xffffu; s_cnt = r >> 16; }
            }
        } else if (compact || tight16) {
            // one CTA: f[fcap], n[leaves + 8] as 16-bit ids, the queue
            const uint32_t fcap = compact ? fcap_c : fcap_t16;
            unsigned char *s_bytes = reinterpret_cast<unsigned char *>(s_heap);
            uint32_t *s_queue = reinterpret_cast<uint32_t *>(s_bytes + (size_t)fcap * 4u + n16_bytes);
            for (int q = t; q < (int)fcap; q += blockDim.x) s_heap[q] = HEAP_FINF;
            for (int q = t; q < 2 * HQ_CAP; q += blockDim.x) s_queue[q] = 0u;
            if (t == 0) s_qdone = 0u;
            __syncthreads();
            if (t < 64) {
                // the addresses pass through an opaque asm so that the compiler keeps them in registers instead of re-deriving the
                // shared window base (S2R + LEA) at every use inside the sift loops
                HeapSmem h;
                const unsigned int fa0 = (unsigned int)__cvta_generic_to_shared(s_heap), na0 = fa0 + 4u * fcap, qa0 = na0 + n16_bytes;
                unsigned int qaddr;
                asm volatile("mov.u32 %0, %3;\n\tmov.u32 %1, %4;\n\tmov.u32 %2, %5;" : "=r"(h.fa), "=r"(h.na), "=r"(qaddr) : "r"(fa0), "r"(na0), "r"(qa0));
                h.n16 = true;
                if (t < 32) {
                    HeapQueue Q; Q.qa = qaddr; Q.done = &s_qdone; Q.t = 0;
                    if (compact) heap2_warp_a<false, true>(h, Q, (int)fcap, leaves, lfreq, D, tk1);
                    else heap2_warp_a<true, true>(h, Q, (int)fcap, leaves, lfreq, D, tk1);
                } else { NodeIdsSmem na; na.h = h; const uint32_t root = heap2_warp_b(na, qaddr, (unsigned)__cvta_generic_to_shared(&s_qdone), false, leaves, true, D); if (t == 32) s_root = root; }
            }
        } else if (pair) {
            uint32_t *s_queue = s_heap, *s_arr = s_heap + 2 * HQ_CAP;                      // rank 0: f[fcap_p]; rank 1: n[fcap_p]
            if (rank == 0) for (int q = t; q < (int)fcap_p; q += blockDim.x) s_arr[q] = HEAP_FINF;
            for (int q = t; q < 2 * HQ_CAP; q += blockDim.x) s_queue[q] = 0u;
            if (t == 0) s_qdone = 0u;
            cluster_sync_all();                                                          // the ring in rank 1 is clear before the first entry
            if (t < 32) {
                const unsigned int qa0 = (unsigned int)__cvta_generic_to_shared(s_queue), aa0 = (unsigned int)__cvta_generic_to_shared(s_arr);
                unsigned int qaddr, aaddr;
                asm volatile("mov.u32 %0, %2;\n\tmov.u32 %1, %3;" : "=r"(qaddr), "=r"(aaddr) : "r"(qa0), "r"(aa0));
                HeapSmem h;
                if (rank == 0) {
                    h.fa = aaddr; h.na = 0;
                    HeapQueue Q; Q.qa = cluster_addr(qaddr, 1u); Q.remote = true; Q.done = &s_qdone; Q.t = 0;
                    heap2_warp_a<true, false>(h, Q, (int)fcap_p, leaves, lfreq, D, tk1);
                    if (Q.dead && t == 0) s_err = 2;
                } else {
                    h.fa = 0; h.na = aaddr;
                    NodeIdsSmem na; na.h = h;
                    const uint32_t root = heap2_warp_b(na, qaddr, cluster_addr((unsigned)__cvta_generic_to_shared(&s_qdone), 0u), true, leaves, false, D);
                    if (t == 0) asm volatile("st.relaxed.cluster.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr((unsigned)__cvta_generic_to_shared(&s_root), 0u)), "r"(root) : "memory");
                }
            }
            if (rank != 0) __threadfence();                                              // the children written by the consumer, for rank 0's passes
            cluster_sync_all();
            if (rank != 0) return;
        } else {
            uint32_t *hg = reinterpret_cast<uint32_t *>(D.heap_g);       // capacity 64-bit words = two 32-bit arrays of that length
            const int gcap = (int)(D.capacity & ~3u);
            for (int q = t; q < gcap; q += blockDim.x) hg[q] = HEAP_FINF;
            __syncthreads();
            if (t == 0) { HeapGmem h; h.f = hg; h.n = hg + gcap; s_root = heap_build_tree(h, gcap, leaves, lfreq, D, tk1); }
        }
    }
    __syncthreads();
    if (s_root >= nodeCount || s_err == 2) {                    // a side of the two-warp build gave up waiting for the other (never seen)
        if (t == 0) { D.out[0] = 0; D.out[1] = 0; D.out[2] = leaves; D.out[3] = 2; }
        return;
    }
    const long long tk2 = clock64();
    // ---- serialised layout (ref: Huffman.c:503-585): [endian:1][L: w*n][R: w*n][C: 4*n][t: n], native endian ----
    const uint32_t w = nodeCount <= 256 ? 1u : (nodeCount <= 65536 ? 2u : 4u);
    unsigned char *tb = D.tree_bytes;
    unsigned char *pl = tb + 1, *pr = pl + (size_t)w * nodeCount, *pc = pr + (size_t)w * nodeCount, *pt = pc + (size_t)4 * nodeCount;
    auto put_w = [&](unsigned char *base, uint32_t idx, uint32_t v) {
        if (w == 1) base[idx] = (unsigned char)v;
        else if (w == 2) { base[2 * idx] = (unsigned char)v; base[2 * idx + 1] = (unsigned char)(v >> 8); }
        else { base[4 * idx] = (unsigned char)v; base[4 * idx + 1] = (unsigned char)(v >> 8); base[4 * idx + 2] = (unsigned char)(v >> 16); base[4 * idx + 3] = (unsigned char)(v >> 24); }
    };
    if (piped_tree) {
        // ---- the pipelined heap left, per node, {parent, 1 + size of the left sibling's subtree if it is a right child, else 1}.  The
        //      pre-order index of a node (pad_tree_*, Huffman.c:443-501) is the sum of the second field along its ancestor chain, its code
        //      (build_code, :122-157) the "is a right child" bits along the chain, deepest bit last.  Every node walks its own chain, in
        //      shared memory (the heap's, free now): no level-by-level passes, no frontier ----
        uint32_t *ups = s_heap;
        uint32_t *pre_g = D.n_pre;                                  // pre-order index per node (the leaf frequencies kept there are done with)
        for (uint32_t n = t; n < nodeCount; n += blockDim.x) ups[n] = D.n_left[n];       // the records of the last few merges and the root
        if (t == 0) tb[0] = 0;
        __syncthreads();
        for (uint32_t k = t; k < s_cnt; k += blockDim.x) hp_apply_log(ups, D.n_right, D.n_sizeL, leaves, k);   // all the others, from the merge log
        __syncthreads();
        for (uint32_t n = t; n < nodeCount; n += blockDim.x) D.n_left[n] = ups[n];
        __syncthreads();
        for (uint32_t n = t; n < nodeCount; n += blockDim.x) {
            uint32_t e = ups[n], pre = 0, len = 0; uint64_t code = 0;
            while (e != HP_ROOT) {
                const uint32_t step = e >> 16;
                pre += step;
                if (len < 64 && step > 1u) code |= 1ULL << len;
                len++;
                e = ups[e & 0xffffu];
            }
            pre_g[n] = pre;
            if (n < leaves) {
                const uint32_t sym = D.n_sym[n];
                if (len > 64) s_err = 1;
                D.code[sym] = code; D.clen[sym] = (uint8_t)len;
            }
        }
        __syncthreads();
        // the serialised arrays are written in THEIR order (consecutive threads, consecutive records): inv[] = node of a pre-order
        // index, rch[] = pre-order index of the right child (the left child's is the record's own + 1), both in shared memory
        uint16_t *inv = reinterpret_cast<uint16_t *>(s_heap), *rch = inv + nodeCount + 1;
        for (uint32_t n = t; n < nodeCount; n += blockDim.x) {
            const uint32_t pre = pre_g[n], e0 = D.n_left[n];        // (the shared copy of the records is being overwritten)
            inv[pre] = (uint16_t)n;
            if (e0 != HP_ROOT && (e0 >> 16) > 1u) rch[pre - (e0 >> 16)] = (uint16_t)pre;
        }
        __syncthreads();
        for (uint32_t q = t; q < nodeCount; q += blockDim.x) {
            const uint32_t n = inv[q];
            uint32_t l = 0, r = 0, c = 0, leaf = 0;
            if (n < leaves) { c = D.n_sym[n]; leaf = 1; } else { l = q + 1; r = rch[q]; }
            put_w(pl, q, l); put_w(pr, q, r);
            pc[4 * q] = (unsigned char)c; pc[4 * q + 1] = (unsigned char)(c >> 8); pc[4 * q + 2] = (unsigned char)(c >> 16); pc[4 * q + 3] = (unsigned char)(c >> 24);
            pt[q] = (unsigned char)leaf;
        }
        __syncthreads();
        if (t == 0) { D.out[0] = 1 + 2 * w * nodeCount + 4 * nodeCount + nodeCount; D.out[1] = nodeCount; D.out[2] = leaves; D.out[3] = s_err;
                      D.out[4] = (uint32_t)((tk1 - tk0) >> 10); D.out[7] = (uint32_t)((clock64() - tk2) >> 10);
                      unsigned long long gt1; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt1));
                      D.out[12] = (uint32_t)(gt1 - gt0); D.out[13] = (uint32_t)((clock64() - tk0) >> 10); }      // wall ns and kcycles of the whole CTA (profiling aid)
        return;
    }
    // ---- phase C (parallel): breadth-first order of the finished tree: order[] = nodes level by level, s_lvl[] = level starts ----
    uint32_t *order = D.frontier;                               // [capacity]
    uint32_t *nsize = D.frontier + D.capacity;                  // [capacity] subtree sizes (in nodes)
    if (t == 0) { order[0] = s_root; s_lvl[0] = 0; s_lvl[1] = 1; s_cnt = 1; }
    __syncthreads();
    int nlev = 1;
    for (; nlev < 70; nlev++) {
        const uint32_t lo = s_lvl[nlev - 1], hi = s_lvl[nlev];
        if (lo == hi) break;
        for (uint32_t q = lo + t; q < hi; q += blockDim.x) {
            const uint32_t n = order[q];
            const uint32_t l = D.n_left[n];
            if (l != 0xffffffffu) { uint32_t o = atomicAdd(&s_cnt, 2u); order[o] = l; order[o + 1] = D.n_right[n]; }
        }
        __syncthreads();
        if (t == 0) s_lvl[nlev + 1] = s_cnt;
        __syncthreads();
    }
    // s_lvl[nlev-1] == s_lvl[nlev]: levels 0 .. nlev-2 are non-empty
    const int last_level = nlev - 2;
    if (t == 0 && s_lvl[nlev - 1] != s_lvl[nlev]) s_err = 1;    // deeper than 68 levels: codes would not fit 64 bits anyway
    // ---- phase D (parallel, bottom-up): subtree sizes ----
    for (int lv = last_level; lv >= 0; lv--) {
        const uint32_t lo = s_lvl[lv], hi = s_lvl[lv + 1];
        for (uint32_t q = lo + t; q < hi; q += blockDim.x) {
            const uint32_t n = order[q];
            const uint32_t l = D.n_left[n];
            nsize[n] = (l == 0xffffffffu) ? 1u : 1u + nsize[l] + nsize[D.n_right[n]];
        }
        __syncthreads();
    }
    // ---- phase E (parallel, top-down): pre-order index (pad_tree_*, Huffman.c:443-501), codes (build_code, :122-157), serialised arrays ----
    if (t == 0) { tb[0] = 0; D.n_pre[s_root] = 0; D.n_code[s_root] = 0; D.n_len[s_root] = 0; }
    __syncthreads();
    for (int lv = 0; lv <= last_level; lv++) {
        const uint32_t lo = s_lvl[lv], hi = s_lvl[lv + 1];
        for (uint32_t q = lo + t; q < hi; q += blockDim.x) {
            const uint32_t n = order[q];
            const uint32_t pre = D.n_pre[n]; const uint64_t code = D.n_code[n]; const uint32_t len = D.n_len[n];
            const uint32_t l = D.n_left[n];
            if (l == 0xffffffffu) {            // leaf
                const uint32_t sym = D.n_sym[n];
                put_w(pl, pre, 0); put_w(pr, pre, 0);
                pc[4 * pre] = (unsigned char)sym; pc[4 * pre + 1] = (unsigned char)(sym >> 8); pc[4 * pre + 2] = (unsigned char)(sym >> 16); pc[4 * pre + 3] = (unsigned char)(sym >> 24);
                pt[pre] = 1;
                if (len > 64) s_err = 1;
                D.code[sym] = code; D.clen[sym] = (uint8_t)len;
            } else {
                const uint32_t r = D.n_right[n];
                const uint32_t pre_l = pre + 1, pre_r = pre + 1 + nsize[l];
                put_w(pl, pre, pre_l); put_w(pr, pre, pre_r);
                pc[4 * pre] = 0; pc[4 * pre + 1] = 0; pc[4 * pre + 2] = 0; pc[4 * pre + 3] = 0;
                pt[pre] = 0;
                D.n_pre[l] = pre_l; D.n_pre[r] = pre_r;
                D.n_code[l] = code << 1; D.n_code[r] = (code << 1) | 1ULL;
                D.n_len[l] = (uint8_t)(len + 1); D.n_len[r] = (uint8_t)(len + 1);
            }
        }
        __syncthreads();
    }
    if (t == 0) { D.out[0] = 1 + 2 * w * nodeCount + 4 * nodeCount + nodeCount; D.out[1] = nodeCount; D.out[2] = leaves; D.out[3] = s_err;
                  D.out[4] = (uint32_t)((tk1 - tk0) >> 10); D.out[7] = (uint32_t)((clock64() - tk2) >> 10);    // phase cycles / 1024 (profiling aid)
                  unsigned long long gt1; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt1));
                  D.out[12] = (uint32_t)(gt1 - gt0); D.out[13] = (uint32_t)((clock64() - tk0) >> 10); }
}

// =====================================================================================================================
// encode: pass 1 = bit length + zero-code count per CTA chunk; pass 2 = bit scatter at absolute bit offsets and
// block-major scatter of the unpredictable values.  BLOCKMAJOR = walk block-major order over a raster code array
// (main type array); otherwise the symbol array is linear (coefficient codes).
// =====================================================================================================================
constexpr int ENC_THREADS = 256;
constexpr int ENC_PER_THREAD = 16;
constexpr int ENC_CHUNK = ENC_THREADS * ENC_PER_THREAD;

template <bool BLOCKMAJOR>
__global__ void __launch_bounds__(ENC_THREADS) enc_len_kernel(const uint16_t *__restrict__ s, uint64_t n, Geom g, const uint8_t *__restrict__ clen,
                                                              uint32_t *chunk_bits, uint32_t *chunk_zeros) {
    const uint64_t p0 = ((uint64_t)blockIdx.x * ENC_THREADS + threadIdx.x) * ENC_PER_THREAD;
    uint32_t bits = 0, zeros = 0;
    if (p0 < n) {
        BMIter it;
        if (BLOCKMAJOR) it.init(g, p0);
        const int cnt = (int)min((uint64_t)ENC_PER_THREAD, n - p0);
        for (int q = 0; q < cnt; q++) {
            uint32_t sym = BLOCKMAJOR ? s[it.raster] : s[p0 + q];
            bits += clen[sym];
            zeros += (sym == 0);
            if (BLOCKMAJOR) it.next(g);
        }
    }
    __shared__ uint32_t sb[ENC_THREADS / 32], sz_[ENC_THREADS / 32];
    for (int o = 16; o > 0; o >>= 1) { bits += __shfl_xor_sync(0xffffffffu, bits, o); zeros += __shfl_xor_sync(0xffffffffu, zeros, o); }
    if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = bits; sz_[threadIdx.x >> 5] = zeros; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t b = 0, z = 0;
        for (int i = 0; i < ENC_THREADS / 32; i++) { b += sb[i]; z += sz_[i]; }
        chunk_bits[blockIdx.x] = b; chunk_zeros[blockIdx.x] = z;
    }
}

template <typename T, bool BLOCKMAJOR>
__global__ void __launch_bounds__(ENC_THREADS) enc_write_kernel(const uint16_t *__restrict__ s, uint64_t n, Geom g,
                                                                const uint64_t *__restrict__ code, const uint8_t *__restrict__ clen,
                                                                const unsigned long long *__restrict__ chunk_bit_off,
                                                                const unsigned long long *__restrict__ chunk_zero_off,
                                                                uint32_t *out_words /*zeroed, 4-byte aligned*/,
                                                                const T *__restrict__ x, T *unpred_out) {
    const uint64_t p0 = ((uint64_t)blockIdx.x * ENC_THREADS + threadIdx.x) * ENC_PER_THREAD;
    uint32_t syms[ENC_PER_THREAD];
    uint64_t rast0 = 0;
    uint32_t bits = 0, zeros = 0;
    int cnt = 0;
    BMIter it0;
    if (p0 < n) {
        BMIter it;
        if (BLOCKMAJOR) { it.init(g, p0); it0 = it; rast0 = it.raster; }
        cnt = (int)min((uint64_t)ENC_PER_THREAD, n - p0);
#pragma unroll
        for (int q = 0; q < ENC_PER_THREAD; q++) {
            if (q < cnt) {
                uint32_t sym = BLOCKMAJOR ? s[it.raster] : s[p0 + q];
                syms[q] = sym; bits += clen[sym]; zeros += (sym == 0);
                if (BLOCKMAJOR) it.next(g);
            } else syms[q] = 0;
        }
    }
    (void)rast0;
    // CTA-wide exclusive prefix of (bits, zeros)
    __shared__ uint32_t wb[ENC_THREADS / 32], wz[ENC_THREADS / 32];
    uint32_t ib = bits, iz = zeros;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t a = __shfl_up_sync(0xffffffffu, ib, o), b = __shfl_up_sync(0xffffffffu, iz, o);
        if (lane >= o) { ib += a; iz += b; }
    }
    if (lane == 31) { wb[wid] = ib; wz[wid] = iz; }
    __syncthreads();
    uint32_t pb = 0, pz = 0;
    for (int i = 0; i < wid; i++) { pb += wb[i]; pz += wz[i]; }
    if (cnt == 0) return;
    uint64_t pos = chunk_bit_off[blockIdx.x] + pb + (ib - bits);
    uint64_t zr = chunk_zero_off[blockIdx.x] + pz + (iz - zeros);
    // bit scatter: MSB-first stream, 32-bit big-endian words, atomicOr on shared boundary words
    uint64_t widx = pos >> 5;
    uint32_t fill = (uint32_t)(pos & 31), cur = 0;
    BMIter it = it0;
#pragma unroll
    for (int q = 0; q < ENC_PER_THREAD; q++) {
        if (q < cnt) {
            uint32_t sym = syms[q];
            uint32_t len = clen[sym];
            uint64_t cb = code[sym];
            while (len > 0) {
                uint32_t take = min(32u - fill, len);
                uint32_t piece = (uint32_t)((cb >> (len - take)) & ((take == 32) ? 0xffffffffULL : ((1ULL << take) - 1ULL)));
                cur |= piece << (32 - fill - take);
                fill += take; len -= take;
                if (fill == 32) { atomicOr(out_words + widx, bswap32(cur)); widx++; cur = 0; fill = 0; }
            }
            if (unpred_out && sym == 0) { unpred_out[zr++] = BLOCKMAJOR ? x[it.raster] : x[p0 + q]; }
            if (BLOCKMAJOR) it.next(g);
        }
    }
    if (fill > 0 && cur != 0) atomicOr(out_words + widx, bswap32(cur));
}

// The coefficient streams (linear symbol arrays, one per coefficient) in ONE launch: bit lengths, their prefix over the chunks by
// decoupled look-back, and the bit scatter.  grid = (chunks, coefficients); chunks are taken by ticket so that every chunk a CTA
// waits for is already running.  state[e][chunk]: bits 63..62 = 1 the chunk's own bit count is in the low bits, 2 the count of all
// chunks up to and including it; zeroed before the launch.
struct CoefEnc { const uint16_t *syms; const uint64_t *code; const uint8_t *clen; uint32_t *out_words; unsigned long long *total_bits; };
struct CoefEncArgs { CoefEnc c[4]; uint64_t n; unsigned long long *state; unsigned int *ticket; uint32_t nchunks; };
__global__ void __launch_bounds__(ENC_THREADS) coef_encode_kernel(CoefEncArgs A) {
    const int e = blockIdx.y, t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const CoefEnc C = A.c[e];
    __shared__ unsigned int s_chunk;
    __shared__ uint32_t wb[ENC_THREADS / 32];
    __shared__ unsigned long long s_excl;
    if (t == 0) s_chunk = atomicAdd(A.ticket + e, 1u);
    __syncthreads();
    const uint32_t chunk = s_chunk;
    unsigned long long *state = A.state + (size_t)e * A.nchunks;
    const uint64_t p0 = ((uint64_t)chunk * ENC_THREADS + t) * ENC_PER_THREAD;
    uint32_t syms[ENC_PER_THREAD], lens[ENC_PER_THREAD];
    const int cnt = p0 < A.n ? (int)min((uint64_t)ENC_PER_THREAD, A.n - p0) : 0;
    uint32_t bits = 0;
#pragma unroll
    for (int q = 0; q < ENC_PER_THREAD; q++) syms[q] = q < cnt ? C.syms[p0 + q] : 0u;
#pragma unroll
    for (int q = 0; q < ENC_PER_THREAD; q++) { lens[q] = q < cnt ? (uint32_t)C.clen[syms[q]] : 0u; bits += lens[q]; }
    uint32_t ib = bits;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t a = __shfl_up_sync(0xffffffffu, ib, o); if (lane >= o) ib += a; }
    if (lane == 31) wb[wid] = ib;
    __syncthreads();
    uint32_t pb = 0, tot = 0;
    for (int i = 0; i < ENC_THREADS / 32; i++) { if (i < wid) pb += wb[i]; tot += wb[i]; }
    if (wid == 0) {
        const unsigned long long F_AGG = 1ull << 62, F_PRE = 2ull << 62, VMASK = (1ull << 62) - 1ull;
        if (lane == 0) { __threadfence(); atomicExch(state + chunk, (chunk == 0 ? F_PRE : F_AGG) | (unsigned long long)tot); }
        unsigned long long excl = 0;
        long long look = (long long)chunk - 1;
        while (look >= 0) {                                 // 32 predecessors at a time, nearest first
            const long long idx = look - lane;
            unsigned long long v = F_PRE;                   // before the first chunk: a prefix of zero
            if (idx >= 0) { do { v = *(volatile unsigned long long *)(state + idx); } while ((v >> 62) == 0ull); }
            const unsigned int pre = __ballot_sync(0xffffffffu, (v >> 62) == 2ull);
            const int stop = pre ? __ffs(pre) - 1 : 32;     // the nearest chunk that already knows its prefix
            unsigned long long part = lane <= stop ? (v & VMASK) : 0ull;
            for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            excl += part;
            if (pre) break;
            look -= 32;
        }
        if (lane == 0) {
            if (chunk != 0) { __threadfence(); atomicExch(state + chunk, F_PRE | (excl + tot)); }
            s_excl = excl;
            if (chunk == A.nchunks - 1) C.total_bits[0] = excl + tot;
        }
    }
    __syncthreads();
    if (cnt == 0) return;
    const uint64_t pos = s_excl + pb + (ib - bits);
    uint64_t widx = pos >> 5;
    uint32_t fill = (uint32_t)(pos & 31), cur = 0;
    uint64_t cd[ENC_PER_THREAD];
#pragma unroll
    for (int q = 0; q < ENC_PER_THREAD; q++) cd[q] = q < cnt ? C.code[syms[q]] : 0ull;
#pragma unroll
    for (int q = 0; q < ENC_PER_THREAD; q++) {
        uint32_t len = lens[q];
        const uint64_t cb = cd[q];
        while (len > 0) {
            const uint32_t take = min(32u - fill, len);
            const uint32_t piece = (uint32_t)((cb >> (len - take)) & ((take == 32) ? 0xffffffffULL : ((1ULL << take) - 1ULL)));
            cur |= piece << (32 - fill - take);
            fill += take; len -= take;
            if (fill == 32) { atomicOr(C.out_words + widx, bswap32(cur)); widx++; cur = 0; fill = 0; }
        }
    }
    if (fill > 0 && cur != 0) atomicOr(C.out_words + widx, bswap32(cur));
}

// =====================================================================================================================
// decode.  The stream carries no chunk index (ref: Huffman.c:310-343 is a bit-serial walk), so the decoder is
// self-synchronising: every thread owns a fixed-size sub-sequence of the bit stream; it first decodes speculatively
// from the sub-sequence start, then repeatedly restarts from the position where its predecessor's last codeword ended
// until no start position changes any more (Jacobi iteration; after m rounds the first m+1 sub-sequences are exact,
// a fixed point is therefore the exact serial decode).  Counts are prefix-summed and a final pass writes symbols.
// =====================================================================================================================
constexpr int DEC_SUB_BITS = 256;

// unpack the serialised tree (arbitrary byte alignment) into u32 child arrays; leaf <=> t != 0.  A child index carries
// DEC_CHILD_LEAF when the child is a leaf, so that the bit-serial walk needs one load per bit, not two.
constexpr uint32_t DEC_CHILD_LEAF = 0x80000000u;
__global__ void dec_tree_unpack_kernel(const unsigned char *__restrict__ tb, uint32_t nodeCount, uint32_t *L, uint32_t *R, uint32_t *C, uint8_t *leaf) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodeCount) return;
    const uint32_t w = nodeCount <= 256 ? 1u : (nodeCount <= 65536 ? 2u : 4u);
    const unsigned char *pl = tb + 1, *pr = pl + (size_t)w * nodeCount, *pc = pr + (size_t)w * nodeCount, *pt = pc + (size_t)4 * nodeCount;
    uint32_t l = 0, r = 0;
    for (uint32_t b = 0; b < w; b++) { l |= (uint32_t)pl[(size_t)w * i + b] << (8 * b); r |= (uint32_t)pr[(size_t)w * i + b] << (8 * b); }
    uint32_t c = (uint32_t)pc[4 * (size_t)i] | ((uint32_t)pc[4 * (size_t)i + 1] << 8) | ((uint32_t)pc[4 * (size_t)i + 2] << 16) | ((uint32_t)pc[4 * (size_t)i + 3] << 24);
    L[i] = l < nodeCount ? (l | (pt[l] ? DEC_CHILD_LEAF : 0u)) : 0; R[i] = r < nodeCount ? (r | (pt[r] ? DEC_CHILD_LEAF : 0u)) : 0;
    C[i] = c; leaf[i] = pt[i] ? 1 : 0;
}

constexpr int DEC_K = 11;                                   // first-level lookup width (2^11 x 4 B = 8 KB of shared memory per CTA)
constexpr uint32_t DEC_LEAF = 0x80000000u;

struct DecTree { const uint32_t *L, *R, *C; const uint8_t *leaf; uint32_t nodeCount; const uint32_t *lut; };

// lut[v], v = the next DEC_K stream bits (MSB first): a codeword of length <= DEC_K that prefixes v -> LEAF | len << 16 | symbol;
// otherwise the index of the internal node reached after DEC_K bits (the walk continues from there bit by bit).
// Behind it, two tables of ALL the codewords that lie completely inside the window (short codes: several symbols per look-up):
//   lutS[v] (u32, at lut + 2^K):   bits consumed | start mask << 4 | zero-symbol start mask << 15   (bit j of a mask: a symbol starts at
//                                  window offset j); 0 when not even the first codeword fits.  For the synchronisation rounds, which
//                                  only need positions and counts.
//   lutW[v] (u64, at lut + 2^(K+1)): k | bits of the first k symbols << 3 | "a zero among them" << 7 | symbols, 8 bits each, from bit 8;
//                                  k <= 6 leading symbols that are all below 256; k = 0: none.  For the write pass.
constexpr int DEC_LUT_WORDS = (1 << DEC_K) * 4;                // u32 words of the three tables together
__global__ void dec_lut_kernel(const uint32_t *__restrict__ L, const uint32_t *__restrict__ R, const uint32_t *__restrict__ C,
                               const uint8_t *__restrict__ leaf, uint32_t *__restrict__ lut) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= (1u << DEC_K)) return;
    uint32_t *lutS = lut + (1u << DEC_K);
    unsigned long long *lutW = reinterpret_cast<unsigned long long *>(lut + (2u << DEC_K));
    uint32_t node = 0, e = 0;
    bool done = false;
    for (int d = 0; d < DEC_K; d++) {
        const uint32_t bit = (v >> (DEC_K - 1 - d)) & 1u;
        const uint32_t nx = bit ? R[node] : L[node];
        node = nx & ~DEC_CHILD_LEAF;
        if ((nx & DEC_CHILD_LEAF) || node == 0) { e = DEC_LEAF | ((uint32_t)(d + 1) << 16) | (C[node] & 0xffffu); done = true; break; }
    }
    lut[v] = done ? e : node;
    // every codeword inside the window, greedily from offset 0
    uint32_t smask = 0, zmask = 0, used = 0, kw = 0, usedw = 0, zw = 0;
    unsigned long long symsw = 0;
    bool wopen = true;
    for (uint32_t off = 0; off < (uint32_t)DEC_K;) {
        uint32_t nd = 0, len = 0; bool hit = false;
        for (uint32_t d = off; d < (uint32_t)DEC_K; d++) {
            const uint32_t bit = (v >> (DEC_K - 1 - d)) & 1u;
            const uint32_t nx = bit ? R[nd] : L[nd];
            nd = nx & ~DEC_CHILD_LEAF;
            if ((nx & DEC_CHILD_LEAF) || nd == 0) { len = d + 1 - off; hit = true; break; }
        }
        if (!hit) break;
        const uint32_t sym = C[nd];
        smask |= 1u << off; if (sym == 0) zmask |= 1u << off;
        used = off + len;
        if (wopen && kw < 6 && sym < 256u) { symsw |= (unsigned long long)sym << (8 * kw); kw++; usedw = used; if (sym == 0) zw = 1; } else wopen = false;
        off += len;
    }
    lutS[v] = smask ? (used | (smask << 4) | (zmask << 15)) : 0u;
    lutW[v] = kw ? ((unsigned long long)kw | ((unsigned long long)usedw << 3) | ((unsigned long long)zw << 7) | (symsw << 8)) : 0ull;
}

// MSB-first bit reader over the 4-byte aligned, zero-padded copy of the payload (big-endian words)
struct BitReader {
    const uint32_t *words; uint64_t wi; uint64_t buf; int nbits; uint64_t pos;
    __device__ __forceinline__ void init(const uint32_t *w, uint64_t p) {
        words = w; pos = p; wi = p >> 5;
        const int off = (int)(p & 31);
        buf = (((uint64_t)bswap32(__ldg(words + wi)) << 32) | (uint64_t)bswap32(__ldg(words + wi + 1))) << off;
        nbits = 64 - off; wi += 2;
    }
    __device__ __forceinline__ void refill() {               // keeps at least 32 valid bits in buf
        if (nbits < 32) { buf |= (uint64_t)bswap32(__ldg(words + wi)) << (32 - nbits); wi++; nbits += 32; }
    }
    __device__ __forceinline__ void consume(int n) { buf <<= n; nbits -= n; pos += (uint64_t)n; }
};

// decode ONE symbol at the reader's position (exactly the reference's bit-serial walk, ref: Huffman.c:310-343)
__device__ __forceinline__ uint32_t dec_symbol(BitReader &br, const uint32_t *s_lut, const DecTree &tr) {
    br.refill();
    const uint32_t e = s_lut[(uint32_t)(br.buf >> (64 - DEC_K))];
    if (e & DEC_LEAF) { br.consume((int)((e >> 16) & 63u)); return e & 0xffffu; }
    uint32_t node = e;
    br.consume(DEC_K);
    for (int guard = 0; guard < 64; guard++) {
        br.refill();
        const uint32_t bit = (uint32_t)(br.buf >> 63);
        br.consume(1);
        const uint32_t nx = bit ? tr.R[node] : tr.L[node];
        node = nx & ~DEC_CHILD_LEAF;
        if ((nx & DEC_CHILD_LEAF) || node == 0) break;
    }
    return tr.C[node];
}

// ---- D1: walk the section headers of a stream body on the device (SURVEY Appendix A from the block_size field on; ref:
// TightDataPointStorageF.c:54-126, szd_float.c:5750-5800) so that the host needs ONE small read-back instead of a dozen
// synchronous ones.  All multi-byte fields are read bytewise (the sections are not aligned).
struct DecHead {
    uint32_t err;                     // 0 ok, 1 truncated, 2 unsupported field
    uint32_t block_size, intervals, tree_size, node_count, use_mean, reg_count;
    unsigned long long eb_bits, mean_bits;                       // raw bits of the T-typed fields (big-endian decoded / native)
    unsigned long long o_tree, o_ind, o_unp, o_bits, total_unpred;
    unsigned long long prec_bits[4], o_ctree[4], o_cbits[4], o_cunp[4], cbits_bytes[4];
    uint32_t crad[4], ctree_size[4], cnode[4], cunp_cnt[4];
};
__device__ __forceinline__ unsigned long long rd_be(const unsigned char *b, int n) { unsigned long long v = 0; for (int i = 0; i < n; i++) v = (v << 8) | b[i]; return v; }
__device__ __forceinline__ unsigned long long rd_le(const unsigned char *b, int n) { unsigned long long v = 0; for (int i = n - 1; i >= 0; i--) v = (v << 8) | b[i]; return v; }
__global__ void __launch_bounds__(256) dec_parse_kernel(const unsigned char *__restrict__ body, unsigned long long body_len, int tsize /*sizeof(T)*/,
                                                        int ncoef, uint32_t nb, DecHead *out) {
    __shared__ DecHead H;
    __shared__ uint32_t s_cnt[8];
    const int t = threadIdx.x;
    unsigned long long o = 0;
    if (t == 0) {
        memset(&H, 0, sizeof(H));
        if (o + 4 + tsize + 12 > body_len) H.err = 1;
        else {
            H.block_size = (uint32_t)rd_be(body, 4); H.eb_bits = rd_be(body + 4, tsize); H.intervals = (uint32_t)rd_be(body + 4 + tsize, 4);
            H.tree_size = (uint32_t)rd_be(body + 8 + tsize, 4); H.node_count = (uint32_t)rd_be(body + 12 + tsize, 4);
            o = 4 + tsize + 12;
            H.o_tree = o; o += H.tree_size;
            const unsigned long long ind_bytes = (nb - 1) / 8 + 1;
            if (o + 1 + tsize + ind_bytes > body_len) H.err = 1;
            else { H.use_mean = body[o]; H.mean_bits = rd_le(body + o + 1, tsize); H.o_ind = o + 1 + tsize; }
        }
    }
    __syncthreads();
    if (H.err) { if (t == 0) *out = H; return; }
    // regression blocks = zero bits of the indicator (MSB-first; the pad bits of the last byte do not count)
    uint32_t zeros = 0;
    const unsigned char *ib = body + H.o_ind;
    for (uint32_t byte = t; byte < (nb + 7) / 8; byte += blockDim.x) {
        unsigned int v = ib[byte];
        const uint32_t valid = min(8u, nb - byte * 8u);
        v >>= (8 - valid);                                   // keep the `valid` leading bits
        zeros += valid - __popc(v);
    }
    for (int d = 16; d > 0; d >>= 1) zeros += __shfl_down_sync(0xffffffffu, zeros, d);
    if ((t & 31) == 0) s_cnt[t >> 5] = zeros;
    __syncthreads();
    if (t == 0) {
        uint32_t rc = 0; for (int w = 0; w < 8; w++) rc += s_cnt[w];
        H.reg_count = rc;
        o = H.o_ind + ((nb - 1) / 8 + 1);
        if (rc > 0) {
            for (int e = 0; e < ncoef && !H.err; e++) {
                if (o + tsize + 12 > body_len) { H.err = 1; break; }
                H.prec_bits[e] = rd_be(body + o, tsize); H.crad[e] = (uint32_t)rd_be(body + o + tsize, 4);
                H.ctree_size[e] = (uint32_t)rd_be(body + o + tsize + 4, 4); H.cnode[e] = (uint32_t)rd_be(body + o + tsize + 8, 4);
                o += tsize + 12;
                H.o_ctree[e] = o; o += H.ctree_size[e];
                if (o + 8 > body_len) { H.err = 1; break; }
                H.cbits_bytes[e] = rd_be(body + o, 8); o += 8;
                H.o_cbits[e] = o; o += H.cbits_bytes[e];
                if (o + 4 > body_len || o < H.o_cbits[e]) { H.err = 1; break; }
                H.cunp_cnt[e] = (uint32_t)rd_be(body + o, 4); o += 4;
                H.o_cunp[e] = o; o += (unsigned long long)H.cunp_cnt[e] * tsize;
            }
        }
        if (!H.err) {
            if (o + 8 > body_len) H.err = 1;
            else {
                H.total_unpred = rd_le(body + o, 8); o += 8;
                H.o_unp = o;
                if (H.total_unpred > body_len) H.err = 1; else o += H.total_unpred * tsize;
                if (o > body_len) H.err = 1;
                H.o_bits = o;
            }
        }
        *out = H;
    }
}

// one GLOBAL Jacobi round.  start_used[s] = start position used by the most recent decode of sub-sequence s.
// Inside the round every CTA iterates over its own 256 consecutive sub-sequences until none of them moves any more (the end
// positions travel through shared memory), so a correction entering a CTA is propagated as far as it goes within ONE launch;
// global rounds are only needed to carry corrections across CTA boundaries -- a handful instead of dozens.
// *changed is raised when the end of a CTA's last sub-sequence differs from what the next CTA saw in the previous round.
// All streams of one decompression (the main code stream and up to four coefficient streams) are synchronised by ONE launch
// per global round: a CTA finds its job from its block index.  (The coefficient streams are a few CTAs each with deep trees;
// launched one after the other they cost more than the main stream.)
struct DecJobDev {
    const uint32_t *words; uint64_t total_bits, nsub; DecTree tr;
    const unsigned long long *end_prev; unsigned long long *end_cur, *start_used; uint32_t *count, *zcount;
    unsigned int *changed; uint32_t cta_begin, active;
    unsigned long long *sym_off; uint64_t nsym; uint16_t *out;          // write pass of the coefficient streams
    unsigned long long *total;                                          // number of symbols the stream really holds (checked by the host)
};
constexpr int DEC_MAX_JOBS = 5;
struct DecJobList { DecJobDev j[DEC_MAX_JOBS]; int n; };
__global__ void __launch_bounds__(256) dec_sync_kernel(const DecJobList *__restrict__ list, int first_round) {
    int jid = 0;
    for (int q = 1; q < list->n; q++) if (blockIdx.x >= list->j[q].cta_begin) jid = q;
    const DecJobDev &J = list->j[jid];
    if (!J.active) return;
    const uint32_t *__restrict__ words = J.words; const uint64_t total_bits = J.total_bits, nsub = J.nsub; const DecTree tr = J.tr;
    const unsigned long long *__restrict__ end_prev = J.end_prev; unsigned long long *end_cur = J.end_cur, *start_used = J.start_used;
    uint32_t *count = J.count, *zcount = J.zcount; unsigned int *changed = J.changed;
    const uint32_t block = blockIdx.x - J.cta_begin;
    // A re-decode from a corrected start usually falls back onto the code-word boundaries of the previous decode after a few
    // symbols (self-synchronisation).  Every thread therefore keeps, for the 256 bit positions of its sub-sequence, one mask
    // of the symbol starts of its latest decode and one of the starts of zero symbols (shared memory).  A re-decode stops at the
    // first position whose old bit is set: from there on everything is known, and the counts follow from the masks.  The masks
    // only live inside one launch; a thread's first decode in a launch always runs to the end.
    __shared__ uint32_t s_lut[1 << DEC_K], s_lutS[1 << DEC_K];
    __shared__ unsigned long long s_end[256];
    __shared__ uint32_t s_bm[256][9], s_zm[256][9];            // 8 words + 1 of padding (bank spread)
    for (int q = threadIdx.x; q < (1 << DEC_K); q += blockDim.x) { s_lut[q] = tr.lut[q]; s_lutS[q] = tr.lut[(1 << DEC_K) + q]; }
    const int tid = threadIdx.x;
    const uint64_t sidx = (uint64_t)block * blockDim.x + tid;
    const bool active = sidx < nsub;
    const uint64_t lo = sidx * DEC_SUB_BITS;
    uint64_t hi = lo + DEC_SUB_BITS;
    const uint64_t limit = total_bits + 64;
    if (hi > limit) hi = limit;
    unsigned long long my_end = (active && !first_round) ? end_prev[sidx] : 0ull;
    unsigned long long my_start = (active && !first_round) ? start_used[sidx] : ~0ull;
    const unsigned long long old_end = my_end;
    // start handed in from the previous CTA: its last end of the previous global round (nothing is known in the first round)
    const unsigned long long ext_start = (sidx == 0) ? 0ull : ((first_round || !active) ? lo : end_prev[sidx - 1]);
    uint32_t cnt = 0, zc = 0;
    bool wrote = false, have_mask = false;
    uint32_t *bm = s_bm[tid], *zm = s_zm[tid];
    s_end[tid] = my_end;
    __syncthreads();
    for (int iter = 0; iter <= 256; iter++) {
        unsigned long long start;
        if (tid == 0 || sidx == 0) start = ext_start;
        else if (first_round && iter == 0) start = lo;
        else start = s_end[tid - 1];
        const bool did = active && start != my_start;
        if (did) {
            BitReader br; br.init(words, start);
            if (!have_mask) {
                // full decode; symbol starts are collected 32 positions at a time in registers
                cnt = 0; zc = 0;
#pragma unroll
                for (int w = 0; w < 8; w++) { bm[w] = 0u; zm[w] = 0u; }
                // two words of start bits at a time (a look-up covers DEC_K positions and may straddle a word)
                unsigned long long accb = 0, accz = 0; int curw = 0;
                while (br.pos < hi) {
                    const uint32_t off = (uint32_t)(br.pos - lo);
                    const int w = (int)(off >> 5);
                    while (curw < w) { bm[curw] = (uint32_t)accb; zm[curw] = (uint32_t)accz; accb >>= 32; accz >>= 32; curw++; }
                    br.refill();
                    const uint32_t e = s_lutS[(uint32_t)(br.buf >> (64 - DEC_K))];
                    const uint32_t sm = (e >> 4) & ((1u << DEC_K) - 1u);
                    if (sm) {
                        // all the code words inside the window at once; those that start at or behind `hi` belong to the next sub-sequence
                        const unsigned long long room = hi - br.pos;
                        const uint32_t m = room < (unsigned long long)DEC_K ? (sm & ((1u << (uint32_t)room) - 1u)) : sm;
                        const uint32_t z = (e >> 15) & m;
                        cnt += __popc(m); zc += __popc(z);
                        accb |= (unsigned long long)m << (off & 31u); accz |= (unsigned long long)z << (off & 31u);
                        const uint32_t rest = sm & ~m;
                        br.consume(rest ? (__ffs((int)rest) - 1) : (int)(e & 15u));
                    } else {
                        const uint32_t sym = dec_symbol(br, s_lut, tr);
                        accb |= 1ull << (off & 31u);
                        if (sym == 0) { accz |= 1ull << (off & 31u); zc++; }
                        cnt++;
                    }
                }
                bm[curw] = (uint32_t)accb; zm[curw] = (uint32_t)accz;
                if (curw + 1 < 8) { bm[curw + 1] = (uint32_t)(accb >> 32); zm[curw + 1] = (uint32_t)(accz >> 32); }
                my_end = br.pos;
                have_mask = true;
            } else {
                // re-decode until the new walk meets the old one.  `rem` / `remz` count the old symbol starts that the new walk
                // replaces: those before the new start, and those inside every newly decoded code word.
                uint32_t ncnt = 0, nzc = 0, rem = 0, remz = 0;
                {
                    const uint32_t so = (uint32_t)min((unsigned long long)(start - lo), (unsigned long long)DEC_SUB_BITS);
                    for (int q = 0; q < 8; q++) {
                        const uint32_t lo_q = (uint32_t)q << 5;
                        if (so <= lo_q) break;
                        const uint32_t m = (so - lo_q >= 32u) ? 0xffffffffu : ((1u << (so - lo_q)) - 1u);
                        rem += __popc(bm[q] & m); remz += __popc(zm[q] & m);
                        bm[q] &= ~m; zm[q] &= ~m;
                    }
                }
                bool synced = false;
                while (br.pos < hi) {
                    const uint32_t off = (uint32_t)(br.pos - lo);
                    const int w = (int)(off >> 5);
                    const uint32_t bit = 1u << (off & 31u);
                    if (bm[w] & bit) { synced = true; break; }
                    const uint32_t sym = dec_symbol(br, s_lut, tr);
                    ncnt++; nzc += (sym == 0);
                    // the new code word covers [off, off1): old starts inside it disappear, its own start is recorded
                    const uint32_t off1 = (uint32_t)min((unsigned long long)(br.pos - lo), (unsigned long long)DEC_SUB_BITS);
                    for (uint32_t o2 = off; o2 < off1;) {
                        const int w2 = (int)(o2 >> 5);
                        const uint32_t upto = min(off1, (uint32_t)(w2 + 1) << 5);
                        const uint32_t span = upto - o2;
                        const uint32_t m = (span == 32u ? 0xffffffffu : ((1u << span) - 1u)) << (o2 & 31u);
                        rem += __popc(bm[w2] & m); remz += __popc(zm[w2] & m);
                        bm[w2] &= ~m; zm[w2] &= ~m;
                        o2 = upto;
                    }
                    bm[w] |= bit; if (sym == 0) zm[w] |= bit;
                }
                cnt = cnt - rem + ncnt; zc = zc - remz + nzc;
                if (!synced) my_end = br.pos;         // walked to the end: every old start was replaced (rem == old cnt)
            }
            my_start = start; wrote = true;
        }
        __syncthreads();                          // everybody has read its predecessor's end
        s_end[tid] = my_end;
        if (!__syncthreads_or(did ? 1 : 0)) break;
    }
    if (active) {
        end_cur[sidx] = my_end;
        if (wrote) { start_used[sidx] = my_start; count[sidx] = cnt; zcount[sidx] = zc; }
        if ((tid == 255 || sidx == nsub - 1) && (first_round || my_end != old_end)) *changed = 1u;
    }
}

// ---- final pass: decode each sub-sequence again and write its symbols (block-major index = sym_off[s] + m) ----
// Main-stream write pass, block-major output.  Every warp's 32 sub-sequences hold one contiguous run of symbols in the
// block-major code array; the warp decodes into shared memory and then writes the run with coalesced stores (a thread storing
// its own symbols one by one touches a different 32-byte sector per lane and per symbol: ~0.7 ms of sector traffic at 512^3).
// Unpredictable values (code 0, rare) are pre-placed into the output field at their raster position.
constexpr int DECW_TB = 128, DECW_CAP = 4096;               // symbols a warp can stage; a longer run (codes < 2 bits on average) is stored directly
constexpr size_t DECW_SMEM = 8u * (1u << DEC_K) + (size_t)(DECW_TB / 32) * DECW_CAP * 2u;   // dynamic shared memory of dec_write_bm_kernel
// A main tree that is one leaf with symbol 0 (ref: Huffman.c:316-321 fills the type array with the root's symbol whatever it is):
// every point is unpredictable, the list of raw values IS the field in block-major order -- each thread places 16 of them.
template <typename T>
__global__ void __launch_bounds__(256) unpred_all_kernel(uint64_t n, Geom g, const T *__restrict__ unpred, uint64_t unpred_count, T *field_out) {
    const uint64_t p0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16u;
    if (p0 >= n) return;
    BMIter it; it.init(g, p0);
    const int cnt = (int)min((uint64_t)16, n - p0);
    for (int q = 0; q < cnt; q++) { if (p0 + q < unpred_count) field_out[it.raster] = unpred[p0 + q]; it.next(g); }
}

template <typename T>
__global__ void __launch_bounds__(DECW_TB) dec_write_bm_kernel(const uint32_t *__restrict__ words, uint64_t total_bits, uint64_t nsub, DecTree tr,
                                                              const unsigned long long *__restrict__ start_used,
                                                              const unsigned long long *__restrict__ sym_off, const unsigned long long *__restrict__ zero_off,
                                                              uint64_t n, Geom g, uint16_t *__restrict__ codes_bm, const T *__restrict__ unpred,
                                                              uint64_t unpred_count, T *field_out) {
    __shared__ uint32_t s_lut[1 << DEC_K];
    extern __shared__ __align__(16) unsigned char s_dyn_w[];    // DECW_SMEM bytes: the multi-symbol table, then the warps' staging areas
    unsigned long long *s_lutW = reinterpret_cast<unsigned long long *>(s_dyn_w);
    uint16_t (*s_sym)[DECW_CAP] = reinterpret_cast<uint16_t (*)[DECW_CAP]>(s_dyn_w + 8u * (1u << DEC_K));
    {
        const unsigned long long *gW = reinterpret_cast<const unsigned long long *>(tr.lut + (2u << DEC_K));
        for (int q = threadIdx.x; q < (1 << DEC_K); q += blockDim.x) { s_lut[q] = tr.lut[q]; s_lutW[q] = gW[q]; }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t sidx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t sfirst = sidx - lane;
    if (sfirst >= nsub) return;                               // whole warp beyond the stream
    const bool have = sidx < nsub;
    uint64_t p = have ? min((uint64_t)sym_off[sidx], n) : n;
    const uint64_t wbase = __shfl_sync(0xffffffffu, p, 0);
    const uint64_t wend = (sfirst + 32 < nsub) ? min((uint64_t)sym_off[sfirst + 32], n) : n;
    const bool staged = (wend - wbase) <= (uint64_t)DECW_CAP;
    uint16_t *ss = s_sym[warp];
    if (have && p < n) {
        uint64_t zr = zero_off[sidx];
        uint64_t hi = (sidx + 1) * DEC_SUB_BITS;
        const uint64_t limit = total_bits + 64;
        if (hi > limit) hi = limit;
        BitReader br; br.init(words, start_used[sidx]);
        // The raster position is only needed where an unpredictable value goes (code 0, rare): the block-major walk is not advanced
        // per symbol but caught up there -- stepped when the previous zero is near, set up anew otherwise.
        BMIter it; uint64_t it_p = ~0ull;
        while (br.pos < hi && p < n) {
            br.refill();
            const unsigned long long e = s_lutW[(uint32_t)(br.buf >> (64 - DEC_K))];
            const uint32_t k = (uint32_t)e & 7u;
            if (k != 0u && !(e & 0x80ull) && br.pos + (uint64_t)DEC_K <= hi && p + k <= n) {
                // up to six short code words in one look-up (none of them the zero code, all of them inside this sub-sequence)
#pragma unroll
                for (uint32_t j = 0; j < 6; j++) {
                    if (j < k) { const uint16_t sy = (uint16_t)((e >> (8u + 8u * j)) & 0xffull); if (staged) ss[p - wbase + j] = sy; else codes_bm[p + j] = sy; }
                }
                p += k;
                br.consume((int)((e >> 3) & 15ull));
                continue;
            }
            const uint32_t sym = dec_symbol(br, s_lut, tr);
            if (staged) ss[p - wbase] = (uint16_t)sym; else codes_bm[p] = (uint16_t)sym;
            if (sym == 0 && field_out) {
                if (it_p == ~0ull || p - it_p > 48u) it.init(g, p); else for (uint64_t q = it_p; q < p; q++) it.next(g);
                it_p = p;
                if (zr < unpred_count) field_out[it.raster] = unpred[zr];
                zr++;
            }
            p++;
        }
    }
    __syncwarp();
    if (staged) {
        const uint32_t cnt = (uint32_t)(wend - wbase);
        for (uint32_t q = lane; q < cnt; q += 32) codes_bm[wbase + q] = ss[q];
    }
}

// prefix sums and write pass of ALL coefficient streams of a call in one launch each (they are a few CTAs apiece; launched
// separately their launch latencies add up to more than the work)
__global__ void __launch_bounds__(1024) dec_coef_scan_kernel(const DecJobList *__restrict__ list) {
    const DecJobDev &J = list->j[blockIdx.x];
    if (!J.active) return;
    __shared__ unsigned long long s_w[32];
    __shared__ unsigned long long s_carry;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) s_carry = 0;
    __syncthreads();
    for (uint64_t base = 0; base < J.nsub; base += 1024) {
        const uint64_t i = base + t;
        const unsigned long long v = i < J.nsub ? (unsigned long long)J.count[i] : 0ull;
        unsigned long long x = v;
        for (int d = 1; d < 32; d <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) s_w[warp] = x;
        __syncthreads();
        unsigned long long before = 0, tot = 0;
        for (int w = 0; w < 32; w++) { const unsigned long long c = s_w[w]; before += (w < warp) ? c : 0ull; tot += c; }
        const unsigned long long carry = s_carry;
        if (i < J.nsub) J.sym_off[i] = carry + before + x - v;
        __syncthreads();
        if (t == 0) s_carry = carry + tot;
        __syncthreads();
    }
    if (t == 0 && J.total) *J.total = s_carry;
}
__global__ void __launch_bounds__(256) dec_coef_write_kernel(const DecJobList *__restrict__ list) {
    int jid = 0;
    for (int q = 1; q < list->n; q++) if (blockIdx.x >= list->j[q].cta_begin) jid = q;
    const DecJobDev &J = list->j[jid];
    if (!J.active) return;
    const DecTree tr = J.tr;
    __shared__ uint32_t s_lut[1 << DEC_K];
    for (int q = threadIdx.x; q < (1 << DEC_K); q += blockDim.x) s_lut[q] = tr.lut[q];
    __syncthreads();
    const uint64_t sidx = (uint64_t)(blockIdx.x - J.cta_begin) * blockDim.x + threadIdx.x;
    if (sidx >= J.nsub) return;
    uint64_t p = J.sym_off[sidx];
    if (p >= J.nsym) return;
    uint64_t hi = (sidx + 1) * DEC_SUB_BITS;
    const uint64_t limit = J.total_bits + 64;
    if (hi > limit) hi = limit;
    BitReader br; br.init(J.words, J.start_used[sidx]);
    while (br.pos < hi && p < J.nsym) { J.out[p] = (uint16_t)dec_symbol(br, s_lut, tr); p++; }
}

// single-leaf tree: every symbol is the root's symbol and the payload is empty (ref: Huffman.c:316-321)
__global__ void fill_u16_kernel(uint16_t *out, uint64_t n, uint16_t v) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) out[i] = v;
}

// coefficient chain decode (S1, decoder side; ref: szd_float.c:5809-5819): coeff_chain_decode_kernel in szb_chain.cuh


// indicator bits (MSB-first, 1 = Lorenzo; ref: TypeManager.c:47-71) <-> byte-per-block array
__global__ void indicator_pack_kernel(const unsigned char *__restrict__ ind, uint32_t nb, unsigned char *out) {
    uint32_t byte = blockIdx.x * blockDim.x + threadIdx.x;
    if (byte >= (nb + 7) / 8) return;
    unsigned int v = 0;
    for (int j = 0; j < 8; j++) { uint32_t b = byte * 8 + j; if (b < nb && ind[b] == 1) v |= 1u << (7 - j); }
    out[byte] = (unsigned char)v;
}
__global__ void indicator_unpack_kernel(const unsigned char *__restrict__ in, uint32_t nb, unsigned char *ind) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb) ind[b] = (in[b >> 3] >> (7 - (b & 7))) & 1;
}

}  // namespace szb
