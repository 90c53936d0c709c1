// heap_pipe_sim.cpp -- development model of the pipelined Huffman heap (szb_huffman.cuh, "lanes as operations").
//
// The reference builds its Huffman tree with a binary heap whose exact array layout decides the ties (ref: sz/src/Huffman.c:76-114,
// :165-185), so the build has to be replayed operation by operation.  On the GPU one warp replays it with up to 32 heap operations
// in flight, one per lane, every lane one level per "tick": a removal's sift-down reads level l+1 and writes level l, an insertion
// (done top-down along the ancestor chain of the slot it is appended at) reads and writes level l.  Operations are launched in
// program order and far enough apart that each one only ever reads levels the older ones have finished with.
//
// This file models that schedule on the CPU, lane by lane and tick by tick (all loads of a tick before all of its stores, as the
// warp executes them), and checks the resulting tree against the plain sequential heap on synthetic and recorded histograms.
// It is a tool, not part of the library.   g++ -O2 -o /tmp/heap_pipe_sim tools/heap_pipe_sim.cpp && /tmp/heap_pipe_sim [hist.bin ...]
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <random>
#include <vector>

struct Ent { uint32_t k, n; };

// ---- the sequential reference ---------------------------------------------------------------------------------
static void ref_build(const std::vector<uint32_t> &freq, std::vector<uint32_t> &left, std::vector<uint32_t> &right, uint32_t &root) {
    const uint32_t L = (uint32_t)freq.size();
    std::vector<Ent> q(L + 2);
    uint32_t qend = 1;
    auto ins = [&](Ent e) { uint32_t i = qend++, j; while ((j = i >> 1)) { if (q[j].k <= e.k) break; q[i] = q[j]; i = j; } q[i] = e; };
    auto rem = [&]() {
        Ent top = q[1]; qend--; Ent m = q[qend]; uint32_t i = 1, l;
        while ((l = 2 * i) < qend) { if (l + 1 < qend && q[l + 1].k < q[l].k) l++; if (m.k > q[l].k) { q[i] = q[l]; i = l; } else break; }
        q[i] = m; return top;
    };
    for (uint32_t i = 0; i < L; i++) ins({freq[i], i});
    left.assign(2 * L, 0xffffffffu); right.assign(2 * L, 0xffffffffu);
    uint32_t nn = L;
    while (qend > 2) { Ent b = rem(); Ent a = rem(); left[nn] = a.n; right[nn] = b.n; ins({a.k + b.k, nn}); nn++; }
    root = q[1].n;
}

// ---- the lane model -----------------------------------------------------------------------------------------------------------
static inline int lvl(uint32_t i) { return 31 - __builtin_clz(i); }
static inline bool anc(uint32_t h, uint32_t x) { const int d = lvl(x) - lvl(h); return d >= 0 && (x >> d) == h; }   // h is x or an ancestor of x

// Only the removals are lane operations.  The insertion of a merge is done in ONE tick by the whole warp (every lane looks at one
// ancestor of the new slot, a vote finds how many of them move down: ref Huffman.c:76-85 is bottom-up), as soon as no removal still
// in flight has its hole on that ancestor chain; the element the insertion would leave in the last slot is handed straight to the
// next merge's first removal, which takes the last slot anyway.
struct Lane {
    bool act = false;
    uint32_t op = 0;                    // sequence number (age)
    uint32_t h = 0, bound = 0;          // hole, last valid slot
    Ent e{0, 0};                        // the element that sinks
    uint32_t x = 0; bool stale = false; // slot the element was read from; an older removal could still have landed there after the read
};

struct Pipe {
    std::vector<Ent> q; Lane ln[32];
    uint64_t ticks = 0, stall_elem = 0, stall_ins = 0, stall_lane = 0, stall_freeze = 0, refetch = 0, refetch_changed = 0;
    Lane snap[32];                                           // the lanes as they were when the tick began (what a vote at the top of the tick sees)
    bool older_on(uint32_t x, uint32_t op, bool at_tick_start = false) const {   // a removal older than `op` could still write slot x
        for (int l = 0; l < 32; l++) { const Lane &a = at_tick_start ? snap[l] : ln[l]; if (a.act && a.op < op && x <= a.bound && anc(a.h, x)) return true; }
        return false;
    }
    bool chain_busy(uint32_t s) const {                      // a removal in flight has its hole on the ancestor chain of slot s
        for (int l = 0; l < 32; l++) { const Lane &a = ln[l]; if (a.act && anc(a.h, s)) return true; }
        return false;
    }
    // one tick; returns true if a removal had to wait for its element (it and everything younger did not move)
    bool step() {
        ticks++;
        for (int l = 0; l < 32; l++) snap[l] = ln[l];
        struct St { uint32_t at; Ent v; };
        std::vector<St> st;
        uint32_t frozen_from = 0xffffffffu;
        // the lanes in age order (the model only; on the GPU the vote below gives the same result)
        int order[32]; int n = 0;
        for (int l = 0; l < 32; l++) if (ln[l].act) order[n++] = l;
        for (int i = 0; i < n; i++) for (int j = i + 1; j < n; j++) if (ln[order[j]].op < ln[order[i]].op) { int t = order[i]; order[i] = order[j]; order[j] = t; }
        for (int i = 0; i < n; i++) {
            Lane &a = ln[order[i]];
            if (a.op >= frozen_from) break;
            const uint32_t c0 = 2 * a.h;
            bool move = false; uint32_t c = 0;
            if (c0 <= a.bound) { const bool r = c0 + 1 <= a.bound && q[c0 + 1].k < q[c0].k; c = c0 + (r ? 1 : 0); move = a.e.k > q[c].k; }
            if (!move && a.stale) {
                // about to stop with an element that may have been replaced: sure only when no older removal can reach the slot any more
                if (older_on(a.x, a.op, true)) { frozen_from = a.op; break; }
                refetch++;
                const Ent t = q[a.x];
                a.stale = false;
                if (t.k != a.e.k || t.n != a.e.n) { refetch_changed++; a.e = t; if (c0 <= a.bound) move = a.e.k > q[c].k; }
            }
            if (move) { st.push_back({a.h, q[c]}); a.h = c; }
            else { st.push_back({a.h, a.e}); a.act = false; }
        }
        for (auto &s : st) q[s.at] = s.v;
        return frozen_from != 0xffffffffu;
    }
    void tick() { while (step()) stall_freeze++; }
    bool any() const { for (int l = 0; l < 32; l++) if (ln[l].act) return true; return false; }
};

static void pipe_build(const std::vector<uint32_t> &freq, std::vector<uint32_t> &left, std::vector<uint32_t> &right, uint32_t &root, Pipe &P, uint32_t mtail, bool late_fix) {
    const uint32_t L = (uint32_t)freq.size();
    P.q.assign(L + 2, {0, 0});
    std::vector<Ent> &q = P.q;
    uint32_t qend = 1;
    auto ins = [&](Ent e) { uint32_t i = qend++, j; while ((j = i >> 1)) { if (q[j].k <= e.k) break; q[i] = q[j]; i = j; } q[i] = e; };
    for (uint32_t i = 0; i < L; i++) ins({freq[i], i});
    left.assign(2 * L, 0xffffffffu); right.assign(2 * L, 0xffffffffu);
    uint32_t m = L, k = 0, op = 0;                           // m: elements in the heap; the last one is held in e1, not in its slot
    Ent e1 = q[m], a = q[1];
    bool have_e1 = m >= mtail;
    for (; m >= mtail; m--, k++) {
        const int la = (2 * k) & 31, lb = (2 * k + 1) & 31;
        // first removal (element and root were left by the previous merge's insertion, or by the leaf phase): launched right
        // before the tick that is its first level (which writes the root)
        while (P.ln[la].act) { P.tick(); P.stall_lane++; }
        { Lane &A = P.ln[la]; A = Lane(); A.act = true; A.op = op++; A.h = 1; A.bound = m - 1; A.e = e1; A.x = m; A.stale = false; }
        P.tick();
        P.tick();                                            // its second level: level 1 is final for the next removal
        // second removal: its element is the last slot's, as soon as no older removal can still land there
        while (P.ln[lb].act) { P.tick(); P.stall_lane++; }
        bool st2 = false;
        if (late_fix) st2 = P.older_on(m - 1, op); else while (P.older_on(m - 1, op)) { P.tick(); P.stall_elem++; }
        const Ent b = q[1]; const Ent e2 = q[m - 1];
        { Lane &B = P.ln[lb]; B = Lane(); B.act = true; B.op = op++; B.h = 1; B.bound = m - 2; B.e = e2; B.x = m - 1; B.stale = st2; }
        P.tick();
        P.tick();
        // the insertion, by the whole warp in one go, once no removal has its hole on the chain; then the next merge
        const uint32_t nn = L + k, s = m - 1;
        left[nn] = b.n; right[nn] = a.n;                     // first removed -> right, second -> left (gcc, Huffman.c:181)
        while (P.chain_busy(s)) { P.tick(); P.stall_ins++; }
        {
            const Ent nw{a.k + b.k, nn};
            uint32_t i = s, j; Ent out = nw; bool first = true;
            while ((j = i >> 1)) { if (q[j].k <= nw.k) break; if (first) { out = q[j]; first = false; } else q[i] = q[j]; i = j; }
            if (!first) q[i] = nw;                            // (when nothing moves the new node itself is what the next removal takes)
            e1 = out; a = q[1];
        }
    }
    while (P.any()) P.tick();
    // the rest sequentially
    if (have_e1) q[m] = e1;
    qend = m + 1;
    auto rem = [&]() {
        Ent top = q[1]; qend--; Ent mm = q[qend]; uint32_t i = 1, l;
        while ((l = 2 * i) < qend) { if (l + 1 < qend && q[l + 1].k < q[l].k) l++; if (mm.k > q[l].k) { q[i] = q[l]; i = l; } else break; }
        q[i] = mm; return top;
    };
    uint32_t nn = L + k;
    while (qend > 2) { Ent b = rem(); Ent a2 = rem(); left[nn] = a2.n; right[nn] = b.n; ins({a2.k + b.k, nn}); nn++; }
    root = q[1].n;
}

static bool check1(const char *name, const std::vector<uint32_t> &freq, uint32_t mtail, bool late_fix) {
    std::vector<uint32_t> l0, r0, l1, r1; uint32_t root0, root1;
    ref_build(freq, l0, r0, root0);
    Pipe P; pipe_build(freq, l1, r1, root1, P, mtail, late_fix);
    const bool ok = root0 == root1 && l0 == l1 && r0 == r1;
    const double merges = (double)freq.size() - 1;
    printf("%-24s %s leaves=%6zu %s  ticks/merge=%.2f  waits/merge: insertion %.3f, element %.3f (while sinking %.3f), lane busy %.3f; re-read %.3f (changed %.4f)\n", name, late_fix ? "late-fix" : "waiting ", freq.size(), ok ? "ok " : "MISMATCH",
           P.ticks / merges, P.stall_ins / merges, P.stall_elem / merges, P.stall_freeze / merges, P.stall_lane / merges, P.refetch / merges, P.refetch_changed / merges);
    return ok;
}

static bool check(const char *name, const std::vector<uint32_t> &freq, uint32_t mtail = 8) { const bool a = check1(name, freq, mtail, false); const bool b = check1(name, freq, mtail, true); return a && b; }

int main(int argc, char **argv) {
    bool ok = true;
    std::mt19937 rng(12345);
    for (int it = 0; it < 60; it++) {
        const uint32_t L = 2 + rng() % 3000;
        std::vector<uint32_t> f(L);
        const int kind = it % 6;
        for (uint32_t i = 0; i < L; i++) {
            if (kind == 0) f[i] = 1;
            else if (kind == 1) f[i] = 1 + rng() % 3;
            else if (kind == 2) f[i] = 1 + rng() % 1000000;
            else if (kind == 3) { const double x = (i - L / 2.0) / (L / 6.0); f[i] = 1 + (uint32_t)(40 * exp(-x * x / 2) + (rng() % 3)); }
            else if (kind == 4) f[i] = 1u << (rng() % 20);
            else f[i] = i + 1;
        }
        char nm[64]; snprintf(nm, sizeof(nm), "synthetic kind %d", kind);
        ok &= check(nm, f, 8 + it % 5);
    }
    for (uint32_t L = 2; L < 40; L++) { std::vector<uint32_t> f(L); for (auto &v : f) v = 1 + rng() % 4; ok &= check("tiny", f); }
    for (int a = 1; a < argc; a++) {
        FILE *fp = fopen(argv[a], "rb"); if (!fp) { perror(argv[a]); return 2; }
        std::vector<uint32_t> h(1 << 20); const size_t n = fread(h.data(), 4, h.size(), fp); fclose(fp);
        std::vector<uint32_t> f; for (size_t i = 0; i < n; i++) if (h[i]) f.push_back(h[i]);
        ok &= check(argv[a], f);
    }
    printf(ok ? "ALL OK\n" : "FAILURES\n");
    return ok ? 0 : 1;
}
