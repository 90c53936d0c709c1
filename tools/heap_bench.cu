// heap_bench.cu -- stand-alone timing and check of the pipelined Huffman heap (sz_b200/csrc/szb_heap_pipe.cuh) on one histogram.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I sz_b200/csrc -o /tmp/heap_bench tools/heap_bench.cu
//   /tmp/heap_bench hist.bin [threads per CTA]        (hist.bin: uint32 counts per symbol; without a file: a synthetic bell)
// Prints cycles per phase and per merge, and compares every node's {parent, pre-order step} record with the sequential heap on the host.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "szb_heap_pipe.cuh"

using namespace szb;

__global__ void bench_kernel(const uint32_t *lfreq, uint32_t leaves, uint32_t *up, uint32_t *loga, uint32_t *logb, uint32_t *out, unsigned long long *cyc) {
    extern __shared__ __align__(16) uint32_t s_heap[];
    if (threadIdx.x < 32) {
        const long long t0 = clock64();
        const uint32_t r = hp_build((unsigned)__cvta_generic_to_shared(s_heap), leaves, lfreq, up, loga, logb, out, t0);
        if (threadIdx.x == 0) { cyc[0] = (unsigned long long)(clock64() - t0); out[0] = r & 0xffffu; out[1] = r >> 16; }
    }
    __syncthreads();
    __threadfence();
    for (uint32_t k = threadIdx.x; k < out[1]; k += blockDim.x) hp_apply_log(up, loga, logb, leaves, k);   // the logged merges -> `up` records
}

struct Ent { uint32_t k, n; };
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

int main(int argc, char **argv) {
    std::vector<uint32_t> f;
    if (argc > 1 && argv[1][0] != '-') {
        FILE *fp = fopen(argv[1], "rb"); if (!fp) { perror(argv[1]); return 2; }
        std::vector<uint32_t> h(1 << 20); const size_t n = fread(h.data(), 4, h.size(), fp); fclose(fp);
        for (size_t i = 0; i < n; i++) if (h[i]) f.push_back(h[i]);
    } else {
        const uint32_t L = 19000; uint32_t s = 12345;
        for (uint32_t i = 0; i < L; i++) { s = s * 1664525u + 1013904223u; const double x = (i - L / 2.0) / (L / 6.0); f.push_back(1 + (uint32_t)(24 * exp(-x * x / 2) + ((s >> 16) % 3))); }
    }
    const int threads = argc > 2 ? atoi(argv[2]) : 1024;
    const uint32_t L = (uint32_t)f.size();
    std::vector<uint32_t> l0, r0; uint32_t root0; ref_build(f, l0, r0, root0);
    // the records hp_build leaves: {parent:16 | 1 (left child) or 1 + nodes of the left sibling's subtree (right child) : 16}
    std::vector<uint32_t> size(2 * L, 1), up0(2 * L, 0xffffffffu);
    for (uint32_t n = L; n + 1 < 2 * L; n++) { size[n] = 1 + size[l0[n]] + size[r0[n]]; up0[l0[n]] = n | (1u << 16); up0[r0[n]] = n | ((1 + size[l0[n]]) << 16); }
    uint32_t *d_f, *d_l, *d_la, *d_lb, *d_out; unsigned long long *d_c;
    cudaMalloc(&d_f, 4 * L); cudaMalloc(&d_l, 8 * L); cudaMalloc(&d_la, 4 * L); cudaMalloc(&d_lb, 4 * L); cudaMalloc(&d_out, 64); cudaMalloc(&d_c, 64);
    cudaMemcpy(d_f, f.data(), 4 * L, cudaMemcpyHostToDevice);
    const size_t smem = 8ull * (L + 4);
    cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int rep = 0; rep < 3; rep++) {
        cudaMemset(d_l, 0xee, 8 * L); cudaMemset(d_out, 0, 64);
        bench_kernel<<<1, threads, smem>>>(d_f, L, d_l, d_la, d_lb, d_out, d_c);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
        std::vector<uint32_t> up1(2 * L); uint32_t out[16]; unsigned long long c[8];
        cudaMemcpy(up1.data(), d_l, 8 * L, cudaMemcpyDeviceToHost);
        cudaMemcpy(out, d_out, 64, cudaMemcpyDeviceToHost); cudaMemcpy(c, d_c, 64, cudaMemcpyDeviceToHost);
        size_t bad = 0; for (uint32_t n = 0; n + 1 < 2 * L; n++) if (up0[n] != up1[n]) bad++;
        const double leafc = out[5] * 1024.0, allc = out[6] * 1024.0;
        printf("leaves %u threads %d: %s (root %u/%u, %zu nodes differ)  leaf phase %.0f cycles (%.1f per leaf), merges %.0f cycles (%.1f per merge), total %llu\n",
               L, threads, bad == 0 && out[0] == root0 ? "ok" : "MISMATCH", out[0], root0, bad, leafc, leafc / L, allc - leafc, (allc - leafc) / (L - 1), c[0]);
#ifdef HP_COUNT
        printf("   ticks %u (%.2f per merge), stall ticks: element %u, chain %u\n", out[8], out[8] / (double)(L - 1), out[9], out[10]);
#endif
    }
    return 0;
}
