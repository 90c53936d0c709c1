// szb_noalloc.cu -- one extra C-ABI entry point that never allocates: szb200_range_noalloc().
//
// Who needs it: host codes that hand every byte of HBM to their own arena at start-up (so a later cudaMalloc can fail or
// fragment the arena), and service / I-O threads that must not trigger the device-wide synchronisation that cudaMalloc and
// cudaFree imply while the solver's kernels are in flight.  Such callers cannot use szb200_range(), whose context keeps a
// grow-only workspace obtained with cudaMalloc.  Here the per-CTA partial results and the completion counter of the same
// reduction kernel (range_kernel, szb_stage1.cuh) live in the fixed file-scope device array below; the caller supplies the
// input and a two-element device output, and may supply the stream.
//
// Cost and limits: the array (64 KiB + 256 B) is part of this library's device image, so every GPU that loads the library
// holds it until the process exits.  One CTA reduces one tile of 512 threads x 16 128-bit loads, and there are 4096 partial
// slots, which bounds one call to 512 MiB of input (2^27 floats or 2^26 doubles); larger requests are refused.
// All calls on a device share the array: it still holds the previous call's partials (harmless, every slot a launch reads
// is rewritten by that launch first) and the completion counter, which is cleared before every launch.  Two calls whose
// kernels overlap on the GPU (different streams, or different host threads) race on it -- serialise them.
// Every system / CUDA header the kernel headers below include is included HERE first, at global scope: the kernel headers are
// pulled into an anonymous namespace further down, and a system header that had not been seen yet would be declared inside it.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "../../include/szb200.h"

// This file is a second translation unit of the same library: the shared kernel headers are pulled in with internal
// linkage so that their non-template kernels are not defined twice at link time.
namespace {
#include "szb_stage1.cuh"
}  // namespace

using namespace szb;

namespace {

constexpr size_t NA_SLOTS = 4096;                 // partial-result slots (one per CTA)
constexpr size_t NA_COUNTER_BYTES = 256;          // completion counter, kept on its own 256-byte line
constexpr size_t NA_SCRATCH_BYTES = NA_COUNTER_BYTES + NA_SLOTS * sizeof(MinMax<double>);
constexpr size_t NA_TILE_VECS = 512 * 16;         // 128-bit loads per CTA
constexpr size_t NA_MAX_INPUT_BYTES = NA_SLOTS * NA_TILE_VECS * 16;   // 512 MiB
constexpr int NA_MAX_DEVICES = 64;

// Device-side storage.  In HOST code this name is only a stand-in for the symbol, never a device address: the address is
// looked up once per device with cudaGetSymbolAddress() below and that pointer is what the kernel receives.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

std::mutex g_na_mu;
void *g_na_ptr[NA_MAX_DEVICES];                   // per-device address of g_noalloc_scratch (nullptr = not looked up yet)

// once per device, with that device current
int na_scratch(void **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { fprintf(stderr, "szb200_range_noalloc: cudaGetDevice failed: %s\n", cudaGetErrorString(e)); return -1; }
    if (dev < 0 || dev >= NA_MAX_DEVICES) { fprintf(stderr, "szb200_range_noalloc: device index %d is outside the supported 0..%d\n", dev, NA_MAX_DEVICES - 1); return -1; }
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (g_na_ptr[dev] == nullptr) {
        void *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            fprintf(stderr, "szb200_range_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            return -1;
        }
        g_na_ptr[dev] = p;
    }
    *out = g_na_ptr[dev];
    return 0;
}

template <typename T>
int na_launch(const T *d_data, size_t n, T *d_out, cudaStream_t stream, unsigned char *scratch) {
    const size_t nvec = (n * sizeof(T) + 15) / 16;
    size_t grid = (nvec + NA_TILE_VECS - 1) / NA_TILE_VECS;
    if (grid == 0) grid = 1;
    unsigned int *done = reinterpret_cast<unsigned int *>(scratch);
    MinMax<T> *partials = reinterpret_cast<MinMax<T> *>(scratch + NA_COUNTER_BYTES);
    // the counter must start at zero; the partial slots need no clearing (each CTA writes its slot before the last CTA reads it)
    cudaError_t e = cudaMemsetAsync(done, 0, NA_COUNTER_BYTES, stream);
    if (e != cudaSuccess) { fprintf(stderr, "szb200_range_noalloc: clearing the counter failed: %s\n", cudaGetErrorString(e)); return -1; }
    range_kernel<T><<<(unsigned int)grid, 512, 0, stream>>>(d_data, (uint64_t)n, partials, done, reinterpret_cast<MinMax<T> *>(d_out));
    e = cudaGetLastError();
    if (e != cudaSuccess) { fprintf(stderr, "szb200_range_noalloc: kernel launch failed: %s\n", cudaGetErrorString(e)); return -1; }
    return 0;
}

}  // namespace

extern "C" int szb200_range_noalloc(int dtype, const void *d_data, size_t n, void *d_minmax_out, void *stream) {
    if (dtype != 0 && dtype != 1) { fprintf(stderr, "szb200_range_noalloc: dtype must be SZ_FLOAT (0) or SZ_DOUBLE (1), got %d\n", dtype); return -3; }
    if (d_data == nullptr || d_minmax_out == nullptr || n == 0) { fprintf(stderr, "szb200_range_noalloc: null pointer or empty input\n"); return -1; }
    const size_t es = dtype == 0 ? 4 : 8;
    if (n > NA_MAX_INPUT_BYTES / es) {
        fprintf(stderr, "szb200_range_noalloc: %zu elements (%zu bytes) exceed the fixed scratch of this entry point; the limit is %zu bytes (%zu elements of this type) per call\n",
                n, n * es, NA_MAX_INPUT_BYTES, NA_MAX_INPUT_BYTES / es);
        return -4;
    }
    void *scratch = nullptr;
    if (na_scratch(&scratch) != 0) return -1;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == 0) return na_launch<float>(static_cast<const float *>(d_data), n, static_cast<float *>(d_minmax_out), s, static_cast<unsigned char *>(scratch));
    return na_launch<double>(static_cast<const double *>(d_data), n, static_cast<double *>(d_minmax_out), s, static_cast<unsigned char *>(scratch));
}
