// noalloc.cu -- nsb_max_mat_noalloc(): the max_mat reduction (src/serial/io.c:122-139 ; find_max_kernel + max_mat_cuda,
// src/parallel/main.cu:799-832) for callers that must not allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Nothing else in the library calls into this file and it calls nothing that allocates.
//
// The contexts' reduction (momentum.cu: update_uv_kernel<.., UPDATE = false>) keeps its per-row partial maxima in a
// cudaMalloc'ed buffer of the context.  This variant keeps them in a FILE-SCOPE device array of fixed size:
//   - every GPU that loads libnsb200.so holds NSB_NOALLOC_SCRATCH_BYTES (128.25 KiB) of it until the process exits;
//   - grids beyond NSB_NOALLOC_MAX_ROWS x NSB_NOALLOC_MAX_COLS interior cells are refused (NSB_ERR_ARG, text in
//     nsb_last_error(NULL));
//   - every call on a device shares the one array: two calls that OVERLAP on that GPU (different streams, or a graph
//     replayed concurrently with itself) race on the partials and on the ticket.  Calls ordered on one stream are safe;
//   - the array keeps the last call's contents.  The only word that must start at zero is the block ticket: it is
//     cleared by a cudaMemsetAsync in front of every launch (a memset node under capture), and the partials are all
//     written before they are read (one per launched block), so stale values are never seen.
// In HOST code the array's name is only a stand-in for the symbol, not a device address: the address is looked up with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernel is given.
#include <math.h>
#include <stdio.h>

#include <mutex>

#include "nsb_internal.h"
#include "nsb_math.cuh"

#define NSB_NOALLOC_MAX_ROWS 16384          // i_max: one partial (8 B) per grid row
#define NSB_NOALLOC_MAX_COLS 16384          // j_max: no scratch per column; stated so that the limit is a grid size
#define NSB_NOALLOC_HEAD 32                 // doubles in front of the partials: the ticket's own 256 bytes
#define NSB_NOALLOC_SCRATCH_BYTES ((NSB_NOALLOC_HEAD + NSB_NOALLOC_MAX_ROWS) * sizeof(double))
#define NSB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double nsb_noalloc_scratch[NSB_NOALLOC_HEAD + NSB_NOALLOC_MAX_ROWS];

// One block per grid row i = 1 .. i_max (the reference's first index), 256 threads over j = 1 .. j_max; the last block
// to take a ticket combines the row maxima and applies the reference's seed: the scan starts from m[0][0] and replaces
// it only by strictly larger values (io.c:124-133).  max is order-independent, so the value equals the contexts'.
__global__ void __launch_bounds__(256) max_mat_noalloc_kernel(const double* __restrict__ m, long long row_stride,
                                                               int i_max, int j_max, double* __restrict__ scratch,
                                                               double* __restrict__ out) {
    __shared__ double sh[32];
    __shared__ int is_last;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
    double* partials = scratch + NSB_NOALLOC_HEAD;
    const double* row = m + (long long)(blockIdx.x + 1) * row_stride;
    double mx = -1.0e300;
    for (int j = 1 + threadIdx.x; j <= j_max; j += blockDim.x) mx = fmax(mx, row[j]);
    const double bm = block_max(mx, sh);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = bm;
        __threadfence();
        const unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    double a = -1.0e300;
    for (int i = threadIdx.x; i < i_max; i += blockDim.x) a = fmax(a, __ldcg(partials + i));
    a = block_max(a, sh);
    if (threadIdx.x == 0) {
        const double m00 = m[0];
        *out = (m00 < a) ? a : m00;
        *ticket = 0u;
    }
}

// Set-up, once per device: the device address of the scratch array.
static int noalloc_scratch(double** ptr) {
    static std::mutex mu;
    static double* table[NSB_NOALLOC_MAX_DEVICES];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return nsb_fail(NULL, NSB_ERR_CUDA, "nsb_max_mat_noalloc: cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= NSB_NOALLOC_MAX_DEVICES)
        return nsb_fail(NULL, NSB_ERR_ARG, "nsb_max_mat_noalloc: device %d is beyond the %d this entry point keeps addresses for", dev, NSB_NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mu);
    if (!table[dev]) {
        double* p = NULL;
        e = cudaGetSymbolAddress((void**)&p, nsb_noalloc_scratch);      // the current device's copy of the array
        if (e != cudaSuccess || !p)
            return nsb_fail(NULL, NSB_ERR_CUDA, "nsb_max_mat_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        table[dev] = p;
    }
    *ptr = table[dev];
    return NSB_OK;
}

extern "C" int nsb_max_mat_noalloc(const double* d_m, long long row_stride, int i_max, int j_max, double* d_max,
                                   void* stream) {
    if (!d_m || !d_max) return nsb_fail(NULL, NSB_ERR_ARG, "nsb_max_mat_noalloc: NULL grid or result pointer");
    if (i_max < 1 || j_max < 1) return nsb_fail(NULL, NSB_ERR_ARG, "nsb_max_mat_noalloc: i_max = %d, j_max = %d must be positive", i_max, j_max);
    if (i_max > NSB_NOALLOC_MAX_ROWS || j_max > NSB_NOALLOC_MAX_COLS)
        return nsb_fail(NULL, NSB_ERR_ARG, "nsb_max_mat_noalloc: grid %d x %d exceeds the static scratch limit of %d x %d cells",
                        i_max, j_max, NSB_NOALLOC_MAX_ROWS, NSB_NOALLOC_MAX_COLS);
    if (row_stride < (long long)j_max + 1)
        return nsb_fail(NULL, NSB_ERR_ARG, "nsb_max_mat_noalloc: row_stride %lld is shorter than j_max + 1 = %d", row_stride, j_max + 1);
    double* scratch = NULL;
    const int r = noalloc_scratch(&scratch);
    if (r != NSB_OK) return r;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(scratch, 0, NSB_NOALLOC_HEAD * sizeof(double), s);   // the ticket starts at zero
    if (e != cudaSuccess) return nsb_fail(NULL, NSB_ERR_CUDA, "nsb_max_mat_noalloc: cudaMemsetAsync failed: %s", cudaGetErrorString(e));
    max_mat_noalloc_kernel<<<i_max, 256, 0, s>>>(d_m, row_stride, i_max, j_max, scratch, d_max);
    e = cudaGetLastError();
    if (e != cudaSuccess) return nsb_fail(NULL, NSB_ERR_CUDA, "nsb_max_mat_noalloc: launch failed: %s", cudaGetErrorString(e));
    return NSB_OK;
}
