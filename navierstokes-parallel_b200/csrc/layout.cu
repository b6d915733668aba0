// layout.cu -- conversion between the reference's grid order and the HBM layout.
//
// The reference stores element (i,j) at row i, column j with j contiguous (memory.c:3-26,
// main.cu:104).  The kernels here want i contiguous.  A host grid in reference order is copied to
// a dense staging grid in HBM and transposed there through a padded 32x32 shared-memory tile
// (both the read and the write are coalesced); the opposite direction likewise.
#include "nsb_internal.h"

#define TT 32

// dense reference-order [(ni+2)][(nj+2)]  ->  pitched x-fastest grid
__global__ void ref_to_grid_kernel(const double* __restrict__ src, double* __restrict__ dst, long long pitch,
                                   int nx, int ny) {   // nx = ni+2 (i extent), ny = nj+2 (j extent)
    __shared__ double tile[TT][TT + 1];
    const int j0 = blockIdx.x * TT, i0 = blockIdx.y * TT;
    // read: j fastest in src
    for (int r = threadIdx.y; r < TT; r += blockDim.y) {
        int i = i0 + r, j = j0 + threadIdx.x;
        if (i < nx && j < ny) tile[r][threadIdx.x] = src[(long long)i * ny + j];
    }
    __syncthreads();
    // write: i fastest in dst
    for (int r = threadIdx.y; r < TT; r += blockDim.y) {
        int j = j0 + r, i = i0 + threadIdx.x;
        if (i < nx && j < ny) dst[(long long)(j + NSB_ROW0) * pitch + NSB_COL0 + i] = tile[threadIdx.x][r];
    }
}

__global__ void grid_to_ref_kernel(const double* __restrict__ src, double* __restrict__ dst, long long pitch,
                                   int nx, int ny) {
    __shared__ double tile[TT][TT + 1];
    const int i0 = blockIdx.x * TT, j0 = blockIdx.y * TT;
    for (int r = threadIdx.y; r < TT; r += blockDim.y) {
        int j = j0 + r, i = i0 + threadIdx.x;
        if (i < nx && j < ny) tile[r][threadIdx.x] = src[(long long)(j + NSB_ROW0) * pitch + NSB_COL0 + i];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < TT; r += blockDim.y) {
        int i = i0 + r, j = j0 + threadIdx.x;
        if (i < nx && j < ny) dst[(long long)i * ny + j] = tile[threadIdx.x][r];
    }
}

cudaError_t nsb_launch_to_xfast(nsb_ctx* c, const double* src_ref, double* dst) {
    const int nx = c->g.ni + 2, ny = c->g.nj + 2;
    dim3 b(TT, 8), g((ny + TT - 1) / TT, (nx + TT - 1) / TT);
    ref_to_grid_kernel<<<g, b, 0, c->stream>>>(src_ref, dst, c->g.pitch, nx, ny);
    return NSB_LAUNCHED(c, 1);
}

cudaError_t nsb_launch_from_xfast(nsb_ctx* c, const double* src, double* dst_ref) {
    const int nx = c->g.ni + 2, ny = c->g.nj + 2;
    dim3 b(TT, 8), g((nx + TT - 1) / TT, (ny + TT - 1) / TT);
    grid_to_ref_kernel<<<g, b, 0, c->stream>>>(src, dst_ref, c->g.pitch, nx, ny);
    return NSB_LAUNCHED(c, 1);
}
