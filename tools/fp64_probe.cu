// Developer probe: FP64 FMA issue rate of one GPU (lane-operations per clock and SM), to tell whether the
// IEEE-division sweep is bound by the FP64 pipe.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/fp64_probe tools/fp64_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void __launch_bounds__(256) dfma_kernel(double* out, double a, double b, int iters) {
    double v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; k++) v[k] = threadIdx.x * 1e-3 + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < ILP; k++) v[k] = __fma_rn(v[k], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < ILP; k++) s += v[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double* out; cudaMalloc(&out, sizeof(double) * 256 * p.multiProcessorCount * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int bps = 1; bps <= 8; bps *= 2) {
        const int iters = 20000, blocks = p.multiProcessorCount * bps;
        dfma_kernel<8><<<blocks, 256>>>(out, 0.999, 1e-3, 1000);
        cudaEventRecord(e0);
        dfma_kernel<8><<<blocks, 256>>>(out, 0.999, 1e-3, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double ops = (double)blocks * 256 * 8 * iters;
        printf("blocks/SM %d warps/SM %2d: %.2f TFMA/s = %.1f DFMA lanes per SM per ns (nominal clock %.2f GHz -> %.1f per clock)\n", bps, bps * 8,
               ops / ms / 1e9, ops / ms / 1e6 / p.multiProcessorCount, clk_khz / 1e6, ops / ms / 1e6 / p.multiProcessorCount / (clk_khz / 1e6));
    }
    return 0;
}
