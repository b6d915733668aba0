// bw_probe.cu -- what HBM bandwidth can plain streaming kernels reach on this GPU for the SOR
// kernel's traffic mix (2 reads : 1 write) and for a copy (1 : 1)?  Developer tool, not product.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/bw_probe tools/bw_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_copy(const double2* __restrict__ a, double2* __restrict__ c, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) c[i] = a[i];
}
__global__ void k_add(const double2* __restrict__ a, const double2* __restrict__ b, double2* __restrict__ c, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        double2 x = a[i], y = b[i];
        c[i] = make_double2(x.x + y.x, x.y + y.y);
    }
}
// 4 independent 16-byte loads per array per thread in flight
__global__ void k_add4(const double2* __restrict__ a, const double2* __restrict__ b, double2* __restrict__ c, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i + 3 * stride < n; i += 4 * stride) {
        double2 x0 = a[i], x1 = a[i + stride], x2 = a[i + 2 * stride], x3 = a[i + 3 * stride];
        double2 y0 = b[i], y1 = b[i + stride], y2 = b[i + 2 * stride], y3 = b[i + 3 * stride];
        c[i] = make_double2(x0.x + y0.x, x0.y + y0.y);
        c[i + stride] = make_double2(x1.x + y1.x, x1.y + y1.y);
        c[i + 2 * stride] = make_double2(x2.x + y2.x, x2.y + y2.y);
        c[i + 3 * stride] = make_double2(x3.x + y3.x, x3.y + y3.y);
    }
}
__global__ void k_read2(const double2* __restrict__ a, const double2* __restrict__ b, double* out, size_t n) {
    double s = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        double2 x = a[i], y = b[i];
        s += x.x + y.x + x.y + y.y;
    }
    if (s == 1.2345) out[0] = s;
}

template <typename F>
static float time_it(F f, int reps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; i++) f();
    cudaEventRecord(e0);
    for (int i = 0; i < reps; i++) f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms / reps;
}

int main() {
    const size_t n = (size_t)8200 * 8200 / 2 * 2;     // doubles per array (~538 MB)
    const size_t n2 = n / 2;
    double2 *a, *b, *c; double* out;
    cudaMalloc(&a, n * 8); cudaMalloc(&b, n * 8); cudaMalloc(&c, n * 8); cudaMalloc(&out, 8);
    cudaMemset(a, 0, n * 8); cudaMemset(b, 0, n * 8); cudaMemset(c, 0, n * 8);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int cfgs[][2] = {{256, 8}, {512, 4}, {1024, 2}, {256, 16}, {128, 16}, {256, 4}, {256, 2}};
    for (auto& cf : cfgs) {
        const int th = cf[0], bl = sms * cf[1];
        float t1 = time_it([&] { k_copy<<<bl, th>>>(a, c, n2); }, 20);
        float t2 = time_it([&] { k_add<<<bl, th>>>(a, b, c, n2); }, 20);
        float t3 = time_it([&] { k_add4<<<bl, th>>>(a, b, c, n2); }, 20);
        float t4 = time_it([&] { k_read2<<<bl, th>>>(a, b, out, n2); }, 20);
        printf("threads %4d blocks/SM %2d : copy %7.1f GB/s | add(2R1W) %7.1f GB/s | add4 %7.1f GB/s | read2 %7.1f GB/s\n", th, cf[1],
               2.0 * n * 8 / t1 / 1e6, 3.0 * n * 8 / t2 / 1e6, 3.0 * n * 8 / t3 / 1e6, 2.0 * n * 8 / t4 / 1e6);
    }
    // one block per (many) elements, no grid-stride loop: the hardware block scheduler balances
    for (int th : {256, 512}) {
        const int bl = (int)((n2 + th - 1) / th);
        float t2 = time_it([&] { k_add<<<bl, th>>>(a, b, c, n2); }, 20);
        float t1 = time_it([&] { k_copy<<<bl, th>>>(a, c, n2); }, 20);
        printf("threads %4d one elem/thread   : copy %7.1f GB/s | add(2R1W) %7.1f GB/s\n", th, 2.0 * n * 8 / t1 / 1e6, 3.0 * n * 8 / t2 / 1e6);
    }
    float tm = time_it([&] { cudaMemcpyAsync(c, a, n * 8, cudaMemcpyDeviceToDevice); }, 20);
    printf("cudaMemcpy D2D: %7.1f GB/s\n", 2.0 * n * 8 / tm / 1e6);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
