// ref_cuda_harness.cu -- TEST INFRASTRUCTURE ONLY (never linked into libnsb200.so).
//
// Field-level access to the reference's OWN CUDA red-black solver.  The reference translation unit
// /root/reference/src/parallel/main.cu is #included below UNMODIFIED, where it lies (nothing is copied into
// this repository; `main` is renamed so the file can live inside a shared library), and the entry points
// call its own functions:
//     allocate_device_memory  (main.cu:46-154)     precalculate_borders      (main.cu:194-215)
//     SOR_UVA_with_shared_memory (main.cu:656-726) max_mat_cuda              (main.cu:799-832)
//     set_noslip_cuda / set_inflow_cuda (928-944)  calculate_{F,G,RHS}_kernel, update_velocities_kernel
// refcuda_run repeats the loop body of the reference's main (main.cu:1046-1103) -- that loop is not callable
// as a function -- with the same calls in the same order, and additionally records what the reference throws
// away: the return value of every SOR call (it+1 / -1) and the dt sequence.
//
// Built by `make -C oracle refcuda` into oracle/_ref/librefcuda.so (git-ignored, travels to the GPU box).
// The reference kernel is only valid for i_max, j_max multiples of the block size: tests use those.
// Grid layout on this boundary: flat row-major [i][j], (i_max+2) x (j_max+2) doubles -- the layout of the
// reference's device arrays (main.cu:103-120) and of oracle/ns_oracle.h.
#define main ref_main
#include "main.cu"
#undef main

static double* refcuda_data(double** table) {           // row 0 of the device row-pointer table = the contiguous block
    double* d = NULL;
    CHECK_CUDA_ERROR(cudaMemcpy(&d, table, sizeof(double*), cudaMemcpyDeviceToHost));
    return d;
}

struct RefGrids {
    double **u, **v, **p, **res, **RHS, **F, **G;
    BoundaryPoint* borders;
    int n_border;
    size_t bytes;
};

static void refcuda_alloc(RefGrids& g, int i_max, int j_max) {
    g.n_border = 2 * (i_max + j_max + 2);                // main.cu:1030
    allocate_device_memory(&g.u, &g.v, &g.p, &g.res, &g.RHS, &g.F, &g.G, i_max, j_max, &g.borders, g.n_border);
    precalculate_borders(i_max, j_max, g.borders);
    g.bytes = (size_t)(i_max + 2) * (size_t)(j_max + 2) * sizeof(double);
}

extern "C" int refcuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

// One call of the reference's SOR on caller data.  p is updated in place; returns what the reference
// returns (it+1 when converged, -1 otherwise).
extern "C" int refcuda_sor(double* p, const double* rhs, int i_max, int j_max, double dx, double dy,
                           double omega, double eps, int max_it, int block_size) {
    RefGrids g;
    BLOCK_SIZE = block_size;
    refcuda_alloc(g, i_max, j_max);
    CHECK_CUDA_ERROR(cudaMemcpy(refcuda_data(g.p), p, g.bytes, cudaMemcpyHostToDevice));
    CHECK_CUDA_ERROR(cudaMemcpy(refcuda_data(g.RHS), rhs, g.bytes, cudaMemcpyHostToDevice));
    const int ret = SOR_UVA_with_shared_memory(g.p, i_max, j_max, dx, dy, g.res, g.RHS, omega, eps, max_it,
                                               g.borders, g.n_border, block_size);
    CHECK_CUDA_ERROR(cudaMemcpy(p, refcuda_data(g.p), g.bytes, cudaMemcpyDeviceToHost));
    free_device_memory(g.u, g.v, g.p, g.res, g.RHS, g.F, g.G, g.borders);
    return ret;
}

// Whole run, main.cu:1046-1103, from zero fields.  Outputs: final u, v, p (flat [i][j]), per step the SOR
// return value and dt (up to max_log), number of steps taken.  max_steps <= 0: until t >= T.
extern "C" int refcuda_run(int problem, double f, int i_max, int j_max, double a, double b, double T, double Re,
                           double g_x, double g_y, double tau, double omega, double epsilon, int max_it,
                           int block_size, int max_steps, double* u_out, double* v_out, double* p_out,
                           int* sor_ret, double* dt_log, int max_log) {
    RefGrids g;
    BLOCK_SIZE = block_size;
    refcuda_alloc(g, i_max, j_max);
    const double delta_x = a / i_max, delta_y = b / j_max;        // main.cu:1021-1022
    double t = 0;
    int n = 0;
    while (t < T && (max_steps <= 0 || n < max_steps)) {           // main.cu:1046
        double u_max = max_mat_cuda(i_max, j_max, g.u);            // :1048-1051
        double v_max = max_mat_cuda(i_max, j_max, g.v);
        double delta_t = tau * n_min(3, Re / 2.0 / (1.0 / delta_x / delta_x + 1.0 / delta_y / delta_y), delta_x / fabs(u_max), delta_y / fabs(v_max));
        double gamma = fmax(u_max * delta_t / delta_x, v_max * delta_t / delta_y);
        if (problem == 1 || problem == 2) {                        // :1054-1064
            set_noslip_cuda(i_max, j_max, g.u, g.v, LEFT, g.borders, g.n_border);
            set_noslip_cuda(i_max, j_max, g.u, g.v, RIGHT, g.borders, g.n_border);
            set_noslip_cuda(i_max, j_max, g.u, g.v, BOTTOM, g.borders, g.n_border);
            set_inflow_cuda(i_max, j_max, g.u, g.v, TOP, problem == 1 ? 1.0 : sin(f * t), 0.0, g.borders, g.n_border);
        }
        dim3 blockDim(16, 16);                                     // :1066-1068
        dim3 gridDim((i_max + blockDim.x - 1) / blockDim.x, (j_max + blockDim.y - 1) / blockDim.y);
        calculate_F_kernel<<<gridDim, blockDim>>>(g.F, g.u, g.v, i_max, j_max, Re, g_x, delta_t, delta_x, delta_y, gamma);
        calculate_G_kernel<<<gridDim, blockDim>>>(g.G, g.u, g.v, i_max, j_max, Re, g_y, delta_t, delta_x, delta_y, gamma);
        CHECK_CUDA_ERROR(cudaDeviceSynchronize());
        calculate_RHS_kernel<<<gridDim, blockDim>>>(g.RHS, g.F, g.G, i_max, j_max, delta_t, delta_x, delta_y);
        CHECK_CUDA_ERROR(cudaDeviceSynchronize());
        const int ret = SOR_UVA_with_shared_memory(g.p, i_max, j_max, delta_x, delta_y, g.res, g.RHS, omega, epsilon, max_it,
                                                   g.borders, g.n_border, BLOCK_SIZE);      // :1087
        update_velocities_kernel<<<gridDim, blockDim>>>(g.u, g.v, g.F, g.G, g.p, i_max, j_max, delta_t, delta_x, delta_y);
        CHECK_CUDA_ERROR(cudaDeviceSynchronize());
        if (n < max_log) {
            if (sor_ret) sor_ret[n] = ret;
            if (dt_log) dt_log[n] = delta_t;
        }
        t += delta_t;                                              // :1101-1102
        n++;
    }
    if (u_out) CHECK_CUDA_ERROR(cudaMemcpy(u_out, refcuda_data(g.u), g.bytes, cudaMemcpyDeviceToHost));
    if (v_out) CHECK_CUDA_ERROR(cudaMemcpy(v_out, refcuda_data(g.v), g.bytes, cudaMemcpyDeviceToHost));
    if (p_out) CHECK_CUDA_ERROR(cudaMemcpy(p_out, refcuda_data(g.p), g.bytes, cudaMemcpyDeviceToHost));
    free_device_memory(g.u, g.v, g.p, g.res, g.RHS, g.F, g.G, g.borders);
    return n;
}
