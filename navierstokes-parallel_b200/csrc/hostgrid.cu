// hostgrid.cu -- C ABI group B: the reference's own function signatures on caller-owned jagged
// `double**` grids (src/serial/memory.c:3-26).  Each call stages the grids through pinned memory,
// uploads, runs the CUDA path and downloads the results.  This is the end-to-end path bench.py
// times (host buffers in, host buffers out).
//
// A context is cached per (i_max, j_max, a, b) so repeated calls from a time loop do not reallocate;
// nsb_host_api_reset() drops it.  This ONE cached context is the only process-global mutable state of the
// library (group A, the context API, has none); group B is therefore not thread-safe, like the reference
// (single-threaded, global BLOCK_SIZE main.cu:26).
// Transfers are stream-ordered: a grid is copied into the scratch grid and transposed on the context's stream,
// the next transfer queues behind it; the host only waits where it must touch a staging buffer itself and once
// at the end of a call's downloads.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nsb_internal.h"

static nsb_ctx* g_ctx = NULL;

extern "C" void nsb_host_api_reset(void) {
    if (g_ctx) nsb_destroy(g_ctx);
    g_ctx = NULL;
}

static nsb_ctx* cached_ctx(int i_max, int j_max, double a, double b, const nsb_params* full) {
    nsb_params P;
    if (full) P = *full;
    else {
        memset(&P, 0, sizeof P);
        P.problem = 1; P.i_max = i_max; P.j_max = j_max; P.a = a; P.b = b;
        P.Re = 1.0; P.tau = 1.0; P.omega = 1.0; P.epsilon = 0.0; P.max_it = 0; P.n_print = 1;
    }
    if (g_ctx && g_ctx->P.i_max == P.i_max && g_ctx->P.j_max == P.j_max && g_ctx->P.a == P.a &&
        g_ctx->P.b == P.b) {
        g_ctx->P = P;
        return g_ctx;
    }
    nsb_host_api_reset();
    if (nsb_create(&P, NULL, &g_ctx) != NSB_OK) {
        fprintf(stderr, "nsb200: %s\n", nsb_last_error(NULL));
        g_ctx = NULL;
    }
    return g_ctx;
}

// delta_x, delta_y as the caller passes them (the reference hands them to SOR / FG verbatim), and what follows from them
static void set_spacing(nsb_ctx* c, double dx, double dy) {
    c->dx = dx; c->dy = dy;
    int e1, e2;
    const double dx2 = dx * dx, dy2 = dy * dy;
    c->recip_exact = frexp(dx2, &e1) == 0.5 && frexp(dy2, &e2) == 0.5;
    c->use_recip = (c->O.div_mode == NSB_DIV_RECIP) || (c->O.div_mode == NSB_DIV_AUTO && c->recip_exact);
    c->rcp_valid = false;
}

static int ensure_stage(nsb_ctx* c) {
    const size_t n = ((size_t)c->g.ni + 2) * ((size_t)c->g.nj + 2);
    if (c->h_stage_elems >= n) return 0;
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = NULL; c->h_stage_elems = 0;
    if (cudaHostAlloc((void**)&c->h_stage, n * sizeof(double), cudaHostAllocDefault) != cudaSuccess) return -1;
    c->h_stage_elems = n;
    return 0;
}

static double* dev_field(nsb_ctx* c, int field) {
    void* p = NULL;
    nsb_device_view(c, field, &p, NULL, NULL, NULL);
    return (double*)p;
}

// rows laid out back to back (one allocation, e.g. a numpy array or a pinned buffer)?
static bool contiguous(double** m, int nrows, int ncols) {
    for (int i = 0; i + 1 < nrows; i++)
        if (m[i + 1] != m[i] + ncols) return false;
    return true;
}

#define HG(call) do { if ((call) != cudaSuccess) { snprintf(c->err, sizeof c->err, "%s failed: %s", #call, cudaGetErrorString(cudaGetLastError())); return NSB_ERR_CUDA; } } while (0)

// jagged rows -> dense reference-order grid in HBM (scratch) -> transposed into the field.
// nrows/ncols: what the caller allocated (u: i_max+1 rows; v: j_max+1 columns; memory.c:3-26).
// Contiguous row blocks are DMA'd straight from the caller's memory (at PCIe speed when it is
// pinned); truly jagged grids are gathered row by row into a pinned staging buffer first.
static int put(nsb_ctx* c, int field, double** m, int nrows, int ncols) {
    const size_t ld = (size_t)c->g.nj + 2;
    const size_t total = ((size_t)c->g.ni + 2) * ld;
    double* dst = dev_field(c, field);
    if (!dst) return NSB_ERR_ARG;
    HG(cudaSetDevice(c->device));
    if (contiguous(m, nrows, ncols)) {
        if ((size_t)nrows * (size_t)ncols != total) HG(cudaMemsetAsync(c->scratch, 0, total * sizeof(double), c->stream));
        HG(cudaMemcpy2DAsync(c->scratch, ld * sizeof(double), m[0], (size_t)ncols * sizeof(double),
                             (size_t)ncols * sizeof(double), (size_t)nrows, cudaMemcpyHostToDevice, c->stream));
    } else {
        if (ensure_stage(c)) return NSB_ERR_CUDA;
        HG(cudaStreamSynchronize(c->stream));                  // an earlier copy may still be reading the staging buffer
        if ((size_t)nrows * (size_t)ncols != total) memset(c->h_stage, 0, total * sizeof(double));
        for (int i = 0; i < nrows; i++) memcpy(c->h_stage + (size_t)i * ld, m[i], (size_t)ncols * sizeof(double));
        HG(cudaMemcpyAsync(c->scratch, c->h_stage, total * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    }
    HG(nsb_launch_to_xfast(c, c->scratch, dst));              // stream-ordered: the next transfer may reuse the scratch grid
    c->red_valid = false;
    return NSB_OK;
}

static int get(nsb_ctx* c, int field, double** m, int nrows, int ncols) {
    const size_t ld = (size_t)c->g.nj + 2;
    const size_t total = ((size_t)c->g.ni + 2) * ld;
    const double* src = dev_field(c, field);
    if (!src) return NSB_ERR_ARG;
    HG(cudaSetDevice(c->device));
    HG(nsb_launch_from_xfast(c, src, c->scratch));
    if (contiguous(m, nrows, ncols)) {
        HG(cudaMemcpy2DAsync(m[0], (size_t)ncols * sizeof(double), c->scratch, ld * sizeof(double),
                             (size_t)ncols * sizeof(double), (size_t)nrows, cudaMemcpyDeviceToHost, c->stream));
        // (no wait here: the callers synchronise once after their last download, get_done)
    } else {
        if (ensure_stage(c)) return NSB_ERR_CUDA;
        HG(cudaMemcpyAsync(c->h_stage, c->scratch, total * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        HG(cudaStreamSynchronize(c->stream));
        for (int i = 0; i < nrows; i++) memcpy(m[i], c->h_stage + (size_t)i * ld, (size_t)ncols * sizeof(double));
    }
    return NSB_OK;
}

static int get_done(nsb_ctx* c) {
    HG(cudaSetDevice(c->device));
    HG(cudaStreamSynchronize(c->stream));
    return NSB_OK;
}

// integration.h:165 with the CUDA reference's iteration report (main.cu:717,725)
extern "C" int nsb_SOR_iters(double** p, int i_max, int j_max, double delta_x, double delta_y, double** res,
                             double** RHS, double omega, double eps, int max_it, int* iters, double* last_res) {
    (void)res;
    nsb_ctx* c = cached_ctx(i_max, j_max, delta_x * i_max, delta_y * j_max, NULL);
    if (!c) return NSB_ERR_CUDA;
    // the reference passes delta_x, delta_y themselves: use them verbatim, not a / i_max
    set_spacing(c, delta_x, delta_y);
    c->P.omega = omega; c->P.epsilon = eps; c->P.max_it = max_it;
    int r;
    if ((r = put(c, NSB_P, p, i_max + 2, j_max + 2))) return r;
    if ((r = put(c, NSB_RHS, RHS, i_max + 2, j_max + 2))) return r;
    int it = 0;
    r = nsb_sor(c, &it, last_res, NULL);
    if (r < 0 && r != NSB_NOT_CONVERGED) return r;
    if (iters) *iters = it;
    int g = get(c, NSB_P, p, i_max + 2, j_max + 2);
    if (!g) g = get_done(c);
    if (g) return g;
    return r;
}

extern "C" int nsb_SOR(double** p, int i_max, int j_max, double delta_x, double delta_y, double** res,
                       double** RHS, double omega, double eps, int max_it) {
    int r = nsb_SOR_iters(p, i_max, j_max, delta_x, delta_y, res, RHS, omega, eps, max_it, NULL, NULL);
    if (r > 0) return 0;                                       // integration.c:165-167
    return r;                                                  // -1 (integration.c:172) or an error
}

// integration.h:136
extern "C" int nsb_FG(double** F, double** G, double** u, double** v, int i_max, int j_max, double Re, double g_x,
                      double g_y, double delta_t, double delta_x, double delta_y, double gamma) {
    nsb_ctx* c = cached_ctx(i_max, j_max, delta_x * i_max, delta_y * j_max, NULL);
    if (!c) return NSB_ERR_CUDA;
    set_spacing(c, delta_x, delta_y);
    c->P.Re = Re; c->P.g_x = g_x; c->P.g_y = g_y;
    int r;
    if ((r = put(c, NSB_U, u, i_max + 1, j_max + 2))) return r;     // u: rows 0..i_max
    if ((r = put(c, NSB_V, v, i_max + 2, j_max + 1))) return r;     // v: columns 0..j_max
    if ((r = put(c, NSB_F, F, i_max + 2, j_max + 2))) return r;     // untouched cells keep caller values
    if ((r = put(c, NSB_G, G, i_max + 2, j_max + 2))) return r;
    if ((r = nsb_fg(c, delta_t, gamma))) return r;
    if ((r = get(c, NSB_F, F, i_max + 2, j_max + 2))) return r;
    if ((r = get(c, NSB_G, G, i_max + 2, j_max + 2))) return r;
    return get_done(c);
}

// io.h:42
extern "C" double nsb_max_mat(int i_max, int j_max, double** m) {
    // The reduction kernel yields max u and max v together; here one grid is asked for, so it is
    // uploaded into the u field and only that maximum is returned.  The reference calls max_mat on u
    // (i_max+1 rows of j_max+2) AND on v (i_max+2 rows of j_max+1, memory.c:15) with the same argument list, and
    // scans m[0][0] and i = 1..i_max, j = 1..j_max (io.c:122-139): rows 0..i_max, columns 0..j_max are what both
    // shapes own, so exactly those are read -- nothing past the end of a v row.
    nsb_ctx* c = cached_ctx(i_max, j_max, 1.0, 1.0, NULL);
    if (!c) return NAN;
    if (put(c, NSB_U, m, i_max + 1, j_max + 1)) return NAN;
    double um = 0.0, vm = 0.0;
    if (nsb_max_uv(c, &um, &vm)) return NAN;
    return um;
}

// main.c:88-145 on host grids
extern "C" int nsb_timestep(const nsb_params* P, double t, double** u, double** v, double** p, double** F,
                            double** G, double** RHS, nsb_step_info* info) {
    if (!P) return NSB_ERR_ARG;
    nsb_ctx* c = cached_ctx(P->i_max, P->j_max, P->a, P->b, P);
    if (!c) return NSB_ERR_CUDA;
    set_spacing(c, P->a / P->i_max, P->b / P->j_max);
    const int ni = P->i_max, nj = P->j_max;
    int r;
    if ((r = put(c, NSB_U, u, ni + 1, nj + 2))) return r;
    if ((r = put(c, NSB_V, v, ni + 2, nj + 1))) return r;
    if ((r = put(c, NSB_P, p, ni + 2, nj + 2))) return r;
    if (F && (r = put(c, NSB_F, F, ni + 2, nj + 2))) return r;
    if (G && (r = put(c, NSB_G, G, ni + 2, nj + 2))) return r;
    if ((r = nsb_step(c, t, info))) return r;
    if ((r = get(c, NSB_U, u, ni + 1, nj + 2))) return r;
    if ((r = get(c, NSB_V, v, ni + 2, nj + 1))) return r;
    if ((r = get(c, NSB_P, p, ni + 2, nj + 2))) return r;
    if (F && (r = get(c, NSB_F, F, ni + 2, nj + 2))) return r;
    if (G && (r = get(c, NSB_G, G, ni + 2, nj + 2))) return r;
    if (RHS && (r = get(c, NSB_RHS, RHS, ni + 2, nj + 2))) return r;
    return get_done(c);
}
