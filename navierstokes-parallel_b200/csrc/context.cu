// context.cu -- the C ABI of include/nsb200.h, group A (context API): allocation, the
// time-step phases in the reference's order, the SOR drive loop, transfers.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nsb_internal.h"

static char g_create_err[512] = "";

static int fail(nsb_ctx* c, int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c ? c->err : g_create_err, 512, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(c, call)                                                                                \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail((c), NSB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                        __FILE__, __LINE__);                                                       \
    } while (0)

static bool is_pow2(double x) {
    if (!(x > 0.0) || isinf(x)) return false;
    int e;
    return frexp(x, &e) == 0.5;
}

extern "C" int nsb_abi_version(void) { return NSB_ABI_VERSION; }

extern "C" const char* nsb_last_error(const nsb_ctx* ctx) { return ctx ? ctx->err : g_create_err; }

extern "C" void nsb_default_options(nsb_options* o) {
    memset(o, 0, sizeof *o);
    const char* d = getenv("NSB_DEVICE");
    o->device = d ? atoi(d) : 0;
    o->div_mode = NSB_DIV_AUTO;
    o->sor_engine = NSB_SOR_AUTO;
}

// io.c:12-59: fgets(256) + sscanf of the first token, 15 lines, fixed order.
extern "C" int nsb_read_params(const char* path, nsb_params* o) {
    FILE* fp = fopen(path, "r");
    if (!fp) return fail(NULL, NSB_ERR_IO, "cannot open parameter file %s", path);
    memset(o, 0, sizeof *o);
    char line[256];
    for (int k = 0; k < 15; k++) {
        if (!fgets(line, sizeof line, fp)) break;
        switch (k) {
            case 0: sscanf(line, "%d", &o->problem); break;
            case 1: sscanf(line, "%lf", &o->f); break;
            case 2: sscanf(line, "%d", &o->i_max); break;
            case 3: sscanf(line, "%d", &o->j_max); break;
            case 4: sscanf(line, "%lf", &o->a); break;
            case 5: sscanf(line, "%lf", &o->b); break;
            case 6: sscanf(line, "%lf", &o->T); break;
            case 7: sscanf(line, "%lf", &o->Re); break;
            case 8: sscanf(line, "%lf", &o->g_x); break;
            case 9: sscanf(line, "%lf", &o->g_y); break;
            case 10: sscanf(line, "%lf", &o->tau); break;
            case 11: sscanf(line, "%lf", &o->omega); break;
            case 12: sscanf(line, "%lf", &o->epsilon); break;
            case 13: sscanf(line, "%d", &o->max_it); break;
            case 14: sscanf(line, "%d", &o->n_print); break;
        }
    }
    fclose(fp);
    return NSB_OK;
}

extern "C" void nsb_destroy(nsb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    double* bufs[] = {c->u, c->v, c->F, c->G, c->rhs, c->pbuf[0], c->pbuf[1], c->scratch,
                      c->d_partials, c->d_red_partials};
    for (double* b : bufs) if (b) cudaFree(b);
    if (c->d_ctl) cudaFree(c->d_ctl);
    if (c->d_red) cudaFree(c->d_red);
    if (c->h_ctl) cudaFreeHost(c->h_ctl);
    if (c->h_red) cudaFreeHost(c->h_red);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    for (int k = 0; k < 2; k++) if (c->ev_chk[k]) cudaEventDestroy(c->ev_chk[k]);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    free(c);
}

extern "C" int nsb_create(const nsb_params* P, const nsb_options* opt, nsb_ctx** out) {
    if (!P || !out) return fail(NULL, NSB_ERR_ARG, "nsb_create: null argument");
    *out = NULL;
    if (P->i_max < 1 || P->j_max < 1) return fail(NULL, NSB_ERR_ARG, "i_max and j_max must be >= 1");
    nsb_ctx* c = (nsb_ctx*)calloc(1, sizeof(nsb_ctx));
    if (!c) return fail(NULL, NSB_ERR_ARG, "out of host memory");
    c->P = *P;
    if (opt) c->O = *opt; else nsb_default_options(&c->O);
    c->device = c->O.device;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev < 1) {
        fail(NULL, NSB_ERR_CUDA, "no usable CUDA device (%s): this library has no CPU path",
             e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        free(c);
        return NSB_ERR_CUDA;
    }
#define CUC(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            fail(NULL, NSB_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__));              \
            nsb_destroy(c);                                                                         \
            return NSB_ERR_CUDA;                                                                    \
        }                                                                                           \
    } while (0)
    CUC(cudaSetDevice(c->device));
    cudaDeviceProp prop;
    CUC(cudaGetDeviceProperties(&prop, c->device));
    c->sm_count = prop.multiProcessorCount;
    if (c->O.stream) { c->stream = (cudaStream_t)c->O.stream; c->own_stream = false; }
    else { CUC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }

    c->g.ni = P->i_max; c->g.nj = P->j_max;
    // pitch: COL0 + (ni+2) columns + one spare group, rounded up to 16 doubles (128 B)
    long long need = NSB_COL0 + (long long)P->i_max + 2 + NSB_GROUP;
    c->g.pitch = (need + 15) / 16 * 16;
    c->g.rows = P->j_max + 2 + NSB_ROW0 + NSB_ROWS_ABOVE;
    c->g.elems = (size_t)c->g.pitch * (size_t)c->g.rows;
    c->dx = P->a / P->i_max;                                   // main.c:72-73
    c->dy = P->b / P->j_max;
    c->recip_exact = is_pow2(c->dx * c->dx) && is_pow2(c->dy * c->dy);
    c->use_recip = (c->O.div_mode == NSB_DIV_RECIP) || (c->O.div_mode == NSB_DIV_AUTO && c->recip_exact);

    const size_t bytes = c->g.elems * sizeof(double);
    double** grids[] = {&c->u, &c->v, &c->F, &c->G, &c->rhs, &c->pbuf[0], &c->pbuf[1], &c->scratch};
    for (double** gp : grids) {
        CUC(cudaMalloc((void**)gp, bytes));
        CUC(cudaMemsetAsync(*gp, 0, bytes, c->stream));        // calloc semantics, memory.c:3-26
    }
    c->pcur = 0;
    nsb_sor_plan(c);
    int np = c->n_partials;
    if (np < 8 * P->j_max + 64) np = 8 * P->j_max + 64;
    CUC(cudaMalloc((void**)&c->d_partials, (size_t)np * sizeof(double)));
    c->n_red_partials = 3 * 4 * P->j_max + 64;
    CUC(cudaMalloc((void**)&c->d_red_partials, (size_t)c->n_red_partials * sizeof(double)));
    CUC(cudaMalloc((void**)&c->d_ctl, sizeof(SorCtl)));
    CUC(cudaMalloc((void**)&c->d_red, sizeof(StepRed)));
    CUC(cudaMemsetAsync(c->d_ctl, 0, sizeof(SorCtl), c->stream));
    CUC(cudaMemsetAsync(c->d_red, 0, sizeof(StepRed), c->stream));
    CUC(cudaHostAlloc((void**)&c->h_ctl, sizeof(SorCtl) * 2, cudaHostAllocMapped));
    CUC(cudaHostAlloc((void**)&c->h_red, sizeof(StepRed), cudaHostAllocDefault));
    memset(c->h_ctl, 0, sizeof(SorCtl) * 2);
    {   // device-side control block points at its host twin
        SorCtl init;
        memset(&init, 0, sizeof init);
        CUC(cudaHostGetDevicePointer((void**)&init.mirror, c->h_ctl, 0));
        CUC(cudaMemcpyAsync(c->d_ctl, &init, sizeof init, cudaMemcpyHostToDevice, c->stream));
    }
    memset(c->h_red, 0, sizeof(StepRed));
    c->red_valid = true;                                       // all-zero fields: maxima 0, sum p^2 0
    CUC(cudaEventCreate(&c->ev0));
    CUC(cudaEventCreate(&c->ev1));
    CUC(cudaEventCreateWithFlags(&c->ev_chk[0], cudaEventDisableTiming));
    CUC(cudaEventCreateWithFlags(&c->ev_chk[1], cudaEventDisableTiming));
    CUC(cudaStreamSynchronize(c->stream));
#undef CUC
    *out = c;
    return NSB_OK;
}

extern "C" int nsb_get_params(const nsb_ctx* c, nsb_params* out) {
    if (!c || !out) return NSB_ERR_ARG;
    *out = c->P;
    return NSB_OK;
}

extern "C" void* nsb_stream(const nsb_ctx* c) { return c ? (void*)c->stream : NULL; }

extern "C" double nsb_sor_seconds(const nsb_ctx* c) { return c ? c->sor_seconds : 0.0; }

// ------------------------------------------------------------------ dt, gamma
static int refresh_reductions(nsb_ctx* c) {
    if (c->red_valid) return NSB_OK;
    CU(c, nsb_launch_reduce_uvp(c));
    CU(c, cudaMemcpyAsync(c->h_red, c->d_red, sizeof(StepRed), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->red_valid = true;
    return NSB_OK;
}

extern "C" int nsb_max_uv(nsb_ctx* c, double* u_max, double* v_max) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    int r = refresh_reductions(c);
    if (r) return r;
    if (u_max) *u_max = c->h_red->u_max;
    if (v_max) *v_max = c->h_red->v_max;
    return NSB_OK;
}

// main.c:91-92 with n_min (io.c:141-161): min starts from the first argument, strict `>` compares.
extern "C" int nsb_dt_gamma(const nsb_ctx* c, double u_max, double v_max, double* dt, double* gamma) {
    if (!c) return NSB_ERR_ARG;
    const double dx = c->dx, dy = c->dy;
    double m = c->P.Re / 2.0 / (1.0 / dx / dx + 1.0 / dy / dy);
    const double cu = dx / fabs(u_max), cv = dy / fabs(v_max);
    if (m > cu) m = cu;
    if (m > cv) m = cv;
    const double d = c->P.tau * m;
    if (dt) *dt = d;
    if (gamma) *gamma = fmax(u_max * d / dx, v_max * d / dy);
    return NSB_OK;
}

// ------------------------------------------------------------------ boundary conditions
extern "C" int nsb_set_inflow(nsb_ctx* c, int side, double u_fix, double v_fix) {
    if (!c) return NSB_ERR_ARG;
    if (side < 0 || side > 3) return -1;                       // boundaries.c:34-35
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_set_inflow(c, side, u_fix, v_fix));
    c->red_valid = false;
    return NSB_OK;
}

extern "C" int nsb_set_noslip(nsb_ctx* c, int side) { return nsb_set_inflow(c, side, 0.0, 0.0); }

static int apply_bc_keep_reductions(nsb_ctx* c, double t) {
    double lid;
    if (c->P.problem == 1) lid = 1.0;                          // main.c:95-99
    else if (c->P.problem == 2) lid = sin(c->P.f * t);         // main.c:100-104
    else return fail(c, NSB_ERR_PROBLEM, "Unknown probem type (see parameters.txt).");
    CU(c, nsb_launch_set_inflow(c, NSB_LEFT, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_RIGHT, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_BOTTOM, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_TOP, lid, 0.0));
    return NSB_OK;
}

extern "C" int nsb_apply_bc(nsb_ctx* c, double t) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    int r = apply_bc_keep_reductions(c, t);
    c->red_valid = false;
    return r;
}

// ------------------------------------------------------------------ F, G, RHS, update
extern "C" int nsb_fg(nsb_ctx* c, double dt, double gamma) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_fg(c, dt, gamma));
    return NSB_OK;
}

extern "C" int nsb_rhs(nsb_ctx* c, double dt) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_rhs(c, dt));
    return NSB_OK;
}

extern "C" int nsb_fg_rhs(nsb_ctx* c, double dt, double gamma) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_fg_rhs(c, dt, gamma));
    return NSB_OK;
}

extern "C" int nsb_update_uv(nsb_ctx* c, double dt) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_update_uv(c, dt));
    CU(c, cudaMemcpyAsync(c->h_red, c->d_red, sizeof(StepRed), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->red_valid = true;
    return NSB_OK;
}

// ------------------------------------------------------------------ SOR drive loop
// The reference's host loop (main.cu:684-719) synchronises six times and copies the residual to
// the host in every iteration.  Here the convergence test runs on the device (the iteration that
// meets the threshold raises a flag; iterations already queued behind it return immediately), and
// the host only looks at the control block once per batch, always one batch behind the queue so
// the GPU never idles.
static int pick_engine(const nsb_ctx* c) {
    int e = c->O.sor_engine;
    if (e == NSB_SOR_AUTO) {
        const char* env = getenv("NSB_SOR_ENGINE");
        if (env) e = atoi(env);
    }
    if (e == NSB_SOR_AUTO) e = nsb_sor_resident_supported(c) ? NSB_SOR_RESIDENT : NSB_SOR_STREAM;
    if (e == NSB_SOR_RESIDENT && !nsb_sor_resident_supported(c)) e = NSB_SOR_STREAM;
    return e;
}

static int sor_drive(nsb_ctx* c, int max_it, int check, int* iters_out, double* res_out, double* normp_out,
                     int* converged_out) {
    int r = refresh_reductions(c);                             // sum p^2 for norm_p (main.cu:675)
    if (r) return r;
    const int engine = pick_engine(c);
    c->last_sor_engine = engine;
    CU(c, cudaEventRecord(c->ev0, c->stream));
    c->h_ctl[0].iters = 0; c->h_ctl[0].converged = 0;          // the begin kernel publishes the same
    CU(c, nsb_launch_sor_begin(c));
    int iters = 0, converged = 0;
    if (engine == NSB_SOR_RESIDENT) {
        CU(c, nsb_launch_sor_resident(c, max_it, check));
        CU(c, cudaMemcpyAsync(c->h_ctl, c->d_ctl, sizeof(SorCtl), cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        iters = c->h_ctl[0].iters;
        converged = c->h_ctl[0].converged;
    } else {
        // The host runs ahead of the GPU by at most `depth` launches and watches the progress the
        // closing block of every iteration publishes to pinned host memory (sor_publish).
        const int depth = c->O.sor_check_batch > 0 ? c->O.sor_check_batch : 24;
        int queued = 0;
        const int start = c->pcur;
        volatile SorCtl* hv = c->h_ctl;                       // written by the device (zero-copy)
        while (queued < max_it) {
            if (engine == NSB_SOR_STREAM)
                CU(c, nsb_launch_sor_stream_iter(c, queued, c->pbuf[(start + queued) & 1],
                                                 c->pbuf[(start + queued + 1) & 1], check));
            else
                CU(c, nsb_launch_sor_simple_iter(c, queued, c->pbuf[start], check));
            queued++;
            if (check) {
                while (!hv->converged && queued - hv->iters > depth) {
                    if (cudaStreamQuery(c->stream) != cudaErrorNotReady) break;
                }
                if (hv->converged) break;
            }
        }
        CU(c, cudaStreamSynchronize(c->stream));
        // the flag may have been raised inside the last batch(es) we did not look at
        CU(c, cudaMemcpy(&c->h_ctl[0], c->d_ctl, sizeof(SorCtl), cudaMemcpyDeviceToHost));
        iters = c->h_ctl[0].iters;
        converged = c->h_ctl[0].converged;
        if (engine == NSB_SOR_STREAM) {
            c->pcur = (start + iters) & 1;
            if (iters > 0) CU(c, nsb_launch_sor_finish_ghosts(c, c->pbuf[c->pcur], c->pbuf[c->pcur ^ 1]));
        }
    }
    CU(c, cudaEventRecord(c->ev1, c->stream));
    CU(c, cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->sor_seconds += (double)ms * 1e-3;
    if (iters_out) *iters_out = iters;
    if (res_out) *res_out = c->h_ctl[0].res;
    if (normp_out) *normp_out = c->h_ctl[0].norm_p;
    if (converged_out) *converged_out = converged;
    return NSB_OK;
}

extern "C" int nsb_sor(nsb_ctx* c, int* iters, double* res, double* norm_p) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    int it = 0, conv = 0;
    int r = sor_drive(c, c->P.max_it, 1, &it, res, norm_p, &conv);
    if (r) return r;
    if (iters) *iters = it;
    return conv ? it : NSB_NOT_CONVERGED;                      // main.cu:717 / :725
}

extern "C" int nsb_sor_iterations(nsb_ctx* c, int n, double* res) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    int it = 0;
    int r = sor_drive(c, n, 0, &it, res, NULL, NULL);
    c->red_valid = false;                                      // p changed outside a time step
    return r;
}

// ------------------------------------------------------------------ whole step / whole run
extern "C" int nsb_step(nsb_ctx* c, double t, nsb_step_info* info) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    double u_max, v_max, dt, gamma;
    int r = nsb_max_uv(c, &u_max, &v_max);                     // main.c:89-90
    if (r) return r;
    nsb_dt_gamma(c, u_max, v_max, &dt, &gamma);                // main.c:91-92
    r = apply_bc_keep_reductions(c, t);                        // main.c:95-108
    if (r) return r;
    CU(c, nsb_launch_fg_rhs(c, dt, gamma));                    // main.c:112-120
    int iters = 0, conv = 0;
    double res = 0.0, norm_p = 0.0;
    r = sor_drive(c, c->P.max_it, 1, &iters, &res, &norm_p, &conv);   // main.c:123
    if (r) return r;
    r = nsb_update_uv(c, dt);                                  // main.c:131-136
    if (r) return r;
    if (info) {
        info->t = t; info->dt = dt; info->gamma = gamma; info->u_max = u_max; info->v_max = v_max;
        info->norm_p = norm_p; info->res = res; info->sor_iters = iters;
        info->sor_ret = conv ? iters : NSB_NOT_CONVERGED;
    }
    return NSB_OK;
}

extern "C" int nsb_run(nsb_ctx* c, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end) {
    if (!c) return NSB_ERR_ARG;
    double t = 0.0;
    int n = 0;
    while (t < c->P.T && (max_steps <= 0 || n < max_steps)) {  // main.c:86
        nsb_step_info info;
        int r = nsb_step(c, t, &info);
        if (r) return r;
        if (log && n < max_log) log[n] = info;
        t += info.dt;                                          // main.c:145
        n++;
    }
    if (steps) *steps = n;
    if (t_end) *t_end = t;
    return NSB_OK;
}

extern "C" int nsb_get_center(nsb_ctx* c, double* u_c, double* v_c) {
    if (!c) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    const size_t o = nsb_index(c->g, c->P.i_max / 2, c->P.j_max / 2);   // main.c:148-149
    double uv[2];
    CU(c, cudaMemcpyAsync(&uv[0], c->u + o, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(&uv[1], c->v + o, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (u_c) *u_c = uv[0];
    if (v_c) *v_c = uv[1];
    return NSB_OK;
}

// ------------------------------------------------------------------ transfers
static double* field_ptr(nsb_ctx* c, int field) {
    switch (field) {
        case NSB_U: return c->u;
        case NSB_V: return c->v;
        case NSB_P: return c->pbuf[c->pcur];
        case NSB_F: return c->F;
        case NSB_G: return c->G;
        case NSB_RHS: return c->rhs;
    }
    return NULL;
}

extern "C" int nsb_device_view(nsb_ctx* c, int field, void** dptr, long long* pitch, int* col0, int* row0) {
    if (!c) return NSB_ERR_ARG;
    double* p = field_ptr(c, field);
    if (!p) return fail(c, NSB_ERR_ARG, "unknown field %d", field);
    if (dptr) *dptr = p;
    if (pitch) *pitch = c->g.pitch;
    if (col0) *col0 = NSB_COL0;
    if (row0) *row0 = NSB_ROW0;
    return NSB_OK;
}

extern "C" int nsb_upload(nsb_ctx* c, int field, const double* host, int layout) {
    if (!c || !host) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    double* dst = field_ptr(c, field);
    if (!dst) return fail(c, NSB_ERR_ARG, "unknown field %d", field);
    const size_t nx = (size_t)c->g.ni + 2, ny = (size_t)c->g.nj + 2;
    if (layout == NSB_LAYOUT_XFAST) {
        CU(c, cudaMemcpy2DAsync(dst + NSB_ROW0 * c->g.pitch + NSB_COL0, c->g.pitch * sizeof(double), host,
                                nx * sizeof(double), nx * sizeof(double), ny, cudaMemcpyHostToDevice, c->stream));
    } else if (layout == NSB_LAYOUT_REF) {
        CU(c, cudaMemcpyAsync(c->scratch, host, nx * ny * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        CU(c, nsb_launch_to_xfast(c, c->scratch, dst));
    } else {
        return fail(c, NSB_ERR_ARG, "unknown layout %d", layout);
    }
    CU(c, cudaStreamSynchronize(c->stream));
    c->red_valid = false;
    return NSB_OK;
}

extern "C" int nsb_download(nsb_ctx* c, int field, double* host, int layout) {
    if (!c || !host) return NSB_ERR_ARG;
    CU(c, cudaSetDevice(c->device));
    const double* src = field_ptr(c, field);
    if (!src) return fail(c, NSB_ERR_ARG, "unknown field %d", field);
    const size_t nx = (size_t)c->g.ni + 2, ny = (size_t)c->g.nj + 2;
    if (layout == NSB_LAYOUT_XFAST) {
        CU(c, cudaMemcpy2DAsync(host, nx * sizeof(double), src + NSB_ROW0 * c->g.pitch + NSB_COL0,
                                c->g.pitch * sizeof(double), nx * sizeof(double), ny, cudaMemcpyDeviceToHost,
                                c->stream));
    } else if (layout == NSB_LAYOUT_REF) {
        CU(c, nsb_launch_from_xfast(c, src, c->scratch));
        CU(c, cudaMemcpyAsync(host, c->scratch, nx * ny * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    } else {
        return fail(c, NSB_ERR_ARG, "unknown layout %d", layout);
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return NSB_OK;
}

// ------------------------------------------------------------------ output(), io.c:61-120
// Row order of the files is j-major with i running inside a row: exactly the x-fastest layout, so
// the three grids are downloaded as they lie and formatted row by row.
extern "C" int nsb_output(nsb_ctx* c, double t, const char* prefix) {
    if (!c || !prefix) return NSB_ERR_ARG;
    const size_t nx = (size_t)c->g.ni + 2, ny = (size_t)c->g.nj + 2;
    double* host = (double*)malloc(nx * ny * sizeof(double));
    if (!host) return fail(c, NSB_ERR_ARG, "out of host memory");
    const char* suffix[3] = {"_u.txt", "_v.txt", "_p.txt"};
    const int fields[3] = {NSB_U, NSB_V, NSB_P};
    FILE* fp[3] = {NULL, NULL, NULL};
    char name[1024];
    for (int k = 0; k < 3; k++) {
        snprintf(name, sizeof name, "%s%s", prefix, suffix[k]);
        fp[k] = fopen(name, "w");
    }
    if (!fp[0] || !fp[1] || !fp[2]) {
        for (int k = 0; k < 3; k++) if (fp[k]) fclose(fp[k]);
        free(host);
        return fail(c, NSB_ERR_IO, "Error opening one ore more output files. Make sure the directory 'out' exists.");
    }
    int rc = NSB_OK;
    for (int k = 0; k < 3 && rc == NSB_OK; k++) {
        rc = nsb_download(c, fields[k], host, NSB_LAYOUT_XFAST);
        if (rc) break;
        fprintf(fp[k], "%.5f\n%.5f\n%.5f\n", t, c->P.a, c->P.b);          // io.c:89-99
        const size_t ncol = (k == 0) ? nx - 1 : nx;                     // u omits i = i_max+1
        const size_t nrow = (k == 1) ? ny - 1 : ny;                     // v omits j = j_max+1 ...
        for (size_t j = 0; j < ny; j++) {
            if (j < nrow)
                for (size_t i = 0; i < ncol; i++) fprintf(fp[k], "%.5f ", host[j * nx + i]);
            fputc('\n', fp[k]);                                         // ... but keeps its newline
        }
    }
    for (int k = 0; k < 3; k++) fclose(fp[k]);
    free(host);
    return rc;
}
