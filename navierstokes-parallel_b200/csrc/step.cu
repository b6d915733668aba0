// step.cu -- the time step in the reference's order (src/serial/main.c:88-145 ; src/parallel/
// main.cu:1046-1103), written as a sequence of SUB-STEPS that only launch work on their own rank.
//
// Why sub-steps: with y-slabs every exchange is "push into the neighbour, flag, neighbour waits" inside
// kernels (slab.cuh).  A sub-step never waits for something that a LATER sub-step of another rank
// produces, so the same code drives
//   - one context (world == 1),
//   - one rank per process (torchrun / one process per GPU): each process runs all sub-steps in order,
//   - a group of ranks in ONE process (driver/main.c with NSB_GPUS, and the lock-step emulation of
//     several slabs on one device that the 1-GPU test-suite uses): sub-step k on every rank before
//     sub-step k+1 on any.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nsb_internal.h"

#define CU(c, call)                                                                                    \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess)                                                                        \
            return nsb_fail((c), NSB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                            __FILE__, __LINE__);                                                       \
    } while (0)

// ------------------------------------------------------------------ reductions: max u, max v, sum p^2
int nsb_ph_reduce_launch(nsb_ctx* c) {
    if (c->red_valid) return NSB_OK;
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_reduce_uvp(c));
    return NSB_OK;
}

int nsb_ph_reduce_finish(nsb_ctx* c) {
    if (c->red_valid) return NSB_OK;
    CU(c, cudaSetDevice(c->device));
    if (c->world > 1) CU(c, nsb_launch_red_close(c));
    CU(c, cudaMemcpyAsync(c->h_red, c->d_red, sizeof(StepRed), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->red_round++;
    c->red_valid = true;
    return NSB_OK;
}

extern "C" int nsb_max_uv(nsb_ctx* c, double* u_max, double* v_max) {
    if (!c) return NSB_ERR_ARG;
    int r = nsb_ph_reduce_launch(c);
    if (r == NSB_OK) r = nsb_ph_reduce_finish(c);
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

int nsb_ph_bc(nsb_ctx* c, double t) {
    double lid;
    if (c->P.problem == 1) lid = 1.0;                          // main.c:95-99
    else if (c->P.problem == 2) lid = sin(c->P.f * t);         // main.c:100-104
    else return nsb_fail(c, NSB_ERR_PROBLEM, "Unknown probem type (see parameters.txt).");
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_set_inflow(c, NSB_LEFT, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_RIGHT, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_BOTTOM, 0.0, 0.0));
    CU(c, nsb_launch_set_inflow(c, NSB_TOP, lid, 0.0));
    return NSB_OK;
}

extern "C" int nsb_apply_bc(nsb_ctx* c, double t) {
    if (!c) return NSB_ERR_ARG;
    int r = nsb_ph_bc(c, t);
    c->red_valid = false;
    return r;
}

// ------------------------------------------------------------------ F, G, RHS, update (single-slab entry points)
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

int nsb_ph_update_launch(nsb_ctx* c, double dt) {
    CU(c, cudaSetDevice(c->device));
    CU(c, nsb_launch_update_uv(c, dt));
    c->red_valid = false;
    return NSB_OK;
}

int nsb_ph_update_finish(nsb_ctx* c) { return nsb_ph_reduce_finish(c); }

extern "C" int nsb_update_uv(nsb_ctx* c, double dt) {
    if (!c) return NSB_ERR_ARG;
    int r = nsb_ph_update_launch(c, dt);
    if (r == NSB_OK) r = nsb_ph_update_finish(c);
    return r;
}

// ------------------------------------------------------------------ SOR drive loop
// The reference's host loop (main.cu:684-719) synchronises six times and copies the residual to
// the host in every iteration.  Here the convergence test runs on the device: the closing of the
// iteration that meets the threshold raises a flag, iterations already queued behind it return at
// once, and the host only watches the control block the closing publishes to pinned host memory,
// running ahead of the device by a bounded number of launches.  `cs` holds the n ranks driven by
// THIS process (n == 1 unless it is a single-process group); all ranks of a decomposition see the
// same residual bits, so they all stop after the same iteration.
static int pick_engine(const nsb_ctx* c) {
    int e = c->O.sor_engine;
    if (e == NSB_SOR_AUTO) {
        const char* env = getenv("NSB_SOR_ENGINE");
        if (env) e = atoi(env);
    }
    if (c->world > 1) {                                        // slab-aware engines: the two streaming ones
        if (e == NSB_SOR_STREAM_LAUNCH || !nsb_sor_persist_supported(c)) return NSB_SOR_STREAM_LAUNCH;
        return NSB_SOR_STREAM;
    }
    if (e == NSB_SOR_SERIAL_ORDER) return e;
    if (e == NSB_SOR_AUTO) e = nsb_sor_resident_supported(c) ? NSB_SOR_RESIDENT : NSB_SOR_STREAM;
    if (e == NSB_SOR_RESIDENT && !nsb_sor_resident_supported(c)) e = NSB_SOR_STREAM;
    if (e == NSB_SOR_STREAM && !nsb_sor_persist_supported(c)) e = NSB_SOR_STREAM_LAUNCH;
    return e;
}

// sub-step: SOR entry.  Needs sum p^2 (reductions valid) -- threshold, flags, p halo push.
int nsb_ph_sor_begin(nsb_ctx* c) {
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaEventRecord(c->ev0, c->stream));
    c->h_ctl[0].iters = 0; c->h_ctl[0].converged = 0;          // the begin kernel publishes the same
    CU(c, nsb_launch_sor_begin(c));
    CU(c, nsb_launch_halo_push(c, NSB_X_P));
    return NSB_OK;
}

static int sor_iterate(nsb_ctx** cs, int n, int max_it, int check) {
    nsb_ctx* c0 = cs[0];
    int engine = pick_engine(c0);
    if (engine == NSB_SOR_RESIDENT) {
        CU(c0, cudaSetDevice(c0->device));
        const cudaError_t e = nsb_launch_sor_resident(c0, max_it, check);
        if (e == cudaSuccess) { c0->last_sor_engine = engine; return NSB_OK; }
        // The device refused the cluster / cooperative launch (nothing was enqueued): remember it in this context
        // and run THIS solve with the streaming sweep instead of failing the step.
        cudaGetLastError();
        c0->resident_state = -1;
        engine = pick_engine(c0);
        if (engine == NSB_SOR_RESIDENT) return nsb_fail(c0, NSB_ERR_CUDA, "resident SOR launch failed: %s", cudaGetErrorString(e));
    }
    for (int r = 0; r < n; r++) cs[r]->last_sor_engine = engine;
    if (engine == NSB_SOR_SERIAL_ORDER) {
        CU(c0, cudaSetDevice(c0->device));
        CU(c0, nsb_launch_sor_lex(c0, max_it, check));
        return NSB_OK;
    }
    if (engine == NSB_SOR_STREAM) {
        // Every iteration of the solve in one launch per rank: max_it / 2 passes of TWO iterations each (temporal
        // blocking), then one single-iteration pass when max_it is odd.  Both launches are enqueued at once: the
        // second looks at the device-side `converged` flag and returns without work if the first one met the threshold.
        for (int r = 0; r < n; r++) {
            nsb_ctx* c = cs[r];
            CU(c, cudaSetDevice(c->device));
            const int k = nsb_sor_pass_k(c);
            const int passes = k > 1 ? max_it / k : 0;
            const int tail = max_it - passes * k;
            c->st_k = k; c->st_passes = passes; c->st_tail = tail; c->st_buf0 = c->pcur;
            if (passes > 0) CU(c, nsb_launch_sor_persist(c, k, passes, 0, check, c->pcur, 0, 0));
            if (tail > 0) CU(c, nsb_launch_sor_persist(c, 1, tail, passes * k, check, (c->pcur + passes) % 3, 0, passes > 0 ? 1 : 0));
        }
        return NSB_OK;
    }
    const int depth = c0->O.sor_check_batch > 0 ? c0->O.sor_check_batch : 24;
    volatile SorCtl* hv = c0->h_ctl;                           // written by the device (zero-copy)
    int queued = 0;
    while (queued < max_it) {
        for (int r = 0; r < n; r++) {
            nsb_ctx* c = cs[r];
            if (n > 1) CU(c, cudaSetDevice(c->device));
            if (engine == NSB_SOR_STREAM_LAUNCH)
                CU(c, nsb_launch_sor_stream_iter(c, queued, c->pbuf[(c->pcur + queued) & 1],
                                                 c->pbuf[(c->pcur + queued + 1) & 1], check));
            else
                CU(c, nsb_launch_sor_simple_iter(c, queued, c->pbuf[c->pcur], check));
        }
        if (c0->world > 1)
            for (int r = 0; r < n; r++) {
                nsb_ctx* c = cs[r];
                if (n > 1) CU(c, cudaSetDevice(c->device));
                CU(c, nsb_launch_sor_close(c, queued, check));
            }
        queued++;
        if (check) {
            while (!hv->converged && queued - hv->iters > depth) {
                if (cudaStreamQuery(c0->stream) != cudaErrorNotReady) break;
            }
            if (hv->converged) break;
        }
    }
    return NSB_OK;
}

// sub-step: SOR exit on one rank -- read the control block, settle the ping-pong buffer and the counters.
int nsb_ph_sor_end(nsb_ctx* c, int unused) {
    (void)unused;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    CU(c, cudaMemcpy(&c->h_ctl[1], c->d_ctl, sizeof(SorCtl), cudaMemcpyDeviceToHost));
    {
        int werr = 0;
        CU(c, cudaMemcpy(&werr, &c->d_sync->error, sizeof werr, cudaMemcpyDeviceToHost));
        if (werr) return nsb_fail(c, NSB_ERR_CUDA, "SOR: a wait inside the sweep kernel timed out (rank %d)", c->rank);
    }
    const int iters = c->h_ctl[1].iters;
    if (c->last_sor_engine == NSB_SOR_STREAM) {
        // Which of the three rotating buffers holds iterate `iters`?  Pass q of the K-iteration launch read buffer
        // buf0 + q and wrote buf0 + q + 1 (mod 3).  If the solve stopped inside a pass (its first sub-iteration met the
        // threshold), that iterate was never stored: the pass's input is still intact (the buffer is rewritten two passes
        // later, and nothing runs that far after convergence), so ONE single-iteration pass from it reproduces the
        // iterate bit for bit (same task shape, same arithmetic); it does not close anything again.
        const int k = c->st_k, in_passes = c->st_passes * k;
        int buf;
        if (iters <= in_passes) {
            const int full = iters / k, rest = iters - full * k;
            buf = (c->st_buf0 + full) % 3;
            if (rest > 0) {
                CU(c, nsb_launch_sor_persist(c, 1, rest, iters - rest, 0, buf, 1, 0));
                CU(c, cudaStreamSynchronize(c->stream));
                buf = (buf + rest) % 3;
            }
        } else {
            buf = (c->st_buf0 + c->st_passes + (iters - in_passes)) % 3;
        }
        // between solves the pressure lives in pbuf[0] or pbuf[1] (the other engines flip between those two):
        // rename the buffers instead of copying.  Every rank of a decomposition sees the same iteration counts,
        // so all ranks rename alike and the peers' tables follow.
        if (buf == 2) {
            const int other = c->pcur ^ 1;
            double* tmp = c->pbuf[2]; c->pbuf[2] = c->pbuf[other]; c->pbuf[other] = tmp;
            for (int r = 0; r < c->world; r++) { double* q = c->peer[r].p[2]; c->peer[r].p[2] = c->peer[r].p[other]; c->peer[r].p[other] = q; }
            buf = other;
        }
        c->pcur = buf;
    } else if (c->last_sor_engine == NSB_SOR_STREAM_LAUNCH) {
        c->pcur = (c->pcur + iters) & 1;
        // per-launch sweeps: the other buffer still holds the previous iterate, which the ghost
        // fix-up reads.  The persistent sweep stores the ghost cells itself (iterations that ran past
        // the converged one may have overwritten the other buffer).
        if (iters > 0)
            CU(c, nsb_launch_sor_finish_ghosts(c, c->pbuf[c->pcur], c->pbuf[c->pcur ^ 1]));
    }
    if (c->last_sor_engine != NSB_SOR_STREAM) c->seq += (unsigned long long)iters;   // (the persistent launch advanced it)
    c->red_round += (unsigned long long)iters;
    CU(c, cudaEventRecord(c->ev1, c->stream));
    CU(c, cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->sor_seconds += (double)ms * 1e-3;
    return NSB_OK;
}

// Whole solve on the ranks of this process.  Sub-step order: [reductions] begin, wait P, iterate, end.
static int sor_solve(nsb_ctx** cs, int n, int max_it, int check) {
    int r;
    for (int k = 0; k < n; k++) if ((r = nsb_ph_reduce_launch(cs[k]))) return r;   // sum p^2 (main.cu:675)
    for (int k = 0; k < n; k++) if ((r = nsb_ph_reduce_finish(cs[k]))) return r;
    for (int k = 0; k < n; k++) if ((r = nsb_ph_sor_begin(cs[k]))) return r;
    for (int k = 0; k < n; k++) {
        CU(cs[k], cudaSetDevice(cs[k]->device));
        CU(cs[k], nsb_launch_halo_wait(cs[k], NSB_X_P));
    }
    if ((r = sor_iterate(cs, n, max_it, check))) return r;
    for (int k = 0; k < n; k++) if ((r = nsb_ph_sor_end(cs[k], 0))) return r;
    return NSB_OK;
}

extern "C" int nsb_sor(nsb_ctx* c, int* iters, double* res, double* norm_p) {
    if (!c) return NSB_ERR_ARG;
    int r = sor_solve(&c, 1, c->P.max_it, 1);
    if (r) return r;
    const SorCtl& k = c->h_ctl[1];
    if (iters) *iters = k.iters;
    if (res) *res = k.res;
    if (norm_p) *norm_p = k.norm_p;
    return k.converged ? k.iters : NSB_NOT_CONVERGED;          // main.cu:717 / :725
}

extern "C" int nsb_sor_iterations(nsb_ctx* c, int n, double* res) {
    if (!c) return NSB_ERR_ARG;
    int r = sor_solve(&c, 1, n, 0);
    c->red_valid = false;                                      // p changed outside a time step
    if (res) *res = c->h_ctl[1].res;
    return r;
}

// ------------------------------------------------------------------ whole step / whole run
// main.c:88-145 on the n ranks of this process, sub-step by sub-step.
static int step_ranks(nsb_ctx** cs, int n, double t, nsb_step_info* info) {
    int r;
    nsb_ctx* c0 = cs[0];
#define ALL(expr)                                            \
    for (int k = 0; k < n; k++) {                            \
        nsb_ctx* c = cs[k];                                  \
        if ((r = (expr)) != NSB_OK) return r;                \
    }
#define ALLCU(call)                                          \
    for (int k = 0; k < n; k++) {                            \
        nsb_ctx* c = cs[k];                                  \
        CU(c, cudaSetDevice(c->device));                     \
        CU(c, call);                                         \
    }
    ALL(nsb_ph_reduce_launch(c));                              // main.c:89-90 (cached by the last update)
    ALL(nsb_ph_reduce_finish(c));
    const double u_max = c0->h_red->u_max, v_max = c0->h_red->v_max;
    double dt, gamma;
    nsb_dt_gamma(c0, u_max, v_max, &dt, &gamma);               // main.c:91-92
    ALL(nsb_ph_bc(c, t));                                      // main.c:95-108
    ALLCU(nsb_launch_halo_push(c, NSB_X_UV));
    ALLCU(nsb_launch_halo_wait(c, NSB_X_UV));
    ALLCU(nsb_launch_fg(c, dt, gamma));                        // main.c:112
    ALLCU(nsb_launch_halo_push(c, NSB_X_G));
    ALLCU(nsb_launch_halo_wait(c, NSB_X_G));
    ALLCU(nsb_launch_rhs(c, dt));                              // main.c:116-120
    ALLCU(nsb_launch_halo_push(c, NSB_X_RHS));
    ALLCU(nsb_launch_halo_wait(c, NSB_X_RHS));
    if ((r = sor_solve(cs, n, c0->P.max_it, 1))) return r;     // main.c:123
    ALL(nsb_ph_update_launch(c, dt));                          // main.c:131-136
    ALL(nsb_ph_update_finish(c));
    if (c0->world > 1)
        for (int k = 0; k < n; k++)
            if (nsb_slab_error(cs[k]) != 0)
                return nsb_fail(cs[k], NSB_ERR_CUDA, "rank %d: a wait on a neighbouring slab timed out", cs[k]->rank);
#undef ALL
#undef ALLCU
    if (info) {
        const SorCtl& k = c0->h_ctl[1];
        info->t = t; info->dt = dt; info->gamma = gamma; info->u_max = u_max; info->v_max = v_max;
        info->norm_p = k.norm_p; info->res = k.res; info->sor_iters = k.iters;
        info->sor_ret = k.converged ? k.iters : NSB_NOT_CONVERGED;
    }
    return NSB_OK;
}

extern "C" int nsb_step(nsb_ctx* c, double t, nsb_step_info* info) {
    if (!c) return NSB_ERR_ARG;
    return step_ranks(&c, 1, t, info);
}

static int run_ranks(nsb_ctx** cs, int n, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end) {
    double t = 0.0;
    int k = 0;
    while (t < cs[0]->P.T && (max_steps <= 0 || k < max_steps)) {  // main.c:86
        nsb_step_info info;
        int r = step_ranks(cs, n, t, &info);
        if (r) return r;
        if (log && k < max_log) log[k] = info;
        t += info.dt;                                          // main.c:145
        k++;
    }
    if (steps) *steps = k;
    if (t_end) *t_end = t;
    return NSB_OK;
}

extern "C" int nsb_run(nsb_ctx* c, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end) {
    if (!c) return NSB_ERR_ARG;
    return run_ranks(&c, 1, max_steps, log, max_log, steps, t_end);
}

// ------------------------------------------------------------------ single-process group of slabs
struct nsb_group {
    int n;
    nsb_ctx* c[NSB_MAX_RANKS];
    cudaStream_t shared_stream;       // all ranks on one device (lock-step emulation): one stream orders them
    char err[512];
};

static char g_group_err[512] = "";

extern "C" const char* nsb_group_last_error(const nsb_group* g) {
    if (!g) return g_group_err;
    for (int k = 0; k < g->n; k++) if (g->c[k] && g->c[k]->err[0]) return g->c[k]->err;
    return g->err;
}

extern "C" void nsb_group_destroy(nsb_group* g) {
    if (!g) return;
    for (int k = 0; k < g->n; k++) if (g->c[k]) nsb_destroy(g->c[k]);
    if (g->shared_stream) cudaStreamDestroy(g->shared_stream);
    free(g);
}

// devices[k] = CUDA device of slab k (NULL: device k).  Equal devices are allowed: the slabs then run
// in lock step on one stream, which is how the decomposition is tested on a single GPU.
extern "C" int nsb_group_create(const nsb_params* P, const nsb_options* opt, int n, const int* devices, nsb_group** out) {
    if (!P || !out || n < 1 || n > NSB_MAX_RANKS) { snprintf(g_group_err, 512, "nsb_group_create: bad argument"); return NSB_ERR_ARG; }
    *out = NULL;
    nsb_group* g = (nsb_group*)calloc(1, sizeof(nsb_group));
    if (!g) return NSB_ERR_ARG;
    g->n = n;
    bool same = true;
    for (int k = 1; k < n; k++) if ((devices ? devices[k] : k) != (devices ? devices[0] : 0)) same = false;
    nsb_options o;
    if (opt) o = *opt; else nsb_default_options(&o);
    if (same && n > 1) {
        if (cudaSetDevice(devices ? devices[0] : 0) != cudaSuccess ||
            cudaStreamCreateWithFlags(&g->shared_stream, cudaStreamNonBlocking) != cudaSuccess) {
            snprintf(g_group_err, 512, "cannot create the shared stream (no usable CUDA device?)");
            free(g);
            return NSB_ERR_CUDA;
        }
        o.stream = (void*)g->shared_stream;
    }
    unsigned char descs[NSB_MAX_RANKS * sizeof(NsbSlabDesc)];
    for (int k = 0; k < n; k++) {
        o.device = devices ? devices[k] : k;
        int r = nsb_create_slab_internal(P, &o, k, n, &g->c[k]);
        if (r != NSB_OK) {
            snprintf(g_group_err, 512, "slab %d: %s", k, nsb_last_error(NULL));
            nsb_group_destroy(g);
            return r;
        }
        nsb_slab_export(g->c[k], descs + (size_t)k * sizeof(NsbSlabDesc));
    }
    for (int k = 0; k < n; k++) {
        int r = nsb_slab_connect(g->c[k], descs);
        if (r != NSB_OK) {
            snprintf(g_group_err, 512, "slab %d: %s", k, nsb_last_error(g->c[k]));
            nsb_group_destroy(g);
            return r;
        }
    }
    *out = g;
    return NSB_OK;
}

extern "C" int nsb_group_size(const nsb_group* g) { return g ? g->n : 0; }
extern "C" nsb_ctx* nsb_group_ctx(nsb_group* g, int k) { return (g && k >= 0 && k < g->n) ? g->c[k] : NULL; }

extern "C" int nsb_group_step(nsb_group* g, double t, nsb_step_info* info) {
    if (!g) return NSB_ERR_ARG;
    return step_ranks(g->c, g->n, t, info);
}

extern "C" int nsb_group_run(nsb_group* g, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end) {
    if (!g) return NSB_ERR_ARG;
    return run_ranks(g->c, g->n, max_steps, log, max_log, steps, t_end);
}

extern "C" int nsb_group_sor(nsb_group* g, int* iters, double* res, double* norm_p) {
    if (!g) return NSB_ERR_ARG;
    int r = sor_solve(g->c, g->n, g->c[0]->P.max_it, 1);
    if (r) return r;
    const SorCtl& k = g->c[0]->h_ctl[1];
    if (iters) *iters = k.iters;
    if (res) *res = k.res;
    if (norm_p) *norm_p = k.norm_p;
    return k.converged ? k.iters : NSB_NOT_CONVERGED;
}

extern "C" int nsb_group_sor_iterations(nsb_group* g, int n_it, double* res) {
    if (!g) return NSB_ERR_ARG;
    int r = sor_solve(g->c, g->n, n_it, 0);
    for (int k = 0; k < g->n; k++) g->c[k]->red_valid = false;
    if (res) *res = g->c[0]->h_ctl[1].res;
    return r;
}

extern "C" double nsb_group_sor_seconds(const nsb_group* g) { return g ? g->c[0]->sor_seconds : 0.0; }

// Fields of the whole domain, host layout as in nsb_upload / nsb_download.  Upload also fills the halo
// rows of every slab, so no exchange is needed before the first kernel.
static int group_xfer(nsb_group* g, int field, double* host, int layout, bool up) {
    nsb_ctx* c0 = g->c[0];
    const size_t nx = (size_t)c0->g.ni + 2, ny = (size_t)c0->J + 2;
    double* x = host;
    if (layout == NSB_LAYOUT_REF) {
        x = (double*)malloc(nx * ny * sizeof(double));
        if (!x) return NSB_ERR_ARG;
        if (up) for (size_t i = 0; i < nx; i++) for (size_t j = 0; j < ny; j++) x[j * nx + i] = host[i * ny + j];
    } else if (layout != NSB_LAYOUT_XFAST) return NSB_ERR_ARG;
    int rc = NSB_OK;
    for (int k = 0; k < g->n && rc == NSB_OK; k++) {
        nsb_ctx* c = g->c[k];
        int lo = up ? 1 - NSB_HALO_ROWS : (c->wall_bottom ? 0 : 1);   // local rows to move
        int hi = up ? c->g.nj + NSB_HALO_ROWS : (c->wall_top ? c->g.nj + 1 : c->g.nj);
        if (c->j_lo + lo < 0) lo = -c->j_lo;
        if (c->j_lo + hi > c->J + 1) hi = c->J + 1 - c->j_lo;
        rc = nsb_xfer_rows(c, field, x + (size_t)(c->j_lo + lo) * nx, lo, hi - lo + 1, up ? 1 : 0);
    }
    if (layout == NSB_LAYOUT_REF) {
        if (!up && rc == NSB_OK) for (size_t i = 0; i < nx; i++) for (size_t j = 0; j < ny; j++) host[i * ny + j] = x[j * nx + i];
        free(x);
    }
    if (up) for (int k = 0; k < g->n; k++) g->c[k]->red_valid = false;
    return rc;
}

extern "C" int nsb_group_upload(nsb_group* g, int field, const double* host, int layout) {
    if (!g || !host) return NSB_ERR_ARG;
    return group_xfer(g, field, (double*)host, layout, true);
}

extern "C" int nsb_group_download(nsb_group* g, int field, double* host, int layout) {
    if (!g || !host) return NSB_ERR_ARG;
    return group_xfer(g, field, host, layout, false);
}

// io.h:31 output() for a decomposition: every slab contributes its rows to the same three files
extern "C" int nsb_group_output(nsb_group* g, double t, const char* prefix) {
    if (!g || !prefix) return NSB_ERR_ARG;
    return nsb_out_run(g->c, g->n, t, prefix, 0);
}
extern "C" int nsb_group_output_async(nsb_group* g, double t, const char* prefix) {
    if (!g || !prefix) return NSB_ERR_ARG;
    return nsb_out_run(g->c, g->n, t, prefix, 1);
}
extern "C" int nsb_group_output_wait(nsb_group* g) {
    if (!g) return NSB_ERR_ARG;
    return nsb_out_wait(g->c[0]);
}

extern "C" int nsb_group_get_center(nsb_group* g, double* u_c, double* v_c) {
    if (!g) return NSB_ERR_ARG;
    const int jc = g->c[0]->J / 2;                              // main.c:148-149 (global row)
    for (int k = 0; k < g->n; k++) {
        nsb_ctx* c = g->c[k];
        const int l = jc - c->j_lo;
        if ((l >= 1 && l <= c->g.nj) || (l == 0 && c->wall_bottom)) return nsb_get_value(c, c->P.i_max / 2, l, u_c, v_c);
    }
    return NSB_ERR_ARG;
}
