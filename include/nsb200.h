/*
 * nsb200.h -- C ABI of the B200-native (sm_100a) lid-driven-cavity time step.
 *
 * This is the drop-in boundary for the hot path of guilherme-webster/NavierStokes-parallel
 * (dt/gamma -> boundary conditions -> F,G -> RHS -> red-black SOR + residual test -> velocity
 * update).  The reference has no FFI of its own; its replaceable surface is the set of C
 * prototypes in src/serial/{io,boundaries,integration,memory}.h that its two `main`s call
 * (src/serial/main.c:86-147, src/parallel/main.cu:1046-1103).  Every entry point below names the
 * reference interface it replaces.  Plain C types only: no CUDA, torch or C++ types cross it.
 *
 * Two groups:
 *   A. context API   -- grids live in HBM for the whole run (the fast path; what driver/main.c,
 *                       the python host mirror and bench.py use).
 *   B. host-grid API -- same names/arguments as the reference functions (`double**` grids owned by
 *                       the caller); each call uploads, runs the CUDA path, downloads.  A reference
 *                       maintainer swaps a call site by adding the `nsb_` prefix.
 *
 * Conventions: every function returns 0 on success unless stated; negative = error, text via
 * nsb_last_error().  No exceptions.  Group A has no global mutable state (the reference's global BLOCK_SIZE,
 * main.cu:26, is not reproduced); group B keeps ONE cached context per process between calls (so a time loop
 * does not reallocate; nsb_host_api_reset() drops it) and is therefore, like the reference, single-threaded.
 * One host thread per context, calls are synchronous from the caller's view.  There is NO CPU fallback: without a usable CUDA
 * device nsb_create fails with NSB_ERR_CUDA.
 */
#ifndef NSB200_H
#define NSB200_H

#ifdef __cplusplus
extern "C" {
#endif

#define NSB_ABI_VERSION 1

/* boundaries.h:16-21 */
enum { NSB_TOP = 0, NSB_BOTTOM = 1, NSB_LEFT = 2, NSB_RIGHT = 3 };
/* grids of memory.h:27 (res is CPU-only scratch in the reference and has no device twin) */
enum { NSB_U = 0, NSB_V = 1, NSB_P = 2, NSB_F = 3, NSB_G = 4, NSB_RHS = 5 };
/* host layouts for nsb_upload / nsb_download */
enum {
    NSB_LAYOUT_REF = 0,    /* reference order: (i_max+2) rows of (j_max+2), element [i][j]   */
    NSB_LAYOUT_XFAST = 1   /* (j_max+2) rows of (i_max+2), element [j][i] (output() row order) */
};
enum {
    NSB_OK = 0,
    NSB_ERR_ARG = -2,
    NSB_ERR_CUDA = -3,
    NSB_ERR_IO = -4,
    NSB_ERR_PROBLEM = -5,   /* problem not 1 or 2: serial main.c:105-108 exits here */
    NSB_NOT_CONVERGED = -1  /* SOR hit max_it: integration.h:165 / main.cu:725 return -1 */
};
/* how x/dx^2-type divisions are evaluated (DESIGN.md "arithmetic") */
enum {
    NSB_DIV_AUTO = 0,   /* multiply by the reciprocal when it is exact (dx^2, dy^2 powers of two:
                           bit-identical to the division), true IEEE division otherwise */
    NSB_DIV_EXACT = 1,  /* always divide, like the reference expression */
    NSB_DIV_RECIP = 2   /* always multiply by the reciprocal (<= 1 ulp per operation off) */
};
/* SOR engines */
enum {
    NSB_SOR_AUTO = 0,
    NSB_SOR_STREAM = 1,   /* fused red+black+residual streaming pass; on a single slab ALL iterations of a
                             solve run in one persistent launch (tasks of iteration n+1 start as soon as
                             their neighbours of iteration n are done), on y-slabs one launch per iteration */
    NSB_SOR_RESIDENT = 2, /* grid held in shared memory (one thread-block cluster, or one CTA per SM for grids
                             up to ~1024^2), all iterations in one launch; falls back to the stream engine */
    NSB_SOR_SIMPLE = 3,   /* one launch per colour + residual launch (validation kernel)        */
    NSB_SOR_STREAM_LAUNCH = 4, /* the streaming pass with one launch per iteration on a single slab too */
    NSB_SOR_SERIAL_ORDER = 5   /* the SERIAL reference's order (integration.c:129-173): one ghost refresh per
                                  iteration, in-place lexicographic sweep run as an anti-diagonal wavefront;
                                  validation only (one CTA), single slab */
};

/* io.h:14 init(): the 15 positional values of a parameter file, same order */
typedef struct nsb_params {
    int problem;
    double f;
    int i_max, j_max;
    double a, b, T, Re, g_x, g_y, tau, omega, epsilon;
    int max_it, n_print;
} nsb_params;

typedef struct nsb_options {
    int device;          /* CUDA device ordinal                                                */
    int div_mode;        /* NSB_DIV_*                                                           */
    int sor_engine;      /* NSB_SOR_*                                                           */
    int sor_chunk_rows;  /* rows per streaming chunk (0 = auto: planned for the grid; with y-slabs it is also the band
                            height of the decomposition, planned for a slab of ceil(j_max / world) rows -- the same on
                            every rank; nsb_slab_info / nsb_sor_plan_info report what was chosen)              */
    int sor_check_batch; /* SOR launches the host may run ahead of the device (0 = auto, 24)      */
    void* stream;        /* cudaStream_t to run on (NULL = library creates its own)              */
    int sor_pass_k;      /* SOR iterations per pass over HBM of the persistent sweep: 1 = one, 2 = two (temporal
                            blocking: half the DRAM traffic, same bits), 0 = chosen by the cost model (two from
                            ~5000^2 up with power-of-two spacing); environment NSB_SOR_K overrides */
    int spin_timeout_ms; /* deadline of every in-kernel wait (flags between tasks / slabs); 0 = 2000 ms.  Raise it
                            under compute-sanitizer or ncu replay, where a kernel runs orders of magnitude slower */
    int reserved[6];
} nsb_options;

/* one record per time step: what the reference's loop computes on the host each step */
typedef struct nsb_step_info {
    double t, dt, gamma, u_max, v_max, norm_p, res;
    int sor_ret;      /* it+1 when converged, -1 otherwise (main.cu:656-726 contract) */
    int sor_iters;    /* iterations performed */
} nsb_step_info;

typedef struct nsb_ctx nsb_ctx;

/* ------------------------------------------------------------------ A. context API */

int nsb_abi_version(void);
const char* nsb_last_error(const nsb_ctx* ctx);     /* ctx may be NULL: last creation error */
void nsb_default_options(nsb_options* o);

/* io.h:14 init().  Missing file: returns NSB_ERR_IO (the reference perror()s and exits, io.c:18-22;
 * driver/main.c does that on this code). */
int nsb_read_params(const char* path, nsb_params* out);

/* memory.h:27 allocate_memory(): all grids zero-initialised in HBM.  opt may be NULL. */
int nsb_create(const nsb_params* params, const nsb_options* opt, nsb_ctx** out);
/* memory.h:43 free_memory() */
void nsb_destroy(nsb_ctx* ctx);
int nsb_get_params(const nsb_ctx* ctx, nsb_params* out);
/* the stream every kernel of this context is launched on (cudaStream_t as void*) */
void* nsb_stream(const nsb_ctx* ctx);

/* io.h:42 max_mat() for u and for v (signed maxima over the interior, seeded with m[0][0]);
 * replaces max_mat_cuda main.cu:799-832 */
int nsb_max_uv(nsb_ctx* ctx, double* u_max, double* v_max);
/* main.c:91-92 incl. io.h:54 n_min(): host scalar arithmetic, exact operation order */
int nsb_dt_gamma(const nsb_ctx* ctx, double u_max, double v_max, double* dt, double* gamma);
/* boundaries.h:47 / :33 ; replaces set_inflow_cuda / set_noslip_cuda main.cu:928-944 */
int nsb_set_inflow(nsb_ctx* ctx, int side, double u_fix, double v_fix);
int nsb_set_noslip(nsb_ctx* ctx, int side);
/* main.c:95-108: LEFT, RIGHT, BOTTOM no-slip then lid (problem 1: u=1; problem 2: u=sin(f t)) */
int nsb_apply_bc(nsb_ctx* ctx, double t);
/* integration.h:136 FG(); replaces calculate_F_kernel/calculate_G_kernel main.cu:352-382 */
int nsb_fg(nsb_ctx* ctx, double dt, double gamma);
/* main.c:116-120; replaces calculate_RHS_kernel main.cu:219-229 */
int nsb_rhs(nsb_ctx* ctx, double dt);
/* F, G and RHS in one call (nsb_fg then nsb_rhs) */
int nsb_fg_rhs(nsb_ctx* ctx, double dt, double gamma);
/* integration.h:165 SOR() in the order of SOR_UVA_with_shared_memory main.cu:656-726.
 * Returns it+1 (>0) when converged, NSB_NOT_CONVERGED after max_it, other negatives on error.
 * iters / res / norm_p may be NULL. */
int nsb_sor(nsb_ctx* ctx, int* iters, double* res, double* norm_p);
/* exactly n red-black iterations, no convergence exit (per-iteration parity + benchmarking);
 * res = residual norm after the last one */
int nsb_sor_iterations(nsb_ctx* ctx, int n, double* res);
/* main.c:131-136; replaces update_velocities_kernel main.cu:231-245 (also refreshes the cached
 * max u, max v and sum p^2 that the next step needs) */
int nsb_update_uv(nsb_ctx* ctx, double dt);
/* whole loop body main.c:88-145 at time t.  info may be NULL. */
int nsb_step(nsb_ctx* ctx, double t, nsb_step_info* info);
/* whole loop main.c:86-147: `while (t < T)`; max_steps <= 0 = no cap.  log may be NULL. */
int nsb_run(nsb_ctx* ctx, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end);
/* main.c:148-149: u[i_max/2][j_max/2], v[i_max/2][j_max/2] */
int nsb_get_center(nsb_ctx* ctx, double* u_c, double* v_c);
/* seconds spent inside nsb_sor so far (CUDA events); what main.c:153 prints */
double nsb_sor_seconds(const nsb_ctx* ctx);
/* kernels this library has enqueued on the context so far (bench.py's gpu_launches) */
long long nsb_launch_count(const nsb_ctx* ctx);
/* what the planner chose for the persistent SOR sweep: iterations per pass (1 / 2), rows per chunk, strips, chunks */
int nsb_sor_plan_info(const nsb_ctx* ctx, int* pass_k, int* chunk_rows, int* n_strips, int* n_chunks);

/* host <-> HBM.  host holds (i_max+2)*(j_max+2) doubles in the given layout. */
int nsb_upload(nsb_ctx* ctx, int field, const double* host, int layout);
int nsb_download(nsb_ctx* ctx, int field, double* host, int layout);
/* io.h:31 output(): writes <prefix>_u.txt, _v.txt, _p.txt in the reference's text format (io.c:61-120), byte for
 * byte.  nsb_output returns when the files are closed.  nsb_output_async snapshots the fields on the device and
 * returns at once: the copy to the host runs on a side stream and the text is produced by host threads while the
 * time loop goes on (the n_print cadence of main.c:138-143 without stalling the solver); nsb_output_wait (also
 * implied by the next dump and by nsb_destroy) returns the status of the dump in flight. */
int nsb_output(nsb_ctx* ctx, double t, const char* prefix);
int nsb_output_async(nsb_ctx* ctx, double t, const char* prefix);
int nsb_output_wait(nsb_ctx* ctx);

/* device pointers for zero-copy interop (e.g. torch): base of the padded grid, its pitch in
 * doubles, the column offset of i = 0 and the row offset of j = 0 */
int nsb_device_view(nsb_ctx* ctx, int field, void** dptr, long long* pitch, int* col0, int* row0);

/* ------------------------------------------------ A2. y-slab decomposition over the GPUs of one box
 * New work (the reference is single-GPU): the domain is cut along y into `world` slabs of
 * whole SOR chunks.  Halo rows of p travel inside the SOR sweep kernel as peer stores over NVLink,
 * the residual / dt reductions are combined on the device in an order that does not depend on the
 * decomposition, so 1, 2, 4 and 8 slabs give the same bits (DESIGN.md "multi-GPU").
 *
 * Two ways to run it:
 *   - one process per GPU (bench.py under torchrun): nsb_create_slab on every rank, exchange the
 *     descriptors (nsb_slab_export -> any all-gather -> nsb_slab_connect), then call nsb_step /
 *     nsb_run / nsb_sor on every rank as if it were alone;
 *   - one process, several GPUs (driver/main.c with NSB_GPUS=n): the nsb_group_* calls below.  Slabs
 *     may share a device: they then run in lock step on one stream (how a 1-GPU box tests this). */
int nsb_slab_partition(int j_max, int chunk_rows, int world, int rank, int* j_lo, int* nj,
                       int* band0, int* nbands_global);            /* host arithmetic only */
int nsb_create_slab(const nsb_params* params, const nsb_options* opt, int rank, int world, nsb_ctx** out);
int nsb_slab_desc_bytes(void);
int nsb_slab_export(const nsb_ctx* ctx, void* desc_out);
int nsb_slab_connect(nsb_ctx* ctx, const void* descs_in_rank_order);
int nsb_slab_info(const nsb_ctx* ctx, int* rank, int* world, int* j_lo, int* nj_local, int* chunk_rows);
int nsb_slab_error(nsb_ctx* ctx);       /* 1 if a wait on a neighbour timed out */

typedef struct nsb_group nsb_group;
int nsb_group_create(const nsb_params* params, const nsb_options* opt, int n, const int* devices, nsb_group** out);
void nsb_group_destroy(nsb_group* g);
const char* nsb_group_last_error(const nsb_group* g);
int nsb_group_size(const nsb_group* g);
nsb_ctx* nsb_group_ctx(nsb_group* g, int k);
int nsb_group_step(nsb_group* g, double t, nsb_step_info* info);
int nsb_group_run(nsb_group* g, int max_steps, nsb_step_info* log, int max_log, int* steps, double* t_end);
int nsb_group_sor(nsb_group* g, int* iters, double* res, double* norm_p);
int nsb_group_sor_iterations(nsb_group* g, int n, double* res);
double nsb_group_sor_seconds(const nsb_group* g);
int nsb_group_upload(nsb_group* g, int field, const double* host, int layout);     /* whole domain */
int nsb_group_download(nsb_group* g, int field, double* host, int layout);
int nsb_group_get_center(nsb_group* g, double* u_c, double* v_c);
/* output() for the whole domain: every slab contributes its rows to the same three files */
int nsb_group_output(nsb_group* g, double t, const char* prefix);
int nsb_group_output_async(nsb_group* g, double t, const char* prefix);
int nsb_group_output_wait(nsb_group* g);

/* ---------------------------------------------------------------- B. host-grid API
 * Same argument lists as the reference functions.  Grids are the reference's jagged `double**`
 * (memory.c:3-26): u has rows 0..i_max, v rows have entries 0..j_max, the rest are
 * (i_max+2) x (j_max+2).  Device selection: environment NSB_DEVICE (default 0). */

/* integration.h:165.  Returns 0 / -1 like the serial SOR; sweep order is the CUDA reference's
 * red-black order (main.cu:684-719).  `res` is accepted and left untouched. */
int nsb_SOR(double** p, int i_max, int j_max, double delta_x, double delta_y, double** res,
            double** RHS, double omega, double eps, int max_it);
/* same, also reporting what SOR_UVA_with_shared_memory returns (it+1 / -1) */
int nsb_SOR_iters(double** p, int i_max, int j_max, double delta_x, double delta_y, double** res,
                  double** RHS, double omega, double eps, int max_it, int* iters, double* last_res);
/* integration.h:136 */
int nsb_FG(double** F, double** G, double** u, double** v, int i_max, int j_max, double Re,
           double g_x, double g_y, double delta_t, double delta_x, double delta_y, double gamma);
/* io.h:42 -- for u (i_max+1 rows of j_max+2) and for v (i_max+2 rows of j_max+1, memory.c:15) alike: reads rows
 * 0..i_max, columns 0..j_max only (what the reference scans, io.c:122-139, plus m[0][0]) */
double nsb_max_mat(int i_max, int j_max, double** m);
/* one whole loop body (main.c:88-145) on host grids: uploads u,v,p, runs the step in HBM,
 * downloads u,v,p,F,G,RHS.  The end-to-end call bench.py times. */
int nsb_timestep(const nsb_params* params, double t, double** u, double** v, double** p,
                 double** F, double** G, double** RHS, nsb_step_info* info);
/* release the cached context the host-grid API keeps between calls */
void nsb_host_api_reset(void);

/* ---------------------------------------------------------------- C. allocation-free entry point
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * max_mat (io.h:42 semantics: the scan starts from m[0][0] and takes strictly larger values of rows 1..i_max, columns
 * 1..j_max) of a grid that is ALREADY on the current device: row i starts at d_m + i * row_stride doubles
 * (row_stride >= j_max + 1; the reference's device grids have row_stride = j_max + 2, main.cu:103-120).  The result is
 * written to the device double *d_max, asynchronously on `stream` (a cudaStream_t; NULL = the default stream).
 * Allocates nothing: the per-row partial maxima live in a static device array of 128.25 KiB that every GPU loading
 * the library holds until exit.  Limits: i_max <= 16384 and j_max <= 16384, beyond that NSB_ERR_ARG with the text in
 * nsb_last_error(NULL).  All calls on one device share that array: calls that overlap on the GPU (two streams) race;
 * calls ordered on one stream do not.  The first call on a device looks the array's address up (make it outside
 * stream capture). */
int nsb_max_mat_noalloc(const double* d_m, long long row_stride, int i_max, int j_max, double* d_max, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NSB200_H */
