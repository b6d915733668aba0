// nsb_internal.h -- context, HBM layout and launch prototypes shared by the .cu files.
//
// HBM layout (DESIGN.md "data layout"): every grid is one pitched 1-D array, x fastest.
//   element (i, j)  ->  base[(j + ROW0) * pitch + (i + COL0)]
// COL0 = 3 puts interior column i = 1 on a 32-byte boundary (4 doubles = one "column group", the
// unit a thread of the streaming kernels owns); pitch is a multiple of 16 doubles (128 B) and
// leaves >= 4 padding doubles to the right of the ghost column i_max+1, so every group load is in
// bounds.  ROW0 = 4: four pad rows lie below the ghost row j = 0 and NSB_ROWS_ABOVE = 5 above the
// ghost row j_max+1 (zero-filled, never written), so the streaming SOR kernel, whose row window
// reaches 5 rows beyond a chunk (two iterations per pass, sor_stream.cu) and which fetches rows in (even, odd)
// pairs, never needs a bounds test on a load; in a y-slab decomposition the same rows hold the neighbour's halo.  This replaces the reference's row-pointer tables (`double**`, main.cu:103-120)
// and its j-contiguous order, which made every warp access strided.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/nsb200.h"

#define NSB_COL0 3
#define NSB_ROW0 4
#define NSB_ROWS_ABOVE 5              // pad rows above the ghost row j_max+1
#define NSB_SOR_KMAX 2                // SOR iterations per pass over HBM of the persistent sweep (temporal blocking)
#define NSB_HALO_ROWS 5               // p rows a slab keeps of each neighbour: 2 K + 1 for K = NSB_SOR_KMAX
#define NSB_GROUP 4                 // doubles per column group (32 B)
#define NSB_STRIP_GROUPS 30         // output groups per warp strip (lanes 1..30; 0 and 31 are halo)
#define NSB_STRIP_COLS (NSB_STRIP_GROUPS * NSB_GROUP)

struct NsbGrid {
    int ni, nj;            // global interior size (i_max, j_max)
    long long pitch;       // doubles per row
    int rows;              // stored rows (nj + 2 + NSB_ROW0 + NSB_ROWS_ABOVE)
    size_t elems;          // pitch * rows
};

// Scalars every SOR kernel needs; built on the host with the reference's operation order.
struct SorCoef {
    double om1;      // 1 - omega
    double w;        // omega / (2 (1/dx2 + 1/dy2))          main.cu:493,502
    double dx2, dy2; // delta_x^2, delta_y^2
    double rdx2, rdy2;
    float qlo_x, qlo_y;  // IEEE-division path: lower edge of the quotient window for dx2 / dy2 (nsb_math.cuh)
    long long cells; // i_max * j_max
};

// Device-side control block of one SOR solve (lives in HBM; mirrored to pinned host memory).
struct SorCtl {
    double thresh;       // eps * (norm_p + 1.5)                main.cu:713
    double norm_p;
    double res;          // residual norm after the last finished iteration
    double sum_r2;
    int iters;           // iterations finished                 (iters, converged: one aligned 64-bit word,
    int converged;       // set by the iteration that met the threshold   read with one load by the persistent sweep)
    unsigned int ticket; // last-block election
    int pad;
    SorCtl* mirror;      // zero-copy twin in pinned host memory (NULL in the twin itself)
};

static_assert(offsetof(SorCtl, iters) % 8 == 0 && offsetof(SorCtl, converged) == offsetof(SorCtl, iters) + 4, "SorCtl: {iters, converged} must form one 64-bit word");

#ifdef __CUDACC__
// Publish the control block to its pinned-host twin (one thread).  The host polls `converged` /
// `iters` there, so the stream carries nothing but kernels between SOR iterations (a D2H copy per
// iteration cost ~12 % of the iteration time at 8192^2, profiles/r1c).
__device__ __forceinline__ void sor_publish(SorCtl* ctl) {
    SorCtl* m = ctl->mirror;
    if (!m) return;
    m->thresh = ctl->thresh; m->norm_p = ctl->norm_p; m->res = ctl->res; m->sum_r2 = ctl->sum_r2;
    *((volatile int*)&m->converged) = ctl->converged;
    __threadfence_system();
    *((volatile int*)&m->iters) = ctl->iters;
}
#endif

// Per-step reductions produced by the velocity update (and by the stand-alone reduce kernel).
struct StepRed {
    double u_max, v_max, sum_p2;
    unsigned int ticket;
    int pad;
};

// ------------------------------------------------------------------ y-slab decomposition
// Rank r of `world` owns the global rows j_lo+1 .. j_lo+nj (whole SOR chunks, so every reduction
// band lies inside one rank).  Local row l is global row j_lo + l.  Local rows 0 .. -4 and
// nj+1 .. nj+5 hold the neighbour's rows (halo) where the slab ends inside the domain, and the
// ghost / pad rows where it ends at a wall.  Everything a peer writes lives in one arena per rank
// (one IPC handle); peers address it through the offsets of NsbSlabDesc.
#define NSB_MAX_RANKS 16
#define NSB_RED_Q 3                  // quantities per reduction round (u max, v max, sum) / (sum r^2)
#define NSB_RED_DEPTH 4              // gather buffers (round & 3): a rank may run two rounds ahead of a slow reader

struct SlabSync {                    // in the arena; written by peers with system-scope stores
    unsigned long long halo_epoch[2];              // [0] by the lower neighbour, [1] by the upper
    unsigned long long red_arrived[NSB_MAX_RANKS]; // reduction rounds rank r has contributed to
    int error;                                     // a wait timed out (set locally)
    int pad;
    unsigned long long timeout_ns;                 // deadline of every in-kernel wait; MUST follow `error` (slab.cuh: spin_timeout_ns)
};
static_assert(offsetof(SlabSync, timeout_ns) == offsetof(SlabSync, error) + 8, "SlabSync: timeout_ns must sit 8 bytes behind error");

struct NsbSlabDesc {                 // what a rank tells its peers (POD, travels through IPC / gloo / NCCL)
    unsigned char ipc[64];           // cudaIpcMemHandle_t of the arena
    unsigned long long base;         // arena device pointer in the owner's address space
    long long pid;                   // owner process
    int device, rank, world, nj, j_lo, n_strips, nbands_global, band0;
    long long pitch;
    long long off_u, off_v, off_G, off_rhs, off_p[2], off_sync, off_sorflag, off_gband;   // byte offsets in the arena
    long long arena_bytes;
    unsigned char uuid[16];          // cudaDeviceProp::uuid of the owner's GPU
    long long off_p2;                // third pressure buffer (rotation of the two-iterations-per-pass sweep)
    unsigned long long reserved[1];
};

struct SlabPeer {                    // a peer as seen from this rank (device pointers valid HERE)
    unsigned char* base;             // mapped arena
    double *u, *v, *G, *rhs, *p[3];
    SlabSync* sync;
    unsigned long long* sorflag;     // [2][n_strips]: [0] written by ITS lower neighbour, [1] by its upper
    double* gband;                   // [NSB_RED_DEPTH][NSB_RED_Q][nbands_global]
    int nj;
    bool mapped_ipc;
};

// By-value kernel argument for the collective reductions.
struct SlabRed {
    int rank, world, band0, nb_local, nb_global;
    SlabSync* sync[NSB_MAX_RANKS];
    double* gband[NSB_MAX_RANKS];
};

struct nsb_ctx {
    nsb_params P;
    nsb_options O;
    NsbGrid g;
    int device;
    int sm_count;
    cudaStream_t stream;
    bool own_stream;
    double dx, dy;
    bool recip_exact;      // dx2 and dy2 are powers of two -> reciprocal multiply is exact
    bool rcp_valid;        // rcp_val = nsb_rcp_refined(rcp_key) (IEEE division path, nsb_make_sor_coef)
    double rcp_key[2], rcp_val[2];
    bool use_recip;        // resolved division mode
    // grids
    double* u; double* v; double* F; double* G; double* rhs;
    double* pbuf[3];       // [2]: third buffer of the persistent sweep's rotation; between solves the pressure is in [0] or [1]
    int pcur;              // pbuf[pcur] holds the current pressure (0 or 1)
    double* scratch;       // one grid of staging space (layout conversion)
    // slab decomposition (world == 1: the whole domain)
    int rank, world;
    int J;                 // global j_max (P.j_max); g.nj is the LOCAL row count
    int j_lo;              // global row of local row l is j_lo + l
    bool wall_bottom, wall_top;
    int band0, nbands_global;
    unsigned char* arena; size_t arena_bytes;
    NsbSlabDesc desc;
    SlabPeer peer[NSB_MAX_RANKS];      // peer[rank] is this rank itself
    bool connected;
    bool peers_distinct;               // every rank of the decomposition sits on a different physical GPU
    SlabSync* d_sync; unsigned long long* d_sorflag; double* d_gband;
    unsigned long long seq;            // SOR halo epochs handed out so far (halo flag values)
    unsigned long long seq_tail;       // epoch the last pass of the previous persistent launch published
    unsigned long long red_round;      // collective reduction rounds completed so far
    unsigned long long epoch[2];       // halo exchanges pushed to [0] the lower, [1] the upper neighbour
    unsigned long long epoch_in[2];    // halo exchanges expected from [0] the lower, [1] the upper neighbour
    // SOR machinery
    SorCtl* d_ctl; SorCtl* h_ctl;       // device / pinned host
    double* d_partials; int n_partials;
    // persistent multi-iteration sweep (sor_stream.cu): claim counter + 4 iteration tickets, one
    // "done" flag per task, double-buffered partials, epoch base of the flags
    unsigned long long* d_claim; unsigned long long* d_done; double* d_partials2; double* d_bandvals2;
    int st_k, st_passes, st_tail, st_buf0;   // what the last persistent solve enqueued (nsb_ph_sor_end settles the buffers)
    unsigned long long flag_base;
    int chunk_rows, n_strips, n_chunks;
    int sor_k;             // iterations per pass of the persistent sweep as planned (nsb_sor_plan)
    // step reductions
    StepRed* d_red; StepRed* h_red;
    double* d_red_partials; int n_red_partials;
    double* d_bandvals;    // [NSB_RED_Q][n_chunks] band values staged for the collective reduction
    bool red_valid;        // h_red holds max u, max v, sum p^2 of the current fields
    double* h_stage; size_t h_stage_elems;   // pinned staging for uploads/downloads
    cudaEvent_t ev0, ev1, ev_chk[2];
    double sor_seconds;
    long long n_launches;  // kernels of this library enqueued so far (nsb_launch_count)
    int last_sor_engine;
    void* out_state; void* out_buffers;   // output.cu: the dump in flight (first context of a decomposition) / staging buffers
    int resident_state;    // sor_resident.cu: 0 unknown, 1 usable, -1 this device refused the cluster / cooperative launch
    char err[512];
};

// every launcher ends with this: counts the kernels it enqueued and returns the launch status
#define NSB_LAUNCHED(c, n) ((c)->n_launches += (n), cudaGetLastError())

static inline size_t nsb_index(const NsbGrid& g, int i, int j) {
    return (size_t)(j + NSB_ROW0) * (size_t)g.pitch + (size_t)(i + NSB_COL0);
}

// ---- launchers (each enqueues on ctx->stream and returns the launch status) -----------------
cudaError_t nsb_launch_set_inflow(nsb_ctx* c, int side, double u_fix, double v_fix);
cudaError_t nsb_launch_fg(nsb_ctx* c, double dt, double gamma);
cudaError_t nsb_launch_rhs(nsb_ctx* c, double dt);
cudaError_t nsb_launch_fg_rhs(nsb_ctx* c, double dt, double gamma);
cudaError_t nsb_launch_update_uv(nsb_ctx* c, double dt);
cudaError_t nsb_launch_reduce_uvp(nsb_ctx* c);     // max u, max v, sum p^2 without updating
cudaError_t nsb_launch_sor_begin(nsb_ctx* c);      // threshold from sum p^2, clears the flags
cudaError_t nsb_launch_sor_stream_iter(nsb_ctx* c, int it, const double* pin, double* pout, int check);
int nsb_sor_persist_supported(const nsb_ctx* c);
int nsb_sor_pass_k(const nsb_ctx* c);                                    // iterations per pass of the persistent sweep (1 or 2)
cudaError_t nsb_launch_sor_persist(nsb_ctx* c, int k, int n_pass, int it0, int check, int buf_in, int noclose, int wait_first);
cudaError_t nsb_launch_sor_lex(nsb_ctx* c, int max_it, int check);       // serial (lexicographic) order, validation
cudaError_t nsb_launch_sor_simple_iter(nsb_ctx* c, int it, double* p, int check);
extern "C" int nsb_plan_chunk_rows(int ni, int nj, long long slots, int recip);   // sor_stream.cu, host arithmetic
extern "C" int nsb_plan_chunk_rows_k(int ni, int nj, long long slots, int recip, int k);
extern "C" int nsb_plan_sor(int ni, int nj, long long slots1, long long slots2, int recip, int* rows);
SorCoef nsb_make_sor_coef(nsb_ctx* c, long long cells);                  // sor_simple.cu; cells = GLOBAL i_max * j_max
cudaError_t nsb_launch_sor_finish_ghosts(nsb_ctx* c, double* pfinal, const double* pprev);
cudaError_t nsb_launch_to_xfast(nsb_ctx* c, const double* src_ref, double* dst);   // [i][j] -> grid
cudaError_t nsb_launch_from_xfast(nsb_ctx* c, const double* src, double* dst_ref); // grid -> [i][j]
int nsb_sor_resident_supported(const nsb_ctx* c);
cudaError_t nsb_launch_sor_resident(nsb_ctx* c, int max_it, int check);

void nsb_sor_plan(nsb_ctx* c);
int nsb_sor_preplan_slabs(const nsb_ctx* c, int ni, int j_max, int world, int* k_out);   // band height of a y-slab decomposition (before nsb_sor_plan)
// output.cu: field dumps in the reference's format for the n slabs cs[0..n-1] of one decomposition (rank order)
int nsb_out_run(nsb_ctx** cs, int n, double t, const char* prefix, int async);
int nsb_out_wait(nsb_ctx* c0);
void nsb_out_release(nsb_ctx* c);

// ---- slab decomposition (slab.cu / step.cu)
SlabRed nsb_slab_red(const nsb_ctx* c);
cudaError_t nsb_launch_sor_close(nsb_ctx* c, int it, int check);            // world > 1: wait + combine + test
cudaError_t nsb_launch_red_close(nsb_ctx* c);                               // world > 1: finish the step reduction
// exchange kinds
enum { NSB_X_UV = 0, NSB_X_G = 1, NSB_X_P = 2, NSB_X_RHS = 3 };
cudaError_t nsb_launch_halo_push(nsb_ctx* c, int kind);
cudaError_t nsb_launch_halo_wait(nsb_ctx* c, int kind);
int nsb_create_slab_internal(const nsb_params* P, const nsb_options* opt, int rank, int world, nsb_ctx** out);
int nsb_fail(nsb_ctx* c, int code, const char* fmt, ...);
int nsb_get_value(nsb_ctx* c, int i, int l, double* u_c, double* v_c);
int nsb_xfer_rows(nsb_ctx* c, int field, double* host, int l0, int nrows, int up);
extern "C" int nsb_slab_error(nsb_ctx* c);
extern "C" int nsb_slab_export(const nsb_ctx* c, void* desc_out);
extern "C" int nsb_slab_connect(nsb_ctx* c, const void* descs);
// phases of one time step; every one is collective over the ranks of a decomposition
int nsb_ph_reduce_launch(nsb_ctx* c);     // max u, max v, sum p^2 when the cache is stale
int nsb_ph_reduce_finish(nsb_ctx* c);
int nsb_ph_bc(nsb_ctx* c, double t);
int nsb_ph_sor_begin(nsb_ctx* c);
int nsb_ph_sor_end(nsb_ctx* c, int iters);
int nsb_ph_update_launch(nsb_ctx* c, double dt);
int nsb_ph_update_finish(nsb_ctx* c);
