// context.cu -- the C ABI of include/nsb200.h, group A (context API): allocation, the
// time-step phases in the reference's order, the SOR drive loop, transfers.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <unistd.h>

#include "nsb_internal.h"

static char g_create_err[512] = "";

int nsb_fail(nsb_ctx* c, int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c ? c->err : g_create_err, 512, fmt, ap);
    va_end(ap);
    return code;
}
#define fail nsb_fail

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
    nsb_out_release(c);                                        // joins a dump still being written (output.cu)
    for (int r = 0; r < c->world && r < NSB_MAX_RANKS; r++)
        if (r != c->rank && c->peer[r].mapped_ipc && c->peer[r].base) cudaIpcCloseMemHandle(c->peer[r].base);
    if (c->arena) cudaFree(c->arena);
    double* bufs[] = {c->d_partials, c->d_red_partials, c->d_bandvals, c->d_partials2, c->d_bandvals2};
    if (c->d_claim) cudaFree(c->d_claim);
    if (c->d_done) cudaFree(c->d_done);
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

// Rows of rank `rank` out of `world`: whole bands (SOR chunks of `chunk_rows` rows), the first
// (nbands % world) ranks get one band more.  Pure host arithmetic (also used by the CPU tests).
extern "C" int nsb_slab_partition(int j_max, int chunk_rows, int world, int rank, int* j_lo, int* nj,
                                  int* band0, int* nbands_global) {
    if (j_max < 1 || chunk_rows < 2 || (chunk_rows & 1) || world < 1 || rank < 0 || rank >= world) return NSB_ERR_ARG;
    const int nb = (j_max + chunk_rows - 1) / chunk_rows;
    if (nb < world) return NSB_ERR_ARG;
    // Whole bands per rank.  The FULL bands are dealt out evenly (the first `rem` ranks hold one more); a ragged last
    // band goes to the last rank on top of its share when that rank is not one of those -- 8192 x 2 rows in 124-row
    // bands: 66 + (66 + 16 rows) instead of 67 + (65 + 16 rows).
    const int nfull = j_max / chunk_rows, ragged = nb - nfull;
    int base = nb / world, rem = nb % world;
    if (ragged && nfull >= world && (nfull % world) != 0) { base = nfull / world; rem = nfull % world; }
    else if (ragged && nfull >= world) { base = nfull / world; rem = 0; }
    const bool tail_extra = ragged && nfull >= world;           // the ragged band is an extra of the last rank
    const int b0 = rank * base + (rank < rem ? rank : rem);
    const int nbr = base + (rank < rem ? 1 : 0) + ((tail_extra && rank == world - 1) ? 1 : 0);
    const int lo = b0 * chunk_rows;
    int hi = (b0 + nbr) * chunk_rows;
    if (hi > j_max) hi = j_max;
    if (j_lo) *j_lo = lo;
    if (nj) *nj = hi - lo;
    if (band0) *band0 = b0;
    if (nbands_global) *nbands_global = nb;
    return NSB_OK;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int nsb_create_slab_internal(const nsb_params* P, const nsb_options* opt, int rank, int world, nsb_ctx** out) {
    if (!P || !out) return fail(NULL, NSB_ERR_ARG, "nsb_create: null argument");
    *out = NULL;
    if (P->i_max < 1 || P->j_max < 1) return fail(NULL, NSB_ERR_ARG, "i_max and j_max must be >= 1");
    if (world < 1 || world > NSB_MAX_RANKS || rank < 0 || rank >= world)
        return fail(NULL, NSB_ERR_ARG, "bad rank %d / world %d (at most %d slabs)", rank, world, NSB_MAX_RANKS);
    nsb_ctx* c = (nsb_ctx*)calloc(1, sizeof(nsb_ctx));
    if (!c) return fail(NULL, NSB_ERR_ARG, "out of host memory");
    c->P = *P;
    if (opt) c->O = *opt; else nsb_default_options(&c->O);
    c->device = c->O.device;
    c->rank = rank; c->world = world; c->J = P->j_max;
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

    c->dx = P->a / P->i_max;                                   // main.c:72-73 (global sizes)
    c->dy = P->b / P->j_max;
    c->recip_exact = is_pow2(c->dx * c->dx) && is_pow2(c->dy * c->dy);
    c->use_recip = (c->O.div_mode == NSB_DIV_RECIP) || (c->O.div_mode == NSB_DIV_AUTO && c->recip_exact);

    // ---- rows of this slab
    c->g.ni = P->i_max;
    if (world == 1) {
        c->g.nj = P->j_max; c->j_lo = 0;
    } else {
        // the chunk height is part of the decomposition: fixed before the rows are dealt out
        int ch;
        if (c->O.sor_chunk_rows > 0) ch = (c->O.sor_chunk_rows + 1) & ~1;
        else {
            // planned for the slab height (8192-row slabs: 124 rows -- 124 + 11 rows = 17 blocks of the two-iteration pass)
            int k = 0;
            ch = nsb_sor_preplan_slabs(c, P->i_max, P->j_max, world, &k);
            if (c->O.sor_pass_k == 0) c->O.sor_pass_k = k;     // the iterations per pass this height was chosen for
        }
        if (ch < 16) ch = 16;
        c->O.sor_chunk_rows = ch;
        int nj = 0;
        if (nsb_slab_partition(P->j_max, ch, world, rank, &c->j_lo, &nj, &c->band0, &c->nbands_global) != NSB_OK) {
            fail(NULL, NSB_ERR_ARG, "j_max = %d gives fewer than %d bands of %d rows: too many slabs", P->j_max, world, ch);
            nsb_destroy(c);
            return NSB_ERR_ARG;
        }
        c->g.nj = nj;
    }
    if (c->g.nj > 131070) {
        // the F/G and RHS kernels put two rows per block into gridDim.y (<= 65535); larger slabs need more y-slabs
        fail(NULL, NSB_ERR_ARG, "a slab of %d rows exceeds the 131070 rows one GPU context handles: use more slabs (NSB_GPUS / nsb_group_create)", c->g.nj);
        nsb_destroy(c);
        return NSB_ERR_ARG;
    }
    c->wall_bottom = (rank == 0);
    c->wall_top = (rank == world - 1);
    // pitch: COL0 + (ni+2) columns + one spare group, rounded up to 16 doubles (128 B)
    long long need = NSB_COL0 + (long long)P->i_max + 2 + NSB_GROUP;
    c->g.pitch = (need + 15) / 16 * 16;
    {   // tuning hook: extra padding of every row (doubles, rounded to 128 B)
        const char* ex = getenv("NSB_PITCH_EXTRA");
        if (ex && atoi(ex) > 0) c->g.pitch += (atoi(ex) + 15) / 16 * 16;
    }
    c->g.rows = c->g.nj + 2 + NSB_ROW0 + NSB_ROWS_ABOVE;
    c->g.elems = (size_t)c->g.pitch * (size_t)c->g.rows;

    nsb_sor_plan(c);                                           // chunk_rows, n_strips, n_chunks, n_partials
    if (world == 1) { c->band0 = 0; c->nbands_global = c->n_chunks; }

    // ---- one arena: 9 grids, the sync block, the SOR halo flags, the reduction gather buffer
    const size_t gbytes = align_up(c->g.elems * sizeof(double), 256);
    size_t off = 0;
    NsbSlabDesc& d = c->desc;
    memset(&d, 0, sizeof d);
    d.off_u = off; off += gbytes;
    d.off_v = off; off += gbytes;
    d.off_G = off; off += gbytes;
    d.off_p[0] = off; off += gbytes;
    d.off_p[1] = off; off += gbytes;
    d.off_p2 = off; off += gbytes;
    const size_t off_F = off; off += gbytes;
    d.off_rhs = off; off += gbytes;
    const size_t off_scratch = off; off += gbytes;
    d.off_sync = off; off += align_up(sizeof(SlabSync), 256);
    d.off_sorflag = off; off += align_up((size_t)2 * c->n_strips * sizeof(unsigned long long), 256);
    d.off_gband = off; off += align_up((size_t)NSB_RED_DEPTH * NSB_RED_Q * c->nbands_global * sizeof(double), 256);
    c->arena_bytes = off;
    CUC(cudaMalloc((void**)&c->arena, c->arena_bytes));
    CUC(cudaMemsetAsync(c->arena, 0, c->arena_bytes, c->stream));   // calloc semantics, memory.c:3-26
    c->u = (double*)(c->arena + d.off_u); c->v = (double*)(c->arena + d.off_v);
    c->G = (double*)(c->arena + d.off_G); c->F = (double*)(c->arena + off_F);
    c->rhs = (double*)(c->arena + d.off_rhs); c->scratch = (double*)(c->arena + off_scratch);
    c->pbuf[0] = (double*)(c->arena + d.off_p[0]); c->pbuf[1] = (double*)(c->arena + d.off_p[1]);
    c->pbuf[2] = (double*)(c->arena + d.off_p2);
    c->d_sync = (SlabSync*)(c->arena + d.off_sync);
    c->d_sorflag = (unsigned long long*)(c->arena + d.off_sorflag);
    c->d_gband = (double*)(c->arena + d.off_gband);
    c->pcur = 0;
    {   // deadline of the in-kernel waits
        long long ms = c->O.spin_timeout_ms > 0 ? c->O.spin_timeout_ms : 2000;
        const char* env = getenv("NSB_SPIN_TIMEOUT_MS");
        if (env && atoll(env) > 0) ms = atoll(env);
        const unsigned long long ns = (unsigned long long)ms * 1000000ull;
        CUC(cudaMemcpyAsync(&c->d_sync->timeout_ns, &ns, sizeof ns, cudaMemcpyHostToDevice, c->stream));
        CUC(cudaStreamSynchronize(c->stream));
    }
    d.base = (unsigned long long)(uintptr_t)c->arena;
    d.pid = (long long)getpid();
    d.device = c->device; d.rank = rank; d.world = world; d.nj = c->g.nj; d.j_lo = c->j_lo;
    d.n_strips = c->n_strips; d.nbands_global = c->nbands_global; d.band0 = c->band0;
    d.pitch = c->g.pitch; d.arena_bytes = (long long)c->arena_bytes;
    memcpy(d.uuid, &prop.uuid, 16);
    if (world > 1) {
        cudaIpcMemHandle_t h;
        CUC(cudaIpcGetMemHandle(&h, c->arena));
        memcpy(d.ipc, &h, sizeof h);
    }
    // this rank seen as a peer of itself
    SlabPeer& me = c->peer[rank];
    me.base = c->arena; me.u = c->u; me.v = c->v; me.G = c->G; me.rhs = c->rhs; me.p[0] = c->pbuf[0]; me.p[1] = c->pbuf[1]; me.p[2] = c->pbuf[2];
    me.sync = c->d_sync; me.sorflag = c->d_sorflag; me.gband = c->d_gband; me.nj = c->g.nj; me.mapped_ipc = false;
    c->connected = (world == 1);

    int np = c->n_partials;
    if (np < 8 * c->g.nj + 64) np = 8 * c->g.nj + 64;
    CUC(cudaMalloc((void**)&c->d_partials, (size_t)np * sizeof(double)));
    CUC(cudaMalloc((void**)&c->d_partials2, (size_t)2 * NSB_SOR_KMAX * c->n_partials * sizeof(double)));
    CUC(cudaMalloc((void**)&c->d_bandvals2, (size_t)2 * NSB_SOR_KMAX * (c->n_chunks + 1) * sizeof(double)));
    CUC(cudaMalloc((void**)&c->d_claim, (8 + 2 * (size_t)c->n_chunks) * sizeof(unsigned long long)));   // [0] claim, [1..4] pass tickets, [8..] band tickets by pass parity
    CUC(cudaMalloc((void**)&c->d_done, (size_t)c->n_partials * sizeof(unsigned long long)));
    CUC(cudaMemsetAsync(c->d_partials2, 0, (size_t)2 * NSB_SOR_KMAX * c->n_partials * sizeof(double), c->stream));
    CUC(cudaMemsetAsync(c->d_claim, 0, (8 + 2 * (size_t)c->n_chunks) * sizeof(unsigned long long), c->stream));
    CUC(cudaMemsetAsync(c->d_done, 0, (size_t)c->n_partials * sizeof(unsigned long long), c->stream));
    c->n_red_partials = 3 * 4 * c->g.nj + 64;
    CUC(cudaMalloc((void**)&c->d_red_partials, (size_t)c->n_red_partials * sizeof(double)));
    CUC(cudaMalloc((void**)&c->d_bandvals, (size_t)NSB_RED_Q * (c->n_chunks + 1) * sizeof(double)));
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

extern "C" int nsb_create(const nsb_params* P, const nsb_options* opt, nsb_ctx** out) {
    return nsb_create_slab_internal(P, opt, 0, 1, out);
}

extern "C" int nsb_get_params(const nsb_ctx* c, nsb_params* out) {
    if (!c || !out) return NSB_ERR_ARG;
    *out = c->P;
    return NSB_OK;
}

extern "C" void* nsb_stream(const nsb_ctx* c) { return c ? (void*)c->stream : NULL; }

extern "C" double nsb_sor_seconds(const nsb_ctx* c) { return c ? c->sor_seconds : 0.0; }

extern "C" long long nsb_launch_count(const nsb_ctx* c) { return c ? c->n_launches : 0; }

// What nsb_sor_plan decided for the persistent sweep of this context.
extern "C" int nsb_sor_plan_info(const nsb_ctx* c, int* pass_k, int* chunk_rows, int* n_strips, int* n_chunks) {
    if (!c) return NSB_ERR_ARG;
    if (pass_k) *pass_k = nsb_sor_pass_k(c);
    if (chunk_rows) *chunk_rows = c->chunk_rows;
    if (n_strips) *n_strips = c->n_strips;
    if (n_chunks) *n_chunks = c->n_chunks;
    return NSB_OK;
}

// u and v at (i, local row l)
int nsb_get_value(nsb_ctx* c, int i, int l, double* u_c, double* v_c) {
    CU(c, cudaSetDevice(c->device));
    const size_t o = nsb_index(c->g, i, l);
    double uv[2];
    CU(c, cudaMemcpyAsync(&uv[0], c->u + o, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(&uv[1], c->v + o, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (u_c) *u_c = uv[0];
    if (v_c) *v_c = uv[1];
    return NSB_OK;
}

// main.c:148-149.  On a slab that does not hold the global row j_max/2 this is NSB_ERR_ARG
// (nsb_group_get_center / the python host pick the right rank).
extern "C" int nsb_get_center(nsb_ctx* c, double* u_c, double* v_c) {
    if (!c) return NSB_ERR_ARG;
    const int l = c->J / 2 - c->j_lo;
    if (!((l >= 1 && l <= c->g.nj) || (l == 0 && c->wall_bottom)))
        return fail(c, NSB_ERR_ARG, "global row %d is not on slab %d", c->J / 2, c->rank);
    return nsb_get_value(c, c->P.i_max / 2, l, u_c, v_c);
}

// `nrows` rows starting at local row l0 between the slab and a dense x-fastest host block
// ((i_max+2) doubles per row).  up != 0: host -> HBM.
int nsb_xfer_rows(nsb_ctx* c, int field, double* host, int l0, int nrows, int up) {
    CU(c, cudaSetDevice(c->device));
    double* dev = NULL;
    switch (field) {
        case NSB_U: dev = c->u; break;
        case NSB_V: dev = c->v; break;
        case NSB_P: dev = c->pbuf[c->pcur]; break;
        case NSB_F: dev = c->F; break;
        case NSB_G: dev = c->G; break;
        case NSB_RHS: dev = c->rhs; break;
    }
    if (!dev) return fail(c, NSB_ERR_ARG, "unknown field %d", field);
    if (nrows <= 0) return NSB_OK;
    const size_t nx = (size_t)c->g.ni + 2;
    double* d = dev + (long long)(l0 + NSB_ROW0) * c->g.pitch + NSB_COL0;
    if (up)
        CU(c, cudaMemcpy2DAsync(d, c->g.pitch * sizeof(double), host, nx * sizeof(double), nx * sizeof(double),
                                (size_t)nrows, cudaMemcpyHostToDevice, c->stream));
    else
        CU(c, cudaMemcpy2DAsync(host, nx * sizeof(double), d, c->g.pitch * sizeof(double), nx * sizeof(double),
                                (size_t)nrows, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
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
        // ONE contiguous transfer over PCIe into the scratch grid, then a device-side 2-D copy into the pitched field:
        // a 2-D host copy is 8194 transfers of 64 KB, which the DMA engines of 4+ ranks copying at once serve at half
        // the rate of a single large one (profiles/r2s_h2d_probe.txt against the r2r e2e leg)
        CU(c, cudaMemcpyAsync(c->scratch, host, nx * ny * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpy2DAsync(dst + NSB_ROW0 * c->g.pitch + NSB_COL0, c->g.pitch * sizeof(double), c->scratch,
                                nx * sizeof(double), nx * sizeof(double), ny, cudaMemcpyDeviceToDevice, c->stream));
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
        CU(c, cudaMemcpy2DAsync(c->scratch, nx * sizeof(double), src + NSB_ROW0 * c->g.pitch + NSB_COL0,
                                c->g.pitch * sizeof(double), nx * sizeof(double), ny, cudaMemcpyDeviceToDevice, c->stream));
        CU(c, cudaMemcpyAsync(host, c->scratch, nx * ny * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    } else if (layout == NSB_LAYOUT_REF) {
        CU(c, nsb_launch_from_xfast(c, src, c->scratch));
        CU(c, cudaMemcpyAsync(host, c->scratch, nx * ny * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    } else {
        return fail(c, NSB_ERR_ARG, "unknown layout %d", layout);
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return NSB_OK;
}

// Developer aid: read the persistent sweep's counters on a side stream while a kernel is running.
// out[0..7] = claim counter + tickets, out[8] = ctl.iters, out[9] = ctl.converged, out[10] = wait-error flag,
// out[11] = smallest done flag, out[12] = largest done flag.
extern "C" int nsb_debug_peek(nsb_ctx* c, unsigned long long* out) {
    if (!c || !out) return NSB_ERR_ARG;
    cudaStream_t st;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) return NSB_ERR_CUDA;
    SorCtl k; int err = 0;
    unsigned long long* done = (unsigned long long*)malloc((size_t)c->n_partials * 8);
    cudaMemcpyAsync(out, c->d_claim, 64, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&k, c->d_ctl, sizeof k, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&err, &c->d_sync->error, 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(done, c->d_done, (size_t)c->n_partials * 8, cudaMemcpyDeviceToHost, st);
    cudaError_t e = cudaStreamSynchronize(st);
    cudaStreamDestroy(st);
    out[8] = (unsigned long long)k.iters; out[9] = (unsigned long long)k.converged; out[10] = (unsigned long long)err;
    unsigned long long lo = ~0ull, hi = 0;
    for (int i = 0; i < c->n_partials; i++) { if (done[i] < lo) lo = done[i]; if (done[i] > hi) hi = done[i]; }
    out[11] = lo; out[12] = hi; out[13] = c->flag_base; out[14] = (unsigned long long)c->n_partials;
    int nlo = 0; for (int i = 0; i < c->n_partials; i++) if (done[i] == lo) nlo++;
    out[15] = (unsigned long long)nlo;
    free(done);
    return e == cudaSuccess ? NSB_OK : NSB_ERR_CUDA;
}
