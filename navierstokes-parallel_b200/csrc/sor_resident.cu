// sor_resident.cu -- resident SOR engine for grids that fit on chip: p and RHS of the whole grid live
// in shared memory, every CTA owns a band of rows, all iterations of a solve run inside one launch.
//   cluster variant : ONE thread-block cluster (up to 16 SMs); boundary rows and residual partials
//                     travel through distributed shared memory, hardware cluster barrier;
//   grid variant    : cooperative launch, up to one CTA per SM (grids up to ~1024^2); boundary rows
//                     and partials travel through L2, grid barrier.
//
// Why: at 256^2 (tests/1.in, 1 MB working set) the solve is pure latency.  The reference spends six
// launches, six device synchronisations and a D2H copy per iteration (src/parallel/main.cu:684-719,
// ~56 us per iteration in its published numbers, 78 us on a B200); the streaming engine still pays a
// task hand-over through L2 flags per iteration.  Here one iteration costs two barriers plus the
// arithmetic: 6.5 us at 256^2, 8.7 us at 512^2, 13.5 us at 1024^2 (profiles/r1g_*).
//
// Order of operations = src/parallel/main.cu:684-719.  The ghost cells are STORED in shared memory;
// the thread that owns a wall-adjacent cell lets the ghost follow its update at once, which is what
// the reference's refresh before the NEXT colour (main.cu:686 / :694) would copy.  The residual of a
// black cell is evaluated right after its update and BEFORE its ghost follows, so it sees the
// pre-black value exactly like the reference (no refresh between colour 1 and the residual kernel):
//   colour 0 | exchange + barrier | colour 1 + residual of the black cells | exchange + barrier |
//   residual of the red cells | partials out.
// The partial sums of iteration n are read after the first barrier of iteration n+1: its colour 0 has
// already run by then, keeps the old values in a roll-back tile and is undone if iteration n
// converged.  Arithmetic per cell is sor_cell / sor_residual of nsb_math.cuh (reference association
// order, no contraction): p has the same bits as every other engine.  The residual is summed per
// thread, per block (fixed tree) and over the CTAs in rank order: deterministic, but not the order of
// the streaming engine -- only the convergence test sees it.
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nsb_internal.h"
#include "nsb_math.cuh"

namespace cg = cooperative_groups;

#define RES_THREADS 512
#define RES_MAX_CLUSTER 16
#define RES_SMEM_LIMIT (220 * 1024)

struct ResidentArgs {
    double* p;                 // current pressure buffer (updated in place)
    const double* rhs;
    long long pitch;
    int ni, nj;
    int W;                     // shared-memory row width in doubles (even, >= ni + 3): column i at index i + 1
    int rows_per_cta;          // R of every CTA but possibly the last
    int max_it, check, it0;
    SorCoef k;
    SorCtl* ctl;
    // cooperative-grid variant (tiles spread over up to all SMs): boundary rows and block partials
    // travel through HBM/L2 instead of distributed shared memory
    double* ghalo;             // [n_cta][2][W]: [.][0] = first interior row, [.][1] = last interior row of the CTA
    double* gparts;            // [n_cta]
};

// Final ghost values as the reference leaves them (the refresh before the LAST black sweep,
// main.cu:694): kept beside the working ghost cells, which follow every update at once.
struct GhostFinal {
    double* w; double* e;      // [rows_per_cta + 2] per tile row l
    double* s; double* n;      // [W] per x index (first / last CTA only)
};

// MODE 0: update the cells of `color` (backup of the old value into B when B != nullptr);
//         with_res: also accumulate r^2 of the cells just updated (second colour: neighbours final).
// MODE 1: residual only.   MODE 2: roll the cells of `color` back from B.
// Thread layout: tx along column PAIRS (columns 2q+1, 2q+2), ty along rows.
// Ghost cells are owned by the thread of the wall-adjacent cell: it refreshes them right after its
// update, which is what the reference's refresh before the next colour would copy; the black
// cell's residual is taken BEFORE its ghost follows, so it sees the pre-black value (main.cu:694-702).
template <bool RECIP, int MODE>
__device__ __forceinline__ double resident_pass(double* __restrict__ P, const double* __restrict__ Q, double* __restrict__ B,
                                                const GhostFinal& gf, int W, int R, int ni, int nj, int jlo, int color,
                                                bool with_res, const SorCoef& kf, int tx, int ty, int TX, int TY) {
    double acc = 0.0;
    const int npairs = (ni + 1) >> 1;
    // the cells of one pass are independent of each other: the row loop is unrolled so that the loads and
    // the FP64 chains of several cells overlap (the pass is latency-bound: a handful of cells per thread)
    for (int q = tx; q < npairs; q += TX)
#pragma unroll 4
        for (int l = 1 + ty; l <= R; l += TY) {
            const int j = jlo + l - 1;                              // global row
            double* row = P + (size_t)l * W;
            const double* rq = Q + (size_t)(l - 1) * W;
            const int i = 2 * q + 1 + ((j + 1 + color) & 1);        // the cell of this pair with (i + j) % 2 == color
            if (i > ni) continue;
            const int x = i + 1;
            if (MODE == 2) {
                const double v = B[(size_t)l * W + x];
                row[x] = v;
                if (i == 1) { row[1] = v; gf.w[l] = v; }
                if (i == ni) { row[ni + 2] = v; gf.e[l] = v; }
                if (j == 1) { row[x - W] = v; gf.s[x] = v; }
                if (j == nj) { row[x + W] = v; gf.n[x] = v; }
                continue;
            }
            const double pc = row[x], pe = row[x + 1], pw = row[x - 1];
            const double pn = row[x + W], ps = row[x - W];
            const double r = rq[x];
            if (MODE == 0) {
                const double v = sor_cell<RECIP>(pc, pe, pw, pn, ps, r, kf);
                if (B) B[(size_t)l * W + x] = pc;
                row[x] = v;
                if (with_res) {
                    const double d = sor_residual<RECIP>(v, pe, pw, pn, ps, r, kf);
                    acc = fma(d, d, acc);
                }
                // ghosts follow; the value the reference's LAST refresh would have copied is the new one for
                // the first colour and the old one for the second
                const double fin = (color == 0) ? v : pc;
                if (i == 1) { row[1] = v; gf.w[l] = fin; }
                if (i == ni) { row[ni + 2] = v; gf.e[l] = fin; }
                if (j == 1) { row[x - W] = v; gf.s[x] = fin; }
                if (j == nj) { row[x + W] = v; gf.n[x] = fin; }
            } else {
                const double d = sor_residual<RECIP>(pc, pe, pw, pn, ps, r, kf);
                acc = fma(d, d, acc);
            }
        }
    return acc;
}

// Neumann ghosts of this CTA's tile (main.cu:249-273), once at the start of the solve.
__device__ __forceinline__ void resident_ghosts(double* P, const GhostFinal& gf, int W, int R, int ni, bool first, bool last) {
    for (int l = 1 + (int)threadIdx.x; l <= R; l += RES_THREADS) {
        double* row = P + (size_t)l * W;
        row[1] = row[2]; gf.w[l] = row[2];                          // p[0][j] = p[1][j]
        row[ni + 2] = row[ni + 1]; gf.e[l] = row[ni + 1];           // p[i_max+1][j] = p[i_max][j]
    }
    if (first)
        for (int i = 1 + (int)threadIdx.x; i <= ni; i += RES_THREADS) { P[i + 1] = P[W + i + 1]; gf.s[i + 1] = P[W + i + 1]; }
    if (last)
        for (int i = 1 + (int)threadIdx.x; i <= ni; i += RES_THREADS) {
            P[(size_t)(R + 1) * W + i + 1] = P[(size_t)R * W + i + 1]; gf.n[i + 1] = P[(size_t)R * W + i + 1];
        }
}

// Boundary rows to the neighbours after a colour pass: only the cells of that colour changed.
// Cluster: straight into the neighbour's halo row (DSMEM).  Grid: into this CTA's global halo slot.
template <bool GRID>
__device__ __forceinline__ void resident_push(const double* P, double* Pdown, double* Pup, double* hmine, int W, int R,
                                              int R_down, int ni, int jlo, int color, bool has_down, bool has_up, int tid) {
    const int i_dn = 1 + ((jlo + 1 + color) & 1);            // first cell of this colour in my row l = 1 (global row jlo)
    const int i_up = 1 + ((jlo + R + color) & 1);            // ... in my row l = R (global row jlo + R - 1)
    if (has_down)
        for (int i = i_dn + 2 * tid; i <= ni; i += 2 * RES_THREADS) {
            if (GRID) hmine[i + 1] = P[(size_t)W + i + 1];
            else Pdown[(size_t)(R_down + 1) * W + i + 1] = P[(size_t)W + i + 1];
        }
    if (has_up)
        for (int i = i_up + 2 * tid; i <= ni; i += 2 * RES_THREADS) {
            if (GRID) hmine[W + i + 1] = P[(size_t)R * W + i + 1];
            else Pup[i + 1] = P[(size_t)R * W + i + 1];
        }
}

// Grid variant, after the barrier: fetch the neighbours' boundary cells of that colour into my halo rows.
__device__ __forceinline__ void resident_pull(double* P, const double* ghalo, int W, int R, int ni, int jlo, int color,
                                              int rank, bool has_down, bool has_up, int tid) {
    if (has_down) {                                          // row below me = global row jlo - 1 = last row of rank - 1
        const double* h = ghalo + ((size_t)(rank - 1) * 2 + 1) * W;
        for (int i = 1 + ((jlo + color) & 1) + 2 * tid; i <= ni; i += 2 * RES_THREADS) P[i + 1] = __ldcg(h + i + 1);
    }
    if (has_up) {                                            // row above me = global row jlo + R = first row of rank + 1
        const double* h = ghalo + (size_t)(rank + 1) * 2 * W;
        for (int i = 1 + ((jlo + R + 1 + color) & 1) + 2 * tid; i <= ni; i += 2 * RES_THREADS)
            P[(size_t)(R + 1) * W + i + 1] = __ldcg(h + i + 1);
    }
}

// Barrier over all CTAs of the solve: hardware cluster barrier, or the cooperative grid barrier.
template <bool GRID>
__device__ __forceinline__ void resident_sync(int C) {
    if (GRID) cg::this_grid().sync();
    else if (C > 1) cg::this_cluster().sync();
    else __syncthreads();
}

// GRID = false: one thread-block cluster (<= 16 CTAs), halos and partials through distributed shared memory.
// GRID = true : cooperative launch over up to one CTA per SM, halos and partials through global memory
//               (grids up to ~1024^2; per iteration two grid barriers instead of two cluster barriers).
template <bool RECIP, bool GRID>
__global__ void __launch_bounds__(RES_THREADS, 1) sor_resident_kernel(const ResidentArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red_sh[32];
    __shared__ double parts[RES_MAX_CLUSTER];
    __shared__ double tot_sh;
    const int C = GRID ? (int)gridDim.x : (int)cg::this_cluster().num_blocks();
    const int rank = GRID ? (int)blockIdx.x : (int)cg::this_cluster().block_rank();
    const int W = a.W, ni = a.ni, nj = a.nj;
    const int jlo = 1 + rank * a.rows_per_cta;                         // first global row of this CTA
    const int R = min(a.rows_per_cta, nj - jlo + 1);                   // >= 1 by construction (resident_plan)
    const bool first = (rank == 0), last = (jlo + R - 1 == nj);
    const size_t tile = (size_t)(a.rows_per_cta + 2) * W;
    double* P = reinterpret_cast<double*>(smem_raw);                  // rows jlo-1 .. jlo+R, x index = i + 1
    double* B = P + tile;                                             // old values of the first colour (roll-back)
    double* Q = B + tile;                                             // RHS rows jlo .. jlo+R-1
    GhostFinal gf;
    gf.w = Q + (size_t)a.rows_per_cta * W; gf.e = gf.w + (a.rows_per_cta + 2);
    gf.s = gf.e + (a.rows_per_cta + 2); gf.n = gf.s + W;
    const int tid = threadIdx.x;

    // ---- load the tile
    const double* gp = a.p + (long long)NSB_ROW0 * a.pitch + NSB_COL0;
    const double* gq = a.rhs + (long long)NSB_ROW0 * a.pitch + NSB_COL0;
    for (int l = 0; l <= R + 1; l++) {
        const long long ro = (long long)(jlo - 1 + l) * a.pitch;
        for (int i = tid; i <= ni + 1; i += RES_THREADS) P[(size_t)l * W + i + 1] = gp[ro + i];
        if (tid == 0) { P[(size_t)l * W] = 0.0; for (int x = ni + 3; x < W; x++) P[(size_t)l * W + x] = 0.0; }
    }
    for (int l = 1; l <= R; l++) {
        const long long ro = (long long)(jlo - 1 + l) * a.pitch;
        for (int i = tid; i <= ni + 1; i += RES_THREADS) Q[(size_t)(l - 1) * W + i + 1] = gq[ro + i];
    }
    // neighbours' tiles through distributed shared memory (cluster variant)
    double* Pdown = (!GRID && rank > 0) ? cg::this_cluster().map_shared_rank(P, rank - 1) : nullptr;
    double* Pup = (!GRID && rank + 1 < C) ? cg::this_cluster().map_shared_rank(P, rank + 1) : nullptr;
    double* hmine = GRID ? a.ghalo + (size_t)rank * 2 * W : nullptr;
    const int R_down = a.rows_per_cta;                                // every CTA below the last one is full
    const bool has_up = (rank + 1 < C) && (jlo + R <= nj);
    const bool has_down = (rank > 0);
    const int npairs = (ni + 1) >> 1;
    int TX = 32;
    while (TX < npairs && TX < RES_THREADS) TX <<= 1;
    const int TY = RES_THREADS / TX;
    const int tx = tid % TX, ty = tid / TX;
    const double thresh = a.ctl->thresh;
    __syncthreads();
    resident_ghosts(P, gf, W, R, ni, first, last);                                       // main.cu:686 of iteration 0
    __syncthreads();
    resident_sync<GRID>(C);

    // Per iteration two cluster barriers.  The partial sums of iteration n become visible at the first
    // barrier of iteration n+1, i.e. after its first colour has already run: that colour keeps the old
    // values in B and is rolled back if iteration n turns out to have converged.
    int it = 0, converged = 0;
    double res = a.ctl->res, total = a.ctl->sum_r2;
#ifdef RES_TIMING
    long long tk[8];
#define TICK(n) tk[n] = clock64()
#else
#define TICK(n)
#endif
    for (;; it++) {
        const bool real = it < a.max_it;
        TICK(0);
        if (real) {
            resident_pass<RECIP, 0>(P, Q, B, gf, W, R, ni, nj, jlo, 0, false, a.k, tx, ty, TX, TY);
            TICK(1);
            __syncthreads();
            resident_push<GRID>(P, Pdown, Pup, hmine, W, R, R_down, ni, jlo, 0, has_down, has_up, tid);
        }
        TICK(2);
        resident_sync<GRID>(C);
        if (GRID && real) { resident_pull(P, a.ghalo, W, R, ni, jlo, 0, rank, has_down, has_up, tid); __syncthreads(); }
        TICK(3);
        if (it > 0) {                                                   // close iteration it-1 (every CTA, same order)
            double t = 0.0;
            if (GRID) {
                if (tid < 32) {
                    double q = 0.0;
                    for (int r = tid; r < C; r += 32) q += __ldcg(a.gparts + r);
                    q = warp_sum(q);
                    if (tid == 0) tot_sh = q;
                }
                __syncthreads();
                t = tot_sh;
            } else {
                for (int r = 0; r < C; r++) t += parts[r];
            }
            total = t;
            res = sqrt(t / (double)a.k.cells);                          // main.cu:588
            converged = (res <= thresh) ? 1 : 0;                        // main.cu:713
        }
        if (!real) break;
        if (it > 0 && converged && a.check) {
            resident_pass<RECIP, 2>(P, Q, B, gf, W, R, ni, nj, jlo, 0, false, a.k, tx, ty, TX, TY);
            __syncthreads();
            break;
        }
        TICK(4);
        double acc = resident_pass<RECIP, 0>(P, Q, nullptr, gf, W, R, ni, nj, jlo, 1, true, a.k, tx, ty, TX, TY);
        TICK(5);
        __syncthreads();
        resident_push<GRID>(P, Pdown, Pup, hmine, W, R, R_down, ni, jlo, 1, has_down, has_up, tid);
        resident_sync<GRID>(C);
        if (GRID) { resident_pull(P, a.ghalo, W, R, ni, jlo, 1, rank, has_down, has_up, tid); __syncthreads(); }
        TICK(6);
        // residual of the cells of colour 0: their neighbours are final now; ghosts next to them hold their own value
        acc += resident_pass<RECIP, 1>(P, Q, nullptr, gf, W, R, ni, nj, jlo, 0, false, a.k, tx, ty, TX, TY);
        const double bsum = block_sum(acc, red_sh);
        if (tid == 0) {
            if (GRID) a.gparts[rank] = bsum;
            else for (int r = 0; r < C; r++) cg::this_cluster().map_shared_rank(parts, r)[rank] = bsum;
        }
#ifdef RES_TIMING
        TICK(7);
        if (it == 7 && tid == 0 && (rank == 0 || rank == C / 2))
            printf("rank %d: red %lld sync+push %lld csync %lld close %lld black %lld sync+push+csync %lld redres+sum %lld total %lld\n", rank,
                   tk[1] - tk[0], tk[2] - tk[1], tk[3] - tk[2], tk[4] - tk[3], tk[5] - tk[4], tk[6] - tk[5], tk[7] - tk[6], tk[7] - tk[0]);
#endif
        // (visible to every CTA after the next cluster barrier; rewritten only two barriers later)
    }
    // `it` iterations are complete and closed

    // ---- write the tile back: interior rows, ghost columns / rows as the reference leaves them
    double* gpw = a.p + (long long)NSB_ROW0 * a.pitch + NSB_COL0;
    for (int l = 1; l <= R; l++) {
        const long long ro = (long long)(jlo - 1 + l) * a.pitch;
        for (int i = 1 + tid; i <= ni; i += RES_THREADS) gpw[ro + i] = P[(size_t)l * W + i + 1];
        if (tid == 0 && it > 0) { gpw[ro] = gf.w[l]; gpw[ro + ni + 1] = gf.e[l]; }
    }
    if (it > 0) {
        if (first) for (int i = 1 + tid; i <= ni; i += RES_THREADS) gpw[i] = gf.s[i + 1];
        if (last) for (int i = 1 + tid; i <= ni; i += RES_THREADS) gpw[(long long)(nj + 1) * a.pitch + i] = gf.n[i + 1];
    }
    if (rank == 0 && tid == 0) {
        SorCtl* ctl = a.ctl;
        if (it > 0) { ctl->sum_r2 = total; ctl->res = res; }
        if (converged && it > 0) ctl->converged = 1;
        __threadfence();
        *((volatile int*)&ctl->iters) = a.it0 + it;
        __threadfence();
        sor_publish(ctl);
    }
}

// ------------------------------------------------------------------------------------------ host
struct ResidentPlan {
    int n_cta, rows_per_cta, W, grid;   // grid = 1: cooperative-grid variant, 0: one cluster
    size_t smem;
};

static size_t resident_bytes(int R, int w) {
    // p tile + roll-back tile (R+2 rows each), RHS (R rows), final ghosts (2 (R+2) + 2 w)
    return ((size_t)(3 * R + 4) * w + 2 * (size_t)(R + 2) + 2 * (size_t)w) * sizeof(double);
}

static int resident_plan(const nsb_ctx* c, ResidentPlan* pl) {
    if (c->world != 1) return 0;
    const int ni = c->g.ni, nj = c->g.nj;
    const int w = (ni + 3 + 1) & ~1;
    const long long cells = (long long)ni * nj;
    pl->W = w;
    const char* mode = getenv("NSB_RES_MODE");                        // tuning: "grid" / "cluster" forces a variant
    const bool force_grid = mode && mode[0] == 'g', force_cluster = mode && mode[0] == 'c';
    // one cluster: up to 16 CTAs
    int C = 1;
    while (C < RES_MAX_CLUSTER && cells / C > 4096) C <<= 1;
    const char* env = getenv("NSB_RES_CLUSTER");                      // tuning: smallest cluster size to try
    if (env && atoi(env) >= 1 && atoi(env) <= RES_MAX_CLUSTER) C = atoi(env);
    for (; C <= RES_MAX_CLUSTER && !force_grid; C <<= 1) {
        const int R = (nj + C - 1) / C;
        if (resident_bytes(R, w) <= RES_SMEM_LIMIT && (long long)(C - 1) * R < nj) {   // every CTA owns at least one row
            pl->n_cta = C; pl->rows_per_cta = R; pl->smem = resident_bytes(R, w); pl->grid = 0;
            return 1;
        }
    }
    if (force_cluster) return 0;
    // cooperative grid: up to one CTA per SM
    int G = c->sm_count < nj ? c->sm_count : nj;
    const long long fit = (long long)(c->g.elems / (size_t)(2 * w + 1));   // halo rows + partials live in the scratch grid
    if (G > fit) G = (int)fit;
    if (G < 1) return 0;
    const int R = (nj + G - 1) / G;
    G = (nj + R - 1) / R;
    if (resident_bytes(R, w) > RES_SMEM_LIMIT) return 0;
    pl->n_cta = G; pl->rows_per_cta = R; pl->smem = resident_bytes(R, w); pl->grid = 1;
    return 1;
}

// c->resident_state (per context, hence per device): 0 unknown, 1 usable, -1 the device refused the launch
// (MIG slice, a GPC with fewer free SMs than the cluster needs, another SKU): the caller falls through to the
// streaming sweep within the same solve (step.cu: sor_iterate).
int nsb_sor_resident_supported(const nsb_ctx* c) {
    if (c->resident_state < 0) return 0;
    ResidentPlan pl;
    return resident_plan(c, &pl);
}

cudaError_t nsb_launch_sor_resident(nsb_ctx* c, int max_it, int check) {
    ResidentPlan pl;
    if (!resident_plan(c, &pl)) return cudaErrorNotSupported;
    if (getenv("NSB_DEBUG_RES_REFUSE")) { c->resident_state = -1; return cudaErrorNotSupported; }   // test hook: a device that refuses
    ResidentArgs a;
    memset(&a, 0, sizeof a);
    a.p = c->pbuf[c->pcur]; a.rhs = c->rhs; a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.W = pl.W; a.rows_per_cta = pl.rows_per_cta; a.max_it = max_it; a.check = check; a.it0 = 0;
    a.k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->J);
    a.ctl = c->d_ctl;
    cudaError_t e;
    if (pl.grid) {
        // halo rows + partials in HBM: the scratch grid is free during a solve
        a.ghalo = c->scratch;
        a.gparts = c->scratch + (size_t)pl.n_cta * 2 * pl.W;
        if (((size_t)pl.n_cta * 2 * pl.W + pl.n_cta) > c->g.elems) return cudaErrorNotSupported;
        const void* kern = c->use_recip ? (const void*)sor_resident_kernel<true, true> : (const void*)sor_resident_kernel<false, true>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
        if (e != cudaSuccess) return e;
        int occ = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, RES_THREADS, pl.smem);
        if (occ * c->sm_count < pl.n_cta) { c->resident_state = -1; return cudaErrorNotSupported; }
        void* args[] = {&a};
        e = cudaLaunchCooperativeKernel(kern, dim3(pl.n_cta), dim3(RES_THREADS), args, pl.smem, c->stream);
        if (e != cudaSuccess) { c->resident_state = -1; cudaGetLastError(); return e; }
        c->resident_state = 1;
        return NSB_LAUNCHED(c, 1);
    }
    auto kern = c->use_recip ? sor_resident_kernel<true, false> : sor_resident_kernel<false, false>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) return e;
    if (pl.n_cta > 8) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(pl.n_cta, 1, 1);
    cfg.blockDim = dim3(RES_THREADS, 1, 1);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = c->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = pl.n_cta; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (c->resident_state == 0) {
        // can this device co-schedule one cluster of that size with that much shared memory at all?
        int n_clusters = 0;
        e = cudaOccupancyMaxActiveClusters(&n_clusters, kern, &cfg);
        if (e != cudaSuccess || n_clusters < 1) { c->resident_state = -1; cudaGetLastError(); return cudaErrorNotSupported; }
    }
    e = cudaLaunchKernelEx(&cfg, kern, a);
    if (e != cudaSuccess) { c->resident_state = -1; cudaGetLastError(); return e; }
    c->resident_state = 1;
    return NSB_LAUNCHED(c, 1);
}
