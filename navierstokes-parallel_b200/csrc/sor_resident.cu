// sor_resident.cu -- cluster-resident SOR engine for grids that fit on chip: p and RHS of the whole
// grid live in the shared memory of ONE thread-block cluster (up to 16 SMs), all iterations of a
// solve run inside one launch, neighbouring CTAs exchange their boundary rows through distributed
// shared memory and synchronise with the hardware cluster barrier.
//
// Why: at 256^2 (tests/1.in, 1 MB working set) the solve is pure latency.  The reference spends six
// launches, six device synchronisations and a D2H copy per iteration (src/parallel/main.cu:684-719,
// ~56 us per iteration in its published numbers); the streaming engine still pays a task hand-over
// through L2 flags per iteration.  Here one iteration costs three cluster barriers plus the arithmetic.
//
// Order of operations = src/parallel/main.cu:684-719, with the ghost cells STORED in shared memory
// and refreshed exactly where the reference refreshes them (before colour 0 and before colour 1, not
// before the residual):
//   refresh ghosts | colour 0 | exchange | refresh ghosts | colour 1 (+ residual of the black cells,
//   whose neighbours are final) | exchange | residual of the red cells | cluster-wide sum | test.
// Arithmetic per cell is sor_cell / sor_residual of nsb_math.cuh (reference association order, no
// contraction): p has the same bits as every other engine.  The residual is summed per thread, per
// block (fixed tree) and over the CTAs in rank order: deterministic, but not the summation order of
// the streaming engine -- only the convergence test sees it.
#include <cooperative_groups.h>
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
};

// One colour pass over the rows 1..R of this CTA's tile.  Thread layout: tx along column PAIRS
// (columns 2q+1, 2q+2), ty along rows.  WITH_RES: also accumulate r^2 of the cells just updated
// (valid for the second colour: all their neighbours are final).
template <bool RECIP, bool UPDATE>
__device__ __forceinline__ double resident_pass(double* __restrict__ P, const double* __restrict__ Q, int W, int R,
                                                int ni, int jlo, int color, bool with_res, const SorCoef& kf,
                                                int tx, int ty, int TX, int TY) {
    double acc = 0.0;
    const int npairs = (ni + 1) >> 1;
    for (int l = 1 + ty; l <= R; l += TY) {
        const int j = jlo + l - 1;                                  // global row
        double* row = P + (size_t)l * W;
        const double* rq = Q + (size_t)(l - 1) * W;
        for (int q = tx; q < npairs; q += TX) {
            // the cell of this pair with (i + j) % 2 == color
            const int i = 2 * q + 1 + ((j + 1 + color) & 1);        // i = 2q+1 if (2q+1+j)%2 == color else 2q+2
            if (i > ni) continue;
            const int x = i + 1;
            const double pc = row[x], pe = row[x + 1], pw = row[x - 1];
            const double pn = row[x + W], ps = row[x - W];
            const double r = rq[x];
            if (UPDATE) {
                const double v = sor_cell<RECIP>(pc, pe, pw, pn, ps, r, kf);
                row[x] = v;
                if (with_res) {
                    const double d = sor_residual<RECIP>(v, pe, pw, pn, ps, r, kf);
                    acc = fma(d, d, acc);
                }
            } else {
                const double d = sor_residual<RECIP>(pc, pe, pw, pn, ps, r, kf);
                acc = fma(d, d, acc);
            }
        }
    }
    return acc;
}

// Neumann ghosts of this CTA's tile (main.cu:249-273): west / east columns for its rows, the bottom
// ghost row on the first CTA, the top ghost row on the last.
__device__ __forceinline__ void resident_ghosts(double* P, int W, int R, int ni, bool first, bool last) {
    for (int l = 1 + (int)threadIdx.x; l <= R; l += RES_THREADS) {
        double* row = P + (size_t)l * W;
        row[1] = row[2];                    // p[0][j] = p[1][j]
        row[ni + 2] = row[ni + 1];          // p[i_max+1][j] = p[i_max][j]
    }
    if (first)
        for (int i = 1 + (int)threadIdx.x; i <= ni; i += RES_THREADS) P[i + 1] = P[W + i + 1];                    // p[i][0] = p[i][1]
    if (last)
        for (int i = 1 + (int)threadIdx.x; i <= ni; i += RES_THREADS) P[(size_t)(R + 1) * W + i + 1] = P[(size_t)R * W + i + 1];
}

template <bool RECIP>
__global__ void __launch_bounds__(RES_THREADS, 1) sor_resident_kernel(const ResidentArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double red_sh[32];
    __shared__ double parts[RES_MAX_CLUSTER];
    __shared__ int stop_flag;
    cg::cluster_group cluster = cg::this_cluster();
    const int C = (int)cluster.num_blocks();
    const int rank = (int)cluster.block_rank();
    const int W = a.W, ni = a.ni;
    const int jlo = 1 + rank * a.rows_per_cta;                         // first global row of this CTA
    int R = min(a.rows_per_cta, a.nj - jlo + 1);
    if (R < 0) R = 0;
    const bool first = (rank == 0), last = (jlo + R - 1 == a.nj) && R > 0;
    double* P = reinterpret_cast<double*>(smem_raw);                  // (rows_per_cta + 2) x W : rows jlo-1 .. jlo+R
    double* Q = P + (size_t)(a.rows_per_cta + 2) * W;                 // rows_per_cta x W        : RHS of rows jlo .. jlo+R-1
    const int tid = threadIdx.x;

    // ---- load the tile (row l of P = global row jlo - 1 + l), x index = i + 1
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
    // neighbours' tiles through distributed shared memory
    double* Pdown = (rank > 0) ? cluster.map_shared_rank(P, rank - 1) : nullptr;
    double* Pup = (rank + 1 < C) ? cluster.map_shared_rank(P, rank + 1) : nullptr;
    const int R_down = a.rows_per_cta;                                // every CTA below the last one is full
    const bool has_up = (rank + 1 < C) && (jlo + R <= a.nj);          // rows exist above this CTA
    const bool has_down = (rank > 0) && R > 0;
    // thread layout for the passes
    const int npairs = (ni + 1) >> 1;
    int TX = 32;
    while (TX < npairs && TX < RES_THREADS) TX <<= 1;
    const int TY = RES_THREADS / TX;
    const int tx = tid % TX, ty = tid / TX;
    const double thresh = a.ctl->thresh;
    if (tid == 0) stop_flag = 0;
    __syncthreads();
    cluster.sync();

    int it = 0, converged = 0;
    double res = a.ctl->res, total = a.ctl->sum_r2;
    for (; it < a.max_it; it++) {
        double acc = 0.0;
        for (int color = 0; color < 2; color++) {
            if (R > 0) resident_ghosts(P, W, R, ni, first, last);                       // main.cu:686 / :694
            __syncthreads();
            if (R > 0) acc += resident_pass<RECIP, true>(P, Q, W, R, ni, jlo, color, color == 1, a.k, tx, ty, TX, TY);
            __syncthreads();
            // boundary rows to the neighbours' halo rows: only the cells of this colour changed
            if (has_down) {
                const int j = jlo;                                                      // my row l = 1 -> their row R_down + 1
                for (int i = 1 + ((j + 1 + color) & 1) + 2 * tid; i <= ni; i += 2 * RES_THREADS)
                    Pdown[(size_t)(R_down + 1) * W + i + 1] = P[(size_t)W + i + 1];
            }
            if (has_up) {
                const int j = jlo + R - 1;                                              // my row l = R -> their row 0
                for (int i = 1 + ((j + 1 + color) & 1) + 2 * tid; i <= ni; i += 2 * RES_THREADS)
                    Pup[i + 1] = P[(size_t)R * W + i + 1];
            }
            cluster.sync();
        }
        // residual of the red cells (their black neighbours are final now), ghosts as refreshed before colour 1
        if (R > 0) acc += resident_pass<RECIP, false>(P, Q, W, R, ni, jlo, 0, false, a.k, tx, ty, TX, TY);
        const double bsum = block_sum(acc, red_sh);
        if (tid == 0) {
            for (int r = 0; r < C; r++) {
                double* dst = cluster.map_shared_rank(parts, r);
                dst[rank] = bsum;
            }
        }
        cluster.sync();
        if (tid == 0) {
            double t = 0.0;
            for (int r = 0; r < C; r++) t += parts[r];
            total = t;
            res = sqrt(t / (double)a.k.cells);                                          // main.cu:588
            stop_flag = (res <= thresh) ? 1 : 0;                                        // main.cu:713
        }
        __syncthreads();
        converged = stop_flag;
        // `parts` is rewritten only after the next iteration's cluster barriers: no race with the reads above
        if (converged && a.check) { it++; break; }
    }

    // ---- write the tile back: interior rows with their ghost columns, ghost rows on the first / last CTA
    double* gpw = a.p + (long long)NSB_ROW0 * a.pitch + NSB_COL0;
    const int l0 = first ? 0 : 1, l1 = last ? R + 1 : R;
    for (int l = l0; l <= l1 && R > 0; l++) {
        const long long ro = (long long)(jlo - 1 + l) * a.pitch;
        const bool ghost_row = (l == 0) || (l == R + 1);
        for (int i = tid; i <= ni + 1; i += RES_THREADS) {
            if (ghost_row && (i == 0 || i == ni + 1)) continue;                         // corners: never written by the reference's reads
            gpw[ro + i] = P[(size_t)l * W + i + 1];
        }
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
static int resident_plan(const nsb_ctx* c, int* cluster, int* rows_per_cta, int* W, size_t* smem) {
    if (c->world != 1) return 0;
    const int ni = c->g.ni, nj = c->g.nj;
    const int w = (ni + 3 + 1) & ~1;
    const long long cells = (long long)ni * nj;
    int C = 1;
    while (C < RES_MAX_CLUSTER && cells / C > 4096) C <<= 1;
    for (; C <= RES_MAX_CLUSTER; C <<= 1) {
        const int R = (nj + C - 1) / C;
        const size_t bytes = (size_t)(2 * R + 2) * w * sizeof(double);
        if (bytes <= RES_SMEM_LIMIT && (long long)(C - 1) * R < nj) {       // every CTA owns at least one row
            *cluster = C; *rows_per_cta = R; *W = w; *smem = bytes;
            return 1;
        }
    }
    return 0;
}

static int g_resident_state = 0;   // 0 unknown, 1 usable, -1 the device refused a cluster launch

int nsb_sor_resident_supported(const nsb_ctx* c) {
    if (g_resident_state < 0) return 0;
    int C, R, W; size_t smem;
    return resident_plan(c, &C, &R, &W, &smem);
}

cudaError_t nsb_launch_sor_resident(nsb_ctx* c, int max_it, int check) {
    int C, R, W; size_t smem;
    if (!resident_plan(c, &C, &R, &W, &smem)) return cudaErrorNotSupported;
    ResidentArgs a;
    memset(&a, 0, sizeof a);
    a.p = c->pbuf[c->pcur]; a.rhs = c->rhs; a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.W = W; a.rows_per_cta = R; a.max_it = max_it; a.check = check; a.it0 = 0;
    const double dx2 = c->dx * c->dx, dy2 = c->dy * c->dy;
    a.k.om1 = 1.0 - c->P.omega;
    a.k.w = c->P.omega / (2.0 * (1.0 / dx2 + 1.0 / dy2));
    a.k.dx2 = dx2; a.k.dy2 = dy2; a.k.rdx2 = 1.0 / dx2; a.k.rdy2 = 1.0 / dy2;
    a.k.cells = (long long)c->g.ni * (long long)c->J;
    a.ctl = c->d_ctl;
    auto kern = c->use_recip ? sor_resident_kernel<true> : sor_resident_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (C > 8) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(C, 1, 1);
    cfg.blockDim = dim3(RES_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, kern, a);
    if (e != cudaSuccess) { g_resident_state = -1; cudaGetLastError(); return e; }
    g_resident_state = 1;
    return NSB_LAUNCHED(c, 1);
}
