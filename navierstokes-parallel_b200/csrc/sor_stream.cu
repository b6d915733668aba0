// sor_stream.cu -- fused red + black + residual SOR iteration as ONE streaming pass over HBM.
//
// Replaces, per iteration, the six launches of src/parallel/main.cu:684-710
// (ghost refresh, colour 0, ghost refresh, colour 1, residual, block-norm reduce) and reproduces
// their order exactly, without ghost storage:
//   * during a sweep a wall-side neighbour is the cell's own pre-update value (ghost refreshed
//     just before each colour, main.cu:686/694);
//   * in the residual the wall-side neighbour of a wall-adjacent RED cell is its current value,
//     of a wall-adjacent BLACK cell its PRE-black value (ghosts are not refreshed after colour 1).
//
// Design (DESIGN.md "SOR streaming kernel"):
//   - ping-pong pressure buffers: read p_in, write p_out, so tiles never see half-updated
//     neighbours and both colours are applied in one pass: algorithmic traffic 24 B per cell per
//     iteration (p in 8 + RHS in 8 + p out 8).
//   - one WARP owns a strip of 30 column groups (120 columns) x chunk_rows rows and marches up in
//     j with a register-resident window of 8 rows; lane L owns 4 adjacent columns (one 32-byte
//     sector), lanes 0 and 31 are halo lanes.  x-neighbours come from warp shuffles, y-neighbours
//     from the register window, so there is no shared memory, no block barrier and every load and
//     store is a fully used, aligned 32-byte sector.
//   - software pipeline inside the window: when row j arrives, row j-1 gets its red update, row
//     j-2 its black update (and is stored), row j-3 its residual.  Loads run 3 rows ahead.
//   - residual: per-warp partial (fixed order), last block of the grid adds the partials in a fixed
//     order, takes sqrt(sum/(i_max*j_max)) (main.cu:588), compares with eps*(norm_p+1.5)
//     (main.cu:713) and raises the device-side converged flag; later launches exit at once.
#include "nsb_internal.h"
#include "nsb_math.cuh"

#define SOR_WARPS 4
#define SOR_THREADS (SOR_WARPS * 32)
#define SOR_MIN_BLOCKS 3
#define SOR_WINDOW 8
#define SOR_AHEAD 3          // p rows loaded ahead of the row being consumed

struct SorStreamArgs {
    const double* __restrict__ pin;
    double* __restrict__ pout;
    const double* __restrict__ rhs;
    long long pitch;
    int ni, nj;
    int chunk_rows, n_strips, n_tasks;
    int n_groups;            // pitch / 4
    int it, check;
    SorCoef k;
    double* partials;
    SorCtl* ctl;
};

// All threads of the LAST block of an iteration: fixed-order sum of the partials, convergence test.
__device__ __forceinline__ void sor_close_iteration(const double* partials, int n, SorCtl* ctl, int it,
                                                    long long cells, double* sh) {
    double s = 0.0;
    for (int q = threadIdx.x; q < n; q += blockDim.x) s += __ldcg(partials + q);
    double total = block_sum(s, sh);
    if (threadIdx.x == 0) {
        double res = sqrt(total / (double)cells);          // main.cu:588
        ctl->sum_r2 = total;
        ctl->res = res;
        ctl->iters = it + 1;
        if (res <= ctl->thresh) ctl->converged = 1;         // main.cu:713
        ctl->ticket = 0;
        __threadfence();
    }
}

__device__ __forceinline__ void load_group(double (&dst)[4], const double* row, bool ok) {
    if (ok) {
        const double2* q = reinterpret_cast<const double2*>(row);
        double2 a = __ldg(q), b = __ldg(q + 1);
        dst[0] = a.x; dst[1] = a.y; dst[2] = b.x; dst[3] = b.y;
    } else {
        dst[0] = dst[1] = dst[2] = dst[3] = 0.0;
    }
}

struct LaneGeom {
    bool wallW;      // column k=0 is i == 1
    bool wallE[4];   // column k is i == i_max
    bool inner[4];   // column k is an interior column
    bool owner;      // lanes 1..30 with an in-bounds group: stores and residual
};

// One pipeline step with compile-time window slot S (row parity == S parity).
template <bool RECIP, int S>
__device__ __forceinline__ void sor_step(double (&P)[SOR_WINDOW][4], double (&R)[SOR_WINDOW][4],
                                         double (&PB)[2][2], const int j, const SorStreamArgs& a,
                                         const LaneGeom& lg, const double* pin_col,
                                         const double* rhs_col, double* pout_col, const bool col_ok,
                                         const int jc0, const int jc1, double& acc) {
    constexpr int sE = S & 7, sD = (S + 7) & 7, sC = (S + 6) & 7, sB = (S + 5) & 7, sA = (S + 4) & 7;
    constexpr int sPF = (S + SOR_AHEAD) & 7, sRF = (S + SOR_AHEAD - 1) & 7;
    constexpr int k0 = (S & 1) ? 1 : 0;     // cells updated this step: k0 and k0 + 2 (see header)
    constexpr int k1 = k0 + 2;
    const unsigned FULL = 0xffffffffu;

    // ---- prefetch: p row j+3, rhs row j+2
    {
        const int rp = j + SOR_AHEAD, rr = j + SOR_AHEAD - 1;
        load_group(P[sPF], pin_col + (long long)rp * a.pitch, col_ok && rp >= 0 && rp <= a.nj + 1 && rp <= jc1 + 3);
        load_group(R[sRF], rhs_col + (long long)rr * a.pitch, col_ok && rr >= 1 && rr <= a.nj && rr <= jc1 + 2);
    }

    // ---- colour 0 (red) on row D = j-1, neighbours: C (j-2, black cells still old), E (j)
    {
        const int r = j - 1;
        const bool row_in = (r >= 1) && (r <= a.nj);
        const bool sWall = (r == 1), nWall = (r == a.nj);
        double (&X)[4] = P[sD];
        // the one neighbour that lives in another lane
        double xw = 0.0, xe = 0.0;
        if (k0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double c0 = X[k0], c1 = X[k1];
        double w0 = (k0 == 0) ? (lg.wallW ? c0 : xw) : X[(k0 + 3) & 3];
        double e0 = lg.wallE[k0] ? c0 : X[k0 + 1];
        double w1 = X[k1 - 1];
        double e1 = lg.wallE[k1] ? c1 : ((k1 == 3) ? xe : X[(k1 + 1) & 3]);
        double n0 = nWall ? c0 : P[sE][k0], s0 = sWall ? c0 : P[sC][k0];
        double n1 = nWall ? c1 : P[sE][k1], s1 = sWall ? c1 : P[sC][k1];
        double u0 = sor_cell<RECIP>(c0, e0, w0, n0, s0, R[sD][k0], a.k);
        double u1 = sor_cell<RECIP>(c1, e1, w1, n1, s1, R[sD][k1], a.k);
        if (row_in && lg.inner[k0]) X[k0] = u0;
        if (row_in && lg.inner[k1]) X[k1] = u1;
    }

    // ---- colour 1 (black) on row C = j-2, neighbours: B (j-3, final), D (j-1, red done)
    {
        const int r = j - 2;
        const bool row_in = (r >= 1) && (r <= a.nj);
        const bool sWall = (r == 1), nWall = (r == a.nj);
        double (&X)[4] = P[sC];
        double xw = 0.0, xe = 0.0;
        if (k0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double c0 = X[k0], c1 = X[k1];
        PB[S & 1][0] = c0;                       // pre-black values, needed by the residual at walls
        PB[S & 1][1] = c1;
        double w0 = (k0 == 0) ? (lg.wallW ? c0 : xw) : X[(k0 + 3) & 3];
        double e0 = lg.wallE[k0] ? c0 : X[k0 + 1];
        double w1 = X[k1 - 1];
        double e1 = lg.wallE[k1] ? c1 : ((k1 == 3) ? xe : X[(k1 + 1) & 3]);
        double n0 = nWall ? c0 : P[sD][k0], s0 = sWall ? c0 : P[sB][k0];
        double n1 = nWall ? c1 : P[sD][k1], s1 = sWall ? c1 : P[sB][k1];
        double u0 = sor_cell<RECIP>(c0, e0, w0, n0, s0, R[sC][k0], a.k);
        double u1 = sor_cell<RECIP>(c1, e1, w1, n1, s1, R[sC][k1], a.k);
        if (row_in && lg.inner[k0]) X[k0] = u0;
        if (row_in && lg.inner[k1]) X[k1] = u1;
        // row j-2 is final: store it
        if (lg.owner && r >= jc0 && r <= jc1) {
            double2* q = reinterpret_cast<double2*>(pout_col + (long long)r * a.pitch);
            q[0] = make_double2(X[0], X[1]);
            q[1] = make_double2(X[2], X[3]);
        }
    }

    // ---- residual on row B = j-3, neighbours A (j-4) and C (j-2), all final
    {
        const int r = j - 3;
        double (&X)[4] = P[sB];
        // black cells of this row are kb0 = 1-k0.. : the set NOT updated this step
        constexpr int kb0 = (S & 1) ? 0 : 1;
        const double xw = __shfl_up_sync(FULL, X[3], 1);
        const double xe = __shfl_down_sync(FULL, X[0], 1);
        if (r >= jc0 && r <= jc1) {     // uniform; rows of this chunk are interior rows
            const bool sWall = (r == 1), nWall = (r == a.nj);
            double l1[4];               // value each cell had when the ghosts were last refreshed
#pragma unroll
            for (int k = 0; k < 4; k++) l1[k] = X[k];
            l1[kb0] = PB[(S + 1) & 1][0];
            l1[kb0 + 2] = PB[(S + 1) & 1][1];
            double part = 0.0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                double w = (k == 0) ? (lg.wallW ? l1[0] : xw) : X[(k + 3) & 3];
                double e = lg.wallE[k] ? l1[k] : ((k == 3) ? xe : X[(k + 1) & 3]);
                double n = nWall ? l1[k] : P[sC][k];
                double s = sWall ? l1[k] : P[sA][k];
                double q = sor_residual<RECIP>(X[k], e, w, n, s, R[sB][k], a.k);
                if (lg.inner[k]) part = fma(q, q, part);
            }
            if (lg.owner) acc += part;
        }
    }
}

template <bool RECIP>
__global__ void __launch_bounds__(SOR_THREADS, SOR_MIN_BLOCKS) sor_stream_kernel(const SorStreamArgs a) {
    __shared__ double sh[32];
    __shared__ int is_last;
    if (a.check && *((volatile int*)&a.ctl->converged)) return;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int task = blockIdx.x * SOR_WARPS + warp;
    if (task < a.n_tasks) {
        const int strip = task % a.n_strips, chunk = task / a.n_strips;
        const int g = strip * NSB_STRIP_GROUPS + lane;       // column group of this lane
        const int i0 = g * NSB_GROUP - NSB_COL0;             // i of column k = 0
        const bool col_ok = g < a.n_groups;
        LaneGeom lg;
        lg.wallW = (i0 == 1);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            lg.wallE[k] = (i0 + k == a.ni);
            lg.inner[k] = (i0 + k >= 1) && (i0 + k <= a.ni);
        }
        lg.owner = col_ok && lane >= 1 && lane <= NSB_STRIP_GROUPS;
        const int jc0 = 1 + chunk * a.chunk_rows;
        const int jc1 = min(jc0 + a.chunk_rows - 1, a.nj);
        const int jstart = jc0 - 3, jend = jc1 + 3;          // jstart is even (chunk_rows is even)
        const long long coff = (long long)NSB_ROW0 * a.pitch + (long long)g * NSB_GROUP;
        const double* pin_col = a.pin + coff;
        const double* rhs_col = a.rhs + coff;
        double* pout_col = a.pout + coff;

        double P[SOR_WINDOW][4], R[SOR_WINDOW][4], PB[2][2];
#pragma unroll
        for (int s = 0; s < SOR_WINDOW; s++)
#pragma unroll
            for (int k = 0; k < 4; k++) { P[s][k] = 0.0; R[s][k] = 0.0; }
        PB[0][0] = PB[0][1] = PB[1][0] = PB[1][1] = 0.0;
        // prologue: p rows jstart .. jstart+2 into slots 0..2, rhs rows jstart, jstart+1
#pragma unroll
        for (int s = 0; s < SOR_AHEAD; s++) {
            const int r = jstart + s;
            load_group(P[s], pin_col + (long long)r * a.pitch, col_ok && r >= 0 && r <= a.nj + 1);
        }
#pragma unroll
        for (int s = 0; s < SOR_AHEAD - 1; s++) {
            const int r = jstart + s;
            load_group(R[s], rhs_col + (long long)r * a.pitch, col_ok && r >= 1 && r <= a.nj);
        }
        double acc = 0.0;
        for (int jb = jstart; jb <= jend; jb += SOR_WINDOW) {
#define NSB_STEP(S)                                                                              \
    if (jb + S <= jend)                                                                          \
        sor_step<RECIP, S>(P, R, PB, jb + S, a, lg, pin_col, rhs_col, pout_col, col_ok, jc0, jc1, acc);
            NSB_STEP(0) NSB_STEP(1) NSB_STEP(2) NSB_STEP(3)
            NSB_STEP(4) NSB_STEP(5) NSB_STEP(6) NSB_STEP(7)
#undef NSB_STEP
        }
        acc = warp_sum(acc);
        if (lane == 0) a.partials[task] = acc;
    }

    // ---- last block closes the iteration
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = atomicAdd(&a.ctl->ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        sor_close_iteration(a.partials, a.n_tasks, a.ctl, a.it, a.k.cells, sh);
    }
}

// ------------------------------------------------------------------------------------------
// Planning: strips are fixed (120 columns); pick an even chunk height that fills whole waves.
void nsb_sor_plan(nsb_ctx* c) {
    const int ni = c->g.ni, nj = c->g.nj;
    c->n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    int blocks_per_sm = 0;
    if (c->use_recip)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, sor_stream_kernel<true>, SOR_THREADS, 0);
    else
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, sor_stream_kernel<false>, SOR_THREADS, 0);
    if (blocks_per_sm < 1) blocks_per_sm = 1;
    const long long slots = (long long)c->sm_count * blocks_per_sm * SOR_WARPS;
    int best = 0;
    double best_cost = 1e300;
    if (c->O.sor_chunk_rows > 0) {
        best = (c->O.sor_chunk_rows + 1) & ~1;
    } else {
        for (int rows = 16; rows <= 1024; rows += 2) {
            long long chunks = (nj + rows - 1) / rows;
            long long tasks = chunks * c->n_strips;
            long long waves = (tasks + slots - 1) / slots;
            double cost = (double)waves * (rows + 6 + 4);      // +4: prologue/epilogue latency
            if (cost < best_cost - 1e-9) { best_cost = cost; best = rows; }
            if (rows >= nj) break;
        }
    }
    if (best < 2) best = 2;
    c->chunk_rows = best;
    c->n_chunks = (nj + best - 1) / best;
    c->n_partials = c->n_chunks * c->n_strips;
}

cudaError_t nsb_launch_sor_stream_iter(nsb_ctx* c, int it, const double* pin, double* pout, int check) {
    SorStreamArgs a;
    a.pin = pin; a.pout = pout; a.rhs = c->rhs;
    a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.chunk_rows = c->chunk_rows; a.n_strips = c->n_strips; a.n_tasks = c->n_partials;
    a.n_groups = (int)(c->g.pitch / NSB_GROUP);
    a.it = it; a.check = check;
    const double dx2 = c->dx * c->dx, dy2 = c->dy * c->dy;
    a.k.om1 = 1.0 - c->P.omega;
    a.k.w = c->P.omega / (2.0 * (1.0 / dx2 + 1.0 / dy2));
    a.k.dx2 = dx2; a.k.dy2 = dy2; a.k.rdx2 = 1.0 / dx2; a.k.rdy2 = 1.0 / dy2;
    a.k.inv_cells_unused = 0.0;
    a.k.cells = (long long)c->g.ni * (long long)c->g.nj;
    a.partials = c->d_partials; a.ctl = c->d_ctl;
    const int blocks = (a.n_tasks + SOR_WARPS - 1) / SOR_WARPS;
    if (c->use_recip) sor_stream_kernel<true><<<blocks, SOR_THREADS, 0, c->stream>>>(a);
    else sor_stream_kernel<false><<<blocks, SOR_THREADS, 0, c->stream>>>(a);
    return cudaGetLastError();
}
