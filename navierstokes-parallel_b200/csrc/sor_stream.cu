// sor_stream.cu -- fused red + black + residual SOR iteration as ONE streaming pass over HBM.
//
// Replaces, per iteration, the six launches of src/parallel/main.cu:684-710
// (ghost refresh, colour 0, ghost refresh, colour 1, residual, block-norm reduce) and reproduces
// their order exactly.  Ghost cells are not stored in HBM; they are materialised in
// REGISTERS and refreshed exactly where the reference refreshes them: before colour 0 and before
// colour 1 (main.cu:686/694), not before the residual -- so the residual sees, next to a wall, the
// value the wall cell had before the black sweep, as the reference does.
//
// Design (DESIGN.md "SOR streaming kernel"):
//   - ping-pong pressure buffers: read p_in, write p_out, so a tile never sees a half-updated
//     neighbour and both colours are applied in one pass: algorithmic traffic 24 B per cell per
//     iteration (p in 8 + RHS in 8 + p out 8).
//   - one WARP owns a strip of 30 column groups (120 columns) x chunk_rows rows and marches up in
//     j.  Lane L owns 4 adjacent columns (one 32-byte sector); lanes 0 and 31 are halo lanes.
//     x-neighbours come from warp shuffles, y-neighbours from a register window of 5 rows, so
//     there is no block barrier anywhere.
//   - rows are fetched by the TMA engine: one elected lane issues `cp.async.bulk` (1 KB of p and
//     1 KB of RHS per row) into a per-warp ring of shared-memory stages, completion on an
//     mbarrier; the ring runs SOR_STAGES rows ahead of the arithmetic, which keeps enough bytes in
//     flight to cover HBM latency without spending registers on prefetch.
//   - software pipeline inside the window: when row j arrives, row j-1 gets its red update, row
//     j-2 its black update (and is stored, two 16-byte stores = one full sector per lane), row
//     j-3 its residual.
//   - residual: per-warp partial (fixed order), last block of the grid adds the partials in a fixed
//     order, takes sqrt(sum/(i_max*j_max)) (main.cu:588), compares with eps*(norm_p+1.5)
//     (main.cu:713) and raises the device-side converged flag; later launches exit at once.
#include "nsb_internal.h"
#include "nsb_math.cuh"

#ifndef SOR_WARPS
#define SOR_WARPS 4
#endif
#define SOR_THREADS (SOR_WARPS * 32)
#ifndef SOR_MIN_BLOCKS
#define SOR_MIN_BLOCKS 3
#endif
#ifndef SOR_STAGES
#define SOR_STAGES 8         // rows in flight per warp (power of two, <= 8)
#endif
#define SOR_ROW_BYTES 1024   // 32 lanes x 32 B
#define SOR_STAGE_BYTES (2 * SOR_ROW_BYTES)
#define SOR_WARP_SMEM (SOR_STAGES * SOR_STAGE_BYTES)
#define SOR_SMEM_BYTES (SOR_WARPS * SOR_WARP_SMEM + SOR_WARPS * SOR_STAGES * 8)

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

// ---------------------------------------------------------------- mbarrier / bulk-copy PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
// global -> shared bulk copy executed by the TMA engine (UBLKCP in SASS), completes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

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

// Per-lane geometry, fixed for the whole task.
struct LaneGeom {
    bool inner[4];   // column k is an interior column (1 <= i <= i_max)
    bool owner;      // lanes 1..30 with an in-bounds group: stores and residual
    bool ghostW;     // this lane's column 3 is the west ghost column i = 0
    int kGhostE;     // column of this lane that is the east ghost i = i_max+1, or -1
    bool warpW;      // warp-uniform: some lane holds the west ghost
    bool warpE;      // warp-uniform: some lane holds the east ghost
};

// Neumann ghost columns of one row, refreshed in registers (main.cu:249-273: p[0][j] = p[1][j],
// p[i_max+1][j] = p[i_max][j]).  Warp-uniform branches: only the first / last strip enter.
__device__ __forceinline__ void refresh_ghost_columns(double (&X)[4], const LaneGeom& lg) {
    const unsigned FULL = 0xffffffffu;
    if (lg.warpW) {
        const double t = __shfl_down_sync(FULL, X[0], 1);       // i = 1 lives in the next lane, column 0
        if (lg.ghostW) X[3] = t;
    }
    if (lg.warpE) {
        const double t = __shfl_up_sync(FULL, X[3], 1);         // i = i_max may live in the previous lane
        if (lg.kGhostE == 0) X[0] = t;
        else if (lg.kGhostE == 1) X[1] = X[0];
        else if (lg.kGhostE == 2) X[2] = X[1];
        else if (lg.kGhostE == 3) X[3] = X[2];
    }
}

struct TaskCtx {
    int jstart, jc0, jc1;
    int nj;
    long long pitch;
    const double* pin_row0;   // &pin[row 0][first column of the strip]
    const double* rhs_row0;
    double* pout_col;         // &pout[row 0][this lane's group]
    uint32_t copy_bytes;      // bytes per row copy (1 KB, clipped at the right edge of the pitch)
    uint32_t smem_base;       // this warp's ring
    uint32_t bar_base;        // this warp's mbarriers
    const double* smem_gen;   // generic pointer to this warp's ring (for LDS)
    int lane;
};

// Elected lane: arm stage `st` and start the copies of row r (p) and row r (RHS) when in range.
__device__ __forceinline__ void issue_row(const TaskCtx& t, int st, int r) {
    const bool p_ok = (r >= 0) && (r <= t.nj + 1) && (r <= t.jc1 + 3);
    const bool r_ok = (r >= 1) && (r <= t.nj) && (r <= t.jc1 + 2);
    const uint32_t bar = t.bar_base + st * 8;
    if (!p_ok) {            // nothing to fetch: complete the phase anyway so parities stay in step
        mbar_arrive(bar);
        return;
    }
    const uint32_t dst = t.smem_base + st * SOR_STAGE_BYTES;
    mbar_expect_tx(bar, r_ok ? 2 * t.copy_bytes : t.copy_bytes);
    bulk_g2s(dst, t.pin_row0 + (long long)r * t.pitch, t.copy_bytes, bar);
    if (r_ok) bulk_g2s(dst + SOR_ROW_BYTES, t.rhs_row0 + (long long)r * t.pitch, t.copy_bytes, bar);
}

// One pipeline step with compile-time window slot S (row parity == S parity).  `step` is the step
// number since the start of the task (step & 7 == S).
template <bool RECIP, int S>
__device__ __forceinline__ void sor_step(double (&P)[8][4], double (&R)[8][4], const int step,
                                         const TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double& acc) {
    constexpr int sE = S & 7, sD = (S + 7) & 7, sC = (S + 6) & 7, sB = (S + 5) & 7, sA = (S + 4) & 7;
    constexpr int st = S & (SOR_STAGES - 1);
    constexpr int k0 = (S & 1) ? 1 : 0;     // cells updated this step: columns k0 and k0 + 2
    constexpr int k1 = k0 + 2;
    const unsigned FULL = 0xffffffffu;
    const int j = t.jstart + step;

    // ---- row j (p and RHS) from the ring into registers, then hand the stage back to the TMA
    {
        const bool has_row = (j >= 0) && (j <= t.nj + 1) && (j <= t.jc1 + 3);
        mbar_wait(t.bar_base + st * 8, (uint32_t)(step / SOR_STAGES) & 1u);
        if (has_row) {
            const double2* q = reinterpret_cast<const double2*>(t.smem_gen + st * (SOR_STAGE_BYTES / 8)) + 2 * t.lane;
            const double2 a = q[0], b = q[1];
            const double2 c = q[SOR_ROW_BYTES / 16], d = q[SOR_ROW_BYTES / 16 + 1];
            P[sE][0] = a.x; P[sE][1] = a.y; P[sE][2] = b.x; P[sE][3] = b.y;
            R[sE][0] = c.x; R[sE][1] = c.y; R[sE][2] = d.x; R[sE][3] = d.y;
        }
        __syncwarp();
        if (t.lane == 0) issue_row(t, st, j + SOR_STAGES);
    }

    // ---- colour 0 (red) on row D = j-1; neighbours C (j-2, black cells still old) and E (j)
    {
        const int r = j - 1;
        const bool row_in = (r >= 1) && (r <= t.nj);
        double (&X)[4] = P[sD];
        refresh_ghost_columns(X, lg);                                   // main.cu:686
        if (r == 1) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sC][k] = X[k];                // p[i][0] = p[i][1]
        }
        if (r == t.nj) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sE][k] = X[k];                // p[i][j_max+1] = p[i][j_max]
        }
        double xw = 0.0, xe = 0.0;
        if (k0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double w0 = (k0 == 0) ? xw : X[(k0 + 3) & 3];
        const double e0 = X[k0 + 1];
        const double w1 = X[k1 - 1];
        const double e1 = (k1 == 3) ? xe : X[(k1 + 1) & 3];
        const double u0 = sor_cell<RECIP>(X[k0], e0, w0, P[sE][k0], P[sC][k0], R[sD][k0], kf);
        const double u1 = sor_cell<RECIP>(X[k1], e1, w1, P[sE][k1], P[sC][k1], R[sD][k1], kf);
        if (row_in && lg.inner[k0]) X[k0] = u0;
        if (row_in && lg.inner[k1]) X[k1] = u1;
    }

    // ---- colour 1 (black) on row C = j-2; neighbours B (j-3, final) and D (j-1, red done)
    {
        const int r = j - 2;
        const bool row_in = (r >= 1) && (r <= t.nj);
        double (&X)[4] = P[sC];
        refresh_ghost_columns(X, lg);                                   // main.cu:694
        if (r == 1) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sB][k] = X[k];
        }
        if (r == t.nj) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sD][k] = X[k];
        }
        double xw = 0.0, xe = 0.0;
        if (k0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double w0 = (k0 == 0) ? xw : X[(k0 + 3) & 3];
        const double e0 = X[k0 + 1];
        const double w1 = X[k1 - 1];
        const double e1 = (k1 == 3) ? xe : X[(k1 + 1) & 3];
        const double u0 = sor_cell<RECIP>(X[k0], e0, w0, P[sD][k0], P[sB][k0], R[sC][k0], kf);
        const double u1 = sor_cell<RECIP>(X[k1], e1, w1, P[sD][k1], P[sB][k1], R[sC][k1], kf);
        if (row_in && lg.inner[k0]) X[k0] = u0;
        if (row_in && lg.inner[k1]) X[k1] = u1;
        // row j-2 is final: store it (one full 32-byte sector per lane)
        if (lg.owner && r >= t.jc0 && r <= t.jc1) {
            double2* q = reinterpret_cast<double2*>(t.pout_col + (long long)r * t.pitch);
            q[0] = make_double2(X[0], X[1]);
            q[1] = make_double2(X[2], X[3]);
        }
    }

    // ---- residual on row B = j-3 with the ghosts as they are (last refreshed before colour 1)
    {
        const int r = j - 3;
        double (&X)[4] = P[sB];
        const double xw = __shfl_up_sync(FULL, X[3], 1);
        const double xe = __shfl_down_sync(FULL, X[0], 1);
        if (r >= t.jc0 && r <= t.jc1) {       // warp-uniform; rows of the chunk are interior rows
            double part = 0.0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const double w = (k == 0) ? xw : X[(k + 3) & 3];
                const double e = (k == 3) ? xe : X[(k + 1) & 3];
                const double q = sor_residual<RECIP>(X[k], e, w, P[sC][k], P[sA][k], R[sB][k], kf);
                if (lg.inner[k]) part = fma(q, q, part);
            }
            if (lg.owner) acc += part;
        }
    }
}

template <bool RECIP>
__global__ void __launch_bounds__(SOR_THREADS, SOR_MIN_BLOCKS) sor_stream_kernel(const SorStreamArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ double sh[32];
    __shared__ int is_last;
    if (a.check && *((volatile int*)&a.ctl->converged)) return;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int task = blockIdx.x * SOR_WARPS + warp;
    if (task < a.n_tasks) {
        const int strip = task % a.n_strips, chunk = task / a.n_strips;
        const int g0 = strip * NSB_STRIP_GROUPS;             // first column group of the strip
        const int g = g0 + lane;                             // column group of this lane
        const int i0 = g * NSB_GROUP - NSB_COL0;             // i of column k = 0
        LaneGeom lg;
#pragma unroll
        for (int k = 0; k < 4; k++) lg.inner[k] = (i0 + k >= 1) && (i0 + k <= a.ni);
        lg.owner = (g < a.n_groups) && lane >= 1 && lane <= NSB_STRIP_GROUPS;
        lg.ghostW = (i0 + 3 == 0);
        const int kg = a.ni + 1 - i0;
        lg.kGhostE = (kg >= 0 && kg <= 3) ? kg : -1;
        lg.warpW = __any_sync(0xffffffffu, lg.ghostW);
        lg.warpE = __any_sync(0xffffffffu, lg.kGhostE >= 0);

        TaskCtx t;
        t.jc0 = 1 + chunk * a.chunk_rows;
        t.jc1 = min(t.jc0 + a.chunk_rows - 1, a.nj);
        t.jstart = t.jc0 - 3;                                // even, because chunk_rows is even
        t.nj = a.nj;
        t.pitch = a.pitch;
        const long long coff = (long long)NSB_ROW0 * a.pitch + (long long)g0 * NSB_GROUP;
        t.pin_row0 = a.pin + coff;
        t.rhs_row0 = a.rhs + coff;
        t.pout_col = a.pout + coff + (long long)lane * NSB_GROUP;
        const int groups_here = min(32, a.n_groups - g0);
        t.copy_bytes = (uint32_t)groups_here * 32u;
        unsigned char* ring = smem_raw + warp * SOR_WARP_SMEM;
        t.smem_gen = reinterpret_cast<const double*>(ring);
        t.smem_base = smem_u32(ring);
        t.bar_base = smem_u32(smem_raw + SOR_WARPS * SOR_WARP_SMEM + warp * SOR_STAGES * 8);
        t.lane = lane;

        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < SOR_STAGES; s++) mbar_init(t.bar_base + s * 8, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
            for (int s = 0; s < SOR_STAGES; s++) issue_row(t, s, t.jstart + s);
        }
        __syncwarp();

        double P[8][4], R[8][4];
#pragma unroll
        for (int s = 0; s < 8; s++)
#pragma unroll
            for (int k = 0; k < 4; k++) { P[s][k] = 0.0; R[s][k] = 0.0; }
        double acc = 0.0;
        const int n_steps = t.jc1 + 3 - t.jstart + 1;        // rows jstart .. jc1+3
        for (int sb = 0; sb < n_steps; sb += 8) {
            sor_step<RECIP, 0>(P, R, sb + 0, t, a.k, lg, acc);
            sor_step<RECIP, 1>(P, R, sb + 1, t, a.k, lg, acc);
            sor_step<RECIP, 2>(P, R, sb + 2, t, a.k, lg, acc);
            sor_step<RECIP, 3>(P, R, sb + 3, t, a.k, lg, acc);
            sor_step<RECIP, 4>(P, R, sb + 4, t, a.k, lg, acc);
            sor_step<RECIP, 5>(P, R, sb + 5, t, a.k, lg, acc);
            sor_step<RECIP, 6>(P, R, sb + 6, t, a.k, lg, acc);
            sor_step<RECIP, 7>(P, R, sb + 7, t, a.k, lg, acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) a.partials[task] = acc;
    }

    // ---- last block closes the iteration
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tk = atomicAdd(&a.ctl->ticket, 1u);
        is_last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        sor_close_iteration(a.partials, a.n_tasks, a.ctl, a.it, a.k.cells, sh);
    }
}

// ------------------------------------------------------------------------------------------
// Planning: strips are fixed (120 columns); pick an even chunk height that fills whole waves.
static int g_attr_done = 0;

void nsb_sor_plan(nsb_ctx* c) {
    const int ni = c->g.ni, nj = c->g.nj;
    if (!g_attr_done) {
        cudaFuncSetAttribute(sor_stream_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES);
        cudaFuncSetAttribute(sor_stream_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES);
        g_attr_done = 1;
    }
    c->n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    int blocks_per_sm = 0;
    if (c->use_recip)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, sor_stream_kernel<true>, SOR_THREADS, SOR_SMEM_BYTES);
    else
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, sor_stream_kernel<false>, SOR_THREADS, SOR_SMEM_BYTES);
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
            double cost = (double)waves * (rows + 6 + 8);      // +8: pipeline fill / drain
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
    if (c->use_recip) sor_stream_kernel<true><<<blocks, SOR_THREADS, SOR_SMEM_BYTES, c->stream>>>(a);
    else sor_stream_kernel<false><<<blocks, SOR_THREADS, SOR_SMEM_BYTES, c->stream>>>(a);
    return cudaGetLastError();
}
