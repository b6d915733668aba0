// sor_stream.cu -- fused red + black + residual SOR iteration as ONE streaming pass over HBM.
//
// Replaces, per iteration, the six launches of src/parallel/main.cu:684-710
// (ghost refresh, colour 0, ghost refresh, colour 1, residual, block-norm reduce) and reproduces
// their order exactly.  Ghost cells are not stored in HBM; they are materialised in
// REGISTERS and refreshed exactly where the reference refreshes them: before colour 0 and before
// colour 1 (main.cu:686/694), not before the residual -- so the residual sees, next to a wall, the
// value the wall cell had before the black sweep, as the reference does.
//
// Design (DESIGN.md "SOR streaming kernel", v3):
//   - ping-pong pressure buffers: read p_in, write p_out, so a tile never sees a half-updated
//     neighbour and both colours are applied in one pass: algorithmic traffic 24 B per cell per
//     iteration (p in 8 + RHS in 8 + p out 8).
//   - one WARP owns a strip of 30 column groups (120 columns) x chunk_rows rows and marches up in
//     j.  Lane L owns 4 adjacent columns (one 32-byte sector); lanes 0 and 31 are halo lanes.
//     x-neighbours come from warp shuffles, y-neighbours from a register window of 5 rows, so
//     there is no block barrier anywhere.
//   - rows are fetched by the TMA engine in (even, odd) row PAIRS: one elected lane issues four
//     `cp.async.bulk` copies (2 x 1 KB of p, 2 x 1 KB of RHS) into a per-warp ring of
//     shared-memory stages, completion on an mbarrier.  The grids carry pad rows (NSB_ROW0,
//     NSB_ROWS_ABOVE), so no load ever needs a bounds test.
//   - software pipeline inside the window: when row j arrives, row j-1 gets its red update, row
//     j-2 its black update (and is stored, two 16-byte stores = one full sector per lane), row
//     j-3 its residual.
//   - the step is compiled in three variants and chosen per block of 8 rows with warp-uniform
//     branches: LEAN (interior strip, interior rows: no predicates, no ghost handling -- the
//     variant > 90 % of the rows run), EY (rows next to the bottom / top wall or to the chunk ends)
//     and EX+EY (first / last strip, which hold the ghost columns).  v2 ran the fully general code
//     everywhere and spent 79 % of its instructions on predication (profiles/r1b_*).
//   - residual: per-warp partial (fixed order), last block of the grid adds the partials in a fixed
//     order, takes sqrt(sum/(i_max*j_max)) (main.cu:588), compares with eps*(norm_p+1.5)
//     (main.cu:713) and raises the device-side converged flag; later launches exit at once.
//   - two kernels share this code: sor_stream_kernel (one launch per iteration: validation, and
//     several slabs emulated on one GPU) and sor_persist_kernel (the product: all iterations of a
//     solve in one launch, tasks claimed from a global counter and ordered by per-task flags; see
//     the comment above it).  Both are instantiated for RECIP (reciprocal product / IEEE division
//     with the refined reciprocal, nsb_math.cuh) and for MULTI (one slab / several): every body
//     that is not needed is left out, because the kernel is co-limited by the instruction cache.
#include <string.h>

#include "nsb_internal.h"
#include "nsb_math.cuh"
#include "slab.cuh"

#ifndef SOR_WARPS
#define SOR_WARPS 4
#endif
#define SOR_THREADS (SOR_WARPS * 32)
#ifndef SOR_MIN_BLOCKS
#define SOR_MIN_BLOCKS 3
#endif
// Steps per unrolled block (= slots of the register window, a compile-time ring) and row-pair stages of the
// per-warp TMA ring.  SOR_BLK must be a multiple of 2 * SOR_NST and larger than SOR_LAG_RES + 1.
#ifndef SOR_BLK
#define SOR_BLK 8
#endif
#ifndef SOR_NST
#define SOR_NST 2
#endif
#define SOR_USES_PER_BLK (SOR_BLK / 2 / SOR_NST)   // times a stage is filled and read per block
static_assert(SOR_BLK % (2 * SOR_NST) == 0, "SOR_BLK must be a multiple of 2 * SOR_NST");
#define SOR_ROW_BYTES 1024     // 32 lanes x 32 B
#define SOR_ROW_DBL (SOR_ROW_BYTES / 8)
#define SOR_STAGE_BYTES (4 * SOR_ROW_BYTES)   // p even row, p odd row, RHS even row, RHS odd row
#define SOR_WARP_SMEM (SOR_NST * SOR_STAGE_BYTES)
#define SOR_SMEM_BYTES (SOR_WARPS * SOR_WARP_SMEM + SOR_WARPS * 2 * SOR_NST * 8)   // rings + FULL/EMPTY mbarriers
// Rows behind the arriving row j at which the black update and the residual run (red is always j-1).
// 2 / 3 is the tightest schedule (each phase uses what the previous one produced in the same step);
// 3 / 5 makes the three phases of a step independent of each other (more instruction-level
// parallelism per warp at the price of two more rows in the register window).
#ifndef SOR_LAG_BLACK
#define SOR_LAG_BLACK 2
#endif
#ifndef SOR_LAG_RES
#define SOR_LAG_RES 3
#endif

struct SorStreamArgs {
    const double* __restrict__ pin;
    double* __restrict__ pout;
    const double* __restrict__ rhs;
    long long pitch;
    int ni, nj;
    int chunk_rows, n_strips, n_tasks, n_chunks;
    int n_groups;            // pitch / 4
    int it, check;
    unsigned int zero;       // always 0; opaque to the compiler
    SorCoef k;
    double* partials;
    double* bandvals;
    SorCtl* ctl;
    // y-slab decomposition (world == 1: unused)
    int world, wall_bottom, wall_top, down_nj;
    unsigned long long seq;                  // SOR iterations executed before this one, all solves
    unsigned long long round;                // reduction round of this iteration
    double* down_pout; double* up_pout;      // the neighbours' output buffers (storage row 0)
    const unsigned long long* flag_from_down; const unsigned long long* flag_from_up;   // in my arena
    unsigned long long* down_flag; unsigned long long* up_flag;                         // in theirs
    int* err;
    SlabRed red;
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
#ifdef SOR_TEST_WAIT
    // non-suspending poll: test_wait returns at once; try_wait may park the warp until a time limit
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
    return;
#endif
#ifdef SOR_WAIT_CPP
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
    return;
#endif
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}"
        ::"r"(bar), "r"(parity) : "memory");
}
// global -> shared bulk copy executed by the TMA engine (UBLKCP in SASS), completes on `bar`
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// One thread: closes iteration `it` with the global sum of r^2.
__device__ __forceinline__ void sor_close_iteration(double total, SorCtl* ctl, int it, long long cells) {
    double res = sqrt(total / (double)cells);              // main.cu:588
    ctl->sum_r2 = total;
    ctl->res = res;
    if (res <= ctl->thresh) ctl->converged = 1;             // main.cu:713
    __threadfence();                                        // `converged` is visible before `iters` says "closed"
    *((volatile int*)&ctl->iters) = it + 1;
    __threadfence();
    sor_publish(ctl);
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
    int rlo, rhi;             // rows that take updates: 1..nj at walls, -1 / nj+2 at a slab boundary
    bool wall_bottom, wall_top;
    bool push_down, push_up;  // this task owns rows of the neighbour's halo
    double* down_col; double* up_col;   // neighbour's pout, this lane's column group, row 0 of its storage
    int down_nj;
    long long pitch;
    double* pout_row;         // &pout[row of the black update of the current step][this lane's group]
    const double* lds_lane;   // generic pointer: this warp's ring + 4 * lane doubles
    uint32_t bar_base;        // this warp's mbarriers (shared address): NST FULL, then NST EMPTY
    int lane;
    uint32_t zero;            // run-time 0 (see the EMPTY-barrier arrive)
    uint32_t full_par, empty_par;   // phase parities the barriers start this task with (persistent sweep: the
                                    // barriers live across tasks; 0 in the one-task-per-warp kernel)
    // producer state (warp-uniform)
    const double* pin_next;   // first row of the next pair to fetch, first column of the strip
    const double* rhs_next;
    int j_next, j_limit;      // pairs whose first row is > j_limit are not fetched
    uint32_t copy_bytes;      // bytes per row copy (1 KB, clipped at the right edge of the pitch)
    uint32_t smem_base;       // this warp's ring (shared address)
};

// One elected lane arms stage `st` and starts the four copies of the next row pair.  Pairs beyond
// the rows this task needs complete their phase without data so that the parities stay in step.
// The producer state is kept warp-uniform (every lane advances it), and the election uses
// elect.sync so that the compiler moves the operands to uniform registers without a lane loop.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .b32 r;\n\t.reg .pred p;\n\t"
        "elect.sync r|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void issue_pair(TaskCtx& t, int st, uint32_t empty_phase, bool wait_empty) {
#ifdef SOR_LANE0_ISSUE
    if (t.lane == 0) {
#else
    if (elect_one()) {
#endif
        const uint32_t bar = t.bar_base + st * 8;
        if (wait_empty) mbar_wait(t.bar_base + (SOR_NST + st) * 8, empty_phase);
        if (t.j_next <= t.j_limit) {
            const uint32_t dst = t.smem_base + st * SOR_STAGE_BYTES;
            mbar_expect_tx(bar, 4u * t.copy_bytes);
            bulk_g2s(dst, t.pin_next, t.copy_bytes, bar);
            bulk_g2s(dst + SOR_ROW_BYTES, t.pin_next + t.pitch, t.copy_bytes, bar);
            bulk_g2s(dst + 2 * SOR_ROW_BYTES, t.rhs_next, t.copy_bytes, bar);
            bulk_g2s(dst + 3 * SOR_ROW_BYTES, t.rhs_next + t.pitch, t.copy_bytes, bar);
        } else {
            mbar_arrive(bar);
        }
    }
    t.j_next += 2;
    t.pin_next += 2 * t.pitch;
    t.rhs_next += 2 * t.pitch;
}

// One pipeline step with compile-time window slot S (row parity == S parity).  sb is the step
// number of the first step of the current block of 8 (a multiple of 8); the row that arrives in
// this step is j = jstart + sb + S.
//   EX: the strip holds ghost / pad columns (first and last strips).
//   EY: the block of 8 steps touches rows that need gating (walls, chunk ends).
//   QM: how quotients are formed (nsb_math.cuh); with NSB_QUOT_TRY `ok` is cleared when a quotient that
//       is USED left the window of the three-operation division (pad / ghost columns of an edge strip may
//       hold anything and are not counted).
template <int QM, bool EX, bool EY, int S>
__device__ __forceinline__ void sor_step(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0, TaskCtx& t,
                                         const SorCoef& kf, const LaneGeom& lg, double& acc, bool& ok) {
    constexpr int LB = SOR_LAG_BLACK, LR = SOR_LAG_RES;
    constexpr int NS = SOR_BLK;                                                    // window slots
    static_assert(LR + 2 <= NS, "register window too small for the residual lag");
    constexpr int sE = S % NS, sD = (S + NS - 1) % NS, sC = (S + NS - 2) % NS;     // rows j, j-1, j-2
    constexpr int sK = (S + 2 * NS - LB) % NS, sKn = (S + 2 * NS + 1 - LB) % NS, sKs = (S + 2 * NS - 1 - LB) % NS;   // black row, north, south
    constexpr int sR = (S + 2 * NS - LR) % NS, sRn = (S + 2 * NS + 1 - LR) % NS, sRs = (S + 2 * NS - 1 - LR) % NS;   // residual row, north, south
    constexpr int st = (S >> 1) % SOR_NST;
    constexpr int k0 = (S & 1) ? 1 : 0;     // red cells of row j-1: columns k0 and k0 + 2
    constexpr int k1 = k0 + 2;
    constexpr int b0 = ((S + LB) & 1) ? 1 : 0;   // black cells of row j-LB
    constexpr int b1 = b0 + 2;
    const unsigned FULL = 0xffffffffu;
    const int j = t.jstart + sb + S;

    // ---- row j (p and RHS) from the ring into registers; after the odd row of a pair the stage
    //      goes back to the TMA engine
    {
        const uint32_t ph = (use0 + (uint32_t)((S >> 1) / SOR_NST)) & 1u;      // use count of the stage, mod 2
        if ((S & 1) == 0) {
            mbar_wait(t.bar_base + st * 8, ph ^ t.full_par);
#ifdef SOR_SYNC_EACH_STEP
            __syncwarp();
#endif
        }
        const double2* q = reinterpret_cast<const double2*>(t.lds_lane + st * (SOR_STAGE_BYTES / 8) + (S & 1) * SOR_ROW_DBL);
        const double2 a = q[0], b = q[1];
        const double2 c = q[SOR_ROW_BYTES / 8], d = q[SOR_ROW_BYTES / 8 + 1];     // + 2 rows = 2 * 1024 B = 128 double2
        P[sE][0] = a.x; P[sE][1] = a.y; P[sE][2] = b.x; P[sE][3] = b.y;
        R[sE][0] = c.x; R[sE][1] = c.y; R[sE][2] = d.x; R[sE][3] = d.y;
        if (S & 1) {
            // Both rows of the pair are read: every lane releases the stage on its EMPTY barrier and
            // the elected lane re-arms the stage only after acquiring that phase (32 arrivals).
            // The arrive must not overtake the LDS above: neither __syncwarp() nor the arrive's
            // release semantics kept the bulk copy from overwriting a row whose LDS was still queued
            // (seen on B200 as sporadically wrong RHS values in the last-loaded 16 bytes).  So the
            // barrier ADDRESS is made data-dependent on all eight LDS of the pair (this step's
            // a, b, c, d and the previous step's, now in window slot D): `zero` is a run-time 0 the
            // compiler cannot fold, and the arrive cannot issue before the loads have landed.
            const int dep = (__double2hiint(a.x) ^ __double2hiint(b.x) ^ __double2hiint(c.x)) ^
                            (__double2hiint(d.x) ^ __double2hiint(P[sD][0]) ^ __double2hiint(P[sD][2])) ^
                            (__double2hiint(R[sD][0]) ^ __double2hiint(R[sD][2]));
            mbar_arrive(t.bar_base + (SOR_NST + st) * 8 + ((uint32_t)dep & t.zero));
            issue_pair(t, st, ph ^ t.empty_par, true);
        }
    }

    // ---- colour 0 (red) on row D = j-1; neighbours C (j-2, black cells still old) and E (j)
    {
        const int r = j - 1;
        const bool row_in = !EY || ((r >= t.rlo) && (r <= t.rhi));
        double (&X)[4] = P[sD];
        if (EX) refresh_ghost_columns(X, lg);                           // main.cu:686
        if (EY) {
            if (r == 1 && t.wall_bottom) {
#pragma unroll
                for (int k = 0; k < 4; k++) P[sC][k] = X[k];            // p[i][0] = p[i][1]
            }
            if (r == t.nj && t.wall_top) {
#pragma unroll
                for (int k = 0; k < 4; k++) P[sE][k] = X[k];            // p[i][j_max+1] = p[i][j_max]
            }
        }
        double xw = 0.0, xe = 0.0;
        if (k0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double w0 = (k0 == 0) ? xw : X[(k0 + 3) & 3];
        const double e0 = X[k0 + 1];
        const double w1 = X[k1 - 1];
        const double e1 = (k1 == 3) ? xe : X[(k1 + 1) & 3];
        bool ok0 = true, ok1 = true;
        const double u0 = sor_cell_q<QM>(X[k0], e0, w0, P[sE][k0], P[sC][k0], R[sD][k0], kf, EX ? ok0 : ok);
        const double u1 = sor_cell_q<QM>(X[k1], e1, w1, P[sE][k1], P[sC][k1], R[sD][k1], kf, EX ? ok1 : ok);
        if (row_in && (!EX || lg.inner[k0])) { X[k0] = u0; if (EX) ok = ok && ok0; }
        if (row_in && (!EX || lg.inner[k1])) { X[k1] = u1; if (EX) ok = ok && ok1; }
    }

    // ---- colour 1 (black) on row K = j-LB; neighbours: south (final) and north (red done, black cells still old)
    {
        const int r = j - LB;
        const bool row_in = !EY || ((r >= t.rlo) && (r <= t.rhi));
        double (&X)[4] = P[sK];
        if (EX) refresh_ghost_columns(X, lg);                           // main.cu:694
        if (EY) {
            if (r == 1 && t.wall_bottom) {
#pragma unroll
                for (int k = 0; k < 4; k++) P[sKs][k] = X[k];
            }
            if (r == t.nj && t.wall_top) {
#pragma unroll
                for (int k = 0; k < 4; k++) P[sKn][k] = X[k];
            }
        }
        double xw = 0.0, xe = 0.0;
        if (b0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
        else xe = __shfl_down_sync(FULL, X[0], 1);
        const double w0 = (b0 == 0) ? xw : X[(b0 + 3) & 3];
        const double e0 = X[b0 + 1];
        const double w1 = X[b1 - 1];
        const double e1 = (b1 == 3) ? xe : X[(b1 + 1) & 3];
        bool ok0 = true, ok1 = true;
        const double u0 = sor_cell_q<QM>(X[b0], e0, w0, P[sKn][b0], P[sKs][b0], R[sK][b0], kf, EX ? ok0 : ok);
        const double u1 = sor_cell_q<QM>(X[b1], e1, w1, P[sKn][b1], P[sKs][b1], R[sK][b1], kf, EX ? ok1 : ok);
        if (row_in && (!EX || lg.inner[b0])) { X[b0] = u0; if (EX) ok = ok && ok0; }
        if (row_in && (!EX || lg.inner[b1])) { X[b1] = u1; if (EX) ok = ok && ok1; }
        if (EX && (!EY || (r >= t.jc0 && r <= t.jc1))) {
            // ghost columns of the final row, as the refresh before colour 1 left them (main.cu:694):
            // the west ghost lives in the halo lane 0 of the first strip, the east ghost in an owner
            // lane (stored below with its group) or in the halo lane 31 of the last strip
            if (lg.ghostW) t.pout_row[3] = X[3];
            if (lg.kGhostE >= 0 && !lg.owner && t.lane == 31) {
                if (lg.kGhostE == 0) t.pout_row[0] = X[0];
                else if (lg.kGhostE == 1) t.pout_row[1] = X[1];
                else if (lg.kGhostE == 2) t.pout_row[2] = X[2];
                else t.pout_row[3] = X[3];
            }
        }
        // the row is final: store it (one full 32-byte sector per lane)
        if (lg.owner && (!EY || (r >= t.jc0 && r <= t.jc1))) {
            double2* q = reinterpret_cast<double2*>(t.pout_row);
            q[0] = make_double2(X[0], X[1]);
            q[1] = make_double2(X[2], X[3]);
            if (EY) {
                // ghost rows j = 0 / j_max+1 as the same refresh left them (copies of row 1 / j_max taken
                // before the black update, held in the window slots B / D)
                if (r == 1 && t.wall_bottom) {
                    double2* h = reinterpret_cast<double2*>(t.pout_row - t.pitch);
                    h[0] = make_double2(P[sKs][0], P[sKs][1]);
                    h[1] = make_double2(P[sKs][2], P[sKs][3]);
                }
                if (r == t.nj && t.wall_top) {
                    double2* h = reinterpret_cast<double2*>(t.pout_row + t.pitch);
                    h[0] = make_double2(P[sKn][0], P[sKn][1]);
                    h[1] = make_double2(P[sKn][2], P[sKn][3]);
                }
                // y-slab halo: the three rows next to a slab boundary also go straight into the
                // neighbour's halo rows over NVLink (peer store), SURVEY 8e
                if (t.push_down && r <= 3) {
                    double2* h = reinterpret_cast<double2*>(t.down_col + (long long)(t.down_nj + r + NSB_ROW0) * t.pitch);
                    h[0] = make_double2(X[0], X[1]);
                    h[1] = make_double2(X[2], X[3]);
                }
                if (t.push_up && r >= t.nj - 2) {
                    double2* h = reinterpret_cast<double2*>(t.up_col + (long long)(r - t.nj + NSB_ROW0) * t.pitch);
                    h[0] = make_double2(X[0], X[1]);
                    h[1] = make_double2(X[2], X[3]);
                }
            }
        }
        t.pout_row += t.pitch;
    }

    // ---- residual on row j-LR (final, as are its neighbours) with the ghosts as they are (last refreshed before colour 1)
    {
        const int r = j - LR;
        double (&X)[4] = P[sR];
        const double xw = __shfl_up_sync(FULL, X[3], 1);
        const double xe = __shfl_down_sync(FULL, X[0], 1);
        if (!EY || (r >= t.jc0 && r <= t.jc1)) {   // warp-uniform; rows of the chunk are interior rows
            double part = 0.0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const double w = (k == 0) ? xw : X[(k + 3) & 3];
                const double e = (k == 3) ? xe : X[(k + 1) & 3];
                bool okr = true;
                const double q = sor_residual_q<QM>(X[k], e, w, P[sRn][k], P[sRs][k], R[sR][k], kf, EX ? okr : ok);
                if (!EX || lg.inner[k]) { part = fma(q, q, part); if (EX) ok = ok && okr; }
            }
            acc += part;                           // non-owner lanes are discarded at the end of the task
        }
    }
}

template <int QM, bool EX, bool EY, int S>
__device__ __forceinline__ void sor_steps_from(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0,
                                               TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double& acc, bool& ok) {
    if constexpr (S < SOR_BLK) {
        sor_step<QM, EX, EY, S>(P, R, sb, use0, t, kf, lg, acc, ok);
        sor_steps_from<QM, EX, EY, S + 1>(P, R, sb, use0, t, kf, lg, acc, ok);
    }
}
// One unrolled block of SOR_BLK steps; use0 = times every ring stage has been used before this block.
template <int QM, bool EX, bool EY>
__device__ __forceinline__ void sor_block8(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0,
                                           TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double& acc, bool& ok) {
    sor_steps_from<QM, EX, EY, 0>(P, R, sb, use0, t, kf, lg, acc, ok);
}

// Steps of a task: rows jstart .. jc1 + LR arrive (the last ones only to drain the pipeline: the residual of
// row jc1 runs LR steps after that row arrived; rows above jc1 + 3 are not fetched).
__device__ __forceinline__ int sor_task_steps(const TaskCtx& t) { return t.jc1 + SOR_LAG_RES - t.jstart + 1; }

// One task = one strip x one chunk of rows, marched bottom to top by one warp.
template <int QM, bool EX>
__device__ __forceinline__ double sor_run_task(TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, bool& ok) {
    double P[SOR_BLK][4], R[SOR_BLK][4];
#pragma unroll
    for (int s = 0; s < SOR_BLK; s++)
#pragma unroll
        for (int k = 0; k < 4; k++) { P[s][k] = 0.0; R[s][k] = 0.0; }
    double acc = 0.0;
    const int n_steps = sor_task_steps(t);
    uint32_t use0 = 0u;
    for (int sb = 0; sb < n_steps; sb += SOR_BLK, use0 += SOR_USES_PER_BLK) {
        const int jb = t.jstart + sb;                        // row arriving in the first step of the block
        // rows touched by the block (B = SOR_BLK): residual jb-LR .. jb+B-1-LR, stores jb-LB .. jb+B-1-LB, loads jb .. jb+B-1
#ifdef SOR_NO_LEAN
        const bool clean = false;
#else
        const bool clean = (jb - SOR_LAG_RES >= t.jc0) && (jb + SOR_BLK - 1 - SOR_LAG_BLACK <= t.jc1) && (jb + SOR_BLK - 1 <= t.nj) &&
                           !(t.push_down && jb - SOR_LAG_BLACK <= 3) && !(t.push_up && jb + SOR_BLK - 1 - SOR_LAG_BLACK >= t.nj - 2);
#endif
        // (a fourth variant <EX, !EY> for the interior rows of the two edge strips was measured: the larger
        //  kernel body costs the LEAN path 10 % -- 5.27 against 5.86 TB/s at 8192^2 -- and was dropped)
        if (EX) sor_block8<QM, true, true>(P, R, sb, use0, t, kf, lg, acc, ok);
        else if (clean && QM != NSB_QUOT_FULL) sor_block8<QM, false, false>(P, R, sb, use0, t, kf, lg, acc, ok);
        else sor_block8<QM, false, true>(P, R, sb, use0, t, kf, lg, acc, ok);
    }
    return lg.owner ? acc : 0.0;
}

// Phase parities of a warp's mbarriers after one task of n_steps rows (see sor_persist_kernel).
__device__ __forceinline__ void sor_advance_parities(int n_steps, uint32_t& full_par, uint32_t& empty_par) {
    const uint32_t pairs_per_stage = (uint32_t)(((n_steps + SOR_BLK - 1) / SOR_BLK) * SOR_USES_PER_BLK);
    full_par ^= (pairs_per_stage + 1u) & 1u;
    empty_par ^= pairs_per_stage & 1u;
}

// A whole task: the first row pairs are requested, the strip is marched.  With the IEEE division the
// first pass uses the three-operation quotients; if any lane saw an operand outside their window
// (non-finite or denormal-range data) the SAME task runs again with out-of-line divisions -- its
// outputs (rows, halo rows, partial) are simply written twice, nothing is published in between.
// `row0` = &array[row jstart][first column of the strip] offsets, as set up by the caller.
template <bool RECIP>
__device__ __forceinline__ double sor_task(TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, bool edge_strip,
                                           uint32_t& full_par, uint32_t& empty_par) {
    constexpr int QM = RECIP ? NSB_QUOT_RECIP : NSB_QUOT_TRY;
    const int n_steps = sor_task_steps(t);
    const double* pin0 = t.pin_next; const double* rhs0 = t.rhs_next; double* pout0 = t.pout_row;
    t.full_par = full_par; t.empty_par = empty_par;
    sor_advance_parities(n_steps, full_par, empty_par);
    __syncwarp();
#pragma unroll
    for (int q = 0; q < SOR_NST; q++) issue_pair(t, q, 0u, false);
    bool ok = true;
    double acc;
    if (edge_strip) acc = sor_run_task<QM, true>(t, kf, lg, ok);
    else acc = sor_run_task<QM, false>(t, kf, lg, ok);
    if (!RECIP) {
        if (!__all_sync(0xffffffffu, ok)) {
            t.pin_next = pin0; t.rhs_next = rhs0; t.pout_row = pout0; t.j_next = t.jstart;
            t.full_par = full_par; t.empty_par = empty_par;
            sor_advance_parities(n_steps, full_par, empty_par);
            __syncwarp();
#pragma unroll
            for (int q = 0; q < SOR_NST; q++) issue_pair(t, q, 0u, false);
            if (edge_strip) acc = sor_run_task<NSB_QUOT_FULL, true>(t, kf, lg, ok);
            else acc = sor_run_task<NSB_QUOT_FULL, false>(t, kf, lg, ok);
        }
    }
    return acc;
}

#ifdef SOR_MAXNREG
#define SOR_KERNEL_ATTR __maxnreg__(SOR_MAXNREG)
#else
#define SOR_KERNEL_ATTR __launch_bounds__(SOR_THREADS, SOR_MIN_BLOCKS)
#endif
// MULTI: several y-slabs (world > 1).  The single-slab instantiation carries no halo pushes, peer flags or slab
// reductions: a smaller body, which the instruction cache rewards (8192^2 265 -> 257 us, 2048^2 45.7 -> 39.6 us).
template <bool RECIP, bool MULTI>
__global__ void SOR_KERNEL_ATTR sor_stream_kernel(const SorStreamArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ double sh[32];
    __shared__ int is_last;
    if (a.check && *((volatile int*)&a.ctl->converged)) return;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int task = blockIdx.x * SOR_WARPS + warp;
    if (task < a.n_tasks) {
        const int strip = task % a.n_strips;
        int chunk = task / a.n_strips;
        if ((MULTI && (a.world > 1)) && a.n_chunks > 1) {
            // slab boundaries first: their rows reach the neighbours a whole launch before they are read
            chunk = (chunk == 0) ? 0 : (chunk == 1) ? a.n_chunks - 1 : chunk - 1;
        }
        const int g0 = strip * NSB_STRIP_GROUPS;             // first column group of the strip
        const int g = g0 + lane;                             // column group of this lane
        const int i0 = g * NSB_GROUP - NSB_COL0;             // i of column k = 0
        LaneGeom lg;
        bool all_inner = (g < a.n_groups);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            lg.inner[k] = (i0 + k >= 1) && (i0 + k <= a.ni);
            all_inner = all_inner && lg.inner[k];
        }
        lg.owner = (g < a.n_groups) && lane >= 1 && lane <= NSB_STRIP_GROUPS;
        lg.ghostW = (i0 + 3 == 0);
        const int kg = a.ni + 1 - i0;
        lg.kGhostE = (kg >= 0 && kg <= 3) ? kg : -1;
        lg.warpW = __any_sync(0xffffffffu, lg.ghostW);
        lg.warpE = __any_sync(0xffffffffu, lg.kGhostE >= 0);
#ifdef SOR_FORCE_EX
        const bool edge_strip = true;
#else
        const bool edge_strip = !__all_sync(0xffffffffu, all_inner);
#endif

        TaskCtx t;
        t.jc0 = 1 + chunk * a.chunk_rows;
        t.jc1 = min(t.jc0 + a.chunk_rows - 1, a.nj);
        t.jstart = t.jc0 - 3;                                // even, because chunk_rows is even
        t.nj = a.nj;
        t.wall_bottom = !(MULTI && (a.wall_bottom == 0)); t.wall_top = !(MULTI && (a.wall_top == 0));
        t.rlo = t.wall_bottom ? 1 : -1;
        t.rhi = t.wall_top ? a.nj : a.nj + 2;
        t.push_down = (MULTI && (!t.wall_bottom && t.jc0 <= 3));
        t.push_up = (MULTI && (!t.wall_top && t.jc1 >= a.nj - 2));
        t.pitch = a.pitch;
        t.lane = lane;
        t.zero = a.zero;
        t.full_par = 0u; t.empty_par = 0u;
        const long long coff = (long long)g0 * NSB_GROUP;    // first column of the strip
        const long long row_first = (long long)(t.jstart + NSB_ROW0) * a.pitch;
        t.pin_next = a.pin + row_first + coff;
        t.rhs_next = a.rhs + row_first + coff;
        t.j_next = t.jstart;
        t.j_limit = min(t.jc1 + 3, a.nj + 3);
        t.pout_row = a.pout + row_first - SOR_LAG_BLACK * a.pitch + coff + (long long)lane * NSB_GROUP;
        t.down_col = a.down_pout + coff + (long long)lane * NSB_GROUP;
        t.up_col = a.up_pout + coff + (long long)lane * NSB_GROUP;
        t.down_nj = a.down_nj;
        const int groups_here = min(32, a.n_groups - g0);
        t.copy_bytes = (uint32_t)groups_here * 32u;
        unsigned char* ring = smem_raw + warp * SOR_WARP_SMEM;
        t.lds_lane = reinterpret_cast<const double*>(ring) + lane * NSB_GROUP;
        t.smem_base = smem_u32(ring);
        t.bar_base = smem_u32(smem_raw + SOR_WARPS * SOR_WARP_SMEM + warp * 2 * SOR_NST * 8);

        // ---- y-slab: the halo rows this task reads were stored by the neighbour's launch seq-1
        if ((MULTI && (a.world > 1)) && a.it > 0) {    // it == 0: the halo came with the p exchange at the solve entry
            const bool need_down = !t.wall_bottom && t.jc0 - 3 <= 0;
            const bool need_up = !t.wall_top && t.jc1 + 3 >= a.nj + 1;
            if (lane < 3) {
                const int sidx = strip - 1 + lane;           // the columns read belong to strips s-1, s, s+1
                if (sidx >= 0 && sidx < a.n_strips) {
                    if (need_down) spin_until_ge(a.flag_from_down + sidx, a.seq, a.err);
                    if (need_up) spin_until_ge(a.flag_from_up + sidx, a.seq, a.err);
                }
            }
            __syncwarp();
            asm volatile("fence.proxy.async;" ::: "memory");   // peer stores (generic) -> TMA reads (async)
        }

        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < SOR_NST; s++) {
                mbar_init(t.bar_base + s * 8, 1);                    // FULL: the producer's expect_tx
                mbar_init(t.bar_base + (SOR_NST + s) * 8, 32);       // EMPTY: every lane has read the stage
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        uint32_t full_par = 0u, empty_par = 0u;
        double acc = sor_task<RECIP>(t, a.k, lg, edge_strip, full_par, empty_par);
        acc = warp_sum(acc);
        if (lane == 0) a.partials[chunk * a.n_strips + strip] = acc;

        if (t.push_down || t.push_up) {                      // halo rows are out: tell the neighbour(s)
            __threadfence_system();
            __syncwarp();
            if (lane == 0) {
                if (t.push_down) st_release_sys(a.down_flag + strip, a.seq + 1ull);
                if (t.push_up) st_release_sys(a.up_flag + strip, a.seq + 1ull);
            }
        }
    }

    // ---- last block: band sums (one band = one chunk of rows, strips added in a fixed order),
    //      contributed to the collective reduction; with one slab it also closes the iteration
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tk = atomicAdd(&a.ctl->ticket, 1u);
        is_last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        for (int b = warp; b < a.n_chunks; b += SOR_WARPS) {
            const double v = warp_ordered_sum(a.partials + (size_t)b * a.n_strips, a.n_strips, lane);
            if (lane == 0) a.bandvals[b] = v;
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) a.ctl->ticket = 0;
        slab_red_contribute(a.red, a.round, 1, a.bandvals);
        if (!MULTI) {
            slab_red_finish(a.red, a.round, 1, 0u, sh);
            if (threadIdx.x == 0) sor_close_iteration(sh[0], a.ctl, a.it, a.k.cells);
        }
    }
}

// ------------------------------------------------------------------------------------------
// Persistent multi-iteration sweep (single slab): ALL iterations of a solve in one launch.
//
// Why: one launch per iteration costs a pipeline drain per iteration -- a task (one warp marching up
// 126 rows) lives ~100-150 us, so the last round of blocks leaves most SMs idle; measured launch time
// = 75 us + 4.03 us per 120-row band at 8192 columns (profiles/r1d_wave_probe.txt), i.e. 22 % of a
// launch at 8192^2 is drain.  Here the warps of a fixed grid claim tasks from one global counter,
// task id = iteration * n_tasks + (chunk * n_strips + strip), and a task of iteration n starts as
// soon as its 3 x 3 neighbourhood of tasks of iteration n-1 is done (their rows are what it reads,
// and they are the only readers of the rows it overwrites in the ping-pong buffer), so the sweep
// of iteration n+1 follows the sweep of iteration n up the grid without a global barrier.
//   - deadlock freedom: ids are claimed in order and a task only waits for smaller ids, which are
//     running on resident warps or finished; every wait still has a deadline (slab.cuh);
//   - residual: per-task partials (double-buffered by iteration parity), the warp that takes the
//     last ticket of an iteration sums them in the same fixed order as the one-launch-per-iteration
//     kernel (same bits), closes the iteration and publishes the control block;
//   - convergence: a task of iteration n waits until iteration n-2 is closed and stops if that or an
//     earlier one converged.  Tasks of iteration m+1 may run after iteration m converged: they write
//     the OTHER ping-pong buffer, the converged iterate stays intact (same contract as the queued
//     launches of the per-iteration kernel).
struct SorPersistArgs {
    SorStreamArgs s;                 // geometry, coefficients, ctl; pin / pout = buffers of iteration 0
    int n_iter;
    unsigned long long* claim;       // [0] task counter, [1..4] tickets of iterations mod 4
    unsigned long long* done;        // [n_tasks] epoch of the last iteration the task finished
    unsigned long long flag_base;    // epoch of "before iteration 0 of this launch"
    double* partials2;               // [2][n_tasks]
    // y-slabs: this rank's two pressure buffers and the neighbours' (index = buffer number), the buffer
    // iteration 0 reads, and the reduction round of iteration 0
    double* own_p[2]; double* down_p[2]; double* up_p[2];
    int buf0;
    unsigned long long round0;
};

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_gpu_s32(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// one lane: wait until *flag >= want (device scope); false on timeout
__device__ __forceinline__ bool spin_gpu_ge(const unsigned long long* flag, unsigned long long want, int* err) {
    if (ld_acquire_gpu(flag) >= want) return true;
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_gpu(flag) < want) {
        if (*((volatile int*)err)) return false;
        if (global_timer_ns() - t0 > NSB_SPIN_TIMEOUT_NS) { *((volatile int*)err) = 1; return false; }
        __nanosleep(64);
    }
    return true;
}

template <bool RECIP, bool MULTI>
__global__ void SOR_KERNEL_ATTR sor_persist_kernel(const SorPersistArgs pa) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const SorStreamArgs& a = pa.s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned FULLM = 0xffffffffu;
    const unsigned long long total = (unsigned long long)pa.n_iter * (unsigned long long)a.n_tasks;
    unsigned char* ring = smem_raw + warp * SOR_WARP_SMEM;
    const uint32_t bar_base = smem_u32(smem_raw + SOR_WARPS * SOR_WARP_SMEM + warp * 2 * SOR_NST * 8);
    // The mbarriers are initialised once and live across the tasks of this warp.  Per task and stage
    // the FULL barrier completes (pairs per stage + 1) phases (the last re-arm is never waited for, it
    // is a plain arrive), the EMPTY barrier (pairs per stage): the parities a task starts with follow.
    uint32_t full_par = 0u, empty_par = 0u;
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < 2 * SOR_NST; q++) mbar_init(bar_base + q * 8, q < SOR_NST ? 1u : 32u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    for (;;) {
        unsigned long long id = 0;
        if (lane == 0) id = atomicAdd(pa.claim, 1ull);
        id = __shfl_sync(FULLM, id, 0);
        if (id >= total) break;
        const int it = (int)(id / (unsigned long long)a.n_tasks);
        const int task = (int)(id - (unsigned long long)it * (unsigned long long)a.n_tasks);
        const int strip = task % a.n_strips;
        int chunk = task / a.n_strips;
        if ((MULTI && (a.world > 1)) && a.n_chunks > 1)     // slab boundaries first: their rows travel while the interior is swept
            chunk = (chunk == 0) ? 0 : (chunk == 1) ? a.n_chunks - 1 : chunk - 1;
        const int tslot = chunk * a.n_strips + strip;      // index of this task's flag and partial

        // ---- may this iteration run?  iteration it-2 must be closed, and nothing before may have converged
        // (the wait also keeps the double-buffered partials and the 4 tickets of iteration it-2 / it-4 safe,
        //  so it does not depend on `check`)
        int stop = 0;
        {
            if (lane == 0) {
                // {iters, converged} are adjacent: ONE 64-bit acquire load per task reads both (two dependent
                // L2 round trips per task were ~2 % of the sweep with the convergence test on)
                const unsigned long long* word = reinterpret_cast<const unsigned long long*>(&a.ctl->iters);
                unsigned long long w = ld_acquire_gpu(word);
                if (it >= 2) {
                    const unsigned long long t0 = global_timer_ns();
                    while ((int)(unsigned)(w & 0xffffffffull) < a.it + it - 1) {
                        if (a.check && (unsigned)(w >> 32)) break;                      // nothing closes after convergence
                        if (*((volatile int*)a.err) || global_timer_ns() - t0 > NSB_SPIN_TIMEOUT_NS) { *((volatile int*)a.err) = 1; stop = 1; break; }
                        __nanosleep(64);
                        w = ld_acquire_gpu(word);
                    }
                }
                if (a.check && (unsigned)(w >> 32)) stop = 1;
            }
            stop = __shfl_sync(FULLM, stop, 0);
        }
        if (stop) break;

        // ---- wait for the 3 x 3 neighbourhood of the previous iteration
        if (it > 0) {
            int ok = 1;
            if (lane < 9) {
                const int cc = chunk + lane / 3 - 1, ss = strip + lane % 3 - 1;
                if (cc >= 0 && cc < a.n_chunks && ss >= 0 && ss < a.n_strips)
                    ok = spin_gpu_ge(pa.done + (size_t)cc * a.n_strips + ss, pa.flag_base + (unsigned long long)it, a.err) ? 1 : 0;
            }
            ok = __all_sync(FULLM, ok);
            if (!ok) break;
            asm volatile("fence.proxy.async;" ::: "memory");   // other warps' generic stores -> this warp's TMA reads
        }

        const int bo = (pa.buf0 + it + 1) & 1;              // buffer this iteration writes
        const double* pin = pa.own_p[bo ^ 1];
        double* pout = pa.own_p[bo];

        const int g0 = strip * NSB_STRIP_GROUPS;
        const int g = g0 + lane;
        const int i0 = g * NSB_GROUP - NSB_COL0;
        LaneGeom lg;
        bool all_inner = (g < a.n_groups);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            lg.inner[k] = (i0 + k >= 1) && (i0 + k <= a.ni);
            all_inner = all_inner && lg.inner[k];
        }
        lg.owner = (g < a.n_groups) && lane >= 1 && lane <= NSB_STRIP_GROUPS;
        lg.ghostW = (i0 + 3 == 0);
        const int kg = a.ni + 1 - i0;
        lg.kGhostE = (kg >= 0 && kg <= 3) ? kg : -1;
        lg.warpW = __any_sync(FULLM, lg.ghostW);
        lg.warpE = __any_sync(FULLM, lg.kGhostE >= 0);
        const bool edge_strip = !__all_sync(FULLM, all_inner);

        TaskCtx t;
        t.jc0 = 1 + chunk * a.chunk_rows;
        t.jc1 = min(t.jc0 + a.chunk_rows - 1, a.nj);
        t.jstart = t.jc0 - 3;
        t.nj = a.nj;
        t.wall_bottom = !(MULTI && (a.wall_bottom == 0)); t.wall_top = !(MULTI && (a.wall_top == 0));
        t.rlo = t.wall_bottom ? 1 : -1;
        t.rhi = t.wall_top ? a.nj : a.nj + 2;
        t.push_down = (MULTI && (!t.wall_bottom && t.jc0 <= 3));
        t.push_up = (MULTI && (!t.wall_top && t.jc1 >= a.nj - 2));
        t.pitch = a.pitch;
        t.lane = lane;
        t.zero = a.zero;
        const long long coff = (long long)g0 * NSB_GROUP;
        const long long row_first = (long long)(t.jstart + NSB_ROW0) * a.pitch;
        t.pin_next = pin + row_first + coff;
        t.rhs_next = a.rhs + row_first + coff;
        t.j_next = t.jstart;
        t.j_limit = min(t.jc1 + 3, a.nj + 3);
        t.pout_row = pout + row_first - SOR_LAG_BLACK * a.pitch + coff + (long long)lane * NSB_GROUP;
        t.down_col = pa.down_p[bo] + coff + (long long)lane * NSB_GROUP;
        t.up_col = pa.up_p[bo] + coff + (long long)lane * NSB_GROUP;
        t.down_nj = a.down_nj;
        const int groups_here = min(32, a.n_groups - g0);
        t.copy_bytes = (uint32_t)groups_here * 32u;
        t.lds_lane = reinterpret_cast<const double*>(ring) + lane * NSB_GROUP;
        t.smem_base = smem_u32(ring);
        t.bar_base = bar_base;

        // ---- y-slab: rows of the previous iteration pushed by the neighbouring ranks (their boundary
        //      tasks of iteration it-1, strips s-1 .. s+1); a.seq = halo epoch of "before iteration 0"
        if ((MULTI && (a.world > 1)) && it > 0) {
            const bool need_down = !t.wall_bottom && t.jc0 - 3 <= 0;
            const bool need_up = !t.wall_top && t.jc1 + 3 >= a.nj + 1;
            int ok2 = 1;
            if (lane < 3) {
                const int sidx = strip - 1 + lane;
                if (sidx >= 0 && sidx < a.n_strips) {
                    if (need_down) ok2 = spin_until_ge(a.flag_from_down + sidx, a.seq + (unsigned long long)it, a.err) ? 1 : 0;
                    if (need_up && ok2) ok2 = spin_until_ge(a.flag_from_up + sidx, a.seq + (unsigned long long)it, a.err) ? 1 : 0;
                }
            }
            ok2 = __all_sync(FULLM, ok2);
            if (!ok2) break;
            asm volatile("fence.proxy.async;" ::: "memory");
        }

        double acc = sor_task<RECIP>(t, a.k, lg, edge_strip, full_par, empty_par);
        acc = warp_sum(acc);
        double* part = pa.partials2 + (size_t)(it & 1) * a.n_tasks;
        if (lane == 0) part[tslot] = acc;

        // ---- publish: rows and partial are out
        __threadfence();
        __syncwarp();
        unsigned long long tk = 0;
        if (t.push_down || t.push_up) {                      // halo rows are out: tell the neighbouring rank(s)
            __threadfence_system();
            __syncwarp();
            if (lane == 0) {
                if (t.push_down) st_release_sys(a.down_flag + strip, a.seq + (unsigned long long)it + 1ull);
                if (t.push_up) st_release_sys(a.up_flag + strip, a.seq + (unsigned long long)it + 1ull);
            }
        }
        if (lane == 0) {
            st_release_gpu(pa.done + tslot, pa.flag_base + (unsigned long long)it + 1ull);
            tk = atomicAdd(pa.claim + 1 + (it & 3), 1ull);
        }
        tk = __shfl_sync(FULLM, tk, 0);
        if (tk == (unsigned long long)a.n_tasks - 1ull) {
            // this warp closes iteration `it`: band sums in the fixed order of the per-iteration kernel
            __threadfence();
            if (lane == 0) pa.claim[1 + (it & 3)] = 0ull;
            // 16 bands at a time, so that 16 loads per lane are in flight: one band after the other
            // (an L2 round trip each) took ~0.5 us per band, 40 us for the 86 bands of 2048^2 -- as long as
            // the iteration itself, and iteration n+2 waits for this closing (ncu source page, 2048^2:
            // half of all warp samples sat in the `iters` poll above).  Per band the additions are those
            // of warp_ordered_sum (lane-strided partials in q order, then the butterfly): same bits.
            for (int b0 = 0; b0 < a.n_chunks; b0 += 16) {
                double sb[16];
#pragma unroll
                for (int u = 0; u < 16; u++) sb[u] = 0.0;
                for (int q = lane; q < a.n_strips; q += 32) {
#pragma unroll
                    for (int u = 0; u < 16; u++)
                        if (b0 + u < a.n_chunks) sb[u] += __ldcg(part + (size_t)(b0 + u) * a.n_strips + q);
                }
#pragma unroll
                for (int u = 0; u < 16; u++) sb[u] = warp_sum(sb[u]);
                if (lane == 0) {
#pragma unroll
                    for (int u = 0; u < 16; u++)
                        if (b0 + u < a.n_chunks) a.bandvals[b0 + u] = sb[u];
                }
            }
            __threadfence();
            __syncwarp();
            // iterations close in order; nothing closes (and no reduction round is spent) after the one
            // that converged -- every rank sees the same residual bits, so every rank decides the same
            int skip = 0;
            if (lane == 0) {
                const unsigned long long t0 = global_timer_ns();
                while (ld_acquire_gpu_s32(&a.ctl->iters) < a.it + it) {
                    if (a.check && ld_acquire_gpu_s32(&a.ctl->converged)) break;
                    if (global_timer_ns() - t0 > NSB_SPIN_TIMEOUT_NS) { *((volatile int*)a.err) = 1; skip = 1; break; }
                    __nanosleep(64);
                }
                if (a.check && ld_acquire_gpu_s32(&a.ctl->converged)) skip = 1;
            }
            skip = __shfl_sync(FULLM, skip, 0);
            if (!skip) {
                double tot;
                if ((MULTI && (a.world > 1))) {
                    const unsigned long long round = pa.round0 + (unsigned long long)it;
                    slab_red_contribute_warp(a.red, round, 1, a.bandvals, lane);
                    slab_red_finish_sum_warp(a.red, round, lane, &tot);
                } else {
                    tot = warp_ordered_sum(a.bandvals, a.n_chunks, lane);
                }
                if (lane == 0) sor_close_iteration(tot, a.ctl, a.it + it, a.k.cells);
            }
            __syncwarp();
        }
    }
}

typedef void (*SorStreamFn)(const SorStreamArgs);
typedef void (*SorPersistFn)(const SorPersistArgs);
static SorStreamFn sor_stream_fn(bool recip, bool multi) {
    return recip ? (multi ? sor_stream_kernel<true, true> : sor_stream_kernel<true, false>)
                 : (multi ? sor_stream_kernel<false, true> : sor_stream_kernel<false, false>);
}
static SorPersistFn sor_persist_fn(bool recip, bool multi) {
    return recip ? (multi ? sor_persist_kernel<true, true> : sor_persist_kernel<true, false>)
                 : (multi ? sor_persist_kernel<false, true> : sor_persist_kernel<false, false>);
}

// y-slabs: only when every rank sits on a GPU of its own -- the kernels of neighbouring ranks wait
// for each other, so they must be able to run at the same time (a 1-GPU box emulates the ranks with
// the one-launch-per-iteration kernel in lock step instead).
int nsb_sor_persist_supported(const nsb_ctx* c) {
    if (c->chunk_rows < 4) return 0;
    return c->world == 1 || (c->connected && c->peers_distinct);
}

// world > 1: the closing of an iteration is its own one-block kernel, so that the sweep kernel never
// waits for another GPU (and so that the ranks can be emulated in lock step on one device).
__global__ void __launch_bounds__(128) sor_close_kernel(SlabRed red, unsigned long long round, SorCtl* ctl, int it,
                                                        long long cells, int check) {
    __shared__ double out[4];
    if (check && *((volatile int*)&ctl->converged)) return;
    slab_red_finish(red, round, 1, 0u, out);
    if (threadIdx.x == 0) sor_close_iteration(out[0], ctl, it, cells);
}

cudaError_t nsb_launch_sor_close(nsb_ctx* c, int it, int check) {
    sor_close_kernel<<<1, 128, 0, c->stream>>>(nsb_slab_red(c), c->red_round + (unsigned long long)it, c->d_ctl, it,
                                               (long long)c->g.ni * (long long)c->J, check);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------------------------------
// Planning: strips are fixed (120 columns); pick an even chunk height that fills whole waves.

// Chunk height of the persistent sweep (pure host arithmetic, also exported for the CPU tests):
// `slots` = resident warps of the sweep kernel on the device, recip = reciprocal-product arithmetic.
extern "C" int nsb_plan_chunk_rows(int ni, int nj, long long slots, int recip) {
    const long long n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    if (slots < 1) slots = 1;
    int best = 16;
    double best_cost = 1e300;
    // Chunk height of the persistent sweep, 16 .. 128 rows: minimise
    //     max(tasks per resident warp and iteration, floor) x (rows + 18),   floor = 1.6
    // -- a task costs its rows plus ~18 row-steps of overhead (3 + 3 halo rows, 8 steps of pipeline
    // fill, claim / flag waits), and below ~1.6 tasks per warp there is nothing to overlap the
    // dependency waits between two iterations with (4096^2: 50 chunks of 82 rows = one task per warp
    // 110 us per iteration, 86 chunks of 48 rows 90 us).  Above 128 rows the tail of a launch and the
    // balance between SMs suffer (8192^2: 240 rows 298 us, 128 rows 265 us).  Measured optima
    // (profiles/r1h_chunk_sweep2.txt): 2048^2 16-24, 3072^2 32, 4096^2 48, 5120^2 48-80, 6144^2 64-112,
    // 7168^2 96-104, 8192^2 120-128; the model picks 16, 28, 50, 78, 112, 128, 128.
    // With the IEEE division the sweep is instruction bound instead of HBM bound and the balance between
    // warps counts for more: the measured optima sit at ~2.7 tasks per warp (4096^2 32-48 rows, 5120^2 48,
    // 6144^2 64-80, 7168^2 90, 8191^2 112; profiles/r1h_chunk_sweep2.txt), hence the higher floor.
    const double floor_tasks = recip ? 1.6 : 2.6;
    for (int rows = 16; rows <= 128; rows += 2) {
        const long long chunks = (nj + rows - 1) / rows;
        const double per_warp = (double)(chunks * n_strips) / (double)slots;
        const double cost = (per_warp > floor_tasks ? per_warp : floor_tasks) * (rows + 18);
        if (cost < best_cost - 1e-9) { best_cost = cost; best = rows; }
        if (rows >= nj) break;
    }
    return best;
}

void nsb_sor_plan(nsb_ctx* c) {
    const int ni = c->g.ni, nj = c->g.nj;
    // function attributes are per device: set them for the device of every context that is planned
    for (int v = 0; v < 4; v++) {
        cudaFuncSetAttribute((const void*)sor_stream_fn(v & 1, v & 2), cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES);
        cudaFuncSetAttribute((const void*)sor_persist_fn(v & 1, v & 2), cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES);
    }
    c->n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    int blocks_per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, (const void*)sor_stream_fn(c->use_recip, c->world > 1), SOR_THREADS,
                                                  SOR_SMEM_BYTES);
    if (blocks_per_sm < 1) blocks_per_sm = 1;
    const long long slots = (long long)c->sm_count * blocks_per_sm * SOR_WARPS;
    int best = 0;
    if (c->O.sor_chunk_rows > 0) {
        best = (c->O.sor_chunk_rows + 1) & ~1;
    } else {
        best = nsb_plan_chunk_rows(ni, nj, slots, c->use_recip ? 1 : 0);
    }
    if (best < 2) best = 2;
    c->chunk_rows = best;
    c->n_chunks = (nj + best - 1) / best;
    c->n_partials = c->n_chunks * c->n_strips;
}

SlabRed nsb_slab_red(const nsb_ctx* c) {
    SlabRed r;
    memset(&r, 0, sizeof r);
    r.rank = c->rank; r.world = c->world; r.band0 = c->band0; r.nb_local = c->n_chunks; r.nb_global = c->nbands_global;
    for (int k = 0; k < c->world; k++) { r.sync[k] = c->peer[k].sync; r.gband[k] = c->peer[k].gband; }
    return r;
}

// `it` = iteration number inside the current solve.  pin / pout are pbuf[x] / pbuf[x ^ 1] of this
// rank; the neighbours' output buffer of the same parity receives the halo rows.
cudaError_t nsb_launch_sor_stream_iter(nsb_ctx* c, int it, const double* pin, double* pout, int check) {
    SorStreamArgs a;
    memset(&a, 0, sizeof a);
    a.pin = pin; a.pout = pout; a.rhs = c->rhs;
    a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.chunk_rows = c->chunk_rows; a.n_strips = c->n_strips; a.n_tasks = c->n_partials; a.n_chunks = c->n_chunks;
    a.n_groups = (int)(c->g.pitch / NSB_GROUP);
    a.it = it; a.check = check; a.zero = 0u;
    a.k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->J);     // main.cu:588: the GLOBAL cell count
    a.partials = c->d_partials; a.bandvals = c->d_bandvals; a.ctl = c->d_ctl;
    a.world = c->world; a.wall_bottom = c->wall_bottom; a.wall_top = c->wall_top;
    a.seq = c->seq + (unsigned long long)it;
    a.round = c->red_round + (unsigned long long)it;
    a.red = nsb_slab_red(c);
    a.err = &c->d_sync->error;
    const int par = (pout == c->pbuf[1]) ? 1 : 0;
    a.down_pout = pout; a.up_pout = pout; a.down_nj = 0;
    a.flag_from_down = c->d_sorflag; a.flag_from_up = c->d_sorflag + c->n_strips;
    a.down_flag = c->d_sorflag; a.up_flag = c->d_sorflag;
    if (c->world > 1) {
        if (!c->connected) return cudaErrorNotReady;
        if (!c->wall_bottom) {
            const SlabPeer& d = c->peer[c->rank - 1];
            a.down_pout = d.p[par]; a.down_nj = d.nj;
            a.down_flag = d.sorflag + c->n_strips;             // I am its UPPER neighbour
        }
        if (!c->wall_top) {
            const SlabPeer& u = c->peer[c->rank + 1];
            a.up_pout = u.p[par];
            a.up_flag = u.sorflag;                             // I am its LOWER neighbour
        }
    }
    const int blocks = (a.n_tasks + SOR_WARPS - 1) / SOR_WARPS;
    sor_stream_fn(c->use_recip, c->world > 1)<<<blocks, SOR_THREADS, SOR_SMEM_BYTES, c->stream>>>(a);
    return NSB_LAUNCHED(c, 1);
}

// All `n_iter` iterations of a solve in one launch (single slab).  Iteration 0 reads pbuf[pcur].
cudaError_t nsb_launch_sor_persist(nsb_ctx* c, int n_iter, int check) {
    if (n_iter <= 0) return cudaSuccess;
    SorPersistArgs pa;
    memset(&pa, 0, sizeof pa);
    SorStreamArgs& a = pa.s;
    a.pin = c->pbuf[c->pcur]; a.pout = c->pbuf[c->pcur ^ 1]; a.rhs = c->rhs;
    a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.chunk_rows = c->chunk_rows; a.n_strips = c->n_strips; a.n_tasks = c->n_partials; a.n_chunks = c->n_chunks;
    a.n_groups = (int)(c->g.pitch / NSB_GROUP);
    a.it = 0; a.check = check; a.zero = 0u;
    a.k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->J);     // main.cu:588: the GLOBAL cell count
    a.partials = c->d_partials; a.bandvals = c->d_bandvals; a.ctl = c->d_ctl;
    a.world = c->world; a.wall_bottom = c->wall_bottom; a.wall_top = c->wall_top;
    a.err = &c->d_sync->error;
    a.seq = c->seq;
    a.red = nsb_slab_red(c);
    pa.round0 = c->red_round;
    pa.buf0 = c->pcur;
    for (int k = 0; k < 2; k++) { pa.own_p[k] = c->pbuf[k]; pa.down_p[k] = c->pbuf[k]; pa.up_p[k] = c->pbuf[k]; }
    a.down_nj = 0;
    a.flag_from_down = c->d_sorflag; a.flag_from_up = c->d_sorflag + c->n_strips;
    a.down_flag = c->d_sorflag; a.up_flag = c->d_sorflag;
    if (c->world > 1) {
        if (!c->connected) return cudaErrorNotReady;
        if (!c->wall_bottom) {
            const SlabPeer& d = c->peer[c->rank - 1];
            pa.down_p[0] = d.p[0]; pa.down_p[1] = d.p[1]; a.down_nj = d.nj;
            a.down_flag = d.sorflag + c->n_strips;             // I am its UPPER neighbour
        }
        if (!c->wall_top) {
            const SlabPeer& u = c->peer[c->rank + 1];
            pa.up_p[0] = u.p[0]; pa.up_p[1] = u.p[1];
            a.up_flag = u.sorflag;                             // I am its LOWER neighbour
        }
    }
    c->seq += (unsigned long long)n_iter + 1ull;               // halo epochs: iterations ATTEMPTED, like flag_base
    pa.n_iter = n_iter;
    pa.claim = c->d_claim; pa.done = c->d_done; pa.partials2 = c->d_partials2;
    pa.flag_base = c->flag_base;
    c->flag_base += (unsigned long long)n_iter + 1ull;
    int bps = 0;
    const SorPersistFn kern = sor_persist_fn(c->use_recip, c->world > 1);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (const void*)kern, SOR_THREADS, SOR_SMEM_BYTES);
    if (bps < 1) bps = 1;
    const long long want = ((long long)n_iter * a.n_tasks + SOR_WARPS - 1) / SOR_WARPS;
    long long blocks = (long long)c->sm_count * bps;
    if (blocks > want) blocks = want;
    // at most two iterations are in flight at any time: more warps than that would only poll flags
    const long long useful = (2LL * a.n_tasks + SOR_WARPS - 1) / SOR_WARPS + 1;
    if (blocks > useful) blocks = useful;
    cudaError_t e = cudaMemsetAsync(c->d_claim, 0, 8 * sizeof(unsigned long long), c->stream);
    if (e != cudaSuccess) return e;
    kern<<<(int)blocks, SOR_THREADS, SOR_SMEM_BYTES, c->stream>>>(pa);
    return NSB_LAUNCHED(c, 1);
}
