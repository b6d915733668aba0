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
#include <stdlib.h>
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
// per-warp TMA ring.  SOR_BLK must be a multiple of 2 * SOR_NST and at least 3 K + 2 (SorK below).
#ifndef SOR_BLK
#define SOR_BLK 8
#endif
#ifndef SOR_NST
#define SOR_NST 2
#endif
#define SOR_USES_PER_BLK (SOR_BLK / 2 / SOR_NST)   // times a stage is filled and read per block
static_assert(SOR_BLK % (2 * SOR_NST) == 0, "SOR_BLK must be a multiple of 2 * SOR_NST");
#define SOR_ROW_BYTES 1024     // 32 lanes x 32 B
// One ring stage = p even row, p odd row, RHS even row, RHS odd row.  With K > 1 sub-iterations per pass a p row
// carries one more column group on either side of the 32 the lanes own (one more input column is needed there, see
// sor_colour): 34 groups = 1088 B.
#define SOR_PROW_BYTES(K) (SOR_ROW_BYTES + ((K) > 1 ? 64 : 0))
// K > 1 (SOR_RS): the RHS rows are NOT copied into the register window -- a row is used in six consecutive steps, and 7 rows
// of RHS are 56 registers, the difference between 8 and 10 resident warps per SM.  They stay in a ring of SOR_RS_PAIRS row
// pairs in shared memory (a pair is loaded with its p pair, under the same FULL barrier, and is dead 7 steps later: with 8
// pairs in the ring a slot is rewritten 11 steps after its last read) and every phase reads the words it needs.
// MEASURED AND NOT KEPT AS THE DEFAULT (profiles/r2y_rs_experiment.txt): at 200 registers it is still 8 warps per SM;
// forced to 168 registers (12 warps, 40 B of spills) the sweep takes 207 us per iteration at 8192^2 against 192 us with the
// RHS in registers -- the shared-memory loads sit right in front of their use (short_scoreboard 2.3 of 8.5 stall cycles
// per issue) and the spills add local-memory traffic (long_scoreboard 1.25).  -DSOR_K2_RHS_SMEM=1 builds it.
#ifndef SOR_K2_RHS_SMEM
#define SOR_K2_RHS_SMEM 0
#endif
#define SOR_RS(K) ((K) > 1 && SOR_K2_RHS_SMEM)
#define SOR_RS_PAIRS 6                                            // 12 warps x (4.25 KB of p stages + 12 KB of RHS ring) = 195 KB per SM
#define SOR_KWARPS(K) (SOR_WARPS)                                 // warps per block
#define SOR_KTHREADS(K) (SOR_KWARPS(K) * 32)
#define SOR_STAGE_BYTES(K) (2 * SOR_PROW_BYTES(K) + (SOR_RS(K) ? 0 : 2 * SOR_ROW_BYTES))
#define SOR_RING_BYTES(K) (SOR_NST * SOR_STAGE_BYTES(K))          // the p (+ RHS) stages of one warp
#define SOR_WARP_SMEM(K) (SOR_RING_BYTES(K) + (SOR_RS(K) ? SOR_RS_PAIRS * 2 * SOR_ROW_BYTES : 0))
#define SOR_BAR_BYTES (2 * SOR_NST * 8 + 16)                      // per warp: FULL/EMPTY mbarriers + the two push distances
#define SOR_SMEM_BYTES(K) (SOR_KWARPS(K) * SOR_WARP_SMEM(K) + SOR_KWARPS(K) * SOR_BAR_BYTES)   // rings + barrier blocks
// Temporal blocking (K sub-iterations per pass over HBM).  A row that arrived in step j goes through
//     red (lag 1), black (lag 2), residual (lag 3)        of sub-iteration 0,
//     red (lag 4), black (lag 5), residual (lag 6)        of sub-iteration 1, ...
// one phase per step: in step j the phase with lag L works on row j - L.  Phases run in the order of
// increasing lag inside a step, so a phase finds its north neighbour (lag L-1) already advanced by this
// step and its south neighbour (lag L+1) not yet -- exactly the state the reference's sequence
// (ghosts, colour 0, ghosts, colour 1, residual; main.cu:684-706) produces row by row.  K = 2 needs rows
// j .. j-7 in the window: the 8 slots the unrolled block has anyway.
// How strips that reach beyond the interior columns are handled when every lane is entirely inside or entirely
// outside the domain (i_max a multiple of 4): 0 = by the general per-column edge body like any other edge strip,
// 1 = by the plain bodies with a per-lane mask evaluated at run time (one body for all strips),
// 2 = by separate masked instantiations of the plain bodies (interior strips carry no mask code).
// The general edge body costs ~3 x the plain one per row, and the edge tasks of pass n+1 wait for the edge tasks of
// pass n: they are the critical path of every pass whenever there is less than ~2.5 tasks per resident warp.
// Which mode is fastest depends on what it does to the register allocation and the instruction footprint of each
// instantiation -- measured on one B200 in one run (profiles/r2m_variants.txt, us per iteration, modes 0 / 1 / 2):
//   one iteration per pass, reciprocal product: 2048^2 36.6 / 32.2 / 42.4   4096^2 77.2 / 75.4 / 89.6   8192^2 250.3 / 253.5 / 253.2
//   one iteration per pass, IEEE division 8191^2: 309.8 / 383.1 / 431.8      (mode 2 spills in the plain body)
//   two iterations per pass, 8192^2: 248.0 / 299.3 / 193.3                   (its edge body is ~100 KB: mode 0 runs at ITS pace)
// hence: two iterations per pass -> 2; reciprocal product -> 1; IEEE division -> 0.  -DSOR_LANEMASK_MODE=n forces one.
__host__ __device__ constexpr int sor_lanemask_mode(int qm, int k) {
#ifdef SOR_LANEMASK_MODE
    return SOR_LANEMASK_MODE;
#else
    return k >= 2 ? 2 : (qm == NSB_QUOT_RECIP ? 1 : 0);
#endif
}
template <int K>
struct SorK {
    static constexpr int HALO = 2 * K + 1;        // rows a task reads below its first / above its last row; rows pushed across a slab boundary
    static constexpr int LAG_STORE = 3 * K - 1;   // the last black update: the row is final and is stored
    static constexpr int LAG_LAST = 3 * K;        // the last residual
};
static_assert(SorK<NSB_SOR_KMAX>::LAG_LAST + 2 <= SOR_BLK, "register window too small for the last residual");
static_assert(SorK<NSB_SOR_KMAX>::HALO <= NSB_ROW0 + 1 && SorK<NSB_SOR_KMAX>::HALO <= NSB_ROWS_ABOVE, "pad rows must cover the task halo");

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
    int dbg_fixed_loads;     // developer probe (WRONG results): loads never leave two rows, i.e. come from L2 -- the sweep's speed without HBM reads
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
__device__ __forceinline__ long long lds_s64(uint32_t addr) {
    long long v;
    asm volatile("ld.shared.s64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_s64(uint32_t addr, long long v) {
    asm volatile("st.shared.s64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}

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
__device__ __forceinline__ void sor_close_iteration(double total, SorCtl* ctl, int it, long long cells, bool publish = true) {
    double res = sqrt(total / (double)cells);              // main.cu:588
    ctl->sum_r2 = total;
    ctl->res = res;
    if (res <= ctl->thresh) ctl->converged = 1;             // main.cu:713
    __threadfence();                                        // `converged` is visible before `iters` says "closed"
    *((volatile int*)&ctl->iters) = it + 1;
    __threadfence();
    if (publish) sor_publish(ctl);
}

// Per-lane geometry, fixed for the whole task.
struct LaneGeom {
    bool inner[4];   // column k is an interior column (1 <= i <= i_max)
    bool li;         // all four columns of this lane are interior columns
    bool masked;     // warp-uniform: the strip holds ghost / pad columns AND every lane is either all-interior or
                     // all-exterior (i_max a multiple of 4): the plain bodies handle it with the per-lane mask `li`
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
    // Slab geometry of the task in ONE register (bit 0 / 1: the slab's bottom / top is a wall of the domain; bit 2 / 3:
    // this task owns rows of the lower / upper neighbour's halo).  Where the pushed rows go is not kept in registers at
    // all: the byte distance from this slab's output row to the same row in the neighbour's halo sits in two shared-memory
    // words of the warp (sor_task_buffers), read only on the <= 5 rows per boundary task that are pushed.  (r2: two
    // pointers, a row count and four flags were live across the hot body of the multi-slab instantiation.)
    uint32_t mf;
    uint32_t push_slot;       // shared address of {down, up} byte distances
    __device__ __forceinline__ bool wall_bottom() const { return (mf & 1u) != 0u; }
    __device__ __forceinline__ bool wall_top() const { return (mf & 2u) != 0u; }
    __device__ __forceinline__ bool push_down() const { return (mf & 4u) != 0u; }
    __device__ __forceinline__ bool push_up() const { return (mf & 8u) != 0u; }
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
    uint32_t copy_bytes;      // bytes per RHS row copy (1 KB, clipped at the right edge of the pitch)
    uint32_t copy_bytes_p;    // bytes per p row copy (K > 1: one more group on either side)
    double ex;                // K > 1: the extra input column of the row that arrived in the previous step (lane 0: west, lane 31: east)
    uint32_t smem_base;       // this warp's ring (shared address)
    uint32_t rs_base;         // SOR_RS: this warp's RHS ring (shared address); rs_lane = generic pointer + 4 * lane doubles
    const double* rs_lane;
    int rs_pair;              // SOR_RS: pairs issued so far in this task (slot = rs_pair & 7)
    int rs_h;                 // SOR_RS: ring slot of the current block's first pair
    long long ld_step;        // 2 * pitch (0 with the developer probe NSB_DEBUG_FIXED_LOADS: every pair re-reads the same two rows)
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

template <int K>
__device__ __forceinline__ void issue_pair(TaskCtx& t, int st, uint32_t empty_phase, bool wait_empty) {
#ifdef SOR_LANE0_ISSUE
    if (t.lane == 0) {
#else
    if (elect_one()) {
#endif
        const uint32_t bar = t.bar_base + st * 8;
        if (wait_empty) mbar_wait(t.bar_base + (SOR_NST + st) * 8, empty_phase);
        if (t.j_next <= t.j_limit) {
            const uint32_t dst = t.smem_base + st * SOR_STAGE_BYTES(K);
            const uint32_t rdst = SOR_RS(K) ? t.rs_base + (uint32_t)(t.rs_pair % SOR_RS_PAIRS) * (2u * SOR_ROW_BYTES)
                                            : dst + 2 * SOR_PROW_BYTES(K);
            mbar_expect_tx(bar, 2u * t.copy_bytes_p + 2u * t.copy_bytes);
            bulk_g2s(dst, t.pin_next, t.copy_bytes_p, bar);
            bulk_g2s(dst + SOR_PROW_BYTES(K), t.pin_next + t.pitch, t.copy_bytes_p, bar);
            bulk_g2s(rdst, t.rhs_next, t.copy_bytes, bar);
            bulk_g2s(rdst + SOR_ROW_BYTES, t.rhs_next + t.pitch, t.copy_bytes, bar);
        } else {
            mbar_arrive(bar);
        }
    }
    t.j_next += 2;
    t.rs_pair += 1;
    t.pin_next += t.ld_step;
    t.rhs_next += t.ld_step;
}

// SOR_RS: the RHS row of the phase with lag L in step S of a block (row j - L), this lane's 4 columns.  The block's first
// pair sits in ring slot t.rs_h = (pairs before the block) mod 6 (warp-uniform, set per block); the pair of this phase is d
// pairs further (d compile-time, -3 .. 3).
template <int S, int L>
__device__ __forceinline__ const double* sor_rs_row(const TaskCtx& t) {
    constexpr int rel = S - L;
    constexpr int d = (rel >= 0) ? rel / 2 : -((-rel + 1) / 2);     // floor(rel / 2): pair relative to the block's first
    constexpr int par = rel - 2 * d;                                 // row of the pair
    int slot = t.rs_h + d;
    if (d < 0) { if (slot < 0) slot += SOR_RS_PAIRS; }
    else if (d > 0) { if (slot >= SOR_RS_PAIRS) slot -= SOR_RS_PAIRS; }
    return t.rs_lane + slot * (2 * SOR_ROW_BYTES / 8) + par * (SOR_ROW_BYTES / 8);
}

// ---- one colour phase: sub-iteration Q's red (L = 3Q+1) or black (L = 3Q+2) update of row j - L.
//   EX: the strip holds ghost / pad columns (first and last strips).
//   EY: the block of steps touches rows that need gating (walls, chunk ends, pushed rows).
//   QM: how quotients are formed (nsb_math.cuh); with NSB_QUOT_TRY `ok` is cleared when a quotient that
//       is USED left the window of the three-operation division (pad / ghost columns of an edge strip may
//       hold anything and are not counted).
// The last black phase (L = 3K-1) leaves the row final and stores it (one full 32-byte sector per lane).
template <int QM, bool EX, bool EY, bool MASK, int K, int S, int L>
__device__ __forceinline__ void sor_colour(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int j, TaskCtx& t,
                                           const SorCoef& kf, const LaneGeom& lg, bool& ok) {
    constexpr int NS = SOR_BLK;
    constexpr int Q = (L - 1) / 3;
    constexpr bool BLACK = (L % 3) == 2;
    constexpr bool STORE = (L == SorK<K>::LAG_STORE);
    constexpr int HALO = SorK<K>::HALO;
    constexpr int sX = (S + 2 * NS - L) % NS, sN = (S + 2 * NS - L + 1) % NS, sS = (S + 2 * NS - L - 1) % NS;   // row, north, south
    // cells of the colour in row j - L: (i + row) % 2 == colour; column k of a lane has i = 1 + k (mod 2), row j has the parity of S
    constexpr int c0 = ((BLACK ? 1 : 0) + 1 + S + L) & 1;
    constexpr int c1 = c0 + 2;
    // rows that take updates: 1 .. nj at walls; beyond a slab boundary sub-iteration Q also advances the
    // margin rows the later sub-iterations of this pass read
    constexpr int margin = 1 + 2 * (K - 1 - Q);
    const unsigned FULL = 0xffffffffu;
    const int r = j - L;
    const bool row_in = !EY || ((r >= (t.wall_bottom() ? 1 : -margin)) && (r <= (t.wall_top() ? t.nj : t.nj + 1 + margin)));
    double (&X)[4] = P[sX];
    constexpr int LM = sor_lanemask_mode(QM, K);
    const bool mask = MASK || (LM == 1 && !EX && lg.masked);
    if (EX || mask) refresh_ghost_columns(X, lg);                       // main.cu:686 / 694
    if (EY) {
        if (r == 1 && t.wall_bottom()) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sS][k] = X[k];                // p[i][0] = p[i][1]
        }
        if (r == t.nj && t.wall_top()) {
#pragma unroll
            for (int k = 0; k < 4; k++) P[sN][k] = X[k];                // p[i][j_max+1] = p[i][j_max]
        }
    }
    double xw = 0.0, xe = 0.0;
    if (c0 == 0) xw = __shfl_up_sync(FULL, X[3], 1);
    else xe = __shfl_down_sync(FULL, X[0], 1);
    if (K > 1 && L == 1) {                    // first red update of the pass: the column beyond the halo lanes (sor_step)
        if (c0 == 0) { if (t.lane == 0) xw = t.ex; }
        else { if (t.lane == 31) xe = t.ex; }
    }
    const double w0 = (c0 == 0) ? xw : X[(c0 + 3) & 3];
    const double e0 = X[c0 + 1];
    const double w1 = X[c1 - 1];
    const double e1 = (c1 == 3) ? xe : X[(c1 + 1) & 3];
    bool ok0 = true, ok1 = true;
    double rh0, rh1;
    if constexpr (SOR_RS(K)) { const double* rr = sor_rs_row<S, L>(t); rh0 = rr[c0]; rh1 = rr[c1]; }
    else { rh0 = R[sX][c0]; rh1 = R[sX][c1]; }
    const double u0 = sor_cell_q<QM>(X[c0], e0, w0, P[sN][c0], P[sS][c0], rh0, kf, EX ? ok0 : ok);
    const double u1 = sor_cell_q<QM>(X[c1], e1, w1, P[sN][c1], P[sS][c1], rh1, kf, EX ? ok1 : ok);
    if (EX) {
        if (row_in && lg.inner[c0]) { X[c0] = u0; ok = ok && ok0; }
        if (row_in && lg.inner[c1]) { X[c1] = u1; ok = ok && ok1; }
    } else {
        // MASK: lanes outside the domain (first / last strip of a grid whose i_max is a multiple of 4) keep
        // their values -- a per-lane select instead of the per-column predicates of the general edge body
        if (row_in && (!(MASK || LM == 1) || lg.li)) { X[c0] = u0; X[c1] = u1; }    // (pad lanes' `ok` is discarded at the end of the task)
    }
    if (STORE) {
        const bool in_chunk = (r >= t.jc0 && r <= t.jc1);     // warp-uniform; the blocks at the ends of a chunk run this body too
        if ((EX || mask) && in_chunk) {
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
        if (lg.owner && in_chunk) {
            double2* q = reinterpret_cast<double2*>(t.pout_row);
            q[0] = make_double2(X[0], X[1]);
            q[1] = make_double2(X[2], X[3]);
            if (EY) {
                // ghost rows j = 0 / j_max+1 as the same refresh left them (copies of row 1 / j_max taken
                // before the black update, held in the window slots south / north of the row)
                if (r == 1 && t.wall_bottom()) {
                    double2* h = reinterpret_cast<double2*>(t.pout_row - t.pitch);
                    h[0] = make_double2(P[sS][0], P[sS][1]);
                    h[1] = make_double2(P[sS][2], P[sS][3]);
                }
                if (r == t.nj && t.wall_top()) {
                    double2* h = reinterpret_cast<double2*>(t.pout_row + t.pitch);
                    h[0] = make_double2(P[sN][0], P[sN][1]);
                    h[1] = make_double2(P[sN][2], P[sN][3]);
                }
                // y-slab halo: the HALO rows next to a slab boundary also go straight into the
                // neighbour's halo rows over NVLink (peer store), SURVEY 8e
                if (t.push_down() && r <= HALO) {
                    double2* h = reinterpret_cast<double2*>(reinterpret_cast<char*>(t.pout_row) + lds_s64(t.push_slot));
                    h[0] = make_double2(X[0], X[1]);
                    h[1] = make_double2(X[2], X[3]);
                }
                if (t.push_up() && r >= t.nj - HALO + 1) {
                    double2* h = reinterpret_cast<double2*>(reinterpret_cast<char*>(t.pout_row) + lds_s64(t.push_slot + 8u));
                    h[0] = make_double2(X[0], X[1]);
                    h[1] = make_double2(X[2], X[3]);
                }
            }
        }
        t.pout_row += t.pitch;
    }
}

// ---- residual of row j - L (final for its sub-iteration, as are its neighbours) with the ghosts as they are
//      (last refreshed before colour 1, main.cu:694-702)
template <int QM, bool EX, bool EY, bool MASK, int K, int S, int L>
__device__ __forceinline__ void sor_resid(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int j, const TaskCtx& t,
                                          const SorCoef& kf, const LaneGeom& lg, double& acc, bool& ok) {
    constexpr int NS = SOR_BLK;
    constexpr int sX = (S + 2 * NS - L) % NS, sN = (S + 2 * NS - L + 1) % NS, sS = (S + 2 * NS - L - 1) % NS;
    const unsigned FULL = 0xffffffffu;
    const int r = j - L;
    double (&X)[4] = P[sX];
    const double xw = __shfl_up_sync(FULL, X[3], 1);
    const double xe = __shfl_down_sync(FULL, X[0], 1);
    if (r >= t.jc0 && r <= t.jc1) {            // warp-uniform; rows of the chunk are interior rows
        double part = 0.0;
        double rh[4];
        if constexpr (SOR_RS(K)) {
            const double2* rr = reinterpret_cast<const double2*>(sor_rs_row<S, L>(t));
            const double2 ra = rr[0], rb = rr[1];
            rh[0] = ra.x; rh[1] = ra.y; rh[2] = rb.x; rh[3] = rb.y;
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) rh[k] = R[sX][k];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const double w = (k == 0) ? xw : X[(k + 3) & 3];
            const double e = (k == 3) ? xe : X[(k + 1) & 3];
            bool okr = true;
            const double q = sor_residual_q<QM>(X[k], e, w, P[sN][k], P[sS][k], rh[k], kf, EX ? okr : ok);
            if (!EX || lg.inner[k]) { part = fma(q, q, part); if (EX) ok = ok && okr; }
        }
        if ((MASK || (sor_lanemask_mode(QM, K) == 1 && !EX)) && !lg.li) part = 0.0;        // lanes outside the domain (masked strips)
        acc += part;                           // non-owner lanes are discarded at the end of the task
    }
}

// the three phases of sub-iteration Q, then those of Q + 1
template <int QM, bool EX, bool EY, bool MASK, int K, int S, int Q>
__device__ __forceinline__ void sor_phases(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int j, TaskCtx& t,
                                           const SorCoef& kf, const LaneGeom& lg, double (&acc)[K], bool& ok) {
    if constexpr (Q < K) {
        sor_colour<QM, EX, EY, MASK, K, S, 3 * Q + 1>(P, R, j, t, kf, lg, ok);
        sor_colour<QM, EX, EY, MASK, K, S, 3 * Q + 2>(P, R, j, t, kf, lg, ok);
        sor_resid<QM, EX, EY, MASK, K, S, 3 * Q + 3>(P, R, j, t, kf, lg, acc[Q], ok);
        sor_phases<QM, EX, EY, MASK, K, S, Q + 1>(P, R, j, t, kf, lg, acc, ok);
    }
}

// One pipeline step with compile-time window slot S (row parity == S parity).  sb is the step
// number of the first step of the current block of 8 (a multiple of 8); the row that arrives in
// this step is j = jstart + sb + S.
template <int QM, bool EX, bool EY, bool MASK, int K, int S>
__device__ __forceinline__ void sor_step(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0, TaskCtx& t,
                                         const SorCoef& kf, const LaneGeom& lg, double (&acc)[K], bool& ok) {
    constexpr int NS = SOR_BLK;                                                    // window slots
    constexpr int sE = S % NS, sD = (S + NS - 1) % NS;                             // rows j, j-1
    constexpr int st = (S >> 1) % SOR_NST;
    const int j = t.jstart + sb + S;

    // ---- row j (p and RHS) from the ring into registers; after the odd row of a pair the stage
    //      goes back to the TMA engine
    double ex_new = 0.0;
    {
        const uint32_t ph = (use0 + (uint32_t)((S >> 1) / SOR_NST)) & 1u;      // use count of the stage, mod 2
        if ((S & 1) == 0) {
            mbar_wait(t.bar_base + st * 8, ph ^ t.full_par);
#ifdef SOR_SYNC_EACH_STEP
            __syncwarp();
#endif
        }
        const double* prow = t.lds_lane + st * (SOR_STAGE_BYTES(K) / 8) + (S & 1) * (SOR_PROW_BYTES(K) / 8);
        const double2* q = reinterpret_cast<const double2*>(prow + (K > 1 ? NSB_GROUP : 0));
        const double2 a = q[0], b = q[1];
        double2 c = make_double2(0.0, 0.0), d = make_double2(0.0, 0.0);
        if constexpr (!SOR_RS(K)) {
            const double* rrow = t.lds_lane + st * (SOR_STAGE_BYTES(K) / 8) + 2 * (SOR_PROW_BYTES(K) / 8) + (S & 1) * (SOR_ROW_BYTES / 8);
            const double2* qr = reinterpret_cast<const double2*>(rrow);
            c = qr[0]; d = qr[1];
        }
        if (K > 1) {
            // The red update of sub-iteration 0 needs ONE input column beyond the 128 the lanes hold: the west
            // neighbour of lane 0's column 0 in rows where that cell is red, the east neighbour of lane 31's column 3 in
            // the others (two sub-iterations and the residual reach 5 columns beyond the 120 owned ones, the halo lanes
            // hold 4).  It sits in the extra group the bulk copy brought; every lane reads the same word.
            const double* e = t.lds_lane - t.lane * NSB_GROUP + st * (SOR_STAGE_BYTES(K) / 8) + (S & 1) * (SOR_PROW_BYTES(K) / 8);
            ex_new = (S & 1) ? e[NSB_GROUP - 1] : e[NSB_GROUP + 32 * NSB_GROUP];
        }
        P[sE][0] = a.x; P[sE][1] = a.y; P[sE][2] = b.x; P[sE][3] = b.y;
        if constexpr (!SOR_RS(K)) { R[sE][0] = c.x; R[sE][1] = c.y; R[sE][2] = d.x; R[sE][3] = d.y; }
        if (S & 1) {
            // Both rows of the pair are read: every lane releases the stage on its EMPTY barrier and
            // the elected lane re-arms the stage only after acquiring that phase (32 arrivals).
            // The arrive must not overtake the LDS above: neither __syncwarp() nor the arrive's
            // release semantics kept the bulk copy from overwriting a row whose LDS was still queued
            // (seen on B200 as sporadically wrong RHS values in the last-loaded 16 bytes).  So the
            // barrier ADDRESS is made data-dependent on all LDS of the pair (this step's
            // a, b, c, d and the previous step's, now in window slot D): `zero` is a run-time 0 the
            // compiler cannot fold, and the arrive cannot issue before the loads have landed.
            int dep = (__double2hiint(a.x) ^ __double2hiint(b.x)) ^ (__double2hiint(P[sD][0]) ^ __double2hiint(P[sD][2]));
            if constexpr (!SOR_RS(K)) dep ^= (__double2hiint(c.x) ^ __double2hiint(d.x)) ^ (__double2hiint(R[sD][0]) ^ __double2hiint(R[sD][2]));
            if (K > 1) dep ^= __double2hiint(ex_new) ^ __double2hiint(t.ex);
            mbar_arrive(t.bar_base + (SOR_NST + st) * 8 + ((uint32_t)dep & t.zero));
            issue_pair<K>(t, st, ph ^ t.empty_par, true);
        }
    }
    sor_phases<QM, EX, EY, MASK, K, S, 0>(P, R, j, t, kf, lg, acc, ok);
    if (K > 1) t.ex = ex_new;                 // used by the red update of this row in the next step
}

template <int QM, bool EX, bool EY, bool MASK, int K, int S>
__device__ __forceinline__ void sor_steps_from(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0,
                                               TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double (&acc)[K], bool& ok) {
    if constexpr (S < SOR_BLK) {
        sor_step<QM, EX, EY, MASK, K, S>(P, R, sb, use0, t, kf, lg, acc, ok);
        sor_steps_from<QM, EX, EY, MASK, K, S + 1>(P, R, sb, use0, t, kf, lg, acc, ok);
    }
}
// One unrolled block of SOR_BLK steps; use0 = times every ring stage has been used before this block.
template <int QM, bool EX, bool EY, bool MASK, int K>
__device__ __forceinline__ void sor_block8(double (&P)[SOR_BLK][4], double (&R)[SOR_BLK][4], const int sb, const uint32_t use0,
                                           TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double (&acc)[K], bool& ok) {
    sor_steps_from<QM, EX, EY, MASK, K, 0>(P, R, sb, use0, t, kf, lg, acc, ok);
}

// Steps of a task: rows jstart .. jc1 + 3K arrive (the last ones only to drain the pipeline: the last residual of
// row jc1 runs 3K steps after that row arrived; rows above jc1 + HALO are not fetched).
template <int K>
__device__ __forceinline__ int sor_task_steps(const TaskCtx& t) { return t.jc1 + SorK<K>::LAG_LAST - t.jstart + 1; }

// One task = one strip x one chunk of rows, marched bottom to top by one warp.  acc[q] = this lane's share of
// sum r^2 of sub-iteration q over the rows of the chunk.
template <int QM, bool EX, int K>
__device__ __forceinline__ void sor_run_task(TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, double (&acc)[K], bool& ok) {
    double P[SOR_BLK][4], R[SOR_BLK][4];
#pragma unroll
    for (int s = 0; s < SOR_BLK; s++)
#pragma unroll
        for (int k = 0; k < 4; k++) { P[s][k] = 0.0; R[s][k] = 0.0; }
#pragma unroll
    for (int q = 0; q < K; q++) acc[q] = 0.0;
    const int n_steps = sor_task_steps<K>(t);
    uint32_t use0 = 0u;
    for (int sb = 0; sb < n_steps; sb += SOR_BLK, use0 += SOR_USES_PER_BLK) {
        const int jb = t.jstart + sb;                        // row arriving in the first step of the block
        if constexpr (SOR_RS(K)) t.rs_h = (sb >> 1) % SOR_RS_PAIRS;
        // rows touched by the block (B = SOR_BLK): residuals jb-3K .. jb+B-4, stores jb-(3K-1) .. jb+B-3K, updates up to jb+B-2, loads jb .. jb+B-1
#ifdef SOR_NO_LEAN
        const bool clean = false;
#else
        // The plain body gates nothing but the stores and the residual rows (by the rows of the chunk): it serves every
        // block whose updates stay off the wall rows 1 / nj (ghost-row mirrors, rows that must not change) and whose
        // stores are not pushed to a neighbouring slab -- i.e. all blocks of all chunks but the first and the last one.
        // (r1: a separate chunk-edge body made the kernel larger and slower; the gates cost ~3 % of the plain body's
        // instructions and keep the edge blocks of short chunks out of the 2.5 x larger wall body.)
        constexpr int LS = SorK<K>::LAG_STORE, HALO = SorK<K>::HALO;
        const bool clean = (!t.wall_bottom() || jb - LS >= 2) && (!t.wall_top() || jb + SOR_BLK - 2 <= t.nj - 1) &&
                           !(t.push_down() && jb - LS <= HALO) && !(t.push_up() && jb + SOR_BLK - 1 - LS >= t.nj - HALO + 1);
#endif
        // (a fourth variant <EX, !EY> for the interior rows of the two edge strips was measured: the larger
        //  kernel body costs the LEAN path 10 % -- 5.27 against 5.86 TB/s at 8192^2 -- and was dropped)
        // Five bodies: the general edge body (per-column predicates), and plain / wall bodies with and without the
        // per-lane mask.  Interior strips -- all but two -- only ever run the two unmasked ones.
        if (EX) sor_block8<QM, true, true, false, K>(P, R, sb, use0, t, kf, lg, acc, ok);
        else if (sor_lanemask_mode(QM, K) == 2 && lg.masked) {
            if (clean && QM != NSB_QUOT_FULL) sor_block8<QM, false, false, true, K>(P, R, sb, use0, t, kf, lg, acc, ok);
            else sor_block8<QM, false, true, true, K>(P, R, sb, use0, t, kf, lg, acc, ok);
        } else {
            if (clean && QM != NSB_QUOT_FULL) sor_block8<QM, false, false, false, K>(P, R, sb, use0, t, kf, lg, acc, ok);
            else sor_block8<QM, false, true, false, K>(P, R, sb, use0, t, kf, lg, acc, ok);
        }
    }
    if (!lg.owner) {
#pragma unroll
        for (int q = 0; q < K; q++) acc[q] = 0.0;
    }
    if (!EX && lg.masked && !lg.li) ok = true; // quotients of lanes outside the domain are never used: they may hold anything
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
template <bool RECIP, int K>
__device__ __forceinline__ void sor_task(TaskCtx& t, const SorCoef& kf, const LaneGeom& lg, bool edge_strip,
                                         uint32_t& full_par, uint32_t& empty_par, double (&acc)[K]) {
    constexpr int QM = RECIP ? NSB_QUOT_RECIP : NSB_QUOT_TRY;
    const int n_steps = sor_task_steps<K>(t);
    const double* pin0 = t.pin_next; const double* rhs0 = t.rhs_next; double* pout0 = t.pout_row;
    t.full_par = full_par; t.empty_par = empty_par;
    sor_advance_parities(n_steps, full_par, empty_par);
    t.rs_pair = 0;
    __syncwarp();
#pragma unroll
    for (int q = 0; q < SOR_NST; q++) issue_pair<K>(t, q, 0u, false);
    bool ok = true;
    if (edge_strip) sor_run_task<QM, true, K>(t, kf, lg, acc, ok);
    else sor_run_task<QM, false, K>(t, kf, lg, acc, ok);
    if (!RECIP) {
        if (!__all_sync(0xffffffffu, ok)) {
            t.pin_next = pin0; t.rhs_next = rhs0; t.pout_row = pout0; t.j_next = t.jstart;
            t.rs_pair = 0;
            t.full_par = full_par; t.empty_par = empty_par;
            sor_advance_parities(n_steps, full_par, empty_par);
            __syncwarp();
#pragma unroll
            for (int q = 0; q < SOR_NST; q++) issue_pair<K>(t, q, 0u, false);
            if (edge_strip) sor_run_task<NSB_QUOT_FULL, true, K>(t, kf, lg, acc, ok);
            else sor_run_task<NSB_QUOT_FULL, false, K>(t, kf, lg, acc, ok);
        }
    }
}

// Geometry of one task (strip, chunk) for a pass of K sub-iterations: everything of TaskCtx that does not depend on
// the buffers.  MULTI: several y-slabs.
template <bool RECIP, bool MULTI, int K>
__device__ __forceinline__ void sor_task_geometry(TaskCtx& t, LaneGeom& lg, bool& edge_strip, const SorStreamArgs& a,
                                                  int strip, int chunk, int lane) {
    constexpr int HALO = SorK<K>::HALO;
    const unsigned FULLM = 0xffffffffu;
    const int g0 = strip * NSB_STRIP_GROUPS;             // first column group of the strip
    const int g = g0 + lane;                             // column group of this lane
    const int i0 = g * NSB_GROUP - NSB_COL0;             // i of column k = 0
    bool all_inner = (g < a.n_groups), none_inner = true;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        lg.inner[k] = (i0 + k >= 1) && (i0 + k <= a.ni);
        all_inner = all_inner && lg.inner[k];
        none_inner = none_inner && !lg.inner[k];
    }
    lg.li = all_inner;
    lg.owner = (g < a.n_groups) && lane >= 1 && lane <= NSB_STRIP_GROUPS;
    lg.ghostW = (i0 + 3 == 0);
    const int kg = a.ni + 1 - i0;
    lg.kGhostE = (kg >= 0 && kg <= 3) ? kg : -1;
    lg.warpW = __any_sync(FULLM, lg.ghostW);
    lg.warpE = __any_sync(FULLM, lg.kGhostE >= 0);
    // Strips that reach beyond the interior columns: when every lane is entirely inside or entirely outside
    // (i_max a multiple of 4: the west ghost is the last column of lane 0, the east ghost the first column of its
    // lane) the plain bodies handle them with a per-lane mask; only otherwise the general per-column edge body runs.
    // (The edge body costs ~3 x the plain one per row and sits on the critical path of every iteration: the edge
    // tasks of iteration n+1 wait for the edge tasks of iteration n.  2048^2 .. 4096^2 ran at exactly their pace.)
    const bool any_out = !__all_sync(FULLM, all_inner);
    const bool uniform = __all_sync(FULLM, all_inner || none_inner || g >= a.n_groups);
#ifdef SOR_FORCE_EX
    edge_strip = true;
    lg.masked = false;
#else
    constexpr int LM = sor_lanemask_mode(RECIP ? NSB_QUOT_RECIP : NSB_QUOT_TRY, K);
    edge_strip = LM == 0 ? any_out : (any_out && !uniform);
    lg.masked = LM != 0 && any_out && uniform;
#endif
    t.jc0 = 1 + chunk * a.chunk_rows;
    t.jc1 = min(t.jc0 + a.chunk_rows - 1, a.nj);
    t.jstart = t.jc0 - HALO;                             // even, because chunk_rows is even and HALO is odd
    t.nj = a.nj;
    {
        const bool wb = !(MULTI && (a.wall_bottom == 0)), wt = !(MULTI && (a.wall_top == 0));
        const bool pd = (MULTI && (!wb && t.jc0 <= HALO)), pu = (MULTI && (!wt && t.jc1 >= a.nj - HALO + 1));
        t.mf = (wb ? 1u : 0u) | (wt ? 2u : 0u) | (pd ? 4u : 0u) | (pu ? 8u : 0u);
    }
    t.pitch = a.pitch;
    t.ld_step = a.dbg_fixed_loads ? 0 : 2 * a.pitch;
    t.lane = lane;
    t.zero = a.zero;
    t.j_next = t.jstart;
    t.j_limit = min(t.jc1 + HALO, a.nj + HALO);
    const int groups_here = min(32, a.n_groups - g0);
    t.copy_bytes = (uint32_t)groups_here * 32u;
    t.copy_bytes_p = K > 1 ? (uint32_t)min(34, a.n_groups - g0 + 1) * 32u : t.copy_bytes;   // from group g0 - 1
    t.ex = 0.0;
}

// Buffers of a task: first rows to fetch, first row to store, the neighbours' halo rows.
// Row r of this slab's output is at pout + (r + ROW0) * pitch + column; the same row in the lower neighbour's upper halo
// at down_pout + (down_nj + r + ROW0) * pitch + column, in the upper neighbour's lower halo at
// up_pout + (r - nj + ROW0) * pitch + column: two byte distances per task, kept in the warp's shared-memory slot.
template <int K>
__device__ __forceinline__ void sor_task_buffers(TaskCtx& t, const double* pin, double* pout, const double* rhs,
                                                 double* down_pout, double* up_pout, int down_nj, uint32_t push_slot,
                                                 int strip, int lane) {
    t.push_slot = push_slot;
    if ((t.mf & 12u) && lane == 0) {
        const long long dd = (long long)(reinterpret_cast<const char*>(down_pout) - reinterpret_cast<const char*>(pout)) +
                             (long long)down_nj * t.pitch * 8;
        const long long du = (long long)(reinterpret_cast<const char*>(up_pout) - reinterpret_cast<const char*>(pout)) -
                             (long long)t.nj * t.pitch * 8;
        sts_s64(push_slot, dd);
        sts_s64(push_slot + 8u, du);
    }
    __syncwarp();
    const long long coff = (long long)strip * NSB_STRIP_GROUPS * NSB_GROUP;    // first column of the strip
    const long long row_first = (long long)(t.jstart + NSB_ROW0) * t.pitch;
    t.pin_next = pin + row_first + coff - (K > 1 ? NSB_GROUP : 0);     // K > 1: the copy starts one group to the left
    t.rhs_next = rhs + row_first + coff;
    t.pout_row = pout + row_first - SorK<K>::LAG_STORE * t.pitch + coff + (long long)lane * NSB_GROUP;
}

#ifdef SOR_MAXNREG
#define SOR_KERNEL_ATTR __maxnreg__(SOR_MAXNREG)
#else
#define SOR_KERNEL_ATTR __launch_bounds__(SOR_THREADS, SOR_MIN_BLOCKS)
#endif
// K = 2 keeps 8 rows of p and 7 of RHS in registers (~230): two blocks per SM
#ifndef SOR_MIN_BLOCKS_K2
#define SOR_MIN_BLOCKS_K2 2
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
        LaneGeom lg;
        TaskCtx t;
        bool edge_strip;
        sor_task_geometry<RECIP, MULTI, 1>(t, lg, edge_strip, a, strip, chunk, lane);
        t.full_par = 0u; t.empty_par = 0u;
        t.bar_base = smem_u32(smem_raw + SOR_WARPS * SOR_WARP_SMEM(1) + warp * SOR_BAR_BYTES);
        sor_task_buffers<1>(t, a.pin, a.pout, a.rhs, a.down_pout, a.up_pout, a.down_nj, t.bar_base + 2 * SOR_NST * 8, strip, lane);
        unsigned char* ring = smem_raw + warp * SOR_WARP_SMEM(1);
        t.lds_lane = reinterpret_cast<const double*>(ring) + lane * NSB_GROUP;
        t.smem_base = smem_u32(ring);
        t.rs_base = 0u; t.rs_lane = NULL; t.rs_h = 0; t.rs_pair = 0;

        // ---- y-slab: the halo rows this task reads were stored by the neighbour's launch seq-1
        if ((MULTI && (a.world > 1)) && a.it > 0) {    // it == 0: the halo came with the p exchange at the solve entry
            const bool need_down = !t.wall_bottom() && t.jc0 - 3 <= 0;
            const bool need_up = !t.wall_top() && t.jc1 + 3 >= a.nj + 1;
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
        double accs[1];
        sor_task<RECIP, 1>(t, a.k, lg, edge_strip, full_par, empty_par, accs);
        const double acc = warp_sum(accs[0]);
        if (lane == 0) a.partials[chunk * a.n_strips + strip] = acc;

        if (t.push_down() || t.push_up()) {                      // halo rows are out: tell the neighbour(s)
            __threadfence_system();
            __syncwarp();
            if (lane == 0) {
                if (t.push_down()) st_release_sys(a.down_flag + strip, a.seq + 1ull);
                if (t.push_up()) st_release_sys(a.up_flag + strip, a.seq + 1ull);
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
// Persistent multi-iteration sweep: ALL iterations of a solve in one launch, K sub-iterations per pass over HBM.
//
// Why persistent: one launch per iteration costs a pipeline drain per iteration -- a task (one warp marching up
// 126 rows) lives ~100-150 us, so the last round of blocks leaves most SMs idle; measured launch time
// = 75 us + 4.03 us per 120-row band at 8192 columns (profiles/r1d_wave_probe.txt), i.e. 22 % of a
// launch at 8192^2 is drain.  Here the warps of a fixed grid claim tasks from one global counter,
// task id = pass * n_tasks + (chunk * n_strips + strip), and a task of pass n starts as
// soon as its 3 x 3 neighbourhood of tasks of pass n-1 is done (their rows are what it reads,
// and together with THEIR 3 x 3 neighbourhoods of pass n-2 they are the only readers of the rows it
// overwrites), so the sweep of pass n+1 follows the sweep of pass n up the grid without a global barrier.
// Why K = 2 (temporal blocking, north_star): the pass reads p and RHS once and writes p once for TWO
// iterations -- 12 B per cell-update instead of 24 -- at the price of 5 + 5 halo rows per chunk instead of 3 + 3.
//   - buffers: three pressure buffers in rotation (pass n reads buffer n mod 3 and writes n+1 mod 3), so the
//     input of pass n is still intact when pass n is closed: if its FIRST sub-iteration met the threshold, the
//     host replays that single iteration from the intact input (K = 1 launch, bit-identical);
//   - deadlock freedom: ids are claimed in order and a task only waits for smaller ids, which are
//     running on resident warps or finished; every wait still has a deadline (slab.cuh);
//   - residual: per-task partials per sub-iteration (double-buffered by pass parity), the warp that takes the
//     last ticket of a pass sums them in the same fixed order as the one-launch-per-iteration
//     kernel (same bits) and closes the K iterations in order, stopping at the first that converged;
//   - convergence: a task of pass n waits until pass n-2 is closed and stops if that or an
//     earlier one converged.  Tasks of pass m+1 may run after pass m converged: they write
//     a buffer that neither holds the converged iterate nor the input of pass m.
struct SorPersistArgs {
    SorStreamArgs s;                 // geometry, coefficients, ctl; s.it = global number of the first iteration of this launch
    int n_pass;                      // passes of K iterations each
    int noclose;                     // replay of an iteration that is already closed: do not close again
    int wait_first;                  // y-slabs: pass 0 reads halo rows pushed by the previous launch (epoch s.seq)
    unsigned long long* claim;       // [0] task counter, [1..4] tickets of passes mod 4, [5] task counter of the edge strips
                                     // (edge_sms > 0), [8 ..] band tickets by pass parity
    int edge_sms;                    // > 0: the tasks of the first and the last strip are claimed from a queue of their own,
                                     // which the warps of SMs 0 .. edge_sms-1 serve first (sor_claim)
    unsigned long long* done;        // [n_tasks] epoch of the last pass the task finished
    unsigned long long flag_base;    // epoch of "before pass 0 of this launch"
    double* partials2;               // [2][NSB_SOR_KMAX][n_tasks]
    double* bandvals2;               // [2][NSB_SOR_KMAX][n_chunks + 1]
    // this rank's three pressure buffers and the neighbours' (index = buffer number), the buffer
    // pass 0 reads, and the reduction round of iteration s.it
    double* own_p[3]; double* down_p[3]; double* up_p[3];
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
        if (global_timer_ns() - t0 > spin_timeout_ns(err)) { *((volatile int*)err) = 1; return false; }
        __nanosleep(64);
    }
    return true;
}

// ---- claiming tasks.  One queue: id = pass * n_tasks + chunk * n_strips + strip.  Two queues (edge_sms > 0): the first
// and the last strip -- whose tasks run bodies of their own (lane-masked / general edge body), rarely executed and
// therefore cold in the 32 KB instruction cache whenever a warp next to the interior bodies runs them: 0.30 against 0.14
// stall samples per instruction, and the interior warps of that SM miss too (profiles/r3v_task_kinds.txt) -- have a
// queue of their own (bit 63 of the returned id), served first by the warps of SMs 0 .. edge_sms-1, which then run nothing
// else; the other SMs serve the interior queue first.  Within a queue tasks are claimed in (pass, chunk) order, a claimed
// task is held by a resident warp and waits only for tasks of earlier passes: no deadlock.  Lane 0 only.
#define SOR_EDGE_QUEUE (1ull << 63)
#define SOR_NO_TASK (~0ull)
__device__ __forceinline__ unsigned long long sor_claim(const SorPersistArgs& pa, bool prefer_edge, unsigned long long total_main,
                                                        unsigned long long total_edge) {
    if (pa.edge_sms <= 0) {
        const unsigned long long id = atomicAdd(pa.claim, 1ull);
        return id < total_main ? id : SOR_NO_TASK;
    }
    if (prefer_edge) {
        unsigned long long id = atomicAdd(pa.claim + 5, 1ull);
        if (id < total_edge) return id | SOR_EDGE_QUEUE;
        id = atomicAdd(pa.claim, 1ull);
        return id < total_main ? id : SOR_NO_TASK;
    }
    unsigned long long id = atomicAdd(pa.claim, 1ull);
    if (id < total_main) return id;
    id = atomicAdd(pa.claim + 5, 1ull);
    return id < total_edge ? (id | SOR_EDGE_QUEUE) : SOR_NO_TASK;
}

#ifdef SOR_TASK_TIMING
// Developer build (tools/task_timing.py): where a task's time goes, summed over all tasks of a launch (lane 0, ns).
// [0] claim  [1] closed-pass / convergence word  [2] 3 x 3 flags  [3] geometry, buffers, neighbour flags  [4] the march
// [5] partials, fence, done flag, ticket  [6] closing a pass (the closing warps only)  [7] tasks
__device__ unsigned long long g_task_t[16];   // [8..10] march time of interior / wall-chunk / edge-strip tasks, [11..13] their counts
extern "C" void nsb_debug_task_timing(unsigned long long* out) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_task_t, sizeof g_task_t);
    unsigned long long z[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    cudaMemcpyToSymbol(g_task_t, z, sizeof z);
}
#define SOR_TT(i)                                                                          \
    do {                                                                                   \
        if (lane == 0) {                                                                   \
            const unsigned long long now__ = global_timer_ns();                            \
            atomicAdd(&g_task_t[i], now__ - tt_prev);                                      \
            tt_prev = now__;                                                               \
        }                                                                                  \
    } while (0)
#else
#define SOR_TT(i) do { } while (0)
#endif

template <bool RECIP, bool MULTI, int K>
// registers: 168 (3 blocks of 4 warps) for one iteration per pass; two: 168 with the RHS ring in shared memory, 255 without (2 blocks of 4)
__global__ void __maxnreg__(K == 1 ? 168 : (SOR_RS(K) ? 168 : 255)) sor_persist_kernel(const SorPersistArgs pa) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const SorStreamArgs& a = pa.s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned FULLM = 0xffffffffu;
    // tasks per pass of the main queue / of the edge-strip queue (sor_claim)
    const int per_main = pa.edge_sms > 0 ? (a.n_strips - 2) * a.n_chunks : a.n_tasks;
    const int per_edge = pa.edge_sms > 0 ? 2 * a.n_chunks : 0;
    const unsigned long long total = (unsigned long long)pa.n_pass * (unsigned long long)per_main;
    const unsigned long long total_edge = (unsigned long long)pa.n_pass * (unsigned long long)per_edge;
    bool prefer_edge = false;
    if (pa.edge_sms > 0) {
        unsigned int smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        prefer_edge = (int)smid < pa.edge_sms;
    }
    unsigned char* ring = smem_raw + warp * SOR_WARP_SMEM(K);
    const uint32_t bar_base = smem_u32(smem_raw + SOR_KWARPS(K) * SOR_WARP_SMEM(K) + warp * SOR_BAR_BYTES);
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

#ifdef SOR_TASK_TIMING
    unsigned long long tt_prev = global_timer_ns();
#endif
    // The first task is claimed here; every further one while the finished task's stores drain (below): the claim's
    // round trip to L2 (~1 us) overlaps the fence instead of preceding the next task.  (Claiming BEFORE the march, one
    // task ahead, kept the claimed id live across the hot loop and was slower: 2048^2 32 -> 42 us)
    unsigned long long id = 0;
    if (lane == 0) id = sor_claim(pa, prefer_edge, total, total_edge);
    id = __shfl_sync(FULLM, id, 0);
    for (;;) {
        if (id == SOR_NO_TASK) break;
        SOR_TT(0);
        int ps, strip, chunk;                                                       // pass, strip, chunk of the claimed task
        if (pa.edge_sms <= 0) {
            ps = (int)(id / (unsigned long long)a.n_tasks);
            const int task = (int)(id - (unsigned long long)ps * (unsigned long long)a.n_tasks);
            strip = task % a.n_strips;
            chunk = task / a.n_strips;
        } else if (id & SOR_EDGE_QUEUE) {
            const unsigned long long e = id & ~SOR_EDGE_QUEUE;
            ps = (int)(e / (unsigned long long)per_edge);
            const int r = (int)(e - (unsigned long long)ps * (unsigned long long)per_edge);
            chunk = r >> 1;
            strip = (r & 1) ? a.n_strips - 1 : 0;
        } else {
            ps = (int)(id / (unsigned long long)per_main);
            const int r = (int)(id - (unsigned long long)ps * (unsigned long long)per_main);
            chunk = r / (a.n_strips - 2);
            strip = 1 + r % (a.n_strips - 2);
        }
        const int it_first = a.it + ps * K;                                        // global number of this pass's first iteration
        // Slabs are swept from both boundaries inwards (chunks 0, last, 1, last-1, ...): the boundary rows travel to the
        // neighbours while the interior is swept, and every chunk a task of the next pass waits for (its own, the one
        // below, the one above) was claimed a whole pass before it.  (r2 claimed 0, last, 1, 2, ...: chunk `last` of pass
        // n+1, second in the order, then waited for chunk last-1 of pass n, the LAST in the order -- one warp per strip
        // idle for a task's duration in every pass.  Measured at 2 x 8192^2: 669.8 against 670.3 GLUPS, no difference --
        // the sweep is HBM bound and the other warps of the SM use the bandwidth; kept because no task waits by design.)
        if ((MULTI && (a.world > 1)) && a.n_chunks > 1)
            chunk = (chunk & 1) ? a.n_chunks - 1 - (chunk >> 1) : (chunk >> 1);
        const int tslot = chunk * a.n_strips + strip;      // index of this task's flag and partials

        // ---- may this task run?  (1) pass ps-2 must be closed and nothing before may have converged (lane 9 reads
        // {iters, converged}, one 64-bit word; the wait also keeps the double-buffered partials, band values and tickets
        // of pass ps-2 safe, so it does not depend on `check`); (2) the 3 x 3 neighbourhood of the previous pass must be
        // done (lanes 0 .. 8, one flag each).  ONE acquire load per round serves both: the two used to be dependent L2
        // round trips, 0.8 + 1.5 us per task (profiles/r3s_task_timing.txt).  A converged solve stops the task before it
        // could wait for flags that never come: tasks of pass ps-1 that saw the convergence never ran.
        int stop = 0;
        {
            const int cc = chunk + lane / 3 - 1, ss = strip + lane % 3 - 1;
            const bool is_ctl = (lane == 9);
            const bool is_flag = ps > 0 && lane < 9 && cc >= 0 && cc < a.n_chunks && ss >= 0 && ss < a.n_strips;
            const unsigned long long* addr = is_ctl ? reinterpret_cast<const unsigned long long*>(&a.ctl->iters)
                                                    : pa.done + (is_flag ? (size_t)cc * a.n_strips + ss : (size_t)0);
            const unsigned long long want = pa.flag_base + (unsigned long long)ps;
            const bool wait_close = (ps >= 2 && !pa.noclose);
            bool pending = is_ctl || is_flag;
            unsigned long long t0 = 0;
            for (int round = 0;; round++) {
                int st = 0;
                if (pending) {
                    const unsigned long long w = ld_acquire_gpu(addr);
                    if (is_ctl) {
                        if (a.check && (unsigned)(w >> 32)) st = 1;                       // nothing closes after convergence
                        else if (!wait_close || (int)(unsigned)(w & 0xffffffffull) >= it_first - K) pending = false;
                    } else if (w >= want) pending = false;
                }
                if (__any_sync(FULLM, st)) { stop = 1; break; }
                if (!__any_sync(FULLM, pending)) break;
                if (round == 0) t0 = global_timer_ns();
                int bad = 0;
                if (lane == 0 && (*((volatile int*)a.err) || global_timer_ns() - t0 > spin_timeout_ns(a.err))) { *((volatile int*)a.err) = 1; bad = 1; }
                if (__any_sync(FULLM, bad)) { stop = 1; break; }
                __nanosleep(64);
            }
        }
        if (stop) break;
        SOR_TT(1);
        asm volatile("fence.proxy.async;" ::: "memory");       // other warps' generic stores -> this warp's TMA reads

        SOR_TT(2);
        const int bi = (pa.buf0 + ps) % 3;                  // buffer this pass reads
        const int bo = (bi + 1) % 3;                        // ... and writes

        LaneGeom lg;
        TaskCtx t;
        bool edge_strip;
        sor_task_geometry<RECIP, MULTI, K>(t, lg, edge_strip, a, strip, chunk, lane);
        sor_task_buffers<K>(t, pa.own_p[bi], pa.own_p[bo], a.rhs, pa.down_p[bo], pa.up_p[bo], a.down_nj, bar_base + 2 * SOR_NST * 8, strip, lane);
        t.lds_lane = reinterpret_cast<const double*>(ring) + lane * NSB_GROUP;
        t.smem_base = smem_u32(ring);
        t.bar_base = bar_base;
        t.rs_base = smem_u32(ring + SOR_RING_BYTES(K));
        t.rs_lane = reinterpret_cast<const double*>(ring + SOR_RING_BYTES(K)) + lane * NSB_GROUP;
        t.rs_h = 0; t.rs_pair = 0;

        // ---- y-slab: rows of the previous pass pushed by the neighbouring ranks (their boundary
        //      tasks of pass ps-1, strips s-1 .. s+1); a.seq = halo epoch of "before pass 0"
        if ((MULTI && (a.world > 1)) && (ps > 0 || pa.wait_first)) {
            const bool need_down = !t.wall_bottom() && t.jc0 - SorK<K>::HALO <= 0;
            const bool need_up = !t.wall_top() && t.jc1 + SorK<K>::HALO >= a.nj + 1;
            int ok2 = 1;
            if (lane < 3) {
                const int sidx = strip - 1 + lane;
                if (sidx >= 0 && sidx < a.n_strips) {
                    if (need_down) ok2 = spin_until_ge(a.flag_from_down + sidx, a.seq + (unsigned long long)ps, a.err) ? 1 : 0;
                    if (need_up && ok2) ok2 = spin_until_ge(a.flag_from_up + sidx, a.seq + (unsigned long long)ps, a.err) ? 1 : 0;
                }
            }
            ok2 = __all_sync(FULLM, ok2);
            if (!ok2) break;
            asm volatile("fence.proxy.async;" ::: "memory");
        }

        double accs[K];
        SOR_TT(3);
#ifdef SOR_TASK_TIMING
        const unsigned long long tt_m0 = global_timer_ns();
#endif
        sor_task<RECIP, K>(t, a.k, lg, edge_strip, full_par, empty_par, accs);
#ifdef SOR_TASK_TIMING
        if (lane == 0) {
            const int kind = (strip == 0 || strip == a.n_strips - 1) ? 2 : ((chunk == 0 || chunk == a.n_chunks - 1) ? 1 : 0);
            atomicAdd(&g_task_t[8 + kind], global_timer_ns() - tt_m0);
            atomicAdd(&g_task_t[11 + kind], 1ull);
        }
#endif
        SOR_TT(4);
        double* part = pa.partials2 + (size_t)(ps & 1) * NSB_SOR_KMAX * a.n_tasks;
#pragma unroll
        for (int q = 0; q < K; q++) {
            const double v = warp_sum(accs[q]);
            if (lane == 0) part[(size_t)q * a.n_tasks + tslot] = v;
        }

        // ---- publish: rows and partials are out
        unsigned long long next_id = 0;
        if (lane == 0) next_id = sor_claim(pa, prefer_edge, total, total_edge);   // the next task of this warp, in flight during the fence
        __threadfence();
        __syncwarp();
        unsigned long long tk = 0;
        if (t.push_down() || t.push_up()) {                      // halo rows are out: tell the neighbouring rank(s)
            __threadfence_system();
            __syncwarp();
            if (lane == 0) {
                if (t.push_down()) st_release_sys(a.down_flag + strip, a.seq + (unsigned long long)ps + 1ull);
                if (t.push_up()) st_release_sys(a.up_flag + strip, a.seq + (unsigned long long)ps + 1ull);
            }
        }
        // Two-level closing.  The LAST task of a band (chunk) sums that band's partials, in the fixed order of the
        // per-iteration kernel (warp_ordered_sum over the strips), and only then takes a ticket of the pass; the last
        // BAND closes the pass.  (r2: one ticket per task, and the closing warp summed all n_tasks partials of the pass --
        // 25-90 us of dependent L2 round trips per pass, on one warp, with the closes of successive passes in series: at
        // 2048^2 that chain, not the sweep, set the pace of an iteration; profiles/r3s_task_timing.txt.)
        // The band tickets and band values are double-buffered by pass parity: no task of pass ps+2 starts before pass ps
        // is closed.
        unsigned long long* band_ticket = pa.claim + 8 + (size_t)(ps & 1) * a.n_chunks + chunk;
        if (lane == 0) {
            st_release_gpu(pa.done + tslot, pa.flag_base + (unsigned long long)ps + 1ull);
            tk = atomicAdd(band_ticket, 1ull);
        }
        tk = __shfl_sync(FULLM, tk, 0);
        id = __shfl_sync(FULLM, next_id, 0);
        SOR_TT(5);
#ifdef SOR_TASK_TIMING
        if (lane == 0) atomicAdd(&g_task_t[7], 1ull);
#endif
        if (tk != (unsigned long long)a.n_strips - 1ull) continue;
        double* bands = pa.bandvals2 + (size_t)(ps & 1) * NSB_SOR_KMAX * (a.n_chunks + 1);
        __threadfence();
        if (lane == 0) *band_ticket = 0ull;
        if (!pa.noclose) {
#pragma unroll
            for (int q = 0; q < K; q++) {
                const double v = warp_ordered_sum(part + (size_t)q * a.n_tasks + (size_t)chunk * a.n_strips, a.n_strips, lane);
                if (lane == 0) bands[(size_t)q * (a.n_chunks + 1) + chunk] = v;
            }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) tk = atomicAdd(pa.claim + 1 + (ps & 3), 1ull);
        tk = __shfl_sync(FULLM, tk, 0);
        if (tk == (unsigned long long)a.n_chunks - 1ull) {
            // this warp closes pass `ps`: the bands in the fixed global order, per sub-iteration
            __threadfence();
            if (lane == 0) pa.claim[1 + (ps & 3)] = 0ull;
            __syncwarp();
            // iterations close in order; nothing closes (and no reduction round is spent) after the one
            // that converged -- every rank sees the same residual bits, so every rank decides the same
            int skip = pa.noclose;
            if (lane == 0 && !skip) {
                const unsigned long long t0 = global_timer_ns();
                while (ld_acquire_gpu_s32(&a.ctl->iters) < it_first) {
                    if (a.check && ld_acquire_gpu_s32(&a.ctl->converged)) break;
                    if (global_timer_ns() - t0 > spin_timeout_ns(a.err)) { *((volatile int*)a.err) = 1; skip = 1; break; }
                    __nanosleep(64);
                }
                if (a.check && ld_acquire_gpu_s32(&a.ctl->converged)) skip = 1;
            }
            skip = __shfl_sync(FULLM, skip, 0);
#pragma unroll
            for (int q = 0; q < K; q++) {
                if (skip) break;
                const double* bq = bands + (size_t)q * (a.n_chunks + 1);
                double tot;
                if ((MULTI && (a.world > 1))) {
                    const unsigned long long round = pa.round0 + (unsigned long long)(ps * K + q);
                    slab_red_contribute_warp(a.red, round, 1, bq, lane);
                    slab_red_finish_sum_warp(a.red, round, lane, &tot);
                } else {
                    tot = warp_ordered_sum(bq, a.n_chunks, lane);
                }
                int conv = 0;
                if (lane == 0) {
                    // (no copy to the pinned-host twin: nobody polls it during this kernel, nsb_ph_sor_end reads the
                    //  device block after the launch -- a system-scope fence per iteration is not free)
                    sor_close_iteration(tot, a.ctl, it_first + q, a.k.cells, false);
                    conv = (a.check && a.ctl->converged) ? 1 : 0;
                }
                skip = __shfl_sync(FULLM, conv, 0);          // the later sub-iterations of a converged pass are not closed
            }
            __syncwarp();
            SOR_TT(6);
        }
    }
}

typedef void (*SorStreamFn)(const SorStreamArgs);
typedef void (*SorPersistFn)(const SorPersistArgs);
static SorStreamFn sor_stream_fn(bool recip, bool multi) {
    return recip ? (multi ? sor_stream_kernel<true, true> : sor_stream_kernel<true, false>)
                 : (multi ? sor_stream_kernel<false, true> : sor_stream_kernel<false, false>);
}
static SorPersistFn sor_persist_fn(bool recip, bool multi, int k) {
    if (k == 2)
        return recip ? (multi ? sor_persist_kernel<true, true, 2> : sor_persist_kernel<true, false, 2>)
                     : (multi ? sor_persist_kernel<false, true, 2> : sor_persist_kernel<false, false, 2>);
    return recip ? (multi ? sor_persist_kernel<true, true, 1> : sor_persist_kernel<true, false, 1>)
                 : (multi ? sor_persist_kernel<false, true, 1> : sor_persist_kernel<false, false, 1>);
}

// y-slabs: only when every rank sits on a GPU of its own -- the kernels of neighbouring ranks wait
// for each other, so they must be able to run at the same time (a 1-GPU box emulates the ranks with
// the one-launch-per-iteration kernel in lock step instead).
int nsb_sor_persist_supported(const nsb_ctx* c) {
    if (c->chunk_rows < 4) return 0;
    return c->world == 1 || (c->connected && c->peers_distinct);
}

// Sub-iterations per pass of the persistent sweep for this context: nsb_sor_plan decided (c->sor_k; nsb_options.sor_pass_k
// 1 / 2 or NSB_SOR_K force it, 0 = by the cost model).  Falls back to 1 where the chunks or the slab are too short for
// the 5-row halo.
int nsb_sor_pass_k(const nsb_ctx* c) {
    int k = c->sor_k >= 1 && c->sor_k <= NSB_SOR_KMAX ? c->sor_k : 1;
    if (c->chunk_rows < 2 * SorK<2>::HALO) k = 1;
    if (c->world > 1 && c->g.nj < 2 * SorK<2>::HALO) k = 1;
    return k;
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
// `slots` = resident warps of the sweep kernel on the device, recip = reciprocal-product arithmetic,
// k = sub-iterations per pass.  *cost_out = modelled cost of ONE PASS in step units (see nsb_plan_sor).
static int plan_chunk_rows(int ni, int nj, long long slots, int recip, int k, double* cost_out) {
    const long long n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    if (slots < 1) slots = 1;
    int best = 10;
    double best_cost = 1e300;
    // A task costs its blocks of 8 steps (rows + halo + drain, rounded up) plus `fill` steps of claim / flag / pipeline-fill
    // latency; an iteration (pass) costs the tasks per resident warp times that, but never less than `floor` tasks:
    // below it there is nothing to overlap the dependency waits between two passes with.
    //   k = 1, reciprocal product (profiles/r2j_chunk_sweep.txt): floor 1.6, fill 12; when ONE wave of warps covers an
    //     iteration (<= 1 task per warp) the latencies are exposed (factor 1.25).  Measured optima / picks, rows:
    //     1024^2 10 / 10, 2048^2 26 / 22, 3072^2 34 / 26, 4096^2 50 / 50, 5120^2 66-98 / 74, 6144^2 114-128 / 106, 8192^2 128 / 128.
    //   k = 1, IEEE division (instruction bound, the balance between warps counts for more; profiles/r1h_chunk_sweep2.txt):
    //     floor 2.6 -- 4096^2 32-48 rows, 5120^2 48, 6144^2 64-80, 7168^2 90, 8191^2 112.
    //   k = 2 (profiles/r2n_k2_sweep.txt; 8 resident warps per SM instead of 12): floor 3.6, fill 19, at most 124 rows
    //     (128 rows + 11 need an 18th block).  Optima / picks: 4096^2 44 / 36, 5120^2 60 / 52, 6144^2 60-76 / 76,
    //     7168^2 92 / 92, 8192^2 116-124 / 124, 16384^2 124 / 124.
    const double floor_tasks = k >= 2 ? 3.6 : (recip ? 1.6 : 2.6);
    const int fill = k >= 2 ? 19 : 12;
    const int drain = 3 * (k >= 2 ? 2 : 1) + (2 * (k >= 2 ? 2 : 1) + 1);        // last residual lag + halo rows below
    const int rows_max = k >= 2 ? 124 : 128;
    for (int rows = 10; rows <= rows_max; rows += 2) {
        const long long chunks = (nj + rows - 1) / rows;
        const double per_warp = (double)(chunks * n_strips) / (double)slots;
        const int steps = (rows + drain + SOR_BLK - 1) / SOR_BLK * SOR_BLK + fill;
        double a = per_warp > floor_tasks ? per_warp : floor_tasks;
        if (k == 1 && recip && per_warp <= 1.0) a = 1.25;
        const double cost = a * steps;
        if (cost < best_cost - 1e-9) { best_cost = cost; best = rows; }
        if (rows >= nj) break;
    }
    if (cost_out) *cost_out = best_cost;
    return best;
}
extern "C" int nsb_plan_chunk_rows_k(int ni, int nj, long long slots, int recip, int k) { return plan_chunk_rows(ni, nj, slots, recip, k, NULL); }
extern "C" int nsb_plan_chunk_rows(int ni, int nj, long long slots, int recip) { return plan_chunk_rows(ni, nj, slots, recip, 1, NULL); }

// Iterations per pass AND chunk height: the two-iteration pass halves the DRAM traffic but runs 8 instead of 12 warps per
// SM and needs ~2.3 x the instructions per pass, so it wins where the one-iteration pass is HBM-bound -- from ~5000^2 up
// (8192^2: 192 against 251 us per iteration, 16384^2: 704 against 985).  One model step costs 0.685 us with k = 1 and
// 0.34 us per ITERATION with k = 2 (both calibrated on 4096^2 .. 8192^2, profiles/r2n_k2_sweep.txt).  IEEE division
// (instruction-bound already) stays with one iteration per pass: 8191^2 309 against 480 us.
// slots1 / slots2 = resident warps of the k = 1 / k = 2 kernels.  Returns k, *rows = chunk height.
extern "C" int nsb_plan_sor(int ni, int nj, long long slots1, long long slots2, int recip, int* rows) {
    double c1 = 0.0, c2 = 0.0;
    const int r1 = plan_chunk_rows(ni, nj, slots1, recip, 1, &c1);
    const int r2 = plan_chunk_rows(ni, nj, slots2, recip, 2, &c2);
    const bool two = recip && nj >= 4 * SorK<2>::HALO && 0.34 * c2 < 0.97 * 0.685 * c1;
    if (rows) *rows = two ? r2 : r1;
    return two ? 2 : 1;
}

// Resident warps of the persistent sweep kernels on this context's device (slots[1], slots[2]: one / two iterations per
// pass).  Function attributes are per device: they are set for the device of every context that is planned.
static void sor_slots(const nsb_ctx* c, long long (&slots)[3]) {
    for (int v = 0; v < 4; v++) {
        cudaFuncSetAttribute((const void*)sor_stream_fn(v & 1, v & 2), cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES(1));
        cudaFuncSetAttribute((const void*)sor_persist_fn(v & 1, v & 2, 1), cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES(1));
        cudaFuncSetAttribute((const void*)sor_persist_fn(v & 1, v & 2, 2), cudaFuncAttributeMaxDynamicSharedMemorySize, SOR_SMEM_BYTES(2));
        // two blocks of the RHS-ring variant need 2 x 102 KB: ask for the largest shared-memory carve-out (the default
        // carve-out fitted ONE block: 5 warps per SM, profiles/r2w)
        cudaFuncSetAttribute((const void*)sor_persist_fn(v & 1, v & 2, 2), cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    }
    slots[0] = 0;
    for (int k = 1; k <= 2; k++) {
        int bps = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (const void*)sor_persist_fn(c->use_recip, c->world > 1, k), SOR_KTHREADS(k), SOR_SMEM_BYTES(k));
        if (bps < 1) bps = 1;
        slots[k] = (long long)c->sm_count * bps * SOR_KWARPS(k);
    }
}

// Iterations per pass the caller asked for: nsb_options.sor_pass_k, overridden by NSB_SOR_K; 0 = by the model.
static int sor_wanted_k(const nsb_ctx* c) {
    int want = c->O.sor_pass_k >= 1 && c->O.sor_pass_k <= NSB_SOR_KMAX ? c->O.sor_pass_k : 0;
    const char* env = getenv("NSB_SOR_K");
    if (env && atoi(env) >= 1 && atoi(env) <= NSB_SOR_KMAX) want = atoi(env);
    return want;
}

// y-slabs: the band height is part of the decomposition, so it is planned BEFORE the rows are dealt out, for a slab of
// ceil(j_max / world) rows -- the same arithmetic on every rank (same grid, same world size, same kind of GPU).  r2 used
// 124-row bands for every decomposition: right for 8192-row slabs, but 16384^2 on 8 GPUs (2048-row slabs, 17 bands) then
// left the two-iteration pass with 2 tasks per warp and the model fell back to one iteration per pass
// (profiles/r3m_scale_summary.txt: 154 us per iteration, 0.56 of 8 x the single GPU).
// Returns the band height; *k_out = iterations per pass that go with it.
int nsb_sor_preplan_slabs(const nsb_ctx* c, int ni, int j_max, int world, int* k_out) {
    long long slots[3];
    sor_slots(c, slots);
    const int nj_slab = (j_max + world - 1) / world;
    const int want = sor_wanted_k(c);
    int rows = 0, k = want;
    if (k == 0) k = nsb_plan_sor(ni, nj_slab, slots[1], slots[2], c->use_recip ? 1 : 0, &rows);
    else rows = plan_chunk_rows(ni, nj_slab, slots[k], c->use_recip ? 1 : 0, k, NULL);
    if (rows < 16) rows = 16;
    while (rows > 16 && (j_max + rows - 1) / rows < world) rows -= 2;      // every slab needs a band
    if (k_out) *k_out = k;
    return rows;
}

void nsb_sor_plan(nsb_ctx* c) {
    const int ni = c->g.ni, nj = c->g.nj;
    c->n_strips = (ni + NSB_STRIP_COLS - 1) / NSB_STRIP_COLS;
    long long slots[3];
    sor_slots(c, slots);
    const int want = sor_wanted_k(c);
    int rows = 0, k = want;
    if (c->O.sor_chunk_rows > 0) {
        rows = (c->O.sor_chunk_rows + 1) & ~1;
        // the caller fixed the chunk height (y-slabs: it is part of the decomposition): choose k for THAT height.
        // Every rank of a decomposition must decide alike: the inputs are the bands of the LARGEST slab (ranks
        // 0 .. rem-1 hold one band more) and the rows of the SMALLEST (the ragged last band may leave a few rows only).
        if (k == 0) {
            const long long bands = ((long long)c->J + rows - 1) / rows;
            const long long bands_rank = (bands + c->world - 1) / c->world;
            const long long tasks = (long long)c->n_strips * bands_rank;
            k = (c->use_recip && rows >= 64 && (double)tasks / (double)slots[2] >= 3.0) ? 2 : 1;
        }
    } else if (k == 0) {
        k = nsb_plan_sor(ni, nj, slots[1], slots[2], c->use_recip ? 1 : 0, &rows);
    } else {
        rows = plan_chunk_rows(ni, nj, slots[k], c->use_recip ? 1 : 0, k, NULL);
    }
    if (rows < 2) rows = 2;
    if (c->world > 1 && k > 1) {
        // two iterations per pass push 5 halo rows: every slab must own at least twice that
        int min_nj = c->g.nj;
        for (int r = 0; r < c->world; r++) {
            int nj_r = 0;
            if (nsb_slab_partition(c->J, rows, c->world, r, NULL, &nj_r, NULL, NULL) == NSB_OK && nj_r < min_nj) min_nj = nj_r;
        }
        if (min_nj < 2 * SorK<2>::HALO) k = 1;
    }
    c->sor_k = k;
    c->chunk_rows = rows;
    c->n_chunks = (nj + rows - 1) / rows;
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
    sor_stream_fn(c->use_recip, c->world > 1)<<<blocks, SOR_THREADS, SOR_SMEM_BYTES(1), c->stream>>>(a);
    return NSB_LAUNCHED(c, 1);
}

// One launch of the persistent sweep: `n_pass` passes of `k` iterations each.  Pass 0 reads pbuf[buf_in] (0..2) and the
// buffers rotate; `it0` = number of the first iteration inside the current solve.  noclose: replay of an iteration that
// is already closed.  wait_first (y-slabs): the halo rows of pbuf[buf_in] are still being pushed by the neighbours'
// previous launch of this solve.
cudaError_t nsb_launch_sor_persist(nsb_ctx* c, int k, int n_pass, int it0, int check, int buf_in, int noclose, int wait_first) {
    if (n_pass <= 0) return cudaSuccess;
    SorPersistArgs pa;
    memset(&pa, 0, sizeof pa);
    SorStreamArgs& a = pa.s;
    a.pin = c->pbuf[buf_in]; a.pout = c->pbuf[(buf_in + 1) % 3]; a.rhs = c->rhs;
    a.pitch = c->g.pitch; a.ni = c->g.ni; a.nj = c->g.nj;
    a.chunk_rows = c->chunk_rows; a.n_strips = c->n_strips; a.n_tasks = c->n_partials; a.n_chunks = c->n_chunks;
    a.n_groups = (int)(c->g.pitch / NSB_GROUP);
    a.it = it0; a.check = check; a.zero = 0u;
    a.dbg_fixed_loads = getenv("NSB_DEBUG_FIXED_LOADS") ? 1 : 0;
    a.k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->J);     // main.cu:588: the GLOBAL cell count
    a.partials = c->d_partials; a.bandvals = c->d_bandvals; a.ctl = c->d_ctl;
    a.world = c->world; a.wall_bottom = c->wall_bottom; a.wall_top = c->wall_top;
    a.err = &c->d_sync->error;
    a.seq = wait_first ? c->seq_tail : c->seq;
    a.red = nsb_slab_red(c);
    pa.round0 = c->red_round + (unsigned long long)it0;
    pa.buf0 = buf_in;
    pa.noclose = noclose; pa.wait_first = wait_first;
    for (int q = 0; q < 3; q++) { pa.own_p[q] = c->pbuf[q]; pa.down_p[q] = c->pbuf[q]; pa.up_p[q] = c->pbuf[q]; }
    a.down_nj = 0;
    a.flag_from_down = c->d_sorflag; a.flag_from_up = c->d_sorflag + c->n_strips;
    a.down_flag = c->d_sorflag; a.up_flag = c->d_sorflag;
    if (c->world > 1) {
        if (!c->connected) return cudaErrorNotReady;
        if (!c->wall_bottom) {
            const SlabPeer& d = c->peer[c->rank - 1];
            for (int q = 0; q < 3; q++) pa.down_p[q] = d.p[q];
            a.down_nj = d.nj;
            a.down_flag = d.sorflag + c->n_strips;             // I am its UPPER neighbour
        }
        if (!c->wall_top) {
            const SlabPeer& u = c->peer[c->rank + 1];
            for (int q = 0; q < 3; q++) pa.up_p[q] = u.p[q];
            a.up_flag = u.sorflag;                             // I am its LOWER neighbour
        }
    }
    // halo epochs: pass q of this launch publishes a.seq + q + 1.  A launch that continues the previous one
    // (wait_first) starts at the epoch its last pass published; any other launch starts above everything published.
    c->seq_tail = a.seq + (unsigned long long)n_pass;
    c->seq = c->seq_tail + 1ull;
    pa.n_pass = n_pass;
    pa.claim = c->d_claim; pa.done = c->d_done; pa.partials2 = c->d_partials2; pa.bandvals2 = c->d_bandvals2;
    pa.flag_base = c->flag_base;
    c->flag_base += (unsigned long long)n_pass + 1ull;
    int bps = 0;
    // (developer probe NSB_DEBUG_FORCE_MULTI: the multi-slab instantiation on a single slab -- what the extra code costs without any exchange)
    const SorPersistFn kern = sor_persist_fn(c->use_recip, c->world > 1 || getenv("NSB_DEBUG_FORCE_MULTI") != NULL, k);
    const int kw = SOR_KWARPS(k);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, (const void*)kern, SOR_KTHREADS(k), SOR_SMEM_BYTES(k));
    if (bps < 1) bps = 1;
    const long long want = ((long long)n_pass * a.n_tasks + kw - 1) / kw;
    long long blocks = (long long)c->sm_count * bps;
    if (blocks > want) blocks = want;
    // at most two passes are in flight at any time: more warps than that would only poll flags
    const long long useful = (2LL * a.n_tasks + kw - 1) / kw + 1;
    if (blocks > useful) blocks = useful;
    // SMs that serve the edge-strip queue first (sor_claim): where the edge strips run bodies of their own -- two iterations
    // per pass (lane-masked instantiations) and IEEE division (general edge body) --, with a full grid of resident blocks
    // and enough strips for the split to matter.  How many: the edge strips' share of the work with an edge task costing
    // w = 2.5 interior ones (it runs ~2.2 x as long even on an SM of its own), plus two -- too few starve every pass
    // (8192^2, two iterations per pass: 6 SMs 281 us, 10 SMs 192, 13 SMs 173, 16 SMs 176, none 194;
    //  IEEE division: 8191^2 318 -> 286 us with 13, 6143^2 206 -> 177 with 16, 4100^2 126 -> 100 with 22; profiles/r4b_edge_sms.txt).
    // NSB_EDGE_SMS=<n> forces a number, 0 switches the split off.
    pa.edge_sms = 0;
    {
        int edge = 0;
        const char* env = getenv("NSB_EDGE_SMS");
        if (env) edge = atoi(env);
        else if (k >= 2 || !c->use_recip) edge = (int)ceil((double)c->sm_count * 5.0 / ((double)(a.n_strips - 2) + 5.0)) + 2;
        if (edge > c->sm_count / 2) edge = c->sm_count / 2;
        if (edge > 0 && a.n_strips >= 8 && blocks == (long long)c->sm_count * bps) pa.edge_sms = edge;
    }
    cudaError_t e = cudaMemsetAsync(c->d_claim, 0, (8 + 2 * (size_t)c->n_chunks) * sizeof(unsigned long long), c->stream);   // claim counters, pass tickets, band tickets
    if (e != cudaSuccess) return e;
    kern<<<(int)blocks, SOR_KTHREADS(k), SOR_SMEM_BYTES(k), c->stream>>>(pa);
    return NSB_LAUNCHED(c, 1);
}
