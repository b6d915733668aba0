// momentum.cu -- everything of the time step that is not the pressure solve:
//   boundary conditions (boundaries.c:7-39 ; set_*_kernel main.cu:838-925),
//   F and G (integration.c:7-96 ; calculate_F/G_kernel main.cu:289-382),
//   Poisson right-hand side (main.c:116-120 ; calculate_RHS_kernel main.cu:219-229),
//   velocity update (main.c:131-136 ; update_velocities_kernel main.cu:231-245) fused with the
//   reductions the NEXT step needs: signed max u, max v (io.c:122-139 ; find_max_kernel
//   main.cu:729-796) and sum p^2 (integration.c:115-124 ; calculate_norm_kernel main.cu:593-622).
//
// All kernels run x-fastest over the pitched grids, one thread per cell, threadIdx.x along i:
// a warp touches 256 contiguous bytes per array (the reference maps threadIdx.x to the strided
// index).  Arithmetic order is the reference's, evaluated with non-contracting
// intrinsics (nsb_math.cuh), so F, G, RHS, u, v are bit-identical to the CPU build.
#include <string.h>

#include "nsb_internal.h"
#include "nsb_math.cuh"
#include "slab.cuh"

#define MB_X 128
#define MB_Y 2

struct MomCoef {
    double dt, gamma;
    double inv_dx, inv_dy;       // 1/delta_x, 1/delta_y            integration.c:14,26
    double g_dx, g_dy;           // gamma/delta_x, gamma/delta_y
    double dx2, dy2, rdx2, rdy2; // delta_x*delta_x ...              integration.c:58
    float qlo_x, qlo_y;          // IEEE division: rdx2 / rdy2 are the refined reciprocals and these the quotient windows (nsb_math.cuh)
    double dx, dy, rdx, rdy;
    double inv_re;               // 1.0/Re                          integration.c:79
    double inv_dt;               // 1.0/delta_t                     main.c:118
    double g_x, g_y;
};

static MomCoef make_mom(const nsb_ctx* c, double dt, double gamma) {
    MomCoef m;
    m.dt = dt; m.gamma = gamma;
    m.dx = c->dx; m.dy = c->dy;
    m.inv_dx = 1 / c->dx; m.inv_dy = 1 / c->dy;
    m.g_dx = gamma / c->dx; m.g_dy = gamma / c->dy;
    m.dx2 = c->dx * c->dx; m.dy2 = c->dy * c->dy;
    m.rdx2 = 1.0 / m.dx2; m.rdy2 = 1.0 / m.dy2;
    m.qlo_x = INFINITY; m.qlo_y = INFINITY;
    m.rdx = 1.0 / c->dx; m.rdy = 1.0 / c->dy;
    m.inv_re = 1.0 / c->P.Re;
    m.inv_dt = 1.0 / dt;
    m.g_x = c->P.g_x; m.g_y = c->P.g_y;
    return m;
}

// ------------------------------------------------------------------ boundary conditions
// boundaries.c:7-39, one side per launch exactly like the reference's call sequence.
__global__ void set_inflow_kernel(double* u, double* v, long long pitch, int ni, int nj, int side,
                                  double u_fix, double v_fix) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x + 1;
    double* U = u + (long long)NSB_ROW0 * pitch + NSB_COL0;
    double* V = v + (long long)NSB_ROW0 * pitch + NSB_COL0;
    if (side == NSB_TOP) {
        if (t <= ni) {
            V[(long long)nj * pitch + t] = v_fix;
            U[(long long)(nj + 1) * pitch + t] = __dsub_rn(__dmul_rn(2.0, u_fix), U[(long long)nj * pitch + t]);
        }
    } else if (side == NSB_BOTTOM) {
        if (t <= ni) {
            V[t] = v_fix;
            U[t] = __dsub_rn(__dmul_rn(2.0, u_fix), U[pitch + t]);
        }
    } else if (side == NSB_LEFT) {
        if (t <= nj) {
            U[(long long)t * pitch] = u_fix;
            V[(long long)t * pitch] = __dsub_rn(__dmul_rn(2.0, v_fix), V[(long long)t * pitch + 1]);
        }
    } else {
        if (t <= nj) {
            U[(long long)t * pitch + ni] = u_fix;
            V[(long long)t * pitch + ni + 1] = __dsub_rn(__dmul_rn(2.0, v_fix), V[(long long)t * pitch + ni]);
        }
    }
}

cudaError_t nsb_launch_set_inflow(nsb_ctx* c, int side, double u_fix, double v_fix) {
    if (side == NSB_TOP && !c->wall_top) return cudaSuccess;          // the lid belongs to the last slab
    if (side == NSB_BOTTOM && !c->wall_bottom) return cudaSuccess;
    const int n = (side == NSB_TOP || side == NSB_BOTTOM) ? c->g.ni : c->g.nj;
    set_inflow_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(c->u, c->v, c->g.pitch, c->g.ni, c->g.nj,
                                                              side, u_fix, v_fix);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------ F and G
// Neighbour naming: c = (i,j), e = (i+1,j), w = (i-1,j), n = (i,j+1), s = (i,j-1).
template <bool RECIP>
__device__ __forceinline__ double lap(double c, double e, double w, double n, double s, const MomCoef& m) {
    // d2/dx2 + d2/dy2, integration.c:57-71:  (e - 2*c + w) / (dx*dx)  +  (n - 2*c + s) / (dy*dy)
    const double two_c = __dmul_rn(2.0, c);
    const double nx = __dadd_rn(__dsub_rn(e, two_c), w), ny = __dadd_rn(__dsub_rn(n, two_c), s);
    if (RECIP) return __dadd_rn(__dmul_rn(nx, m.rdx2), __dmul_rn(ny, m.rdy2));
    // IEEE division by dx^2, dy^2 as in the sweeps: three operations with the refined reciprocal, the real division
    // when a quotient leaves their window (r2: __ddiv_rn inline, ~10 FP64 operations and a MUFU per quotient, four
    // per cell -- 777 us against 467 us for the reciprocal product at 8191^2 / 8192^2)
    bool ok = true;
    double lx = nsb_div_try(nx, m.dx2, m.rdx2, m.qlo_x, ok);
    double ly = nsb_div_try(ny, m.dy2, m.rdy2, m.qlo_y, ok);
    if (!ok) { lx = nsb_div_full(nx, m.dx2); ly = nsb_div_full(ny, m.dy2); }
    return __dadd_rn(lx, ly);
}

// integration.c:7-15 (du2_dx) and :29-37 (dv2_dy): same shape along one axis.
//   a = 0.5(c + p), b = 0.5(q + c);  1/h (a*a - b*b) + gamma/h (|a| 0.5 (c - p) - |b| 0.5 (q - c))
__device__ __forceinline__ double donor_self(double c, double p, double q, double inv_h, double g_h) {
    double a = __dmul_rn(0.5, __dadd_rn(c, p));
    double b = __dmul_rn(0.5, __dadd_rn(q, c));
    double da = __dmul_rn(__dmul_rn(fabs(a), 0.5), __dsub_rn(c, p));
    double db = __dmul_rn(__dmul_rn(fabs(b), 0.5), __dsub_rn(q, c));
    return __dadd_rn(__dmul_rn(inv_h, __dsub_rn(__dmul_rn(a, a), __dmul_rn(b, b))),
                     __dmul_rn(g_h, __dsub_rn(da, db)));
}

// integration.c:17-27 (duv_dy for F) and :39-51 (duv_dx for G): transported quantity `x` by the
// cross velocity.  t1a+t1b is the cross-velocity pair on the far face, t2a+t2b on the near face;
// xc the cell value, xp the far neighbour, xq the near neighbour.
__device__ __forceinline__ double donor_cross(double t1a, double t1b, double t2a, double t2b, double xc,
                                              double xp, double xq, double inv_h, double g_h) {
    double a = __dmul_rn(0.5, __dadd_rn(t1a, t1b));
    double b = __dmul_rn(0.5, __dadd_rn(t2a, t2b));
    double fa = __dmul_rn(__dmul_rn(a, 0.5), __dadd_rn(xc, xp));
    double fb = __dmul_rn(__dmul_rn(b, 0.5), __dadd_rn(xq, xc));
    double da = __dmul_rn(__dmul_rn(fabs(a), 0.5), __dsub_rn(xc, xp));
    double db = __dmul_rn(__dmul_rn(fabs(b), 0.5), __dsub_rn(xq, xc));
    return __dadd_rn(__dmul_rn(inv_h, __dsub_rn(fa, fb)), __dmul_rn(g_h, __dsub_rn(da, db)));
}

// ---- F and G of one cell from the 3 x 3 neighbourhoods of u and v (index 0 / 1 / 2 = column i-1 / i / i+1; s, c, n =
// row j-1 / j / j+1), with the y-face terms of the two donor-cell fluxes that run along y carried from row to row.
//
// The flux through the face between rows j and j+1 enters row j as its far face (a, fa, da below) and row j+1 as its near
// face (b, fb, db in donor_self / donor_cross): the same operands in the same operations, additions commuted at most --
// the same bits.  A thread that marches up a column therefore computes each y-face once: 90 instead of 104 FP64
// operations per cell (integration.c:17-37 evaluates both faces for every cell), and 6 instead of 18 loads.
struct FgFaces {
    double f_fa, f_da;       // duv_dy (F):  (a * 0.5) * (u_j + u_j+1),  (|a| * 0.5) * (u_j - u_j+1),  a = 0.5 (v_i + v_i+1) on row j
    double g_aa, g_da;       // dv2_dy (G):  a * a,                      (|a| * 0.5) * (v_j - v_j+1),  a = 0.5 (v_j + v_j+1)
};

// faces between the rows `lo` (j) and `hi` (j+1)
__device__ __forceinline__ FgFaces fg_faces(const double (&u_lo)[3], const double (&u_hi)[3], const double (&v_lo)[3],
                                            const double (&v_hi)[3]) {
    FgFaces f;
    const double a = __dmul_rn(0.5, __dadd_rn(v_lo[1], v_lo[2]));
    f.f_fa = __dmul_rn(__dmul_rn(a, 0.5), __dadd_rn(u_lo[1], u_hi[1]));
    f.f_da = __dmul_rn(__dmul_rn(fabs(a), 0.5), __dsub_rn(u_lo[1], u_hi[1]));
    const double g = __dmul_rn(0.5, __dadd_rn(v_lo[1], v_hi[1]));
    f.g_aa = __dmul_rn(g, g);
    f.g_da = __dmul_rn(__dmul_rn(fabs(g), 0.5), __dsub_rn(v_lo[1], v_hi[1]));
    return f;
}

// near = faces below row j (from the previous step), returns F and G of (i, j) and leaves the faces above row j in `near`
template <bool RECIP>
__device__ __forceinline__ void fg_cell(const double (&us)[3], const double (&uc)[3], const double (&un)[3],
                                        const double (&vs)[3], const double (&vc)[3], const double (&vn)[3],
                                        FgFaces& near, const MomCoef& m, double& Fo, double& Go) {
    const FgFaces far = fg_faces(uc, un, vc, vn);
    {   // integration.c:73-82
        const double visc = __dmul_rn(m.inv_re, lap<RECIP>(uc[1], uc[2], uc[0], un[1], us[1], m));
        const double conv_x = donor_self(uc[1], uc[2], uc[0], m.inv_dx, m.g_dx);                    // du2_dx
        const double conv_y = __dadd_rn(__dmul_rn(m.inv_dy, __dsub_rn(far.f_fa, near.f_fa)),        // duv_dy
                                        __dmul_rn(m.g_dy, __dsub_rn(far.f_da, near.f_da)));
        const double rhs = __dadd_rn(__dsub_rn(__dsub_rn(visc, conv_x), conv_y), m.g_x);
        Fo = __dadd_rn(uc[1], __dmul_rn(m.dt, rhs));
    }
    {   // integration.c:84-90
        const double visc = __dmul_rn(m.inv_re, lap<RECIP>(vc[1], vc[2], vc[0], vn[1], vs[1], m));
        const double conv_x = donor_cross(uc[1], un[1], uc[0], un[0], vc[1], vc[2], vc[0], m.inv_dx, m.g_dx);   // duv_dx
        const double conv_y = __dadd_rn(__dmul_rn(m.inv_dy, __dsub_rn(far.g_aa, near.g_aa)),        // dv2_dy
                                        __dmul_rn(m.g_dy, __dsub_rn(far.g_da, near.g_da)));
        const double rhs = __dadd_rn(__dsub_rn(__dsub_rn(visc, conv_x), conv_y), m.g_y);
        Go = __dadd_rn(vc[1], __dmul_rn(m.dt, rhs));
    }
    near = far;
}

#ifndef FG_BX
#define FG_BX 128           // columns per block, one per thread
#endif
#ifndef FG_ROWS
#define FG_ROWS 3           // rows marched by a thread (a multiple of 3: the three-row window rotates by renaming)
#endif

__device__ __forceinline__ void fg_load3(double (&x)[3], const double* __restrict__ p) {
    x[0] = p[-1]; x[1] = p[0]; x[2] = p[1];
}

// One thread per column i, marching FG_ROWS rows upwards (r1/r2: one thread per cell, 18 loads and both faces of every
// flux per cell -- 586 us at 8192^2, FP64 pipe at 64 %, profiles/r2u_ncu_summary.txt).  Measured at 8192^2
// (profiles/r3_fg_variants.txt, tools/mom_probe.py): 24 rows 463 us, 12 rows 443, 6 rows 427-431, 3 rows 408-425; 256
// columns per block +60 us; more registers per thread are slower at every height (56 -> 72 by a register cap: +40 us;
// 78 with the next row's loads issued a step ahead: +45 us) -- the kernel wants resident warps, not longer marches:
// with 3 rows it still shares 2 of 3 y-faces (94 operations per cell) and loads 10 values per cell.
template <bool RECIP>
__global__ void __launch_bounds__(FG_BX) fg_kernel(const double* __restrict__ u, const double* __restrict__ v,
                                                    double* __restrict__ F, double* __restrict__ G, long long pitch,
                                                    int ni, int nj, int g_hi, MomCoef m) {
    const int i = blockIdx.x * FG_BX + threadIdx.x + 1;
    const int j0 = blockIdx.y * FG_ROWS + 1;
    if (i > ni) return;
    const int jend = min(j0 + FG_ROWS - 1, nj);
    long long o = (long long)(j0 + NSB_ROW0) * pitch + NSB_COL0 + i;            // (i, j0)
    double ua[3], ub[3], uc[3], va[3], vb[3], vc[3];
    fg_load3(ua, u + o - pitch); fg_load3(va, v + o - pitch);
    fg_load3(ub, u + o); fg_load3(vb, v + o);
    fg_load3(uc, u + o + pitch); fg_load3(vc, v + o + pitch);
    FgFaces fc = fg_faces(ua, ub, va, vb);
    const bool f_col = (i <= ni - 1);                                           // integration.c:76
#define FG_STEP(US, UC, UN, VS, VC, VN, J)                                                          \
    {                                                                                               \
        double Fo, Go;                                                                              \
        fg_cell<RECIP>(US, UC, UN, VS, VC, VN, fc, m, Fo, Go);                                      \
        if (f_col) F[o] = Fo;                                                                       \
        if ((J) <= g_hi) G[o] = Go;                          /* global j <= j_max-1 (integration.c:84) */ \
        o += pitch;                                                                                 \
        if ((J) + 1 > jend) break;                                                                  \
        fg_load3(US, u + o + pitch); fg_load3(VS, v + o + pitch);   /* row J+2 replaces row J-1 */  \
    }
    for (int j = j0;; j += 3) {
        FG_STEP(ua, ub, uc, va, vb, vc, j)
        FG_STEP(ub, uc, ua, vb, vc, va, j + 1)
        FG_STEP(uc, ua, ub, vc, va, vb, j + 2)
    }
#undef FG_STEP
}

cudaError_t nsb_launch_fg(nsb_ctx* c, double dt, double gamma) {
    MomCoef m = make_mom(c, dt, gamma);
    dim3 b(FG_BX), g((c->g.ni + FG_BX - 1) / FG_BX, (c->g.nj + FG_ROWS - 1) / FG_ROWS);
    const int g_hi = c->wall_top ? c->g.nj - 1 : c->g.nj;
    if (!c->use_recip) {
        const SorCoef k = nsb_make_sor_coef(c, 0);           // refined reciprocals of dx^2, dy^2 (cached per geometry)
        m.rdx2 = k.rdx2; m.rdy2 = k.rdy2; m.qlo_x = k.qlo_x; m.qlo_y = k.qlo_y;
    }
    if (c->use_recip) fg_kernel<true><<<g, b, 0, c->stream>>>(c->u, c->v, c->F, c->G, c->g.pitch, c->g.ni, c->g.nj, g_hi, m);
    else fg_kernel<false><<<g, b, 0, c->stream>>>(c->u, c->v, c->F, c->G, c->g.pitch, c->g.ni, c->g.nj, g_hi, m);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------ RHS (main.c:116-120)
//   RHS = 1.0/dt * ((F - F_w)/dx + (G - G_s)/dy)
template <bool RECIPD>
__device__ __forceinline__ double rhs_cell(double f, double fw, double g, double gs, const MomCoef& m) {
    double a = nsb_div<RECIPD>(__dsub_rn(f, fw), m.dx, m.rdx);
    double b = nsb_div<RECIPD>(__dsub_rn(g, gs), m.dy, m.rdy);
    return __dmul_rn(m.inv_dt, __dadd_rn(a, b));
}

template <bool RECIPD>
__global__ void __launch_bounds__(MB_X* MB_Y) rhs_kernel(const double* __restrict__ F, const double* __restrict__ G,
                                                          double* __restrict__ rhs, long long pitch, int ni, int nj,
                                                          MomCoef m) {
    const int i = blockIdx.x * MB_X + threadIdx.x + 1;
    const int j = blockIdx.y * MB_Y + threadIdx.y + 1;
    if (i > ni || j > nj) return;
    const long long o = (long long)(j + NSB_ROW0) * pitch + NSB_COL0 + i;
    rhs[o] = rhs_cell<RECIPD>(F[o], F[o - 1], G[o], G[o - pitch], m);
}

// exactness of 1/dx, 1/dy as reciprocals is decided on the host (nsb_ctx::recip_exact covers
// dx2, dy2; dx, dy powers of two follow from that only when the exponents are even, so the host
// passes a separate flag)
static bool pow2(double x) {
    if (!(x > 0.0)) return false;
    int e;
    return frexp(x, &e) == 0.5;
}

cudaError_t nsb_launch_rhs(nsb_ctx* c, double dt) {
    MomCoef m = make_mom(c, dt, 0.0);
    dim3 b(MB_X, MB_Y), g((c->g.ni + MB_X - 1) / MB_X, (c->g.nj + MB_Y - 1) / MB_Y);
    const bool rd = (c->O.div_mode == NSB_DIV_RECIP) ||
                    (c->O.div_mode == NSB_DIV_AUTO && pow2(c->dx) && pow2(c->dy));
    if (rd) rhs_kernel<true><<<g, b, 0, c->stream>>>(c->F, c->G, c->rhs, c->g.pitch, c->g.ni, c->g.nj, m);
    else rhs_kernel<false><<<g, b, 0, c->stream>>>(c->F, c->G, c->rhs, c->g.pitch, c->g.ni, c->g.nj, m);
    return NSB_LAUNCHED(c, 1);
}

// F, G, then RHS.  (A fused single-pass kernel -- 128 x 16 tile, F and G parked in shared memory with a
// recomputed halo column / row, 40 instead of 56 B per cell -- was measured at 928 us against 589 + 278 us
// for the two kernels at 8192^2: the stencils are instruction-issue bound (~200 instructions per cell
// in the non-contracting reference arithmetic), not HBM bound, and the fused body's register pressure
// halves the occupancy.  profiles/r1g_ncu_summary.txt.)
cudaError_t nsb_launch_fg_rhs(nsb_ctx* c, double dt, double gamma) {
    cudaError_t e = nsb_launch_fg(c, dt, gamma);
    if (e != cudaSuccess) return e;
    return nsb_launch_rhs(c, dt);
}

// ------------------------------------------------------------------ velocity update + reductions
//   u = F - dt * ((pE - p)/dx)   i <= i_max-1        main.c:133 with dp_dx integration.c:101-103
//   v = G - dt * ((pN - p)/dy)   j <= j_max-1        main.c:134 with dp_dy integration.c:108-110
// and, over the interior after the update: signed max u, signed max v (seeded with m[0][0] like
// io.c:124), sum p^2.
struct UvRedArgs {
    double* partials;        // [3][blocks]
    double* bandvals;        // [3][n_chunks]
    StepRed* red;
    SlabRed sr;
    unsigned long long round;
    int chunk_rows, n_chunks, v_hi, seed00, world;
};

// One thread: the step reductions are complete.
__device__ __forceinline__ void uv_red_store(StepRed* red, const double* out) {
    red->u_max = out[0];
    red->v_max = out[1];
    red->sum_p2 = out[2];
}

// One block per grid row (1-D grid: any number of rows), 256 threads, TWO adjacent cells per thread and pass:
// interior column i = 1 sits on a 32-byte boundary, so cells (i, i+1) with odd i are one aligned double2 of every
// array -- F, G, p, p one row up in, u, v out; 72 B in flight per thread and pass (r1: one cell per thread, five scalar
// loads, a block per quarter row: 4.16 TB/s with 39 stall cycles per issue on the loads, profiles/r1j_ncu_summary.txt).
#ifndef UV_THREADS
#define UV_THREADS 256
#endif
#ifndef UV_UNROLL
#define UV_UNROLL 2
#endif
#define UV_PRAGMA(x) _Pragma(#x)
#define UV_UNROLL_PRAGMA(n) UV_PRAGMA(unroll n)
template <bool RECIPD, bool UPDATE>
__global__ void __launch_bounds__(UV_THREADS) update_uv_kernel(double* __restrict__ u, double* __restrict__ v,
                                                         const double* __restrict__ F, const double* __restrict__ G,
                                                         const double* __restrict__ p, long long pitch, int ni,
                                                         int nj, MomCoef m, UvRedArgs ra) {
    __shared__ double sh[32];
    __shared__ int is_last;
    const int j = blockIdx.x + 1;
    const long long ro = (long long)(j + NSB_ROW0) * pitch + NSB_COL0;
    const bool v_row = (j <= ra.v_hi);                           // global j <= j_max-1 (main.c:134)
    double mu = -1.0e300, mv = -1.0e300, sp = 0.0;
    UV_UNROLL_PRAGMA(UV_UNROLL)
    for (int i = 1 + 2 * threadIdx.x; i <= ni; i += 2 * UV_THREADS) {
        const long long o = ro + i;
        const bool two = (i + 1 <= ni);                          // the second cell of the pair is an interior cell
        const double2 pc = *reinterpret_cast<const double2*>(p + o);
        double2 un, vn;
        if (UPDATE) {
            // u, v are only READ where the update does not reach (last column / last row): the rest
            // of the pass reads F, G, p and writes u, v -- 40 B per cell
            const double2 f = *reinterpret_cast<const double2*>(F + o);
            const double pe2 = p[o + 2];
            un.x = (i <= ni - 1) ? __dsub_rn(f.x, __dmul_rn(m.dt, nsb_div<RECIPD>(__dsub_rn(pc.y, pc.x), m.dx, m.rdx))) : u[o];
            un.y = (i + 1 <= ni - 1) ? __dsub_rn(f.y, __dmul_rn(m.dt, nsb_div<RECIPD>(__dsub_rn(pe2, pc.y), m.dx, m.rdx))) : (two ? u[o + 1] : 0.0);
            if (i + 1 <= ni - 1) *reinterpret_cast<double2*>(u + o) = un;
            else if (i <= ni - 1) u[o] = un.x;
            if (v_row) {
                const double2 g = *reinterpret_cast<const double2*>(G + o);
                const double2 pn = *reinterpret_cast<const double2*>(p + o + pitch);
                vn.x = __dsub_rn(g.x, __dmul_rn(m.dt, nsb_div<RECIPD>(__dsub_rn(pn.x, pc.x), m.dy, m.rdy)));
                vn.y = __dsub_rn(g.y, __dmul_rn(m.dt, nsb_div<RECIPD>(__dsub_rn(pn.y, pc.y), m.dy, m.rdy)));
                if (two) *reinterpret_cast<double2*>(v + o) = vn;
                else v[o] = vn.x;
            } else {
                vn.x = v[o]; vn.y = two ? v[o + 1] : 0.0;
            }
        } else {
            un.x = u[o]; vn.x = v[o];
            un.y = two ? u[o + 1] : 0.0; vn.y = two ? v[o + 1] : 0.0;
        }
        mu = fmax(mu, un.x); mv = fmax(mv, vn.x); sp = fma(pc.x, pc.x, sp);
        if (two) { mu = fmax(mu, un.y); mv = fmax(mv, vn.y); sp = fma(pc.y, pc.y, sp); }
    }
    const int nblk = gridDim.x;
    const int bid = blockIdx.x;
    double bu = block_max(mu, sh);
    double bv = block_max(mv, sh);
    double bs = block_sum(sp, sh);
    if (threadIdx.x == 0) {
        ra.partials[bid] = bu;
        ra.partials[nblk + bid] = bv;
        ra.partials[2 * nblk + bid] = bs;
        __threadfence();
        unsigned int t = atomicAdd(&ra.red->ticket, 1u);
        is_last = (t == (unsigned)nblk - 1);
    }
    __syncthreads();
    if (is_last) {
        // band values: one band = the rows of one SOR chunk (never split between slabs), its block
        // partials combined in a fixed order; the bands are then combined over ALL slabs in an order
        // that depends on the global band count only (slab.cuh)
        __threadfence();
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gx = 1;
        for (int b = warp; b < ra.n_chunks; b += UV_THREADS / 32) {
            const int r0 = b * ra.chunk_rows;
            const int rows = min(ra.chunk_rows, nj - r0);
            const double a = warp_ordered_max(ra.partials + (size_t)r0 * gx, rows * gx, lane);
            const double c = warp_ordered_max(ra.partials + nblk + (size_t)r0 * gx, rows * gx, lane);
            const double s = warp_ordered_sum(ra.partials + 2 * (size_t)nblk + (size_t)r0 * gx, rows * gx, lane);
            if (lane == 0) {
                double aa = a, cc = c;
                if (b == 0 && ra.seed00) {                       // io.c:124: the scan starts from m[0][0]
                    const double u00 = u[(long long)NSB_ROW0 * pitch + NSB_COL0];
                    const double v00 = v[(long long)NSB_ROW0 * pitch + NSB_COL0];
                    aa = (u00 < a) ? a : u00;
                    cc = (v00 < c) ? c : v00;
                }
                ra.bandvals[b] = aa;
                ra.bandvals[ra.n_chunks + b] = cc;
                ra.bandvals[2 * ra.n_chunks + b] = s;
            }
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) ra.red->ticket = 0;
        slab_red_contribute(ra.sr, ra.round, 3, ra.bandvals);
        if (ra.world == 1) {
            slab_red_finish(ra.sr, ra.round, 3, 0x3u, sh);
            if (threadIdx.x == 0) uv_red_store(ra.red, sh);
        }
    }
}

__global__ void __launch_bounds__(128) red_close_kernel(SlabRed sr, unsigned long long round, StepRed* red) {
    __shared__ double out[4];
    slab_red_finish(sr, round, 3, 0x3u, out);
    if (threadIdx.x == 0) uv_red_store(red, out);
}

cudaError_t nsb_launch_red_close(nsb_ctx* c) {
    red_close_kernel<<<1, 128, 0, c->stream>>>(nsb_slab_red(c), c->red_round, c->d_red);
    return NSB_LAUNCHED(c, 1);
}

static void uv_grid(const nsb_ctx* c, dim3& g) {
    g = dim3(c->g.nj);                                         // one block per row
}

static UvRedArgs uv_red_args(nsb_ctx* c) {
    UvRedArgs ra;
    memset(&ra, 0, sizeof ra);
    ra.partials = c->d_red_partials; ra.bandvals = c->d_bandvals; ra.red = c->d_red;
    ra.sr = nsb_slab_red(c);
    ra.round = c->red_round;
    ra.chunk_rows = c->chunk_rows; ra.n_chunks = c->n_chunks;
    ra.v_hi = c->wall_top ? c->g.nj - 1 : c->g.nj;
    ra.seed00 = c->wall_bottom ? 1 : 0;
    ra.world = c->world;
    return ra;
}

// Both launchers use reduction round c->red_round; the caller advances it (nsb_ph_*_finish).
cudaError_t nsb_launch_update_uv(nsb_ctx* c, double dt) {
    MomCoef m = make_mom(c, dt, 0.0);
    dim3 g;
    uv_grid(c, g);
    const bool rd = (c->O.div_mode == NSB_DIV_RECIP) ||
                    (c->O.div_mode == NSB_DIV_AUTO && pow2(c->dx) && pow2(c->dy));
    double* p = c->pbuf[c->pcur];
    UvRedArgs ra = uv_red_args(c);
    if (rd) update_uv_kernel<true, true><<<g, UV_THREADS, 0, c->stream>>>(c->u, c->v, c->F, c->G, p, c->g.pitch, c->g.ni, c->g.nj, m, ra);
    else update_uv_kernel<false, true><<<g, UV_THREADS, 0, c->stream>>>(c->u, c->v, c->F, c->G, p, c->g.pitch, c->g.ni, c->g.nj, m, ra);
    return NSB_LAUNCHED(c, 1);
}

cudaError_t nsb_launch_reduce_uvp(nsb_ctx* c) {
    MomCoef m = make_mom(c, 1.0, 0.0);
    dim3 g;
    uv_grid(c, g);
    update_uv_kernel<true, false><<<g, UV_THREADS, 0, c->stream>>>(c->u, c->v, c->F, c->G, c->pbuf[c->pcur], c->g.pitch,
                                                            c->g.ni, c->g.nj, m, uv_red_args(c));
    return NSB_LAUNCHED(c, 1);
}
