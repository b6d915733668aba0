/*
 * ns_oracle.h -- CPU restatement of the lid-driven-cavity time step.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the CUDA library
 * libnsb200.so, the C driver, the python host mirror) may include, link or
 * call this.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or as
 * the timed CPU baseline.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle.py)
 *   - against the reference's own golden vectors tests/der{,2,3,4}.txt
 *     (committed gzip'd under tests/golden/), and
 *   - bit-for-bit against the unmodified reference serial sources compiled
 *     in place into oracle/_ref/ (oracle/Makefile), function by function and
 *     end to end, plus committed fixtures generated from that build
 *     (tests/golden/make_golden.py).
 *
 * Layout: every grid is a flat row-major array of (i_max+2)*(j_max+2)
 * doubles, element (i,j) at [i*(j_max+2)+j]  (i = x index, j = y index; the
 * reference's `double**` grids have the same index meaning,
 * src/serial/memory.c:3-26).  u's row i_max+1 and v's column j_max+1 exist
 * here but are never touched, like the cells the reference does not allocate.
 */
#ifndef NS_ORACLE_H
#define NS_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_TOP = 0, ORC_BOTTOM = 1, ORC_LEFT = 2, ORC_RIGHT = 3 }; /* boundaries.h:16-21 */

typedef struct {
    int problem;        /* 1 = lid-driven cavity, 2 = oscillating lid   */
    double f;           /* lid frequency for problem 2                  */
    int i_max, j_max;
    double a, b, T, Re, g_x, g_y, tau, omega, epsilon;
    int max_it, n_print;
} orc_params;

/* per-time-step record written by orc_run (the log the GPU path must reproduce) */
typedef struct {
    double t, dt, u_max, v_max, gamma, norm_p, res;
    int sor_ret;        /* what the SOR variant returned                */
    int sor_iters;      /* iterations actually performed                */
} orc_step_log;

/* io.c:12-59  -- 15 positional lines, first token of each */
int orc_read_params(const char* path, orc_params* out);

/* io.c:122-139 -- SIGNED max over interior, seeded with m(0,0) */
double orc_max_mat(int i_max, int j_max, const double* m);
/* io.c:141-161 specialised to the one call site main.c:91 */
double orc_min3(double x, double y, double z);
/* main.c:91-92 */
void orc_dt_gamma(const orc_params* P, double dx, double dy, double u_max, double v_max,
                  double* dt, double* gamma);

/* boundaries.c:3-39 */
int orc_set_inflow(int i_max, int j_max, double* u, double* v, int side, double u_fix, double v_fix);
int orc_set_noslip(int i_max, int j_max, double* u, double* v, int side);
/* main.c:95-108 : LEFT, RIGHT, BOTTOM, TOP(lid).  returns -1 for unknown problem */
int orc_apply_bc(const orc_params* P, double t, double* u, double* v);

/* integration.c:7-71 -- the 8 stencils */
double orc_du2_dx(const double* u, int ld, int i, int j, double dx, double gamma);
double orc_duv_dy(const double* u, const double* v, int ld, int i, int j, double dy, double gamma);
double orc_dv2_dy(const double* v, int ld, int i, int j, double dy, double gamma);
double orc_duv_dx(const double* u, const double* v, int ld, int i, int j, double dx, double gamma);
double orc_lap_x(const double* m, int ld, int i, int j, double dx);   /* d2u_dx2 / d2v_dx2 */
double orc_lap_y(const double* m, int ld, int i, int j, double dy);   /* d2u_dy2 / d2v_dy2 */

/* integration.c:73-96 */
void orc_FG(double* F, double* G, const double* u, const double* v, int i_max, int j_max,
            double Re, double g_x, double g_y, double dt, double dx, double dy, double gamma);
/* main.c:116-120 */
void orc_RHS(double* RHS, const double* F, const double* G, int i_max, int j_max,
             double dt, double dx, double dy);
/* integration.c:115-124 */
double orc_L2(const double* m, int i_max, int j_max);
/* integration.c:129-173 : lexicographic sweep.  returns 0 / -1; *iters = sweeps done */
int orc_SOR_lex(double* p, int i_max, int j_max, double dx, double dy, double* res,
                const double* RHS, double omega, double eps, int max_it,
                int* iters, double* last_res, double* norm_p0);
/* src/parallel/main.cu:656-726 order: ghosts, colour 0, ghosts, colour 1, residual, test.
 * returns it+1 / -1 like the CUDA reference; *iters = sweeps done */
int orc_SOR_rb(double* p, int i_max, int j_max, double dx, double dy,
               const double* RHS, double omega, double eps, int max_it,
               int* iters, double* last_res, double* norm_p0);
/* one red-black iteration only (ghosts, c0, ghosts, c1) + residual norm; for per-iteration parity */
double orc_SOR_rb_iter(double* p, int i_max, int j_max, double dx, double dy,
                       const double* RHS, double omega);
/* main.c:131-136 */
void orc_update_uv(double* u, double* v, const double* F, const double* G, const double* p,
                   int i_max, int j_max, double dt, double dx, double dy);

/* whole time loop main.c:86-147.  sor_order: 0 = lexicographic (serial), 1 = red-black (CUDA).
 * fields are caller-allocated, zero-initialised.  logs up to max_log steps.
 * max_steps <= 0 means "until t >= T".  returns number of steps taken, -1 on unknown problem. */
int orc_run(const orc_params* P, int sor_order, int max_steps,
            double* u, double* v, double* p, double* F, double* G, double* RHS,
            orc_step_log* log, int max_log, double* t_end, double* sor_seconds);

/* io.c:61-120 -- writes <prefix>_{u,v,p}.txt; returns 0, -1 if a file cannot be opened */
int orc_output(int i_max, int j_max, const double* u, const double* v, const double* p,
               double t, double a, double b, const char* prefix);

/* multi-threaded (OpenMP) red-black iterations for the all-cores CPU baseline.  Same arithmetic
 * as orc_SOR_rb; the residual sum is reduced per row then over rows in row order. */
int orc_SOR_rb_omp(double* p, int i_max, int j_max, double dx, double dy,
                   const double* RHS, double omega, double eps, int max_it,
                   int* iters, double* last_res, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
