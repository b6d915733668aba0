/*
 * ns_oracle.c -- CPU restatement of the reference hot path (see ns_oracle.h).
 * TEST INFRASTRUCTURE ONLY: never linked into the product.
 *
 * Every function cites the reference lines it restates (paths relative to
 * the reference root).  Expressions keep the reference's association order so
 * that, compiled without FMA contraction (gcc x86-64 default), results are
 * bit-identical to the reference serial build; tests/test_oracle.py checks it.
 */
#include "ns_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define AT(m, i, j) ((m)[(size_t)(i) * ld + (j)])

/* ------------------------------------------------------------------ params */

/* src/serial/io.c:12-59: one value per line, first token, rest of line ignored.
 * Line order: problem f i_max j_max a b T Re g_x g_y tau omega epsilon max_it n_print. */
int orc_read_params(const char* path, orc_params* o)
{
    FILE* fp = fopen(path, "r");
    if (!fp) return -1;
    char line[256];
    double* dslots[15] = {0, &o->f, 0, 0, &o->a, &o->b, &o->T, &o->Re, &o->g_x, &o->g_y,
                          &o->tau, &o->omega, &o->epsilon, 0, 0};
    int* islots[15] = {&o->problem, 0, &o->i_max, &o->j_max, 0, 0, 0, 0, 0, 0, 0, 0, 0,
                       &o->max_it, &o->n_print};
    memset(o, 0, sizeof *o);
    for (int k = 0; k < 15; k++) {
        if (!fgets(line, sizeof line, fp)) break;   /* the reference ignores short files too */
        if (islots[k]) sscanf(line, "%d", islots[k]);
        else sscanf(line, "%lf", dslots[k]);
    }
    fclose(fp);
    return 0;
}

/* ------------------------------------------------------------ dt reductions */

/* src/serial/io.c:122-139 */
double orc_max_mat(int i_max, int j_max, const double* m)
{
    const int ld = j_max + 2;
    double best = AT(m, 0, 0);
    for (int i = 1; i <= i_max; i++)
        for (int j = 1; j <= j_max; j++)
            if (best < AT(m, i, j)) best = AT(m, i, j);
    return best;
}

/* src/serial/io.c:141-161 with num == 3 (main.c:91) */
double orc_min3(double x, double y, double z)
{
    double m = x;
    if (m > y) m = y;
    if (m > z) m = z;
    return m;
}

/* src/serial/main.c:91-92 */
void orc_dt_gamma(const orc_params* P, double dx, double dy, double u_max, double v_max,
                  double* dt, double* gamma)
{
    double visc = P->Re / 2.0 / (1.0 / dx / dx + 1.0 / dy / dy);
    *dt = P->tau * orc_min3(visc, dx / fabs(u_max), dy / fabs(v_max));
    *gamma = fmax(u_max * *dt / dx, v_max * *dt / dy);
}

/* ------------------------------------------------------ boundary conditions */

/* src/serial/boundaries.c:7-39 */
int orc_set_inflow(int i_max, int j_max, double* u, double* v, int side, double u_fix, double v_fix)
{
    const int ld = j_max + 2;
    if (side == ORC_TOP) {
        for (int i = 1; i <= i_max; i++) {
            AT(v, i, j_max) = v_fix;
            AT(u, i, j_max + 1) = 2 * u_fix - AT(u, i, j_max);
        }
    } else if (side == ORC_BOTTOM) {
        for (int i = 1; i <= i_max; i++) {
            AT(v, i, 0) = v_fix;
            AT(u, i, 0) = 2 * u_fix - AT(u, i, 1);
        }
    } else if (side == ORC_LEFT) {
        for (int j = 1; j <= j_max; j++) {
            AT(u, 0, j) = u_fix;
            AT(v, 0, j) = 2 * v_fix - AT(v, 1, j);
        }
    } else if (side == ORC_RIGHT) {
        for (int j = 1; j <= j_max; j++) {
            AT(u, i_max, j) = u_fix;
            AT(v, i_max + 1, j) = 2 * v_fix - AT(v, i_max, j);
        }
    } else {
        return -1;
    }
    return 0;
}

/* src/serial/boundaries.c:3-5 */
int orc_set_noslip(int i_max, int j_max, double* u, double* v, int side)
{
    return orc_set_inflow(i_max, j_max, u, v, side, 0, 0);
}

/* src/serial/main.c:95-108 */
int orc_apply_bc(const orc_params* P, double t, double* u, double* v)
{
    double lid;
    if (P->problem == 1) lid = 1.0;
    else if (P->problem == 2) lid = sin(P->f * t);
    else return -1;
    orc_set_noslip(P->i_max, P->j_max, u, v, ORC_LEFT);
    orc_set_noslip(P->i_max, P->j_max, u, v, ORC_RIGHT);
    orc_set_noslip(P->i_max, P->j_max, u, v, ORC_BOTTOM);
    orc_set_inflow(P->i_max, P->j_max, u, v, ORC_TOP, lid, 0.0);
    return 0;
}

/* ----------------------------------------------------------------- stencils */

/* src/serial/integration.c:7-15 */
double orc_du2_dx(const double* u, int ld, int i, int j, double dx, double gamma)
{
    double ue = 0.5 * (AT(u, i, j) + AT(u, i + 1, j));
    double uw = 0.5 * (AT(u, i - 1, j) + AT(u, i, j));
    double de = fabs(ue) * 0.5 * (AT(u, i, j) - AT(u, i + 1, j));
    double dw = fabs(uw) * 0.5 * (AT(u, i - 1, j) - AT(u, i, j));
    return 1 / dx * (ue * ue - uw * uw) + gamma / dx * (de - dw);
}

/* src/serial/integration.c:17-27 */
double orc_duv_dy(const double* u, const double* v, int ld, int i, int j, double dy, double gamma)
{
    double vn = 0.5 * (AT(v, i, j) + AT(v, i + 1, j));
    double vs = 0.5 * (AT(v, i, j - 1) + AT(v, i + 1, j - 1));
    double fn = vn * 0.5 * (AT(u, i, j) + AT(u, i, j + 1));
    double fs = vs * 0.5 * (AT(u, i, j - 1) + AT(u, i, j));
    double dn = fabs(vn) * 0.5 * (AT(u, i, j) - AT(u, i, j + 1));
    double ds = fabs(vs) * 0.5 * (AT(u, i, j - 1) - AT(u, i, j));
    return 1 / dy * (fn - fs) + gamma / dy * (dn - ds);
}

/* src/serial/integration.c:29-37 */
double orc_dv2_dy(const double* v, int ld, int i, int j, double dy, double gamma)
{
    double vn = 0.5 * (AT(v, i, j) + AT(v, i, j + 1));
    double vs = 0.5 * (AT(v, i, j - 1) + AT(v, i, j));
    double dn = fabs(vn) * 0.5 * (AT(v, i, j) - AT(v, i, j + 1));
    double ds = fabs(vs) * 0.5 * (AT(v, i, j - 1) - AT(v, i, j));
    return 1 / dy * (vn * vn - vs * vs) + gamma / dy * (dn - ds);
}

/* src/serial/integration.c:39-51 */
double orc_duv_dx(const double* u, const double* v, int ld, int i, int j, double dx, double gamma)
{
    double ue = 0.5 * (AT(u, i, j) + AT(u, i, j + 1));
    double uw = 0.5 * (AT(u, i - 1, j) + AT(u, i - 1, j + 1));
    double fe = ue * 0.5 * (AT(v, i, j) + AT(v, i + 1, j));
    double fw = uw * 0.5 * (AT(v, i - 1, j) + AT(v, i, j));
    double de = fabs(ue) * 0.5 * (AT(v, i, j) - AT(v, i + 1, j));
    double dw = fabs(uw) * 0.5 * (AT(v, i - 1, j) - AT(v, i, j));
    return 1 / dx * (fe - fw) + gamma / dx * (de - dw);
}

/* src/serial/integration.c:57-59, 65-67 */
double orc_lap_x(const double* m, int ld, int i, int j, double dx)
{
    return (AT(m, i + 1, j) - 2 * AT(m, i, j) + AT(m, i - 1, j)) / (dx * dx);
}

/* src/serial/integration.c:61-63, 69-71 */
double orc_lap_y(const double* m, int ld, int i, int j, double dy)
{
    return (AT(m, i, j + 1) - 2 * AT(m, i, j) + AT(m, i, j - 1)) / (dy * dy);
}

/* ------------------------------------------------------------ F, G, RHS, uv */

/* src/serial/integration.c:73-96.  F(0,.), F(i_max,.), G(.,0), G(.,j_max) are left untouched
 * (zero from allocation) exactly like the reference. */
void orc_FG(double* F, double* G, const double* u, const double* v, int i_max, int j_max,
            double Re, double g_x, double g_y, double dt, double dx, double dy, double gamma)
{
    const int ld = j_max + 2;
    for (int i = 1; i <= i_max; i++) {
        for (int j = 1; j <= j_max; j++) {
            if (i <= i_max - 1) {
                AT(F, i, j) = AT(u, i, j) +
                    dt * (1.0 / Re * (orc_lap_x(u, ld, i, j, dx) + orc_lap_y(u, ld, i, j, dy)) -
                          orc_du2_dx(u, ld, i, j, dx, gamma) -
                          orc_duv_dy(u, v, ld, i, j, dy, gamma) +
                          g_x);
            }
            if (j <= j_max - 1) {
                AT(G, i, j) = AT(v, i, j) +
                    dt * (1.0 / Re * (orc_lap_x(v, ld, i, j, dx) + orc_lap_y(v, ld, i, j, dy)) -
                          orc_duv_dx(u, v, ld, i, j, dx, gamma) -
                          orc_dv2_dy(v, ld, i, j, dy, gamma) +
                          g_y);
            }
        }
    }
}

/* src/serial/main.c:116-120 */
void orc_RHS(double* RHS, const double* F, const double* G, int i_max, int j_max,
             double dt, double dx, double dy)
{
    const int ld = j_max + 2;
    for (int i = 1; i <= i_max; i++)
        for (int j = 1; j <= j_max; j++)
            AT(RHS, i, j) = 1.0 / dt * ((AT(F, i, j) - AT(F, i - 1, j)) / dx +
                                        (AT(G, i, j) - AT(G, i, j - 1)) / dy);
}

/* src/serial/main.c:131-136 with dp_dx / dp_dy from integration.c:101-110 */
void orc_update_uv(double* u, double* v, const double* F, const double* G, const double* p,
                   int i_max, int j_max, double dt, double dx, double dy)
{
    const int ld = j_max + 2;
    for (int i = 1; i <= i_max; i++) {
        for (int j = 1; j <= j_max; j++) {
            if (i <= i_max - 1)
                AT(u, i, j) = AT(F, i, j) - dt * ((AT(p, i + 1, j) - AT(p, i, j)) / dx);
            if (j <= j_max - 1)
                AT(v, i, j) = AT(G, i, j) - dt * ((AT(p, i, j + 1) - AT(p, i, j)) / dy);
        }
    }
}

/* ---------------------------------------------------------------------- SOR */

/* src/serial/integration.c:115-124 */
double orc_L2(const double* m, int i_max, int j_max)
{
    const int ld = j_max + 2;
    double acc = 0.0;
    for (int i = 1; i <= i_max; i++)
        for (int j = 1; j <= j_max; j++)
            acc += AT(m, i, j) * AT(m, i, j);
    return sqrt(acc / i_max / j_max);
}

/* Neumann ghost refresh: src/serial/integration.c:138-146 (and main.cu:249-273 for the
 * non-corner points; the corner copies there are racy and never read). */
static void refresh_ghosts(double* p, int i_max, int j_max)
{
    const int ld = j_max + 2;
    for (int j = 1; j <= j_max; j++) {
        AT(p, 0, j) = AT(p, 1, j);
        AT(p, i_max + 1, j) = AT(p, i_max, j);
    }
    for (int i = 1; i <= i_max; i++) {
        AT(p, i, 0) = AT(p, i, 1);
        AT(p, i, j_max + 1) = AT(p, i, j_max);
    }
}

/* src/serial/integration.c:129-173 */
int orc_SOR_lex(double* p, int i_max, int j_max, double dx, double dy, double* res,
                const double* RHS, double omega, double eps, int max_it,
                int* iters, double* last_res, double* norm_p0)
{
    const int ld = j_max + 2;
    double dydy = dy * dy;
    double dxdx = dx * dx;
    int it = 0;
    double norm_p = orc_L2(p, i_max, j_max);
    double r = 0.0;
    if (norm_p0) *norm_p0 = norm_p;
    while (it < max_it) {
        refresh_ghosts(p, i_max, j_max);
        for (int i = 1; i <= i_max; i++)
            for (int j = 1; j <= j_max; j++)
                AT(p, i, j) = (1.0 - omega) * AT(p, i, j) +
                    omega / (2.0 * (1.0 / dxdx + 1.0 / dydy)) *
                    ((AT(p, i + 1, j) + AT(p, i - 1, j)) / dxdx +
                     (AT(p, i, j + 1) + AT(p, i, j - 1)) / dydy - AT(RHS, i, j));
        for (int i = 1; i <= i_max; i++)
            for (int j = 1; j <= j_max; j++)
                AT(res, i, j) = (AT(p, i + 1, j) - 2.0 * AT(p, i, j) + AT(p, i - 1, j)) / dxdx +
                                (AT(p, i, j + 1) - 2.0 * AT(p, i, j) + AT(p, i, j - 1)) / dydy -
                                AT(RHS, i, j);
        r = orc_L2(res, i_max, j_max);
        if (r <= eps * (norm_p + 1.5)) {
            if (iters) *iters = it + 1;
            if (last_res) *last_res = r;
            return 0;
        }
        it++;
    }
    if (iters) *iters = it;
    if (last_res) *last_res = r;
    return -1;
}

/* one colour of src/parallel/main.cu:384-511 (update expression 491-506) */
static void rb_sweep(double* p, int i_max, int j_max, double dx2, double dy2,
                     const double* RHS, double omega, int color)
{
    const int ld = j_max + 2;
    double coeff = 2.0 * (1.0 / dx2 + 1.0 / dy2);
    for (int i = 1; i <= i_max; i++) {
        for (int j = 1; j <= j_max; j++) {
            if ((i + j) % 2 != color) continue;
            double p_old = AT(p, i, j);
            AT(p, i, j) = (1.0 - omega) * p_old +
                omega / coeff *
                ((AT(p, i + 1, j) + AT(p, i - 1, j)) / dx2 +
                 (AT(p, i, j + 1) + AT(p, i, j - 1)) / dy2 -
                 AT(RHS, i, j));
        }
    }
}

/* src/parallel/main.cu:515-561 + 564-590: sum of squared residuals, sqrt(sum/(i_max*j_max)).
 * The CUDA reference sums in a block-tree order; this restatement sums in index order, which
 * moves the result by O(1e-16) relative -- only the convergence test sees it. */
static double rb_residual(const double* p, int i_max, int j_max, double dx2, double dy2,
                          const double* RHS)
{
    const int ld = j_max + 2;
    double acc = 0.0;
    for (int i = 1; i <= i_max; i++) {
        for (int j = 1; j <= j_max; j++) {
            double r = (AT(p, i + 1, j) - 2.0 * AT(p, i, j) + AT(p, i - 1, j)) / dx2 +
                       (AT(p, i, j + 1) - 2.0 * AT(p, i, j) + AT(p, i, j - 1)) / dy2 -
                       AT(RHS, i, j);
            acc += r * r;
        }
    }
    return sqrt(acc / ((double)i_max * (double)j_max));  /* main.cu:588 (int product there) */
}

/* src/parallel/main.cu:684-703 for one `it` */
double orc_SOR_rb_iter(double* p, int i_max, int j_max, double dx, double dy,
                       const double* RHS, double omega)
{
    double dx2 = dx * dx, dy2 = dy * dy;
    refresh_ghosts(p, i_max, j_max);                          /* :686 */
    rb_sweep(p, i_max, j_max, dx2, dy2, RHS, omega, 0);       /* :690 */
    refresh_ghosts(p, i_max, j_max);                          /* :694 */
    rb_sweep(p, i_max, j_max, dx2, dy2, RHS, omega, 1);       /* :698 */
    return rb_residual(p, i_max, j_max, dx2, dy2, RHS);       /* :702-706, ghosts of :694 */
}

/* src/parallel/main.cu:656-726 */
int orc_SOR_rb(double* p, int i_max, int j_max, double dx, double dy,
               const double* RHS, double omega, double eps, int max_it,
               int* iters, double* last_res, double* norm_p0)
{
    const int ld = j_max + 2;
    double acc = 0.0;                                         /* :675 -> :593-653, :588 */
    for (int i = 1; i <= i_max; i++)
        for (int j = 1; j <= j_max; j++)
            acc += AT(p, i, j) * AT(p, i, j);
    double norm_p = sqrt(acc / ((double)i_max * (double)j_max));
    double r = 0.0;
    if (norm_p0) *norm_p0 = norm_p;
    for (int it = 0; it < max_it; it++) {
        r = orc_SOR_rb_iter(p, i_max, j_max, dx, dy, RHS, omega);
        if (r <= eps * (norm_p + 1.5)) {                      /* :713 */
            if (iters) *iters = it + 1;
            if (last_res) *last_res = r;
            return it + 1;
        }
    }
    if (iters) *iters = max_it;
    if (last_res) *last_res = r;
    return -1;                                                /* :725 */
}

/* All-cores variant for the CPU baseline: identical per-cell arithmetic (each colour sweep is
 * order-independent), rows of i distributed over threads; residual summed per i-row, then rows
 * added in index order. */
int orc_SOR_rb_omp(double* p, int i_max, int j_max, double dx, double dy,
                   const double* RHS, double omega, double eps, int max_it,
                   int* iters, double* last_res, int nthreads)
{
    const int ld = j_max + 2;
    double dx2 = dx * dx, dy2 = dy * dy;
    double coeff = 2.0 * (1.0 / dx2 + 1.0 / dy2);
    double* rowsum = (double*)malloc(sizeof(double) * (size_t)(i_max + 2));
    double acc = 0.0, r = 0.0;
    int done = -1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
    for (int i = 1; i <= i_max; i++)
        for (int j = 1; j <= j_max; j++)
            acc += AT(p, i, j) * AT(p, i, j);
    double norm_p = sqrt(acc / ((double)i_max * (double)j_max));
    for (int it = 0; it < max_it && done < 0; it++) {
        for (int color = 0; color < 2; color++) {
            refresh_ghosts(p, i_max, j_max);
#pragma omp parallel for schedule(static)
            for (int i = 1; i <= i_max; i++) {
                for (int j = 1 + ((i + 1 + color) & 1); j <= j_max; j += 2) {
                    double p_old = AT(p, i, j);
                    AT(p, i, j) = (1.0 - omega) * p_old +
                        omega / coeff *
                        ((AT(p, i + 1, j) + AT(p, i - 1, j)) / dx2 +
                         (AT(p, i, j + 1) + AT(p, i, j - 1)) / dy2 -
                         AT(RHS, i, j));
                }
            }
        }
#pragma omp parallel for schedule(static)
        for (int i = 1; i <= i_max; i++) {
            double s = 0.0;
            for (int j = 1; j <= j_max; j++) {
                double q = (AT(p, i + 1, j) - 2.0 * AT(p, i, j) + AT(p, i - 1, j)) / dx2 +
                           (AT(p, i, j + 1) - 2.0 * AT(p, i, j) + AT(p, i, j - 1)) / dy2 -
                           AT(RHS, i, j);
                s += q * q;
            }
            rowsum[i] = s;
        }
        acc = 0.0;
        for (int i = 1; i <= i_max; i++) acc += rowsum[i];
        r = sqrt(acc / ((double)i_max * (double)j_max));
        if (r <= eps * (norm_p + 1.5)) done = it + 1;
    }
    free(rowsum);
    if (iters) *iters = done > 0 ? done : max_it;
    if (last_res) *last_res = r;
    return done > 0 ? done : -1;
}

/* ---------------------------------------------------------------- time loop */

/* src/serial/main.c:86-147 (sor_order 0) / src/parallel/main.cu:1046-1103 (sor_order 1) */
int orc_run(const orc_params* P, int sor_order, int max_steps,
            double* u, double* v, double* p, double* F, double* G, double* RHS,
            orc_step_log* log, int max_log, double* t_end, double* sor_seconds)
{
    const int i_max = P->i_max, j_max = P->j_max;
    double dx = P->a / i_max;                                 /* main.c:72-73 */
    double dy = P->b / j_max;
    double* res = NULL;
    double t = 0.0, sor_s = 0.0;
    int n = 0;
    if (sor_order == 0) res = (double*)calloc((size_t)(i_max + 2) * (j_max + 2), sizeof(double));
    while (t < P->T && (max_steps <= 0 || n < max_steps)) {
        double u_max = orc_max_mat(i_max, j_max, u);
        double v_max = orc_max_mat(i_max, j_max, v);
        double dt, gamma, norm_p = 0.0, r = 0.0;
        int its = 0, ret;
        orc_dt_gamma(P, dx, dy, u_max, v_max, &dt, &gamma);
        if (orc_apply_bc(P, t, u, v) != 0) { free(res); return -1; }
        orc_FG(F, G, u, v, i_max, j_max, P->Re, P->g_x, P->g_y, dt, dx, dy, gamma);
        orc_RHS(RHS, F, G, i_max, j_max, dt, dx, dy);
        clock_t c0 = clock();
        if (sor_order == 0)
            ret = orc_SOR_lex(p, i_max, j_max, dx, dy, res, RHS, P->omega, P->epsilon, P->max_it,
                              &its, &r, &norm_p);
        else
            ret = orc_SOR_rb(p, i_max, j_max, dx, dy, RHS, P->omega, P->epsilon, P->max_it,
                             &its, &r, &norm_p);
        sor_s += (double)(clock() - c0) / CLOCKS_PER_SEC;
        orc_update_uv(u, v, F, G, p, i_max, j_max, dt, dx, dy);
        if (log && n < max_log) {
            log[n].t = t; log[n].dt = dt; log[n].u_max = u_max; log[n].v_max = v_max;
            log[n].gamma = gamma; log[n].norm_p = norm_p; log[n].res = r;
            log[n].sor_ret = ret; log[n].sor_iters = its;
        }
        t += dt;
        n++;
    }
    free(res);
    if (t_end) *t_end = t;
    if (sor_seconds) *sor_seconds = sor_s;
    return n;
}

/* ------------------------------------------------------------- field output */

/* src/serial/io.c:61-120.  u omits column i_max+1, v omits row j_max+1 (its newline is kept). */
int orc_output(int i_max, int j_max, const double* u, const double* v, const double* p,
               double t, double a, double b, const char* prefix)
{
    const int ld = j_max + 2;
    char name[3][512];
    FILE* fp[3];
    const char* suffix[3] = {"_u.txt", "_v.txt", "_p.txt"};
    for (int k = 0; k < 3; k++) {
        snprintf(name[k], sizeof name[k], "%s%s", prefix, suffix[k]);
        fp[k] = fopen(name[k], "w");
    }
    if (!fp[0] || !fp[1] || !fp[2]) {
        for (int k = 0; k < 3; k++) if (fp[k]) fclose(fp[k]);
        return -1;
    }
    for (int k = 0; k < 3; k++) fprintf(fp[k], "%.5f\n%.5f\n%.5f\n", t, a, b);
    for (int j = 0; j < j_max + 2; j++) {
        for (int i = 0; i < i_max + 2; i++) {
            if (i < i_max + 1) fprintf(fp[0], "%.5f ", AT(u, i, j));
            if (j < j_max + 1) fprintf(fp[1], "%.5f ", AT(v, i, j));
            fprintf(fp[2], "%.5f ", AT(p, i, j));
        }
        for (int k = 0; k < 3; k++) fprintf(fp[k], "\n");
    }
    for (int k = 0; k < 3; k++) fclose(fp[k]);
    return 0;
}
