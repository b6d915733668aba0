// sor_simple.cu -- SOR bookkeeping kernels and the launch-per-phase validation engine.
//
// NSB_SOR_SIMPLE restates src/parallel/main.cu:684-706 literally with stored ghosts and one launch
// per phase (ghost refresh, colour 0, ghost refresh, colour 1, residual).  It exists to check the
// fused engines against an independent GPU implementation at sizes the CPU oracle cannot reach
// (tests/test_gpu_parity.py); it is coalesced (x fastest) but makes no attempt to be fast.
#include "nsb_internal.h"
#include <math.h>
#include <string.h>

#include "nsb_math.cuh"

// ---- SOR entry: threshold from the cached sum p^2 (replaces calculate_L2_norm_device,
//      main.cu:626-653 and the host test of main.cu:713)
__global__ void sor_begin_kernel(SorCtl* ctl, const StepRed* red, double eps, long long cells) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double norm_p = sqrt(red->sum_p2 / (double)cells);      // main.cu:588
        ctl->norm_p = norm_p;
        ctl->thresh = eps * (norm_p + 1.5);
        ctl->res = 0.0;
        ctl->sum_r2 = 0.0;
        ctl->iters = 0;
        ctl->converged = 0;
        ctl->ticket = 0;
        sor_publish(ctl);
    }
}

cudaError_t nsb_launch_sor_begin(nsb_ctx* c) {
    sor_begin_kernel<<<1, 32, 0, c->stream>>>(c->d_ctl, c->d_red, c->P.epsilon,
                                              (long long)c->g.ni * (long long)c->J);      // global cell count
    return NSB_LAUNCHED(c, 1);
}

// ---- Neumann ghosts, main.cu:249-273 (corner copies there are racy and never read: skipped)
__global__ void ghost_refresh_kernel(double* p, long long pitch, int ni, int nj, const SorCtl* ctl) {
    if (ctl && *((volatile const int*)&ctl->converged)) return;
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    double* base = p + (long long)NSB_ROW0 * pitch + NSB_COL0;
    if (t < nj) {
        int j = t + 1;
        base[(long long)j * pitch + 0] = base[(long long)j * pitch + 1];
        base[(long long)j * pitch + ni + 1] = base[(long long)j * pitch + ni];
    }
    if (t < ni) {
        int i = t + 1;
        base[i] = base[pitch + i];
        base[(long long)(nj + 1) * pitch + i] = base[(long long)nj * pitch + i];
    }
}

template <bool RECIP>
__global__ void sor_color_kernel(double* p, const double* __restrict__ rhs, long long pitch, int ni,
                                 int nj, int color, SorCoef k, const SorCtl* ctl) {
    if (ctl && *((volatile const int*)&ctl->converged)) return;
    // thread -> one cell of the colour: x index over half-rows
    int half = blockIdx.x * blockDim.x + threadIdx.x;     // 0 .. ceil(ni/2)-1
    int j = blockIdx.y + 1;
    int i = 2 * half + 1 + ((1 + j + color) & 1);         // first i with (i + j) % 2 == color
    if (i > ni) return;
    double* row = p + (long long)(j + NSB_ROW0) * pitch + NSB_COL0;
    double pc = row[i];
    double v = sor_cell<RECIP>(pc, row[i + 1], row[i - 1], row[i + pitch], row[i - pitch],
                               rhs[(long long)(j + NSB_ROW0) * pitch + NSB_COL0 + i], k);
    row[i] = v;
}

template <bool RECIP>
__global__ void sor_residual_kernel(const double* __restrict__ p, const double* __restrict__ rhs,
                                    long long pitch, int ni, int nj, SorCoef k, double* partials,
                                    SorCtl* ctl, int it, int check) {
    __shared__ double sh[32];
    __shared__ int is_last;
    if (check && *((volatile int*)&ctl->converged)) return;
    int j = blockIdx.y + 1;
    const double* row = p + (long long)(j + NSB_ROW0) * pitch + NSB_COL0;
    const double* rrow = rhs + (long long)(j + NSB_ROW0) * pitch + NSB_COL0;
    double acc = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x + 1; i <= ni; i += gridDim.x * blockDim.x) {
        double q = sor_residual<RECIP>(row[i], row[i + 1], row[i - 1], row[i + pitch], row[i - pitch],
                                       rrow[i], k);
        acc = fma(q, q, acc);
    }
    double tot = block_sum(acc, sh);
    const int nblk = gridDim.x * gridDim.y;
    if (threadIdx.x == 0) {
        partials[blockIdx.y * gridDim.x + blockIdx.x] = tot;
        __threadfence();
        unsigned int t = atomicAdd(&ctl->ticket, 1u);
        is_last = (t == (unsigned)nblk - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double s = 0.0;
        for (int q = threadIdx.x; q < nblk; q += blockDim.x) s += __ldcg(partials + q);
        double total = block_sum(s, sh);
        if (threadIdx.x == 0) {
            double res = sqrt(total / (double)k.cells);
            ctl->sum_r2 = total;
            ctl->res = res;
            ctl->iters = it + 1;
            if (res <= ctl->thresh) ctl->converged = 1;
            ctl->ticket = 0;
            sor_publish(ctl);
        }
    }
}

// ---- the scalars of a solve.  With IEEE division the kernels need nsb_rcp_refined(dx2 / dy2)
// (nsb_math.cuh): a one-thread kernel computes them once per (dx, dy) and the values are cached.
__device__ double g_rcp_out[2];
__global__ void rcp_refine_kernel(double dx2, double dy2) {
    g_rcp_out[0] = nsb_rcp_refined(dx2);
    g_rcp_out[1] = nsb_rcp_refined(dy2);
}

// Lower edge of the quotient window of nsb_div_try for divisor d, as the binary32 pattern its compare uses on the
// HIGH WORD of the quotient: above the compiler's own edge (high word 0x00100000) and above 2^-968 / d, so that a
// quotient inside the window comes from an |x| >= 2^-969 (the compiler's edge for the dividend).  A divisor outside
// 2^-500 .. 2^500 gets +inf: every quotient then takes the real division.
static float quot_window_lo(double d) {
    uint32_t h = 0x00100000u;
    if (!(d > ldexp(1.0, -500) && d < ldexp(1.0, 500))) return INFINITY;
    const double lo = ldexp(1.0, -968) / d;
    uint64_t bits;
    memcpy(&bits, &lo, sizeof bits);
    const uint32_t hi = (uint32_t)(bits >> 32) + 1u;
    if (hi > h) h = hi;
    float f;
    memcpy(&f, &h, sizeof f);
    return f;
}

SorCoef nsb_make_sor_coef(nsb_ctx* c, long long cells) {
    SorCoef k;
    const double dx2 = c->dx * c->dx, dy2 = c->dy * c->dy;
    k.om1 = 1.0 - c->P.omega;
    k.w = c->P.omega / (2.0 * (1.0 / dx2 + 1.0 / dy2));
    k.dx2 = dx2; k.dy2 = dy2; k.rdx2 = 1.0 / dx2; k.rdy2 = 1.0 / dy2;
    k.qlo_x = quot_window_lo(dx2); k.qlo_y = quot_window_lo(dy2);
    k.cells = cells;
    if (!c->use_recip) {
        if (!c->rcp_valid || c->rcp_key[0] != dx2 || c->rcp_key[1] != dy2) {
            rcp_refine_kernel<<<1, 1, 0, c->stream>>>(dx2, dy2);
            double out[2] = {0.0, 0.0};
            cudaError_t e = cudaMemcpyFromSymbolAsync(out, g_rcp_out, sizeof out, 0, cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
            // a failed refinement (a CUDA error, which the next launch reports) closes the quotient window
            // below: every quotient then takes the real division
            if (e == cudaSuccess && out[0] != 0.0 && out[1] != 0.0) {
                c->rcp_key[0] = dx2; c->rcp_key[1] = dy2;
                c->rcp_val[0] = out[0]; c->rcp_val[1] = out[1];
                c->rcp_valid = true;
            }
        }
        if (c->rcp_valid) { k.rdx2 = c->rcp_val[0]; k.rdy2 = c->rcp_val[1]; }
        else { k.qlo_x = INFINITY; k.qlo_y = INFINITY; }
    }
    return k;
}

// One full iteration in place on p.  With `check`, every kernel of an iteration queued after the
// converged flag was raised returns at once, so p stays the converged iterate.
cudaError_t nsb_launch_sor_simple_iter(nsb_ctx* c, int it, double* p, int check) {
    const SorCtl* ck = check ? c->d_ctl : nullptr;
    const int ni = c->g.ni, nj = c->g.nj;
    const SorCoef k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->g.nj);
    const int nb = (max(ni, nj) + 255) / 256;
    const int half = (ni + 1) / 2;
    dim3 cgrid((half + 127) / 128, nj);
    dim3 rgrid(min((ni + 255) / 256, 8), nj);
    // partials: one per residual block; the streaming plan always allocates >= 8 * nj + 64
    for (int color = 0; color < 2; color++) {
        ghost_refresh_kernel<<<nb, 256, 0, c->stream>>>(p, c->g.pitch, ni, nj, ck);
        if (c->use_recip) sor_color_kernel<true><<<cgrid, 128, 0, c->stream>>>(p, c->rhs, c->g.pitch, ni, nj, color, k, ck);
        else sor_color_kernel<false><<<cgrid, 128, 0, c->stream>>>(p, c->rhs, c->g.pitch, ni, nj, color, k, ck);
    }
    if (c->use_recip) sor_residual_kernel<true><<<rgrid, 256, 0, c->stream>>>(p, c->rhs, c->g.pitch, ni, nj, k, c->d_partials, c->d_ctl, it, check);
    else sor_residual_kernel<false><<<rgrid, 256, 0, c->stream>>>(p, c->rhs, c->g.pitch, ni, nj, k, c->d_partials, c->d_ctl, it, check);
    return NSB_LAUNCHED(c, 5);
}

// ---- ghosts of the final pressure for the fused engines.  The reference leaves in the ghost
// cells what the refresh BEFORE the last black sweep copied (main.cu:694): the current value of a
// wall-adjacent red cell, the pre-black value of a wall-adjacent black cell.  The pre-black value
// of the last iteration is what the other ping-pong buffer still holds.
__global__ void sor_finish_ghosts_kernel(double* pf, const double* __restrict__ pv, long long pitch,
                                         int ni, int nj, int wall_bottom, int wall_top) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    double* f = pf + (long long)NSB_ROW0 * pitch + NSB_COL0;
    const double* v = pv + (long long)NSB_ROW0 * pitch + NSB_COL0;
    if (t < nj) {
        int j = t + 1;
        long long r = (long long)j * pitch;
        f[r + 0] = ((1 + j) & 1) ? v[r + 1] : f[r + 1];
        f[r + ni + 1] = ((ni + j) & 1) ? v[r + ni] : f[r + ni];
    }
    if (t < ni) {
        int i = t + 1;
        if (wall_bottom) f[i] = ((i + 1) & 1) ? v[pitch + i] : f[pitch + i];
        long long r = (long long)nj * pitch;
        if (wall_top) f[r + pitch + i] = ((i + nj) & 1) ? v[r + i] : f[r + i];   // slab boundary: the neighbour's row stays
    }
}

cudaError_t nsb_launch_sor_finish_ghosts(nsb_ctx* c, double* pfinal, const double* pprev) {
    const int n = max(c->g.ni, c->g.nj);
    sor_finish_ghosts_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(pfinal, pprev, c->g.pitch, c->g.ni, c->g.nj,
                                                                     c->wall_bottom ? 1 : 0, c->wall_top ? 1 : 0);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------------------------------
// Serial-order engine (validation only): the reference's CPU solver sweeps the grid
// lexicographically in place, i outer / j inner (src/serial/integration.c:149-154), refreshes the
// ghost cells ONCE per iteration (138-146) and evaluates the residual with those same ghosts
// (156-160).  Cell (i, j) then depends on the NEW values of (i-1, j) and (i, j-1) and the OLD values
// of (i+1, j) and (i, j+1): all cells of one anti-diagonal i + j = d are independent, and sweeping
// the diagonals in increasing d reproduces the lexicographic result bit for bit.  One CTA walks the
// diagonals with a block barrier between them -- low throughput by construction; it
// exists so that whole runs can be compared with the reference's SERIAL binary including its
// iteration counts.  norm_p = sqrt(sum / i_max / j_max) as in L2() (integration.c:115-124).
#define LEX_THREADS 1024

template <bool RECIP>
__global__ void __launch_bounds__(LEX_THREADS, 1) sor_lex_kernel(double* p, const double* __restrict__ rhs, long long pitch,
                                                                  int ni, int nj, SorCoef k, double eps, const StepRed* red,
                                                                  SorCtl* ctl, int max_it, int check) {
    __shared__ double sh[32];
    __shared__ double res_sh;
    double* P = p + (long long)NSB_ROW0 * pitch + NSB_COL0;
    const double* Q = rhs + (long long)NSB_ROW0 * pitch + NSB_COL0;
    const int tid = threadIdx.x;
    const double norm_p = sqrt(red->sum_p2 / (double)ni / (double)nj);
    const double thresh = eps * (norm_p + 1.5);
    int it = 0, converged = 0;
    double res = 0.0, total = 0.0;
    while (it < max_it) {
        for (int j = 1 + tid; j <= nj; j += LEX_THREADS) {                    // integration.c:138-142
            P[(long long)j * pitch] = P[(long long)j * pitch + 1];
            P[(long long)j * pitch + ni + 1] = P[(long long)j * pitch + ni];
        }
        for (int i = 1 + tid; i <= ni; i += LEX_THREADS) {                    // integration.c:143-146
            P[i] = P[pitch + i];
            P[(long long)(nj + 1) * pitch + i] = P[(long long)nj * pitch + i];
        }
        __syncthreads();
        for (int d = 2; d <= ni + nj; d++) {                                  // integration.c:149-154 as a wavefront
            const int ilo = max(1, d - nj), ihi = min(ni, d - 1);
            for (int i = ilo + tid; i <= ihi; i += LEX_THREADS) {
                const long long o = (long long)(d - i) * pitch + i;
                P[o] = sor_cell<RECIP>(P[o], P[o + 1], P[o - 1], P[o + pitch], P[o - pitch], Q[o], k);
            }
            __syncthreads();
        }
        double acc = 0.0;                                                     // integration.c:156-164
        for (int j = 1 + (tid >> 5); j <= nj; j += LEX_THREADS / 32)
            for (int i = 1 + (tid & 31); i <= ni; i += 32) {
                const long long o = (long long)j * pitch + i;
                const double r = sor_residual<RECIP>(P[o], P[o + 1], P[o - 1], P[o + pitch], P[o - pitch], Q[o], k);
                acc = fma(r, r, acc);
            }
        const double t = block_sum(acc, sh);
        if (tid == 0) { total = t; res_sh = sqrt(t / (double)ni / (double)nj); }
        __syncthreads();
        res = res_sh;
        it++;
        if (res <= thresh) { converged = 1; if (check) break; }               // integration.c:164-166
        __syncthreads();
    }
    if (tid == 0) {
        ctl->norm_p = norm_p; ctl->thresh = thresh;
        if (it > 0) { ctl->res = res; ctl->sum_r2 = total; }
        ctl->converged = converged;
        __threadfence();
        *((volatile int*)&ctl->iters) = it;
        __threadfence();
        sor_publish(ctl);
    }
}

cudaError_t nsb_launch_sor_lex(nsb_ctx* c, int max_it, int check) {
    const SorCoef k = nsb_make_sor_coef(c, (long long)c->g.ni * (long long)c->g.nj);
    double* p = c->pbuf[c->pcur];
    if (c->use_recip) sor_lex_kernel<true><<<1, LEX_THREADS, 0, c->stream>>>(p, c->rhs, c->g.pitch, c->g.ni, c->g.nj, k, c->P.epsilon, c->d_red, c->d_ctl, max_it, check);
    else sor_lex_kernel<false><<<1, LEX_THREADS, 0, c->stream>>>(p, c->rhs, c->g.pitch, c->g.ni, c->g.nj, k, c->P.epsilon, c->d_red, c->d_ctl, max_it, check);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------------------------------
// Self-check of nsb_div_try / nsb_div_full against __ddiv_rn (development / test aid, not in the header):
// n quotients x / d per mode, x from a counter-based generator.  mode 0: random significand,
// exponent in [-64, 64] (field data); 1: random 64-bit patterns (all exponents, NaN, Inf,
// denormals); 2: zeros, denormals and the edges of the exponent window.  NaN == NaN here.
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__global__ void divcheck_kernel(double d1, double d2, double r1, double r2, float lo1, float lo2, long long n, unsigned long long seed, int mode,
                                unsigned long long* bad, double* first_bad) {
    unsigned long long local = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long h = mix64(seed + 2ull * (unsigned long long)i), h2 = mix64(h);
        double x1, x2;
        if (mode == 0) {
            const long long e1 = 1023 - 64 + (long long)((h >> 52) % 129ull), e2 = 1023 - 64 + (long long)((h2 >> 52) % 129ull);
            x1 = __longlong_as_double((long long)((h & 0x800fffffffffffffull) | ((unsigned long long)e1 << 52)));
            x2 = __longlong_as_double((long long)((h2 & 0x800fffffffffffffull) | ((unsigned long long)e2 << 52)));
        } else if (mode == 1) {
            x1 = __longlong_as_double((long long)h); x2 = __longlong_as_double((long long)h2);
        } else {
            const unsigned long long sel = h >> 61;
            const unsigned long long m = h2 & 0x000fffffffffffffull;
            const unsigned long long sgn = h & 0x8000000000000000ull >> 0;
            unsigned long long b;
            if (sel == 0) b = 0ull;                                        // +-0
            else if (sel == 1) b = m;                                      // denormal
            else if (sel == 2) b = (0x035ull << 52) | m;                   // just below the window
            else if (sel == 3) b = (0x036ull << 52) | m;                   // lower edge
            else if (sel == 4) b = (0x7f7ull << 52) | m;                   // upper edge
            else if (sel == 5) b = (0x7feull << 52) | m;                   // largest finite
            else if (sel == 6) b = (0x001ull << 52) | m;                   // smallest normal
            else b = (0x7ffull << 52) | (m & 1ull ? m : 0ull);             // Inf / NaN
            x1 = __longlong_as_double((long long)(b | (sgn & 0x8000000000000000ull)));
            x2 = -x1;
        }
        // inside the window (w1 / w2 stay true) the three-operation quotient must BE the division;
        // outside, the callers fall back to nsb_div_full
        bool w1 = true, w2 = true;
        double q1 = nsb_div_try(x1, d1, r1, lo1, w1), q2 = nsb_div_try(x2, d2, r2, lo2, w2);
        if (!w1) q1 = nsb_div_full(x1, d1);
        if (!w2) q2 = nsb_div_full(x2, d2);
        const double t1 = __ddiv_rn(x1, d1), t2 = __ddiv_rn(x2, d2);
        const bool ok1 = (__double_as_longlong(q1) == __double_as_longlong(t1)) || (q1 != q1 && t1 != t1);
        const bool ok2 = (__double_as_longlong(q2) == __double_as_longlong(t2)) || (q2 != q2 && t2 != t2);
        if (mode == 0 && !(w1 && w2)) atomicAdd(bad + 1, 1ull << 32);   // field-like operands must never leave the window
        if (!ok1 || !ok2) {
            local++;
            if ((atomicAdd(bad + 1, 1ull) & 0xffffffffull) == 0ull) { first_bad[0] = ok1 ? x2 : x1; first_bad[1] = ok1 ? d2 : d1; }
        }
    }
    if (local) atomicAdd(bad, local);
}

extern "C" int nsb_debug_divcheck(double d1, double d2, long long n, unsigned long long seed, int mode,
                                  long long* mismatches, double* first_bad_x, double* first_bad_d) {
    unsigned long long* bad = nullptr;
    double* fb = nullptr;
    if (cudaMalloc(&bad, 16) != cudaSuccess || cudaMalloc(&fb, 16) != cudaSuccess) return NSB_ERR_CUDA;
    cudaMemset(bad, 0, 16); cudaMemset(fb, 0, 16);
    rcp_refine_kernel<<<1, 1>>>(d1, d2);
    double r[2] = {0.0, 0.0};
    cudaMemcpyFromSymbol(r, g_rcp_out, sizeof r);
    divcheck_kernel<<<148 * 8, 256>>>(d1, d2, r[0], r[1], quot_window_lo(d1), quot_window_lo(d2), n, seed, mode, bad, fb);
    unsigned long long hb[2] = {0, 0};
    double hf[2] = {0.0, 0.0};
    cudaError_t e = cudaMemcpy(hb, bad, 16, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(hf, fb, 16, cudaMemcpyDeviceToHost);
    cudaFree(bad); cudaFree(fb);
    if (e != cudaSuccess) return NSB_ERR_CUDA;
    if (mismatches) *mismatches = (long long)hb[0] + (long long)(hb[1] >> 32);   // + field-like operands outside the window
    if (first_bad_x) *first_bad_x = hf[0];
    if (first_bad_d) *first_bad_d = hf[1];
    return NSB_OK;
}
