// slab.cuh -- device-side building blocks of the y-slab decomposition:
// system-scope flags over NVLink peer memory and the decomposition-independent reduction.
//
// The reference has no multi-GPU code at all; this is new work.  The rules it follows:
//   - data crosses GPUs only as plain stores into the neighbour's arena (halo rows, band sums),
//     issued by the kernel that produced the data, followed by a system-scope release flag;
//   - a consumer polls its OWN memory (acquire, system scope), never the peer's;
//   - every wait has a deadline (SlabSync::timeout_ns, 2 s unless nsb_options.spin_timeout_ms / NSB_SPIN_TIMEOUT_MS say
//     otherwise -- raise it under compute-sanitizer or ncu replay) and reports through SlabSync::error instead of
//     hanging the device;
//   - reductions are summed per BAND (one SOR chunk of rows, always inside one rank) and the band
//     values are combined in an order that depends only on the global band count, so 1, 2, 4 or 8
//     slabs produce the same bits.
#pragma once

#include "nsb_internal.h"
#include "nsb_math.cuh"

// `err` always points at SlabSync::error of this rank; the deadline of the waits sits right behind it
__device__ __forceinline__ unsigned long long spin_timeout_ns(const int* err) {
    return *reinterpret_cast<const volatile unsigned long long*>(err + 2);
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// One thread: wait until *flag >= want.  Returns false (and raises *err) on timeout.
__device__ __forceinline__ bool spin_until_ge(const unsigned long long* flag, unsigned long long want, int* err) {
    if (ld_acquire_sys(flag) >= want) return true;
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(flag) < want) {
        if (*((volatile int*)err)) return false;             // something already timed out: do not pile up
        if (global_timer_ns() - t0 > spin_timeout_ns(err)) {
            *((volatile int*)err) = 1;
            return false;
        }
        __nanosleep(100);
    }
    return true;
}

// Fixed-order sum of n values by one warp: lane-strided sequential partials, then a butterfly.
// The order depends on n only.
__device__ __forceinline__ double warp_ordered_sum(const double* v, int n, int lane) {
    double s = 0.0;
    for (int q = lane; q < n; q += 32) s += __ldcg(v + q);
    return warp_sum(s);
}
__device__ __forceinline__ double warp_ordered_max(const double* v, int n, int lane) {
    double s = -1.0e300;
    for (int q = lane; q < n; q += 32) s = fmax(s, __ldcg(v + q));
    return warp_max(s);
}

// ---- collective reduction, part 1 (any block, all threads): contribute this rank's band values.
// vals[q * nb_local + b] for q < nq.  Writes them into every rank's gather buffer (parity
// round & 3), then tells every rank that this rank has contributed to `round`.
__device__ __forceinline__ void slab_red_contribute(const SlabRed& sr, unsigned long long round, int nq,
                                                    const double* vals) {
    const int par = (int)(round & (unsigned long long)(NSB_RED_DEPTH - 1));
    for (int r = 0; r < sr.world; r++) {
        double* dst = sr.gband[r] + (size_t)par * NSB_RED_Q * sr.nb_global;
        for (int x = threadIdx.x; x < nq * sr.nb_local; x += blockDim.x) {
            const int q = x / sr.nb_local, b = x - q * sr.nb_local;
            dst[(size_t)q * sr.nb_global + sr.band0 + b] = vals[x];
        }
    }
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < sr.world)
        st_release_sys(&sr.sync[threadIdx.x]->red_arrived[sr.rank], round + 1ull);
}

// ---- collective reduction, part 2 (one block, >= 96 threads): wait for every rank, then combine the
// global band values: quantity q is a MAX if bit q of max_mask is set, else a SUM.  Result in
// out[q] (shared or global memory), valid after the trailing __syncthreads().
__device__ __forceinline__ void slab_red_finish(const SlabRed& sr, unsigned long long round, int nq,
                                                unsigned max_mask, double* out) {
    if ((int)threadIdx.x < sr.world)
        spin_until_ge(&sr.sync[sr.rank]->red_arrived[threadIdx.x], round + 1ull, &sr.sync[sr.rank]->error);
    __syncthreads();
    const int par = (int)(round & (unsigned long long)(NSB_RED_DEPTH - 1));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < nq) {
        const double* src = sr.gband[sr.rank] + ((size_t)par * NSB_RED_Q + warp) * sr.nb_global;
        const double v = ((max_mask >> warp) & 1u) ? warp_ordered_max(src, sr.nb_global, lane)
                                                   : warp_ordered_sum(src, sr.nb_global, lane);
        if (lane == 0) out[warp] = v;
    }
    __syncthreads();
}


// ---- the same collective reduction driven by ONE WARP (persistent sweep: the warp that closes an
// iteration).  Same buffers, same flags, same summation order as the block versions above.
__device__ __forceinline__ void slab_red_contribute_warp(const SlabRed& sr, unsigned long long round, int nq,
                                                         const double* vals, int lane) {
    const int par = (int)(round & (unsigned long long)(NSB_RED_DEPTH - 1));
    for (int r = 0; r < sr.world; r++) {
        double* dst = sr.gband[r] + (size_t)par * NSB_RED_Q * sr.nb_global;
        for (int x = lane; x < nq * sr.nb_local; x += 32) {
            const int q = x / sr.nb_local, b = x - q * sr.nb_local;
            dst[(size_t)q * sr.nb_global + sr.band0 + b] = __ldcg(vals + x);
        }
    }
    __threadfence_system();
    __syncwarp();
    if (lane < sr.world) st_release_sys(&sr.sync[lane]->red_arrived[sr.rank], round + 1ull);
}

// Returns false if a rank did not arrive in time.  out = combined value of quantity 0 (a SUM).
__device__ __forceinline__ bool slab_red_finish_sum_warp(const SlabRed& sr, unsigned long long round, int lane,
                                                         double* out) {
    int ok = 1;
    if (lane < sr.world)
        ok = spin_until_ge(&sr.sync[sr.rank]->red_arrived[lane], round + 1ull, &sr.sync[sr.rank]->error) ? 1 : 0;
    ok = __all_sync(0xffffffffu, ok);
    const int par = (int)(round & (unsigned long long)(NSB_RED_DEPTH - 1));
    const double* src = sr.gband[sr.rank] + (size_t)par * NSB_RED_Q * sr.nb_global;
    *out = warp_ordered_sum(src, sr.nb_global, lane);
    return ok != 0;
}
