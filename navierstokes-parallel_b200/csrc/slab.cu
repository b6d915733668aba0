// slab.cu -- y-slab decomposition across the GPUs of one NVSwitch box: time-step-level
// halo exchange kernels, wiring of the peers (same process: direct peer access; other processes:
// CUDA IPC), and the descriptor exchange entry points of the C ABI.
//
// The reference is single-GPU (no NCCL/MPI/P2P anywhere); this is new work defined by
// BASELINE.json.  No collective library is on the data path: every transfer is a store into the
// neighbour's arena issued by the kernel that owns the data, every synchronisation a system-scope
// flag (slab.cuh).  NCCL / gloo (through torch.distributed in the python host, or any other channel
// in a C host) only carries the 256-byte descriptors once, at start-up.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "nsb_internal.h"
#include "slab.cuh"

// ------------------------------------------------------------------ halo rows of u, v, G, p
struct HaloCopy {
    const double* src;    // first source row (already offset to the row)
    double* dst;          // first destination row in the peer's arena
    int rows;
};
struct HaloPushArgs {
    HaloCopy cp[4];
    int n;
    long long pitch;
    unsigned long long* flag;     // peer's halo_epoch[dir]
    unsigned long long epoch;     // value to publish
};

// One block per direction (blockIdx.x = 0: to the lower neighbour, 1: to the upper one).
__global__ void __launch_bounds__(1024) halo_push_kernel(HaloPushArgs down, HaloPushArgs up) {
    const HaloPushArgs& a = blockIdx.x == 0 ? down : up;
    if (a.n == 0) return;
    const long long n2 = a.pitch / 2;                      // double2 per row (pitch is a multiple of 16)
    for (int k = 0; k < a.n; k++) {
        const double2* s = reinterpret_cast<const double2*>(a.cp[k].src);
        double2* d = reinterpret_cast<double2*>(a.cp[k].dst);
        for (long long x = threadIdx.x; x < n2 * a.cp[k].rows; x += blockDim.x) d[x] = s[x];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(a.flag, a.epoch);
}

__global__ void halo_wait_kernel(const unsigned long long* from_down, unsigned long long want_down,
                                 const unsigned long long* from_up, unsigned long long want_up, int* err) {
    if (threadIdx.x == 0 && from_down) spin_until_ge(from_down, want_down, err);
    if (threadIdx.x == 1 && from_up) spin_until_ge(from_up, want_up, err);
}

static inline double* row_ptr(double* base, const NsbGrid& g, int l) {
    return base + (long long)(l + NSB_ROW0) * g.pitch;
}

// kind NSB_X_UV: u, v rows 1 -> lower's nj'+1 ; rows nj -> upper's 0          (after the BCs)
//      NSB_X_G : G row nj -> upper's 0                                          (after F, G)
//      NSB_X_P : p rows 1..5 -> lower's nj'+1..nj'+5 ; rows nj-4..nj -> upper's -4..0   (SOR entry; NSB_HALO_ROWS)
//      NSB_X_RHS: RHS rows 1..4 -> lower's nj'+1..nj'+4 ; rows nj-3..nj -> upper's -3..0 (after the RHS:
//                 the fused two-iteration sweep also updates the four halo rows next to a slab boundary)
cudaError_t nsb_launch_halo_push(nsb_ctx* c, int kind) {
    if (c->world == 1) return cudaSuccess;
    if (!c->connected) return cudaErrorNotReady;
    HaloPushArgs dn, up;
    memset(&dn, 0, sizeof dn);
    memset(&up, 0, sizeof up);
    dn.pitch = up.pitch = c->g.pitch;
    const int nj = c->g.nj;
    const int pc = c->pcur;
    if (!c->wall_bottom) {
        SlabPeer& d = c->peer[c->rank - 1];
        NsbGrid gd = c->g; gd.nj = d.nj;
        if (kind == NSB_X_UV) {
            dn.cp[0] = {row_ptr(c->u, c->g, 1), row_ptr(d.u, gd, d.nj + 1), 1};
            dn.cp[1] = {row_ptr(c->v, c->g, 1), row_ptr(d.v, gd, d.nj + 1), 1};
            dn.n = 2;
        } else if (kind == NSB_X_P) {
            dn.cp[0] = {row_ptr(c->pbuf[pc], c->g, 1), row_ptr(d.p[pc], gd, d.nj + 1), NSB_HALO_ROWS};
            dn.n = 1;
        } else if (kind == NSB_X_RHS) {
            dn.cp[0] = {row_ptr(c->rhs, c->g, 1), row_ptr(d.rhs, gd, d.nj + 1), NSB_HALO_ROWS - 1};
            dn.n = 1;
        }
        if (dn.n) { dn.flag = &d.sync->halo_epoch[1]; dn.epoch = ++c->epoch[0]; }   // I am its upper neighbour
    }
    if (!c->wall_top) {
        SlabPeer& u = c->peer[c->rank + 1];
        NsbGrid gu = c->g; gu.nj = u.nj;
        if (kind == NSB_X_UV) {
            up.cp[0] = {row_ptr(c->u, c->g, nj), row_ptr(u.u, gu, 0), 1};
            up.cp[1] = {row_ptr(c->v, c->g, nj), row_ptr(u.v, gu, 0), 1};
            up.n = 2;
        } else if (kind == NSB_X_G) {
            up.cp[0] = {row_ptr(c->G, c->g, nj), row_ptr(u.G, gu, 0), 1};
            up.n = 1;
        } else if (kind == NSB_X_P) {
            up.cp[0] = {row_ptr(c->pbuf[pc], c->g, nj - NSB_HALO_ROWS + 1), row_ptr(u.p[pc], gu, 1 - NSB_HALO_ROWS), NSB_HALO_ROWS};
            up.n = 1;
        } else if (kind == NSB_X_RHS) {
            up.cp[0] = {row_ptr(c->rhs, c->g, nj - NSB_HALO_ROWS + 2), row_ptr(u.rhs, gu, 2 - NSB_HALO_ROWS), NSB_HALO_ROWS - 1};
            up.n = 1;
        }
        if (up.n) { up.flag = &u.sync->halo_epoch[0]; up.epoch = ++c->epoch[1]; }   // I am its lower neighbour
    }
    if (dn.n == 0 && up.n == 0) return cudaSuccess;
    halo_push_kernel<<<2, 1024, 0, c->stream>>>(dn, up);
    return NSB_LAUNCHED(c, 1);
}

cudaError_t nsb_launch_halo_wait(nsb_ctx* c, int kind) {
    if (c->world == 1) return cudaSuccess;
    const unsigned long long* fd = NULL;
    const unsigned long long* fu = NULL;
    // what the neighbours push to me (mirror image of nsb_launch_halo_push)
    if (!c->wall_bottom) {                                      // the lower neighbour pushes UV, G and P upward
        fd = &c->d_sync->halo_epoch[0];
        ++c->epoch_in[0];
    }
    if (!c->wall_top && kind != NSB_X_G) {                      // the upper neighbour pushes UV and P downward
        fu = &c->d_sync->halo_epoch[1];
        ++c->epoch_in[1];
    }
    if (!fd && !fu) return cudaSuccess;
    halo_wait_kernel<<<1, 32, 0, c->stream>>>(fd, c->epoch_in[0], fu, c->epoch_in[1], &c->d_sync->error);
    return NSB_LAUNCHED(c, 1);
}

// ------------------------------------------------------------------ wiring
extern "C" int nsb_create_slab(const nsb_params* params, const nsb_options* opt, int rank, int world, nsb_ctx** out) {
    return nsb_create_slab_internal(params, opt, rank, world, out);
}

extern "C" int nsb_slab_desc_bytes(void) { return (int)sizeof(NsbSlabDesc); }

extern "C" int nsb_slab_export(const nsb_ctx* c, void* desc_out) {
    if (!c || !desc_out) return NSB_ERR_ARG;
    memcpy(desc_out, &c->desc, sizeof(NsbSlabDesc));
    return NSB_OK;
}

// descs: `world` descriptors in rank order (this rank's own included).
extern "C" int nsb_slab_connect(nsb_ctx* c, const void* descs) {
    if (!c || !descs) return NSB_ERR_ARG;
    if (c->world == 1) { c->connected = true; return NSB_OK; }
    const NsbSlabDesc* D = (const NsbSlabDesc*)descs;
    if (cudaSetDevice(c->device) != cudaSuccess) return nsb_fail(c, NSB_ERR_CUDA, "cudaSetDevice failed");
    for (int r = 0; r < c->world; r++) {
        const NsbSlabDesc& d = D[r];
        if (d.rank != r || d.world != c->world || d.pitch != c->g.pitch || d.n_strips != c->n_strips ||
            d.nbands_global != c->nbands_global)
            return nsb_fail(c, NSB_ERR_ARG, "descriptor %d does not match this decomposition", r);
        if (r == c->rank) continue;
        SlabPeer& p = c->peer[r];
        if (d.pid == (long long)getpid()) {
            // same process: the pointer is valid here; other device -> enable peer access
            p.base = (unsigned char*)(uintptr_t)d.base;
            p.mapped_ipc = false;
            if (d.device != c->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, c->device, d.device);
                if (!can) return nsb_fail(c, NSB_ERR_CUDA, "device %d cannot access device %d", c->device, d.device);
                cudaError_t e = cudaDeviceEnablePeerAccess(d.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                    return nsb_fail(c, NSB_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", d.device, cudaGetErrorString(e));
                cudaGetLastError();
            }
        } else {
            cudaIpcMemHandle_t h;
            memcpy(&h, d.ipc, sizeof h);
            void* ptr = NULL;
            cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess)
                return nsb_fail(c, NSB_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
            p.base = (unsigned char*)ptr;
            p.mapped_ipc = true;
        }
        p.u = (double*)(p.base + d.off_u); p.v = (double*)(p.base + d.off_v); p.G = (double*)(p.base + d.off_G);
        p.rhs = (double*)(p.base + d.off_rhs);
        p.p[0] = (double*)(p.base + d.off_p[0]); p.p[1] = (double*)(p.base + d.off_p[1]); p.p[2] = (double*)(p.base + d.off_p2);
        p.sync = (SlabSync*)(p.base + d.off_sync);
        p.sorflag = (unsigned long long*)(p.base + d.off_sorflag);
        p.gband = (double*)(p.base + d.off_gband);
        p.nj = d.nj;
    }
    c->peers_distinct = true;
    for (int r = 0; r < c->world; r++)
        for (int q = r + 1; q < c->world; q++)
            if (memcmp(D[r].uuid, D[q].uuid, 16) == 0) c->peers_distinct = false;
    if (getenv("NSB_SLAB_LOCKSTEP")) c->peers_distinct = false;    // force the one-launch-per-iteration sweep
    c->connected = true;
    return NSB_OK;
}

extern "C" int nsb_slab_info(const nsb_ctx* c, int* rank, int* world, int* j_lo, int* nj_local, int* chunk_rows) {
    if (!c) return NSB_ERR_ARG;
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    if (j_lo) *j_lo = c->j_lo;
    if (nj_local) *nj_local = c->g.nj;
    if (chunk_rows) *chunk_rows = c->chunk_rows;
    return NSB_OK;
}

// a wait timed out somewhere on this rank?
extern "C" int nsb_slab_error(nsb_ctx* c) {
    if (!c) return NSB_ERR_ARG;
    if (c->world == 1) return 0;
    int e = 0;
    cudaSetDevice(c->device);
    if (cudaMemcpy(&e, &c->d_sync->error, sizeof e, cudaMemcpyDeviceToHost) != cudaSuccess) return NSB_ERR_CUDA;
    return e;
}
