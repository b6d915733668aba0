// sor_resident.cu -- cluster-resident SOR engine (grid held on-chip for all iterations).
// Placeholder until the engine lands: reports "unsupported" so NSB_SOR_AUTO picks the streaming pass.
#include "nsb_internal.h"

int nsb_sor_resident_supported(const nsb_ctx* c) { (void)c; return 0; }

cudaError_t nsb_launch_sor_resident(nsb_ctx* c, int max_it, int check) {
    (void)c; (void)max_it; (void)check;
    return cudaErrorNotSupported;
}
