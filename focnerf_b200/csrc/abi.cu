// abi.cu -- ABI version and the API-compatibility no-ops of include/foc_b200.h.
#include "common.cuh"
#include <stddef.h>

FOC_API int foc_abi_version(void) { return 1; }

// ffmlp.h:12-13 (allocate_splitk / free_splitk) create and destroy the reference's process-global side streams
// and events for its split-K weight-gradient GEMMs (ffmlp.cu:711-730).  The sm_100a backward reduces weight
// gradients inside one kernel, so there is nothing to allocate; the symbols exist so FFMLP.__init__ /
// cleanup keep their call sites.
FOC_API int foc_allocate_splitk(size_t size) { (void)size; return 0; }
FOC_API int foc_free_splitk(void) { return 0; }
