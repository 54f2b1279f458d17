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

// ---- measurement aid (bench.py): time stamps around library calls INSIDE a captured CUDA graph.
// A plain cudaEventRecord during stream capture becomes an internal dependency node whose time cannot be read.  Recorded with
// cudaEventRecordExternal it is an event-record node: every replay of the graph stamps the event, and cudaEventElapsedTime reads
// the stamps after a synchronize.  So the per-entry-point times bench.py reports come from the graph-replayed step that `value`
// times, not from a separate eager pass whose host launch gaps leak into the intervals.
#include <new>
struct FocTimer { int n; cudaEvent_t* ev; };

FOC_API int foc_timer_create(int n, void** handle) {
    if (n <= 0 || handle == nullptr) return FOC_ERR_UNSUPPORTED;
    FocTimer* t = new (std::nothrow) FocTimer{n, new (std::nothrow) cudaEvent_t[n]};
    if (t == nullptr || t->ev == nullptr) return (int)cudaErrorMemoryAllocation;
    for (int i = 0; i < n; ++i) {
        cudaError_t e = cudaEventCreate(&t->ev[i]);
        if (e != cudaSuccess) return (int)e;
    }
    *handle = t;
    return 0;
}
FOC_API int foc_timer_mark(void* handle, int slot, void* stream) {
    FocTimer* t = static_cast<FocTimer*>(handle);
    if (t == nullptr || slot < 0 || slot >= t->n) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaError_t e = cudaStreamIsCapturing(st, &cs);
    if (e != cudaSuccess) return (int)e;
    e = cudaEventRecordWithFlags(t->ev[slot], st, cs == cudaStreamCaptureStatusActive ? cudaEventRecordExternal : cudaEventRecordDefault);
    return (int)e;
}
FOC_API int foc_timer_elapsed(void* handle, int a, int b, float* ms) {
    FocTimer* t = static_cast<FocTimer*>(handle);
    if (t == nullptr || a < 0 || b < 0 || a >= t->n || b >= t->n || ms == nullptr) return FOC_ERR_UNSUPPORTED;
    return (int)cudaEventElapsedTime(ms, t->ev[a], t->ev[b]);
}
FOC_API int foc_timer_destroy(void* handle) {
    FocTimer* t = static_cast<FocTimer*>(handle);
    if (t == nullptr) return 0;
    for (int i = 0; i < t->n; ++i) cudaEventDestroy(t->ev[i]);
    delete[] t->ev;
    delete t;
    return 0;
}
