// capi_noalloc.cu -- dirt_rasterise_forward_noalloc (include/dirt_b200.h): the forward pass for callers
// that may not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
//
// The forward kernels need scratch (depth|face keys, projected vertices, counters, lists: the
// "workspace" of dirt_rasterise_forward).  Here that scratch is a file-scope device array of a fixed,
// modest size: every GPU that loads the library holds DIRT_NOALLOC_WORKSPACE_BYTES of it until the
// process exits.  In host code the array's name is only a stand-in, not a device address, so its
// address is resolved once per device -- with that device current -- by cudaGetSymbolAddress, and the
// resolved pointer is what the kernels receive.  No cudaMalloc / cudaFree / allocating library routine
// is called on this path: the kernels' launches and the cudaMemsetAsync calls that clear what must
// start empty (keys := 0xFF.., counters := 0, done by the forward pass itself before every use) are all.
//
// Every call on a device shares the one array, and it keeps the last call's contents.  Calls that
// overlap on the GPU (two streams, two threads) therefore race; serialise them or use
// dirt_rasterise_forward with one workspace per stream.
#include <cstdio>
#include <mutex>
#include "../../include/dirt_b200.h"
#include <cuda_runtime.h>

namespace dirt { int set_error(int status, const char *what, cudaError_t e); }    // capi.cu

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[DIRT_NOALLOC_WORKSPACE_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void *g_scratch_ptr[kMaxDevices] = {};

// Device address of the scratch array on the current device (resolved on first use per device).
int scratch_address(void **out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "no usable CUDA device (there is no CPU path)", e);
    if (dev < 0 || dev >= kMaxDevices) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "device ordinal out of range", cudaSuccess);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        void *ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return dirt::set_error(DIRT_STATUS_CUDA_ERROR, "cudaGetSymbolAddress(static scratch)", e);
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return DIRT_STATUS_OK;
}

}  // namespace

extern "C" {

size_t dirt_noalloc_workspace_limit(void) { return (size_t)DIRT_NOALLOC_WORKSPACE_BYTES; }

int dirt_rasterise_forward_noalloc(const DirtForwardArgs *a, void *stream)
{
    if (!a) return dirt::set_error(DIRT_STATUS_INVALID_ARGUMENT, "null arguments", cudaSuccess);
    const size_t need = dirt_workspace_bytes(&a->shape);
    if (need == 0) {
        // an invalid shape: let the ordinary entry point produce its message (it fails before touching the GPU)
        DirtForwardArgs d = *a;
        d.workspace = nullptr; d.workspace_bytes = 0;
        return dirt_rasterise_forward(&d, stream);
    }
    if (need > (size_t)DIRT_NOALLOC_WORKSPACE_BYTES) {
        char msg[256];
        snprintf(msg, sizeof(msg),
                 "dirt_rasterise_forward_noalloc: this shape needs %zu bytes of scratch, the static scratch holds %zu "
                 "(DIRT_NOALLOC_WORKSPACE_BYTES); use dirt_rasterise_forward with a caller-owned workspace",
                 need, (size_t)DIRT_NOALLOC_WORKSPACE_BYTES);
        return dirt::set_error(DIRT_STATUS_UNSUPPORTED, msg, cudaSuccess);
    }
    void *scratch = nullptr;
    if (int rc = scratch_address(&scratch)) return rc;
    DirtForwardArgs d = *a;
    d.workspace = scratch;
    d.workspace_bytes = (size_t)DIRT_NOALLOC_WORKSPACE_BYTES;
    return dirt_rasterise_forward(&d, stream);
}

}  // extern "C"
