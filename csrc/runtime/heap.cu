// N1: symmetric peer heap over CUDA IPC.
//
// Each rank allocates the same regions (payload slots, flags, gradient buffer), exports one IPC
// handle per allocation and maps every peer's allocation into its own address space; kernels then
// address peer memory with ordinary pointers and NVLink/NVSwitch does the rest.  This replaces the
// smcuda BTL's IPC handshake and `get` path (ompi:opal/mca/btl/smcuda/btl_smcuda.c:1081), the
// driver shim that resolves cuIpc* by dlopen (ompi:opal/mca/common/cuda/common_cuda.c:449-480,
// 1006-1152) and OSHMEM's symmetric heap (oshmem/mca/memheap).  No host staging anywhere.
#include <map>
#include <mutex>

#include "../common.cuh"

static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");

extern "C" int ewdml_heap_alloc(void** base, size_t bytes) {
    EWDML_CHECK_ARG(base && bytes > 0);
    void* p = nullptr;
    EWDML_CUDA(cudaMalloc(&p, bytes));
    EWDML_CUDA(cudaMemset(p, 0, bytes));
    EWDML_CUDA(cudaDeviceSynchronize());
    *base = p;
    return 0;
}

extern "C" int ewdml_heap_free(void* base) {
    if (base) EWDML_CUDA(cudaFree(base));
    return 0;
}

extern "C" int ewdml_heap_export(void* base, uint8_t handle[64]) {
    EWDML_CHECK_ARG(base && handle);
    cudaIpcMemHandle_t h;
    EWDML_CUDA(cudaIpcGetMemHandle(&h, base));
    memcpy(handle, &h, 64);
    return 0;
}

extern "C" int ewdml_heap_import(const uint8_t handle[64], void** mapped) {
    EWDML_CHECK_ARG(handle && mapped);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void* p = nullptr;
    EWDML_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *mapped = p;
    return 0;
}

extern "C" int ewdml_heap_unmap(void* mapped) {
    if (mapped) EWDML_CUDA(cudaIpcCloseMemHandle(mapped));
    return 0;
}

extern "C" int ewdml_can_access_peer(int dev, int peer, int* out) {
    EWDML_CHECK_ARG(out);
    if (dev == peer) { *out = 1; return 0; }
    EWDML_CUDA(cudaDeviceCanAccessPeer(out, dev, peer));
    return 0;
}
