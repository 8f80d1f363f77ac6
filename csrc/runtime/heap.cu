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

// ---------------------------------------------------------------------------------------------
// Symmetric-heap sub-allocator (OSHMEM memheap analogue, oshmem/mca/memheap/{buddy,ptmalloc}): hands out OFFSETS inside
// one peer-mapped region.  It is deterministic -- the same sequence of alloc/free calls yields the same offsets on every
// rank -- so after the region itself has been exchanged once, symmetric allocations need no further rendezvous: rank r's
// copy of an object lives at mapped_base[r] + offset.  First fit by lowest offset, neighbours coalesce on free.
// ---------------------------------------------------------------------------------------------
#include <map>

struct ewdml_symheap {
    int64_t bytes, align, used, peak;
    std::map<int64_t, int64_t> free_;     // offset -> size
    std::map<int64_t, int64_t> live_;     // offset -> size
};

extern "C" int ewdml_symheap_create(ewdml_symheap** out, int64_t bytes, int64_t align) {
    EWDML_CHECK_ARG(out && bytes > 0 && align > 0 && (align & (align - 1)) == 0);
    ewdml_symheap* h = new ewdml_symheap();
    h->bytes = bytes / align * align; h->align = align; h->used = 0; h->peak = 0;
    h->free_[0] = h->bytes;
    *out = h;
    return 0;
}

extern "C" int64_t ewdml_symheap_alloc(ewdml_symheap* h, int64_t bytes) {
    if (!h || bytes <= 0) return -1;
    const int64_t need = (bytes + h->align - 1) / h->align * h->align;
    for (auto it = h->free_.begin(); it != h->free_.end(); ++it) {
        if (it->second < need) continue;
        const int64_t off = it->first, rest = it->second - need;
        h->free_.erase(it);
        if (rest > 0) h->free_[off + need] = rest;
        h->live_[off] = need;
        h->used += need;
        if (h->used > h->peak) h->peak = h->used;
        return off;
    }
    ewdml_set_error("ewdml_symheap_alloc: %lld bytes do not fit (%lld of %lld in use)", (long long)need, (long long)h->used,
                    (long long)h->bytes);
    return -1;
}

extern "C" int ewdml_symheap_free(ewdml_symheap* h, int64_t offset) {
    EWDML_CHECK_ARG(h);
    auto it = h->live_.find(offset);
    if (it == h->live_.end()) { ewdml_set_error("ewdml_symheap_free: offset %lld is not a live allocation", (long long)offset); return EWDML_ERR_ARG; }
    int64_t off = it->first, size = it->second;
    h->live_.erase(it);
    h->used -= size;
    auto nxt = h->free_.lower_bound(off);
    if (nxt != h->free_.end() && off + size == nxt->first) { size += nxt->second; nxt = h->free_.erase(nxt); }
    if (nxt != h->free_.begin()) {
        auto prv = std::prev(nxt);
        if (prv->first + prv->second == off) { off = prv->first; size += prv->second; h->free_.erase(prv); }
    }
    h->free_[off] = size;
    return 0;
}

extern "C" int64_t ewdml_symheap_used(const ewdml_symheap* h) { return h ? h->used : -1; }
extern "C" int64_t ewdml_symheap_peak(const ewdml_symheap* h) { return h ? h->peak : -1; }
extern "C" int64_t ewdml_symheap_largest_free(const ewdml_symheap* h) {
    if (!h) return -1;
    int64_t best = 0;
    for (auto& kv : h->free_) best = kv.second > best ? kv.second : best;
    return best;
}
extern "C" int ewdml_symheap_destroy(ewdml_symheap* h) { delete h; return 0; }
