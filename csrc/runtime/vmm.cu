// N1 (second transport): symmetric peer heap on the CUDA virtual-memory-management API + NVLS multicast objects.
//
// cudaMalloc + cudaIpc* (heap.cu) cannot be bound to a multicast object, so when the device supports it the symmetric
// heap is built from cuMemCreate allocations instead: every rank creates one physical allocation, exports it as a POSIX
// file descriptor, passes the descriptor to its peers over a UNIX-domain socket (SCM_RIGHTS), and maps every peer's
// allocation into its own address space (cuMemImportFromShareableHandle + cuMemMap + cuMemSetAccess).  Rank 0 then creates
// ONE multicast object per buffer (cuMulticastCreate), the other ranks import it, every rank adds its device and binds its
// own physical allocation; mapping the multicast object gives an address through which
//     a store         is replicated by the NVSwitch into the same offset of every rank's buffer  (multimem.st), and
//     a reduce-load   returns the sum over all ranks' copies of that offset, computed in the switch (multimem.ld_reduce).
// That is what replaces W-1 unicast stores per vector in the K1/K3 pushes and the peer loads of the two-shot allreduce.
//
// Plays the role of: smcuda BTL IPC handshake (ompi:opal/mca/btl/smcuda/btl_smcuda.c:1081), common/cuda handle
// get/open (ompi:opal/mca/common/cuda/common_cuda.c:1006-1152), PMIx modex of the handles (here: fd passing), and the
// OSHMEM symmetric heap segment creation (oshmem/mca/sshmem).  Driver entry points are resolved at run time through
// cudaGetDriverEntryPoint, so the library neither links libcuda nor fails to load on a box without a driver.
#include <cuda.h>
#include <errno.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <mutex>
#include <vector>

#include "../common.cuh"

namespace {

struct Driver {
    bool ok = false;
    decltype(&cuMemCreate) MemCreate = nullptr;
    decltype(&cuMemRelease) MemRelease = nullptr;
    decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
    decltype(&cuMemAddressFree) MemAddressFree = nullptr;
    decltype(&cuMemMap) MemMap = nullptr;
    decltype(&cuMemUnmap) MemUnmap = nullptr;
    decltype(&cuMemSetAccess) MemSetAccess = nullptr;
    decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
    decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
    decltype(&cuMemGetAllocationGranularity) MemGranularity = nullptr;
    decltype(&cuMulticastCreate) McCreate = nullptr;
    decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
    decltype(&cuMulticastBindMem) McBindMem = nullptr;
    decltype(&cuMulticastUnbind) McUnbind = nullptr;
    decltype(&cuMulticastGetGranularity) McGranularity = nullptr;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
    decltype(&cuDeviceGet) DeviceGet = nullptr;
    decltype(&cuGetErrorString) GetErrorString = nullptr;
};

template <typename F>
bool resolve(const char* name, F& fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        cudaGetLastError();
        return false;
    }
    fn = reinterpret_cast<F>(p);
    return true;
}

Driver* driver() {
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        if (cudaFree(0) != cudaSuccess) { cudaGetLastError(); return; }       // make sure the primary context exists
        bool ok = true;
        ok &= resolve("cuMemCreate", d.MemCreate);
        ok &= resolve("cuMemRelease", d.MemRelease);
        ok &= resolve("cuMemAddressReserve", d.MemAddressReserve);
        ok &= resolve("cuMemAddressFree", d.MemAddressFree);
        ok &= resolve("cuMemMap", d.MemMap);
        ok &= resolve("cuMemUnmap", d.MemUnmap);
        ok &= resolve("cuMemSetAccess", d.MemSetAccess);
        ok &= resolve("cuMemExportToShareableHandle", d.MemExport);
        ok &= resolve("cuMemImportFromShareableHandle", d.MemImport);
        ok &= resolve("cuMemGetAllocationGranularity", d.MemGranularity);
        ok &= resolve("cuDeviceGetAttribute", d.DeviceGetAttribute);
        ok &= resolve("cuDeviceGet", d.DeviceGet);
        ok &= resolve("cuGetErrorString", d.GetErrorString);
        // multicast entry points are optional (older drivers): their absence only disables NVLS
        resolve("cuMulticastCreate", d.McCreate);
        resolve("cuMulticastAddDevice", d.McAddDevice);
        resolve("cuMulticastBindMem", d.McBindMem);
        resolve("cuMulticastUnbind", d.McUnbind);
        resolve("cuMulticastGetGranularity", d.McGranularity);
        d.ok = ok;
    });
    return &d;
}

const char* cu_err(CUresult r) {
    const char* s = nullptr;
    Driver* d = driver();
    if (d->GetErrorString && d->GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown driver error";
}

#define EWDML_CU(expr)                                                                                  \
    do {                                                                                                \
        CUresult _r = (expr);                                                                           \
        if (_r != CUDA_SUCCESS) {                                                                       \
            ewdml_set_error("%s:%d: %s -> %s (%d)", __FILE__, __LINE__, #expr, cu_err(_r), (int)_r);    \
            return EWDML_ERR_CUDA;                                                                      \
        }                                                                                               \
    } while (0)

int current_device(CUdevice* dev, int* ordinal) {
    int o = 0;
    if (cudaGetDevice(&o) != cudaSuccess) { ewdml_set_error("vmm: no current device"); return EWDML_ERR_CUDA; }
    EWDML_CU(driver()->DeviceGet(dev, o));
    if (ordinal) *ordinal = o;
    return 0;
}

CUmemAllocationProp mem_prop(int ordinal) {
    CUmemAllocationProp p;
    memset(&p, 0, sizeof(p));
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = ordinal;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

int map_rw(CUdeviceptr* va, size_t bytes, CUmemGenericAllocationHandle h, int ordinal, size_t align) {
    Driver* d = driver();
    EWDML_CU(d->MemAddressReserve(va, bytes, align, 0, 0));
    CUresult r = d->MemMap(*va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        d->MemAddressFree(*va, bytes);
        ewdml_set_error("vmm: cuMemMap failed: %s (%d)", cu_err(r), (int)r);
        return EWDML_ERR_CUDA;
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = ordinal;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d->MemSetAccess(*va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        d->MemUnmap(*va, bytes);
        d->MemAddressFree(*va, bytes);
        ewdml_set_error("vmm: cuMemSetAccess failed: %s (%d)", cu_err(r), (int)r);
        return EWDML_ERR_CUDA;
    }
    return 0;
}

}  // namespace

// One VMM object: a physical allocation (own or imported) or a multicast object, plus its mapping in this process.
struct ewdml_vmm_buf {
    CUmemGenericAllocationHandle handle = 0;
    CUdeviceptr va = 0;
    size_t bytes = 0;
    int kind = 0;          // 0 = own memory, 1 = imported peer memory, 2 = multicast object
    int bound = 0;         // multicast: local memory bound
    int ordinal = 0;
};

// {vmm usable, multicast usable, posix-fd handles usable} on the current device
extern "C" int ewdml_vmm_query(int out[3]) {
    EWDML_CHECK_ARG(out);
    out[0] = out[1] = out[2] = 0;
    Driver* d = driver();
    if (!d->ok) return 0;
    CUdevice dev;
    if (current_device(&dev, nullptr)) return 0;
    int v = 0;
    if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev) == CUDA_SUCCESS) out[0] = v;
    v = 0;
    if (d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS) out[2] = v;
    v = 0;
    if (d->McCreate && d->McAddDevice && d->McBindMem && d->McGranularity &&
        d->DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS)
        out[1] = v;
    return 0;
}

// Allocation granularity that satisfies both cuMemCreate and (when `ndev` > 1 and multicast is usable) cuMulticastBindMem.
extern "C" int ewdml_vmm_granularity(int ndev, size_t* out) {
    EWDML_CHECK_ARG(out);
    Driver* d = driver();
    if (!d->ok) { ewdml_set_error("vmm: driver entry points unavailable"); return EWDML_ERR_STATE; }
    CUdevice dev; int ord = 0;
    int rc = current_device(&dev, &ord);
    if (rc) return rc;
    CUmemAllocationProp p = mem_prop(ord);
    size_t g = 0;
    EWDML_CU(d->MemGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    int q[3];
    ewdml_vmm_query(q);
    if (ndev > 1 && q[1]) {
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = (unsigned)ndev;
        mp.size = g;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (d->McGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS && mg > g) g = (mg + g - 1) / g * g;
    }
    *out = g;
    return 0;
}

// Own allocation of `bytes` (a multiple of ewdml_vmm_granularity), zero-filled, mapped read/write; *fd is the exported
// descriptor to hand to the peers (close it with ewdml_vmm_close_fd once everybody has imported it).
extern "C" int ewdml_vmm_alloc(ewdml_vmm_buf** out, size_t bytes, void** base, int* fd) {
    EWDML_CHECK_ARG(out && base && fd && bytes > 0);
    Driver* d = driver();
    if (!d->ok) { ewdml_set_error("vmm: driver entry points unavailable"); return EWDML_ERR_STATE; }
    CUdevice dev; int ord = 0;
    int rc = current_device(&dev, &ord);
    if (rc) return rc;
    CUmemAllocationProp p = mem_prop(ord);
    ewdml_vmm_buf* b = new ewdml_vmm_buf();
    b->bytes = bytes; b->ordinal = ord; b->kind = 0;
    CUresult r = d->MemCreate(&b->handle, bytes, &p, 0);
    if (r != CUDA_SUCCESS) { ewdml_set_error("vmm: cuMemCreate(%zu) failed: %s (%d)", bytes, cu_err(r), (int)r); delete b; return EWDML_ERR_CUDA; }
    rc = map_rw(&b->va, bytes, b->handle, ord, 0);
    if (rc) { d->MemRelease(b->handle); delete b; return rc; }
    int f = -1;
    r = d->MemExport(&f, b->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        ewdml_set_error("vmm: cuMemExportToShareableHandle failed: %s (%d)", cu_err(r), (int)r);
        d->MemUnmap(b->va, bytes); d->MemAddressFree(b->va, bytes); d->MemRelease(b->handle); delete b;
        return EWDML_ERR_CUDA;
    }
    if (cudaMemset((void*)b->va, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        ewdml_set_error("vmm: zero-fill failed: %s", cudaGetErrorString(cudaGetLastError()));
        return EWDML_ERR_CUDA;
    }
    *out = b; *base = (void*)b->va; *fd = f;
    return 0;
}

// Map a peer's allocation (descriptor received over the fd channel) for access from the current device.
extern "C" int ewdml_vmm_import(ewdml_vmm_buf** out, int fd, size_t bytes, void** mapped) {
    EWDML_CHECK_ARG(out && mapped && fd >= 0 && bytes > 0);
    Driver* d = driver();
    if (!d->ok) { ewdml_set_error("vmm: driver entry points unavailable"); return EWDML_ERR_STATE; }
    CUdevice dev; int ord = 0;
    int rc = current_device(&dev, &ord);
    if (rc) return rc;
    ewdml_vmm_buf* b = new ewdml_vmm_buf();
    b->bytes = bytes; b->ordinal = ord; b->kind = 1;
    CUresult r = d->MemImport(&b->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) { ewdml_set_error("vmm: cuMemImportFromShareableHandle failed: %s (%d)", cu_err(r), (int)r); delete b; return EWDML_ERR_CUDA; }
    rc = map_rw(&b->va, bytes, b->handle, ord, 0);
    if (rc) { d->MemRelease(b->handle); delete b; return rc; }
    *out = b; *mapped = (void*)b->va;
    return 0;
}

extern "C" int ewdml_vmm_close_fd(int fd) {
    if (fd >= 0) close(fd);
    return 0;
}

// Multicast object over `ndev` devices for buffers of `bytes`; root only.  *fd is the descriptor to broadcast.
extern "C" int ewdml_mc_create(ewdml_vmm_buf** out, int ndev, size_t bytes, int* fd) {
    EWDML_CHECK_ARG(out && fd && ndev >= 2 && bytes > 0);
    Driver* d = driver();
    if (!d->ok || !d->McCreate) { ewdml_set_error("multicast: not supported by this driver"); return EWDML_ERR_STATE; }
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)ndev;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    ewdml_vmm_buf* b = new ewdml_vmm_buf();
    b->bytes = bytes; b->kind = 2;
    cudaGetDevice(&b->ordinal);
    CUresult r = d->McCreate(&b->handle, &mp);
    if (r != CUDA_SUCCESS) { ewdml_set_error("multicast: cuMulticastCreate(%d devices, %zu bytes) failed: %s (%d)", ndev, bytes, cu_err(r), (int)r); delete b; return EWDML_ERR_CUDA; }
    int f = -1;
    r = d->MemExport(&f, b->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) { ewdml_set_error("multicast: export failed: %s (%d)", cu_err(r), (int)r); d->MemRelease(b->handle); delete b; return EWDML_ERR_CUDA; }
    *out = b; *fd = f;
    return 0;
}

extern "C" int ewdml_mc_import(ewdml_vmm_buf** out, int fd, size_t bytes) {
    EWDML_CHECK_ARG(out && fd >= 0 && bytes > 0);
    Driver* d = driver();
    if (!d->ok || !d->McCreate) { ewdml_set_error("multicast: not supported by this driver"); return EWDML_ERR_STATE; }
    ewdml_vmm_buf* b = new ewdml_vmm_buf();
    b->bytes = bytes; b->kind = 2;
    cudaGetDevice(&b->ordinal);
    CUresult r = d->MemImport(&b->handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    if (r != CUDA_SUCCESS) { ewdml_set_error("multicast: import failed: %s (%d)", cu_err(r), (int)r); delete b; return EWDML_ERR_CUDA; }
    *out = b;
    return 0;
}

// Every rank, before ANY rank binds memory (put a host barrier between the two phases).
extern "C" int ewdml_mc_add_device(ewdml_vmm_buf* mc) {
    EWDML_CHECK_ARG(mc && mc->kind == 2);
    CUdevice dev; int ord = 0;
    int rc = current_device(&dev, &ord);
    if (rc) return rc;
    EWDML_CU(driver()->McAddDevice(mc->handle, dev));
    return 0;
}

// Bind this rank's allocation at offset 0 of the multicast object and map the object; *mc_base is the multicast address
// whose offset x reaches offset x of every rank's buffer.
extern "C" int ewdml_mc_bind_map(ewdml_vmm_buf* mc, ewdml_vmm_buf* mem, void** mc_base) {
    EWDML_CHECK_ARG(mc && mem && mc_base && mc->kind == 2 && mem->kind == 0 && mem->bytes == mc->bytes);
    Driver* d = driver();
    EWDML_CU(d->McBindMem(mc->handle, 0, mem->handle, 0, mem->bytes, 0));
    mc->bound = 1;
    size_t g = 0;
    ewdml_vmm_granularity(2, &g);
    int rc = map_rw(&mc->va, mc->bytes, mc->handle, mem->ordinal, g);
    if (rc) return rc;
    *mc_base = (void*)mc->va;
    return 0;
}

extern "C" int ewdml_vmm_free(ewdml_vmm_buf* b) {
    if (!b) return 0;
    Driver* d = driver();
    if (b->va) {
        d->MemUnmap(b->va, b->bytes);
        d->MemAddressFree(b->va, b->bytes);
    }
    if (b->handle) d->MemRelease(b->handle);
    delete b;
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// File-descriptor channel between the ranks of one node: one abstract-namespace UNIX datagram socket per rank,
// descriptors travel as SCM_RIGHTS ancillary data with a (tag, source rank) header.  Messages that arrive before they are
// asked for are kept in a stash, sends never block on a full receiver queue without draining our own (no send/send
// deadlock at any world size).
// ---------------------------------------------------------------------------------------------------------------------
struct FdMsg { uint64_t tag; int src; int fd; };
struct ewdml_fdx {
    int sock = -1, rank = 0, world = 0;
    char job[64];
    std::vector<FdMsg> stash;
};

static socklen_t fdx_addr(const ewdml_fdx* x, int rank, sockaddr_un* a) {
    memset(a, 0, sizeof(*a));
    a->sun_family = AF_UNIX;
    int n = snprintf(a->sun_path + 1, sizeof(a->sun_path) - 1, "ewdml-%s-%d", x->job, rank);    // sun_path[0] = 0: abstract
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

extern "C" int ewdml_fdx_open(ewdml_fdx** out, const char* job, int rank, int world) {
    EWDML_CHECK_ARG(out && job && rank >= 0 && rank < world && strlen(job) < 48);
    ewdml_fdx* x = new ewdml_fdx();
    x->rank = rank; x->world = world;
    snprintf(x->job, sizeof(x->job), "%s", job);
    x->sock = socket(AF_UNIX, SOCK_DGRAM | SOCK_CLOEXEC, 0);
    if (x->sock < 0) { ewdml_set_error("fdx: socket: %s", strerror(errno)); delete x; return EWDML_ERR_IO; }
    sockaddr_un a;
    socklen_t len = fdx_addr(x, rank, &a);
    if (bind(x->sock, (sockaddr*)&a, len) != 0) {
        ewdml_set_error("fdx: bind(@%s): %s", a.sun_path + 1, strerror(errno));
        close(x->sock); delete x;
        return EWDML_ERR_IO;
    }
    *out = x;
    return 0;
}

static int fdx_drain_one(ewdml_fdx* x) {      // 1 = stashed a message, 0 = nothing there, <0 = error
    uint64_t hdr[2];
    iovec iov{hdr, sizeof(hdr)};
    char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr m;
    memset(&m, 0, sizeof(m));
    m.msg_iov = &iov; m.msg_iovlen = 1; m.msg_control = ctrl; m.msg_controllen = sizeof(ctrl);
    ssize_t n = recvmsg(x->sock, &m, MSG_DONTWAIT | MSG_CMSG_CLOEXEC);
    if (n < 0) return (errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR) ? 0 : -1;
    if (n != (ssize_t)sizeof(hdr)) return 0;
    int fd = -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&m); c; c = CMSG_NXTHDR(&m, c))
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) memcpy(&fd, CMSG_DATA(c), sizeof(int));
    x->stash.push_back(FdMsg{hdr[0], (int)hdr[1], fd});
    return 1;
}

extern "C" int ewdml_fdx_send(ewdml_fdx* x, int dst, int fd, uint64_t tag, double timeout_s) {
    EWDML_CHECK_ARG(x && dst >= 0 && dst < x->world && fd >= 0);
    uint64_t hdr[2] = {tag, (uint64_t)x->rank};
    iovec iov{hdr, sizeof(hdr)};
    char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    sockaddr_un a;
    socklen_t len = fdx_addr(x, dst, &a);
    msghdr m;
    memset(&m, 0, sizeof(m));
    m.msg_name = &a; m.msg_namelen = len; m.msg_iov = &iov; m.msg_iovlen = 1; m.msg_control = ctrl; m.msg_controllen = sizeof(ctrl);
    cmsghdr* c = CMSG_FIRSTHDR(&m);
    c->cmsg_level = SOL_SOCKET; c->cmsg_type = SCM_RIGHTS; c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    double waited = 0;
    while (true) {
        if (sendmsg(x->sock, &m, MSG_DONTWAIT) == (ssize_t)sizeof(hdr)) return 0;
        // receiver queue full (EAGAIN) or receiver not bound yet (ECONNREFUSED / ENOENT): drain our own queue, back off
        if (!(errno == EAGAIN || errno == EWOULDBLOCK || errno == ECONNREFUSED || errno == ENOENT || errno == EINTR ||
              errno == ENOBUFS)) {
            ewdml_set_error("fdx: sendmsg to rank %d: %s", dst, strerror(errno));
            return EWDML_ERR_IO;
        }
        while (fdx_drain_one(x) > 0) {}
        usleep(2000);
        waited += 0.002;
        if (waited > timeout_s) { ewdml_set_error("fdx: send to rank %d timed out", dst); return EWDML_ERR_TIMEOUT; }
    }
}

// Receive the descriptor sent with `tag` by rank `src` (src < 0: any rank; *src_out tells which).
extern "C" int ewdml_fdx_recv(ewdml_fdx* x, uint64_t tag, int src, int* fd, int* src_out, double timeout_s) {
    EWDML_CHECK_ARG(x && fd);
    double waited = 0;
    while (true) {
        for (size_t i = 0; i < x->stash.size(); ++i) {
            if (x->stash[i].tag == tag && (src < 0 || x->stash[i].src == src)) {
                *fd = x->stash[i].fd;
                if (src_out) *src_out = x->stash[i].src;
                x->stash.erase(x->stash.begin() + i);
                return 0;
            }
        }
        int r = fdx_drain_one(x);
        if (r < 0) { ewdml_set_error("fdx: recvmsg: %s", strerror(errno)); return EWDML_ERR_IO; }
        if (r > 0) continue;
        pollfd p{x->sock, POLLIN, 0};
        poll(&p, 1, 50);
        waited += 0.05;
        if (waited > timeout_s) { ewdml_set_error("fdx: receive (tag %llu) timed out", (unsigned long long)tag); return EWDML_ERR_TIMEOUT; }
    }
}

extern "C" int ewdml_fdx_close(ewdml_fdx* x) {
    if (!x) return 0;
    for (auto& m : x->stash) if (m.fd >= 0) close(m.fd);
    if (x->sock >= 0) close(x->sock);
    delete x;
    return 0;
}
