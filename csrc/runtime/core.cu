// Error reporting, counters/tracing (N5) and the device error word.
//
// Counters replace Open MPI's software performance counters (SPC_RECORD, ompi:ompi/runtime/
// ompi_spc.h:86-159) and the pml/coll monitoring interposers (per-peer byte matrix,
// ompi:ompi/mca/common/monitoring/README:13-40); the trace switch replaces libompitrace
// (ompi:ompi/contrib/libompitrace): one line per native call on stderr.
#include <atomic>
#include <cstdarg>
#include <cstdlib>
#include <mutex>

#include "../common.cuh"

static thread_local char g_err[1024] = "";
static std::atomic<int64_t> g_ctr[EWDML_CTR_COUNT];
static std::atomic<int64_t> g_peer_bytes[EWDML_MAX_PEERS];
static std::atomic<int> g_trace{-1};
static std::atomic<uint64_t> g_spin_limit{0};

void ewdml_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    if (g_trace.load() > 0) fprintf(stderr, "[ewdml] error: %s\n", g_err);
}

void ewdml_count_launch(int n) { g_ctr[EWDML_CTR_LAUNCHES].fetch_add(n, std::memory_order_relaxed); }

extern "C" const char* ewdml_last_error(void) { return g_err; }
extern "C" int ewdml_version(void) { return 100; }

extern "C" int64_t ewdml_counter_get(int id) {
    return (id >= 0 && id < EWDML_CTR_COUNT) ? g_ctr[id].load() : -1;
}
extern "C" void ewdml_counter_add(int id, int64_t v) {
    if (id >= 0 && id < EWDML_CTR_COUNT) g_ctr[id].fetch_add(v, std::memory_order_relaxed);
    if (g_trace.load() > 0 && id != EWDML_CTR_LAUNCHES && id != EWDML_CTR_COMPUTE_LAUNCHES)
        fprintf(stderr, "[ewdml] counter %d += %lld\n", id, (long long)v);
}
extern "C" void ewdml_counters_reset(void) {
    for (auto& c : g_ctr) c.store(0);
    for (auto& c : g_peer_bytes) c.store(0);
}
extern "C" void ewdml_peer_matrix_add(int peer, int64_t bytes) {
    if (peer >= 0 && peer < EWDML_MAX_PEERS) g_peer_bytes[peer].fetch_add(bytes, std::memory_order_relaxed);
}
extern "C" int ewdml_peer_matrix_get(int64_t out[EWDML_MAX_PEERS]) {
    for (int i = 0; i < EWDML_MAX_PEERS; ++i) out[i] = g_peer_bytes[i].load();
    return 0;
}
extern "C" void ewdml_set_trace(int on) { g_trace.store(on ? 1 : 0); }

extern "C" uint64_t ewdml_spin_limit() {
    uint64_t v = g_spin_limit.load();
    if (v == 0) {
        const char* e = getenv("EWDML_SPIN_LIMIT");
        v = e ? strtoull(e, nullptr, 10) : (1ull << 26);
        if (v == 0) v = 1ull << 26;
        g_spin_limit.store(v);
        if (g_trace.load() < 0) {
            const char* t = getenv("EWDML_TRACE");
            g_trace.store(t && atoi(t) ? 1 : 0);
        }
    }
    return v;
}
extern "C" int ewdml_set_spin_limit(uint64_t iters) {
    g_spin_limit.store(iters ? iters : (1ull << 26));
    return 0;
}

// SM count of the current device, queried once per device (launch geometry of every persistent / grid-stride kernel).
extern "C" int ewdml_sm_count() {
    static std::atomic<int> cache[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = cache[dev].load(std::memory_order_relaxed);
    if (v <= 0) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cache[dev].store(v, std::memory_order_relaxed);
    }
    return v;
}

// One error word per device, allocated on first use and never freed (process lifetime).
static std::mutex g_err_mu;
static uint32_t* g_err_word[64] = {nullptr};

extern "C" uint32_t* ewdml_device_error_word() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(g_err_mu);
    if (!g_err_word[dev]) {
        uint32_t* p = nullptr;
        if (cudaMalloc(&p, 256) != cudaSuccess) { ewdml_set_error("cannot allocate device error word"); return nullptr; }
        cudaMemset(p, 0, 256);
        g_err_word[dev] = p;
    }
    return g_err_word[dev];
}

extern "C" int ewdml_get_device_error(uint32_t* out, void* stream) {
    EWDML_CHECK_ARG(out);
    uint32_t* w = ewdml_device_error_word();
    if (!w) return EWDML_ERR_CUDA;
    EWDML_CUDA(cudaMemcpyAsync(out, w, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    EWDML_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (*out) EWDML_CUDA(cudaMemsetAsync(w, 0, 4, (cudaStream_t)stream));
    return 0;
}
