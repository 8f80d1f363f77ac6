// Shared device/host helpers for the ewdml sm_100a kernels.
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../include/ewdml.h"

// ---------------------------------------------------------------------------------------------
// host-side error plumbing
// ---------------------------------------------------------------------------------------------
void ewdml_set_error(const char* fmt, ...);
void ewdml_count_launch(int n = 1);
extern "C" uint32_t* ewdml_device_error_word();      // device pointer, lazily allocated per device
extern "C" uint64_t ewdml_spin_limit();
extern "C" int ewdml_sm_count();                     // SMs of the current device (cached per device)

#define EWDML_CUDA(expr)                                                                   \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            ewdml_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                                \
        }                                                                                  \
    } while (0)

#define EWDML_CHECK_ARG(cond)                                                              \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            ewdml_set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond);         \
            return EWDML_ERR_ARG;                                                          \
        }                                                                                  \
    } while (0)

#define EWDML_LAUNCH_CHECK()                                                               \
    do {                                                                                   \
        cudaError_t _e = cudaGetLastError();                                               \
        if (_e != cudaSuccess) {                                                           \
            ewdml_set_error("%s:%d: launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return (int)_e;                                                                \
        }                                                                                  \
        ewdml_count_launch();                                                              \
    } while (0)

// ---------------------------------------------------------------------------------------------
// memory-ordering primitives for the cross-GPU flag protocol (signal/wait over NVLink)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_na_v4(void* p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
// plain (coherent) 16-byte load: payload slots are written by peers, never use .nc on them
__device__ __forceinline__ uint4 ld_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------
// NVLS multicast (multimem) primitives.  `p` is an address inside a mapped multicast object (runtime/vmm.cu): a store is
// replicated by the NVSwitch into the same offset of every bound GPU's buffer, a reduce-load returns the in-switch sum
// of all GPUs' copies.  multimem.st moves bits unchanged, so raw payload words travel through the .f32 vector form.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mc_st_v4(void* p, uint4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void mc_st_b64(void* p, uint2 v) {
    const unsigned long long x = ((unsigned long long)v.y << 32) | v.x;
    asm volatile("multimem.st.weak.global.b64 [%0], %1;" ::"l"(p), "l"(x) : "memory");
}
__device__ __forceinline__ void mc_st_b32(void* p, uint32_t v) {
    asm volatile("multimem.st.weak.global.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 mc_ld_reduce_add_v4(const void* p) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
// unicast or multicast store of one 16-byte / 8-byte / 4-byte unit (the branch is warp-uniform)
__device__ __forceinline__ void push_v4(void* p, uint4 v, bool mc) { if (mc) mc_st_v4(p, v); else st_na_v4(p, v); }
__device__ __forceinline__ void push_b64(void* p, uint2 v, bool mc) { if (mc) mc_st_b64(p, v); else *reinterpret_cast<uint2*>(p) = v; }
__device__ __forceinline__ void push_b32(void* p, uint32_t v, bool mc) { if (mc) mc_st_b32(p, v); else *reinterpret_cast<uint32_t*>(p) = v; }

// Spin until *flag >= seq (sequence numbers only grow).  Returns false on timeout and records it.
// The first EWDML_FAST_POLLS polls are back to back (a local acquire load is ~0.2 us, so that is the first
// millisecond, where every exchange of a healthy job completes); after that each poll sleeps ~1 us, which turns
// `limit` into a time budget (2^26 polls = about a minute) that does not depend on how fast an L2 hit is, and
// keeps hundreds of spinning CTAs from hammering the memory system while a peer is slow (first step, graph
// capture, checkpoint).  Round 1 counted raw polls: 2^24 of them lasted ~4 s, not the documented 20 s, and an
// 8-rank ResNet-50 parameter-server job tripped it while the workers were still capturing their graphs.
#define EWDML_FAST_POLLS 4096
__device__ __forceinline__ bool wait_flag_ge(const uint32_t* flag, uint32_t seq, uint64_t limit,
                                             uint32_t* err_word, uint32_t err_code) {
    uint64_t it = 0;
    while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
        if (++it > limit) {
            atomicExch(err_word, err_code);
            return false;
        }
        if (it > EWDML_FAST_POLLS) __nanosleep(1000);
        else if ((it & 63) == 0) __nanosleep(64);
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Random123); must match ewdml/compress/rng.py bit for bit.
// counter = (idx/4 low, stream, idx/4 high, 0); key = (seed, step); lane idx%4; u = (x>>8)*2^-24
// ---------------------------------------------------------------------------------------------
struct Philox4 {
    uint32_t v[4];
};
__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    Philox4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}
__device__ __forceinline__ float philox_to_uniform(uint32_t x) {
    return (float)(x >> 8) * 5.9604644775390625e-08f;   // 2^-24, exact
}

// ---------------------------------------------------------------------------------------------
// block-level helpers (blockDim.x threads, power of two multiple of 32, <= 1024)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}
// exclusive prefix of v over the block; *total receives the block sum.  smem: >= 33 words.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* smem, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint32_t inc = warp_incl_scan(v);
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < nw ? smem[lane] : 0;
        uint32_t wi = warp_incl_scan(w);
        smem[lane] = wi - w;
        if (lane == 31) smem[32] = wi;
    }
    __syncthreads();
    uint32_t res = smem[warp] + inc - v;
    *total = smem[32];
    __syncthreads();
    return res;
}
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}
// deterministic block sum (fixed tree); result valid in all threads.  smem: >= 32 T.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = warp_sum(v);
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    T r = lane < nw ? smem[lane] : T(0);
    r = warp_sum(r);
    __syncthreads();
    return r;
}

__host__ __device__ __forceinline__ int64_t round_up_i64(int64_t x, int64_t a) { return (x + a - 1) / a * a; }
__host__ __device__ __forceinline__ uint32_t align16(uint32_t x) { return (x + 15u) & ~15u; }
