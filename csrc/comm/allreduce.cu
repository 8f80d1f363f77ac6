// K4: dense fused-scale allreduce over NVLink peer memory, in place, as ONE kernel per call.
//
// Replaces hvd.allreduce(op=Average, gradient_predivide_factor) (horvod_pytorch.py:194-201) and,
// underneath it, Open MPI's host-side ring (ompi:ompi/mca/coll/base/coll_base_allreduce.c:341:
// 2(N-1) latency-bound send/recv steps + CPU AVX reduction, ompi:ompi/mca/op/avx/op_avx_functions.c,
// with GPU buffers staged through the host, ompi:ompi/mca/coll/cuda/coll_cuda_allreduce.c:30-74).
// On NVSwitch every peer is one hop away at full bandwidth, so rings and trees buy nothing:
//
//   two-shot (default):  rank r owns shard r.  For each owned 16-byte vector it loads the W peer
//       copies (W-1 over NVLink), sums them in rank order, scales, and stores the result into all
//       W buffers.  Only rank r touches shard r of any buffer during the kernel, so the operation
//       is in place with no intermediate barrier; traffic per rank is (W-1)/W * N in and out.
//   one-shot (small N):  every rank reads all W buffers entirely and writes only its own copy
//       into a private staging area, then copies back after the exit barrier (latency-optimal).
//
// Entry barrier: "my buffer is final" -> all peers.  Exit barrier: "I am done with your buffer".
// Both are sequence-numbered flag stores with release/acquire at system scope; the exit wait is
// done by the last CTA so kernel completion implies the buffer may be overwritten by the next
// backward pass.
#include "plan.h"

// flag block layout (uint32 words, per rank): [0..P) entry flags from peer p, [P..2P) exit flags,
// [2P..3P) barrier flags.
#define FLAG_ENTRY(p) (p)
#define FLAG_EXIT(p) (EWDML_MAX_PEERS + (p))
#define FLAG_BAR(p) (2 * EWDML_MAX_PEERS + (p))
#define SEQ(a) ((a).state ? (a).state[0] : (a).seq)

struct ArArgs {
    ewdml_comm comm;
    int64_t n4;             // number of float4 vectors
    float scale;
    uint32_t seq;
    const uint32_t* state;  // optional device step state (seq at [0])
    uint32_t* work_counter; // device counter used to elect the last CTA (self-resetting)
    uint32_t* err_word;
    unsigned long long spin_limit;
};

__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

__device__ __forceinline__ void entry_barrier(const ArArgs& a, int* s_ok) {
    const int tid = threadIdx.x, W = a.comm.world, me = a.comm.rank;
    if (blockIdx.x == 0 && tid < W) {
        __threadfence_system();
        st_release_sys(a.comm.peer_flags[tid] + FLAG_ENTRY(me), SEQ(a));
    }
    if (tid == 0) *s_ok = 1;
    __syncthreads();
    if (tid < W)
        if (!wait_flag_ge(a.comm.flags + FLAG_ENTRY(tid), SEQ(a), a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT))
            *s_ok = 0;
    __syncthreads();
}

__device__ __forceinline__ void exit_barrier(const ArArgs& a) {
    __shared__ int s_last;
    const int tid = threadIdx.x, W = a.comm.world, me = a.comm.rank;
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        s_last = atomicAdd(a.work_counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    if (tid < W) {
        __threadfence_system();
        st_release_sys(a.comm.peer_flags[tid] + FLAG_EXIT(me), SEQ(a));
        wait_flag_ge(a.comm.flags + FLAG_EXIT(tid), SEQ(a), a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT);
    }
    if (tid == 0) *a.work_counter = 0;
}

__global__ void __launch_bounds__(512) k_allreduce_twoshot(const ArArgs a) {
    __shared__ int s_ok;
    entry_barrier(a, &s_ok);
    if (s_ok) {
        const int W = a.comm.world, me = a.comm.rank;
        const int64_t per = (a.n4 + W - 1) / W;
        const int64_t lo = per * me, hi = min(a.n4, lo + per);
        for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
            float4 v[EWDML_MAX_PEERS];
#pragma unroll
            for (int r = 0; r < EWDML_MAX_PEERS; ++r)
                if (r < W) {
                    uint4 u = ld_v4(reinterpret_cast<const float4*>(a.comm.bufs[r]) + i);
                    v[r] = make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w));
                }
            float4 s = v[0];
#pragma unroll
            for (int r = 1; r < EWDML_MAX_PEERS; ++r)
                if (r < W) s = add4(s, v[r]);
            s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
            uint4 o = make_uint4(__float_as_uint(s.x), __float_as_uint(s.y), __float_as_uint(s.z), __float_as_uint(s.w));
#pragma unroll
            for (int r = 0; r < EWDML_MAX_PEERS; ++r)
                if (r < W) st_na_v4(reinterpret_cast<float4*>(a.comm.bufs[r]) + i, o);
        }
    }
    exit_barrier(a);
}

// NVLS: the same ownership rule as two-shot (only rank r touches shard r, so in place and no middle barrier), but the W-way
// sum happens inside the NVSwitch (multimem.ld_reduce: one reduced vector comes back over the link instead of W-1 peer
// copies) and the result leaves as ONE multicast store instead of W unicast stores: link traffic per rank drops from
// 2(W-1)/W * N to about N in each direction, and the SM issues 2 memory operations per vector instead of 2W.
__global__ void __launch_bounds__(512) k_allreduce_nvls(const ArArgs a) {
    __shared__ int s_ok;
    entry_barrier(a, &s_ok);
    if (s_ok) {
        const int W = a.comm.world, me = a.comm.rank;
        const int64_t per = (a.n4 + W - 1) / W;
        const int64_t lo = per * me, hi = min(a.n4, lo + per);
        float4* mc = reinterpret_cast<float4*>(a.comm.mc_buf);
        const int64_t stride = (int64_t)gridDim.x * blockDim.x;
        int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < hi; i += 4 * stride) {          // four reduce-loads in flight per thread
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = mc_ld_reduce_add_v4(mc + i + u * stride);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4 s = v[u];
                mc_st_v4(mc + i + u * stride, make_uint4(__float_as_uint(s.x * a.scale), __float_as_uint(s.y * a.scale),
                                                         __float_as_uint(s.z * a.scale), __float_as_uint(s.w * a.scale)));
            }
        }
        for (; i < hi; i += stride) {
            const float4 s = mc_ld_reduce_add_v4(mc + i);
            mc_st_v4(mc + i, make_uint4(__float_as_uint(s.x * a.scale), __float_as_uint(s.y * a.scale),
                                        __float_as_uint(s.z * a.scale), __float_as_uint(s.w * a.scale)));
        }
    }
    exit_barrier(a);
}

// one-shot: result is accumulated into registers from all peers and written to the local buffer
// only AFTER every rank has finished reading it (exit barrier inside), via a staging buffer that
// lives right behind the data in the same symmetric allocation (bufs[me] + n4*4 floats).
__global__ void __launch_bounds__(512) k_allreduce_oneshot(const ArArgs a, float* __restrict__ staging) {
    __shared__ int s_ok;
    entry_barrier(a, &s_ok);
    if (s_ok) {
        const int W = a.comm.world;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n4; i += (int64_t)gridDim.x * blockDim.x) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < EWDML_MAX_PEERS; ++r)
                if (r < W) {
                    uint4 u = ld_v4(reinterpret_cast<const float4*>(a.comm.bufs[r]) + i);
                    s = add4(s, make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w)));
                }
            s.x *= a.scale; s.y *= a.scale; s.z *= a.scale; s.w *= a.scale;
            reinterpret_cast<float4*>(staging)[i] = s;
        }
    }
    exit_barrier(a);
}

__global__ void __launch_bounds__(512) k_copy4(float4* __restrict__ dst, const float4* __restrict__ src, int64_t n4) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

__global__ void k_barrier(const ArArgs a) {
    const int tid = threadIdx.x, W = a.comm.world, me = a.comm.rank;
    if (tid < W) {
        __threadfence_system();
        st_release_sys(a.comm.peer_flags[tid] + FLAG_BAR(me), SEQ(a));
        wait_flag_ge(a.comm.flags + FLAG_BAR(tid), SEQ(a), a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT);
    }
}

static int sm_count() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms > 0 ? sms : 148;
}

static int fill_args(ArArgs& a, const ewdml_comm* comm, int64_t n, float scale, uint32_t seq, uint32_t* work_counter) {
    memset(&a, 0, sizeof(a));
    a.comm = *comm; a.n4 = n / 4; a.scale = scale; a.seq = seq; a.work_counter = work_counter;
    a.err_word = ewdml_device_error_word();
    a.spin_limit = ewdml_spin_limit();
    return a.err_word ? 0 : EWDML_ERR_CUDA;
}

extern "C" int ewdml_allreduce_state(const ewdml_comm* comm, int64_t n, float scale, int algo, uint32_t seq, const uint32_t* state,
                                     uint32_t* work_counter, void* stream);
extern "C" int ewdml_allreduce(const ewdml_comm* comm, int64_t n, float scale, int algo, uint32_t seq,
                               uint32_t* work_counter, void* stream) {
    return ewdml_allreduce_state(comm, n, scale, algo, seq, nullptr, work_counter, stream);
}
extern "C" int ewdml_allreduce_state(const ewdml_comm* comm, int64_t n, float scale, int algo, uint32_t seq, const uint32_t* state,
                                     uint32_t* work_counter, void* stream) {
    EWDML_CHECK_ARG(comm && comm->world >= 1 && comm->world <= EWDML_MAX_PEERS && comm->rank >= 0 &&
                    comm->rank < comm->world && n > 0 && n % 4 == 0 && work_counter);
    ArArgs a;
    if (fill_args(a, comm, n, scale, seq, work_counter)) return EWDML_ERR_CUDA;
    a.state = state;
    cudaStream_t st = (cudaStream_t)stream;
    if (algo == EWDML_AR_AUTO)
        algo = (n * 4 <= (1 << 18)) ? EWDML_AR_ONESHOT : (comm->mc_buf && comm->world > 1 ? EWDML_AR_NVLS : EWDML_AR_TWOSHOT);
    if (algo == EWDML_AR_NVLS && !comm->mc_buf) {
        ewdml_set_error("ewdml_allreduce: NVLS requested but the buffer has no multicast mapping");
        return EWDML_ERR_ARG;
    }
    const int sms = sm_count();
    if (algo == EWDML_AR_NVLS) {
        int64_t per = (a.n4 + comm->world - 1) / comm->world;
        int grid = (int)std::min<int64_t>((per + 511) / 512, (int64_t)sms);
        if (grid < 1) grid = 1;
        k_allreduce_nvls<<<grid, 512, 0, st>>>(a);
        EWDML_LAUNCH_CHECK();
    } else if (algo == EWDML_AR_TWOSHOT) {
        int64_t per = (a.n4 + comm->world - 1) / comm->world;
        int grid = (int)std::min<int64_t>((per + 511) / 512, (int64_t)sms);
        if (grid < 1) grid = 1;
        k_allreduce_twoshot<<<grid, 512, 0, st>>>(a);
        EWDML_LAUNCH_CHECK();
    } else {
        // staging lives behind the payload: callers reserve 2*n floats per rank for one-shot
        float* staging = comm->bufs[comm->rank] + n;
        int grid = (int)std::min<int64_t>((a.n4 + 511) / 512, (int64_t)sms);
        if (grid < 1) grid = 1;
        k_allreduce_oneshot<<<grid, 512, 0, st>>>(a, staging);
        EWDML_LAUNCH_CHECK();
        k_copy4<<<grid, 512, 0, st>>>(reinterpret_cast<float4*>(comm->bufs[comm->rank]),
                                     reinterpret_cast<const float4*>(staging), a.n4);
        EWDML_LAUNCH_CHECK();
    }
    ewdml_counter_add(EWDML_CTR_ALLREDUCE_CALLS, 1);
    ewdml_counter_add(EWDML_CTR_BYTES_SENT, (int64_t)(n * 4.0 * (comm->world - 1) / comm->world));
    return 0;
}

extern "C" int ewdml_barrier(const ewdml_comm* comm, uint32_t seq, void* stream) {
    EWDML_CHECK_ARG(comm && comm->world >= 1 && comm->world <= EWDML_MAX_PEERS);
    ArArgs a;
    uint32_t dummy = 0;
    if (fill_args(a, comm, 4, 1.f, seq, &dummy)) return EWDML_ERR_CUDA;
    k_barrier<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    EWDML_LAUNCH_CHECK();
    return 0;
}
