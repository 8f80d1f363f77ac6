// K2: receive side -- wait for the senders' flags, unpack + dequantise every payload with its own
// norm, accumulate in rank order, divide by the number of contributors and apply momentum-SGD (or
// write the dense average / assign weights) in ONE pass over the parameters.
//
// Replaces: QSGD decompress (Compresssor/qsgd.py:34-40), top-k desparsify (TopK.py:13-17), the
// server's sum and divide (sync_replicas_master_nn.py:189,216), the per-tensor broadcast receive
// (distributed_worker.py:269-279) and the optimizer step (optim/sgd.py:59-92) -- six passes over
// the model collapsed into one; the fp32 master weights, the momentum buffer and the bf16 shadow
// weights used by the forward pass are each read/written exactly once.
#include "plan.h"

struct ApplyArgs {
    const TileDesc* tiles;
    SlotLayout layout;
    ewdml_sources srcs[2];             // [parity]
    const uint32_t* state;             // optional device step state {seq, step, first}
    int mode, val, code_bits, levels, ntiles, out_mode;
    uint32_t seq;
    float divisor;
    ewdml_sgd sgd;
    float* param;
    float* momentum;
    __nv_bfloat16* param_bf16;
    float* out;
    uint32_t* err_word;
    unsigned long long spin_limit;
};

__device__ __forceinline__ uint32_t get_bits(const uint32_t* words, uint32_t bitpos, int nbits) {
    uint32_t w = bitpos >> 5, sh = bitpos & 31;
    uint32_t lo = words[w] >> sh;
    if (sh + nbits > 32) lo |= words[w + 1] << (32 - sh);
    return nbits == 32 ? lo : (lo & ((1u << nbits) - 1u));
}

__device__ __forceinline__ float decode_value(uint32_t code, int cw, int val, float dq) {
    if (val == EWDML_VAL_RAW32) return __uint_as_float(code);
    const uint32_t sign = code >> (cw - 1);
    const float lvl = (float)((code & ((1u << (cw - 1)) - 1u)) + 1u);
    return __fmul_rn(dq, sign ? -lvl : lvl);
}

// Momentum-SGD coefficients resolved once per kernel so the per-element update is branch-free:
//   d   = g + wd * p
//   buf = cm * m + cd * d          (first step: cm = 0, cd = 1  ->  buf = d, optim/sgd.py:79-82)
//   d2  = cn1 * d + cn2 * buf      (nesterov: d + mu * buf; plain: buf; no momentum: d)
//   p  -= lr * d2
struct SgdCoef {
    float wd, cm, cd, cn1, cn2, lr;
    bool use_m;
};
__device__ __forceinline__ SgdCoef make_coef(const ApplyArgs& a, bool first_step) {
    SgdCoef c;
    c.use_m = a.sgd.momentum != 0.f && a.momentum != nullptr;
    c.wd = a.sgd.weight_decay;
    c.cm = first_step ? 0.f : a.sgd.momentum;
    c.cd = first_step ? 1.f : 1.f - a.sgd.dampening;
    c.cn1 = c.use_m ? (a.sgd.nesterov ? 1.f : 0.f) : 1.f;
    c.cn2 = c.use_m ? (a.sgd.nesterov ? a.sgd.momentum : 1.f) : 0.f;
    c.lr = a.sgd.lr;
    return c;
}
__device__ __forceinline__ void sgd_update4(const ApplyArgs& a, const SgdCoef& c, size_t idx, float4 g) {
    const float4 p = *reinterpret_cast<const float4*>(a.param + idx);
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c.use_m && c.cm != 0.f) m = *reinterpret_cast<const float4*>(a.momentum + idx);
    float gg[4] = {g.x, g.y, g.z, g.w}, pp[4] = {p.x, p.y, p.z, p.w}, mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float d = fmaf(c.wd, pp[q], gg[q]);
        const float buf = fmaf(c.cd, d, c.cm * mm[q]);
        mm[q] = buf;
        const float d2 = fmaf(c.cn2, buf, c.cn1 * d);
        pp[q] = fmaf(-c.lr, d2, pp[q]);
    }
    *reinterpret_cast<float4*>(a.param + idx) = make_float4(pp[0], pp[1], pp[2], pp[3]);
    if (c.use_m) *reinterpret_cast<float4*>(a.momentum + idx) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    if (a.param_bf16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(pp[0], pp[1]), hi = __floats2bfloat162_rn(pp[2], pp[3]);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(a.param_bf16 + idx) = pk;
    }
}

__global__ void __launch_bounds__(256) k_recv_apply(const ApplyArgs a) {
    __shared__ __align__(16) float acc[EWDML_TILE];
    __shared__ uint32_t scan_smem[40];
    __shared__ int s_ok;
    const int tid = threadIdx.x;
    const uint32_t seq = a.state ? a.state[0] : a.seq;
    const ewdml_sources& S = a.srcs[a.state ? (seq & 1u) : 0u];
    // ---- acquire: every CTA waits until all sources have published this sequence number -------
    if (tid == 0) s_ok = 1;
    __syncthreads();
    if (tid < S.n && S.flag[tid] != nullptr) {
        if (!wait_flag_ge(S.flag[tid], seq, a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT)) s_ok = 0;
    }
    __syncthreads();
    if (!s_ok) return;
    const float inv_scale = a.divisor;
    // a power-of-two contributor count (1, 2, 4, 8 GPUs) divides exactly by multiplying with its reciprocal
    const bool pow2 = (__float_as_uint(inv_scale) & 0x007FFFFFu) == 0u;
    const float rcp = __fdiv_rn(1.f, inv_scale);
    const bool first_step = a.state ? (a.state[2] != 0) : (a.sgd.first_step != 0);
    const SgdCoef coef = make_coef(a, first_step);
    for (int t = blockIdx.x; t < a.ntiles; t += gridDim.x) {
        const TileDesc td = a.tiles[t];
        const uint32_t n4 = (td.len + 3) >> 2;
        if (a.mode != EWDML_MODE_DENSE) {
            for (uint32_t i = tid; i < n4; i += 256) reinterpret_cast<float4*>(acc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            __syncthreads();
            for (int s = 0; s < S.n; ++s) {
                const char* slot = (const char*)S.slot[s];
                const uint2 de = *reinterpret_cast<const uint2*>(slot + a.layout.dir + (size_t)t * 8);
                const uint32_t cnt = de.y;
                const float norm = reinterpret_cast<const float*>(slot + a.layout.norms)[td.seg];
                const float dq = __fdiv_rn(norm, (float)a.levels);
                const char* chunk = slot + a.layout.data + de.x;
                if (a.mode == EWDML_MODE_IDX16) {
                    const uint16_t* idx = reinterpret_cast<const uint16_t*>(chunk);
                    const uint32_t* codes = reinterpret_cast<const uint32_t*>(chunk + align16(cnt * 2));
                    for (uint32_t j = tid; j < cnt; j += 256) {
                        const uint32_t e = idx[j];
                        const uint32_t code = get_bits(codes, j * a.code_bits, a.code_bits);
                        if (e < td.len) acc[e] += decode_value(code, a.code_bits, a.val, dq);   // e unique per source
                    }
                } else {  // BITMAP (thread-major, see encode.cu): word t, bit 4j+q  <->  element (t + 256 j) * 4 + q
                    const uint32_t word = reinterpret_cast<const uint32_t*>(slot + a.layout.bitmap)[(size_t)t * 256 + tid];
                    uint32_t total;
                    uint32_t pos = block_excl_scan(__popc(word), scan_smem, &total);
                    const uint32_t* codes = reinterpret_cast<const uint32_t*>(chunk);
                    uint32_t m = word;
                    while (m) {
                        const int b = __ffs(m) - 1;
                        m &= m - 1;
                        if (pos < cnt) {
                            const uint32_t code = get_bits(codes, pos * a.code_bits, a.code_bits);
                            acc[((tid + 256 * (b >> 2)) << 2) + (b & 3)] += decode_value(code, a.code_bits, a.val, dq);
                        }
                        ++pos;
                    }
                }
                __syncthreads();
            }
        }
        // ---- one pass over the parameters of this tile ---------------------------------------
        for (uint32_t i = tid; i < n4; i += 256) {
            float4 g;
            if (a.mode == EWDML_MODE_DENSE) {
                g = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int s = 0; s < S.n; ++s) {
                    const float4 v = *reinterpret_cast<const float4*>(
                        (const char*)S.slot[s] + a.layout.data + ((size_t)td.start + i * 4) * 4);
                    g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
                }
            } else {
                g = reinterpret_cast<const float4*>(acc)[i];
            }
            if (pow2) {
                g.x *= rcp; g.y *= rcp; g.z *= rcp; g.w *= rcp;
            } else {
                g.x = __fdiv_rn(g.x, inv_scale); g.y = __fdiv_rn(g.y, inv_scale);
                g.z = __fdiv_rn(g.z, inv_scale); g.w = __fdiv_rn(g.w, inv_scale);
            }
            const size_t idx = (size_t)td.start + i * 4;
            if (a.out_mode == EWDML_OUT_SGD) {
                sgd_update4(a, coef, idx, g);
            } else if (a.out_mode == EWDML_OUT_GRAD) {
                *reinterpret_cast<float4*>(a.out + idx) = g;
            } else {  // ASSIGN: weights pulled from the server
                *reinterpret_cast<float4*>(a.param + idx) = g;
                if (a.param_bf16) {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(g.x, g.y), hi = __floats2bfloat162_rn(g.z, g.w);
                    uint2 pk;
                    pk.x = *reinterpret_cast<uint32_t*>(&lo);
                    pk.y = *reinterpret_cast<uint32_t*>(&hi);
                    *reinterpret_cast<uint2*>(a.param_bf16 + idx) = pk;
                }
            }
        }
        __syncthreads();
    }
}

// Local fused SGD (no exchange): optim/sgd.py:59-92 over the flat buffers.
__global__ void __launch_bounds__(256) k_sgd_apply(const float* __restrict__ grad, ApplyArgs a, int64_t n4, float gscale) {
    const SgdCoef coef = make_coef(a, a.state ? (a.state[2] != 0) : (a.sgd.first_step != 0));
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
        float4 g = *reinterpret_cast<const float4*>(grad + i * 4);
        g.x *= gscale; g.y *= gscale; g.z *= gscale; g.w *= gscale;
        sgd_update4(a, coef, (size_t)i * 4, g);
    }
}

static int num_sms() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

static int recv_impl(ewdml_plan* p, const ewdml_sources* srcs, const ewdml_sources* srcs_odd, const uint32_t* state, uint32_t seq,
                     float divisor, int out_mode, const ewdml_sgd* sgd, float* param, float* momentum, void* param_bf16, float* out,
                     void* stream) {
    EWDML_CHECK_ARG(p && srcs && srcs->n >= 1 && srcs->n <= EWDML_MAX_PEERS && divisor != 0.f);
    EWDML_CHECK_ARG(out_mode != EWDML_OUT_SGD || (sgd && param));
    EWDML_CHECK_ARG(out_mode != EWDML_OUT_GRAD || out);
    EWDML_CHECK_ARG(out_mode != EWDML_OUT_ASSIGN || param);
    ApplyArgs a;
    memset(&a, 0, sizeof(a));
    a.tiles = p->d_tile; a.layout = p->layout; a.srcs[0] = *srcs; a.srcs[1] = srcs_odd ? *srcs_odd : *srcs;
    a.state = state;
    a.mode = p->codec.mode; a.val = p->codec.val; a.code_bits = p->codec.code_bits; a.levels = p->codec.levels;
    a.ntiles = p->ntiles; a.out_mode = out_mode; a.seq = seq; a.divisor = divisor;
    if (sgd) a.sgd = *sgd;
    a.param = param; a.momentum = momentum; a.param_bf16 = (__nv_bfloat16*)param_bf16; a.out = out;
    a.err_word = ewdml_device_error_word();
    a.spin_limit = ewdml_spin_limit();
    if (!a.err_word) return EWDML_ERR_CUDA;
    int grid = p->ntiles < 4 * num_sms() ? p->ntiles : 4 * num_sms();
    k_recv_apply<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    EWDML_LAUNCH_CHECK();
    ewdml_counter_add(EWDML_CTR_RECV_CALLS, 1);
    return 0;
}

extern "C" int ewdml_recv_apply(ewdml_plan* p, const ewdml_sources* srcs, uint32_t seq, float divisor, int out_mode,
                                const ewdml_sgd* sgd, float* param, float* momentum, void* param_bf16, float* out,
                                void* stream) {
    return recv_impl(p, srcs, nullptr, nullptr, seq, divisor, out_mode, sgd, param, momentum, param_bf16, out, stream);
}

extern "C" int ewdml_recv_apply_state(ewdml_plan* p, const ewdml_sources* srcs_even, const ewdml_sources* srcs_odd,
                                      const uint32_t* state, float divisor, int out_mode, const ewdml_sgd* sgd, float* param,
                                      float* momentum, void* param_bf16, float* out, void* stream) {
    EWDML_CHECK_ARG(state && srcs_odd);
    return recv_impl(p, srcs_even, srcs_odd, state, 0, divisor, out_mode, sgd, param, momentum, param_bf16, out, stream);
}

// Device step state {seq, step, first, applies}.  Called once at the start of every optimizer application:
// step += 1 (RNG stream position), first = (no update applied yet), applies += 1, and for exchanges seq += 1
// (flag sequence number / slot parity).  Everything downstream reads the state from memory, so a captured
// CUDA graph replays correctly.
__global__ void k_state_advance(uint32_t* st, int bump_seq) {
    if (bump_seq) st[0] = st[0] + 1;
    st[1] = st[1] + 1;
    st[2] = st[3] == 0 ? 1u : 0u;
    st[3] = st[3] + 1;
}
extern "C" int ewdml_state_advance(uint32_t* state, int bump_seq, void* stream) {
    EWDML_CHECK_ARG(state);
    k_state_advance<<<1, 1, 0, (cudaStream_t)stream>>>(state, bump_seq);
    EWDML_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_sgd_apply_state(const float* grad, float* param, float* momentum, void* param_bf16, int64_t n,
                                     const ewdml_sgd* sgd, float grad_scale, const uint32_t* state, void* stream);
extern "C" int ewdml_sgd_apply(const float* grad, float* param, float* momentum, void* param_bf16, int64_t n,
                               const ewdml_sgd* sgd, float grad_scale, void* stream) {
    return ewdml_sgd_apply_state(grad, param, momentum, param_bf16, n, sgd, grad_scale, nullptr, stream);
}
extern "C" int ewdml_sgd_apply_state(const float* grad, float* param, float* momentum, void* param_bf16, int64_t n,
                                     const ewdml_sgd* sgd, float grad_scale, const uint32_t* state, void* stream) {
    EWDML_CHECK_ARG(grad && param && sgd && n > 0 && n % 4 == 0);
    ApplyArgs a;
    memset(&a, 0, sizeof(a));
    a.sgd = *sgd; a.param = param; a.momentum = momentum; a.param_bf16 = (__nv_bfloat16*)param_bf16; a.state = state;
    int64_t n4 = n / 4;
    int64_t want = (n4 + 255) / 256;
    int grid = (int)(want < 8 * num_sms() ? want : 8 * num_sms());
    k_sgd_apply<<<grid, 256, 0, (cudaStream_t)stream>>>(grad, a, n4, grad_scale);
    EWDML_LAUNCH_CHECK();
    return 0;
}
