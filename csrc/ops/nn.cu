// K5 (operand builders), K7, K8, K9: the memory-bound layers around the tcgen05 GEMM.
//
// Activations are bf16, NHWC.  A convolution pass is
//   fprop : col = im2col(x)            ; y  = col  * W^T      (GEMM, K-major/K-major)
//   wgrad : dW += dy^T * col                                   (GEMM, MN-major/MN-major, split-K, fp32 atomics
//                                                               straight into the flat gradient buffer)
//   dgrad : dcol = dy * W               ; dx = col2im(dcol)    (GEMM, K-major/MN-major)
// and BatchNorm / ReLU / residual add / 2x2 max-pool are fused into one statistics pass plus one
// normalise-activate(-pool) pass forward, and one reduce pass plus one apply pass backward.
// Reference sites: nn.Conv2d / BatchNorm2d / ReLU / MaxPool2d / avg_pool2d / Dropout / CrossEntropyLoss in
// src/model_ops/{lenet,vgg,resnet}.py and distributed_worker.py:103 -- all ATen library calls there.
#include <mutex>

#include "../common.cuh"

namespace {

struct ConvGeom {
    int N, H, W, C;          // input
    int KH, KW, stride, pad;
    int HO, WO;              // output spatial
    int K, Kp;               // K = KH*KW*C, Kp = row pitch of col (multiple of 8)
};

__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xFFFF0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&p);
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
    f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// ---------------------------------------------------------------------------------------------
// im2col / col2im
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_im2col_vec8(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g, int64_t total) {
    const int c8n = g.C >> 3, taps = g.KH * g.KW;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c8 = (int)(t % c8n);
        const int64_t r = t / c8n;
        const int tap = (int)(r % taps);
        const int64_t m = r / taps;
        const int wo = (int)(m % g.WO), ho = (int)((m / g.WO) % g.HO), n = (int)(m / ((int64_t)g.WO * g.HO));
        const int h = ho * g.stride - g.pad + tap / g.KW, w = wo * g.stride - g.pad + tap % g.KW;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (h >= 0 && h < g.H && w >= 0 && w < g.W)
            v = *reinterpret_cast<const uint4*>(x + (((int64_t)n * g.H + h) * g.W + w) * g.C + c8 * 8);
        *reinterpret_cast<uint4*>(col + m * g.Kp + (int64_t)tap * g.C + c8 * 8) = v;
    }
}

// any C (1, 3, 20 ...): one thread builds 8 consecutive k entries (one 16-byte store), zero padding of [K, Kp)
__global__ void __launch_bounds__(256) k_im2col_scalar(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, ConvGeom g, int64_t total) {
    const int kgn = g.Kp >> 3;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int kg = (int)(t % kgn);
        const int64_t m = t / kgn;
        const int wo = (int)(m % g.WO), ho = (int)((m / g.WO) % g.HO), n = (int)(m / ((int64_t)g.WO * g.HO));
        const int h0 = ho * g.stride - g.pad, w0 = wo * g.stride - g.pad;
        const __nv_bfloat16* img = x + (int64_t)n * g.H * g.W * g.C;
        unsigned short v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int kidx = kg * 8 + i;
            v[i] = 0;
            if (kidx < g.K) {
                const int tap = kidx / g.C, c = kidx - tap * g.C;
                const int kh = tap / g.KW, kw = tap - kh * g.KW;
                const int h = h0 + kh, w = w0 + kw;
                if (h >= 0 && h < g.H && w >= 0 && w < g.W)
                    v[i] = *reinterpret_cast<const unsigned short*>(img + ((int64_t)h * g.W + w) * g.C + c);
            }
        }
        uint4 pk;
        pk.x = v[0] | ((uint32_t)v[1] << 16); pk.y = v[2] | ((uint32_t)v[3] << 16);
        pk.z = v[4] | ((uint32_t)v[5] << 16); pk.w = v[6] | ((uint32_t)v[7] << 16);
        *reinterpret_cast<uint4*>(col + m * g.Kp + kg * 8) = pk;
    }
}

// dx[n,h,w,c] = sum over taps of dcol[m(n,ho,wo)][tap][c]   (gather form: no atomics)
template <int VEC>
__global__ void __launch_bounds__(256) k_col2im(const __nv_bfloat16* __restrict__ dcol, __nv_bfloat16* __restrict__ dx, ConvGeom g, int64_t total) {
    const int cvn = g.C / VEC;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int cv = (int)(t % cvn);
        const int64_t p = t / cvn;
        const int w = (int)(p % g.W), h = (int)((p / g.W) % g.H), n = (int)(p / ((int64_t)g.W * g.H));
        float acc[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
        for (int kh = 0; kh < g.KH; ++kh) {
            const int hn = h + g.pad - kh;
            if (hn < 0 || hn % g.stride) continue;
            const int ho = hn / g.stride;
            if (ho >= g.HO) continue;
            for (int kw = 0; kw < g.KW; ++kw) {
                const int wn = w + g.pad - kw;
                if (wn < 0 || wn % g.stride) continue;
                const int wo = wn / g.stride;
                if (wo >= g.WO) continue;
                const int64_t m = ((int64_t)n * g.HO + ho) * g.WO + wo;
                const __nv_bfloat16* src = dcol + m * g.Kp + (int64_t)(kh * g.KW + kw) * g.C + cv * VEC;
                if (VEC == 8) {
                    float f[8];
                    unpack8(*reinterpret_cast<const uint4*>(src), f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[i] += f[i];
                } else {
                    acc[0] += __bfloat162float(src[0]);
                }
            }
        }
        __nv_bfloat16* dst = dx + p * g.C + cv * VEC;
        if (VEC == 8) {
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = acc[i];
            *reinterpret_cast<uint4*>(dst) = pack8(f);
        } else {
            dst[0] = __float2bfloat16_rn(acc[0]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// First-layer convolution (C_in <= 4, e.g. RGB): K = KH*KW*C_in is far too small for a tensor-core tile, and
// an im2col buffer would be 10x the size of the image.  Direct form: one thread computes 8 output channels of one
// output pixel with the filter bank in shared memory; reads the image once (fp32 or bf16), writes y once.
// ---------------------------------------------------------------------------------------------
template <typename TIn>
__global__ void __launch_bounds__(256) k_conv_small_fwd(const TIn* __restrict__ x, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
                                                        __nv_bfloat16* __restrict__ y, ConvGeom g, int Cout, int64_t total) {
    extern __shared__ float s_w[];                      // [K][Cout]
    for (int i = threadIdx.x; i < g.K * Cout; i += 256) {
        const int co = i / g.K, k = i - co * g.K;
        s_w[k * Cout + co] = __bfloat162float(w[i]);
    }
    __syncthreads();
    const int c8n = Cout >> 3;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c8 = (int)(t % c8n);
        const int64_t m = t / c8n;
        const int wo = (int)(m % g.WO), ho = (int)((m / g.WO) % g.HO), n = (int)(m / ((int64_t)g.WO * g.HO));
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = bias ? bias[c8 * 8 + i] : 0.f;
        for (int kh = 0; kh < g.KH; ++kh) {
            const int h = ho * g.stride - g.pad + kh;
            if (h < 0 || h >= g.H) continue;
            for (int kw = 0; kw < g.KW; ++kw) {
                const int wi = wo * g.stride - g.pad + kw;
                if (wi < 0 || wi >= g.W) continue;
                const TIn* px = x + (((int64_t)n * g.H + h) * g.W + wi) * g.C;
                for (int c = 0; c < g.C; ++c) {
                    // the GEMM path rounds the input to bf16 first; do the same so both paths agree bit for bit on inputs
                    const float xv = __bfloat162float(__float2bfloat16_rn((float)px[c]));
                    const float* wr = s_w + ((kh * g.KW + kw) * g.C + c) * Cout + c8 * 8;
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[i] = fmaf(xv, wr[i], acc[i]);
                }
            }
        }
        *reinterpret_cast<uint4*>(y + m * Cout + c8 * 8) = pack8(acc);
    }
}

// dW[co][k] += sum_m dy[m][co] * x_patch[m][k]: thread = (co, tap group); pixels are staged in batches of 64
template <typename TIn>
__global__ void __launch_bounds__(256) k_conv_small_wgrad(const TIn* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw, ConvGeom g,
                                                          int Cout, int64_t M) {
    constexpr int PB = 64, KMAX = 36;
    __shared__ float s_x[PB][KMAX + 1];
    __shared__ float s_dy[PB][128];
    const int groups = 256 / Cout;                       // Cout in {64, 128}
    const int co = threadIdx.x % Cout, grp = threadIdx.x / Cout;
    float acc[20];
#pragma unroll
    for (int i = 0; i < 20; ++i) acc[i] = 0.f;
    const int64_t nb = (M + PB - 1) / PB;
    for (int64_t b = blockIdx.x; b < nb; b += gridDim.x) {
        const int64_t m0 = b * PB;
        for (int i = threadIdx.x; i < PB * g.K; i += 256) {
            const int p = i / g.K, k = i - p * g.K;
            const int64_t m = m0 + p;
            float v = 0.f;
            if (m < M) {
                const int c = k % g.C, tap = k / g.C;
                const int wo = (int)(m % g.WO), ho = (int)((m / g.WO) % g.HO), n = (int)(m / ((int64_t)g.WO * g.HO));
                const int h = ho * g.stride - g.pad + tap / g.KW, wi = wo * g.stride - g.pad + tap % g.KW;
                if (h >= 0 && h < g.H && wi >= 0 && wi < g.W)
                    v = __bfloat162float(__float2bfloat16_rn((float)x[(((int64_t)n * g.H + h) * g.W + wi) * g.C + c]));
            }
            s_x[p][k] = v;
        }
        for (int i = threadIdx.x; i < PB * Cout; i += 256) {
            const int p = i / Cout, c = i - p * Cout;
            s_dy[p][c] = (m0 + p < M) ? __bfloat162float(dy[(m0 + p) * Cout + c]) : 0.f;
        }
        __syncthreads();
        for (int p = 0; p < PB; ++p) {
            const float d = s_dy[p][co];
#pragma unroll
            for (int j = 0; j < 20; ++j) {
                const int k = grp + j * groups;
                if (k < g.K) acc[j] = fmaf(d, s_x[p][k], acc[j]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 20; ++j) {
        const int k = grp + j * groups;
        if (k < g.K) atomicAdd(&dw[co * g.K + k], acc[j]);
    }
}

// compact NHWC -> zero-bordered [N][H+2][W+2][C] (operand layout of the implicit-GEMM 3x3 convolution)
__global__ void __launch_bounds__(256) k_pad_nhwc(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int H, int W, int C, int64_t total) {
    const int c8n = C >> 3, WP = W + 2, HP = H + 2;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c8 = (int)(t % c8n);
        const int64_t r = t / c8n;
        const int wp = (int)(r % WP), hp = (int)((r / WP) % HP), n = (int)(r / ((int64_t)WP * HP));
        uint4 v = make_uint4(0, 0, 0, 0);
        if (hp >= 1 && hp <= H && wp >= 1 && wp <= W)
            v = *reinterpret_cast<const uint4*>(x + (((int64_t)n * H + hp - 1) * W + wp - 1) * C + c8 * 8);
        *reinterpret_cast<uint4*>(out + r * C + c8 * 8) = v;
    }
}

// ---------------------------------------------------------------------------------------------
// per-channel column statistics of a [M, C] bf16 matrix: sum and (optionally) sum of squares, or the
// two BatchNorm-backward sums.  Thread = one channel pair; a CTA covers `rows_par` rows at a time.
// ---------------------------------------------------------------------------------------------
struct ActGeom {
    int64_t M;               // rows of the un-pooled activation
    int C;
    int pool;                // 1: dz_src / output are 2x2-pooled (H, W even)
    int N, H, W;             // needed when pool == 1
    int relu;
    int has_affine;          // scale/shift present (BatchNorm); else identity
};

// z value before activation for element (row, c): scale*y + shift (+ residual)
__device__ __forceinline__ float pre_act(float y, float sc, float sh, float res) { return fmaf(sc, y, sh) + res; }

// Fold NV per-thread partial sums into the CTA's shared accumulator red[lane*16 + i].  Threads whose index differs by a
// multiple of `lanes` hold partials of the same channels; when lanes < 32 those threads sit in the same warp, so they are
// combined with shuffles first and only one thread per (warp, channel group) touches shared memory.
template <int NV>
__device__ __forceinline__ void fold_partials(float (&v)[NV], float* red, int lanes, int lane, int base) {
    const bool shuffle = lanes < 32 && (lanes & (lanes - 1)) == 0;
    if (shuffle) {
        for (int d = lanes; d < 32; d <<= 1)
#pragma unroll
            for (int i = 0; i < NV; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], d);
        if ((int)(threadIdx.x & 31) >= lanes) return;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) atomicAdd(&red[lane * 16 + base + i], v[i]);
}

// 8 channels per thread (one 16-byte load per row), `lanes` threads along channels, 256/lanes rows in flight
// per CTA; per-CTA reduction in shared memory, then one global atomic per channel per CTA.
__global__ void __launch_bounds__(256) k_colstats_v8(const __nv_bfloat16* __restrict__ y, float* __restrict__ sums, int64_t M, int C, int want_sq,
                                                     int rows_per_cta) {
    __shared__ float red[256 * 16];
    const int c8n = C >> 3;
    const int lanes = c8n < 256 ? c8n : 256;
    const int rows_par = 256 / lanes;
    const int lane = threadIdx.x % lanes, rsub = threadIdx.x / lanes;
    const int c8 = blockIdx.x * lanes + lane;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < M ? r0 + rows_per_cta : M;
    for (int i = threadIdx.x; i < lanes * 16; i += 256) red[i] = 0.f;
    __syncthreads();
    {
        float s[8], q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
        if (c8 < c8n && rsub < rows_par) {
            for (int64_t r = r0 + rsub; r < r1; r += rows_par) {
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(y + r * C + c8 * 8), f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
            }
        }
        if (lanes * rows_par == 256 || (c8 < c8n && rsub < rows_par)) {
            fold_partials<8>(s, red, lanes, lane, 0);
            if (want_sq) fold_partials<8>(q, red, lanes, lane, 8);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < lanes * 16; i += 256) {
        const int l = i >> 4, j = i & 15, c = (blockIdx.x * lanes + l) * 8 + (j & 7);
        if (c < C && (j < 8 || want_sq)) atomicAdd(&sums[(j < 8 ? 0 : C) + c], red[i]);
    }
}

__global__ void __launch_bounds__(256) k_colstats(const __nv_bfloat16* __restrict__ y, float* __restrict__ sums, int64_t M, int C, int want_sq,
                                                  int rows_per_cta) {
    const int cp_n = C >> 1;                                  // channel pairs
    const int lanes = cp_n < 256 ? cp_n : 256;                // threads along channels in this CTA
    const int rows_par = 256 / lanes;
    const int cp = blockIdx.x * lanes + (threadIdx.x % lanes);
    const int rsub = threadIdx.x / lanes;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < M ? r0 + rows_per_cta : M;
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
    if (cp < cp_n && rsub < rows_par) {
        for (int64_t r = r0 + rsub; r < r1; r += rows_par) {
            const uint32_t v = *reinterpret_cast<const uint32_t*>(y + r * C + cp * 2);
            const float a = bf16lo(v), b = bf16hi(v);
            s0 += a; s1 += b;
            q0 = fmaf(a, a, q0); q1 = fmaf(b, b, q1);
        }
        atomicAdd(&sums[cp * 2], s0);
        atomicAdd(&sums[cp * 2 + 1], s1);
        if (want_sq) {
            atomicAdd(&sums[C + cp * 2], q0);
            atomicAdd(&sums[C + cp * 2 + 1], q1);
        }
    }
}

// split-K epilogue: fp32 accumulator [M][C] (+bias) -> bf16 [M][C], optionally accumulating the BatchNorm column
// statistics of the ROUNDED values in the same pass (so the separate statistics read disappears)
__global__ void __launch_bounds__(256) k_f32_to_bf16_stats(const float* __restrict__ src, const float* __restrict__ bias, __nv_bfloat16* __restrict__ dst,
                                                           float* __restrict__ sums, int64_t M, int C, int rows_per_cta) {
    __shared__ float red[256 * 16];
    const int c8n = C >> 3;
    const int lanes = c8n < 256 ? c8n : 256;
    const int rows_par = 256 / lanes;
    const int lane = threadIdx.x % lanes, rsub = threadIdx.x / lanes;
    const int c8 = blockIdx.x * lanes + lane;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < M ? r0 + rows_per_cta : M;
    if (sums) {
        for (int i = threadIdx.x; i < lanes * 16; i += 256) red[i] = 0.f;
        __syncthreads();
    }
    if (c8 < c8n && rsub < rows_par) {
        float b[8], s[8], q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { b[i] = bias ? bias[c8 * 8 + i] : 0.f; s[i] = 0.f; q[i] = 0.f; }
        for (int64_t r = r0 + rsub; r < r1; r += rows_par) {
            const float4 a0 = *reinterpret_cast<const float4*>(src + r * C + c8 * 8);
            const float4 a1 = *reinterpret_cast<const float4*>(src + r * C + c8 * 8 + 4);
            float f[8] = {a0.x + b[0], a0.y + b[1], a0.z + b[2], a0.w + b[3], a1.x + b[4], a1.y + b[5], a1.z + b[6], a1.w + b[7]};
            const uint4 pk = pack8(f);
            *reinterpret_cast<uint4*>(dst + r * C + c8 * 8) = pk;
            if (sums) {
                float fr[8];
                unpack8(pk, fr);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += fr[i]; q[i] = fmaf(fr[i], fr[i], q[i]); }
            }
        }
        if (sums) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { atomicAdd(&red[lane * 16 + i], s[i]); atomicAdd(&red[lane * 16 + 8 + i], q[i]); }
        }
    }
    if (sums) {
        __syncthreads();
        for (int i = threadIdx.x; i < lanes * 16; i += 256) {
            const int l = i >> 4, j = i & 15, c = (blockIdx.x * lanes + l) * 8 + (j & 7);
            if (c < C) atomicAdd(&sums[(j < 8 ? 0 : C) + c], red[i]);
        }
    }
}

// mean/var -> scale/shift, running statistics update (momentum, unbiased variance), like nn.BatchNorm2d
__global__ void k_bn_finalize(const float* __restrict__ sums, const float* __restrict__ gamma, const float* __restrict__ beta,
                              float* __restrict__ running_mean, float* __restrict__ running_var, float* __restrict__ stats /* [4][C]: mean, rstd, scale, shift */,
                              int C, float inv_m, float unbias, float momentum, float eps, long long* __restrict__ num_batches) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && num_batches) *num_batches += 1;
    if (c >= C) return;
    const float mean = sums[c] * inv_m;
    float var = fmaf(-mean, mean, sums[C + c] * inv_m);
    var = var > 0.f ? var : 0.f;
    const float rstd = rsqrtf(var + eps);
    const float sc = gamma[c] * rstd;
    stats[c] = mean; stats[C + c] = rstd; stats[2 * C + c] = sc; stats[3 * C + c] = fmaf(-mean, sc, beta[c]);
    if (running_mean) {
        running_mean[c] = fmaf(momentum, mean - running_mean[c], running_mean[c]);
        running_var[c] = fmaf(momentum, var * unbias - running_var[c], running_var[c]);
    }
}

// forward: out = act(scale*y + shift (+res)), optionally 2x2 max-pooled.  8 channels per thread.
__global__ void __launch_bounds__(256) k_bn_act_fwd(const __nv_bfloat16* __restrict__ y, const float* __restrict__ stats, const __nv_bfloat16* __restrict__ res,
                                                    __nv_bfloat16* __restrict__ out, ActGeom g, int64_t total) {
    const int c8n = g.C >> 3;
    const float* scale = stats + 2 * g.C;
    const float* shift = stats + 3 * g.C;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c8 = (int)(t % c8n);
        const int64_t p = t / c8n;
        float sc[8], sh[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            sc[i] = g.has_affine ? scale[c8 * 8 + i] : 1.f;
            sh[i] = g.has_affine ? shift[c8 * 8 + i] : 0.f;
        }
        float o[8];
        if (!g.pool) {
            float f[8], r[8];
            unpack8(*reinterpret_cast<const uint4*>(y + p * g.C + c8 * 8), f);
            if (res) unpack8(*reinterpret_cast<const uint4*>(res + p * g.C + c8 * 8), r);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float z = pre_act(f[i], sc[i], sh[i], res ? r[i] : 0.f);
                o[i] = g.relu ? fmaxf(z, 0.f) : z;
            }
        } else {
            const int WP = g.W >> 1, HP = g.H >> 1;
            const int wp = (int)(p % WP), hp = (int)((p / WP) % HP), n = (int)(p / ((int64_t)WP * HP));
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = -INFINITY;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t row = ((int64_t)n * g.H + hp * 2 + (q >> 1)) * g.W + wp * 2 + (q & 1);
                float f[8];
                unpack8(*reinterpret_cast<const uint4*>(y + row * g.C + c8 * 8), f);
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], pre_act(f[i], sc[i], sh[i], 0.f));
            }
            if (g.relu) {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaxf(o[i], 0.f);
            }
        }
        *reinterpret_cast<uint4*>(out + p * g.C + c8 * 8) = pack8(o);
    }
}

// scalar-channel fallback (C % 8 != 0: LeNet's 20 / 50 channels)
__global__ void __launch_bounds__(256) k_bn_act_fwd_scalar(const __nv_bfloat16* __restrict__ y, const float* __restrict__ stats, const __nv_bfloat16* __restrict__ res,
                                                           __nv_bfloat16* __restrict__ out, ActGeom g, int64_t total) {
    const float* scale = stats + 2 * g.C;
    const float* shift = stats + 3 * g.C;
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c = (int)(t % g.C);
        const int64_t p = t / g.C;
        const float sc = g.has_affine ? scale[c] : 1.f, sh = g.has_affine ? shift[c] : 0.f;
        float o;
        if (!g.pool) {
            o = pre_act(__bfloat162float(y[p * g.C + c]), sc, sh, res ? __bfloat162float(res[p * g.C + c]) : 0.f);
        } else {
            const int WP = g.W >> 1, HP = g.H >> 1;
            const int wp = (int)(p % WP), hp = (int)((p / WP) % HP), n = (int)(p / ((int64_t)WP * HP));
            o = -INFINITY;
            for (int q = 0; q < 4; ++q) {
                const int64_t row = ((int64_t)n * g.H + hp * 2 + (q >> 1)) * g.W + wp * 2 + (q & 1);
                o = fmaxf(o, pre_act(__bfloat162float(y[row * g.C + c]), sc, sh, 0.f));
            }
        }
        if (g.relu) o = fmaxf(o, 0.f);
        out[t] = __float2bfloat16_rn(o);
    }
}

// gradient w.r.t. the pre-activation z of un-pooled element (row, c), recomputed from y:
//   no pool: dz = dout * [z > 0]        pool: dz = dout(window) * [this element is the window's first max] * [z > 0]
__device__ __forceinline__ float dz_elem(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                         const ActGeom& g, int64_t row, int c, float sc, float sh, float yv) {
    if (!g.pool) {
        const float d = __bfloat162float(dout[row * g.C + c]);
        if (!g.relu) return d;
        const float z = pre_act(yv, sc, sh, res ? __bfloat162float(res[row * g.C + c]) : 0.f);
        return z > 0.f ? d : 0.f;
    }
    const int w = (int)(row % g.W), h = (int)((row / g.W) % g.H), n = (int)(row / ((int64_t)g.W * g.H));
    const int hp = h >> 1, wp = w >> 1, me = ((h & 1) << 1) | (w & 1);
    float best = -INFINITY;
    int arg = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int64_t r2 = ((int64_t)n * g.H + hp * 2 + (q >> 1)) * g.W + wp * 2 + (q & 1);
        const float z = pre_act(q == me ? yv : __bfloat162float(y[r2 * g.C + c]), sc, sh, 0.f);
        if (z > best) { best = z; arg = q; }
    }
    if (arg != me || (g.relu && !(best > 0.f))) return 0.f;
    return __bfloat162float(dout[(((int64_t)n * (g.H >> 1) + hp) * (g.W >> 1) + wp) * g.C + c]);
}

// backward pass 1: per channel sum(dz) and sum(dz * xhat)   (thread = channel, CTA covers rows_par rows)
__global__ void __launch_bounds__(256) k_bn_bwd_reduce(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                                       const float* __restrict__ stats, float* __restrict__ sums, ActGeom g, int rows_per_cta) {
    const int lanes = g.C < 256 ? g.C : 256;
    const int rows_par = 256 / lanes;
    const int c = blockIdx.x * lanes + (threadIdx.x % lanes);
    const int rsub = threadIdx.x / lanes;
    if (c >= g.C || rsub >= rows_par) return;
    const float mean = g.has_affine ? stats[c] : 0.f, rstd = g.has_affine ? stats[g.C + c] : 1.f;
    const float sc = g.has_affine ? stats[2 * g.C + c] : 1.f, sh = g.has_affine ? stats[3 * g.C + c] : 0.f;
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < g.M ? r0 + rows_per_cta : g.M;
    float s = 0.f, sx = 0.f;
    for (int64_t r = r0 + rsub; r < r1; r += rows_par) {
        const float yv = __bfloat162float(y[r * g.C + c]);
        const float dz = dz_elem(y, res, dout, g, r, c, sc, sh, yv);
        s += dz;
        sx = fmaf(dz, (yv - mean) * rstd, sx);
    }
    atomicAdd(&sums[c], s);
    atomicAdd(&sums[g.C + c], sx);
}

// backward pass 2: dy = scale * (dz - mean(dz) - xhat * mean(dz*xhat)); dres = dz; also emits dgamma/dbeta
__global__ void __launch_bounds__(256) k_bn_bwd_apply(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                                      const float* __restrict__ stats, const float* __restrict__ sums, __nv_bfloat16* __restrict__ dy,
                                                      __nv_bfloat16* __restrict__ dres, ActGeom g, float inv_m, int64_t total) {
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c = (int)(t % g.C);
        const int64_t r = t / g.C;
        const float mean = g.has_affine ? stats[c] : 0.f, rstd = g.has_affine ? stats[g.C + c] : 1.f;
        const float sc = g.has_affine ? stats[2 * g.C + c] : 1.f, sh = g.has_affine ? stats[3 * g.C + c] : 0.f;
        const float yv = __bfloat162float(y[t]);
        const float dz = dz_elem(y, res, dout, g, r, c, sc, sh, yv);
        float d = dz;
        if (g.has_affine) d = sc * (dz - sums[c] * inv_m - (yv - mean) * rstd * sums[g.C + c] * inv_m);
        dy[t] = __float2bfloat16_rn(d);
        if (dres) dres[t] = __float2bfloat16_rn(dz);
    }
}

// ---- vectorised backward (C % 8 == 0): a work unit is one row (no pool) or one 2x2 window (pool) x 8 channels ----
// dz for the unit, recomputed from the saved pre-BN output y: every y value is read exactly once per pass.
template <bool POOL>
__device__ __forceinline__ void unit_dz(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                        const ActGeom& g, int64_t unit, int c8, const float (&sc)[8], const float (&sh)[8],
                                        float (&yv)[POOL ? 4 : 1][8], float (&dz)[POOL ? 4 : 1][8], int64_t (&rows)[POOL ? 4 : 1]) {
    if (!POOL) {
        rows[0] = unit;
        float d[8], r[8];
        unpack8(*reinterpret_cast<const uint4*>(y + unit * g.C + c8 * 8), yv[0]);
        unpack8(*reinterpret_cast<const uint4*>(dout + unit * g.C + c8 * 8), d);
        if (res) unpack8(*reinterpret_cast<const uint4*>(res + unit * g.C + c8 * 8), r);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float z = pre_act(yv[0][i], sc[i], sh[i], res ? r[i] : 0.f);
            dz[0][i] = (!g.relu || z > 0.f) ? d[i] : 0.f;
        }
    } else {
        const int WP = g.W >> 1, HP = g.H >> 1;
        const int wp = (int)(unit % WP), hp = (int)((unit / WP) % HP), n = (int)(unit / ((int64_t)WP * HP));
        float d[8], best[8];
        int arg[8];
        unpack8(*reinterpret_cast<const uint4*>(dout + unit * g.C + c8 * 8), d);
#pragma unroll
        for (int i = 0; i < 8; ++i) { best[i] = -INFINITY; arg[i] = 0; }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            rows[q] = ((int64_t)n * g.H + hp * 2 + (q >> 1)) * g.W + wp * 2 + (q & 1);
            unpack8(*reinterpret_cast<const uint4*>(y + rows[q] * g.C + c8 * 8), yv[q]);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float z = pre_act(yv[q][i], sc[i], sh[i], 0.f);
                if (z > best[i]) { best[i] = z; arg[i] = q; }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int i = 0; i < 8; ++i) dz[q][i] = (arg[i] == q && (!g.relu || best[i] > 0.f)) ? d[i] : 0.f;
    }
}

template <bool POOL>
__global__ void __launch_bounds__(256) k_bn_bwd_reduce_v8(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                                          const float* __restrict__ stats, float* __restrict__ sums, ActGeom g, int64_t units, int units_per_cta) {
    __shared__ float red[256 * 16];
    constexpr int R = POOL ? 4 : 1;
    const int c8n = g.C >> 3;
    const int lanes = c8n < 256 ? c8n : 256;
    const int par = 256 / lanes;
    const int lane = threadIdx.x % lanes, usub = threadIdx.x / lanes;
    const int c8 = blockIdx.x * lanes + lane;
    for (int i = threadIdx.x; i < lanes * 16; i += 256) red[i] = 0.f;
    __syncthreads();
    if (c8 < c8n && usub < par) {
        float mean[8], rstd[8], sc[8], sh[8], s[8], sx[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int c = c8 * 8 + i;
            mean[i] = stats[c]; rstd[i] = stats[g.C + c]; sc[i] = stats[2 * g.C + c]; sh[i] = stats[3 * g.C + c];
            s[i] = 0.f; sx[i] = 0.f;
        }
        const int64_t u0 = (int64_t)blockIdx.y * units_per_cta;
        const int64_t u1 = u0 + units_per_cta < units ? u0 + units_per_cta : units;
        for (int64_t u = u0 + usub; u < u1; u += par) {
            float yv[R][8], dz[R][8];
            int64_t rows[R];
            unit_dz<POOL>(y, res, dout, g, u, c8, sc, sh, yv, dz, rows);
#pragma unroll
            for (int q = 0; q < R; ++q)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    s[i] += dz[q][i];
                    sx[i] = fmaf(dz[q][i], (yv[q][i] - mean[i]) * rstd[i], sx[i]);
                }
        }
        if (lanes * par != 256) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { atomicAdd(&red[lane * 16 + i], s[i]); atomicAdd(&red[lane * 16 + 8 + i], sx[i]); }
        } else {
            fold_partials<8>(s, red, lanes, lane, 0);
            fold_partials<8>(sx, red, lanes, lane, 8);
        }
    } else if (lanes * par == 256) {
        // (unreachable when the CTA is full: every thread has c8 < c8n or the whole warp is out of range)
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        fold_partials<8>(z, red, lanes, lane, 0);
        fold_partials<8>(z, red, lanes, lane, 8);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < lanes * 16; i += 256) {
        const int l = i >> 4, j = i & 15, c = (blockIdx.x * lanes + l) * 8 + (j & 7);
        if (c < g.C) atomicAdd(&sums[(j < 8 ? 0 : g.C) + c], red[i]);
    }
}

template <bool POOL>
__global__ void __launch_bounds__(256) k_bn_bwd_apply_v8(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res, const __nv_bfloat16* __restrict__ dout,
                                                         const float* __restrict__ stats, const float* __restrict__ sums, __nv_bfloat16* __restrict__ dy,
                                                         __nv_bfloat16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, ActGeom g,
                                                         float inv_m, int64_t units, int64_t total) {
    constexpr int R = POOL ? 4 : 1;
    const int c8n = g.C >> 3;
    if (blockIdx.x == 0 && dgamma && dbeta && g.has_affine)            // parameter gradients into the flat fp32 buffer
        for (int c = threadIdx.x; c < g.C; c += 256) { dbeta[c] += sums[c]; dgamma[c] += sums[g.C + c]; }
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int c8 = (int)(t % c8n);
        const int64_t u = t / c8n;
        float mean[8], rstd[8], sc[8], sh[8], m1[8], m2[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int c = c8 * 8 + i;
            mean[i] = g.has_affine ? stats[c] : 0.f; rstd[i] = g.has_affine ? stats[g.C + c] : 1.f;
            sc[i] = g.has_affine ? stats[2 * g.C + c] : 1.f; sh[i] = g.has_affine ? stats[3 * g.C + c] : 0.f;
            m1[i] = g.has_affine ? sums[c] * inv_m : 0.f; m2[i] = g.has_affine ? sums[g.C + c] * inv_m : 0.f;
        }
        float yv[R][8], dz[R][8];
        int64_t rows[R];
        unit_dz<POOL>(y, res, dout, g, u, c8, sc, sh, yv, dz, rows);
#pragma unroll
        for (int q = 0; q < R; ++q) {
            float o[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                o[i] = g.has_affine ? sc[i] * (dz[q][i] - m1[i] - (yv[q][i] - mean[i]) * rstd[i] * m2[i]) : dz[q][i];
            *reinterpret_cast<uint4*>(dy + rows[q] * g.C + c8 * 8) = pack8(o);
            if (dres) *reinterpret_cast<uint4*>(dres + rows[q] * g.C + c8 * 8) = pack8(dz[q]);
        }
    }
}

// dgamma += sum(dz*xhat), dbeta += sum(dz)  (into the flat fp32 gradient buffer)
__global__ void k_bn_param_grads(const float* __restrict__ sums, float* __restrict__ dgamma, float* __restrict__ dbeta, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    dbeta[c] += sums[c];
    dgamma[c] += sums[C + c];
}

__global__ void k_add_vec(const float* __restrict__ src, float* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}

// ---------------------------------------------------------------------------------------------
// small elementwise / head kernels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_cast_f32_bf16(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) y[i] = __float2bfloat16_rn(x[i]);
}

// copy a [rows, cols] bf16 matrix into a wider pitch (zero padded): conv weights with K % 8 != 0
__global__ void __launch_bounds__(256) k_pad_rows(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols, int pitch) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < (int64_t)rows * pitch; i += (int64_t)gridDim.x * 256) {
        const int c = (int)(i % pitch);
        const int64_t r = i / pitch;
        dst[i] = c < cols ? src[r * cols + c] : __float2bfloat16(0.f);
    }
}

// y = relu(x) in place variant is done in the GEMM epilogue; backward mask: dx = dy * [y > 0]
__global__ void __launch_bounds__(256) k_relu_bwd(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ dx, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        dx[i] = __bfloat162float(y[i]) > 0.f ? dy[i] : __float2bfloat16(0.f);
}
__global__ void __launch_bounds__(256) k_relu_fwd(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        y[i] = __bfloat162float(x[i]) > 0.f ? x[i] : __float2bfloat16(0.f);
}

// inverted dropout with a counter-based mask (recomputed in backward, nothing stored)
__global__ void __launch_bounds__(256) k_dropout(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n, float p, float scale,
                                                 uint32_t seed, uint32_t offset, const uint32_t* __restrict__ step_ptr) {
    if (step_ptr) offset += step_ptr[1] * 0x9E3779B1u;      // device step counter: fresh mask on every graph replay
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        Philox4 r = philox4x32_10((uint32_t)(i >> 2), offset, 0x9E37u, 0u, seed, 0xD409u);
        const uint32_t li = (uint32_t)i & 3u;
        const uint32_t rx = li == 0 ? r.v[0] : (li == 1 ? r.v[1] : (li == 2 ? r.v[2] : r.v[3]));
        const float u = philox_to_uniform(rx);
        y[i] = u >= p ? __float2bfloat16_rn(__bfloat162float(x[i]) * scale) : __float2bfloat16(0.f);
    }
}

// [N, HW, C] -> [N, C] mean over HW (F.avg_pool2d(out, 4) on a 4x4 map, resnet.py:93) and its backward
__global__ void __launch_bounds__(256) k_avgpool_fwd(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW, int C) {
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < (int64_t)N * C; t += (int64_t)gridDim.x * 256) {
        const int c = (int)(t % C), n = (int)(t / C);
        float s = 0.f;
        for (int i = 0; i < HW; ++i) s += __bfloat162float(x[((int64_t)n * HW + i) * C + c]);
        y[t] = __float2bfloat16_rn(s / HW);
    }
}
__global__ void __launch_bounds__(256) k_avgpool_bwd(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int N, int HW, int C) {
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < (int64_t)N * HW * C; t += (int64_t)gridDim.x * 256) {
        const int c = (int)(t % C), n = (int)(t / ((int64_t)HW * C));
        dx[t] = __float2bfloat16_rn(__bfloat162float(dy[(int64_t)n * C + c]) / HW);
    }
}

// softmax cross-entropy, mean over the batch: one warp per row.  Writes loss (atomic mean), dlogits = (p - onehot)/B
// and the top-1 hit count (accuracy, distributed_worker.py:27-39).
__global__ void __launch_bounds__(256) k_xent(const __nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ target, float* __restrict__ loss_acc /* [2]: loss, correct */,
                                              __nv_bfloat16* __restrict__ dlogits, int B, int C, int ld) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= B) return;
    const __nv_bfloat16* x = logits + (int64_t)row * ld;
    float mx = -INFINITY;
    int amax = 0;
    for (int c = lane; c < C; c += 32) {
        float v = __bfloat162float(x[c]);
        if (v > mx) { mx = v; amax = c; }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        float o = __shfl_xor_sync(0xffffffffu, mx, d);
        int oa = __shfl_xor_sync(0xffffffffu, amax, d);
        if (o > mx || (o == mx && oa < amax)) { mx = o; amax = oa; }
    }
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(__bfloat162float(x[c]) - mx);
    se = warp_sum(se);
    const int tgt = (int)target[row];
    const float lse = __logf(se) + mx;
    const float invB = 1.f / B;
    for (int c = lane; c < C; c += 32) {
        const float p = __expf(__bfloat162float(x[c]) - lse);
        dlogits[(int64_t)row * ld + c] = __float2bfloat16_rn((p - (c == tgt ? 1.f : 0.f)) * invB);
    }
    if (lane == 0) {
        atomicAdd(&loss_acc[0], (lse - __bfloat162float(x[tgt])) * invB);
        if (amax == tgt) atomicAdd(&loss_acc[1], 1.f);
    }
}

// ---------------------------------------------------------------------------------------------
// Streaming forms of the four BatchNorm passes (un-pooled, C % 8 == 0, C/8 a power of two <= 256).
//
// Round 1's kernels re-derived the channel group of every 16-byte vector (`t % c8n`) and re-loaded up to 48 per-channel
// floats for it on every iteration, one vector in flight per thread: 16-25 % of the HBM roofline on ResNet-50, where they
// were 55 % of the step (profiles/launch_list_resnet50_r2.txt).  Here the activation is walked as a flat array of 16-byte
// vectors with a grid-wide stride that is a multiple of C/8, so a thread keeps ONE channel group for its whole life: the
// per-channel constants are loaded once into registers and the loop body is four independent vector loads (eight or
// twelve with dout / residual), the arithmetic, and the stores.  Accesses stay perfectly coalesced (consecutive threads =
// consecutive vectors).  Statistics use a per-channel shift K (row 0 of the tensor): sums of (y - K) and (y - K)^2, so the
// variance S2/M - (S1/M)^2 does not cancel when |mean| >> std.
// ---------------------------------------------------------------------------------------------
#define BN_UNROLL 4

__device__ __forceinline__ void load_ch8(const float* __restrict__ p, int c8, float (&o)[8]) {
    const float4 a = reinterpret_cast<const float4*>(p)[c8 * 2], b = reinterpret_cast<const float4*>(p)[c8 * 2 + 1];
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

// CTA-level fold of per-thread partials (threads with equal tid % c8n hold the same channels) + one global atomic per
// channel and quantity per CTA.  red: c8n * 16 floats of shared memory, zeroed by the caller before use.
__device__ __forceinline__ void fold16_to_global(float (&s)[8], float (&q)[8], float* red, int c8n, int C, float* __restrict__ sums) {
    const int lane = threadIdx.x & (c8n - 1);
    fold_partials<8>(s, red, c8n < 256 ? c8n : 256, lane, 0);
    fold_partials<8>(q, red, c8n < 256 ? c8n : 256, lane, 8);
    __syncthreads();
    const int c8base = (blockIdx.x * 256) & (c8n - 1);      // = 0 when c8n <= 256 (it is): kept for clarity
    for (int i = threadIdx.x; i < c8n * 16; i += 256) {
        const int l = i >> 4, j = i & 15, c = ((c8base + l) & (c8n - 1)) * 8 + (j & 7);
        atomicAdd(&sums[(j < 8 ? 0 : C) + c], red[i]);
    }
}

// sums[c] += sum(y - K), sums[C + c] += sum((y - K)^2) with K = shift_row[c] (or 0 when shift_row == NULL)
__global__ void __launch_bounds__(256) k_colstats_stream(const __nv_bfloat16* __restrict__ y, float* __restrict__ sums, int64_t V, int c8n, int C,
                                                         const __nv_bfloat16* __restrict__ shift_row) {
    __shared__ float red[256 * 16];
    for (int i = threadIdx.x; i < c8n * 16; i += 256) red[i] = 0.f;
    __syncthreads();
    const int64_t T = (int64_t)gridDim.x * 256;
    int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(v & (c8n - 1));
    float K[8], s[8], q[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { K[i] = 0.f; s[i] = 0.f; q[i] = 0.f; }
    if (shift_row) unpack8(*reinterpret_cast<const uint4*>(shift_row + c8 * 8), K);
    const uint4* yv = reinterpret_cast<const uint4*>(y);
    for (; v + (BN_UNROLL - 1) * T < V; v += BN_UNROLL * T) {
        uint4 a[BN_UNROLL];
#pragma unroll
        for (int u = 0; u < BN_UNROLL; ++u) a[u] = yv[v + u * T];
#pragma unroll
        for (int u = 0; u < BN_UNROLL; ++u) {
            float f[8];
            unpack8(a[u], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) { const float d = f[i] - K[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
        }
    }
    for (; v < V; v += T) {
        float f[8];
        unpack8(yv[v], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float d = f[i] - K[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
    }
    fold16_to_global(s, q, red, c8n, C, sums);
}

// as k_bn_finalize, for sums taken relative to K = shift_row (bf16 [C], may be NULL = no shift)
__global__ void k_bn_finalize_shifted(const float* __restrict__ sums, const __nv_bfloat16* __restrict__ shift_row, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
                                      float* __restrict__ stats, int C, float inv_m, float unbias, float momentum, float eps,
                                      long long* __restrict__ num_batches) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && num_batches) *num_batches += 1;
    if (c >= C) return;
    const float K = shift_row ? __bfloat162float(shift_row[c]) : 0.f;
    const float m1 = sums[c] * inv_m;
    const float mean = K + m1;
    float var = fmaf(-m1, m1, sums[C + c] * inv_m);
    var = var > 0.f ? var : 0.f;
    const float rstd = rsqrtf(var + eps);
    const float sc = gamma[c] * rstd;
    stats[c] = mean; stats[C + c] = rstd; stats[2 * C + c] = sc; stats[3 * C + c] = fmaf(-mean, sc, beta[c]);
    if (running_mean) {
        running_mean[c] = fmaf(momentum, mean - running_mean[c], running_mean[c]);
        running_var[c] = fmaf(momentum, var * unbias - running_var[c], running_var[c]);
    }
}

// out = act(scale*y + shift (+res))
template <bool RES>
__global__ void __launch_bounds__(256) k_bn_act_fwd_stream(const __nv_bfloat16* __restrict__ y, const float* __restrict__ stats,
                                                           const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ out, int64_t V, int c8n,
                                                           int C, int relu, int has_affine) {
    const int64_t T = (int64_t)gridDim.x * 256;
    int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(v & (c8n - 1));
    float sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { sc[i] = 1.f; sh[i] = 0.f; }
    if (has_affine) { load_ch8(stats + 2 * C, c8, sc); load_ch8(stats + 3 * C, c8, sh); }
    const uint4* yv = reinterpret_cast<const uint4*>(y);
    const uint4* rv = reinterpret_cast<const uint4*>(res);
    uint4* ov = reinterpret_cast<uint4*>(out);
    const float lo = relu ? 0.f : -INFINITY;
    for (; v + (BN_UNROLL - 1) * T < V; v += BN_UNROLL * T) {
        uint4 a[BN_UNROLL], r[BN_UNROLL];
#pragma unroll
        for (int u = 0; u < BN_UNROLL; ++u) { a[u] = yv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
        for (int u = 0; u < BN_UNROLL; ++u) {
            float f[8], g[8], o[8];
            unpack8(a[u], f);
            if (RES) unpack8(r[u], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = fmaxf(pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f), lo);
            ov[v + u * T] = pack8(o);
        }
    }
    for (; v < V; v += T) {
        float f[8], g[8], o[8];
        unpack8(yv[v], f);
        if (RES) unpack8(rv[v], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaxf(pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f), lo);
        ov[v] = pack8(o);
    }
}

// sums[c] += sum(dz), sums[C + c] += sum(dz * xhat), dz = dout masked by the recomputed ReLU
template <bool RES>
__global__ void __launch_bounds__(256) k_bn_bwd_reduce_stream(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res,
                                                              const __nv_bfloat16* __restrict__ dout, const float* __restrict__ stats,
                                                              float* __restrict__ sums, int64_t V, int c8n, int C, int relu) {
    __shared__ float red[256 * 16];
    for (int i = threadIdx.x; i < c8n * 16; i += 256) red[i] = 0.f;
    __syncthreads();
    const int64_t T = (int64_t)gridDim.x * 256;
    int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(v & (c8n - 1));
    float rstd[8], nm[8], sc[8], sh[8], s[8], sx[8];
    load_ch8(stats, c8, nm); load_ch8(stats + C, c8, rstd); load_ch8(stats + 2 * C, c8, sc); load_ch8(stats + 3 * C, c8, sh);
#pragma unroll
    for (int i = 0; i < 8; ++i) { nm[i] = -nm[i] * rstd[i]; s[i] = 0.f; sx[i] = 0.f; }      // xhat = y * rstd + nm
    const uint4* yv = reinterpret_cast<const uint4*>(y);
    const uint4* rv = reinterpret_cast<const uint4*>(res);
    const uint4* dv = reinterpret_cast<const uint4*>(dout);
    const float thr = relu ? 0.f : -INFINITY;
    constexpr int U = RES ? 2 : BN_UNROLL;
    for (; v + (U - 1) * T < V; v += U * T) {
        uint4 a[U], d[U], r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { a[u] = yv[v + u * T]; d[u] = dv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float f[8], g[8], e[8];
            unpack8(a[u], f); unpack8(d[u], e);
            if (RES) unpack8(r[u], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
                const float dz = z > thr ? e[i] : 0.f;
                s[i] += dz;
                sx[i] = fmaf(dz, fmaf(f[i], rstd[i], nm[i]), sx[i]);
            }
        }
    }
    for (; v < V; v += T) {
        float f[8], g[8], e[8];
        unpack8(yv[v], f); unpack8(dv[v], e);
        if (RES) unpack8(rv[v], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
            const float dz = z > thr ? e[i] : 0.f;
            s[i] += dz;
            sx[i] = fmaf(dz, fmaf(f[i], rstd[i], nm[i]), sx[i]);
        }
    }
    fold16_to_global(s, sx, red, c8n, C, sums);
}

// dy = sc * (dz - m1 - xhat * m2) = sc*dz + B*y + Cc ; dres = dz
template <bool RES, bool DRES>
__global__ void __launch_bounds__(256) k_bn_bwd_apply_stream(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res,
                                                             const __nv_bfloat16* __restrict__ dout, const float* __restrict__ stats,
                                                             const float* __restrict__ sums, __nv_bfloat16* __restrict__ dy,
                                                             __nv_bfloat16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                             int64_t V, int c8n, int C, int relu, int has_affine, float inv_m) {
    if (blockIdx.x == 0 && dgamma && dbeta && has_affine)              // parameter gradients into the flat fp32 buffer
        for (int c = threadIdx.x; c < C; c += 256) { dbeta[c] += sums[c]; dgamma[c] += sums[C + c]; }
    const int64_t T = (int64_t)gridDim.x * 256;
    int64_t v = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(v & (c8n - 1));
    float sc[8], sh[8], B[8], Cc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { sc[i] = 1.f; sh[i] = 0.f; B[i] = 0.f; Cc[i] = 0.f; }
    if (has_affine) {
        float mean[8], rstd[8], m1[8], m2[8];
        load_ch8(stats, c8, mean); load_ch8(stats + C, c8, rstd); load_ch8(stats + 2 * C, c8, sc); load_ch8(stats + 3 * C, c8, sh);
        load_ch8(sums, c8, m1); load_ch8(sums + C, c8, m2);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float k2 = sc[i] * m2[i] * inv_m;                   // sc * mean(dz * xhat)
            B[i] = -k2 * rstd[i];
            Cc[i] = -sc[i] * m1[i] * inv_m + k2 * mean[i] * rstd[i];
        }
    }
    const uint4* yv = reinterpret_cast<const uint4*>(y);
    const uint4* rv = reinterpret_cast<const uint4*>(res);
    const uint4* dv = reinterpret_cast<const uint4*>(dout);
    uint4* dyv = reinterpret_cast<uint4*>(dy);
    uint4* drv = reinterpret_cast<uint4*>(dres);
    const float thr = relu ? 0.f : -INFINITY;
    constexpr int U = RES ? 2 : BN_UNROLL;
    for (; v + (U - 1) * T < V; v += U * T) {
        uint4 a[U], d[U], r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { a[u] = yv[v + u * T]; d[u] = dv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float f[8], g[8], e[8], o[8];
            unpack8(a[u], f); unpack8(d[u], e);
            if (RES) unpack8(r[u], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
                e[i] = z > thr ? e[i] : 0.f;
                o[i] = fmaf(sc[i], e[i], fmaf(B[i], f[i], Cc[i]));
            }
            dyv[v + u * T] = pack8(o);
            if (DRES) drv[v + u * T] = pack8(e);
        }
    }
    for (; v < V; v += T) {
        float f[8], g[8], e[8], o[8];
        unpack8(yv[v], f); unpack8(dv[v], e);
        if (RES) unpack8(rv[v], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
            e[i] = z > thr ? e[i] : 0.f;
            o[i] = fmaf(sc[i], e[i], fmaf(B[i], f[i], Cc[i]));
        }
        dyv[v] = pack8(o);
        if (DRES) drv[v] = pack8(e);
    }
}

// ---------------------------------------------------------------------------------------------
// Fused BatchNorm passes: statistics + normalise in ONE cooperative kernel, backward reduce + apply in ONE.
//
// A training-mode BatchNorm needs the per-channel sums of the whole tensor before it can touch a single element, which
// is why round 1 had separate kernels (colstats -> finalize -> act; reduce -> apply).  With every CTA co-resident
// (cooperative launch, grid = SMs x occupancy) the dependency becomes a grid-wide barrier inside one kernel: phase 1
// streams the tensor from HBM and reduces, the barrier publishes the sums, phase 2 streams the same tensor again -- now
// largely out of the 126 MB L2 -- and writes the result.  Per layer that removes two launches forward and one backward
// (ResNet-50: 159 launches per step) and up to a third of the HBM traffic; the tiny finalize kernel disappears into
// phase 2 (every thread derives scale/shift for its own 8 channels, CTA 0 stores the saved statistics and updates the
// running ones).  The kernels keep the fixed-channel-group-per-thread structure of the streaming kernels above.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// All CTAs of the (cooperatively launched) grid arrive; `bar` is a zero-initialised word used once per launch.
__device__ __forceinline__ void grid_barrier_once(uint32_t* bar) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        while (ld_acquire_gpu_u32(bar) < gridDim.x) __nanosleep(20);
    }
    __syncthreads();
}
__device__ __forceinline__ void ldcg_ch8(const float* p, int c8, float (&o)[8]) {      // L2 loads: written by other CTAs in phase 1
    const float4 a = __ldcg(reinterpret_cast<const float4*>(p) + c8 * 2), b = __ldcg(reinterpret_cast<const float4*>(p) + c8 * 2 + 1);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}

struct BnFwdArgs {
    const __nv_bfloat16* y; const __nv_bfloat16* res; __nv_bfloat16* out;
    float* sums;            // [2C] zero on entry (have_sums = 0) or finished un-shifted sums (have_sums = 1)
    uint32_t* bar;          // zero on entry
    const float* gamma; const float* beta;
    float* running_mean; float* running_var; long long* num_batches;
    float* stats;           // out: [mean | rstd | scale | shift]
    int64_t V; int c8n, C, relu, have_sums;
    float inv_m, unbias, momentum, eps;
};

#define BN_U1 8      // vectors in flight per thread, one input stream

template <bool RES>
__global__ void __launch_bounds__(256) k_bn_fwd_coop(const BnFwdArgs a) {
    __shared__ float red[256 * 16];
    const int64_t T = (int64_t)gridDim.x * 256;
    const int64_t g0 = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(g0 & (a.c8n - 1));
    const uint4* yv = reinterpret_cast<const uint4*>(a.y);
    float K[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) K[i] = 0.f;
    if (!a.have_sums) {
        // ---- phase 1: sums of (y - K), (y - K)^2 with K = row 0 ---------------------------------------
        for (int i = threadIdx.x; i < a.c8n * 16; i += 256) red[i] = 0.f;
        __syncthreads();
        unpack8(yv[c8], K);
        float s[8], q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
        int64_t v = g0;
        for (; v + (BN_U1 - 1) * T < a.V; v += BN_U1 * T) {
            uint4 x[BN_U1];
#pragma unroll
            for (int u = 0; u < BN_U1; ++u) x[u] = yv[v + u * T];
#pragma unroll
            for (int u = 0; u < BN_U1; ++u) {
                float f[8];
                unpack8(x[u], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { const float d = f[i] - K[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
            }
        }
        for (; v < a.V; v += T) {
            float f[8];
            unpack8(yv[v], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) { const float d = f[i] - K[i]; s[i] += d; q[i] = fmaf(d, d, q[i]); }
        }
        fold16_to_global(s, q, red, a.c8n, a.C, a.sums);
        grid_barrier_once(a.bar);
    }
    // ---- phase 2: scale / shift for this thread's channels, then normalise ------------------------------
    float sc[8], sh[8];
    {
        float s1[8], s2[8], ga[8], be[8];
        ldcg_ch8(a.sums, c8, s1); ldcg_ch8(a.sums + a.C, c8, s2);
        load_ch8(a.gamma, c8, ga); load_ch8(a.beta, c8, be);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float m1 = s1[i] * a.inv_m;
            const float mean = K[i] + m1;
            float var = fmaf(-m1, m1, s2[i] * a.inv_m);
            var = var > 0.f ? var : 0.f;
            sc[i] = ga[i] * rsqrtf(var + a.eps);
            sh[i] = fmaf(-mean, sc[i], be[i]);
        }
    }
    if (blockIdx.x == 0) {          // saved statistics for the backward pass + running statistics (nn.BatchNorm2d semantics)
        if (threadIdx.x == 0 && a.num_batches) *a.num_batches += 1;
        for (int c = threadIdx.x; c < a.C; c += 256) {
            const float Kc = a.have_sums ? 0.f : __bfloat162float(a.y[c]);
            const float m1 = __ldcg(a.sums + c) * a.inv_m;
            const float mean = Kc + m1;
            float var = fmaf(-m1, m1, __ldcg(a.sums + a.C + c) * a.inv_m);
            var = var > 0.f ? var : 0.f;
            const float rstd = rsqrtf(var + a.eps);
            const float s = a.gamma[c] * rstd;
            a.stats[c] = mean; a.stats[a.C + c] = rstd; a.stats[2 * a.C + c] = s; a.stats[3 * a.C + c] = fmaf(-mean, s, a.beta[c]);
            if (a.running_mean) {
                a.running_mean[c] = fmaf(a.momentum, mean - a.running_mean[c], a.running_mean[c]);
                a.running_var[c] = fmaf(a.momentum, var * a.unbias - a.running_var[c], a.running_var[c]);
            }
        }
    }
    const uint4* rv = reinterpret_cast<const uint4*>(a.res);
    uint4* ov = reinterpret_cast<uint4*>(a.out);
    const float lo = a.relu ? 0.f : -INFINITY;
    constexpr int U = RES ? 4 : BN_U1;
    int64_t v = g0;
    for (; v + (U - 1) * T < a.V; v += U * T) {
        uint4 x[U], r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { x[u] = yv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float f[8], g[8], o[8];
            unpack8(x[u], f);
            if (RES) unpack8(r[u], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = fmaxf(pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f), lo);
            ov[v + u * T] = pack8(o);
        }
    }
    for (; v < a.V; v += T) {
        float f[8], g[8], o[8];
        unpack8(yv[v], f);
        if (RES) unpack8(rv[v], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaxf(pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f), lo);
        ov[v] = pack8(o);
    }
}

struct BnBwdArgs {
    const __nv_bfloat16* y; const __nv_bfloat16* res; const __nv_bfloat16* dout;
    const float* stats;
    float* sums;            // [2C] zero on entry
    uint32_t* bar;          // zero on entry
    __nv_bfloat16* dy; __nv_bfloat16* dres;
    float* dgamma; float* dbeta;
    int64_t V; int c8n, C, relu;
    float inv_m;
};

template <bool RES, bool DRES>
__global__ void __launch_bounds__(256) k_bn_bwd_coop(const BnBwdArgs a) {
    __shared__ float red[256 * 16];
    const int64_t T = (int64_t)gridDim.x * 256;
    const int64_t g0 = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int c8 = (int)(g0 & (a.c8n - 1));
    const uint4* yv = reinterpret_cast<const uint4*>(a.y);
    const uint4* rv = reinterpret_cast<const uint4*>(a.res);
    const uint4* dv = reinterpret_cast<const uint4*>(a.dout);
    const float thr = a.relu ? 0.f : -INFINITY;
    float rstd[8], nm[8], sc[8], sh[8];
    load_ch8(a.stats, c8, nm); load_ch8(a.stats + a.C, c8, rstd); load_ch8(a.stats + 2 * a.C, c8, sc); load_ch8(a.stats + 3 * a.C, c8, sh);
#pragma unroll
    for (int i = 0; i < 8; ++i) nm[i] = -nm[i] * rstd[i];               // xhat = y * rstd + nm
    constexpr int U = RES ? 2 : 4;
    // ---- phase 1: sum(dz), sum(dz * xhat) ---------------------------------------------------------------
    for (int i = threadIdx.x; i < a.c8n * 16; i += 256) red[i] = 0.f;
    __syncthreads();
    {
        float s[8], sx[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i] = 0.f; sx[i] = 0.f; }
        int64_t v = g0;
        for (; v + (U - 1) * T < a.V; v += U * T) {
            uint4 x[U], d[U], r[U];
#pragma unroll
            for (int u = 0; u < U; ++u) { x[u] = yv[v + u * T]; d[u] = dv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float f[8], g[8], e[8];
                unpack8(x[u], f); unpack8(d[u], e);
                if (RES) unpack8(r[u], g);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
                    const float dz = z > thr ? e[i] : 0.f;
                    s[i] += dz;
                    sx[i] = fmaf(dz, fmaf(f[i], rstd[i], nm[i]), sx[i]);
                }
            }
        }
        for (; v < a.V; v += T) {
            float f[8], g[8], e[8];
            unpack8(yv[v], f); unpack8(dv[v], e);
            if (RES) unpack8(rv[v], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
                const float dz = z > thr ? e[i] : 0.f;
                s[i] += dz;
                sx[i] = fmaf(dz, fmaf(f[i], rstd[i], nm[i]), sx[i]);
            }
        }
        fold16_to_global(s, sx, red, a.c8n, a.C, a.sums);
    }
    grid_barrier_once(a.bar);
    // ---- phase 2: dy = sc*dz + B*y + Cc --------------------------------------------------------------------
    if (blockIdx.x == 0 && a.dgamma && a.dbeta)
        for (int c = threadIdx.x; c < a.C; c += 256) { a.dbeta[c] += __ldcg(a.sums + c); a.dgamma[c] += __ldcg(a.sums + a.C + c); }
    float B[8], Cc[8];
    {
        float m1[8], m2[8];
        ldcg_ch8(a.sums, c8, m1); ldcg_ch8(a.sums + a.C, c8, m2);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float k2 = sc[i] * m2[i] * a.inv_m;
            B[i] = -k2 * rstd[i];
            Cc[i] = -sc[i] * m1[i] * a.inv_m - k2 * nm[i];              // nm = -mean * rstd
        }
    }
    uint4* dyv = reinterpret_cast<uint4*>(a.dy);
    uint4* drv = reinterpret_cast<uint4*>(a.dres);
    int64_t v = g0;
    for (; v + (U - 1) * T < a.V; v += U * T) {
        uint4 x[U], d[U], r[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { x[u] = yv[v + u * T]; d[u] = dv[v + u * T]; if (RES) r[u] = rv[v + u * T]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float f[8], g[8], e[8], o[8];
            unpack8(x[u], f); unpack8(d[u], e);
            if (RES) unpack8(r[u], g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
                e[i] = z > thr ? e[i] : 0.f;
                o[i] = fmaf(sc[i], e[i], fmaf(B[i], f[i], Cc[i]));
            }
            dyv[v + u * T] = pack8(o);
            if (DRES) drv[v + u * T] = pack8(e);
        }
    }
    for (; v < a.V; v += T) {
        float f[8], g[8], e[8], o[8];
        unpack8(yv[v], f); unpack8(dv[v], e);
        if (RES) unpack8(rv[v], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float z = pre_act(f[i], sc[i], sh[i], RES ? g[i] : 0.f);
            e[i] = z > thr ? e[i] : 0.f;
            o[i] = fmaf(sc[i], e[i], fmaf(B[i], f[i], Cc[i]));
        }
        dyv[v] = pack8(o);
        if (DRES) drv[v] = pack8(e);
    }
}

int sm_count_cached() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

// grid of a cooperative streaming kernel: every CTA resident at once (SMs x occupancy), fewer when the tensor is small
template <typename K>
int coop_grid(K kernel, int64_t V, int per_thread) {
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 256, 0) != cudaSuccess || occ < 1) { cudaGetLastError(); occ = 1; }
    const int64_t cap = (int64_t)sm_count_cached() * occ;
    int64_t g = (V + (int64_t)256 * per_thread - 1) / ((int64_t)256 * per_thread);
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// launch geometry of the streaming kernels: whole CTAs of 256 threads, >= `min_per_thread` vectors per thread when the
// tensor is big enough, at most `cap_per_sm` CTAs per SM.  c8n <= 256 and a power of two, so grid * 256 % c8n == 0.
bool stream_ok(int C) {
    const int c8n = C >> 3;
    return C % 8 == 0 && c8n >= 1 && c8n <= 256 && (c8n & (c8n - 1)) == 0;
}
int stream_grid(int64_t V, int min_per_thread, int cap_per_sm);

int grid_for(int64_t work_items, int per_cta = 256) {
    int64_t g = (work_items + per_cta - 1) / per_cta;
    const int64_t cap = (int64_t)ewdml_sm_count() * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

int stream_grid(int64_t V, int min_per_thread, int cap_per_sm) {
    int dev = 0, sms = 0;
    static int cached_sms = 0;
    if (!cached_sms) {
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cached_sms = sms > 0 ? sms : 148;
    }
    int64_t g = (V + (int64_t)256 * min_per_thread - 1) / ((int64_t)256 * min_per_thread);
    const int64_t cap = (int64_t)cached_sms * cap_per_sm;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

ConvGeom make_geom(int N, int H, int W, int C, int KH, int KW, int stride, int pad) {
    ConvGeom g;
    g.N = N; g.H = H; g.W = W; g.C = C; g.KH = KH; g.KW = KW; g.stride = stride; g.pad = pad;
    g.HO = (H + 2 * pad - KH) / stride + 1;
    g.WO = (W + 2 * pad - KW) / stride + 1;
    g.K = KH * KW * C;
    g.Kp = (g.K + 7) / 8 * 8;
    return g;
}

#define NN_LAUNCH_CHECK()                                                                                 \
    do {                                                                                                  \
        cudaError_t _e = cudaGetLastError();                                                              \
        if (_e != cudaSuccess) { ewdml_set_error("%s:%d: launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); return (int)_e; } \
        ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);                                                 \
    } while (0)

}  // namespace

// ---- C ABI -----------------------------------------------------------------------------------------
extern "C" int ewdml_im2col(const void* x, void* col, int N, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(x && col && N > 0 && C > 0);
    ConvGeom g = make_geom(N, H, W, C, KH, KW, stride, pad);
    const int64_t M = (int64_t)N * g.HO * g.WO;
    cudaStream_t st = (cudaStream_t)stream;
    if (C % 8 == 0) {
        const int64_t total = M * KH * KW * (C / 8);
        k_im2col_vec8<<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, g, total);
    } else {
        const int64_t total = M * (g.Kp / 8);
        k_im2col_scalar<<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, g, total);
    }
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_col2im(const void* dcol, void* dx, int N, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(dcol && dx && N > 0 && C > 0);
    ConvGeom g = make_geom(N, H, W, C, KH, KW, stride, pad);
    cudaStream_t st = (cudaStream_t)stream;
    if (C % 8 == 0) {
        const int64_t total = (int64_t)N * H * W * (C / 8);
        k_col2im<8><<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)dcol, (__nv_bfloat16*)dx, g, total);
    } else {
        const int64_t total = (int64_t)N * H * W * C;
        k_col2im<1><<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)dcol, (__nv_bfloat16*)dx, g, total);
    }
    NN_LAUNCH_CHECK();
    return 0;
}

// sums: fp32 [2*C], must be zero on entry.  want_sq: also accumulate sum of squares in sums[C..2C).
extern "C" int ewdml_colstats(const void* y, float* sums, int64_t M, int C, int want_sq, void* stream) {
    EWDML_CHECK_ARG(y && sums && M > 0 && C > 0 && C % 2 == 0);
    if (stream_ok(C)) {
        // want_sq == 2: sums relative to the per-channel shift K = y[0][:] (finalize with ewdml_bn_finalize3(..., shift_row = y))
        const int c8n = C / 8;
        const int64_t V = M * c8n;
        k_colstats_stream<<<stream_grid(V, 8 * (c8n > 32 ? c8n / 32 : 1), 4), 256, 0, (cudaStream_t)stream>>>(
            (const __nv_bfloat16*)y, sums, V, c8n, C, want_sq == 2 ? (const __nv_bfloat16*)y : nullptr);
        NN_LAUNCH_CHECK();
        return 0;
    }
    if (want_sq == 2) { ewdml_set_error("ewdml_colstats: shifted sums need C/8 to be a power of two <= 256 (C = %d)", C); return EWDML_ERR_ARG; }
    if (C % 8 == 0) {
        const int c8n = C / 8, lanes = c8n < 256 ? c8n : 256, rows_par = 256 / lanes;
        const int gx = (c8n + lanes - 1) / lanes;
        int gy = (int)((4 * ewdml_sm_count() + gx - 1) / gx);
        int64_t rows_per = (M + gy - 1) / gy;
        if (rows_per < 8 * rows_par) rows_per = 8 * rows_par;
        gy = (int)((M + rows_per - 1) / rows_per);
        k_colstats_v8<<<dim3(gx, gy), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, sums, M, C, want_sq, (int)rows_per);
        NN_LAUNCH_CHECK();
        return 0;
    }
    const int cp = C / 2, lanes = cp < 256 ? cp : 256;
    const int gx = (cp + lanes - 1) / lanes;
    int gy = (int)((2 * ewdml_sm_count() + gx - 1) / gx);
    int64_t rows_per = (M + gy - 1) / gy;
    const int rows_par = 256 / lanes;
    if (rows_per < 4 * rows_par) rows_per = 4 * rows_par;
    gy = (int)((M + rows_per - 1) / rows_per);
    k_colstats<<<dim3(gx, gy), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, sums, M, C, want_sq, (int)rows_per);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_bn_finalize2(const float* sums, const float* gamma, const float* beta, float* running_mean, float* running_var,
                                  float* stats, int C, int64_t M, float momentum, float eps, int64_t* num_batches, void* stream);
extern "C" int ewdml_bn_finalize(const float* sums, const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 float* stats, int C, int64_t M, float momentum, float eps, void* stream) {
    return ewdml_bn_finalize2(sums, gamma, beta, running_mean, running_var, stats, C, M, momentum, eps, nullptr, stream);
}
// as ewdml_bn_finalize2 for sums gathered with ewdml_colstats(..., want_sq = 2): shift_row = the tensor whose row 0 was the shift
extern "C" int ewdml_bn_finalize3(const float* sums, const void* shift_row, const float* gamma, const float* beta, float* running_mean,
                                  float* running_var, float* stats, int C, int64_t M, float momentum, float eps, int64_t* num_batches,
                                  void* stream) {
    EWDML_CHECK_ARG(sums && gamma && beta && stats && C > 0 && M > 0);
    const float unbias = M > 1 ? (float)((double)M / (double)(M - 1)) : 1.f;
    k_bn_finalize_shifted<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(sums, (const __nv_bfloat16*)shift_row, gamma, beta, running_mean,
                                                                             running_var, stats, C, (float)(1.0 / (double)M), unbias, momentum,
                                                                             eps, (long long*)num_batches);
    NN_LAUNCH_CHECK();
    return 0;
}
// as above, and increments *num_batches (nn.BatchNorm2d.num_batches_tracked) in the same launch
extern "C" int ewdml_bn_finalize2(const float* sums, const float* gamma, const float* beta, float* running_mean, float* running_var,
                                  float* stats, int C, int64_t M, float momentum, float eps, int64_t* num_batches, void* stream) {
    EWDML_CHECK_ARG(sums && gamma && beta && stats && C > 0 && M > 0);
    const float unbias = M > 1 ? (float)((double)M / (double)(M - 1)) : 1.f;
    k_bn_finalize<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(sums, gamma, beta, running_mean, running_var, stats, C,
                                                                     (float)(1.0 / (double)M), unbias, momentum, eps, (long long*)num_batches);
    NN_LAUNCH_CHECK();
    return 0;
}

static ActGeom make_act(int64_t M, int C, int pool, int N, int H, int W, int relu, int has_affine) {
    ActGeom g;
    g.M = M; g.C = C; g.pool = pool; g.N = N; g.H = H; g.W = W; g.relu = relu; g.has_affine = has_affine;
    return g;
}

// out = act(scale*y+shift (+res)), optionally 2x2 max-pooled; stats = [mean|rstd|scale|shift] or NULL (identity)
extern "C" int ewdml_bn_act_fwd(const void* y, const float* stats, const void* res, void* out, int N, int H, int W, int C, int relu, int pool,
                                void* stream) {
    EWDML_CHECK_ARG(y && out && (!pool || (H % 2 == 0 && W % 2 == 0 && !res)));
    const int64_t M = (int64_t)N * H * W;
    ActGeom g = make_act(M, C, pool, N, H, W, relu, stats != nullptr);
    const int64_t rows_out = pool ? M / 4 : M;
    cudaStream_t st = (cudaStream_t)stream;
    if (!pool && stream_ok(C)) {
        const int c8n = C / 8;
        const int64_t V = M * c8n;
        const int grid = stream_grid(V, 4, 8);
        if (res)
            k_bn_act_fwd_stream<true><<<grid, 256, 0, st>>>((const __nv_bfloat16*)y, stats, (const __nv_bfloat16*)res, (__nv_bfloat16*)out, V, c8n, C,
                                                            relu, stats != nullptr);
        else
            k_bn_act_fwd_stream<false><<<grid, 256, 0, st>>>((const __nv_bfloat16*)y, stats, nullptr, (__nv_bfloat16*)out, V, c8n, C, relu,
                                                             stats != nullptr);
        NN_LAUNCH_CHECK();
        return 0;
    }
    if (C % 8 == 0) {
        const int64_t total = rows_out * (C / 8);
        k_bn_act_fwd<<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)y, stats, (const __nv_bfloat16*)res, (__nv_bfloat16*)out, g, total);
    } else {
        const int64_t total = rows_out * C;
        k_bn_act_fwd_scalar<<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)y, stats, (const __nv_bfloat16*)res, (__nv_bfloat16*)out, g, total);
    }
    NN_LAUNCH_CHECK();
    return 0;
}

// Training-mode BatchNorm forward in one cooperative launch (statistics -> barrier -> normalise [+res] [+ReLU]); un-pooled
// tensors with C/8 a power of two <= 256.  sums: fp32 [2*C + 32]: 2*C zeroed accumulators + one zeroed barrier word at
// sums[2*C] (have_sums = 1: sums[0..2C) already hold finished UN-shifted sums, e.g. from ewdml_f32_to_bf16_stats; no
// statistics phase, no barrier).  Writes stats = [mean | rstd | scale | shift] and updates the running statistics.
// Returns EWDML_ERR_STATE (nothing launched) when the shape is not covered: call colstats + finalize + bn_act_fwd instead.
extern "C" int ewdml_bn_fwd_fused(const void* y, float* sums, int have_sums, const float* gamma, const float* beta, float* running_mean,
                                  float* running_var, int64_t* num_batches, float* stats, const void* res, void* out, int N, int H,
                                  int W, int C, int relu, float momentum, float eps, void* stream) {
    EWDML_CHECK_ARG(y && sums && gamma && beta && stats && out && N > 0 && C > 0);
    if (!stream_ok(C)) return EWDML_ERR_STATE;
    const int64_t M = (int64_t)N * H * W;
    BnFwdArgs a;
    a.y = (const __nv_bfloat16*)y; a.res = (const __nv_bfloat16*)res; a.out = (__nv_bfloat16*)out;
    a.sums = sums; a.bar = reinterpret_cast<uint32_t*>(sums + 2 * C); a.gamma = gamma; a.beta = beta;
    a.running_mean = running_mean; a.running_var = running_var; a.num_batches = (long long*)num_batches; a.stats = stats;
    a.c8n = C / 8; a.V = M * a.c8n; a.C = C; a.relu = relu; a.have_sums = have_sums;
    a.inv_m = (float)(1.0 / (double)M); a.unbias = M > 1 ? (float)((double)M / (double)(M - 1)) : 1.f;
    a.momentum = momentum; a.eps = eps;
    void* args[] = {&a};
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e;
    if (res) {
        const int grid = coop_grid(k_bn_fwd_coop<true>, a.V, 4);
        e = have_sums ? (k_bn_fwd_coop<true><<<grid, 256, 0, st>>>(a), cudaGetLastError())
                      : cudaLaunchCooperativeKernel((void*)k_bn_fwd_coop<true>, dim3(grid), dim3(256), args, 0, st);
    } else {
        const int grid = coop_grid(k_bn_fwd_coop<false>, a.V, 8);
        e = have_sums ? (k_bn_fwd_coop<false><<<grid, 256, 0, st>>>(a), cudaGetLastError())
                      : cudaLaunchCooperativeKernel((void*)k_bn_fwd_coop<false>, dim3(grid), dim3(256), args, 0, st);
    }
    if (e != cudaSuccess) { ewdml_set_error("ewdml_bn_fwd_fused: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);
    return 0;
}

// Backward of a BatchNorm block in one cooperative launch (reduce -> barrier -> apply).  Same contract as ewdml_bn_act_bwd
// except that sums holds 2*C + 32 floats (all zero on entry; sums[2*C] is the barrier word) and stats must be given.
// Returns EWDML_ERR_STATE (nothing launched) for shapes it does not cover (pooled blocks, C/8 not a power of two <= 256).
extern "C" int ewdml_bn_bwd_fused(const void* y, const float* stats, const void* res, const void* dout, float* sums, void* dy, void* dres,
                                  float* dgamma, float* dbeta, int N, int H, int W, int C, int relu, void* stream) {
    EWDML_CHECK_ARG(y && dout && sums && dy && stats);
    if (!stream_ok(C)) return EWDML_ERR_STATE;
    const int64_t M = (int64_t)N * H * W;
    BnBwdArgs a;
    a.y = (const __nv_bfloat16*)y; a.res = (const __nv_bfloat16*)res; a.dout = (const __nv_bfloat16*)dout; a.stats = stats;
    a.sums = sums; a.bar = reinterpret_cast<uint32_t*>(sums + 2 * C); a.dy = (__nv_bfloat16*)dy; a.dres = (__nv_bfloat16*)dres;
    a.dgamma = dgamma; a.dbeta = dbeta; a.c8n = C / 8; a.V = M * a.c8n; a.C = C; a.relu = relu; a.inv_m = (float)(1.0 / (double)M);
    void* args[] = {&a};
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e;
#define BWD_COOP(RES_, DRES_, PT_)                                                                                        \
    e = cudaLaunchCooperativeKernel((void*)k_bn_bwd_coop<RES_, DRES_>, dim3(coop_grid(k_bn_bwd_coop<RES_, DRES_>, a.V, PT_)), \
                                    dim3(256), args, 0, st)
    if (res && dres) BWD_COOP(true, true, 2);
    else if (res) BWD_COOP(true, false, 2);
    else if (dres) BWD_COOP(false, true, 4);
    else BWD_COOP(false, false, 4);
#undef BWD_COOP
    if (e != cudaSuccess) { ewdml_set_error("ewdml_bn_bwd_fused: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);
    return 0;
}

// Backward of the fused block.  sums: fp32 [2*C] zero on entry (scratch).  dgamma/dbeta (fp32, accumulated) may be
// NULL when stats == NULL.  dres (may be NULL) receives the gradient of the residual input.
extern "C" int ewdml_bn_act_bwd(const void* y, const float* stats, const void* res, const void* dout, float* sums, void* dy, void* dres,
                                float* dgamma, float* dbeta, int N, int H, int W, int C, int relu, int pool, void* stream) {
    EWDML_CHECK_ARG(y && dout && sums && dy);
    const int64_t M = (int64_t)N * H * W;
    ActGeom g = make_act(M, C, pool, N, H, W, relu, stats != nullptr);
    cudaStream_t st = (cudaStream_t)stream;
    if (!pool && stream_ok(C)) {
        const int c8n = C / 8;
        const int64_t V = M * c8n;
        const __nv_bfloat16 *yb = (const __nv_bfloat16*)y, *rb = (const __nv_bfloat16*)res, *db = (const __nv_bfloat16*)dout;
        if (stats) {
            const int grid = stream_grid(V, 8 * (c8n > 32 ? c8n / 32 : 1), 4);
            if (res) k_bn_bwd_reduce_stream<true><<<grid, 256, 0, st>>>(yb, rb, db, stats, sums, V, c8n, C, relu);
            else k_bn_bwd_reduce_stream<false><<<grid, 256, 0, st>>>(yb, nullptr, db, stats, sums, V, c8n, C, relu);
            NN_LAUNCH_CHECK();
        }
        const int grid = stream_grid(V, 4, 8);
        const float inv_m = (float)(1.0 / (double)M);
        const int aff = stats != nullptr;
#define BN_APPLY(RES_, DRES_)                                                                                                        \
        k_bn_bwd_apply_stream<RES_, DRES_><<<grid, 256, 0, st>>>(yb, rb, db, stats, sums, (__nv_bfloat16*)dy, (__nv_bfloat16*)dres, dgamma, \
                                                                 dbeta, V, c8n, C, relu, aff, inv_m)
        if (res && dres) BN_APPLY(true, true);
        else if (res) BN_APPLY(true, false);
        else if (dres) BN_APPLY(false, true);
        else BN_APPLY(false, false);
#undef BN_APPLY
        NN_LAUNCH_CHECK();
        return 0;
    }
    if (C % 8 == 0) {
        const int64_t units = pool ? M / 4 : M;
        if (stats) {
            const int c8n = C / 8, lanes = c8n < 256 ? c8n : 256, par = 256 / lanes;
            const int gx = (c8n + lanes - 1) / lanes;
            int gy = (int)((4 * ewdml_sm_count() + gx - 1) / gx);
            int64_t per = (units + gy - 1) / gy;
            if (per < 4 * par) per = 4 * par;
            gy = (int)((units + per - 1) / per);
            if (pool)
                k_bn_bwd_reduce_v8<true><<<dim3(gx, gy), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout,
                                                                       stats, sums, g, units, (int)per);
            else
                k_bn_bwd_reduce_v8<false><<<dim3(gx, gy), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout,
                                                                        stats, sums, g, units, (int)per);
            NN_LAUNCH_CHECK();
        }
        const int64_t total = units * (C / 8);
        if (pool)
            k_bn_bwd_apply_v8<true><<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout, stats, sums,
                                                                     (__nv_bfloat16*)dy, (__nv_bfloat16*)dres, dgamma, dbeta, g, (float)(1.0 / (double)M), units, total);
        else
            k_bn_bwd_apply_v8<false><<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout, stats, sums,
                                                                      (__nv_bfloat16*)dy, (__nv_bfloat16*)dres, dgamma, dbeta, g, (float)(1.0 / (double)M), units, total);
        NN_LAUNCH_CHECK();
        return 0;
    }
    if (stats) {
        const int lanes = C < 256 ? C : 256;
        const int gx = (C + lanes - 1) / lanes;
        int gy = (int)((2 * ewdml_sm_count() + gx - 1) / gx);
        int64_t rows_per = (M + gy - 1) / gy;
        const int rows_par = 256 / lanes;
        if (rows_per < 4 * rows_par) rows_per = 4 * rows_par;
        gy = (int)((M + rows_per - 1) / rows_per);
        k_bn_bwd_reduce<<<dim3(gx, gy), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout, stats, sums, g,
                                                       (int)rows_per);
        NN_LAUNCH_CHECK();
        if (dgamma && dbeta) {
            k_bn_param_grads<<<(C + 127) / 128, 128, 0, st>>>(sums, dgamma, dbeta, C);
            NN_LAUNCH_CHECK();
        }
    }
    const int64_t total = M * C;
    k_bn_bwd_apply<<<grid_for(total), 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)res, (const __nv_bfloat16*)dout, stats, sums,
                                                    (__nv_bfloat16*)dy, (__nv_bfloat16*)dres, g, (float)(1.0 / (double)M), total);
    NN_LAUNCH_CHECK();
    return 0;
}

// dst (bf16 [M][C]) = src (fp32 [M][C]) + bias; if sums != NULL also accumulates per-channel sum / sum of squares of the
// rounded output into sums[2C] (zero on entry).  C % 8 == 0.
extern "C" int ewdml_f32_to_bf16_stats(const float* src, const float* bias, void* dst, float* sums, int64_t M, int C, void* stream) {
    EWDML_CHECK_ARG(src && dst && M > 0 && C > 0 && C % 8 == 0);
    const int c8n = C / 8, lanes = c8n < 256 ? c8n : 256, rows_par = 256 / lanes;
    const int gx = (c8n + lanes - 1) / lanes;
    int gy = (int)((4 * ewdml_sm_count() + gx - 1) / gx);
    int64_t rows_per = (M + gy - 1) / gy;
    if (rows_per < 4 * rows_par) rows_per = 4 * rows_par;
    gy = (int)((M + rows_per - 1) / rows_per);
    k_f32_to_bf16_stats<<<dim3(gx, gy), 256, 0, (cudaStream_t)stream>>>(src, bias, (__nv_bfloat16*)dst, sums, M, C, (int)rows_per);
    NN_LAUNCH_CHECK();
    return 0;
}

// First-layer convolution, direct form.  x: NHWC, fp32 (x_is_f32 = 1) or bf16; w: bf16 [C_out][KH*KW*C_in]; y: bf16 [N*HO*WO][C_out].
// C_in <= 4, KH*KW*C_in <= 36, C_out a multiple of 8 and <= 256.
extern "C" int ewdml_conv_small_fwd(const void* x, int x_is_f32, const void* w, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                                    int KH, int KW, int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(x && w && y && Cin >= 1 && Cin <= 4 && KH * KW * Cin <= 36 && Cout % 8 == 0 && Cout <= 256);
    ConvGeom g = make_geom(N, H, W, Cin, KH, KW, stride, pad);
    const int64_t total = (int64_t)N * g.HO * g.WO * (Cout / 8);
    const size_t smem = sizeof(float) * g.K * Cout;
    cudaStream_t st = (cudaStream_t)stream;
    if (x_is_f32) k_conv_small_fwd<float><<<grid_for(total), 256, smem, st>>>((const float*)x, (const __nv_bfloat16*)w, bias, (__nv_bfloat16*)y, g, Cout, total);
    else k_conv_small_fwd<__nv_bfloat16><<<grid_for(total), 256, smem, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, bias, (__nv_bfloat16*)y, g, Cout, total);
    NN_LAUNCH_CHECK();
    return 0;
}
// dw (fp32 [C_out][KH*KW*C_in], accumulated) += dy^T patches(x).  C_out in {64, 128}.
extern "C" int ewdml_conv_small_wgrad(const void* x, int x_is_f32, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                      int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(x && dy && dw && Cin >= 1 && Cin <= 4 && KH * KW * Cin <= 36 && (Cout == 64 || Cout == 128));
    ConvGeom g = make_geom(N, H, W, Cin, KH, KW, stride, pad);
    const int64_t M = (int64_t)N * g.HO * g.WO;
    EWDML_CHECK_ARG((g.K + 256 / Cout - 1) / (256 / Cout) <= 20);
    const int64_t nb = (M + 63) / 64;
    const int grid = (int)(nb < 2 * ewdml_sm_count() ? nb : 2 * ewdml_sm_count());
    cudaStream_t st = (cudaStream_t)stream;
    if (x_is_f32) k_conv_small_wgrad<float><<<grid, 256, 0, st>>>((const float*)x, (const __nv_bfloat16*)dy, dw, g, Cout, M);
    else k_conv_small_wgrad<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, dw, g, Cout, M);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_pad_nhwc(const void* x, void* out, int N, int H, int W, int C, void* stream) {
    EWDML_CHECK_ARG(x && out && C % 8 == 0);
    const int64_t total = (int64_t)N * (H + 2) * (W + 2) * (C / 8);
    k_pad_nhwc<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, N, H, W, C, total);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_add_vec_f32(const float* src, float* dst, int n, void* stream) {
    EWDML_CHECK_ARG(src && dst && n > 0);
    k_add_vec<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(src, dst, n);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_cast_f32_bf16(const float* x, void* y, int64_t n, void* stream) {
    EWDML_CHECK_ARG(x && y && n > 0);
    k_cast_f32_bf16<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(x, (__nv_bfloat16*)y, n);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_pad_rows_bf16(const void* src, void* dst, int rows, int cols, int pitch, void* stream) {
    EWDML_CHECK_ARG(src && dst && rows > 0 && cols > 0 && pitch >= cols);
    k_pad_rows<<<grid_for((int64_t)rows * pitch), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, rows, cols, pitch);
    NN_LAUNCH_CHECK();
    return 0;
}

extern "C" int ewdml_relu_fwd(const void* x, void* y, int64_t n, void* stream) {
    k_relu_fwd<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n);
    NN_LAUNCH_CHECK();
    return 0;
}
extern "C" int ewdml_relu_bwd(const void* dy, const void* y, void* dx, int64_t n, void* stream) {
    k_relu_bwd<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y, (__nv_bfloat16*)dx, n);
    NN_LAUNCH_CHECK();
    return 0;
}
extern "C" int ewdml_dropout(const void* x, void* y, int64_t n, float p, uint32_t seed, uint32_t offset, const uint32_t* step_state, void* stream) {
    EWDML_CHECK_ARG(p >= 0.f && p < 1.f);
    k_dropout<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n, p, 1.f / (1.f - p), seed, offset, step_state);
    NN_LAUNCH_CHECK();
    return 0;
}
extern "C" int ewdml_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
    k_avgpool_fwd<<<grid_for((int64_t)N * C), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, HW, C);
    NN_LAUNCH_CHECK();
    return 0;
}
extern "C" int ewdml_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
    k_avgpool_bwd<<<grid_for((int64_t)N * HW * C), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, HW, C);
    NN_LAUNCH_CHECK();
    return 0;
}
// loss_acc: fp32 [2] zero on entry -> {mean loss, #correct}; dlogits has the same pitch as logits
extern "C" int ewdml_xent(const void* logits, const int64_t* target, float* loss_acc, void* dlogits, int B, int C, int ld, void* stream) {
    EWDML_CHECK_ARG(logits && target && loss_acc && dlogits && B > 0 && C > 0);
    k_xent<<<(B + 7) / 8, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits, target, loss_acc, (__nv_bfloat16*)dlogits, B, C, ld);
    NN_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// BatchNorm batch statistics without any allocation (no counterpart in the reference).
//
// ewdml_colstats + ewdml_bn_finalize need a zeroed fp32 scratch vector of 2*C partial sums, which the
// Python layer takes from torch's caching allocator.  Some users of this C ABI cannot do that: a plain
// C/C++ host loop that links libewdml.so without torch has no caching allocator, and calling cudaMalloc
// per step from it would synchronise the whole device and can fail outright when another component of
// the process has already reserved the GPU's memory.  For them the scratch lives in the library image:
// a file-scope device array of EWDML_BN_STATIC_MAX_C * 2 floats (32 KiB) that every GPU which loads the
// library keeps until the process exits.
//
// Rules that follow from sharing ONE array:
//   * C must be <= EWDML_BN_STATIC_MAX_C (4096); larger sizes are rejected with EWDML_ERR_ARG.
//   * the array keeps the previous call's sums, so each call clears the 2*C floats it uses first
//     (stream-ordered memset on `stream`);
//   * two calls whose work overlaps on the GPU (different streams, or several host threads) race on the
//     array -- serialise them on one stream;
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress, with that
//     device current).  That lookup is part of the first call, so make the first call on a device before
//     starting any stream capture on it.
// ---------------------------------------------------------------------------------------------------
#define EWDML_BN_STATIC_MAX_C 4096
static __device__ __align__(256) float g_bn_static_sums[2 * EWDML_BN_STATIC_MAX_C];

static float* bn_static_scratch_for_current_device() {
    // In host code `g_bn_static_sums` is only a handle for the symbol, never a usable address: the real
    // device address differs per GPU and has to be queried with that GPU current.
    static float* addr[64] = {nullptr};
    static std::mutex mu;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        ewdml_set_error("ewdml_bn_stats_static: cannot determine the current device");
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_bn_static_sums);
        if (e != cudaSuccess || !p) {
            ewdml_set_error("ewdml_bn_stats_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        addr[dev] = p;
    }
    return addr[dev];
}

// Same result as ewdml_colstats(y, sums, M, C, 1) followed by ewdml_bn_finalize(sums, ...): writes
// stats = [mean | rstd | scale | shift] (4*C floats) and updates the running statistics when given.
// Inputs and outputs are the caller's; nothing is allocated or freed here.
extern "C" int ewdml_bn_stats_static(const void* y, const float* gamma, const float* beta, float* running_mean, float* running_var,
                                     float* stats, int C, int64_t M, float momentum, float eps, void* stream) {
    EWDML_CHECK_ARG(y && gamma && beta && stats && M > 0 && C > 0 && C % 2 == 0);
    if (C > EWDML_BN_STATIC_MAX_C) {
        ewdml_set_error("ewdml_bn_stats_static: C = %d exceeds the built-in scratch (max %d channels); use ewdml_colstats + "
                        "ewdml_bn_finalize with your own 2*C-float buffer", C, EWDML_BN_STATIC_MAX_C);
        return EWDML_ERR_ARG;
    }
    float* sums = bn_static_scratch_for_current_device();
    if (!sums) return EWDML_ERR_CUDA;
    cudaStream_t st = (cudaStream_t)stream;
    EWDML_CUDA(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * (size_t)C, st));       // the previous call's sums are still there
    const int cp = C / 2, lanes = cp < 256 ? cp : 256;
    const int gx = (cp + lanes - 1) / lanes;
    int gy = (int)((2 * 148 + gx - 1) / gx);
    int64_t rows_per = (M + gy - 1) / gy;
    const int rows_par = 256 / lanes;
    if (rows_per < 4 * rows_par) rows_per = 4 * rows_par;
    gy = (int)((M + rows_per - 1) / rows_per);
    k_colstats<<<dim3(gx, gy), 256, 0, st>>>((const __nv_bfloat16*)y, sums, M, C, 1, (int)rows_per);
    NN_LAUNCH_CHECK();
    const float unbias = M > 1 ? (float)((double)M / (double)(M - 1)) : 1.f;
    k_bn_finalize<<<(C + 127) / 128, 128, 0, st>>>(sums, gamma, beta, running_mean, running_var, stats, C, (float)(1.0 / (double)M), unbias,
                                                   momentum, eps, nullptr);
    NN_LAUNCH_CHECK();
    return 0;
}
