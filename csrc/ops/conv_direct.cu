// K5, small-channel part: direct convolution kernels for shapes the tcgen05 implicit GEMM does not take -- channel counts
// that are not multiples of 8 (LeNet's 1 -> 20 -> 50 channel 5x5 convolutions, src/model_ops/lenet.py:18-19), any kernel
// size, stride and padding.  These layers are far too small for tensor-core tiles (conv1: 25 MACs per output, conv2: 500),
// so each pass is a plain CUDA-core kernel over NHWC bf16 activations with fp32 accumulation:
//
//   fprop : one thread per output element, output channel fastest (the 25..500 activation loads of neighbouring threads are
//           the same addresses -> broadcast; the weight row of each thread is contiguous)
//   dgrad : one thread per input element, input channel fastest
//   wgrad : one warp per weight element, lanes stride over the N*HO*WO output positions, shuffle reduction, the warp's
//           sum is added straight into the flat fp32 gradient buffer (no atomics: one warp owns one element)
//
// Weights are [C_out][KH][KW][C_in] bf16 (channels-last storage = the K-major GEMM operand the other layers use).
// With these, a LeNet step runs no cuDNN kernel (ewdml/ops/native.py conv_block).
#include "../common.cuh"

namespace {

struct DGeom {
    int N, H, W, Cin, Cout, KH, KW, stride, pad, HO, WO;
};

__global__ void __launch_bounds__(256) k_conv_direct_fwd(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                                         const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, DGeom g, int64_t total) {
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int co = (int)(t % g.Cout);
        int64_t p = t / g.Cout;
        const int wo = (int)(p % g.WO); p /= g.WO;
        const int ho = (int)(p % g.HO);
        const int n = (int)(p / g.HO);
        float acc = bias ? bias[co] : 0.f;
        const __nv_bfloat16* wr = w + (int64_t)co * g.KH * g.KW * g.Cin;
        for (int kh = 0; kh < g.KH; ++kh) {
            const int hi = ho * g.stride - g.pad + kh;
            if (hi < 0 || hi >= g.H) continue;
            for (int kw = 0; kw < g.KW; ++kw) {
                const int wi = wo * g.stride - g.pad + kw;
                if (wi < 0 || wi >= g.W) continue;
                const __nv_bfloat16* xr = x + (((int64_t)n * g.H + hi) * g.W + wi) * g.Cin;
                const __nv_bfloat16* wk = wr + (kh * g.KW + kw) * g.Cin;
                for (int ci = 0; ci < g.Cin; ++ci) acc = fmaf(__bfloat162float(xr[ci]), __bfloat162float(wk[ci]), acc);
            }
        }
        y[t] = __float2bfloat16_rn(acc);
    }
}

__global__ void __launch_bounds__(256) k_conv_direct_dgrad(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ w,
                                                           __nv_bfloat16* __restrict__ dx, DGeom g, int64_t total) {
    for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
        const int ci = (int)(t % g.Cin);
        int64_t p = t / g.Cin;
        const int wi = (int)(p % g.W); p /= g.W;
        const int hi = (int)(p % g.H);
        const int n = (int)(p / g.H);
        float acc = 0.f;
        for (int kh = 0; kh < g.KH; ++kh) {
            const int hn = hi + g.pad - kh;
            if (hn < 0 || hn % g.stride) continue;
            const int ho = hn / g.stride;
            if (ho >= g.HO) continue;
            for (int kw = 0; kw < g.KW; ++kw) {
                const int wn = wi + g.pad - kw;
                if (wn < 0 || wn % g.stride) continue;
                const int wo = wn / g.stride;
                if (wo >= g.WO) continue;
                const __nv_bfloat16* dr = dy + (((int64_t)n * g.HO + ho) * g.WO + wo) * g.Cout;
                const __nv_bfloat16* wk = w + ((int64_t)kh * g.KW + kw) * g.Cin + ci;
                const int64_t wstride = (int64_t)g.KH * g.KW * g.Cin;
                for (int co = 0; co < g.Cout; ++co) acc = fmaf(__bfloat162float(dr[co]), __bfloat162float(wk[co * wstride]), acc);
            }
        }
        dx[t] = __float2bfloat16_rn(acc);
    }
}

__global__ void __launch_bounds__(256) k_conv_direct_wgrad(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                                           float* __restrict__ dw, DGeom g, int64_t nweights) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * 256) >> 5;
    const int64_t positions = (int64_t)g.N * g.HO * g.WO;
    for (int64_t e = warp; e < nweights; e += nwarps) {
        const int ci = (int)(e % g.Cin);
        int64_t r = e / g.Cin;
        const int kw = (int)(r % g.KW); r /= g.KW;
        const int kh = (int)(r % g.KH);
        const int co = (int)(r / g.KH);
        float acc = 0.f;
        for (int64_t p = lane; p < positions; p += 32) {
            const int wo = (int)(p % g.WO);
            const int64_t q = p / g.WO;
            const int ho = (int)(q % g.HO);
            const int n = (int)(q / g.HO);
            const int hi = ho * g.stride - g.pad + kh, wi = wo * g.stride - g.pad + kw;
            if (hi < 0 || hi >= g.H || wi < 0 || wi >= g.W) continue;
            acc = fmaf(__bfloat162float(dy[p * g.Cout + co]), __bfloat162float(x[(((int64_t)n * g.H + hi) * g.W + wi) * g.Cin + ci]), acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) dw[e] += acc;
    }
}

DGeom make(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
    DGeom g;
    g.N = N; g.H = H; g.W = W; g.Cin = Cin; g.Cout = Cout; g.KH = KH; g.KW = KW; g.stride = stride; g.pad = pad;
    g.HO = (H + 2 * pad - KH) / stride + 1;
    g.WO = (W + 2 * pad - KW) / stride + 1;
    return g;
}

int grid_of(int64_t items) {
    int64_t b = (items + 255) / 256;
    return (int)(b < 1 ? 1 : (b > (int64_t)ewdml_sm_count() * 16 ? (int64_t)ewdml_sm_count() * 16 : b));
}

#define DC_LAUNCH_CHECK()                                                                                                    \
    do {                                                                                                                     \
        cudaError_t _e = cudaGetLastError();                                                                                 \
        if (_e != cudaSuccess) { ewdml_set_error("%s:%d: launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); return (int)_e; } \
        ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);                                                                    \
    } while (0)

}  // namespace

// y[N,HO,WO,Cout] (bf16) = conv(x[N,H,W,Cin] bf16, w[Cout,KH,KW,Cin] bf16) + bias (fp32, may be NULL)
extern "C" int ewdml_conv_direct_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                                     int KH, int KW, int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(x && w && y && N > 0 && Cin > 0 && Cout > 0 && KH > 0 && KW > 0 && stride > 0 && pad >= 0);
    DGeom g = make(N, H, W, Cin, Cout, KH, KW, stride, pad);
    EWDML_CHECK_ARG(g.HO > 0 && g.WO > 0);
    const int64_t total = (int64_t)N * g.HO * g.WO * Cout;
    k_conv_direct_fwd<<<grid_of(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, bias,
                                                                        (__nv_bfloat16*)y, g, total);
    DC_LAUNCH_CHECK();
    return 0;
}

// dx[N,H,W,Cin] (bf16) = conv_transpose(dy[N,HO,WO,Cout], w)
extern "C" int ewdml_conv_direct_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                       int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(dy && w && dx && N > 0 && Cin > 0 && Cout > 0);
    DGeom g = make(N, H, W, Cin, Cout, KH, KW, stride, pad);
    const int64_t total = (int64_t)N * H * W * Cin;
    k_conv_direct_dgrad<<<grid_of(total), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)w,
                                                                          (__nv_bfloat16*)dx, g, total);
    DC_LAUNCH_CHECK();
    return 0;
}

// dw[Cout,KH,KW,Cin] (fp32, ACCUMULATED: the flat gradient buffer) += dy^T * x
extern "C" int ewdml_conv_direct_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                       int stride, int pad, void* stream) {
    EWDML_CHECK_ARG(x && dy && dw && N > 0 && Cin > 0 && Cout > 0);
    DGeom g = make(N, H, W, Cin, Cout, KH, KW, stride, pad);
    const int64_t nweights = (int64_t)Cout * KH * KW * Cin;
    k_conv_direct_wgrad<<<grid_of(nweights * 32), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, dw, g,
                                                                                  nweights);
    DC_LAUNCH_CHECK();
    return 0;
}
