// Adasum pairwise combine on the device (Horovod's op=Adasum, horvod_pytorch.py:35,173-180,200).
//
//     adasum(a, b) = (1 - <a,b> / (2 |a|^2)) a + (1 - <a,b> / (2 |b|^2)) b          per tensor (segment)
//
// Horovod reduces W gradients with a binary tree of these combines, each tensor with its own dot product and norms.
// Here `a` and `b` are flat fp32 buffers holding every tensor of the bucket (they may live in peer memory: the gathered
// window of the collective engine), the segment table says where each tensor starts, and two launches do one tree node
// for ALL tensors: k_adasum_dots accumulates {<a,b>, |a|^2, |b|^2} per segment in fp64 (block reduction, one atomic per
// block and quantity), k_adasum_combine applies the two coefficients.  A zero-norm side contributes with coefficient 1
// (the other side's projection term vanishes), which is also what the host formulation in ewdml/parallel/hvd.py does.
#include <algorithm>

#include "../common.cuh"

namespace {

__global__ void __launch_bounds__(256) k_adasum_dots(const float* __restrict__ a, const float* __restrict__ b,
                                                     const int64_t* __restrict__ seg_off, const int64_t* __restrict__ seg_n,
                                                     double* __restrict__ acc) {
    __shared__ double red[32];
    const int seg = blockIdx.y;
    const int64_t off = seg_off[seg], n = seg_n[seg];
    double dab = 0.0, daa = 0.0, dbb = 0.0;
    // segments start on 32-element boundaries (128 B): vector loads are aligned; the tail is scalar
    const int64_t n4 = n >> 2;
    const float4* a4 = reinterpret_cast<const float4*>(a + off);
    const float4* b4 = reinterpret_cast<const float4*>(b + off);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 x = a4[i], y = b4[i];
        dab += (double)(x.x * y.x + x.y * y.y) + (double)(x.z * y.z + x.w * y.w);
        daa += (double)(x.x * x.x + x.y * x.y) + (double)(x.z * x.z + x.w * x.w);
        dbb += (double)(y.x * y.x + y.y * y.y) + (double)(y.z * y.z + y.w * y.w);
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) {
            const float x = a[off + i], y = b[off + i];
            dab += (double)x * y; daa += (double)x * x; dbb += (double)y * y;
        }
    dab = block_sum(dab, red);
    daa = block_sum(daa, red);
    dbb = block_sum(dbb, red);
    if (threadIdx.x == 0) {
        atomicAdd(acc + 3 * seg + 0, dab);
        atomicAdd(acc + 3 * seg + 1, daa);
        atomicAdd(acc + 3 * seg + 2, dbb);
    }
}

__global__ void __launch_bounds__(256) k_adasum_combine(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out,
                                                        const int64_t* __restrict__ seg_off, const int64_t* __restrict__ seg_n,
                                                        const double* __restrict__ acc) {
    const int seg = blockIdx.y;
    const int64_t off = seg_off[seg], n = seg_n[seg];
    const double dab = acc[3 * seg], daa = acc[3 * seg + 1], dbb = acc[3 * seg + 2];
    const float ca = daa > 0.0 ? (float)(1.0 - dab / (2.0 * daa)) : 1.f;
    const float cb = dbb > 0.0 ? (float)(1.0 - dab / (2.0 * dbb)) : 1.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[off + i] = ca * a[off + i] + cb * b[off + i];
}

}  // namespace

// One tree node for every segment: out = adasum(a, b).  seg_off / seg_n / acc are DEVICE arrays (nseg int64, nseg int64,
// 3 * nseg doubles of scratch); `out` may alias `a` or `b`.  Enqueues on `stream`; allocates nothing.
extern "C" int ewdml_adasum_pair(const float* a, const float* b, float* out, int nseg, const int64_t* seg_off,
                                 const int64_t* seg_n, int64_t max_seg_numel, double* acc, void* stream) {
    EWDML_CHECK_ARG(a && b && out && nseg > 0 && seg_off && seg_n && acc && max_seg_numel > 0);
    cudaStream_t st = (cudaStream_t)stream;
    EWDML_CUDA(cudaMemsetAsync(acc, 0, sizeof(double) * 3 * (size_t)nseg, st));
    int bx = (int)std::min<int64_t>((max_seg_numel / 4 + 255) / 256, 64);
    if (bx < 1) bx = 1;
    dim3 grid(bx, nseg);
    k_adasum_dots<<<grid, 256, 0, st>>>(a, b, seg_off, seg_n, acc);
    EWDML_LAUNCH_CHECK();
    k_adasum_combine<<<grid, 256, 0, st>>>(a, b, out, seg_off, seg_n, acc);
    EWDML_LAUNCH_CHECK();
    return 0;
}
