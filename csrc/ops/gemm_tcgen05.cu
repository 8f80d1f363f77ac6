// K6 / core of K5: bf16 GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), fed by TMA.
//
//   C[M,N] (+)= A[M,K] * B[N,K]^T        A, B: bf16, K contiguous ("K-major");  fp32 accumulation
//
// Used for nn.Linear forward/backward (lenet.py:20-21, vgg.py:24-29, resnet.py:77) and, through
// im2col operands (conv.cu), for every convolution pass (vgg.py:53, resnet.py:19-55).  The reference
// reaches these through ATen -> cuBLAS/cuDNN library calls; here the kernel is written out:
//
//   warp 0   TMA producer: cp.async.bulk.tensor 2D loads of a 128x64 A tile and a BNx64 B tile
//            (128-byte swizzle) into a STAGES-deep shared-memory ring, completion on mbarriers
//   warp 1   allocates TMEM, then ONE elected thread issues tcgen05.mma (M=128, N=BN, K=16) four
//            times per stage with shared-memory descriptors; tcgen05.commit releases the stage and
//            finally signals the epilogue
//   warps 2-5  epilogue: tcgen05.ld 32 lanes x 32 columns per warp from TMEM -> registers ->
//            bias / convert -> global (bf16 store, fp32 store, or fp32 atomic add for split-K and
//            for weight gradients accumulated straight into the flat gradient buffer)
//
// One output tile per CTA, up to two CTAs per SM (3 stages x 32 KB + 128 TMEM columns each), so one
// CTA's epilogue overlaps the other's main loop.  grid = (M/128, N/BN, split_k).
#include <cuda.h>

#include <map>
#include <mutex>
#include <tuple>

#include "../common.cuh"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;              // 64 bf16 = 128 bytes = one swizzle row
constexpr int UMMA_K = 16;

// Implicit-GEMM convolution (3x3, stride 1, pad 1) without an im2col buffer.  Activations live in a zero-bordered
// layout [N][H+2][W+2][C]; for a tile of consecutive padded pixel rows, tap (dh, dw) of the filter reads the SAME matrix
// shifted by the constant row offset dh*(W+2)+dw, so every operand tile is still one plain 2-D TMA box:
//   fprop : y[r]  = sum_tap  X[r + off(tap)] * W_tap^T          k-block -> (tap, 64-channel block)
//   dgrad : dx[r] = sum_tap DY[r - off(tap)] * W_tap            k-block -> (tap, 64-channel block of C_out)
//   wgrad : dW_tap += DY^T * X[. + off(tap)]                    n-tile  -> tap   (reduction over padded rows)
// Rows that are border positions produce values that are simply not stored (epilogue maps padded -> compact rows).
enum { CONV_NONE = 0, CONV_FPROP = 1, CONV_DGRAD = 2, CONV_WGRAD = 3 };
struct ConvMap {
    int mode;
    int cblocks;            // 64-wide k-blocks per tap (fprop: C_in/64, dgrad: C_out/64)
    int cin;                // C_in (dgrad: column offset of a tap in W; wgrad: n-tile -> tap)
    int H, W;               // un-padded spatial size (for the padded -> compact row map)
    int remap;              // 1: epilogue stores only interior rows, at their compact row index
    int off[9];             // row offset of each tap in the padded layout
};

struct GemmArgs {
    ConvMap cm;
    void* C;
    const float* bias;
    int M, N, K, ldc;
    int m_store;                    // rows of C (== M except for the padded->compact conv epilogue)
    int out_mode;                   // 0: bf16 store, 1: fp32 store, 2: fp32 atomic add
    int kb_per_split, kb_total;
    float alpha;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra LAB_DONE;\n"
        "bra LAB_WAIT;\n"
        "LAB_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* tmap, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tcgen05_alloc(uint32_t smem_slot, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_slot), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tcgen05_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tcgen05_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte swizzle shared-memory matrix descriptor (rows of 128 B, 8-row groups 1024 B apart)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);          // start address        bits [0,14)
    d |= (uint64_t)1 << 16;                               // leading byte offset  (unused for SW128 K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                     // stride byte offset   bits [32,46)
    d |= (uint64_t)1 << 46;                               // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                               // layout: SWIZZLE_128B
    return d;
}
// MN-major, 128-byte swizzle: the tile is stored as [MN/64] slabs of [64 k-rows][64 mn] (128 B per k-row);
// 8 k-rows x 128 B form one 1024-byte swizzle atom.  Canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in
// 16-byte units: LBO = distance between 64-wide MN slabs (64 rows x 128 B), SBO = distance between atoms
// along K (1024 B).
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((BK * 128) >> 4) << 16;               // LBO: next 64-element MN slab
    d |= (uint64_t)(1024 >> 4) << 32;                     // SBO: next group of 8 k-rows
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, M=128, N=BN; bit 15/16 = A/B is MN-major
__host__ __device__ constexpr uint32_t make_idesc(int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

template <int BN, int STAGES>
struct Smem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int BIAS_OFF = BAR_OFF + 8 * (2 * STAGES + 1) + 16;
    static constexpr int TOTAL = BIAS_OFF + BN * 4 + 1024;                     // + alignment slack
};

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192) k_gemm_tcgen05(const __grid_constant__ CUtensorMap tmA,
                                                         const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
    using S = Smem<BN, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar_base = smem_base + S::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8 * s; };
    auto empty_bar = [&](int s) { return bar_base + 8 * (STAGES + s); };
    const uint32_t tmem_full_bar = bar_base + 8 * (2 * STAGES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::BAR_OFF + 8 * (2 * STAGES + 1));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tile = blockIdx.x, n_tile = blockIdx.y;
    const int kb0 = blockIdx.z * g.kb_per_split;
    const int kb1 = min(g.kb_total, kb0 + g.kb_per_split);
    const int nkb = kb1 - kb0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) tcgen05_alloc(smem_u32(tmem_slot), BN);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

    if (warp == 0) {
        if (lane == 0 && nkb > 0) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                mbar_wait(empty_bar(s), ph ^ 1);
                mbar_expect_tx(full_bar(s), S::STAGE_BYTES);
                const uint32_t a_dst = smem_base + s * S::STAGE_BYTES;
                const int kb = kb0 + i;
                const int k0 = kb * BK;
                int a_row = m_tile * BM, a_col = k0, b_row = n_tile * BN, b_col = k0;      // (row, col) = (outer, inner) of a K-major operand
                if (g.cm.mode == CONV_FPROP) {
                    const int tap = kb / g.cm.cblocks;
                    a_row += g.cm.off[tap];
                    a_col = (kb - tap * g.cm.cblocks) * BK;
                } else if (g.cm.mode == CONV_DGRAD) {
                    const int tap = kb / g.cm.cblocks;
                    a_row -= g.cm.off[tap];
                    a_col = (kb - tap * g.cm.cblocks) * BK;
                    b_col = a_col;                          // B is MN-major here: (inner = tap*C_in + n0, outer = C_out block)
                    b_row = tap * g.cm.cin + n_tile * BN;
                } else if (g.cm.mode == CONV_WGRAD) {
                    const int tap = (n_tile * BN) / g.cm.cin;
                    b_row = n_tile * BN - tap * g.cm.cin;   // channel offset inside the tap
                    b_col = k0 + g.cm.off[tap];             // shifted padded row (reduction index)
                }
                if constexpr (A_MN) {            // box = [64 k-rows][64 mn]: one slab per 64 rows of the tile
#pragma unroll
                    for (int h = 0; h < BM / 64; ++h) tma_load_2d(a_dst + h * (BK * 128), &tmA, full_bar(s), a_row + h * 64, a_col);
                } else {
                    tma_load_2d(a_dst, &tmA, full_bar(s), a_col, a_row);
                }
                if constexpr (B_MN) {
#pragma unroll
                    for (int h = 0; h < BN / 64; ++h)
                        tma_load_2d(a_dst + S::A_BYTES + h * (BK * 128), &tmB, full_bar(s), b_row + h * 64, b_col);
                } else {
                    tma_load_2d(a_dst + S::A_BYTES, &tmB, full_bar(s), b_col, b_row);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && nkb > 0) {
            constexpr uint32_t idesc = make_idesc(BN, A_MN, B_MN);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                mbar_wait(full_bar(s), ph);
                tcgen05_fence_after();
                const uint32_t a_addr = smem_base + s * S::STAGE_BYTES;
                const uint64_t a_desc = A_MN ? make_smem_desc_mn(a_addr) : make_smem_desc(a_addr);
                const uint64_t b_desc = B_MN ? make_smem_desc_mn(a_addr + S::A_BYTES) : make_smem_desc(a_addr + S::A_BYTES);
                // per UMMA_K = 16 step: K-major advances 32 B inside the swizzle row (+2 in addr>>4 units),
                // MN-major advances 16 k-rows = two 1024-byte atoms (+128)
                constexpr uint64_t a_step = A_MN ? 128 : 2, b_step = B_MN ? 128 : 2;
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) {
                    tcgen05_mma_f16(tmem_base, a_desc + a_step * k, b_desc + b_step * k, idesc, (i > 0 || k > 0) ? 1u : 0u);
                }
                tcgen05_commit(empty_bar(s));          // frees the smem stage when these MMAs retire
            }
            tcgen05_commit(tmem_full_bar);             // accumulator complete
        }
    } else {
        // ---- epilogue: warps 2..5 own TMEM lane quadrant (warp % 4) -----------------------------
        const int quad = warp & 3;
        // stage this tile's bias slice in shared memory while the main loop runs (no global loads on the critical path)
        float* s_bias = reinterpret_cast<float*>(smem + S::BIAS_OFF);
        const bool use_bias = g.bias != nullptr && blockIdx.z == 0;
        if (use_bias) {
            for (int j = threadIdx.x - 64; j < BN; j += 128) {
                const int c = n_tile * BN + j;
                s_bias[j] = c < g.N ? g.bias[c] : 0.f;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");       // epilogue warps only
        }
        int row = m_tile * BM + quad * 32 + lane;
        if (g.cm.remap) {                       // padded pixel index -> compact pixel index, border rows are dropped
            const int WP = g.cm.W + 2, HP = g.cm.H + 2;
            const int wp = row % WP, t = row / WP, hp = t % HP, n = t / HP;
            const bool interior = hp >= 1 && hp <= g.cm.H && wp >= 1 && wp <= g.cm.W;
            row = interior ? (n * g.cm.H + hp - 1) * g.cm.W + (wp - 1) : g.m_store;
        }
        if (nkb > 0) {
            mbar_wait(tmem_full_bar, 0);
            tcgen05_fence_after();
        }
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            uint32_t v[32];
            if (nkb > 0) {
                tcgen05_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(c * 32), v);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0;
            }
            const int col0 = n_tile * BN + c * 32;
            if (row >= g.m_store || col0 >= g.N) continue;
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * g.alpha;
            if (use_bias) {
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] += s_bias[c * 32 + j];
            }
            const bool full = col0 + 32 <= g.N;
            if (g.out_mode == 0) {
                __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(g.C) + (size_t)row * g.ldc + col0;
                if (full && (g.ldc % 8 == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        uint4 pk;
                        __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]), p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
                        __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]), p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
                        pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
                        pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
                        *reinterpret_cast<uint4*>(out + j) = pk;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (col0 + j < g.N) out[j] = __float2bfloat16_rn(f[j]);
                }
            } else if (g.out_mode == 1) {
                float* out = reinterpret_cast<float*>(g.C) + (size_t)row * g.ldc + col0;
                if (full && (g.ldc % 4 == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(out + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (col0 + j < g.N) out[j] = f[j];
                }
            } else {
                float* out = reinterpret_cast<float*>(g.C) + (size_t)row * g.ldc + col0;
                if (full && (g.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)       // 16-byte vector reduction: 4x fewer L2 atomic operations
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + j), "f"(f[j]), "f"(f[j + 1]), "f"(f[j + 2]),
                                     "f"(f[j + 3])
                                     : "memory");
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (col0 + j < g.N) atomicAdd(out + j, f[j]);
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) tcgen05_dealloc(tmem_base, BN);
}

// ---- EXPERIMENTAL persistent variant (EWDML_GEMM_PERSIST=1; compiled and resource-checked, not yet validated on a GPU) -------
// ncu shows the one-tile-per-CTA kernel above spending most of its time in per-tile fixed costs (TMEM allocation, barrier and
// descriptor set-up, TMEM -> register -> global epilogue) around K loops of only 9-72 k-blocks.  Here one CTA per SM walks a list
// of work items (m fastest, so neighbouring CTAs share the B tile in L2), the shared-memory ring runs across tile boundaries, and
// TMEM holds TWO accumulators so that the epilogue of item i overlaps the main loop of item i+1.  All waits are bounded (trap).
__device__ __forceinline__ uint32_t mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
    for (uint32_t it = 0; !mbar_try(bar, parity); ++it)
        if (it > (1u << 26)) __trap();
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

template <int BN, int STAGES>
struct SmemP {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int NBARS = 2 * STAGES + 4;                               // full, empty, tmem_full[2], tmem_empty[2]
    static constexpr int BIAS_OFF = BAR_OFF + 8 * (NBARS + 1) + 16;
    static constexpr int TOTAL = BIAS_OFF + 2 * BN * 4 + 1024;
};

struct WorkDims { int m_tiles, n_tiles, total; };

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192) k_gemm_persist(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                         const GemmArgs g, const WorkDims wd) {
    using S = SmemP<BN, STAGES>;
    static_assert(2 * BN <= 512, "two accumulators must fit the 512 TMEM columns");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar_base = smem_base + S::BAR_OFF;
    auto full_bar = [&](int s) { return bar_base + 8 * s; };
    auto empty_bar = [&](int s) { return bar_base + 8 * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8 * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8 * (2 * STAGES + 2 + a); };
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::BAR_OFF + 8 * S::NBARS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }      // 4 epilogue warps arrive
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) tcgen05_alloc(smem_u32(tmem_slot), 2 * BN);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

    // work item w -> (m_tile, n_tile, split z); every split is non-empty by construction
    auto decode = [&](int w, int& m_tile, int& n_tile, int& kb0, int& nkb, int& z) {
        m_tile = w % wd.m_tiles;
        const int rest = w / wd.m_tiles;
        n_tile = rest % wd.n_tiles;
        z = rest / wd.n_tiles;
        kb0 = z * g.kb_per_split;
        nkb = min(g.kb_total, kb0 + g.kb_per_split) - kb0;
    };

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;                                     // k-blocks issued so far: ring position across items
            for (int w = blockIdx.x; w < wd.total; w += gridDim.x) {
                int m_tile, n_tile, kb0, nkb, z;
                decode(w, m_tile, n_tile, kb0, nkb, z);
                for (int i = 0; i < nkb; ++i, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait_bounded(empty_bar(s), ph ^ 1);
                    mbar_expect_tx(full_bar(s), S::STAGE_BYTES);
                    const uint32_t a_dst = smem_base + s * S::STAGE_BYTES;
                    const int kb = kb0 + i;
                    const int k0 = kb * BK;
                    int a_row = m_tile * BM, a_col = k0, b_row = n_tile * BN, b_col = k0;
                    if (g.cm.mode == CONV_FPROP) {
                        const int tap = kb / g.cm.cblocks;
                        a_row += g.cm.off[tap];
                        a_col = (kb - tap * g.cm.cblocks) * BK;
                    } else if (g.cm.mode == CONV_DGRAD) {
                        const int tap = kb / g.cm.cblocks;
                        a_row -= g.cm.off[tap];
                        a_col = (kb - tap * g.cm.cblocks) * BK;
                        b_col = a_col;
                        b_row = tap * g.cm.cin + n_tile * BN;
                    } else if (g.cm.mode == CONV_WGRAD) {
                        const int tap = (n_tile * BN) / g.cm.cin;
                        b_row = n_tile * BN - tap * g.cm.cin;
                        b_col = k0 + g.cm.off[tap];
                    }
                    if constexpr (A_MN) {
#pragma unroll
                        for (int h = 0; h < BM / 64; ++h) tma_load_2d(a_dst + h * (BK * 128), &tmA, full_bar(s), a_row + h * 64, a_col);
                    } else {
                        tma_load_2d(a_dst, &tmA, full_bar(s), a_col, a_row);
                    }
                    if constexpr (B_MN) {
#pragma unroll
                        for (int h = 0; h < BN / 64; ++h)
                            tma_load_2d(a_dst + S::A_BYTES + h * (BK * 128), &tmB, full_bar(s), b_row + h * 64, b_col);
                    } else {
                        tma_load_2d(a_dst + S::A_BYTES, &tmB, full_bar(s), b_col, b_row);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BN, A_MN, B_MN);
            constexpr uint64_t a_step = A_MN ? 128 : 2, b_step = B_MN ? 128 : 2;
            uint32_t it = 0;
            int item = 0;
            for (int w = blockIdx.x; w < wd.total; w += gridDim.x, ++item) {
                int m_tile, n_tile, kb0, nkb, z;
                decode(w, m_tile, n_tile, kb0, nkb, z);
                const int acc = item & 1;
                mbar_wait_bounded(tempty_bar(acc), ((item >> 1) & 1) ^ 1);      // the epilogue has drained this accumulator
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
                for (int i = 0; i < nkb; ++i, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait_bounded(full_bar(s), ph);
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_base + s * S::STAGE_BYTES;
                    const uint64_t a_desc = A_MN ? make_smem_desc_mn(a_addr) : make_smem_desc(a_addr);
                    const uint64_t b_desc = B_MN ? make_smem_desc_mn(a_addr + S::A_BYTES) : make_smem_desc(a_addr + S::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        tcgen05_mma_f16(d_tmem, a_desc + a_step * k, b_desc + b_step * k, idesc, (i > 0 || k > 0) ? 1u : 0u);
                    tcgen05_commit(empty_bar(s));
                }
                tcgen05_commit(tfull_bar(acc));
            }
        }
    } else {
        const int quad = warp & 3;
        float* s_bias_all = reinterpret_cast<float*>(smem + S::BIAS_OFF);
        int item = 0;
        for (int w = blockIdx.x; w < wd.total; w += gridDim.x, ++item) {
            int m_tile, n_tile, kb0, nkb, z;
            decode(w, m_tile, n_tile, kb0, nkb, z);
            const int acc = item & 1;
            float* s_bias = s_bias_all + acc * BN;
            const bool use_bias = g.bias != nullptr && z == 0;
            if (use_bias) {
                for (int j = threadIdx.x - 64; j < BN; j += 128) {
                    const int c = n_tile * BN + j;
                    s_bias[j] = c < g.N ? g.bias[c] : 0.f;
                }
            }
            // every item: also orders the reuse of s_bias[acc] two items later behind the slowest epilogue warp of this one
            asm volatile("bar.sync 1, 128;" ::: "memory");
            int row = m_tile * BM + quad * 32 + lane;
            if (g.cm.remap) {
                const int WP = g.cm.W + 2, HP = g.cm.H + 2;
                const int wp = row % WP, t = row / WP, hp = t % HP, n = t / HP;
                const bool interior = hp >= 1 && hp <= g.cm.H && wp >= 1 && wp <= g.cm.W;
                row = interior ? (n * g.cm.H + hp - 1) * g.cm.W + (wp - 1) : g.m_store;
            }
            mbar_wait_bounded(tfull_bar(acc), (item >> 1) & 1);
            tcgen05_fence_after();
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tcgen05_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + c * 32), v);
                const int col0 = n_tile * BN + c * 32;
                if (row >= g.m_store || col0 >= g.N) continue;
                float f[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * g.alpha;
                if (use_bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] += s_bias[c * 32 + j];
                }
                const bool full = col0 + 32 <= g.N;
                if (g.out_mode == 0) {
                    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(g.C) + (size_t)row * g.ldc + col0;
                    if (full && (g.ldc % 8 == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            uint4 pk;
                            __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]), p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
                            __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]), p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
                            pk.x = *reinterpret_cast<uint32_t*>(&p0); pk.y = *reinterpret_cast<uint32_t*>(&p1);
                            pk.z = *reinterpret_cast<uint32_t*>(&p2); pk.w = *reinterpret_cast<uint32_t*>(&p3);
                            *reinterpret_cast<uint4*>(out + j) = pk;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (col0 + j < g.N) out[j] = __float2bfloat16_rn(f[j]);
                    }
                } else if (g.out_mode == 1) {
                    float* out = reinterpret_cast<float*>(g.C) + (size_t)row * g.ldc + col0;
                    if (full && (g.ldc % 4 == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(out + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (col0 + j < g.N) out[j] = f[j];
                    }
                } else {
                    float* out = reinterpret_cast<float*>(g.C) + (size_t)row * g.ldc + col0;
                    if (full && (g.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + j), "f"(f[j]), "f"(f[j + 1]), "f"(f[j + 2]),
                                         "f"(f[j + 3])
                                         : "memory");
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (col0 + j < g.N) atomicAdd(out + j, f[j]);
                    }
                }
            }
            // this warp's reads of the accumulator are complete (tcgen05.wait::ld inside the load): hand it back to the MMA thread
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) tcgen05_dealloc(tmem_base, 2 * BN);
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && p)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2-D bf16 tensor map over a row-major matrix [outer][inner] with row pitch ld elements.
// K-major operand: inner = K, outer = rows (M or N), box {64, box_rows}.
// MN-major operand: inner = MN, outer = K (reduction), box {64, 64}.
int make_tmap(CUtensorMap* tm, const void* base, int64_t rows, int64_t k, int64_t ld, int box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) { ewdml_set_error("cuTensorMapEncodeTiled not available"); return EWDML_ERR_CUDA; }
    cuuint64_t dims[2] = {(cuuint64_t)k, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        ewdml_set_error("cuTensorMapEncodeTiled failed (%d) base=%p rows=%lld k=%lld ld=%lld", (int)r, base, (long long)rows,
                        (long long)k, (long long)ld);
        return EWDML_ERR_CUDA;
    }
    return 0;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
int launch(const CUtensorMap& tA, const CUtensorMap& tB, const GemmArgs& g, int split, cudaStream_t st) {
    using S = Smem<BN, STAGES>;
    static bool attr = false;
    if (!attr) {
        EWDML_CUDA(cudaFuncSetAttribute(k_gemm_tcgen05<BN, STAGES, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
        attr = true;
    }
    dim3 grid((g.M + BM - 1) / BM, (g.N + BN - 1) / BN, split);
    k_gemm_tcgen05<BN, STAGES, A_MN, B_MN><<<grid, 192, S::TOTAL, st>>>(tA, tB, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { ewdml_set_error("gemm launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);
    return 0;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
int launch_persist(const CUtensorMap& tA, const CUtensorMap& tB, const GemmArgs& g, int split, cudaStream_t st) {
    using S = SmemP<BN, STAGES>;
    static bool attr = false;
    if (!attr) {
        EWDML_CUDA(cudaFuncSetAttribute(k_gemm_persist<BN, STAGES, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
        attr = true;
    }
    WorkDims wd;
    wd.m_tiles = (g.M + BM - 1) / BM; wd.n_tiles = (g.N + BN - 1) / BN; wd.total = wd.m_tiles * wd.n_tiles * split;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = wd.total < sms ? wd.total : sms;             // one CTA per SM (the ring takes most of the shared memory)
    k_gemm_persist<BN, STAGES, A_MN, B_MN><<<grid, 192, S::TOTAL, st>>>(tA, tB, g, wd);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { ewdml_set_error("persistent gemm launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    ewdml_counter_add(EWDML_CTR_COMPUTE_LAUNCHES, 1);
    return 0;
}

struct TmapDims { int64_t a_outer, a_inner, b_outer, b_inner; };   // rows / contiguous width of the matrices the maps describe

template <bool A_MN, bool B_MN>
int gemm_dispatch(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int out_mode,
                  const float* bias, float alpha, int split_k, void* stream, const ConvMap* cm = nullptr,
                  const TmapDims* td = nullptr, int bn_force = 0, int m_store = -1, int kb_total = -1) {
    EWDML_CHECK_ARG(A && B && C && M > 0 && N > 0 && K > 0);
    EWDML_CHECK_ARG(lda % 8 == 0 && ldb % 8 == 0 && ((uintptr_t)A % 16) == 0 && ((uintptr_t)B % 16) == 0);
    EWDML_CHECK_ARG(out_mode >= 0 && out_mode <= 2);
    int bn = bn_force ? bn_force : (N <= 64 ? 64 : 128);
    // EXPERIMENTAL, off unless EWDML_GEMM_BN=256: 128x256 tiles (N = 256 UMMA, 256 TMEM columns, 4 x 48 KB stages, one CTA
    // per SM) halve the A-operand traffic of layers with >= 256 output channels -- the L2 -> smem traffic is what ncu shows
    // these GEMMs waiting on (profiles/ncu_prof_gemm.txt).  Compiled and resource-checked, NOT yet validated on a GPU.
    static const int bn_env = getenv("EWDML_GEMM_BN") ? atoi(getenv("EWDML_GEMM_BN")) : 0;
    if (!bn_force && bn_env == 256 && !A_MN && N % 256 == 0) bn = 256;
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    if (cm) g.cm = *cm;
    g.C = C; g.bias = bias; g.M = M; g.N = N; g.K = K; g.ldc = ldc; g.out_mode = out_mode; g.alpha = alpha;
    g.m_store = m_store >= 0 ? m_store : M;
    g.kb_total = kb_total > 0 ? kb_total : (K + BK - 1) / BK;
    int tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
    int split = split_k;
    if (split <= 0) {
        split = 1;
        if (out_mode == 2) {                       // only the accumulate mode may be split
            int target = 2 * ewdml_sm_count();
            while (tiles * split * 2 <= target && split * 2 <= g.kb_total / 2) split *= 2;
        }
    }
    if (split > 1 && out_mode != 2) { ewdml_set_error("split-K needs out_mode 2 (atomic accumulate)"); return EWDML_ERR_ARG; }
    if (split > g.kb_total) split = g.kb_total;
    g.kb_per_split = (g.kb_total + split - 1) / split;
    split = (g.kb_total + g.kb_per_split - 1) / g.kb_per_split;     // no empty splits
    CUtensorMap tA, tB;
    // K-major: matrix is [rows][K]; MN-major: matrix is [K][rows] (rows contiguous)
    int rc;
    if (td) {
        rc = make_tmap(&tA, A, td->a_outer, td->a_inner, lda, A_MN ? 64 : BM);
        if (rc) return rc;
        rc = make_tmap(&tB, B, td->b_outer, td->b_inner, ldb, B_MN ? 64 : bn);
    } else {
        rc = A_MN ? make_tmap(&tA, A, K, M, lda, 64) : make_tmap(&tA, A, M, K, lda, BM);
        if (rc) return rc;
        rc = B_MN ? make_tmap(&tB, B, K, N, ldb, 64) : make_tmap(&tB, B, N, K, ldb, bn);
    }
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // Pipeline depth: a CTA's k-block rate is (stages in flight) / (TMA round trip).  When the grid fits one CTA per SM
    // there is no second resident CTA to hide that latency, so use the whole shared memory for a deep ring instead.
    const int ctas = tiles * split;
    const bool deep = ctas <= 160;
    if constexpr (!A_MN) {
        if (bn == 256) return launch<256, 4, A_MN, B_MN>(tA, tB, g, split, st);
    }
    // EXPERIMENTAL, off unless EWDML_GEMM_PERSIST=1 (see k_gemm_persist)
    static const int persist_env = getenv("EWDML_GEMM_PERSIST") ? atoi(getenv("EWDML_GEMM_PERSIST")) : 0;
    if (persist_env == 1)
        return bn == 64 ? launch_persist<64, 8, A_MN, B_MN>(tA, tB, g, split, st) : launch_persist<128, 6, A_MN, B_MN>(tA, tB, g, split, st);
    if (bn == 64) return deep ? launch<64, 8, A_MN, B_MN>(tA, tB, g, split, st) : launch<64, 4, A_MN, B_MN>(tA, tB, g, split, st);
    return deep ? launch<128, 6, A_MN, B_MN>(tA, tB, g, split, st) : launch<128, 3, A_MN, B_MN>(tA, tB, g, split, st);
}

}  // namespace

// C[M,N] (op)= alpha * op(A) op(B) (+ bias[N]), bf16 operands, fp32 accumulation in TMEM.
//   a_mn = 0: A is [M][K] (K contiguous);  a_mn = 1: A is [K][M] (M contiguous)
//   b_mn = 0: B is [N][K] (K contiguous);  b_mn = 1: B is [K][N] (N contiguous)
// so (0,0) = x W^T (forward), (0,1) = dy W (data gradient), (1,1) = dy^T x (weight gradient) without
// any transposed copies.  Leading dimensions in elements (multiples of 8), bases 16-byte aligned.
// out_mode 0: bf16 store, 1: fp32 store, 2: fp32 atomic accumulate (split-K).  split_k <= 0: automatic.
extern "C" int ewdml_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int a_mn, int b_mn, int out_mode, const float* bias, float alpha, int split_k, void* stream) {
    if (!a_mn && !b_mn) return gemm_dispatch<false, false>(A, B, C, M, N, K, lda, ldb, ldc, out_mode, bias, alpha, split_k, stream);
    if (!a_mn && b_mn) return gemm_dispatch<false, true>(A, B, C, M, N, K, lda, ldb, ldc, out_mode, bias, alpha, split_k, stream);
    if (a_mn && b_mn) return gemm_dispatch<true, true>(A, B, C, M, N, K, lda, ldb, ldc, out_mode, bias, alpha, split_k, stream);
    return gemm_dispatch<true, false>(A, B, C, M, N, K, lda, ldb, ldc, out_mode, bias, alpha, split_k, stream);
}

extern "C" int ewdml_gemm_bf16_tn(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                  int out_mode, const float* bias, float alpha, int split_k, void* stream) {
    return ewdml_gemm_bf16(A, B, C, M, N, K, lda, ldb, ldc, 0, 0, out_mode, bias, alpha, split_k, stream);
}

// Implicit-GEMM 3x3 / stride 1 / pad 1 convolution passes on zero-bordered NHWC activations [N][H+2][W+2][C] (bf16).
//   pass 1 (fprop): A = padded x,   B = W [C_out][9*C_in],          C = y  compact [N*H*W][C_out]
//   pass 2 (dgrad): A = padded dy,  B = W [C_out][9*C_in],          C = dx compact [N*H*W][C_in]
//   pass 3 (wgrad): A = padded dy,  B = padded x,                   C = dW fp32 [C_out][9*C_in] (accumulated; out_mode forced to 2)
// C_in and C_out must be multiples of 64.  out_mode as in ewdml_gemm_bf16 (2 = fp32 accumulate, allows split-K).
extern "C" int ewdml_conv3x3_igemm(int pass, const void* A, const void* B, void* C, int N, int H, int W, int Cin, int Cout,
                                   int out_mode, const float* bias, int split_k, void* stream) {
    EWDML_CHECK_ARG(pass >= 1 && pass <= 3 && Cin % 64 == 0 && Cout % 64 == 0 && N > 0 && H > 0 && W > 0);
    const int WP = W + 2, HP = H + 2;
    const int64_t Mp64 = (int64_t)N * HP * WP, Mc64 = (int64_t)N * H * W;
    EWDML_CHECK_ARG(Mp64 < (1ll << 31));
    const int Mp = (int)Mp64, Mc = (int)Mc64, K9 = 9 * Cin;
    ConvMap cm;
    memset(&cm, 0, sizeof(cm));
    for (int kh = 0; kh < 3; ++kh)
        for (int kw = 0; kw < 3; ++kw) cm.off[kh * 3 + kw] = (kh - 1) * WP + (kw - 1);
    cm.H = H; cm.W = W; cm.cin = Cin;
    TmapDims td;
    if (pass == 1) {
        cm.mode = CONV_FPROP; cm.cblocks = Cin / 64; cm.remap = 1;
        td = {Mp, Cin, Cout, K9};
        return gemm_dispatch<false, false>(A, B, C, Mp, Cout, K9, Cin, K9, Cout, out_mode, bias, 1.f, split_k, stream, &cm, &td, 0, Mc,
                                           9 * (Cin / 64));
    }
    if (pass == 2) {
        cm.mode = CONV_DGRAD; cm.cblocks = Cout / 64; cm.remap = 1;
        td = {Mp, Cout, Cout, K9};                 // B: matrix [C_out rows][9*C_in contiguous], used MN-major
        return gemm_dispatch<false, true>(A, B, C, Mp, Cin, 9 * Cout, Cout, K9, Cin, out_mode, bias, 1.f, split_k, stream, &cm, &td,
                                          Cin == 64 ? 64 : 128, Mc, 9 * (Cout / 64));
    }
    cm.mode = CONV_WGRAD; cm.remap = 0;
    td = {Mp, Cout, Mp, Cin};                      // both MN-major: [padded rows][channels]
    return gemm_dispatch<true, true>(A, B, C, Cout, K9, Mp, Cout, Cin, K9, 2, nullptr, 1.f, split_k, stream, &cm, &td,
                                     Cin == 64 ? 64 : 128, Cout, (Mp + BK - 1) / BK);
}
