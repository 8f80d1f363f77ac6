// K1: fused select + QSGD quantise + bit-pack + peer store ("sparse allgather send").
//
// One CTA per 8192-element tile (full read #2 of the gradient).  The tile is staged in shared
// memory so that thread t owns elements [32t, 32t+32): its selection mask IS bitmap word t, a block
// scan of the popcounts gives ordered output positions, and the tile's chunk
//     [u16 local index x cnt]  (IDX16 mode only)   [bit-packed codes x cnt]
// is assembled in shared memory, given a byte range of the payload by ONE atomicAdd, and written
// with 16-byte stores straight into the receive slot of every destination over NVLink (or local
// memory for the own slot).  The last CTA to finish publishes header + per-destination flag with
// release semantics; there is no separate pack buffer, copy kernel or collective call.
//
// Replaces: Compresssor/qsgd.py:12-32 (8 elementwise kernels + RNG), Compresssor/TopK.py:5-10
// (topk + gather), the D2H copy and dist.gather per tensor (distributed_worker.py:317-351), and on
// the Horovod path compress + allreduce (horovod_compression.py:16-35, horvod_pytorch.py:194-201).
#include "plan.h"

#define KEY(v) (__float_as_uint(v) & 0x7FFFFFFFu)
#define PAD(e) ((e) + ((e) >> 5))          // conflict-free: thread t reads 33t + j

struct EncodeArgs {
    const float* grad;
    const TileDesc* tiles;
    const SegDesc* segs;
    SegInfo* info;
    Counters* ctr;
    uint32_t* stats;
    SlotLayout layout;
    ewdml_dests dests[2];              // [parity]; both equal when no device step state is used
    const uint32_t* state;             // optional device step state {seq, step, first}: makes the launch graph-replayable
    int mode, sel, val, code_bits, levels, nseg, ntiles;
    uint32_t seq, seed_lo, seed_hi, step, rng_stream;
    uint32_t* err_word;
};

__device__ __forceinline__ void put_bits(uint32_t* words, uint32_t bitpos, uint32_t code, int nbits) {
    uint32_t w = bitpos >> 5, sh = bitpos & 31;
    atomicOr(&words[w], code << sh);
    if (sh + nbits > 32) atomicOr(&words[w + 1], code >> (32 - sh));
}

template <int MODE>
__global__ void __launch_bounds__(256) k_encode_push(const EncodeArgs a) {
    extern __shared__ __align__(16) uint32_t smem[];
    // layout: tile values/codes [8448 words] | idx16 staging [4096 words] | code staging [cw*256 words (+4)]
    uint32_t* s_tile = smem;
    uint16_t* s_idx = reinterpret_cast<uint16_t*>(smem + 8448);
    uint32_t* s_code = smem + 8448 + 4096;
    __shared__ uint32_t scan_smem[40];
    __shared__ uint32_t s_off, s_last;

    const TileDesc td = a.tiles[blockIdx.x];
    const int tid = threadIdx.x;
    const uint32_t seq = a.state ? a.state[0] : a.seq;
    const uint32_t step = a.state ? a.state[1] : a.step;
    const ewdml_dests& D = a.dests[a.state ? (seq & 1u) : 0u];

    if constexpr (MODE == EWDML_MODE_DENSE) {
        // plain tile copy into every destination (uncompressed push / allgather)
        const float* src = a.grad + td.start;
        const uint32_t n4 = (td.len + 3) >> 2;
        for (int d = 0; d < D.n; ++d) {
            char* dst = (char*)D.slot[d] + a.layout.data + (size_t)td.start * 4;
            for (uint32_t i = tid; i < n4; i += 256)
                st_na_v4(dst + (size_t)i * 16, *reinterpret_cast<const uint4*>(src + i * 4));
        }
    } else {
        const SegInfo si = a.info[td.seg];
        const int cw = a.code_bits;
        const int code_words = (cw * 256 + 3);           // <= cw*8192/32 words, +padding for 16 B copies
        // ---- stage the tile ------------------------------------------------------------------
        const float* src = a.grad + td.start;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t e = (tid + 256 * j) * 4;
            if (e < td.len) {
                float4 v = *reinterpret_cast<const float4*>(src + e);
                s_tile[PAD(e)] = __float_as_uint(v.x);
                s_tile[PAD(e + 1)] = __float_as_uint(v.y);
                s_tile[PAD(e + 2)] = __float_as_uint(v.z);
                s_tile[PAD(e + 3)] = __float_as_uint(v.w);
            }
        }
        for (int i = tid; i < code_words; i += 256) s_code[i] = 0;
        __syncthreads();
        // ---- select + quantise: thread owns elements [32 tid, 32 tid + 32) --------------------
        uint32_t mask = 0;
        const uint32_t e0 = tid * 32;
        const float levels_f = (float)a.levels;
        uint32_t cached_group = 0xFFFFFFFFu;
        Philox4 rnd;
        rnd.v[0] = rnd.v[1] = rnd.v[2] = rnd.v[3] = 0;
        if (e0 < td.len) {
#pragma unroll 4
            for (int j = 0; j < 32; ++j) {
                const uint32_t e = e0 + j;
                if (e >= td.len) break;
                const uint32_t bits = s_tile[PAD(e)];
                const uint32_t key = bits & 0x7FFFFFFFu;
                bool sel = true;
                if (a.sel == EWDML_SEL_TOPK) {
                    sel = key > si.thr_key;
                    if (!sel && key == si.thr_key && si.need_eq)
                        sel = atomicAdd(&a.info[td.seg].eq_taken, 1u) < si.need_eq;
                }
                uint32_t code = bits;                       // RAW32: the fp32 bit pattern
                if (sel && a.val == EWDML_VAL_QSGD) {
                    const uint32_t gidx = td.start + e;     // global element index -> RNG counter
                    const uint32_t group = gidx >> 2;
                    if (group != cached_group) {
                        rnd = philox4x32_10(group, a.rng_stream, 0u, 0u, a.seed_lo, step);
                        cached_group = group;
                    }
                    const uint32_t ln = gidx & 3;
                    const uint32_t rx = ln == 0 ? rnd.v[0] : (ln == 1 ? rnd.v[1] : (ln == 2 ? rnd.v[2] : rnd.v[3]));
                    const float u = philox_to_uniform(rx);
                    const float lf = __fmul_rn(si.scale, __uint_as_float(key));
                    const float prev = floorf(lf);
                    float lvl = prev + ((u < __fsub_rn(lf, prev)) ? 1.f : 0.f);
                    lvl = fminf(lvl, levels_f);
                    sel = lvl >= 1.f;                        // level 0 carries no information
                    code = ((bits >> 31) << (cw - 1)) | ((uint32_t)lvl - 1u);
                }
                if (sel) {
                    mask |= 1u << j;
                    s_tile[PAD(e)] = code;                  // keep the code in place of the value
                }
            }
        }
        const uint32_t cnt = __popc(mask);
        uint32_t total;
        const uint32_t pos0 = block_excl_scan(cnt, scan_smem, &total);
        // ---- assemble the chunk in shared memory --------------------------------------------
        {
            uint32_t m = mask, p = pos0;
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t code = s_tile[PAD(e0 + j)];
                if (MODE == EWDML_MODE_IDX16) s_idx[p] = (uint16_t)(e0 + j);
                if (cw == 32) s_code[p] = code;
                else put_bits(s_code, p * cw, code, cw);
                ++p;
            }
        }
        const uint32_t idx_bytes = MODE == EWDML_MODE_IDX16 ? align16(total * 2) : 0;
        const uint32_t code_bytes = align16((total * cw + 7) >> 3);
        if (tid == 0) {
            unsigned long long off = atomicAdd(&a.ctr->data_used, (unsigned long long)(idx_bytes + code_bytes));
            if (off + idx_bytes + code_bytes > a.layout.data_cap) {
                atomicExch(&a.ctr->overflow, 1u);
                off = 0;
            }
            s_off = (uint32_t)off;
            atomicAdd(&a.ctr->entries, total);
        }
        __syncthreads();
        const uint32_t off = s_off;
        const bool overflow = a.ctr->overflow != 0;
        // ---- push to every destination ------------------------------------------------------
        for (int d = 0; d < D.n; ++d) {
            char* slot = (char*)D.slot[d];
            if (tid == 0) {
                uint2 de = make_uint2(off, overflow ? 0u : total);
                *reinterpret_cast<uint2*>(slot + a.layout.dir + (size_t)blockIdx.x * 8) = de;
            }
            if (MODE == EWDML_MODE_BITMAP)
                reinterpret_cast<uint32_t*>(slot + a.layout.bitmap)[(size_t)blockIdx.x * 256 + tid] = mask;
            if (!overflow) {
                char* dst = slot + a.layout.data + off;
                const uint4* si4 = reinterpret_cast<const uint4*>(s_idx);
                for (uint32_t i = tid; i < (idx_bytes >> 4); i += 256) st_na_v4(dst + (size_t)i * 16, si4[i]);
                const uint4* sc4 = reinterpret_cast<const uint4*>(s_code);
                for (uint32_t i = tid; i < (code_bytes >> 4); i += 256)
                    st_na_v4(dst + idx_bytes + (size_t)i * 16, sc4[i]);
            }
        }
    }
    // ---- completion: last CTA publishes header, norms and flags ------------------------------
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        s_last = atomicAdd(&a.ctr->ctas_done, 1u) == (uint32_t)(a.ntiles - 1);
    }
    __syncthreads();
    if (s_last) {
        __threadfence_system();
        const uint32_t entries = a.ctr->entries;
        const uint32_t data_bytes = (uint32_t)a.ctr->data_used;
        const uint32_t overflow = a.ctr->overflow;
        for (int d = 0; d < D.n; ++d) {
            char* slot = (char*)D.slot[d];
            float* norms = reinterpret_cast<float*>(slot + a.layout.norms);
            if (MODE != EWDML_MODE_DENSE)
                for (int s = tid; s < a.nseg; s += 256) norms[s] = a.info[s].norm;
            if (tid == 0) {
                uint32_t* hdr = reinterpret_cast<uint32_t*>(slot);
                hdr[HDR_ENTRIES] = entries;
                hdr[HDR_DATA_BYTES] = data_bytes;
                hdr[HDR_MODE] = (uint32_t)MODE | ((uint32_t)a.code_bits << 8) | (overflow << 31);
                hdr[HDR_SEQ] = seq;
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            for (int d = 0; d < D.n; ++d)
                if (D.flag[d]) st_release_sys(D.flag[d], seq);
            a.stats[0] = entries;
            a.stats[1] = data_bytes;
            a.stats[3] = overflow ? (uint32_t)EWDML_ERR_OVERFLOW : 0u;
            if (overflow) atomicExch(a.err_word, (uint32_t)EWDML_ERR_OVERFLOW);
            a.ctr->data_used = 0;
            a.ctr->entries = 0;
            a.ctr->ctas_done = 0;
            a.ctr->overflow = 0;
        }
    }
}

int ewdml_launch_select(ewdml_plan* p, const float* grad, cudaStream_t st);

static size_t encode_smem_bytes(const ewdml_codec& c) {
    if (c.mode == EWDML_MODE_DENSE) return 0;
    return (size_t)(8448 + 4096 + c.code_bits * 256 + 4) * 4;
}

static int send_impl(ewdml_plan* p, const float* grad, const ewdml_dests* dests, const ewdml_dests* dests_odd,
                     const uint32_t* state, uint32_t seq, uint64_t seed, uint32_t step, uint32_t rng_stream, void* stream) {
    EWDML_CHECK_ARG(p && grad && dests && dests->n >= 1 && dests->n <= EWDML_MAX_PEERS);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = ewdml_launch_select(p, grad, st);
    if (rc) return rc;
    EncodeArgs a;
    memset(&a, 0, sizeof(a));
    a.grad = grad; a.tiles = p->d_tile; a.segs = p->d_seg; a.info = p->d_seginfo; a.ctr = p->d_ctr;
    a.stats = p->d_stats; a.layout = p->layout; a.dests[0] = *dests; a.dests[1] = dests_odd ? *dests_odd : *dests;
    a.state = state;
    a.mode = p->codec.mode; a.sel = p->codec.sel; a.val = p->codec.val; a.code_bits = p->codec.code_bits;
    a.levels = p->codec.levels; a.nseg = p->nseg; a.ntiles = p->ntiles;
    a.seq = seq; a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32); a.step = step;
    a.rng_stream = rng_stream;
    a.err_word = ewdml_device_error_word();
    if (!a.err_word) return EWDML_ERR_CUDA;
    const size_t smem = encode_smem_bytes(p->codec);
    static bool attr_set[3] = {false, false, false};
    switch (p->codec.mode) {
        case EWDML_MODE_DENSE:
            k_encode_push<EWDML_MODE_DENSE><<<p->ntiles, 256, 0, st>>>(a);
            break;
        case EWDML_MODE_IDX16:
            if (!attr_set[1]) {
                EWDML_CUDA(cudaFuncSetAttribute(k_encode_push<EWDML_MODE_IDX16>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
                attr_set[1] = true;
            }
            k_encode_push<EWDML_MODE_IDX16><<<p->ntiles, 256, smem, st>>>(a);
            break;
        default:
            if (!attr_set[2]) {
                EWDML_CUDA(cudaFuncSetAttribute(k_encode_push<EWDML_MODE_BITMAP>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
                attr_set[2] = true;
            }
            k_encode_push<EWDML_MODE_BITMAP><<<p->ntiles, 256, smem, st>>>(a);
            break;
    }
    EWDML_LAUNCH_CHECK();
    ewdml_counter_add(EWDML_CTR_SEND_CALLS, 1);
    return 0;
}

extern "C" int ewdml_send(ewdml_plan* p, const float* grad, const ewdml_dests* dests, uint32_t seq, uint64_t seed,
                          uint32_t step, uint32_t rng_stream, void* stream) {
    return send_impl(p, grad, dests, nullptr, nullptr, seq, seed, step, rng_stream, stream);
}

// Graph-replayable form: sequence number, RNG step and slot parity come from the device step state
// (ewdml_state_advance), dests_even/dests_odd are the destination sets for even/odd sequence numbers.
extern "C" int ewdml_send_state(ewdml_plan* p, const float* grad, const ewdml_dests* dests_even, const ewdml_dests* dests_odd,
                                const uint32_t* state, uint64_t seed, uint32_t rng_stream, void* stream) {
    EWDML_CHECK_ARG(state && dests_odd);
    return send_impl(p, grad, dests_even, dests_odd, state, 0, seed, 0, rng_stream, stream);
}
