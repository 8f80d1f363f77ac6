// K1: fused select + QSGD quantise + bit-pack + peer store ("sparse allgather send").
//
// Persistent CTAs (2 per SM) walk the 8192-element tiles (full read #2 of the gradient), prefetching the next tile
// into registers.  Thread t owns the 16-byte vectors t, t+256, ... of a tile (eight coalesced loads), so the
// selection test on its registers IS bitmap word t (bit 4j+q <-> element (t+256j)*4+q; the decoder uses the same
// "thread-major" mapping).  A block scan of the word popcounts gives ordered output positions.  Dense-ish tiles
// (BITMAP mode) pack their codes behind the bitmap; IDX16 tiles under top-k (about 1 % of the elements) take a
// sparse path that gathers (index, value) pairs in shared memory with shared-memory atomics instead of running
// 32 predicated pack blocks per warp.  The tile's chunk
//     [u16 local index x cnt]  (IDX16 mode only)   [bit-packed codes x cnt]
// goes to a FIXED slot (tile * chunk_stride, worst-case sized: no global allocator atomic; only the bytes produced
// are written or read) with 16-byte stores straight into the receive slot of every destination over NVLink (or
// local memory for the own slot).  The last CTA to finish publishes header + per-destination flag with release
// semantics; there is no separate pack buffer, copy kernel or collective call.
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

template <int MODE, int SEL, int VAL>
__global__ void __launch_bounds__(256, 2) k_encode_push(const EncodeArgs a) {
    extern __shared__ __align__(16) uint32_t smem[];
    // layout: bitmap [256 words] | offsets [256 words] | idx16 staging [4096 words] | code staging [cw*256 words (+4)]
    uint32_t* s_bitmap = smem;
    uint32_t* s_offs = smem + 256;
    uint16_t* s_idx = reinterpret_cast<uint16_t*>(smem + 512);
    uint32_t* s_code = smem + 512 + 4096;
    __shared__ uint32_t scan_smem[40];
    __shared__ uint32_t s_off, s_last;

    int tile = blockIdx.x;
    TileDesc td = a.tiles[tile];
    const int tid = threadIdx.x;
    const uint32_t seq = a.state ? a.state[0] : a.seq;
    const uint32_t step = a.state ? a.state[1] : a.step;
    const ewdml_dests& D = a.dests[a.state ? (seq & 1u) : 0u];
    const bool mc = D.mc_slot != nullptr;          // NVLS: payload goes out once through the multicast mapping
    const int npush = mc ? 1 : D.n;

    if constexpr (MODE == EWDML_MODE_DENSE) {
        // plain tile copy into every destination (uncompressed push / allgather)
        const float* src = a.grad + td.start;
        const uint32_t n4 = (td.len + 3) >> 2;
        if (mc) {           // one multicast store per vector: the switch replicates it into every rank's slot
            char* dst = (char*)D.mc_slot + a.layout.data + (size_t)td.start * 4;
            for (uint32_t i = tid; i < n4; i += 256)
                mc_st_v4(dst + (size_t)i * 16, ld_nc_v4(src + i * 4));
        } else {
            for (int d = 0; d < D.n; ++d) {
                char* dst = (char*)D.slot[d] + a.layout.data + (size_t)td.start * 4;
                for (uint32_t i = tid; i < n4; i += 256)
                    st_na_v4(dst + (size_t)i * 16, *reinterpret_cast<const uint4*>(src + i * 4));
            }
        }
    } else {
        // Thread t holds elements (t + 256 j) * 4 + q, j = 0..7, q = 0..3 (eight coalesced 16-byte loads) and builds its own
        // 32-bit selection word (bit 4j+q).  Word t IS the tile's bitmap word t (a thread-major bit order that needs no
        // data movement); entries are ordered by (thread, bit), positions come from a block scan of the popcounts.
        const int cw = a.code_bits;
        const int code_words = (cw * 256 + 3);
        uint32_t v[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t e = (tid + 256 * j) * 4;
            uint4 u = make_uint4(0, 0, 0, 0);
            if (e < td.len) u = *reinterpret_cast<const uint4*>(a.grad + td.start + e);
            v[j][0] = u.x; v[j][1] = u.y; v[j][2] = u.z; v[j][3] = u.w;
        }
        // persistent loop: the next tile's loads are issued before this tile is processed
        uint32_t my_entries = 0;
        unsigned long long my_bytes = 0;
        while (true) {
        const int next = tile + gridDim.x;
        TileDesc ntd = td;
        uint4 nv[8];
        if (next < a.ntiles) {
            ntd = a.tiles[next];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t e = (tid + 256 * j) * 4;
                nv[j] = e < ntd.len ? *reinterpret_cast<const uint4*>(a.grad + ntd.start + e) : make_uint4(0, 0, 0, 0);
            }
        }
        const SegInfo si = a.info[td.seg];
        for (int i = tid; i < code_words; i += 256) s_code[i] = 0;
        // ---- selection word: branch-free compares, ties handled afterwards (they are rare) ----
        uint32_t mask = 0;
        if (SEL == EWDML_SEL_TOPK) {
            // selected <=> key > thr, evaluated as the sign of (thr - key) and shifted into the word from the top: four
            // integer instructions per element.  thr = 0xFFFFFFFF (-1) selects everything.
            const uint32_t thr = si.thr_key;
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    mask = (mask >> 1) | ((thr - (v[j][q] & 0x7FFFFFFFu)) & 0x80000000u);
            uint32_t valid = 0xFFFFFFFFu;
            if (td.len != EWDML_TILE) {                              // partial tile: drop the out-of-range lanes
                valid = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if ((tid + 256 * j) * 4 + q < td.len) valid |= 1u << (4 * j + q);
                mask &= valid;
            }
            if (si.need_eq) {                                        // genuine tie at the threshold (rare): take tickets
                uint32_t meq = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        meq |= ((v[j][q] & 0x7FFFFFFFu) == thr ? 1u : 0u) << (4 * j + q);
                meq &= valid;
                while (meq) {
                    const int b = __ffs(meq) - 1;
                    meq &= meq - 1;
                    if (atomicAdd(&a.info[td.seg].eq_taken, 1u) < si.need_eq) mask |= 1u << b;
                }
            }
        } else {
            mask = 0xFFFFFFFFu;
            if (td.len != EWDML_TILE) {
                mask = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if ((tid + 256 * j) * 4 + q < td.len) mask |= 1u << (4 * j + q);
            }
        }
        uint32_t total;
        uint32_t myword = mask;
        if constexpr (MODE == EWDML_MODE_IDX16 && SEL == EWDML_SEL_TOPK) {
            // Sparse path (density <= 1/16): selected elements are rare, so do not drag 32 predicated code blocks through
            // every warp.  Stage 1: each selected element is re-read from global memory (an L2 hit) and appended to a
            // shared list.  Stage 2: the list is processed densely, one entry per thread: QSGD level, drop level 0,
            // append (index, code) to the chunk.  Entry order inside a chunk is arbitrary -- the receiver does not care.
            uint32_t* s_cnt = s_bitmap;                               // [0]: list length, [1]: chunk entries
            uint16_t* s_le = reinterpret_cast<uint16_t*>(s_code + code_words + 1);
            uint32_t* s_lv = reinterpret_cast<uint32_t*>(s_le + EWDML_TILE);
            if (tid < 2) s_cnt[tid] = 0;
            __syncthreads();
            uint32_t m = mask;
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t e = ((tid + 256 * (b >> 2)) << 2) + (b & 3);
                const uint32_t bits = __float_as_uint(a.grad[td.start + e]);
                const uint32_t slot = atomicAdd(&s_cnt[0], 1u);
                s_le[slot] = (uint16_t)e;
                s_lv[slot] = bits;
            }
            __syncthreads();
            const uint32_t n1 = s_cnt[0];
            const float levels_f = (float)a.levels;
            for (uint32_t i = tid; i < n1; i += 256) {
                const uint32_t e = s_le[i], bits = s_lv[i];
                uint32_t code = bits;
                bool keep = true;
                if (VAL == EWDML_VAL_QSGD) {
                    const uint32_t gidx = td.start + e;
                    const Philox4 rnd = philox4x32_10(gidx >> 2, a.rng_stream, 0u, 0u, a.seed_lo, step);
                    const uint32_t ln = gidx & 3;
                    const uint32_t rx = ln == 0 ? rnd.v[0] : (ln == 1 ? rnd.v[1] : (ln == 2 ? rnd.v[2] : rnd.v[3]));
                    const float u = philox_to_uniform(rx);
                    const float lf = __fmul_rn(si.scale, __uint_as_float(bits & 0x7FFFFFFFu));
                    const float prev = floorf(lf);
                    float lvl = prev + ((u < __fsub_rn(lf, prev)) ? 1.f : 0.f);
                    lvl = fminf(lvl, levels_f);
                    keep = lvl >= 1.f;                                // level 0 carries no information
                    code = ((bits >> 31) << (cw - 1)) | ((uint32_t)lvl - 1u);
                }
                if (keep) {
                    const uint32_t p = atomicAdd(&s_cnt[1], 1u);
                    s_idx[p] = (uint16_t)e;
                    if (cw == 32) s_code[p] = code;
                    else put_bits(s_code, p * cw, code, cw);
                }
            }
            __syncthreads();
            total = s_cnt[1];
        } else {
            // ---- QSGD: stochastic level for the selected elements; level 0 is dropped from the selection ----
            if (VAL == EWDML_VAL_QSGD && mask) {
                const float levels_f = (float)a.levels;
    #pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if ((mask >> (4 * j)) & 0xFu) {
                        const uint32_t gidx0 = td.start + (tid + 256 * j) * 4;       // multiple of 4: one Philox block per float4
                        const Philox4 rnd = philox4x32_10(gidx0 >> 2, a.rng_stream, 0u, 0u, a.seed_lo, step);
    #pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            if (mask & (1u << (4 * j + q))) {
                                const uint32_t bits = v[j][q];
                                const float u = philox_to_uniform(rnd.v[q]);
                                const float lf = __fmul_rn(si.scale, __uint_as_float(bits & 0x7FFFFFFFu));
                                const float prev = floorf(lf);
                                float lvl = prev + ((u < __fsub_rn(lf, prev)) ? 1.f : 0.f);
                                lvl = fminf(lvl, levels_f);
                                if (lvl >= 1.f) v[j][q] = ((bits >> 31) << (cw - 1)) | ((uint32_t)lvl - 1u);
                                else mask &= ~(1u << (4 * j + q));
                            }
                        }
                    }
                }
            }
            const uint32_t pos0 = block_excl_scan(__popc(mask), scan_smem, &total);
            myword = mask;
            // ---- assemble the chunk in shared memory (only threads that own selected elements do work) ----
            if (mask) {
                uint32_t p = pos0;
    #pragma unroll
                for (int j = 0; j < 8; ++j) {
    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (mask & (1u << (4 * j + q))) {
                            if (MODE == EWDML_MODE_IDX16) s_idx[p] = (uint16_t)((tid + 256 * j) * 4 + q);
                            if (cw == 32) s_code[p] = v[j][q];
                            else put_bits(s_code, p * cw, v[j][q], cw);
                            ++p;
                        }
                    }
                }
            }
        }
        const uint32_t idx_bytes = MODE == EWDML_MODE_IDX16 ? align16(total * 2) : 0;
        const uint32_t code_bytes = align16((total * cw + 7) >> 3);
        my_entries += total;                             // (identical in every thread; thread 0 publishes them once per CTA)
        my_bytes += idx_bytes + code_bytes;
        __syncthreads();
        const uint32_t off = (uint32_t)tile * a.layout.chunk_stride;
        const bool overflow = false;
        // ---- push to every destination ------------------------------------------------------
        for (int d = 0; d < npush; ++d) {
            char* slot = mc ? (char*)D.mc_slot : (char*)D.slot[d];
            if (tid == 0) {
                uint2 de = make_uint2(off, overflow ? 0u : total);
                push_b64(slot + a.layout.dir + (size_t)tile * 8, de, mc);
            }
            if (MODE == EWDML_MODE_BITMAP)
                push_b32(reinterpret_cast<uint32_t*>(slot + a.layout.bitmap) + (size_t)tile * 256 + tid, myword, mc);
            if (!overflow) {
                char* dst = slot + a.layout.data + off;
                const uint4* si4 = reinterpret_cast<const uint4*>(s_idx);
                for (uint32_t i = tid; i < (idx_bytes >> 4); i += 256) push_v4(dst + (size_t)i * 16, si4[i], mc);
                const uint4* sc4 = reinterpret_cast<const uint4*>(s_code);
                for (uint32_t i = tid; i < (code_bytes >> 4); i += 256)
                    push_v4(dst + idx_bytes + (size_t)i * 16, sc4[i], mc);
            }
        }
        if (next >= a.ntiles) break;
        __syncthreads();                                 // staging buffers are reused by the next tile
#pragma unroll
        for (int j = 0; j < 8; ++j) { v[j][0] = nv[j].x; v[j][1] = nv[j].y; v[j][2] = nv[j].z; v[j][3] = nv[j].w; }
        td = ntd;
        tile = next;
        }   // persistent tile loop
        if (tid == 0) {
            atomicAdd(&a.ctr->entries, my_entries);
            atomicAdd(&a.ctr->data_used, my_bytes);
        }
    }
    // ---- completion: last CTA publishes header, norms and flags ------------------------------
    __syncthreads();
    if (tid == 0) {
        // device-scope fence is enough here: this CTA's stores are ordered before its counter increment, the last CTA
        // observes every increment and issues the system-scope fence + release before any peer can see the flag
        __threadfence();
        s_last = atomicAdd(&a.ctr->ctas_done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence_system();
        const uint32_t entries = a.ctr->entries;
        const uint32_t data_bytes = (uint32_t)a.ctr->data_used;
        const uint32_t overflow = a.ctr->overflow;
        for (int d = 0; d < npush; ++d) {
            char* slot = mc ? (char*)D.mc_slot : (char*)D.slot[d];
            float* norms = reinterpret_cast<float*>(slot + a.layout.norms);
            if (MODE != EWDML_MODE_DENSE)
                for (int s = tid; s < a.nseg; s += 256) push_b32(norms + s, __float_as_uint(a.info[s].norm), mc);
            if (tid == 0) {
                uint32_t h[4];
                h[HDR_ENTRIES] = entries;
                h[HDR_DATA_BYTES] = data_bytes;
                h[HDR_MODE] = (uint32_t)MODE | ((uint32_t)a.code_bits << 8) | (overflow << 31);
                h[HDR_SEQ] = seq;
                push_v4(slot, make_uint4(h[0], h[1], h[2], h[3]), mc);
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
    size_t words = 512 + 4096 + (size_t)c.code_bits * 256 + 4;
    if (c.mode == EWDML_MODE_IDX16 && c.sel == EWDML_SEL_TOPK) words += 4 + EWDML_TILE / 2 + EWDML_TILE;   // sparse-path lists
    return words * 4;
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
#define ENC_LAUNCH(MODE_, SEL_, VAL_)                                                                                   \
    do {                                                                                                              \
        EWDML_CUDA(cudaFuncSetAttribute(k_encode_push<MODE_, SEL_, VAL_>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                        160 * 1024));                                                                 \
        k_encode_push<MODE_, SEL_, VAL_><<<p->ntiles < 2 * ewdml_sm_count() ? p->ntiles : 2 * ewdml_sm_count(), 256, smem, st>>>(a);                                            \
    } while (0)
    const int sel = p->codec.sel, val = p->codec.val;
    if (p->codec.mode == EWDML_MODE_DENSE) {
        k_encode_push<EWDML_MODE_DENSE, 0, 0><<<p->ntiles, 256, 0, st>>>(a);
    } else if (p->codec.mode == EWDML_MODE_IDX16) {
        if (sel == EWDML_SEL_TOPK && val == EWDML_VAL_QSGD) ENC_LAUNCH(EWDML_MODE_IDX16, EWDML_SEL_TOPK, EWDML_VAL_QSGD);
        else if (sel == EWDML_SEL_TOPK) ENC_LAUNCH(EWDML_MODE_IDX16, EWDML_SEL_TOPK, EWDML_VAL_RAW32);
        else if (val == EWDML_VAL_QSGD) ENC_LAUNCH(EWDML_MODE_IDX16, EWDML_SEL_ALL, EWDML_VAL_QSGD);
        else ENC_LAUNCH(EWDML_MODE_IDX16, EWDML_SEL_ALL, EWDML_VAL_RAW32);
    } else {
        if (sel == EWDML_SEL_TOPK && val == EWDML_VAL_QSGD) ENC_LAUNCH(EWDML_MODE_BITMAP, EWDML_SEL_TOPK, EWDML_VAL_QSGD);
        else if (sel == EWDML_SEL_TOPK) ENC_LAUNCH(EWDML_MODE_BITMAP, EWDML_SEL_TOPK, EWDML_VAL_RAW32);
        else if (val == EWDML_VAL_QSGD) ENC_LAUNCH(EWDML_MODE_BITMAP, EWDML_SEL_ALL, EWDML_VAL_QSGD);
        else ENC_LAUNCH(EWDML_MODE_BITMAP, EWDML_SEL_ALL, EWDML_VAL_RAW32);
    }
#undef ENC_LAUNCH
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
