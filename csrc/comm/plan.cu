// Exchange plan construction (host side): segment/tile/sample tables, slot layout, scratch.
//
// Replaces the per-step Python bookkeeping of the reference's receive buffers
// (ModelBuffer, distributed_worker.py:41-59; GradientAccumulator, sync_replicas_master_nn.py:49-68):
// the layout is computed once, lives on the device, and every kernel of the exchange indexes it.
#include <vector>

#include "plan.h"

static int sample_stride_lines(int64_t numel) {
    // Aim for >= 32768 sampled elements per segment; stride counts 32-element lines.
    int64_t s = numel / 32768;
    int stride = 1;
    while (stride * 2 <= s && stride < 64) stride *= 2;
    return stride;
}

extern "C" int ewdml_plan_create(ewdml_plan** out, int nseg, const int64_t* seg_offset,
                                 const int64_t* seg_numel, const ewdml_codec* codec) {
    EWDML_CHECK_ARG(out && nseg > 0 && seg_offset && seg_numel && codec);
    EWDML_CHECK_ARG(codec->mode >= EWDML_MODE_DENSE && codec->mode <= EWDML_MODE_BITMAP);
    if (codec->mode != EWDML_MODE_DENSE) {
        EWDML_CHECK_ARG(codec->val == EWDML_VAL_RAW32 ? codec->code_bits == 32
                                                      : (codec->code_bits >= 2 && codec->code_bits <= 16));
        EWDML_CHECK_ARG(codec->val == EWDML_VAL_RAW32 || (codec->levels >= 1 && codec->levels <= 32768));
        EWDML_CHECK_ARG(codec->sel == EWDML_SEL_ALL || (codec->topk_ratio > 0.0 && codec->topk_ratio <= 1.0));
        // a QSGD code holds sign + (level-1) in code_bits-1 bits
        EWDML_CHECK_ARG(codec->val == EWDML_VAL_RAW32 || (1 << (codec->code_bits - 1)) >= codec->levels);
    }
    ewdml_plan* p = new ewdml_plan();
    memset(p, 0, sizeof(*p));
    p->nseg = nseg;
    p->codec = *codec;
    cudaGetDevice(&p->device);

    std::vector<SegDesc> segs(nseg);
    std::vector<TileDesc> tiles;
    std::vector<SampleItem> samples;
    int64_t total = 0, emax = 0;
    for (int i = 0; i < nseg; ++i) {
        if (seg_offset[i] % 32 != 0 || seg_numel[i] <= 0 || seg_offset[i] < total ||
            seg_offset[i] + seg_numel[i] > 0xFFFFFFFFll) {
            delete p;
            ewdml_set_error("ewdml_plan_create: segment %d (offset %lld, numel %lld) must be 32-aligned, "
                            "ordered, non-overlapping and below 2^32", i, (long long)seg_offset[i],
                            (long long)seg_numel[i]);
            return EWDML_ERR_ARG;
        }
        SegDesc& s = segs[i];
        memset(&s, 0, sizeof(s));
        s.offset = (uint32_t)seg_offset[i];
        s.numel = (uint32_t)seg_numel[i];
        int64_t k = (int64_t)((double)seg_numel[i] * codec->topk_ratio);
        if (k < 1) k = 1;
        if (k > seg_numel[i]) k = seg_numel[i];
        s.k = codec->sel == EWDML_SEL_TOPK ? (uint32_t)k : s.numel;
        s.tile0 = (uint32_t)tiles.size();
        s.ntiles = (uint32_t)((seg_numel[i] + EWDML_TILE - 1) / EWDML_TILE);
        s.stride = (uint32_t)sample_stride_lines(seg_numel[i]);
        for (uint32_t t = 0; t < s.ntiles; ++t) {
            TileDesc td;
            td.seg = (uint32_t)i;
            td.start = s.offset + t * EWDML_TILE;
            td.len = (uint32_t)std::min<int64_t>(EWDML_TILE, seg_numel[i] - (int64_t)t * EWDML_TILE);
            td.pad = 0;
            tiles.push_back(td);
        }
        // sampled lines: line indices 0, stride, 2*stride, ... ; 256 sampled lines per CTA
        int64_t nlines = (seg_numel[i] + 31) / 32;
        int64_t nsampled = (nlines + s.stride - 1) / s.stride;
        for (int64_t l0 = 0; l0 < nsampled; l0 += 256) {
            SampleItem it;
            it.seg = (uint32_t)i;
            it.line0 = (uint32_t)l0;
            it.nlines = (uint32_t)std::min<int64_t>(256, nsampled - l0);
            it.pad = 0;
            samples.push_back(it);
        }
        emax += s.k;
        total = round_up_i64(seg_offset[i] + seg_numel[i], 32);
    }
    p->total = total;
    p->ntiles = (int)tiles.size();
    p->nsample = (int)samples.size();

    SlotLayout& L = p->layout;
    L.norms = 16;
    L.dir = (uint32_t)round_up_i64(L.norms + 4ll * nseg, 16);
    L.bitmap = (uint32_t)round_up_i64(L.dir + 8ll * p->ntiles, 16);
    int64_t bitmap_bytes = codec->mode == EWDML_MODE_BITMAP ? (int64_t)p->ntiles * (EWDML_TILE / 8) : 0;
    int64_t data_off = round_up_i64(L.bitmap + bitmap_bytes, 128);
    if (data_off > 0xFFFFFFFFll) { delete p; ewdml_set_error("plan too large"); return EWDML_ERR_ARG; }
    L.data = (uint32_t)data_off;
    L.chunk_stride = 0;
    if (codec->mode == EWDML_MODE_DENSE) {
        L.data_cap = (uint64_t)total * 4;
    } else {
        // Every tile owns a fixed, worst-case sized chunk (all 8192 elements selected).  Only the bytes actually
        // produced are written or read, so the slack costs address space, not bandwidth -- and the encode kernel needs
        // no global allocator (a same-address atomic with return value per tile serialises the whole grid).
        const uint32_t idx_max = codec->mode == EWDML_MODE_IDX16 ? EWDML_TILE * 2 : 0;
        const uint32_t code_max = (uint32_t)(((int64_t)EWDML_TILE * codec->code_bits + 7) / 8 + 15) & ~15u;
        L.chunk_stride = idx_max + code_max;
        L.data_cap = (uint64_t)L.chunk_stride * (uint64_t)p->ntiles;
        if (L.data_cap > 0xFFFF0000ull) { delete p; ewdml_set_error("plan too large for 32-bit chunk offsets"); return EWDML_ERR_ARG; }
    }
    L.total = (uint64_t)round_up_i64((int64_t)(L.data + L.data_cap), 256);

    auto up = [&](void** d, const void* h, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(d, bytes ? bytes : 16);
        if (e != cudaSuccess) return e;
        p->scratch_bytes += (int64_t)bytes;
        if (h) return cudaMemcpy(*d, h, bytes, cudaMemcpyHostToDevice);
        return cudaMemset(*d, 0, bytes ? bytes : 16);
    };
    cudaError_t e = cudaSuccess;
    if ((e = up((void**)&p->d_seg, segs.data(), sizeof(SegDesc) * nseg)) != cudaSuccess ||
        (e = up((void**)&p->d_tile, tiles.data(), sizeof(TileDesc) * tiles.size())) != cudaSuccess ||
        (e = up((void**)&p->d_sample, samples.data(), sizeof(SampleItem) * samples.size())) != cudaSuccess ||
        (e = up((void**)&p->d_seginfo, nullptr, sizeof(SegInfo) * nseg)) != cudaSuccess ||
        (e = up((void**)&p->d_hist, nullptr, sizeof(uint32_t) * EWDML_HIST_BINS * (size_t)nseg)) != cudaSuccess ||
        (e = up((void**)&p->d_tile_sumsq, nullptr, sizeof(double) * tiles.size())) != cudaSuccess ||
        (e = up((void**)&p->d_ctr, nullptr, sizeof(Counters))) != cudaSuccess ||
        (e = up((void**)&p->d_stats, nullptr, sizeof(uint32_t) * 8)) != cudaSuccess ||
        (e = up((void**)&p->d_cand, nullptr,
                codec->sel == EWDML_SEL_TOPK ? sizeof(uint32_t) * (size_t)total : 16)) != cudaSuccess) {
        ewdml_set_error("ewdml_plan_create: device allocation failed: %s", cudaGetErrorString(e));
        ewdml_plan_destroy(p);
        return (int)e;
    }
    *out = p;
    return 0;
}

extern "C" int ewdml_plan_destroy(ewdml_plan* p) {
    if (!p) return 0;
    cudaFree(p->d_seg); cudaFree(p->d_tile); cudaFree(p->d_sample); cudaFree(p->d_seginfo);
    cudaFree(p->d_hist); cudaFree(p->d_cand); cudaFree(p->d_tile_sumsq); cudaFree(p->d_ctr);
    cudaFree(p->d_stats);
    delete p;
    return 0;
}

extern "C" int64_t ewdml_plan_slot_bytes(const ewdml_plan* p) { return p ? (int64_t)p->layout.total : -1; }
extern "C" int64_t ewdml_plan_num_tiles(const ewdml_plan* p) { return p ? p->ntiles : -1; }
extern "C" int64_t ewdml_plan_scratch_bytes(const ewdml_plan* p) { return p ? p->scratch_bytes : -1; }
extern "C" int ewdml_plan_slot_layout(const ewdml_plan* p, int64_t out[4]) {
    EWDML_CHECK_ARG(p && out);
    out[0] = p->layout.norms; out[1] = p->layout.dir; out[2] = p->layout.bitmap; out[3] = p->layout.data;
    return 0;
}

// Segments whose sampled band missed the k-th magnitude (exact-scan fallback in k_refine) since the plan was created; syncs.
extern "C" int ewdml_plan_fallbacks(ewdml_plan* p, int64_t* out, void* stream) {
    EWDML_CHECK_ARG(p && out);
    uint32_t h = 0;
    EWDML_CUDA(cudaMemcpyAsync(&h, p->d_stats + 4, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    EWDML_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    *out = h;
    return 0;
}

extern "C" int ewdml_plan_read_stats(ewdml_plan* p, int64_t out[4], void* stream) {
    EWDML_CHECK_ARG(p && out);
    uint32_t h[8];
    EWDML_CUDA(cudaMemcpyAsync(h, p->d_stats, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    EWDML_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out[0] = h[0];
    out[1] = h[1];
    out[2] = p->codec.mode == EWDML_MODE_DENSE ? (int64_t)p->layout.data + (int64_t)p->layout.data_cap
                                              : (int64_t)p->layout.data + (int64_t)h[1];
    out[3] = h[3];
    return 0;
}

extern "C" int ewdml_plan_read_seginfo(ewdml_plan* p, uint32_t* out, void* stream) {
    EWDML_CHECK_ARG(p && out);
    std::vector<SegInfo> h(p->nseg);
    EWDML_CUDA(cudaMemcpyAsync(h.data(), p->d_seginfo, sizeof(SegInfo) * p->nseg, cudaMemcpyDeviceToHost,
                               (cudaStream_t)stream));
    EWDML_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    for (int i = 0; i < p->nseg; ++i) {
        out[4 * i + 0] = h[i].thr_key;
        out[4 * i + 1] = h[i].need_eq;
        memcpy(&out[4 * i + 2], &h[i].norm, 4);
        out[4 * i + 3] = h[i].ncand | (h[i].fallback << 31);
    }
    return 0;
}
