// Exact per-segment top-k threshold selection + L2 norms, in two full reads of the gradient.
//
// Replaces torch.topk(abs(t), k, sorted=False) (Compresssor/TopK.py:5-10) and tensor.norm()
// (Compresssor/qsgd.py:17) for every parameter tensor of the model in one launch chain:
//
//   k_sample_hist  11-bit histogram of |g| exponent/mantissa-top bits over a strided sample
//   k_bounds       per segment: a key band [lo, hi) that contains the k-th largest key with
//                  overwhelming probability (exact when the sample is the whole segment)
//   k_band         full read #1: count/sum-of-squares of keys >= hi, compact keys in [lo, hi)
//   k_refine       per segment: exact radix select among the candidates (3 digits of 11/11/9
//                  bits); falls back to an exact scan of the raw segment if the band missed;
//                  computes the norm of the selected values and the QSGD scale
//   (k_sumsq, k_norm_finish: norm only, for QSGD without top-k)
//
// Full read #2 is the encode kernel (encode.cu).  Keys are the fp32 bit patterns with the sign
// cleared, so integer order == magnitude order.
#include <stdlib.h>

#include "plan.h"

#define KEY(v) (__float_as_uint(v) & 0x7FFFFFFFu)
// selection test shared with encode.cu: keys are < 2^31, thr may be 0xFFFFFFFF (= -1, "select everything")
#define SEL_GT(key, thr) ((int32_t)((thr) - (key)) < 0)

__global__ void __launch_bounds__(256) k_sample_hist(const float* __restrict__ grad,
                                                     const SegDesc* __restrict__ segs,
                                                     const SampleItem* __restrict__ items,
                                                     uint32_t* __restrict__ hist) {
    __shared__ uint32_t h[EWDML_HIST_BINS];
    const SampleItem it = items[blockIdx.x];
    const SegDesc sg = segs[it.seg];
    for (int b = threadIdx.x; b < EWDML_HIST_BINS; b += 256) h[b] = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* base = grad + sg.offset;
    // 8 warps; each takes 4 sampled lines per iteration so 4 loads are in flight per lane
    for (uint32_t li = warp * 4; li < it.nlines; li += 32) {
        float v[4];
        bool ok[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint64_t e = (uint64_t)(it.line0 + li + j) * sg.stride * 32 + lane;
            ok[j] = (li + j) < it.nlines && e < sg.numel;
            v[j] = ok[j] ? __ldg(base + e) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (ok[j]) atomicAdd(&h[KEY(v[j]) >> 20], 1u);
    }
    __syncthreads();
    uint32_t* gh = hist + (size_t)it.seg * EWDML_HIST_BINS;
    for (int b = threadIdx.x; b < EWDML_HIST_BINS; b += 256)
        if (h[b]) atomicAdd(&gh[b], h[b]);
}

__global__ void __launch_bounds__(256) k_bounds(const SegDesc* __restrict__ segs, uint32_t* __restrict__ hist,
                                                SegInfo* __restrict__ info, double sigmas) {
    __shared__ uint32_t scan_smem[40];
    __shared__ int s_hi, s_lo;
    const SegDesc sg = segs[blockIdx.x];
    uint32_t* gh = hist + (size_t)blockIdx.x * EWDML_HIST_BINS;
    uint32_t c[8], local = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        c[j] = gh[threadIdx.x * 8 + j];
        gh[threadIdx.x * 8 + j] = 0;                 // self-cleaning for the next step
        local += c[j];
    }
    if (threadIdx.x == 0) { s_hi = -1; s_lo = 0; }
    uint32_t m;
    uint32_t below = block_excl_scan(local, scan_smem, &m);   // samples in bins < 8*tid
    const bool exact = sg.stride == 1;
    const double frac = (double)sg.k / (double)sg.numel;
    const double rs = (double)sg.k * ((double)m / (double)sg.numel);
    long long r_lo, r_hi;
    if (exact) {
        r_lo = r_hi = (long long)sg.k;
    } else {
        double var = rs * (1.0 - frac);
        double delta = sigmas * sqrt(var > 1.0 ? var : 1.0) + 2.0;
        r_lo = (long long)floor(rs - delta);
        r_hi = (long long)ceil(rs + delta);
    }
    long long cge = (long long)m - (long long)below;           // samples in bins >= 8*tid
    int my_hi = -1, my_lo = -1;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        int b = threadIdx.x * 8 + j;
        if (cge > r_lo) my_hi = b;                 // largest bin whose suffix count still exceeds r_lo
        if (cge >= r_hi) my_lo = b;                // largest bin whose suffix count reaches r_hi
        cge -= c[j];
    }
    if (my_hi >= 0) atomicMax(&s_hi, my_hi);
    if (my_lo >= 0) atomicMax(&s_lo, my_lo);
    __syncthreads();
    if (threadIdx.x == 0) {
        SegInfo& si = info[blockIdx.x];
        int ubin = s_hi + 1;                       // keys in bins >= ubin are "definitely selected"
        si.key_hi = ubin >= EWDML_HIST_BINS ? 0xFFFFFFFFu : ((uint32_t)ubin << 20);
        si.key_lo = (uint32_t)s_lo << 20;
        si.cnt_hi = 0;
        si.ncand = 0;
        si.eq_taken = 0;
        si.fallback = 0;
    }
}

// full read #1.  Persistent CTAs: the next tile's eight 16-byte loads per thread are issued before the current tile is
// processed, so global loads stay in flight through the scan / compaction phases.
__global__ void __launch_bounds__(256, 2) k_band(const float* __restrict__ grad, const TileDesc* __restrict__ tiles,
                                                 SegInfo* __restrict__ info, uint32_t* __restrict__ cand,
                                                 const SegDesc* __restrict__ segs, double* __restrict__ tile_sumsq, int ntiles) {
    __shared__ uint32_t scan_smem[40];
    __shared__ float ss_smem[8];
    __shared__ uint32_t s_base;
    const int tid = threadIdx.x;
    int tile = blockIdx.x;
    if (tile >= ntiles) return;
    TileDesc td = tiles[tile];
    uint4 v[8], nv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint32_t e = (tid + 256 * j) * 4;
        v[j] = e < td.len ? *reinterpret_cast<const uint4*>(grad + td.start + e) : make_uint4(0, 0, 0, 0);
    }
    while (true) {
        const int next = tile + gridDim.x;
        TileDesc ntd = td;
        if (next < ntiles) {
            ntd = tiles[next];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t e = (tid + 256 * j) * 4;
                nv[j] = e < ntd.len ? *reinterpret_cast<const uint4*>(grad + ntd.start + e) : make_uint4(0, 0, 0, 0);
            }
        }
        const uint32_t key_lo = info[td.seg].key_lo, key_hi = info[td.seg].key_hi;
        // masks by subtraction: bit i of mge_x = (key_i >= x) via the sign of (key - x); built by shift-insert from the top
        uint32_t mge_hi = 0, mge_lo = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t x[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint32_t key = x[q] & 0x7FFFFFFFu;
                mge_hi = (mge_hi >> 1) | (~(key - key_hi) & 0x80000000u);
                mge_lo = (mge_lo >> 1) | (~(key - key_lo) & 0x80000000u);
            }
        }
        if (key_hi > 0x7FFFFFFFu) mge_hi = 0;           // key_hi = 0xFFFFFFFF: nothing is "definitely selected"
        uint32_t mhi = mge_hi, mcand = mge_lo & ~mge_hi;
        float ss = 0.f;
        if (td.len != EWDML_TILE) {                     // partial tile: out-of-range lanes were loaded as +0.0f
            uint32_t valid = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if ((tid + 256 * j) * 4 + q < td.len) valid |= 1u << (4 * j + q);
            mhi &= valid; mcand &= valid;
        }
        // the few elements at or above the band are re-read from global memory (L2 hits) instead of dragging 32
        // predicated blocks through every warp
        for (uint32_t m = mhi; m; m &= m - 1) {
            const int b = __ffs(m) - 1;
            const float f = grad[td.start + (((tid + 256 * (b >> 2)) << 2) + (b & 3))];
            ss = fmaf(f, f, ss);
        }
        ss = warp_sum(ss);
        if ((tid & 31) == 0) ss_smem[tid >> 5] = ss;
        uint32_t tot;
        const uint32_t packed = __popc(mcand) | (__popc(mhi) << 16);
        const uint32_t excl = block_excl_scan(packed, scan_smem, &tot) & 0xFFFFu;
        if (tid == 0) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += (double)ss_smem[w];
            tile_sumsq[tile] = t;
            const uint32_t ctot = tot & 0xFFFFu, hitot = tot >> 16;
            if (hitot) atomicAdd(&info[td.seg].cnt_hi, hitot);
            s_base = ctot ? atomicAdd(&info[td.seg].ncand, ctot) : 0;
        }
        __syncthreads();
        if (mcand) {
            uint32_t* out = cand + segs[td.seg].offset + s_base + excl;
            for (uint32_t m = mcand; m; m &= m - 1) {
                const int b = __ffs(m) - 1;
                *out++ = __float_as_uint(grad[td.start + (((tid + 256 * (b >> 2)) << 2) + (b & 3))]) & 0x7FFFFFFFu;
            }
        }
        if (next >= ntiles) break;
        __syncthreads();                                // s_base / ss_smem are reused by the next tile
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = nv[j];
        td = ntd;
        tile = next;
    }
}

// One radix round over `n` keys: histogram the digit of keys whose higher bits equal `prefix`,
// then locate the digit holding rank r (1-based from the top).  All threads return the same
// digit; *r is reduced by the number of matching keys in higher digits.
template <bool RAW>
__device__ uint32_t radix_round(const uint32_t* keys, const float* raw, uint32_t n, int shift, uint32_t digit_mask,
                                int prefix_shift, uint32_t prefix, uint32_t* r, uint32_t* h, uint32_t* scan_smem,
                                uint32_t* s_digit) {
    for (int b = threadIdx.x; b < EWDML_HIST_BINS; b += blockDim.x) h[b] = 0;
    __syncthreads();
    // Warp-aggregated histogram: the candidates of a band share their leading bits, so in the first round nearly every
    // lane of a warp hits the SAME bin -- 32 serialised shared-memory atomics per warp and ~200 us for a 2.4 M-element
    // tensor in round 1.  Lanes with equal digits elect one leader that adds their count once.  (Uniform trip count so
    // the full-warp match is legal.)
    const int lane = threadIdx.x & 31;
    for (uint32_t base = 0; base < n; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        uint32_t d = 0xFFFFFFFFu;
        if (i < n) {
            const uint32_t key = RAW ? KEY(raw[i]) : keys[i];
            if (prefix_shift >= 32 || (key >> prefix_shift) == prefix) d = (key >> shift) & digit_mask;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (d != 0xFFFFFFFFu && lane == __ffs(peers) - 1) atomicAdd(&h[d], (uint32_t)__popc(peers));
    }
    __syncthreads();
    // 1024 threads x 2 bins, suffix sums via a forward scan
    uint32_t c0 = h[threadIdx.x * 2], c1 = h[threadIdx.x * 2 + 1], tot;
    uint32_t below = block_excl_scan(c0 + c1, scan_smem, &tot);
    uint32_t ge0 = tot - below;          // count in digits >= 2*tid
    uint32_t ge1 = ge0 - c0;             // count in digits >= 2*tid+1
    uint32_t ge2 = ge1 - c1;             // count in digits >= 2*tid+2
    const uint32_t rr = *r;
    __syncthreads();
    if (ge1 >= rr && ge2 < rr) { *s_digit = threadIdx.x * 2 + 1; *r = rr - ge2; }
    else if (ge0 >= rr && ge1 < rr) { *s_digit = threadIdx.x * 2; *r = rr - ge1; }
    __syncthreads();
    return *s_digit;
}

template <bool RAW>
__device__ void exact_select(const uint32_t* keys, const float* raw, uint32_t n, uint32_t rank, uint32_t* h,
                             uint32_t* scan_smem, uint32_t* s_digit, uint32_t* s_r, uint32_t* thr, uint32_t* need_eq, uint32_t* n_eq) {
    if (threadIdx.x == 0) *s_r = rank;
    __syncthreads();
    uint32_t d0 = radix_round<RAW>(keys, raw, n, 20, 0x7FF, 32, 0, s_r, h, scan_smem, s_digit);
    uint32_t d1 = radix_round<RAW>(keys, raw, n, 9, 0x7FF, 20, d0, s_r, h, scan_smem, s_digit);
    uint32_t p1 = (d0 << 11) | d1;
    uint32_t d2 = radix_round<RAW>(keys, raw, n, 0, 0x1FF, 9, p1, s_r, h, scan_smem, s_digit);
    *thr = (p1 << 9) | d2;
    *need_eq = *s_r;
    *n_eq = h[d2];                 // last round's bins are exact keys: how many keys equal the threshold
    __syncthreads();
}

__global__ void __launch_bounds__(1024) k_refine(const float* __restrict__ grad, const SegDesc* __restrict__ segs,
                                                 SegInfo* __restrict__ info, const uint32_t* __restrict__ cand,
                                                 const double* __restrict__ tile_sumsq, int want_norm, float levels,
                                                 uint32_t* __restrict__ fallback_total) {
    __shared__ uint32_t h[EWDML_HIST_BINS];
    __shared__ uint32_t scan_smem[40];
    __shared__ double red_smem[32];
    __shared__ uint32_t s_digit, s_r;
    const SegDesc sg = segs[blockIdx.x];
    SegInfo& si = info[blockIdx.x];
    const uint32_t cnt_hi = si.cnt_hi, ncand = si.ncand;
    const uint32_t* keys = cand + sg.offset;
    const float* raw = grad + sg.offset;
    uint32_t thr, need_eq, n_eq = 0;
    bool fallback = false;
    if (cnt_hi <= sg.k && sg.k - cnt_hi <= ncand) {
        uint32_t r = sg.k - cnt_hi;
        if (r == 0) {
            // everything at or above key_hi is selected and nothing else (key_hi == 0 -> thr = -1 -> "all", see SEL_GT)
            thr = si.key_hi - 1; need_eq = 0;
        }
        else exact_select<false>(keys, raw, ncand, r, h, scan_smem, &s_digit, &s_r, &thr, &need_eq, &n_eq);
    } else {
        fallback = true;                           // the sampled band missed: exact scan of the raw segment
        exact_select<true>(keys, raw, sg.numel, sg.k, h, scan_smem, &s_digit, &s_r, &thr, &need_eq, &n_eq);
    }
    // The usual case is a unique k-th magnitude (or all of its duplicates wanted): fold the ties into the strict
    // comparison by lowering the threshold one ulp, so the encode kernel needs a single compare per element and no
    // tickets.  Only a genuine tie (more equal keys than slots left) keeps need_eq != 0.
    if (need_eq != 0 && need_eq == n_eq) { thr -= 1; need_eq = 0; }
    double sumsq = 0.0;
    if (want_norm) {
        double acc = 0.0;
        if (!fallback) {
            for (uint32_t t = threadIdx.x; t < sg.ntiles; t += blockDim.x) acc += tile_sumsq[sg.tile0 + t];
            for (uint32_t i = threadIdx.x; i < ncand; i += blockDim.x) {
                uint32_t key = keys[i];
                if (SEL_GT(key, thr)) { double f = (double)__uint_as_float(key); acc += f * f; }
            }
        } else {
            for (uint32_t i = threadIdx.x; i < sg.numel; i += blockDim.x) {
                uint32_t key = KEY(raw[i]);
                if (SEL_GT(key, thr)) { double f = (double)__uint_as_float(key); acc += f * f; }
            }
        }
        sumsq = block_sum<double>(acc, red_smem);
        double t = need_eq ? (double)__uint_as_float(thr) : 0.0;
        sumsq += (double)need_eq * t * t;
    }
    if (threadIdx.x == 0) {
        si.thr_key = thr;
        si.need_eq = need_eq;
        si.eq_taken = 0;
        si.fallback = fallback ? 1u : 0u;
        if (fallback) atomicAdd(fallback_total, 1u);          // running count since plan creation (ewdml_plan_fallbacks)
        float norm = (float)sqrt(sumsq);
        bool ok = want_norm && norm > 0.f && isfinite(norm);
        si.norm = want_norm ? norm : 0.f;
        si.scale = ok ? __fdiv_rn(levels, norm) : 0.f;
    }
}

// ---- norm only (QSGD without top-k) ----------------------------------------------------------
__global__ void __launch_bounds__(256) k_sumsq(const float* __restrict__ grad, const TileDesc* __restrict__ tiles,
                                               double* __restrict__ tile_sumsq) {
    __shared__ double red_smem[32];
    const TileDesc td = tiles[blockIdx.x];
    const float* base = grad + td.start;
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t e = (threadIdx.x + 256 * j) * 4;
        if (e < td.len) {
            float4 v = *reinterpret_cast<const float4*>(base + e);
            const float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (e + q < td.len) ss += x[q] * x[q];
        }
    }
    double tot = block_sum<double>((double)ss, red_smem);
    if (threadIdx.x == 0) tile_sumsq[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(256) k_norm_finish(const SegDesc* __restrict__ segs, SegInfo* __restrict__ info,
                                                     const double* __restrict__ tile_sumsq, float levels) {
    __shared__ double red_smem[32];
    const SegDesc sg = segs[blockIdx.x];
    double acc = 0.0;
    for (uint32_t t = threadIdx.x; t < sg.ntiles; t += blockDim.x) acc += tile_sumsq[sg.tile0 + t];
    double sumsq = block_sum<double>(acc, red_smem);
    if (threadIdx.x == 0) {
        SegInfo& si = info[blockIdx.x];
        float norm = (float)sqrt(sumsq);
        bool ok = norm > 0.f && isfinite(norm);
        si.norm = norm;
        si.scale = ok ? __fdiv_rn(levels, norm) : 0.f;
        si.thr_key = 0; si.need_eq = 0; si.eq_taken = 0; si.fallback = 0; si.ncand = 0; si.cnt_hi = 0;
    }
}

// Width of the sampled band in standard deviations of the sampled rank (EWDML_BAND_SIGMA, default 12).  The estimate assumes
// an independent sample; the strided sample of a weight tensor is not quite one (per-channel scales), so a wider band trades a
// few more candidates for fewer exact-scan fallbacks, which one CTA pays with three passes over the whole tensor.  Measured on
// ResNet-50 (1 GPU, 209 exchanges x 161 tensors): 6 sigma = 65 fallbacks and 8.18 ms per step, 12 sigma = 0 and 7.57 ms, 24 sigma = 0
// and 7.62 ms; with W ranks any rank's fallback delays every rank, which was the multi-GPU gap to the dense path.
static double band_sigmas() {
    static double v = 0.0;
    if (v == 0.0) {
        const char* e = getenv("EWDML_BAND_SIGMA");
        v = e ? atof(e) : 12.0;
        if (!(v >= 1.0 && v <= 1000.0)) v = 12.0;
    }
    return v;
}

// Host: enqueue the selection chain for a plan.
int ewdml_launch_select(ewdml_plan* p, const float* grad, cudaStream_t st) {
    const ewdml_codec& c = p->codec;
    if (c.mode == EWDML_MODE_DENSE) return 0;
    const int want_norm = c.val == EWDML_VAL_QSGD;
    if (c.sel == EWDML_SEL_TOPK) {
        k_sample_hist<<<p->nsample, 256, 0, st>>>(grad, p->d_seg, p->d_sample, p->d_hist);
        EWDML_LAUNCH_CHECK();
        k_bounds<<<p->nseg, 256, 0, st>>>(p->d_seg, p->d_hist, p->d_seginfo, band_sigmas());
        EWDML_LAUNCH_CHECK();
        k_band<<<p->ntiles < 2 * ewdml_sm_count() ? p->ntiles : 2 * ewdml_sm_count(), 256, 0, st>>>(grad, p->d_tile, p->d_seginfo, p->d_cand, p->d_seg,
                                                                      p->d_tile_sumsq, p->ntiles);
        EWDML_LAUNCH_CHECK();
        k_refine<<<p->nseg, 1024, 0, st>>>(grad, p->d_seg, p->d_seginfo, p->d_cand, p->d_tile_sumsq, want_norm,
                                           (float)c.levels, p->d_stats + 4);
        EWDML_LAUNCH_CHECK();
    } else if (want_norm) {
        k_sumsq<<<p->ntiles, 256, 0, st>>>(grad, p->d_tile, p->d_tile_sumsq);
        EWDML_LAUNCH_CHECK();
        k_norm_finish<<<p->nseg, 256, 0, st>>>(p->d_seg, p->d_seginfo, p->d_tile_sumsq, (float)c.levels);
        EWDML_LAUNCH_CHECK();
    }
    return 0;
}
