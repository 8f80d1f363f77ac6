// Exchange plan: device tables + scratch shared by select.cu / encode.cu / apply.cu.
#pragma once

#include "../common.cuh"

// per-segment selection state (device). Written by the bounds/refine/norm kernels, read by encode.
struct SegInfo {
    uint32_t key_lo;      // band [key_lo, key_hi): candidates; >= key_hi: definitely selected
    uint32_t key_hi;
    uint32_t cnt_hi;      // #elements with key >= key_hi          (atomic, reset by k_bounds)
    uint32_t ncand;       // #candidates appended                  (atomic, reset by k_bounds)
    uint32_t thr_key;     // final threshold: select key > thr_key, plus `need_eq` with key == thr_key
    uint32_t need_eq;
    uint32_t eq_taken;    // ticket counter for ties               (atomic, reset by k_refine)
    float norm;           // L2 norm of the selected values (QSGD)
    float scale;          // s / norm (0 when norm == 0)
    uint32_t fallback;    // 1 when the exact slow path ran for this segment (diagnostics)
    uint32_t pad0, pad1;
};

struct SegDesc {
    uint32_t offset;      // element offset in the flat buffer (multiple of 32)
    uint32_t numel;
    uint32_t k;           // top-k count
    uint32_t tile0;       // first tile of this segment
    uint32_t ntiles;
    uint32_t stride;      // sampling stride in 32-element lines (power of two)
    uint32_t pad0, pad1;
};

struct TileDesc {
    uint32_t seg;
    uint32_t start;       // element offset in the flat buffer
    uint32_t len;         // <= EWDML_TILE
    uint32_t pad;
};

struct SampleItem {       // one CTA of the sampling histogram kernel
    uint32_t seg;
    uint32_t line0;       // first sampled line index (in units of `stride` lines) within the segment
    uint32_t nlines;      // number of sampled 32-element lines handled by this CTA
    uint32_t pad;
};

struct SlotLayout {       // byte offsets inside a payload slot
    uint32_t norms;       // f32[nseg]
    uint32_t dir;         // {u32 off, u32 cnt}[ntiles]
    uint32_t bitmap;      // u32[ntiles * TILE/32]   (BITMAP mode)
    uint32_t data;        // chunk area
    uint64_t data_cap;    // bytes available in the chunk area
    uint32_t chunk_stride; // compressed modes: tile t's chunk lives at data + t * chunk_stride (worst-case size, so no allocator)
    uint32_t pad_;
    uint64_t total;       // slot size
};

// header words at the start of a slot
enum { HDR_SEQ = 0, HDR_ENTRIES = 1, HDR_DATA_BYTES = 2, HDR_MODE = 3, HDR_WORDS = 4 };

struct Counters {         // device, reset by the last CTA of k_encode_push
    unsigned long long data_used;   // bytes allocated in the chunk area
    unsigned int entries;
    unsigned int ctas_done;
    unsigned int overflow;
    unsigned int pad;
};

struct ewdml_plan {
    int nseg, ntiles, nsample;
    int64_t total;                  // elements spanned by the flat buffer
    ewdml_codec codec;
    SlotLayout layout;
    // device tables
    SegDesc* d_seg;
    TileDesc* d_tile;
    SampleItem* d_sample;
    // device scratch
    SegInfo* d_seginfo;
    uint32_t* d_hist;               // [nseg][EWDML_HIST_BINS]
    uint32_t* d_cand;               // [total] candidate keys, per segment at its element offset
    double* d_tile_sumsq;           // [ntiles] partial sums of squares
    Counters* d_ctr;
    uint32_t* d_stats;              // [4]: entries, data bytes, payload bytes lo, error
    int64_t scratch_bytes;
    int device;
};
