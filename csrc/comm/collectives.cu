// N3: collective engine -- bcast / gather / scatter / allgather / alltoall / reduce / reduce_scatter /
// allreduce / send / recv / barrier on arbitrary device buffers, over a symmetric "window" in peer memory.
//
// Replaces the coll framework's per-communicator function table (ompi:ompi/mca/coll/coll.h:142-255), the
// algorithm zoo of coll/base (ompi:ompi/mca/coll/base/coll_base_{allreduce,allgather,bcast,gather,
// reduce_scatter,alltoall,barrier}.c), the size-based decision layer of coll/tuned
// (ompi:ompi/mca/coll/tuned/coll_tuned_decision_fixed.c:55-217, forced-algorithm variables
// coll_tuned_allreduce_decision.c:37-46) and pml/ob1 tagged point-to-point
// (ompi:ompi/mca/pml/ob1/pml_ob1_{isend,irecv}.c).  None of their algorithms is kept: on an NVSwitch every
// peer is one hop away at full bandwidth, so every operation here is ONE kernel with two phases,
//
//   put   every rank stores its contribution straight into the destination ranks' windows (16-byte remote
//         stores over NVLink), the last CTA writes a 16-byte slot header and publishes a sequence-numbered
//         flag with release semantics at system scope;
//   wait  acquire-spin on the flags of the expected sources, then copy (gather-type operations) or sum in
//         rank order and scale (reduce-type operations) out of the LOCAL window, and acknowledge.
//
// The operations differ only in who puts what where and in how the window is consumed, so one kernel
// serves all of them.  What coll/tuned decides with rule tables is, here, one choice: allreduce as one-shot
// (allgather-style puts, W x N bytes in, latency optimal) or two-shot (reduce_scatter + allgather,
// 2 (W-1)/W x N), switched on message size with an environment / API override.
//
// Window of one rank: [2 regions: collective, p2p][2 parities][W slots][slot_stride bytes]; slot s of a rank's
// window is written only by rank s.  Operation number q uses parity q & 1.  A writer may overwrite a slot only
// after the owner has consumed what the writer put there two operations earlier; the owner says so through a
// per-writer acknowledgement flag, and the writer's kernel waits for ack >= (its last put to that peer with
// this parity) before storing.  Flags only grow, so nothing is ever reset.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "../common.cuh"

namespace {

enum { PH_PUT = 1, PH_WAIT = 2 };
enum { CONS_NONE = 0, CONS_COPY = 1, CONS_SUM = 2 };
// flag block (uint32 words): per region 2*MAX_PEERS words: DATA[src], ACK[dst]
constexpr int kFlagWordsPerRegion = 2 * EWDML_MAX_PEERS;
constexpr int kHeaderBytes = 16;

struct CollArgs {
    int rank, world;
    uint8_t* win[EWDML_MAX_PEERS];
    uint32_t* peer_flags[EWDML_MAX_PEERS];
    uint32_t* flags;
    uint32_t* counters;          // 2 words, CTA election for the two phases (self-resetting)
    int64_t slot_stride;         // bytes per slot including the header
    int64_t region_off;          // byte offset of the region inside a window
    int flag_base;               // word offset of the region's flags
    uint32_t seq;
    uint32_t op_tag;             // op | tag << 8, checked by the consumer
    int phases;
    // put
    uint32_t put_mask;
    const uint8_t* src;
    int64_t src_stride;          // contribution for peer p starts at src + p * src_stride (0: same bytes for everyone)
    uint32_t ack_need[EWDML_MAX_PEERS];
    // wait / consume
    uint32_t src_mask;
    uint8_t* dst;
    int64_t dst_stride;          // CONS_COPY: slot s goes to dst + s * dst_stride
    int consume;
    float scale;
    int64_t nbytes;              // payload bytes per slot (multiple of 16)
    uint32_t* err_word;
    unsigned long long spin_limit;
};

__device__ __forceinline__ uint8_t* slot_ptr(const CollArgs& a, int owner, int writer) {
    return a.win[owner] + a.region_off + ((int64_t)((a.seq & 1u) * a.world + writer)) * a.slot_stride;
}

__device__ __forceinline__ bool elect_last(uint32_t* counter, int* s_flag) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        *s_flag = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    return *s_flag != 0;
}

// grid-stride copy of n16 16-byte vectors with four loads in flight per thread
__device__ __forceinline__ void copy16x4(uint8_t* __restrict__ d, const uint8_t* __restrict__ s, int64_t n16, int64_t gtid, int64_t gsize) {
    for (int64_t i0 = gtid; i0 < n16; i0 += 4 * gsize) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t i = i0 + u * gsize;
            if (i < n16) v[u] = ld_v4(s + (i << 4));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t i = i0 + u * gsize;
            if (i < n16) st_na_v4(d + (i << 4), v[u]);
        }
    }
}

__global__ void __launch_bounds__(512) k_coll(const CollArgs a) {
    __shared__ int s_ok, s_last;
    const int tid = threadIdx.x, W = a.world, me = a.rank;
    const int64_t n16 = a.nbytes >> 4;
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + tid, gsize = (int64_t)gridDim.x * blockDim.x;
    uint32_t* my_data = a.flags + a.flag_base;                         // DATA[src]: written by src
    uint32_t* my_ack = a.flags + a.flag_base + EWDML_MAX_PEERS;        // ACK[dst]: written by dst

    if ((a.phases & PH_PUT) && a.put_mask) {
        if (tid == 0) s_ok = 1;
        __syncthreads();
        // the destination must have consumed what I stored in this slot two operations ago
        if (tid < W && ((a.put_mask >> tid) & 1u) && a.ack_need[tid] != 0)
            if (!wait_flag_ge(my_ack + tid, a.ack_need[tid], a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT)) s_ok = 0;
        __syncthreads();
        if (s_ok) {
            // four independent 16-byte loads in flight per thread before the first store (the copy is latency bound otherwise)
            if (a.src_stride == 0) {
                for (int64_t i0 = gtid; i0 < n16; i0 += 4 * gsize) {
                    uint4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int64_t i = i0 + u * gsize;
                        if (i < n16) v[u] = ld_v4(a.src + (i << 4));
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int64_t i = i0 + u * gsize;
                        if (i < n16) {
#pragma unroll
                            for (int p = 0; p < EWDML_MAX_PEERS; ++p)
                                if (p < W && ((a.put_mask >> p) & 1u)) st_na_v4(slot_ptr(a, p, me) + kHeaderBytes + (i << 4), v[u]);
                        }
                    }
                }
            } else {
                for (int p = 0; p < W; ++p) {
                    if (!((a.put_mask >> p) & 1u)) continue;
                    const uint8_t* s = a.src + (int64_t)p * a.src_stride;
                    uint8_t* d = slot_ptr(a, p, me) + kHeaderBytes;
                    copy16x4(d, s, n16, gtid, gsize);
                }
            }
        }
        if (elect_last(a.counters + 0, &s_last)) {
            if (tid < W && ((a.put_mask >> tid) & 1u)) {
                uint8_t* h = slot_ptr(a, tid, me);
                st_na_v4(h, make_uint4(a.seq, a.op_tag, (uint32_t)a.nbytes, (uint32_t)(a.nbytes >> 32)));
                __threadfence_system();
                st_release_sys(a.peer_flags[tid] + a.flag_base + me, a.seq);
            }
            __syncthreads();
            if (tid == 0) a.counters[0] = 0;
        }
    }

    if ((a.phases & PH_WAIT) && a.src_mask) {
        if (tid == 0) s_ok = 1;
        __syncthreads();
        if (tid < W && ((a.src_mask >> tid) & 1u)) {
            if (!wait_flag_ge(my_data + tid, a.seq, a.spin_limit, a.err_word, (uint32_t)EWDML_ERR_TIMEOUT)) s_ok = 0;
            else if (blockIdx.x == 0) {
                // sequence / operation check: a rank that called a different collective, or the same one with another
                // size, is reported instead of silently mixing payloads (the debug aid of SURVEY 5.2)
                const uint4 h = ld_v4(slot_ptr(a, me, tid));
                if (h.x != a.seq || h.y != a.op_tag || h.z != (uint32_t)a.nbytes) atomicExch(a.err_word, (uint32_t)EWDML_ERR_STATE);
            }
        }
        __syncthreads();
        if (s_ok && a.consume == CONS_COPY) {
            for (int s = 0; s < W; ++s) {
                if (!((a.src_mask >> s) & 1u)) continue;
                const uint8_t* p = slot_ptr(a, me, s) + kHeaderBytes;
                uint8_t* d = a.dst + (int64_t)s * a.dst_stride;
                copy16x4(d, p, n16, gtid, gsize);
            }
        } else if (s_ok && a.consume == CONS_SUM) {
            for (int64_t i = gtid; i < n16; i += gsize) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int s = 0; s < EWDML_MAX_PEERS; ++s)
                    if (s < W && ((a.src_mask >> s) & 1u)) {
                        const uint4 u = ld_v4(slot_ptr(a, me, s) + kHeaderBytes + (i << 4));
                        acc.x += __uint_as_float(u.x); acc.y += __uint_as_float(u.y);
                        acc.z += __uint_as_float(u.z); acc.w += __uint_as_float(u.w);
                    }
                acc.x *= a.scale; acc.y *= a.scale; acc.z *= a.scale; acc.w *= a.scale;
                *reinterpret_cast<float4*>(a.dst + (i << 4)) = acc;
            }
        }
        if (elect_last(a.counters + 1, &s_last)) {
            if (tid < W && ((a.src_mask >> tid) & 1u))
                st_release_sys(a.peer_flags[tid] + a.flag_base + EWDML_MAX_PEERS + me, a.seq);
            __syncthreads();
            if (tid == 0) a.counters[1] = 0;
        }
    }
}

int sm_count_() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms > 0 ? sms : 148;
}

const char* const kOpNames[EWDML_COLL_OP_COUNT] = {"bcast", "gather", "scatter", "allgather", "alltoall", "reduce",
                                                    "reduce_scatter", "allreduce", "send", "recv", "barrier"};

}  // namespace

struct ewdml_engine {
    int rank, world;
    uint8_t* win[EWDML_MAX_PEERS];
    uint32_t* peer_flags[EWDML_MAX_PEERS];
    uint32_t* counters;            // 2 local words inside the own flag block (behind the regions)
    int64_t slot_stride, slot_payload;
    uint32_t seq;                  // collective region: one sequence for all operations
    uint32_t last_put[EWDML_MAX_PEERS][2];
    uint32_t send_seq[EWDML_MAX_PEERS], recv_seq[EWDML_MAX_PEERS];
    uint32_t p2p_last_put[EWDML_MAX_PEERS][2];
    int forced_algo[EWDML_COLL_OP_COUNT];
    int64_t oneshot_max_bytes;
    int64_t calls[EWDML_COLL_OP_COUNT];
    int sms;
};

extern "C" int64_t ewdml_engine_window_bytes(int world, int64_t slot_payload) {
    if (world < 1 || world > EWDML_MAX_PEERS || slot_payload < 16) return -1;
    const int64_t stride = round_up_i64(slot_payload, 256) + 256;      // 16-byte header, slots stay 256-byte aligned
    return 2 /*regions*/ * 2 /*parities*/ * (int64_t)world * stride;
}
extern "C" int64_t ewdml_engine_flag_bytes(void) { return (2 * kFlagWordsPerRegion + 16) * 4; }

extern "C" int ewdml_engine_create(ewdml_engine** out, int rank, int world, void* const* win, uint32_t* const* flags,
                                   int64_t slot_payload) {
    EWDML_CHECK_ARG(out && win && flags && world >= 1 && world <= EWDML_MAX_PEERS && rank >= 0 && rank < world &&
                    slot_payload >= 16 && slot_payload % 16 == 0);
    ewdml_engine* e = new ewdml_engine();
    memset(e, 0, sizeof(*e));
    e->rank = rank; e->world = world;
    for (int r = 0; r < world; ++r) {
        if (!win[r] || !flags[r]) { delete e; ewdml_set_error("ewdml_engine_create: null window/flag pointer for rank %d", r); return EWDML_ERR_ARG; }
        e->win[r] = (uint8_t*)win[r];
        e->peer_flags[r] = flags[r];
    }
    e->slot_payload = round_up_i64(slot_payload, 256);
    e->slot_stride = e->slot_payload + 256;
    e->counters = flags[rank] + 2 * kFlagWordsPerRegion;
    e->sms = sm_count_();
    e->oneshot_max_bytes = 1 << 18;
    if (const char* v = getenv("EWDML_COLL_ONESHOT_MAX_BYTES")) e->oneshot_max_bytes = atoll(v);
    for (int i = 0; i < EWDML_COLL_OP_COUNT; ++i) e->forced_algo[i] = EWDML_ALGO_AUTO;
    // forced-algorithm variable, the analogue of coll_tuned_allreduce_algorithm (coll_tuned_allreduce_decision.c:37-46)
    if (const char* v = getenv("EWDML_COLL_ALLREDUCE_ALGO")) {
        if (!strcmp(v, "oneshot")) e->forced_algo[EWDML_COLL_ALLREDUCE] = EWDML_ALGO_ONESHOT;
        else if (!strcmp(v, "twoshot")) e->forced_algo[EWDML_COLL_ALLREDUCE] = EWDML_ALGO_TWOSHOT;
    }
    *out = e;
    return 0;
}

extern "C" int ewdml_engine_destroy(ewdml_engine* e) { delete e; return 0; }

extern "C" int ewdml_engine_set_algo(ewdml_engine* e, int op, int algo) {
    EWDML_CHECK_ARG(e && op >= 0 && op < EWDML_COLL_OP_COUNT && algo >= EWDML_ALGO_AUTO && algo <= EWDML_ALGO_TWOSHOT);
    e->forced_algo[op] = algo;
    return 0;
}

// decision table (coll/tuned analogue): everything but allreduce has a single algorithm on a switch
extern "C" int ewdml_engine_select_algo(const ewdml_engine* e, int op, int64_t nbytes) {
    if (!e || op < 0 || op >= EWDML_COLL_OP_COUNT) return EWDML_ERR_ARG;
    if (op != EWDML_COLL_ALLREDUCE) return EWDML_ALGO_ONESHOT;
    if (e->forced_algo[op] != EWDML_ALGO_AUTO) return e->forced_algo[op];
    if (e->world <= 2) return EWDML_ALGO_ONESHOT;                       // W x N == 2 (W-1)/W x N x 2 at W = 2
    return nbytes <= e->oneshot_max_bytes ? EWDML_ALGO_ONESHOT : EWDML_ALGO_TWOSHOT;
}

extern "C" int64_t ewdml_engine_calls(const ewdml_engine* e, int op) {
    return (e && op >= 0 && op < EWDML_COLL_OP_COUNT) ? e->calls[op] : -1;
}
extern "C" int64_t ewdml_engine_slot_payload(const ewdml_engine* e) { return e ? e->slot_payload : -1; }

// One kernel launch = one (chunk of an) operation.
static int run(ewdml_engine* e, int op, int tag, int region, uint32_t seq, uint32_t put_mask, const void* src, int64_t src_stride,
               uint32_t src_mask, void* dst, int64_t dst_stride, int consume, float scale, int64_t nbytes, int phases,
               cudaStream_t st) {
    CollArgs a;
    memset(&a, 0, sizeof(a));
    a.rank = e->rank; a.world = e->world;
    for (int r = 0; r < e->world; ++r) { a.win[r] = e->win[r]; a.peer_flags[r] = e->peer_flags[r]; }
    a.flags = e->peer_flags[e->rank];
    a.counters = e->counters;
    a.slot_stride = e->slot_stride;
    a.region_off = (int64_t)region * 2 * e->world * e->slot_stride;
    a.flag_base = region * kFlagWordsPerRegion;
    a.seq = seq; a.op_tag = (uint32_t)op | ((uint32_t)tag << 8);
    a.phases = phases;
    a.put_mask = (phases & PH_PUT) ? put_mask : 0; a.src = (const uint8_t*)src; a.src_stride = src_stride;
    a.src_mask = (phases & PH_WAIT) ? src_mask : 0; a.dst = (uint8_t*)dst; a.dst_stride = dst_stride;
    a.consume = consume; a.scale = scale; a.nbytes = nbytes;
    a.err_word = ewdml_device_error_word();
    a.spin_limit = ewdml_spin_limit();
    if (!a.err_word) return EWDML_ERR_CUDA;
    uint32_t (*lp)[2] = region == 0 ? e->last_put : e->p2p_last_put;
    for (int p = 0; p < e->world; ++p)
        if ((a.put_mask >> p) & 1u) {
            a.ack_need[p] = lp[p][seq & 1u];
            lp[p][seq & 1u] = seq;
        }
    if (!a.put_mask && !a.src_mask) return 0;
    const int64_t n16 = nbytes >> 4;
    // <= 2 CTAs per SM: every CTA of the grid is resident, so CTAs spinning in the wait phase never keep a CTA that still
    // has to put from being scheduled
    int grid = (int)std::min<int64_t>(std::max<int64_t>((n16 + 4 * 512 - 1) / (4 * 512), 1), 2 * (int64_t)e->sms);
    k_coll<<<grid, 512, 0, st>>>(a);
    EWDML_LAUNCH_CHECK();
    int64_t out_bytes = 0;
    for (int p = 0; p < e->world; ++p)
        if (((a.put_mask >> p) & 1u) && p != e->rank) { out_bytes += nbytes; ewdml_peer_matrix_add(p, nbytes); }
    if (out_bytes) ewdml_counter_add(EWDML_CTR_BYTES_SENT, out_bytes);
    return 0;
}

static inline uint32_t all_mask(int world) { return world >= 32 ? 0xFFFFFFFFu : ((1u << world) - 1u); }

// Splits a per-peer message of `n` bytes into slot-sized chunks; each chunk is one operation number.
extern "C" int ewdml_coll(ewdml_engine* e, int op, const void* src, void* dst, int64_t n, int root, float scale, int phases,
                          void* stream) {
    EWDML_CHECK_ARG(e && op >= 0 && op < EWDML_COLL_OP_COUNT && op != EWDML_COLL_SEND && op != EWDML_COLL_RECV);
    EWDML_CHECK_ARG(n >= 0 && n % 16 == 0 && (phases & ~3) == 0 && phases != 0);
    EWDML_CHECK_ARG(root >= 0 && root < e->world);
    EWDML_CHECK_ARG(op == EWDML_COLL_BARRIER || (n > 0 && dst && ((uintptr_t)dst & 15) == 0 && ((uintptr_t)src & 15) == 0));
    EWDML_CHECK_ARG(op == EWDML_COLL_BARRIER || op == EWDML_COLL_BCAST || src);
    const int me = e->rank, W = e->world;
    const uint32_t all = all_mask(W), rootbit = 1u << root, others = all & ~rootbit;
    const bool split = phases != (PH_PUT | PH_WAIT);
    if (split && n > e->slot_payload) {
        ewdml_set_error("ewdml_coll: split put/wait phases (single-device emulation) need n <= slot payload (%lld)",
                        (long long)e->slot_payload);
        return EWDML_ERR_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (op == EWDML_COLL_ALLREDUCE) {
        const int algo = ewdml_engine_select_algo(e, op, n);
        if (algo == EWDML_ALGO_TWOSHOT && !split && n % (16 * W) == 0) {
            // reduce_scatter into my shard of dst, then allgather the shards in place
            const int64_t shard = n / W;
            for (int64_t off = 0; off < shard; off += e->slot_payload) {
                const int64_t len = std::min<int64_t>(e->slot_payload, shard - off);
                uint32_t q = ++e->seq;
                int rc = run(e, EWDML_COLL_REDUCE_SCATTER, 0, 0, q, all, (const uint8_t*)src + off, shard, all,
                             (uint8_t*)dst + (int64_t)me * shard + off, 0, CONS_SUM, scale, len, phases, st);
                if (rc) return rc;
                q = ++e->seq;
                rc = run(e, EWDML_COLL_ALLGATHER, 0, 0, q, all, (const uint8_t*)dst + (int64_t)me * shard + off, 0, all,
                         (uint8_t*)dst + off, shard, CONS_COPY, 1.f, len, phases, st);
                if (rc) return rc;
            }
            e->calls[op]++;
            ewdml_counter_add(EWDML_CTR_ALLREDUCE_CALLS, 1);
            return 0;
        }
    }
    const bool begin = (phases & PH_PUT) != 0;       // a wait-only call completes the operation the put-only call began
    int64_t off = 0;
    do {
        const int64_t len = std::min<int64_t>(e->slot_payload, n - off);
        const uint32_t q = begin ? ++e->seq : e->seq;
        const uint8_t* s = (const uint8_t*)src + off;
        uint8_t* d = (uint8_t*)dst + off;
        int rc = 0;
        switch (op) {
        case EWDML_COLL_BCAST:       // buffer `dst` is the root's source and everybody else's destination
            rc = run(e, op, root, 0, q, me == root ? others : 0, d, 0, me == root ? 0 : rootbit, d, 0, CONS_COPY, 1.f, len, phases, st);
            // CONS_COPY writes dst + s * dst_stride with s = root and stride 0
            break;
        case EWDML_COLL_GATHER:
            rc = run(e, op, root, 0, q, rootbit, s, 0, me == root ? all : 0, d, n, CONS_COPY, 1.f, len, phases, st);
            break;
        case EWDML_COLL_SCATTER:
            rc = run(e, op, root, 0, q, me == root ? all : 0, s, n, rootbit, d, 0, CONS_COPY, 1.f, len, phases, st);
            break;
        case EWDML_COLL_ALLGATHER:
            rc = run(e, op, 0, 0, q, all, s, 0, all, d, n, CONS_COPY, 1.f, len, phases, st);
            break;
        case EWDML_COLL_ALLTOALL:
            rc = run(e, op, 0, 0, q, all, s, n, all, d, n, CONS_COPY, 1.f, len, phases, st);
            break;
        case EWDML_COLL_REDUCE:
            rc = run(e, op, root, 0, q, rootbit, s, 0, me == root ? all : 0, d, 0, CONS_SUM, scale, len, phases, st);
            break;
        case EWDML_COLL_REDUCE_SCATTER:
            rc = run(e, op, 0, 0, q, all, s, n, all, d, 0, CONS_SUM, scale, len, phases, st);
            break;
        case EWDML_COLL_ALLREDUCE:   // one-shot
            rc = run(e, op, 0, 0, q, all, s, 0, all, d, 0, CONS_SUM, scale, len, phases, st);
            break;
        case EWDML_COLL_BARRIER:
            rc = run(e, op, 0, 0, q, all, nullptr, 0, all, nullptr, 0, CONS_NONE, 1.f, 0, phases, st);
            break;
        default:
            return EWDML_ERR_ARG;
        }
        if (rc) return rc;
        off += len;
    } while (off < n);
    if (begin) {
        e->calls[op]++;
        if (op == EWDML_COLL_ALLREDUCE) ewdml_counter_add(EWDML_CTR_ALLREDUCE_CALLS, 1);
    }
    return 0;
}

// Tagged point-to-point: messages between one (sender, receiver) pair match in order; the tag travels in the slot
// header and a mismatch sets the device error word (EWDML_ERR_STATE).
extern "C" int ewdml_coll_send(ewdml_engine* e, const void* src, int64_t n, int peer, int tag, void* stream) {
    EWDML_CHECK_ARG(e && src && ((uintptr_t)src & 15) == 0 && n > 0 && n % 16 == 0 && peer >= 0 && peer < e->world && peer != e->rank && tag >= 0 && tag < (1 << 24));
    for (int64_t off = 0; off < n; off += e->slot_payload) {
        const int64_t len = std::min<int64_t>(e->slot_payload, n - off);
        const uint32_t q = ++e->send_seq[peer];
        int rc = run(e, EWDML_COLL_SEND, tag, 1, q, 1u << peer, (const uint8_t*)src + off, 0, 0, nullptr, 0, CONS_NONE, 1.f, len,
                     PH_PUT, (cudaStream_t)stream);
        if (rc) return rc;
    }
    e->calls[EWDML_COLL_SEND]++;
    ewdml_counter_add(EWDML_CTR_SEND_CALLS, 1);
    return 0;
}

extern "C" int ewdml_coll_recv(ewdml_engine* e, void* dst, int64_t n, int peer, int tag, void* stream) {
    EWDML_CHECK_ARG(e && dst && ((uintptr_t)dst & 15) == 0 && n > 0 && n % 16 == 0 && peer >= 0 && peer < e->world && peer != e->rank && tag >= 0 && tag < (1 << 24));
    for (int64_t off = 0; off < n; off += e->slot_payload) {
        const int64_t len = std::min<int64_t>(e->slot_payload, n - off);
        const uint32_t q = ++e->recv_seq[peer];
        // the header was written with op = SEND by the peer
        int rc = run(e, EWDML_COLL_SEND, tag, 1, q, 0, nullptr, 0, 1u << peer, (uint8_t*)dst + off, 0, CONS_COPY, 1.f,
                     len, PH_WAIT, (cudaStream_t)stream);
        if (rc) return rc;
    }
    e->calls[EWDML_COLL_RECV]++;
    ewdml_counter_add(EWDML_CTR_RECV_CALLS, 1);
    ewdml_counter_add(EWDML_CTR_BYTES_RECV, n);
    return 0;
}

// MPI_Sendrecv: pieces of the outgoing and incoming message alternate, so a ring shift of messages longer than the
// two slots a pair owns cannot dead-lock (send alone has rendezvous semantics beyond 2 x slot payload: its third piece
// waits until the receiver has consumed the first).
extern "C" int ewdml_coll_sendrecv(ewdml_engine* e, const void* src, int64_t n_send, int to, void* dst, int64_t n_recv, int from,
                                   int tag, void* stream) {
    EWDML_CHECK_ARG(e && src && dst && ((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0 && n_send > 0 && n_recv > 0 &&
                    n_send % 16 == 0 && n_recv % 16 == 0 && to >= 0 && to < e->world && from >= 0 && from < e->world &&
                    to != e->rank && from != e->rank && tag >= 0 && tag < (1 << 24));
    int64_t so = 0, ro = 0;
    while (so < n_send || ro < n_recv) {
        if (so < n_send) {
            const int64_t len = std::min<int64_t>(e->slot_payload, n_send - so);
            int rc = run(e, EWDML_COLL_SEND, tag, 1, ++e->send_seq[to], 1u << to, (const uint8_t*)src + so, 0, 0, nullptr, 0, CONS_NONE,
                         1.f, len, PH_PUT, (cudaStream_t)stream);
            if (rc) return rc;
            so += len;
        }
        if (ro < n_recv) {
            const int64_t len = std::min<int64_t>(e->slot_payload, n_recv - ro);
            int rc = run(e, EWDML_COLL_SEND, tag, 1, ++e->recv_seq[from], 0, nullptr, 0, 1u << from, (uint8_t*)dst + ro, 0, CONS_COPY, 1.f,
                         len, PH_WAIT, (cudaStream_t)stream);
            if (rc) return rc;
            ro += len;
        }
    }
    e->calls[EWDML_COLL_SEND]++; e->calls[EWDML_COLL_RECV]++;
    ewdml_counter_add(EWDML_CTR_SEND_CALLS, 1);
    ewdml_counter_add(EWDML_CTR_RECV_CALLS, 1);
    ewdml_counter_add(EWDML_CTR_BYTES_RECV, n_recv);
    return 0;
}

extern "C" const char* ewdml_coll_op_name(int op) { return (op >= 0 && op < EWDML_COLL_OP_COUNT) ? kOpNames[op] : "?"; }
