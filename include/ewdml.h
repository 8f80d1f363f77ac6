/* ewdml -- C ABI of the native runtime and the sm_100a kernels.
 *
 * Everything here is callable through ctypes (ewdml/runtime/lib.py) or from C/C++.  All entry
 * points return 0 on success or a negative ewdml error code / positive cudaError_t value; the text
 * of the last error on the calling thread is available from ewdml_last_error().
 *
 * Reference citations name what each group replaces in
 * AnirudhKaushik10/Efficient-Workers-in-Distributed-Machine-Learning (paths relative to its root;
 * "ompi:" = openmpi-4.1.0/).
 */
#ifndef EWDML_H_
#define EWDML_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EWDML_MAX_PEERS 16
#define EWDML_TILE 8192          /* elements per tile (one CTA) in the select/encode/apply kernels */
#define EWDML_HIST_BINS 2048     /* 11-bit radix digit                                             */

/* wire modes of a payload slot (see DESIGN.md "wire format") */
enum { EWDML_MODE_DENSE = 0, EWDML_MODE_IDX16 = 1, EWDML_MODE_BITMAP = 2 };
/* element selection */
enum { EWDML_SEL_ALL = 0, EWDML_SEL_TOPK = 1 };
/* value coding: QSGD level codes of `code_bits` bits, or raw fp32 */
enum { EWDML_VAL_QSGD = 0, EWDML_VAL_RAW32 = 1 };
/* what the receive kernel does with the averaged gradient */
enum { EWDML_OUT_SGD = 0,        /* momentum-SGD update of param/momentum (+bf16 shadow)          */
       EWDML_OUT_GRAD = 1,       /* write the dense averaged gradient                              */
       EWDML_OUT_ASSIGN = 2 };   /* param = average (weight pull in parameter-server mode)         */

enum { EWDML_ERR_ARG = -1, EWDML_ERR_CUDA = -2, EWDML_ERR_TIMEOUT = -3, EWDML_ERR_OVERFLOW = -4,
       EWDML_ERR_IO = -5, EWDML_ERR_STATE = -6 };

const char* ewdml_last_error(void);
int ewdml_version(void);

/* ------------------------------------------------------------------------------------------
 * Exchange plan: static description of one gradient group (segments = parameter tensors).
 * Replaces the per-tensor Python loops of distributed_worker.py:317-351 / :269-279 and
 * sync_replicas_master_nn.py:218-232, and the compressor objects of Compresssor/{qsgd,TopK}.py.
 * ------------------------------------------------------------------------------------------ */
typedef struct ewdml_plan ewdml_plan;

typedef struct {
    int mode;            /* EWDML_MODE_*                                                        */
    int sel;             /* EWDML_SEL_*                                                         */
    int val;             /* EWDML_VAL_*                                                         */
    int levels;          /* QSGD s (qsgd.py:9 default 128)                                      */
    int code_bits;       /* bits per transmitted code (sign + ceil(log2 s)); 32 for RAW32       */
    double topk_ratio;   /* k = max(1, (int)(numel * ratio)) per segment (TopK.py:7)            */
} ewdml_codec;

/* seg_offset/seg_numel: element offsets (multiples of 32) and sizes of each segment in the flat
 * fp32 buffer.  Allocates the device tables and the scratch (histograms, candidates, counters). */
int ewdml_plan_create(ewdml_plan** out, int nseg, const int64_t* seg_offset, const int64_t* seg_numel,
                      const ewdml_codec* codec);
int ewdml_plan_destroy(ewdml_plan* plan);
/* bytes a receive slot must provide for one sender's payload of this plan */
int64_t ewdml_plan_slot_bytes(const ewdml_plan* plan);
int64_t ewdml_plan_num_tiles(const ewdml_plan* plan);
int64_t ewdml_plan_scratch_bytes(const ewdml_plan* plan);
/* byte offsets of the regions inside a slot: [0]=norms [1]=tile directory [2]=bitmap [3]=data */
int ewdml_plan_slot_layout(const ewdml_plan* plan, int64_t out[4]);

typedef struct {
    int n;                                   /* number of destination slots                       */
    void* slot[EWDML_MAX_PEERS];             /* peer-mapped (or local) slot base addresses        */
    uint32_t* flag[EWDML_MAX_PEERS];         /* peer-mapped flag words, set to `seq` when done    */
    void* mc_slot;                           /* NVLS: multicast address of the same slot in EVERY rank's mailbox (or NULL).
                                              * When set, the payload is written once through it (multimem.st, replicated
                                              * by the NVSwitch) instead of once per slot[]; the flags are still published
                                              * per destination through flag[].  Zero-initialise the struct.          */
} ewdml_dests;

typedef struct {
    int n;                                   /* number of source slots (local memory)             */
    const void* slot[EWDML_MAX_PEERS];
    const uint32_t* flag[EWDML_MAX_PEERS];   /* local flag words to wait on (>= seq); may be NULL */
} ewdml_sources;

/* K1: select (top-k radix select) + QSGD quantise + bit-pack + store into every destination slot
 * and publish the per-destination flag.  `grad` is this rank's flat gradient.  `rng_stream`
 * separates the random draws of different senders (use the rank).  Enqueues on `stream`.        */
int ewdml_send(ewdml_plan* plan, const float* grad, const ewdml_dests* dests, uint32_t seq,
               uint64_t seed, uint32_t step, uint32_t rng_stream, void* stream);

typedef struct {
    float lr, momentum, dampening, weight_decay;
    int nesterov;
    int first_step;       /* 1: momentum buffer := gradient (optim/sgd.py:79-82)                 */
} ewdml_sgd;

/* K2: wait for the sources' flags, unpack + dequantise each payload with its own norm,
 * accumulate, divide by `divisor`, then per `out_mode` update param/momentum (and the optional
 * bf16 shadow copy) or write the dense average to `out`.                                        */
int ewdml_recv_apply(ewdml_plan* plan, const ewdml_sources* srcs, uint32_t seq, float divisor,
                     int out_mode, const ewdml_sgd* sgd, float* param, float* momentum,
                     void* param_bf16, float* out, void* stream);

/* Local fused momentum-SGD on flat buffers (optim/sgd.py:59-92), no exchange.                   */
int ewdml_sgd_apply(const float* grad, float* param, float* momentum, void* param_bf16, int64_t n,
                    const ewdml_sgd* sgd, float grad_scale, void* stream);

/* {entries, data bytes, total payload bytes, error word} of the plan's last ewdml_send; syncs the
 * stream.                                                                                        */
int ewdml_plan_read_stats(ewdml_plan* plan, int64_t out[4], void* stream);
/* Debug/test access to the per-segment selection results of the last send:
 * out[4*i+0]=threshold key, +1=ties taken, +2=norm bits, +3=candidate count                     */
int ewdml_plan_read_seginfo(ewdml_plan* plan, uint32_t* out, void* stream);
/* running count of segments that needed the exact-scan fallback of the top-k selection (sampled band missed); syncs */
int ewdml_plan_fallbacks(ewdml_plan* plan, int64_t* out, void* stream);

/* Graph-replayable forms.  `state` is a device array of 4 uint32 {seq, step, first, applies} advanced by
 * ewdml_state_advance (one tiny launch at the start of every optimizer application; bump_seq = 1 for
 * exchanges).  Sequence number, RNG step, slot parity and the first-step flag are read from it on the
 * device, so a CUDA graph that captured these launches can be replayed every step.  *_even / *_odd are the
 * slot sets used when seq is even / odd.                                                               */
int ewdml_state_advance(uint32_t* state, int bump_seq, void* stream);
int ewdml_send_state(ewdml_plan* plan, const float* grad, const ewdml_dests* dests_even,
                     const ewdml_dests* dests_odd, const uint32_t* state, uint64_t seed, uint32_t rng_stream,
                     void* stream);
int ewdml_recv_apply_state(ewdml_plan* plan, const ewdml_sources* srcs_even, const ewdml_sources* srcs_odd,
                           const uint32_t* state, float divisor, int out_mode, const ewdml_sgd* sgd,
                           float* param, float* momentum, void* param_bf16, float* out, void* stream);
int ewdml_sgd_apply_state(const float* grad, float* param, float* momentum, void* param_bf16, int64_t n,
                          const ewdml_sgd* sgd, float grad_scale, const uint32_t* state, void* stream);

/* ------------------------------------------------------------------------------------------
 * K4: dense fused-scale allreduce over peer memory (replaces hvd.allreduce(Average) and
 * ompi:ompi/mca/coll/base/coll_base_allreduce.c:341 + ompi:ompi/mca/op/avx/op_avx_functions.c).
 * `bufs[r]` is rank r's buffer as mapped in THIS process; `flags` is this rank's local flag block
 * (>= 4*EWDML_MAX_PEERS words) and `peer_flags[r]` the same block of rank r mapped here.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int rank, world;
    float* bufs[EWDML_MAX_PEERS];
    uint32_t* peer_flags[EWDML_MAX_PEERS];
    uint32_t* flags;
    float* mc_buf;     /* NVLS: multicast address of the buffer (all ranks' copies), or NULL.  Zero-initialise the struct. */
} ewdml_comm;

/* NVLS = in-switch reduction: rank r reads its shard with multimem.ld_reduce.add.f32 (the NVSwitch sums the W copies)
 * and writes the scaled result to every rank with one multimem.st; needs comm->mc_buf.  AUTO picks it when available
 * and the message is above the one-shot size.                                                                        */
enum { EWDML_AR_AUTO = 0, EWDML_AR_ONESHOT = 1, EWDML_AR_TWOSHOT = 2, EWDML_AR_NVLS = 3 };
/* in-place: every rank ends with scale * sum_r bufs[r][0..n) in its own buffer                  */
int ewdml_allreduce(const ewdml_comm* comm, int64_t n, float scale, int algo, uint32_t seq,
                    uint32_t* work_counter, void* stream);
/* device-side barrier across ranks (flag exchange)                                              */
int ewdml_allreduce_state(const ewdml_comm* comm, int64_t n, float scale, int algo, uint32_t seq,
                          const uint32_t* state, uint32_t* work_counter, void* stream);
int ewdml_barrier(const ewdml_comm* comm, uint32_t seq, void* stream);
/* Adasum tree node on the device (horvod_pytorch.py:35,200 op=hvd.Adasum): for every segment (tensor) of two flat fp32
 * buffers, out = (1 - <a,b>/(2|a|^2)) a + (1 - <a,b>/(2|b|^2)) b.  seg_off / seg_n / acc are device arrays (acc: 3*nseg
 * doubles of scratch); a and b may be peer-mapped; out may alias either.                                            */
int ewdml_adasum_pair(const float* a, const float* b, float* out, int nseg, const int64_t* seg_off,
                      const int64_t* seg_n, int64_t max_seg_numel, double* acc, void* stream);
/* device error word (0 = ok) set by kernels whose flag spin exceeded the limit                  */
int ewdml_get_device_error(uint32_t* out, void* stream);
/* poll cap of every flag wait: 4096 back-to-back polls, then ~1 us per poll; 0 = default 2^26 (about a minute) */
int ewdml_set_spin_limit(uint64_t iters);

/* ------------------------------------------------------------------------------------------
 * N3: collective engine on arbitrary device buffers (replaces the coll framework + coll/base + coll/tuned,
 * ompi:ompi/mca/coll/coll.h:142-255, ompi:ompi/mca/coll/tuned/coll_tuned_decision_fixed.c:55-217, and pml/ob1
 * tagged send/recv, ompi:ompi/mca/pml/ob1/pml_ob1_isend.c:147).  Every operation is one kernel: remote stores
 * into the destination ranks' symmetric windows + release flag, then acquire-wait + copy or ordered sum out of
 * the local window.  `win[r]` / `flags[r]` are rank r's window and flag block as mapped in this process
 * (ewdml_engine_window_bytes / ewdml_engine_flag_bytes bytes, zero-initialised).  Sizes are bytes, multiples
 * of 16; buffers 16-byte aligned; reductions are fp32 sums in rank order times `scale`.  One engine is used
 * from one stream at a time.
 * ------------------------------------------------------------------------------------------ */
typedef struct ewdml_engine ewdml_engine;
enum { EWDML_COLL_BCAST = 0, EWDML_COLL_GATHER, EWDML_COLL_SCATTER, EWDML_COLL_ALLGATHER, EWDML_COLL_ALLTOALL,
       EWDML_COLL_REDUCE, EWDML_COLL_REDUCE_SCATTER, EWDML_COLL_ALLREDUCE, EWDML_COLL_SEND, EWDML_COLL_RECV,
       EWDML_COLL_BARRIER, EWDML_COLL_OP_COUNT };
enum { EWDML_ALGO_AUTO = 0, EWDML_ALGO_ONESHOT = 1, EWDML_ALGO_TWOSHOT = 2 };
enum { EWDML_PHASE_PUT = 1, EWDML_PHASE_WAIT = 2, EWDML_PHASE_BOTH = 3 };
int64_t ewdml_engine_window_bytes(int world, int64_t slot_payload);
int64_t ewdml_engine_flag_bytes(void);
int ewdml_engine_create(ewdml_engine** out, int rank, int world, void* const* win, uint32_t* const* flags,
                        int64_t slot_payload);
int ewdml_engine_destroy(ewdml_engine* e);
/* decision layer: forced algorithm per operation (also EWDML_COLL_ALLREDUCE_ALGO=oneshot|twoshot and
 * EWDML_COLL_ONESHOT_MAX_BYTES in the environment), and the query the engine itself uses            */
int ewdml_engine_set_algo(ewdml_engine* e, int op, int algo);
int ewdml_engine_select_algo(const ewdml_engine* e, int op, int64_t nbytes);
int64_t ewdml_engine_calls(const ewdml_engine* e, int op);
int64_t ewdml_engine_slot_payload(const ewdml_engine* e);
const char* ewdml_coll_op_name(int op);
/* n = bytes per rank-sized piece: bcast/allreduce/reduce: the whole message; gather/allgather: each rank's
 * contribution (dst holds world * n at the root / everywhere); scatter/alltoall/reduce_scatter: each piece
 * (src holds world * n).  bcast uses `dst` on every rank (source at the root).  Messages larger than the slot
 * payload are pipelined in slot-sized pieces.  phases = EWDML_PHASE_BOTH in production; PUT then WAIT as two
 * calls (n <= slot payload) lets several emulated ranks share one device without kernels waiting on each other. */
int ewdml_coll(ewdml_engine* e, int op, const void* src, void* dst, int64_t n, int root, float scale, int phases,
               void* stream);
/* tagged point-to-point; messages between a pair match in order, a tag/size mismatch sets the device error word */
int ewdml_coll_send(ewdml_engine* e, const void* src, int64_t n, int peer, int tag, void* stream);
int ewdml_coll_recv(ewdml_engine* e, void* dst, int64_t n, int peer, int tag, void* stream);
/* a pair owns two slots: a send longer than 2 x slot payload completes only once the receiver consumes (rendezvous);
 * sendrecv alternates outgoing and incoming pieces so ring shifts of any length cannot dead-lock (MPI_Sendrecv) */
int ewdml_coll_sendrecv(ewdml_engine* e, const void* src, int64_t n_send, int to, void* dst, int64_t n_recv, int from,
                        int tag, void* stream);

/* ------------------------------------------------------------------------------------------
 * N1: symmetric peer heap (replaces BTL smcuda + common_cuda.c IPC paths,
 * ompi:opal/mca/common/cuda/common_cuda.c:1006-1152, and OSHMEM memheap).
 * ------------------------------------------------------------------------------------------ */
int ewdml_heap_alloc(void** base, size_t bytes);                 /* zero-initialised            */
int ewdml_heap_free(void* base);
int ewdml_heap_export(void* base, uint8_t handle[64]);           /* cudaIpcMemHandle_t          */
int ewdml_heap_import(const uint8_t handle[64], void** mapped);  /* enables peer access lazily  */
int ewdml_heap_unmap(void* mapped);
int ewdml_can_access_peer(int dev, int peer, int* out);
int ewdml_sm_count(void);                                        /* SMs of the current device, cached per device */

/* Second heap transport: CUDA virtual-memory-management allocations + NVLS multicast objects (what replaces the smcuda BTL's
 * IPC handshake, ompi:opal/mca/btl/smcuda/btl_smcuda.c:1081, and common/cuda's handle get / open,
 * ompi:opal/mca/common/cuda/common_cuda.c:1006-1152, when stores must be replicated by the NVSwitch).  Driver entry points are
 * resolved at run time (no libcuda link dependency).  Sequence per symmetric buffer: every rank ewdml_vmm_alloc -> send the fd to
 * every peer (ewdml_fdx_send) -> ewdml_vmm_import each received fd; rank 0 ewdml_mc_create -> fd to all; the others
 * ewdml_mc_import; all ewdml_mc_add_device; HOST BARRIER; all ewdml_mc_bind_map.  A store to mc_base + x is replicated into
 * offset x of every rank's allocation (multimem.st); multimem.ld_reduce from it returns the in-switch sum of all copies.        */
typedef struct ewdml_vmm_buf ewdml_vmm_buf;
typedef struct ewdml_fdx ewdml_fdx;
int ewdml_vmm_query(int out[3]);                                 /* {VMM usable, multicast usable, POSIX-fd handles usable} */
int ewdml_vmm_granularity(int ndev, size_t* out);                /* allocation size quantum (ndev > 1: also the multicast one) */
int ewdml_vmm_alloc(ewdml_vmm_buf** out, size_t bytes, void** base, int* fd);      /* zero-filled; bytes % granularity == 0 */
int ewdml_vmm_import(ewdml_vmm_buf** out, int fd, size_t bytes, void** mapped);    /* a peer's allocation, mapped read/write */
int ewdml_vmm_close_fd(int fd);
int ewdml_vmm_free(ewdml_vmm_buf* buf);                          /* unmap + release (own, imported or multicast object)   */
int ewdml_mc_create(ewdml_vmm_buf** out, int ndev, size_t bytes, int* fd);         /* root only                               */
int ewdml_mc_import(ewdml_vmm_buf** out, int fd, size_t bytes);
int ewdml_mc_add_device(ewdml_vmm_buf* mc);                      /* current device; every rank, before any bind            */
int ewdml_mc_bind_map(ewdml_vmm_buf* mc, ewdml_vmm_buf* mem, void** mc_base);
/* file-descriptor channel between the ranks of one node (abstract UNIX datagram sockets, SCM_RIGHTS, tagged, with a stash for
 * messages that arrive early; sends never block without draining the own queue)                                              */
int ewdml_fdx_open(ewdml_fdx** out, const char* job, int rank, int world);
int ewdml_fdx_send(ewdml_fdx* x, int dst, int fd, uint64_t tag, double timeout_s);
int ewdml_fdx_recv(ewdml_fdx* x, uint64_t tag, int src, int* fd, int* src_out, double timeout_s);
int ewdml_fdx_close(ewdml_fdx* x);
/* Deterministic offset allocator inside one peer-mapped region (OSHMEM memheap analogue): the same alloc/free sequence
 * gives the same offsets on every rank, so symmetric objects need no rendezvous after the region was exchanged.      */
typedef struct ewdml_symheap ewdml_symheap;
int ewdml_symheap_create(ewdml_symheap** out, int64_t bytes, int64_t align);
int64_t ewdml_symheap_alloc(ewdml_symheap* h, int64_t bytes);      /* offset, or -1 when it does not fit */
int ewdml_symheap_free(ewdml_symheap* h, int64_t offset);
int64_t ewdml_symheap_used(const ewdml_symheap* h);
int64_t ewdml_symheap_peak(const ewdml_symheap* h);
int64_t ewdml_symheap_largest_free(const ewdml_symheap* h);
int ewdml_symheap_destroy(ewdml_symheap* h);

/* N1: rendezvous key-value store over TCP (replaces the PMIx modex / ORTE oob,
 * ompi:ompi/runtime/ompi_mpi_init.c:657-724).  Rank 0 hosts the server thread.                  */
typedef struct ewdml_store ewdml_store;
int ewdml_store_create(ewdml_store** out, const char* addr, int port, int rank, int world,
                       double timeout_s);
int ewdml_store_set(ewdml_store* s, const char* key, const void* val, size_t len);
int ewdml_store_get(ewdml_store* s, const char* key, void* val, size_t cap, size_t* len); /* blocks */
int ewdml_store_barrier(ewdml_store* s, const char* name);
int ewdml_store_destroy(ewdml_store* s);

/* N4: launcher + watchdog (replaces orterun/plm/odls/errmgr: start local ranks with the
 * RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* contract, tear the job down when one exits non-zero).
 * Returns the job's exit status.                                                                 */
int ewdml_launch(int nproc, const char* master_addr, int master_port, char* const argv[],
                 double grace_s, int* failed_rank);

/* N5: counters (replaces SPC_RECORD / pml+coll monitoring, ompi:ompi/runtime/ompi_spc.h).       */
enum { EWDML_CTR_LAUNCHES = 0, EWDML_CTR_SEND_CALLS, EWDML_CTR_RECV_CALLS, EWDML_CTR_ALLREDUCE_CALLS,
       EWDML_CTR_BYTES_SENT, EWDML_CTR_BYTES_RECV, EWDML_CTR_COMPUTE_LAUNCHES, EWDML_CTR_COUNT };
int64_t ewdml_counter_get(int id);
void ewdml_counter_add(int id, int64_t v);
void ewdml_counters_reset(void);
void ewdml_peer_matrix_add(int peer, int64_t bytes);             /* per-peer byte matrix row     */
int ewdml_peer_matrix_get(int64_t out[EWDML_MAX_PEERS]);
void ewdml_set_trace(int on);                                    /* libompitrace analogue        */

/* ------------------------------------------------------------------------------------------
 * Compute kernels (K5-K9).  Activations are NHWC bf16; weights are bf16 copies of the fp32 master
 * parameters in [C_out][kh][kw][C_in] order; all entry points enqueue on `stream`.
 * ------------------------------------------------------------------------------------------ */

/* K6: C[M,N] (op)= alpha * A[M,K] * B[N,K]^T (+ bias[N]) on tcgen05/TMEM with TMA-fed operands.
 * A, B bf16 with K contiguous, lda/ldb multiples of 8.  out_mode 0 = bf16 store, 1 = fp32 store,
 * 2 = fp32 atomic accumulate (split-K; weight gradients accumulate straight into the flat gradient
 * buffer).  split_k <= 0 chooses automatically.  Replaces nn.Linear / the GEMM inside nn.Conv2d
 * (lenet.py:18-21, vgg.py:24-29,53, resnet.py:19-77), which the reference runs through ATen.      */
int ewdml_gemm_bf16_tn(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                       int ldc, int out_mode, const float* bias, float alpha, int split_k, void* stream);

/* General form: a_mn / b_mn = 1 means the operand is stored with its M / N dimension contiguous
 * ([K][M] / [K][N]), so dgrad (dy W) and wgrad (dy^T x) need no transposed copies.                     */
int ewdml_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                    int a_mn, int b_mn, int out_mode, const float* bias, float alpha, int split_k, void* stream);

/* K5 operand builders (nn.Conv2d, vgg.py:53, resnet.py:19-55): NHWC bf16 x -> col[M][Kp], Kp = round_up(KH*KW*C, 8);
 * and the data-gradient gather dcol -> dx.                                                              */
int ewdml_im2col(const void* x, void* col, int N, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream);
/* Implicit-GEMM 3x3 / stride 1 / pad 1 convolution (no im2col buffer): activations in the zero-bordered layout
 * [N][H+2][W+2][C] built by ewdml_pad_nhwc; pass 1 = fprop (A = padded x, B = W, C = y compact), pass 2 = dgrad
 * (A = padded dy, B = W, C = dx compact), pass 3 = wgrad (A = padded dy, B = padded x, C = dW fp32 accumulated).
 * C_in, C_out multiples of 64.  Every operand tile is a plain 2-D TMA box shifted by a per-tap row offset.           */
int ewdml_conv3x3_igemm(int pass, const void* A, const void* B, void* C, int N, int H, int W, int Cin, int Cout,
                        int out_mode, const float* bias, int split_k, void* stream);
/* First-layer convolution in direct form (C_in <= 4: an im2col/GEMM tile would be almost empty).  x is NHWC fp32 or bf16. */
int ewdml_conv_small_fwd(const void* x, int x_is_f32, const void* w, const float* bias, void* y, int N, int H, int W,
                         int Cin, int Cout, int KH, int KW, int stride, int pad, void* stream);
int ewdml_conv_small_wgrad(const void* x, int x_is_f32, const void* dy, float* dw, int N, int H, int W, int Cin,
                           int Cout, int KH, int KW, int stride, int pad, void* stream);
/* direct small-channel convolutions (C_out % 8 != 0: LeNet's 1 -> 20 -> 50 channel 5x5 layers, lenet.py:18-19): NHWC bf16
 * activations, [C_out][KH][KW][C_in] bf16 weights, fp32 accumulation; wgrad ACCUMULATES into the fp32 gradient            */
int ewdml_conv_direct_fwd(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int Cin, int Cout,
                          int KH, int KW, int stride, int pad, void* stream);
int ewdml_conv_direct_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                            int stride, int pad, void* stream);
int ewdml_conv_direct_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                            int stride, int pad, void* stream);
int ewdml_pad_nhwc(const void* x, void* out, int N, int H, int W, int C, void* stream);
int ewdml_col2im(const void* dcol, void* dx, int N, int H, int W, int C, int KH, int KW, int stride, int pad, void* stream);

/* K7/K8 (nn.BatchNorm2d train mode, ReLU, residual add, MaxPool2d(2,2); vgg.py:51-55, resnet.py:20-63):
 * colstats: per-channel sum (and sum of squares) of a [M][C] bf16 matrix into zeroed fp32 sums[2C];
 * bn_finalize: stats = [mean | rstd | scale | shift] + running-stat update; bn_act_fwd: out = [pool2x2]
 * relu?(scale*y + shift (+res)); bn_act_bwd: dy (and dres), dgamma/dbeta accumulated in fp32.             */
int ewdml_colstats(const void* y, float* sums, int64_t M, int C, int want_sq, void* stream);
int ewdml_bn_finalize(const float* sums, const float* gamma, const float* beta, float* running_mean,
                      float* running_var, float* stats, int C, int64_t M, float momentum, float eps, void* stream);
int ewdml_bn_finalize2(const float* sums, const float* gamma, const float* beta, float* running_mean,
                       float* running_var, float* stats, int C, int64_t M, float momentum, float eps,
                       int64_t* num_batches_tracked, void* stream);
/* want_sq = 2 in ewdml_colstats gathers the sums relative to a per-channel shift K = y[0][:] (sum(y-K), sum((y-K)^2):
 * no cancellation in the variance when |mean| >> std; needs C/8 a power of two <= 256); finalize those with
 * ewdml_bn_finalize3(sums, shift_row = y, ...).  shift_row == NULL makes finalize3 identical to finalize2.            */
int ewdml_bn_finalize3(const float* sums, const void* shift_row, const float* gamma, const float* beta, float* running_mean,
                       float* running_var, float* stats, int C, int64_t M, float momentum, float eps,
                       int64_t* num_batches_tracked, void* stream);
/* One-launch (cooperative, grid barrier inside) training BatchNorm for un-pooled blocks with C/8 a power of two <= 256:
 * forward = statistics + finalize + normalise [+res] [+ReLU]; backward = reduce + apply (+ dgamma/dbeta).  `sums` holds
 * 2*C + 32 zeroed floats (sums[2*C] is the barrier word); have_sums = 1 means sums[0..2C) are finished un-shifted sums
 * (split-K epilogue) and no statistics phase runs.  EWDML_ERR_STATE = shape not covered, nothing launched.            */
int ewdml_bn_fwd_fused(const void* y, float* sums, int have_sums, const float* gamma, const float* beta, float* running_mean,
                       float* running_var, int64_t* num_batches_tracked, float* stats, const void* res, void* out, int N,
                       int H, int W, int C, int relu, float momentum, float eps, void* stream);
int ewdml_bn_bwd_fused(const void* y, const float* stats, const void* res, const void* dout, float* sums, void* dy, void* dres,
                       float* dgamma, float* dbeta, int N, int H, int W, int C, int relu, void* stream);
/* split-K epilogue: bf16 dst = fp32 src + bias, optionally accumulating BatchNorm column statistics in the same pass */
int ewdml_f32_to_bf16_stats(const float* src, const float* bias, void* dst, float* sums, int64_t M, int C, void* stream);
int ewdml_bn_act_fwd(const void* y, const float* stats, const void* res, void* out, int N, int H, int W, int C,
                     int relu, int pool, void* stream);
int ewdml_bn_act_bwd(const void* y, const float* stats, const void* res, const void* dout, float* sums, void* dy,
                     void* dres, float* dgamma, float* dbeta, int N, int H, int W, int C, int relu, int pool,
                     void* stream);
/* No counterpart in the reference.  BatchNorm batch statistics for callers that must not allocate (a C/C++
 * host loop without torch's caching allocator; a process whose GPU memory is already fully reserved): same
 * result as ewdml_colstats + ewdml_bn_finalize, but the 2*C-float scratch is a device array built into the
 * library (32 KiB per GPU, held until exit).  C <= 4096 or EWDML_ERR_ARG.  The array is shared by all calls:
 * it is cleared at the start of each call, and calls that overlap on the GPU race.  The first call on a
 * device also resolves the array's address, so make it before any stream capture on that device.        */
int ewdml_bn_stats_static(const void* y, const float* gamma, const float* beta, float* running_mean,
                          float* running_var, float* stats, int C, int64_t M, float momentum, float eps,
                          void* stream);
int ewdml_add_vec_f32(const float* src, float* dst, int n, void* stream);
int ewdml_cast_f32_bf16(const float* x, void* y, int64_t n, void* stream);
int ewdml_pad_rows_bf16(const void* src, void* dst, int rows, int cols, int pitch, void* stream);
int ewdml_relu_fwd(const void* x, void* y, int64_t n, void* stream);
int ewdml_relu_bwd(const void* dy, const void* y, void* dx, int64_t n, void* stream);
/* nn.Dropout (vgg.py:23-26): inverted dropout, counter-based mask; step_state (optional) = device step state */
int ewdml_dropout(const void* x, void* y, int64_t n, float p, uint32_t seed, uint32_t offset,
                  const uint32_t* step_state, void* stream);
/* F.avg_pool2d(out, 4) on the final 4x4 map (resnet.py:93)                                               */
int ewdml_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int ewdml_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
/* K9: softmax cross-entropy, mean over batch (distributed_worker.py:103): loss_acc[2] = {loss, #top-1 hits} */
int ewdml_xent(const void* logits, const int64_t* target, float* loss_acc, void* dlogits, int B, int C, int ld,
               void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EWDML_H_ */
