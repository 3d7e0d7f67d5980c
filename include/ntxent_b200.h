/*
 * ntxent_b200.h -- C-ABI of the B200-native NT-Xent loss engine.
 *
 * This is the drop-in boundary for the reference's NT-Xent hot path
 * (sanowl/CUDA-NT-Xent-MPI-NCCL-SimCLR).  Plain pointers and sizes only: no
 * torch types, no C++ types, nothing allocated inside.  Every pointer marked
 * "device" is CUDA device memory on the CURRENT device; every call is
 * asynchronous on `stream` and never synchronises the host.
 *
 * Conventions
 *   z        [M, D] row-major, M = 2N stacked views: rows 0..N-1 are view 1,
 *            rows N..2N-1 view 2; the positive of row i is (i + N) mod M.
 *            (The reference takes z[B, D] and duplicates it itself,
 *            src/ntxent_kernel.cu:146-148,161; here the caller stacks.)
 *   dtype    storage type of z and dz: NTXENT_F32 or NTXENT_BF16.
 *   operands the similarity contraction runs on bf16 operands with fp32
 *            accumulation when dtype is BF16 or NTXENT_FLAG_MIXED_PRECISION is
 *            set (the reference's `use_mixed_precision`, include/ntxent_kernel.cuh:34,51,
 *            which it never reads); otherwise on fp32 operands: as tf32 on the
 *            tensor cores when M >= 1024, D is 64 or 128 and tau is not too
 *            cold (log2(e)/(tau sqrt(D)) <= 1.7, i.e. tau >= 0.075 at D = 128;
 *            loss within 1e-5, dz within 1e-4 of fp64), else in exact fp32 FMA
 *            arithmetic (also what NTXENT_FLAG_FORCE_SIMT selects).
 *   return   0 on success; negative NTXENT_ERR_* for argument errors; positive
 *            values are cudaError_t codes.  Nothing throws.
 */
#ifndef NTXENT_B200_H
#define NTXENT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same type as cudaStream_t, spelled so that this header needs no CUDA include. */
typedef struct CUstream_st* ntxent_stream_t;

enum { NTXENT_F32 = 0, NTXENT_BF16 = 1 };

#define NTXENT_FLAG_MIXED_PRECISION 1u /* bf16 operands even for fp32 z                         */
#define NTXENT_FLAG_FORCE_SIMT      2u /* run the fp32-FMA tile kernels instead of tcgen05      */
#define NTXENT_FLAG_ONLINE_MAX      4u /* force the running-max logsumexp (default: automatic)  */
#define NTXENT_FLAG_CROSS_VIEW      8u /* the denominators run over the OTHER view only: row i of
                                          view 1 (rows 0..N-1) against the N rows of view 2 and the
                                          other way round.  With z = [q; k] this is the symmetric
                                          two-tower InfoNCE (CLIP):  (L(q->k) + L(k->q)) / 2  -- the
                                          SURVEY.md section 8 f4 "Q != K" variant; no counterpart in
                                          the reference.  Single-device entry points only (forward,
                                          backward, forward_backward and their _pos / _tau / _static
                                          forms); tcgen05 tiles when N is a multiple of 128, the
                                          fp32-FMA tiles otherwise                                */

#define NTXENT_OK                    0
#define NTXENT_ERR_INVALID_ARGUMENT (-1)
#define NTXENT_ERR_UNSUPPORTED_DEVICE (-2)
#define NTXENT_ERR_WORKSPACE_TOO_SMALL (-3)
#define NTXENT_ERR_UNSUPPORTED_SHAPE (-4)
#define NTXENT_ERR_STATIC_SCRATCH_EXCEEDED (-5)
#define NTXENT_ERR_COLLECTIVE (-6)      /* NCCL missing, or a NCCL call failed (NTXENT_DEBUG=1 names it) */
#define NTXENT_ERR_NO_PEER_ACCESS (-7)  /* the devices of a communicator cannot map each other's memory */

/* Backward formulations of the row-partitioned path. */
enum {
    NTXENT_BWD_SYMMETRIC = 0, /* dU_I = (G + G^T)_{I,:} U / tau; needs l for ALL rows        */
    NTXENT_BWD_CONTRACT  = 1  /* row term G_{I,:} U and column term G_{I,:}^T U_I, the latter
                                 to be reduce-scattered; needs l for rows I only             */
};

/* Replaces ntxent::utils::check_tensor_core_support (include/ntxent_kernel.cuh:99-110,
 * exported to Python at src/binding_new.cpp:19).  1 if the current device is sm_100. */
int ntxent_b200_device_supported(void);

/* Message for a return code of this library (static storage). */
const char* ntxent_b200_strerror(int code);

/* Library version string, e.g. "ntxent_b200 0.1 (sm_100a)". */
const char* ntxent_b200_version(void);

/* Bytes of device scratch the single-device entry points below need for an
 * [M, D] problem.  The reference allocates its temporaries per call through
 * ATen (src/ntxent_kernel.cu:154-158,217-218); here the caller owns them. */
size_t ntxent_b200_workspace_bytes(int64_t M, int64_t D, int dtype, unsigned flags);

/* Forward: replaces ntxent::ntxent_forward_cuda (include/ntxent_kernel.cuh:31-35,
 * src/ntxent_kernel.cu:138-203).  loss_out: device float[1] (mean over the M
 * rows).  lse_out: device float[M], the row log-sum-exp l_i that backward
 * consumes in place of the reference's [2B,2B] softmax tensor. */
int ntxent_b200_forward(const void* z, int64_t M, int64_t D, float tau, int dtype, unsigned flags,
                        void* workspace, size_t workspace_bytes,
                        float* loss_out, float* lse_out, ntxent_stream_t stream);

/* Backward: replaces ntxent::ntxent_backward_cuda (include/ntxent_kernel.cuh:46-52,
 * src/ntxent_kernel.cu:205-239).  lse: device float[M] from forward.
 * grad_out: device float[1] (NULL means 1).  dz_out: device [M, D] in `dtype`.
 * pos_prob_out: device float[M] or NULL; receives P_{i,pos(i)}, the small
 * stand-in for the reference's second return value ("grad_internal",
 * include/ntxent_kernel.cuh:43). */
int ntxent_b200_backward(const void* z, const float* lse, const float* grad_out,
                         int64_t M, int64_t D, float tau, int dtype, unsigned flags,
                         void* workspace, size_t workspace_bytes,
                         void* dz_out, float* pos_prob_out, ntxent_stream_t stream);

/* Backward that also yields the gradient of the loss with respect to the
 * temperature (learnable tau; no counterpart in the
 * reference, whose T is a plain double).  tau_grad_rows_out: device float[M];
 * d loss / d tau = (*grad_out) * sum_i tau_grad_rows_out[i]  (the caller sums:
 * M floats, in whatever order it wants to be deterministic in).  It costs
 * nothing extra: sum_a dU_a . u_a = -2 tau dL/dtau, and that dot product is
 * already formed for the normalisation Jacobian. */
int ntxent_b200_backward_tau(const void* z, const float* lse, const float* grad_out, int64_t M,
                             int64_t D, float tau, int dtype, unsigned flags, void* workspace,
                             size_t workspace_bytes, void* dz_out, float* tau_grad_rows_out,
                             ntxent_stream_t stream);

/* Forward and backward in one call (one normalisation pass); the benchmark's
 * timed unit (src/benchmark.cpp:30-39 times the forward only). */
int ntxent_b200_forward_backward(const void* z, const float* grad_out,
                                 int64_t M, int64_t D, float tau, int dtype, unsigned flags,
                                 void* workspace, size_t workspace_bytes,
                                 float* loss_out, float* lse_out, void* dz_out,
                                 float* pos_prob_out, ntxent_stream_t stream);

/* Positives from an index vector (supervised pairs, any
 * pairing of the rows; no counterpart in the reference, which hard-wires the
 * positive, src/ntxent_kernel.cu:117).  pos[i] (device int32[M], != i) is the
 * positive of row i; everything else is as in ntxent_b200_forward_backward: the
 * denominators still run over all j != i.  The gradient needs the inverse
 * relation too -- inv_idx[inv_ptr[a] .. inv_ptr[a+1]) (device int32[M+1] /
 * int32[M]) lists the rows b with pos[b] == a, in any fixed order -- because
 * row a's gradient holds  -u_pos[a] - sum of those u_b.  The default positives
 * are the involution pos[i] = (i + M/2) mod M, for which both are u_pos[i].
 * dz_out == NULL: forward only (inv_* may then be NULL).  M need not split
 * into two views here, but must still be even. */
int ntxent_b200_forward_backward_pos(const void* z, const int32_t* pos, const int32_t* inv_ptr,
                                     const int32_t* inv_idx, const float* grad_out, int64_t M, int64_t D,
                                     float tau, int dtype, unsigned flags, void* workspace,
                                     size_t workspace_bytes, float* loss_out, float* lse_out, void* dz_out,
                                     ntxent_stream_t stream);

/* ---- Row-partitioned (multi-GPU) building blocks ---------------------------
 * The reference names NCCL/MPI only as link options (CMakeLists.txt:13-14,
 * 41-47,115-121) and has no call sites; these are the pieces a data-parallel
 * caller composes around its collectives (all-gather of U, reduce-scatter of
 * the column term or all-gather of l).  Rank r owns rows
 * [row_begin, row_begin + row_count) of the global M.
 */

/* Size in bytes of one operand element for (dtype, flags): 2 (bf16) or 4 (fp32). */
int ntxent_b200_operand_size(int dtype, unsigned flags);

/* L2-normalise `rows` rows of z: u_out [rows, D] in the operand type,
 * inv_norm_out float[rows] = 1 / max(|z_i|, 1e-12). */
int ntxent_b200_normalize(const void* z, int64_t rows, int64_t D, int dtype, unsigned flags,
                          void* u_out, float* inv_norm_out, ntxent_stream_t stream);

/* Scratch bytes for the two calls below. */
size_t ntxent_b200_rows_workspace_bytes(int64_t M, int64_t D, int64_t row_count,
                                        int dtype, unsigned flags);

/* Forward for a row block against all M columns of the gathered u_all [M, D]
 * (operand type).  lse_rows_out float[row_count]; loss_partial_out float[1] =
 * sum over the block of (l_i - S_{i,pos(i)}) (NOT divided by M). */
int ntxent_b200_forward_rows(const void* u_all, int64_t M, int64_t D,
                             int64_t row_begin, int64_t row_count, float tau,
                             int dtype, unsigned flags, void* workspace, size_t workspace_bytes,
                             float* lse_rows_out, float* loss_partial_out, ntxent_stream_t stream);

/* Backward for a row block.  lse_all float[M] (CONTRACT mode reads only the
 * block's entries).  du_rows_out float[row_count, D]: SYMMETRIC -> the complete
 * dU of the block; CONTRACT -> its row term.  du_cols_out float[M, D]: CONTRACT
 * only, the block's column term for every global row (sum over ranks, e.g. by
 * reduce-scatter, then add to the row term); NULL in SYMMETRIC mode. */
int ntxent_b200_backward_rows(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                              int64_t row_begin, int64_t row_count, float tau, int mode,
                              int dtype, unsigned flags, void* workspace, size_t workspace_bytes,
                              float* du_rows_out, float* du_cols_out, ntxent_stream_t stream);

/* ---- Peer-memory ("push") variants: the collective fused into the kernel ----
 * On an NVSwitch box every GPU can map its peers' buffers (symmetric memory).
 * Instead of producing a local result and all-gathering it with NCCL, these
 * variants store the result straight into every rank's copy: `peers` lists the
 * device pointers of the SAME buffer on each rank (this rank's own included),
 * as mapped into this process.  No counterpart in the reference (it has no
 * communication code at all, CMakeLists.txt:13-14,41-47).
 *
 * Producer and consumers are ordered either by a cross-rank barrier the caller
 * puts between them (pass NULL for the signal/wait arguments), or by the
 * kernels themselves: every rank keeps an array of per-producer flags in
 * symmetric memory; a producing call publishes its call number (`epoch`, a
 * device-side counter, so a replayed CUDA graph needs no new arguments) to slot
 * `rank` of every peer's array once the grid that did the peer stores has
 * completed (a one-block kernel at the end of the call), and a consuming
 * kernel waits -- tile by tile, only for the ranks whose rows that tile reads --
 * until those slots have reached the consumer's own call number.  All ranks
 * must make the same sequence of calls.  The flag arrays and `epoch` start at
 * zero and belong to the library afterwards.
 *
 * A slow peer is not an error: a rank that stalls on its host (a checkpoint
 * write, a data-loader hiccup, a debugger) simply keeps the others waiting, as
 * it would inside a NCCL collective.  Only after the timeout set with
 * ntxent_b200_set_peer_wait_timeout_ms (default 600 000 ms, 0 = wait for ever)
 * does a waiting kernel give up: it stores 1 + the rank it was waiting for in
 * the wait's `status` word (device memory, zero-initialised by the caller) and
 * runs to completion on whatever is in the buffers -- the context stays usable,
 * the step's results are void, and the host finds out by reading the word.
 * With status == NULL the kernel traps at that point instead. */
#define NTXENT_MAX_PEERS 16
typedef struct {
    void* ptr[NTXENT_MAX_PEERS];
    int count;
} ntxent_peer_list;
typedef struct {
    ntxent_peer_list flags; /* uint32_t[NTXENT_MAX_PEERS] on every rank (own included) */
    uint32_t* epoch;        /* this rank's call counter for this signal (device memory) */
    int rank;               /* slot this rank publishes to */
    uint32_t* ticket;       /* optional device words, zero like the flags before the first call and
                               the library's afterwards; they turn kernel boundaries into "the last
                               block to finish does it".  normalize_push: ONE word -- the kernel
                               publishes the signal itself instead of a one-block kernel after it.
                               forward_rows_push: 1 + ceil(row_count / 128) words -- the positive
                               logits, l_i, the loss and the signal come out of the tile kernel
                               itself (the last CTA to finish a row tile, then the last row tile,
                               do the work): no finalize launch. */
} ntxent_peer_signal;
typedef struct {
    const uint32_t* flags;  /* this rank's copy of the flag array */
    const uint32_t* epoch;  /* the counter of the matching signal on this rank */
    int ranks;
    int64_t rows_per_rank;  /* rank r produced rows [r*rows_per_rank, (r+1)*rows_per_rank) */
    uint32_t* status;       /* device word, 0 = fine; set by a wait that timed out; may be NULL */
} ntxent_peer_wait;

/* Process-wide bound on how long a kernel waits for a peer's flag (see above).
 * Takes effect for calls made after it returns. */
int ntxent_b200_set_peer_wait_timeout_ms(uint64_t milliseconds);

/* ntxent_b200_normalize whose rows land at row `row_begin` of every rank's
 * gathered u_all [M, D] (replaces normalise + all-gather of U). */
int ntxent_b200_normalize_push(const void* z, int64_t rows, int64_t D, int dtype, unsigned flags,
                               const ntxent_peer_list* u_all_peers, int64_t row_begin,
                               float* inv_norm_out, const ntxent_peer_signal* u_signal,
                               ntxent_stream_t stream);

/* ntxent_b200_forward_rows whose l_i land at element row_begin + i of every
 * rank's lse_all float[M] and whose partial loss lands at element `loss_index`
 * (the rank) of every rank's loss_parts float[ranks] (replaces the all-gather
 * of l and the all-reduce of the loss).  `u_ready` (optional) makes the tile
 * kernel wait for the peers' normalize_push; `lse_signal` (optional) publishes
 * this call's l and partial loss. */
int ntxent_b200_forward_rows_push(const void* u_all, int64_t M, int64_t D, int64_t row_begin,
                                  int64_t row_count, float tau, int dtype, unsigned flags,
                                  void* workspace, size_t workspace_bytes,
                                  const ntxent_peer_list* lse_all_peers,
                                  const ntxent_peer_list* loss_parts_peers, int loss_index,
                                  const ntxent_peer_wait* u_ready, const ntxent_peer_signal* lse_signal,
                                  ntxent_stream_t stream);

/* One tiny kernel that waits for every slot of `ready` (for consumers that are
 * not this library's kernels, e.g. the sum of loss_parts after a forward-only
 * call). */
int ntxent_b200_peer_wait(const ntxent_peer_wait* ready, ntxent_stream_t stream);

/* Symmetric-form backward for rows [row_begin, row_begin+row_count) straight to
 * dz (ntxent_b200_backward_rows + ntxent_b200_finalize_dz in two launches
 * instead of three).  `lse_ready` (optional) makes the tile kernel wait for the
 * peers' forward_rows_push, block of columns by block of columns.
 * loss_parts / loss_parts_count / loss_out (optional, NULL / 0 / NULL): the
 * epilogue also writes loss_out[0] = (sum of loss_parts[0..count)) / M -- the
 * global mean loss from the ranks' partial sums that forward_rows_push stored,
 * all of which have landed once this call's tile kernel has waited on
 * `lse_ready` -- which saves the caller a kernel of its own. */
int ntxent_b200_backward_rows_dz(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                                 int64_t row_begin, int64_t row_count, float tau, int dtype,
                                 unsigned flags, void* workspace, size_t workspace_bytes,
                                 const void* z_rows, const float* inv_norm_rows, const float* grad_out,
                                 void* dz_out, const ntxent_peer_wait* lse_ready,
                                 const float* loss_parts, int loss_parts_count, float* loss_out,
                                 ntxent_stream_t stream);

/* Normalisation Jacobian: dz_i = g (dU_i - (dU_i . u_i) u_i) * inv_norm_i with
 * u_i = z_i * inv_norm_i.  z_rows, dz_out in `dtype`; g = *grad_out (NULL: 1). */
int ntxent_b200_finalize_dz(const float* du_rows, const void* z_rows, const float* inv_norm_rows,
                            const float* grad_out, int64_t rows, int64_t D, int dtype,
                            void* dz_out, ntxent_stream_t stream);

/* One term of the CONTRACT-form backward at a time, so that a caller can put a
 * collective between them (the column term goes into a reduce-scatter that can
 * run while the row term is being computed): part 0 -> out float[row_count, D],
 * the row term G_{I,:} U; part 1 -> out float[M, D], the column term
 * G_{I,:}^T U_I of every global row.  Same arithmetic as
 * ntxent_b200_backward_rows(mode = NTXENT_BWD_CONTRACT). */
int ntxent_b200_backward_rows_part(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                                   int64_t row_begin, int64_t row_count, float tau, int part,
                                   int dtype, unsigned flags, void* workspace, size_t workspace_bytes,
                                   float* out, ntxent_stream_t stream);

/* ---- Communicator: the row-partitioned step as ONE call per rank --------------
 * The reference names NCCL and MPI as build options (CMakeLists.txt:13-14,41-47,
 * 115-121: ENABLE_NCCL / ENABLE_MPI) and never calls either; north_star asks for
 * "the MPI/NCCL communicator setup".  A communicator is created for a fixed
 * problem (global M, D, dtype, flags) and owns everything a rank needs between
 * its inputs and outputs: the gathered-operand buffers (two rotating sets, mapped
 * on every peer -- cudaMalloc + peer access inside one process, CUDA IPC handles
 * across processes), the flag arrays and call counters of the push path, the row
 * workspace, and -- when libnccl.so.2 can be loaded -- a NCCL communicator for
 * the collective-based exchanges.  Creation and destruction allocate and
 * synchronise; ntxent_b200_dist_forward_backward does neither.
 *
 *   ntxent_b200_comm_init_all   single process, one communicator per listed
 *                               device (drive each from its own host thread, or
 *                               from one thread, a device at a time: the calls
 *                               only enqueue work).
 *   ntxent_b200_comm_unique_id + ntxent_b200_comm_init_rank
 *                               one process per rank: rank 0 creates the 128-byte
 *                               id, the caller hands it to every rank by its own
 *                               means (MPI_Bcast, a torch.distributed broadcast, a
 *                               file), every rank calls init_rank with its device
 *                               current.  Needs NCCL (the bootstrap runs over it).
 * MPI itself is not used anywhere: it would only carry the 128 bytes. */
typedef struct ntxent_comm ntxent_comm_t;

enum {
    NTXENT_EXCHANGE_PUSH = 0, /* kernels store into the peers' buffers, ordered by flags
                                 (the fused form; no collective launches)                  */
    NTXENT_EXCHANGE_NCCL = 1  /* ncclAllGather of U (+ of l, or ncclReduceScatter of the
                                 column term in CONTRACT mode, overlapped with the row pass
                                 on a second stream): north_star's formulation             */
};

int ntxent_b200_comm_init_all(int ndev, const int* devices, int64_t M, int64_t D, int dtype,
                              unsigned flags, ntxent_comm_t** comms_out);
int ntxent_b200_comm_unique_id(void* id128_out);
int ntxent_b200_comm_init_rank(const void* id128, int nranks, int rank, int64_t M, int64_t D,
                               int dtype, unsigned flags, ntxent_comm_t** comm_out);
int ntxent_b200_comm_destroy(ntxent_comm_t* comm);
/* rank, number of ranks, rows per rank, and whether the NCCL exchanges are available */
int ntxent_b200_comm_info(const ntxent_comm_t* comm, int* rank, int* nranks, int64_t* rows_per_rank,
                          int* has_nccl);

/* Fused forward+backward of the GLOBAL loss from this rank's shard.
 * z_rows: device [M/nranks, D] in the communicator's dtype (rows rank*M/nranks ..).
 * grad_out: device float[1] or NULL (= 1).  loss_out: device float[1], the global
 * mean loss (identical on every rank).  dz_rows_out: device [M/nranks, D].
 * exchange / bwd_mode: NTXENT_EXCHANGE_* / NTXENT_BWD_* (PUSH runs the SYMMETRIC
 * form).  Every rank must make the same sequence of calls.  Asynchronous on
 * `stream`; capturable into a CUDA graph once per buffer set (two consecutive
 * calls).  After a peer-wait timeout (see ntxent_b200_set_peer_wait_timeout_ms)
 * the results are void and ntxent_b200_comm_status returns the rank that was
 * waited for. */
int ntxent_b200_dist_forward_backward(ntxent_comm_t* comm, const void* z_rows, const float* grad_out,
                                      float tau, int exchange, int bwd_mode, float* loss_out,
                                      void* dz_rows_out, ntxent_stream_t stream);
/* 0: fine; r + 1: a kernel of this rank gave up waiting for rank r.  Synchronises. */
int ntxent_b200_comm_status(ntxent_comm_t* comm);

/* Forward pass for callers that are not allowed to allocate device memory.
 *
 * This entry point has no counterpart in the reference, which allocates its
 * temporaries through ATen on every call.  It exists for hosts that hand this
 * library a stream and buffers but own no allocator of their own and may not
 * call cudaMalloc: a plugin running inside another framework's memory pool, a
 * service whose device memory was carved up at start-up, an embedded or
 * real-time loop in which allocation is forbidden after start-up.  Such a
 * caller cannot provide the `workspace` that ntxent_b200_forward asks for.
 *
 * Same results as ntxent_b200_forward (same kernels, same order), but the
 * scratch lives in an array of NTXENT_STATIC_SCRATCH_BYTES that is built into
 * the library.  The caller still passes z, loss_out and lse_out.
 *
 *   - Limit: problems whose forward scratch exceeds NTXENT_STATIC_SCRATCH_BYTES
 *     (8 MiB: about 2N = 8192 at D = 128 with fp32 operands, twice that with
 *     bf16) are refused with NTXENT_ERR_STATIC_SCRATCH_EXCEEDED and nothing is
 *     launched.  Every GPU that loads the library keeps the 8 MiB until exit.
 *   - The first call on each device looks up the array's device address (one
 *     cudaGetSymbolAddress with that device current); make that first call
 *     before starting any stream capture.  No call allocates or frees.  The
 *     address is remembered together with the driver context it belongs to,
 *     and looked up again when a call finds another context current (after
 *     cudaDeviceReset(), or in a host that switches contexts on one device).
 *   - All calls on a device share the one array, and it keeps whatever the last
 *     call wrote.  The pass writes every part of it before reading it, so
 *     back-to-back calls on one stream are fine; calls that can run on the GPU
 *     at the same time (different streams, different host threads) race on it
 *     and must be serialised by the caller. */
#define NTXENT_STATIC_SCRATCH_BYTES (8u << 20)
int ntxent_b200_forward_static(const void* z, int64_t M, int64_t D, float tau, int dtype,
                               unsigned flags, float* loss_out, float* lse_out,
                               ntxent_stream_t stream);

/* ---- Measurement aids (the reference times with host clocks around the op,
 * src/benchmark.cpp:30-39; these give the benchmark per-kernel device times) ----
 * Between _begin and _end every launch of the two tile kernels (forward pass,
 * backward pass) is bracketed by CUDA events on its stream.  _end synchronises
 * on them and returns the summed milliseconds and launch counts.  Not
 * thread-safe; meant for bench.py only. */
int ntxent_b200_profile_begin(void);
int ntxent_b200_profile_end(float* fwd_ms, int* fwd_launches, float* bwd_ms, int* bwd_launches);

/* Debug aid for tools/trace_bwd.py: while a device buffer of 5*64*8 uint64 is
 * set, CTA 0 of the tcgen05 backward kernel stamps clock64() at its pipeline
 * hand-offs into it.  NULL switches it off (the default).  The stamps exist only
 * in the trace build of the library (-DNTXENT_TRACE=1, libntxent_b200_trace.so);
 * the product build refuses a non-NULL buffer with NTXENT_ERR_INVALID_ARGUMENT. */
int ntxent_b200_debug_set_trace(void* dev_buffer);

/* Which kernel family and how many column splits a call with these arguments
 * will use (reported by the benchmark next to its numbers). */
int ntxent_b200_plan_info(int64_t M, int64_t D, int64_t row_count, int dtype, unsigned flags,
                          int* uses_tcgen05, int* fwd_splits, int* bwd_splits);

#ifdef __cplusplus
}
#endif
#endif /* NTXENT_B200_H */
