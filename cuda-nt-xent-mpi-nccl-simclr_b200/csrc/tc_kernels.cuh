// tc_kernels.cuh -- interface of the tcgen05/TMEM/TMA tile kernels (tc_kernels.cu).
#pragma once
#include "common.cuh"

namespace ntxent_b200 {

// True when the tcgen05 tile kernels cover this problem: bf16 operands with D in {64,128,192,256}
// (.kind::f16), or fp32 operands with D in {64,128} and M >= 1024 (.kind::tf32).
bool tc_supported(int64_t M, int64_t D, bool op_bf16);

SplitPlan tc_plan_fwd(int64_t M, int64_t D, int64_t row_count, bool op_bf16, int num_sms);
SplitPlan tc_plan_bwd(int64_t M, int64_t D, int64_t a_count, int64_t b_count, int num_sms);

// What the forward tile kernel does after its last tile when it is given tickets: the CTA that
// finishes a row tile LAST (of its column splits) merges the tile's partials, forms the positive
// logits, stores l_i (to every peer) and the tile's loss partial; the CTA that finishes the last row
// tile sums the loss partials in a fixed order, stores the loss and publishes the signal.  No
// finalize kernel is launched then.  tickets: 1 + row_tiles device words, zero before the call (and
// again after it); nullptr: the partials are left for the finalize kernel.
struct FwdTail {
    unsigned* tickets;
    const int* pos_index;  // positives from an index vector, or nullptr
    PeerList lse_outs;
    int64_t lse_offset;
    float* block_partials;
    PeerList loss_outs;
    int loss_index;
    float loss_scale;
    PeerSignal signal;
};

cudaError_t tc_launch_fwd(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                          int64_t row_count, float tau, unsigned flags, const SplitPlan& plan,
                          float* part_m, float* part_s, const PeerWait& u_ready,
                          cudaStream_t stream, bool cross_view = false, const FwdTail* tail = nullptr);

// Backward epilogue inside the tile kernel (bf16 operands, D <= 128, symmetric form, default
// positives): every unit leaves its partial and takes a ticket on its row tile; the CTA that
// takes the last one adds the partials in slot order, subtracts the positive-pair term, applies
// the normalisation Jacobian and stores dz -- the work of bwd_epilogue_kernel without its launch,
// spread over the kernel's run instead of 3.5 latency-bound waves at its end.
// tickets: one device word per row tile, zero before the launch; the kernel leaves them zero.
struct BwdTail {
    unsigned* tickets;
    const void* z_rows;     // [a_count, D], bf16 or fp32
    int z_bf16;
    const float* inv_norm;  // [a_count]
    const float* grad_out;  // device scalar or nullptr (= 1)
    void* dz_out;           // [a_count, D] in z's type
    float pos_coeff;        // 2 for the symmetric form
    float scale;            // 1 / (tau M)
    const float* loss_parts;  // optional: CTA 0 adds loss_n partials in order -> loss_out
    int loss_n;
    float loss_scale;
    float* loss_out;
};
bool tc_bwd_tail_supported(int64_t D, bool op_bf16);

cudaError_t tc_launch_bwd(const void* u_all, bool op_bf16, const float* lse_all, int64_t M, int64_t D,
                          int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count,
                          float tau, int mode, unsigned flags, const SplitPlan& plan, float* acc,
                          float* ut_scratch, const PeerWait& lse_ready, cudaStream_t stream,
                          bool cross_view = false, const BwdTail* tail = nullptr);

// ---- forward over the upper triangle (tc_fwd_sym.cu): whole-batch calls on one device, bf16
// operands, fixed-max regime (log2(e)/tau <= 50), M >= kSymMinRows.
constexpr int64_t kSymMinRows = 4096;
struct SymPlan {
    int T;      // row tiles = column tiles
    int L;      // tiles per chunk (one CTA per chunk)
    int items;  // CTAs
    int slots;  // row-sum partials per row (chunks of the longest row)
    size_t cpart_floats;  // column partials: T (T-1)/2 tile pairs x 4 quadrants x 128 columns
};
bool tc_sym_supported(int64_t M, int64_t D, bool op_bf16, float tau, unsigned flags);
SymPlan tc_sym_plan(int64_t M, int64_t D, int num_sms);
// part_s [slots][M], cpart [cpart_floats], pos_t [M] (positive logits, base-2 units)
cudaError_t tc_launch_fwd_sym(const void* u_all, int64_t M, int64_t D, float tau, const SymPlan& sp,
                              float* part_s, float* cpart, float* pos_t, cudaStream_t stream);
// -> lse_out [M], block_partials [T] (per-row-tile sums of l_i - S_i,pos(i)); with a ticket
// `counter` (zero before the call, zero after it) also loss_out[0] = loss_scale * their sum
cudaError_t tc_launch_fwd_sym_finalize(int64_t M, const SymPlan& sp, const float* part_s, const float* cpart,
                                       const float* pos_t, float* lse_out, float* block_partials,
                                       cudaStream_t stream, unsigned* counter = nullptr, float* loss_out = nullptr,
                                       float loss_scale = 1.f);

// Scratch the tf32 backward needs for its transposed copy of U ([D, M rounded up to 4] fp32);
// zero for bf16 operands.
size_t tc_bwd_scratch_bytes(int64_t M, int64_t D, bool op_bf16);
int64_t tc_ut_cols(int64_t M);

// Debug aid: CTA (0,0) of the backward tile kernel stamps its pipeline hand-offs into this
// device buffer ([3 roles][64 tiles][8 events] of clock64 values); nullptr turns it off.
bool tc_set_trace_buffer(unsigned long long* dev_ptr);  // false: time stamps not compiled in

}  // namespace ntxent_b200
