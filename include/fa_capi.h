/*
 * fa_capi.h -- the C ABI of libflashattn_b200.so, part 1: the FlashAttention-2 style entry points.
 *
 * These are the symbols the reference exports from libflashattn.so (reference:
 * csrc/capi/flash_attn.h) and that PaddlePaddle binds through dlopen/dlsym.  Names, argument
 * order, argument meaning and the error convention are identical, so the shared object is a
 * drop-in: every function returns true on success; on failure it returns false and the message
 * is available from flash_attn_error() on the calling thread.
 *
 * Conventions (reference: csrc/capi/flash_attn.cu:98-229, 576-684):
 *   - every pointer is a DEVICE pointer owned by the caller; work is enqueued on `stream` and the
 *     call returns immediately; the caller keeps the right device current;
 *   - element type is bf16 when is_bf16, else fp16; accumulation is fp32;
 *   - all strides are in ELEMENTS; the innermost (head_size) stride is 1;
 *   - the caller pre-allocates softmax_lse (b, h, seqlen_q) fp32, and for backward
 *     softmax_d (b, h, seqlen_q_rounded) fp32 and dq_accum (b, h, seqlen_q_rounded, head_size_rounded)
 *     fp32, with seqlen_*_rounded = round_up(seqlen, 128), head_size_rounded = round_up(d, 32);
 *   - causal masking is TOP-LEFT aligned; rows with no visible key give out = 0, lse = +inf;
 *   - backward with GQA/MQA: dk/dv carry `num_heads` (expanded) heads, the caller sums groups.
 *
 * Layout notes: q (b, seqlen_q, h, d); k, v (b, seqlen_k, h_k, d); varlen: (total, h, d) + cu_seqlens (b+1).
 */
#ifndef FA_B200_CAPI_H_
#define FA_B200_CAPI_H_

#include <stdint.h>
#include <stdbool.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

/* replaces csrc/capi/flash_attn.h:11-55 */
bool flash_attn_fwd(
    const void *q, const void *k, const void *v,
    void *rng_state, void *out, void *softmax_ptr, void *softmax_lse_ptr,
    int batch_size, int seqlen_q, int seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
    int num_heads, int num_heads_k, int head_size, int head_size_rounded,
    float p_dropout, float softmax_scale, float softmax_unscale,
    bool is_causal, bool return_softmax, bool is_bf16,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const int64_t *mask_dims,
    const void *flashmask_downstart_ptr, const int64_t *flashmask_dims,
    const void *flashmask_upend_ptr, const void *flashmask_downend_ptr,
    const void *flashmask_upstart_ptr, const void *flashmask_maxmin_ptr,
    int q_row_stride, int k_row_stride, int v_row_stride,
    int q_head_stride, int k_head_stride, int v_head_stride,
    int o_row_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride);

/* replaces csrc/capi/flash_attn.h:57-98 */
bool flash_attn_varlen_fwd(
    const void *q, const void *k, const void *v,
    const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k,
    void *rng_state, void *out, void *softmax_ptr, void *softmax_lse_ptr,
    int batch_size, int max_seqlen_q, int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
    int num_heads, int num_heads_k, int head_size, int head_size_rounded,
    float p_dropout, float softmax_scale, float softmax_unscale,
    bool is_causal, bool return_softmax, bool is_bf16,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const void *mask_dims,
    int q_row_stride, int k_row_stride, int v_row_stride,
    int q_head_stride, int k_head_stride, int v_head_stride,
    int o_row_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride,
    bool varlen_padded_input);

/* replaces csrc/capi/flash_attn.h:100-161 */
bool flash_attn_bwd(
    const void *dout, const void *q, const void *k, const void *v, const void *out,
    const void *softmax_d, const void *softmax_lse, void *rng_state,
    void *dq, void *dk, void *dv, void *dq_accum,
    int batch_size, int seqlen_q, int seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
    int num_heads, int num_heads_k, int head_size, int head_size_rounded,
    float p_dropout, float softmax_scale, float softmax_unscale,
    bool is_causal, bool is_bf16, int num_splits,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const int64_t *mask_dims,
    const void *flashmask_downstart_ptr, const int64_t *flashmask_dims,
    const void *flashmask_upend_ptr, const void *flashmask_downend_ptr,
    const void *flashmask_upstart_ptr, const void *flashmask_maxmin_ptr,
    int q_row_stride, int k_row_stride, int v_row_stride,
    int q_head_stride, int k_head_stride, int v_head_stride,
    int o_row_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride,
    int dq_row_stride, int dk_row_stride, int dv_row_stride,
    int dq_head_stride, int dk_head_stride, int dv_head_stride,
    int do_row_stride, int do_head_stride,
    int dq_batch_stride, int dk_batch_stride, int dv_batch_stride, int do_batch_stride);

/* replaces csrc/capi/flash_attn.h:163-221 */
bool flash_attn_varlen_bwd(
    const void *dout, const void *q, const void *k, const void *v, const void *out,
    const void *softmax_d, const void *softmax_lse,
    const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state,
    void *dq, void *dk, void *dv, void *dq_accum,
    int batch_size, int max_seqlen_q, int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
    int num_heads, int num_heads_k, int head_size, int head_size_rounded,
    float p_dropout, float softmax_scale, float softmax_unscale,
    bool is_causal, bool is_bf16, int num_splits,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const int64_t *mask_dims,
    int q_row_stride, int k_row_stride, int v_row_stride,
    int q_head_stride, int k_head_stride, int v_head_stride,
    int o_row_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride,
    int dq_row_stride, int dk_row_stride, int dv_row_stride,
    int dq_head_stride, int dk_head_stride, int dv_head_stride,
    int do_row_stride, int do_head_stride,
    int dq_batch_stride, int dk_batch_stride, int dv_batch_stride, int do_batch_stride,
    bool varlen_padded_input);

/* replaces csrc/capi/flash_attn.h:223-247.  reduced_scores[b,h,k] = sum_q softmax(QK^T)[q,k]. */
bool calc_reduced_attn_scores(
    const void *q, const void *k, const void *softmax_lse, void *reduced_scores, void *softmax_ptr,
    int batch_size, int seqlen_q, int seqlen_k, int num_heads, int num_heads_k, int head_size,
    float softmax_scale, bool return_softmax, bool is_bf16, int num_splits, cudaStream_t stream,
    int q_row_stride, int k_row_stride, int o_row_stride,
    int q_head_stride, int k_head_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int o_batch_stride);

/* replaces csrc/capi/flash_attn.h:249-311 (legacy FA1 path with an additive mask and an additive bias + its gradient;
 * csrc/flash_attn_with_bias_and_mask/flash_attn_with_bias_mask.cu:268-541).  Packed (total, h, d) tensors + cu_seqlens,
 * no GQA, no causal; attn_mask is (b, 1|h, 1|max_seqlen_q, max_seqlen_k), attn_bias (n, h, max_seqlen_q, max_seqlen_k)
 * read as bias[b % n]; softmax_scale must be 1 when either is given; softmax_lse / dsoftmax are
 * (b, h, round_up(max_seqlen_q, 16)) fp32; dbias is (b, h, max_seqlen_q, max_seqlen_k) in the element type, zero-filled
 * by the call; out == NULL only reports *workspace_size (0: no workspace is needed, workspace_ptr is ignored).
 * Mask / bias rows are strided by max_seqlen_k (the reference strides by each sequence's own length, the same thing
 * for the uniform batches its callers pass). */
bool flash_attn_fwd_with_bias_and_mask(
    const void *q, const void *k, const void *v, void *out,
    const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k,
    int total_q, int total_k, int batch_size, int num_heads, int head_size,
    int max_seqlen_q_, int max_seqlen_k_, float p_dropout, float softmax_scale,
    bool zero_tensors, bool is_bf16, int num_splits,
    void *softmax_lse_ptr, void *workspace_ptr, uint64_t *workspace_size,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const void *attn_bias, const int64_t *mask_dims, const int64_t *bias_dims);

bool flash_attn_bwd_with_bias_and_mask(
    const void *q, const void *k, const void *v, void *dq, void *dk, void *dv,
    const void *out, const void *dout,
    const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k,
    int total_q, int total_k, int batch_size, int num_heads, int head_size,
    int max_seqlen_q_, int max_seqlen_k_, float p_dropout, float softmax_scale,
    bool zero_tensors, bool is_bf16, int num_splits,
    const void *softmax_lse_ptr, void *dsoftmax_ptr, void *dbias_ptr,
    void *workspace_ptr, uint64_t *workspace_size,
    cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const void *attn_bias, const int64_t *mask_dims, const int64_t *bias_dims);

/* replaces csrc/capi/flash_attn.h:313-315; message storage is thread-local (flash_attn.cu:45-60) */
void flash_attn_set_error(const char *msg);
const char *flash_attn_error(void);

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_CAPI_H_ */
