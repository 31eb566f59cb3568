/*
 * fa_handle.h -- the C ABI of libflashattn_b200.so, part 3: the setter-style params handles of the reference's
 * FlashMask-v2 and FA3 libraries.
 *
 * Replaces csrc/flashmask_v2/flash_api.h:14-266 (prefix flashmaskv2_) and csrc/flash_attn_v3/flash_api.h (prefix fa3_,
 * the same list without the FlashMask fields): create / clear / cast / destroy a heap params object, one getter and
 * one setter per field, run_mha_fwd / run_mha_bwd / run_mha_fwd_combine and the three heuristics getters.  This is the
 * surface PaddlePaddle drives (`flashmask_attention`, `flash_attn_v3`), so a caller of libflashmaskv2.so /
 * libflashattnv3.so can dlopen this library instead.
 *
 * The handle is opaque to callers (they only ever hold the pointer); every field of the reference's struct has a
 * slot here, the run functions translate the slots into fa_b200_params (fa_b200.h) and launch the same kernels.
 * Semantics on this surface are the v2/v3 ones: causal and window masks are bottom-right aligned
 * (csrc/flashmask_v2/mask.h:85-118), rows with no visible key give lse = -inf (csrc/flashmask_v2/softmax.h:153),
 * varlen lse is (h, total_q) (csrc/flashmask_v2/flash_fwd_launch_template.h:161-164).
 * Errors are thrown C++ exceptions (std::runtime_error), as in the reference (flash_api.cu:14-25).
 *
 * Decode fields are honoured on run_mha_fwd: page_table (+ page_size, num_pages, page_table_batch_stride), kv_batch_idx
 * (+ b_k), leftpad_k, knew / vnew (appended at seqused_k, then attended), rotary_cos / rotary_sin (+ rotary_dim,
 * is_rotary_interleaved; applied to q and knew), softcap (forward and backward).
 * seqused_q limits the query rows of each batch element (forward and backward).
 * is_e4m3 (+ q/k/v_descale_ptr and their strides): FP8 forward, e4m3 q/k/v bytes in, bf16 out (flash_api.cu:107-131);
 * plain / causal / local / varlen calls, head dim a multiple of 16; anything else with is_e4m3 throws.
 * Fields accepted but requiring an unsupported feature make run_* throw: qv, cu_seqlens_knew, p_dropout < 1, dv != d.
 */
#ifndef FA_B200_HANDLE_H_
#define FA_B200_HANDLE_H_

#include <stdint.h>
#include <stdbool.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct Flash_fwd_params Flash_fwd_params;
typedef struct Flash_bwd_params Flash_bwd_params;

/* X-lists of (type, member): forward fields (also present on the backward handle), FlashMask fields, backward fields */
#define FA_HANDLE_FWD_FIELDS(X) \
  X(void *, q_ptr) X(void *, k_ptr) X(void *, v_ptr) \
  X(int64_t, q_batch_stride) X(int64_t, k_batch_stride) X(int64_t, v_batch_stride) \
  X(int64_t, q_row_stride) X(int64_t, k_row_stride) X(int64_t, v_row_stride) \
  X(int64_t, q_head_stride) X(int64_t, k_head_stride) X(int64_t, v_head_stride) X(int64_t, v_dim_stride) \
  X(int, h) X(int, h_k) \
  X(void *, o_ptr) X(void *, oaccum_ptr) \
  X(int64_t, o_batch_stride) X(int64_t, o_row_stride) X(int64_t, o_head_stride) \
  X(void *, softmax_lse_ptr) X(void *, softmax_lseaccum_ptr) \
  X(float *, q_descale_ptr) X(float *, k_descale_ptr) X(float *, v_descale_ptr) \
  X(int64_t, q_descale_batch_stride) X(int64_t, q_descale_head_stride) \
  X(int64_t, k_descale_batch_stride) X(int64_t, k_descale_head_stride) \
  X(int64_t, v_descale_batch_stride) X(int64_t, v_descale_head_stride) \
  X(int, b) X(int, seqlen_q) X(int, seqlen_k) X(int, seqlen_knew) X(int, d) \
  X(int, seqlen_q_rounded) X(int, seqlen_k_rounded) X(int, d_rounded) X(int, rotary_dim) \
  X(int, total_q) X(int, total_k) X(int, total_knew) X(int, b_k) X(int, dv) X(int, dv_rounded) \
  X(float, scale_softmax) X(float, softcap) \
  X(int *, cu_seqlens_q) X(int *, cu_seqlens_k) X(int *, cu_seqlens_knew) X(int *, leftpad_k) \
  X(int *, seqused_q) X(int *, seqused_k) \
  X(int64_t, oaccum_split_stride) X(int64_t, oaccum_batch_stride) X(int64_t, oaccum_row_stride) \
  X(int64_t, oaccum_head_stride) X(int64_t, lseaccum_split_stride) X(int64_t, lseaccum_batch_stride) \
  X(int64_t, lseaccum_head_stride) \
  X(void *, knew_ptr) X(void *, vnew_ptr) \
  X(int64_t, knew_batch_stride) X(int64_t, vnew_batch_stride) X(int64_t, knew_row_stride) \
  X(int64_t, vnew_row_stride) X(int64_t, knew_head_stride) X(int64_t, vnew_head_stride) \
  X(void *, qv_ptr) X(int64_t, qv_batch_stride) X(int64_t, qv_row_stride) X(int64_t, qv_head_stride) \
  X(void *, rotary_cos_ptr) X(void *, rotary_sin_ptr) X(int *, kv_batch_idx) X(int *, page_table) \
  X(int64_t, page_table_batch_stride) X(int, page_size) X(int, num_pages) X(bool, pagedkv_tma) \
  X(float, p_dropout) X(uint8_t, p_dropout_in_uint8_t) X(float, rp_dropout) \
  X(int, window_size_left) X(int, window_size_right) X(uint64_t *, rng_state) \
  X(bool, is_bf16) X(bool, is_fp32) X(bool, is_e4m3) X(bool, is_causal) X(bool, is_local) \
  X(bool, is_rotary_interleaved) X(int, num_splits) X(bool, pack_gqa) \
  X(int *, tile_count_semaphore) X(int *, num_splits_dynamic_ptr) X(bool, skip_scheduler_metadata_computation) \
  X(int, arch) X(int, num_sm)

#define FA_HANDLE_MASK_FIELDS(X) \
  X(int, h_flashmask) X(int, h_h_flashmask_ratio) \
  X(int32_t *, lt_start_ptr) X(int32_t *, lt_end_ptr) X(int32_t *, ut_start_ptr) X(int32_t *, ut_end_ptr) \
  X(int32_t *, flashmask_maxmin_ptr)

#define FA_HANDLE_BWD_FIELDS(X) \
  X(void *, do_ptr) X(void *, dq_ptr) X(void *, dk_ptr) X(void *, dv_ptr) \
  X(void *, dq_accum_ptr) X(void *, dk_accum_ptr) X(void *, dv_accum_ptr) \
  X(int64_t, do_batch_stride) X(int64_t, do_row_stride) X(int64_t, do_head_stride) \
  X(int64_t, dq_batch_stride) X(int64_t, dk_batch_stride) X(int64_t, dv_batch_stride) \
  X(int64_t, dq_row_stride) X(int64_t, dk_row_stride) X(int64_t, dv_row_stride) \
  X(int64_t, dq_head_stride) X(int64_t, dk_head_stride) X(int64_t, dv_head_stride) \
  X(void *, dsoftmax_sum) X(void *, softmax_lse_log2_ptr) \
  X(int *, dq_semaphore) X(int *, dk_semaphore) X(int *, dv_semaphore) \
  X(bool, deterministic) X(int64_t, dq_accum_split_stride)

#define FA_HANDLE_BWD_MASK_FIELDS(X) \
  X(int32_t *, lt_start_nblockmax) X(int32_t *, lt_start_nblockmin) X(int32_t *, lt_end_nblockmax) \
  X(int32_t *, lt_end_nblockmin) X(int32_t *, ut_start_nblockmax) X(int32_t *, ut_start_nblockmin) \
  X(int32_t *, ut_end_nblockmax) X(int32_t *, ut_end_nblockmin)

#define FA_HANDLE_DECLARE_LIFECYCLE(P) \
  Flash_fwd_params *P##_create_fwd_params_handle(void); \
  Flash_bwd_params *P##_create_bwd_params_handle(void); \
  void P##_clear_fwd_params_handle(Flash_fwd_params *params_handle); \
  void P##_clear_bwd_params_handle(Flash_bwd_params *params_handle); \
  Flash_fwd_params *P##_cast_to_fwd_params_handle(Flash_bwd_params *params_handle); \
  void P##_destroy_fwd_params_handle(Flash_fwd_params *params_handle); \
  void P##_destroy_bwd_params_handle(Flash_bwd_params *params_handle); \
  void P##_run_mha_fwd_combine(Flash_fwd_params *params_handle, cudaStream_t stream, bool enable_pdl); \
  void P##_run_mha_fwd(Flash_fwd_params *params_handle, cudaStream_t stream); \
  bool P##_get_pagedkv_tma(Flash_fwd_params *params_handle); \
  bool P##_get_pack_gqa(Flash_fwd_params *params_handle); \
  int P##_get_num_splits(Flash_fwd_params *params_handle); \
  void P##_run_mha_bwd(Flash_bwd_params *params_handle, cudaStream_t stream);

FA_HANDLE_DECLARE_LIFECYCLE(flashmaskv2)
FA_HANDLE_DECLARE_LIFECYCLE(fa3)

#define FA_HANDLE_DECL_BOTH_V2(type, member) \
  type flashmaskv2_fwd_params_get_##member(const Flash_fwd_params *params_handle); \
  void flashmaskv2_fwd_params_set_##member(Flash_fwd_params *params_handle, type value); \
  type flashmaskv2_bwd_params_get_##member(const Flash_bwd_params *params_handle); \
  void flashmaskv2_bwd_params_set_##member(Flash_bwd_params *params_handle, type value);
#define FA_HANDLE_DECL_BWD_V2(type, member) \
  type flashmaskv2_bwd_params_get_##member(const Flash_bwd_params *params_handle); \
  void flashmaskv2_bwd_params_set_##member(Flash_bwd_params *params_handle, type value);
#define FA_HANDLE_DECL_BOTH_FA3(type, member) \
  type fa3_fwd_params_get_##member(const Flash_fwd_params *params_handle); \
  void fa3_fwd_params_set_##member(Flash_fwd_params *params_handle, type value); \
  type fa3_bwd_params_get_##member(const Flash_bwd_params *params_handle); \
  void fa3_bwd_params_set_##member(Flash_bwd_params *params_handle, type value);
#define FA_HANDLE_DECL_BWD_FA3(type, member) \
  type fa3_bwd_params_get_##member(const Flash_bwd_params *params_handle); \
  void fa3_bwd_params_set_##member(Flash_bwd_params *params_handle, type value);

FA_HANDLE_FWD_FIELDS(FA_HANDLE_DECL_BOTH_V2)
FA_HANDLE_MASK_FIELDS(FA_HANDLE_DECL_BOTH_V2)
FA_HANDLE_BWD_FIELDS(FA_HANDLE_DECL_BWD_V2)
FA_HANDLE_BWD_MASK_FIELDS(FA_HANDLE_DECL_BWD_V2)
FA_HANDLE_FWD_FIELDS(FA_HANDLE_DECL_BOTH_FA3)
FA_HANDLE_BWD_FIELDS(FA_HANDLE_DECL_BWD_FA3)

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_HANDLE_H_ */
