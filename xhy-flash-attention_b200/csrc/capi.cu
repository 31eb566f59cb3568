// capi.cu -- the flat FlashAttention-2 style symbols of include/fa_capi.h.
//
// Each function translates the reference's argument list (csrc/capi/flash_attn.h) into fa_b200_params
// the way set_params_fprop_strided / set_params_dgrad_strided do (csrc/capi/flash_attn.cu:98-229,231-300),
// then takes the same dispatcher as the descriptor entry points.  FA2-surface semantics are fixed here:
// top-left causal alignment, +inf LSE for rows without a visible key, per-q-head (expanded) dK/dV.
#include <cstring>

#include "common.h"
#include "../../include/fa_capi.h"

namespace {

using fa::Params;

Params base_params(int batch_size, int seqlen_q, int seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
                   int num_heads, int num_heads_k, int head_size, int head_size_rounded, float softmax_scale,
                   bool is_causal, bool is_bf16) {
  Params p;
  memset(&p, 0, sizeof(p));
  p.abi_version = FA_B200_ABI_VERSION;
  p.is_bf16 = is_bf16 ? 1 : 0;
  p.b = batch_size; p.seqlen_q = seqlen_q; p.seqlen_k = seqlen_k;
  p.h = num_heads; p.h_k = num_heads_k; p.d = head_size;
  p.seqlen_q_rounded = seqlen_q_rounded; p.seqlen_k_rounded = seqlen_k_rounded; p.d_rounded = head_size_rounded;
  p.total_q = p.total_k = -1;
  p.softmax_scale = softmax_scale;
  p.is_causal = is_causal ? 1 : 0;
  p.causal_bottom_right = 0;
  p.window_left = p.window_right = -1;
  p.lse_empty_is_neg_inf = 0;
  return p;
}

fa_b200_tensor tensor(const void *ptr, int batch_stride, int row_stride, int head_stride) {
  fa_b200_tensor t;
  t.ptr = const_cast<void *>(ptr);
  t.batch_stride = batch_stride; t.row_stride = row_stride; t.head_stride = head_stride;
  return t;
}

// dense additive mask: mask_dims = (batch, 1-or-heads, 1-or-seqlen_q, seqlen_k)  (csrc/capi/flash_attn.cu:78-83)
void set_attn_mask(Params &p, const void *attn_mask, const int64_t *mask_dims, float softmax_unscale) {
  if (!attn_mask) return;
  FA_CHECK(mask_dims != nullptr, "mask_dims must be given with attn_mask");
  FA_CHECK(mask_dims[0] == p.b && mask_dims[3] == p.seqlen_k, "mask_dims must be (batch, 1 or heads, 1 or seqlen_q, seqlen_k)");
  FA_CHECK(mask_dims[1] == 1 || mask_dims[1] == p.h, "mask_dims[1] must be 1 or num_heads");
  FA_CHECK(mask_dims[2] == 1 || mask_dims[2] == p.seqlen_q, "mask_dims[2] must be 1 or seqlen_q");
  p.attn_mask = attn_mask;
  p.mask_head_mod = (int)mask_dims[1]; p.mask_seq_mod = (int)mask_dims[2];
  p.unscale_softmax = softmax_unscale;
}

void set_flashmask(Params &p, const void *downstart, const int64_t *dims, const void *upend, const void *downend,
                   const void *upstart, const void *maxmin) {
  if (!downstart && !upend && !downend && !upstart) return;
  FA_CHECK(downstart != nullptr, "flashmask_downstart_ptr is mandatory when any flashmask vector is given");
  FA_CHECK(dims != nullptr, "flashmask_dims must be given");
  // dims = (batch, mask heads, seqlen_k): csrc/capi/flash_attn.cu:623
  FA_CHECK(dims[0] == p.b, "flashmask_dims[0] must equal batch_size");
  FA_CHECK(dims[2] == p.seqlen_k, "flashmask_dims[2] must equal seqlen_k");
  // csrc/capi/flash_attn.cu:204-210: the workspace is required whenever a vector is given
  FA_CHECK(maxmin != nullptr, "flashmask_maxmin_ptr workspace must be given with flashmask vectors");
  p.h_mask = (int)dims[1];
  p.lt_start = (const int32_t *)downstart;
  p.lt_end = (const int32_t *)downend;
  p.ut_start = (const int32_t *)upstart;
  p.ut_end = (const int32_t *)upend;
  p.flashmask_maxmin = (int32_t *)const_cast<void *>(maxmin);
}

// The flat varlen symbols carry no total row counts, which the TMA descriptors of the tcgen05 kernels need as tensor
// extents (the tail tile of the last sequence relies on the out-of-bounds zero fill).  total_q / total_k stay -1 and
// the launchers patch the row extent of their tensor maps ON THE DEVICE from cu_seqlens[b] (tma_host.cu:
// patch_tma_rows): the call only enqueues work, as the reference's does (csrc/capi/flash_attn.cu:576-684), and stays on
// the tcgen05 kernels under CUDA-graph capture.
void resolve_varlen_totals(Params &, cudaStream_t) {}

// shared argument translation of the two FA1 symbols (flash_attn_with_bias_mask.cu:301-350,433-491)
Params fa1_params(int batch_size, int num_heads, int head_size, int max_seqlen_q_, int max_seqlen_k_, int total_q,
                  int total_k, float p_dropout, float softmax_scale, bool is_bf16, uint64_t seed, uint64_t offset,
                  const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, const void *attn_mask, const void *attn_bias,
                  const int64_t *mask_dims, const int64_t *bias_dims) {
  FA_CHECK(batch_size > 0, "batch_size must be positive");
  FA_CHECK(head_size % 8 == 0 && head_size <= 128, "head_size must be a multiple of 8 and <= 128");
  FA_CHECK(cu_seqlens_q && cu_seqlens_k, "cu_seqlens_q and cu_seqlens_k must be given");
  FA_CHECK(p_dropout >= 0.f && p_dropout < 1.f, "p_dropout must be in [0, 1)");
  if (attn_bias || attn_mask)   // the reference adds both before its own scaling and requires scale 1 (ibid.:344-347)
    FA_CHECK(softmax_scale == 1.0f, "softmax_scale must be 1 when attn_bias or attn_mask is given");
  const int sq_r = (max_seqlen_q_ + 15) / 16 * 16;   // row stride of softmax_lse / dsoftmax (ibid.:318)
  Params p = base_params(batch_size, sq_r, max_seqlen_k_, sq_r, (max_seqlen_k_ + 127) / 128 * 128, num_heads, num_heads,
                         head_size, (head_size + 31) / 32 * 32, softmax_scale, false, is_bf16);
  p.total_q = total_q; p.total_k = total_k;
  p.cu_seqlens_q = cu_seqlens_q; p.cu_seqlens_k = cu_seqlens_k;
  p.p_dropout = p_dropout; p.philox_seed = seed; p.philox_offset = offset;
  p.lse_empty_is_neg_inf = 1;                        // FA1 leaves -inf for rows without keys (ibid.:355)
  p.force_generic = 1;
  p.unscale_softmax = 1.f;
  p.bias_rows = max_seqlen_q_;
  p.bias_batch_mod = 1;
  if (attn_mask) {
    FA_CHECK(mask_dims != nullptr, "mask_dims must be given with attn_mask");
    FA_CHECK(mask_dims[0] == batch_size && mask_dims[3] == max_seqlen_k_, "mask_dims must be (batch, 1 or heads, 1 or max_seqlen_q, max_seqlen_k)");
    FA_CHECK(mask_dims[1] == 1 || mask_dims[1] == num_heads, "mask_dims[1] must be 1 or num_heads");
    FA_CHECK(mask_dims[2] == 1 || mask_dims[2] == max_seqlen_q_, "mask_dims[2] must be 1 or max_seqlen_q");
    p.attn_mask = attn_mask;
    p.mask_head_mod = (int)mask_dims[1]; p.mask_seq_mod = (int)mask_dims[2];
  }
  if (attn_bias) {
    FA_CHECK(bias_dims != nullptr, "bias_dims must be given with attn_bias");
    FA_CHECK(bias_dims[1] == num_heads, "bias_dims[1] must equal num_heads");
    FA_CHECK(bias_dims[0] >= 1, "bias_dims[0] must be positive");   // batch element b reads bias[b % bias_dims[0]]
    FA_CHECK(bias_dims[2] == max_seqlen_q_ && bias_dims[3] == max_seqlen_k_, "bias_dims must end in (max_seqlen_q, max_seqlen_k)");
    p.attn_bias = attn_bias;
    p.bias_batch_mod = (int)bias_dims[0];
  }
  return p;
}

}  // namespace

#define FA_BEGIN try {
#define FA_END                                                \
  }                                                           \
  catch (const std::exception &e) {                           \
    flash_attn_set_error(e.what());                           \
    return false;                                             \
  }                                                           \
  catch (...) {                                               \
    flash_attn_set_error(nullptr);                            \
    return false;                                             \
  }                                                           \
  return true;

extern "C" {

bool flash_attn_fwd(
    const void *q, const void *k, const void *v, void *rng_state, void *out, void *softmax_ptr,
    void *softmax_lse_ptr, int batch_size, int seqlen_q, int seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
    int num_heads, int num_heads_k, int head_size, int head_size_rounded, float p_dropout, float softmax_scale,
    float softmax_unscale, bool is_causal, bool return_softmax, bool is_bf16, cudaStream_t stream, uint64_t seed,
    uint64_t offset, const void *attn_mask, const int64_t *mask_dims, const void *flashmask_downstart_ptr,
    const int64_t *flashmask_dims, const void *flashmask_upend_ptr, const void *flashmask_downend_ptr,
    const void *flashmask_upstart_ptr, const void *flashmask_maxmin_ptr, int q_row_stride, int k_row_stride,
    int v_row_stride, int q_head_stride, int k_head_stride, int v_head_stride, int o_row_stride, int o_head_stride,
    int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride) {
  FA_BEGIN
  Params p = base_params(batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded, num_heads, num_heads_k,
                         head_size, head_size_rounded, softmax_scale, is_causal, is_bf16);
  set_attn_mask(p, attn_mask, mask_dims, softmax_unscale);
  p.p_dropout = p_dropout; p.philox_seed = seed; p.philox_offset = offset;   // csrc/capi/flash_attn.cu:672-677
  if (return_softmax) {
    FA_CHECK(softmax_ptr != nullptr, "return_softmax needs softmax_ptr");
    p.s_dmask = softmax_ptr;
  }
  p.q = tensor(q, q_batch_stride, q_row_stride, q_head_stride);
  p.k = tensor(k, k_batch_stride, k_row_stride, k_head_stride);
  p.v = tensor(v, v_batch_stride, v_row_stride, v_head_stride);
  p.o = tensor(out, o_batch_stride, o_row_stride, o_head_stride);
  p.softmax_lse = (float *)softmax_lse_ptr;
  set_flashmask(p, flashmask_downstart_ptr, flashmask_dims, flashmask_upend_ptr, flashmask_downend_ptr,
                flashmask_upstart_ptr, flashmask_maxmin_ptr);
  if (rng_state) {  // the reference records (seed, offset) for backward replay (flash_attn.cu:672-677)
    uint64_t st[2] = {seed, offset};
    FA_CUDA(cudaMemcpyAsync(rng_state, st, sizeof(st), cudaMemcpyHostToDevice, stream));
  }
  (void)softmax_ptr;
  fa::run_fwd(p, stream);
  FA_END
}

bool flash_attn_varlen_fwd(
    const void *q, const void *k, const void *v, const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k,
    void *rng_state, void *out, void *softmax_ptr, void *softmax_lse_ptr, int batch_size, int max_seqlen_q,
    int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size,
    int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale, bool is_causal,
    bool return_softmax, bool is_bf16, cudaStream_t stream, uint64_t seed, uint64_t offset, const void *attn_mask,
    const void *mask_dims, int q_row_stride, int k_row_stride, int v_row_stride, int q_head_stride, int k_head_stride,
    int v_head_stride, int o_row_stride, int o_head_stride, int q_batch_stride, int k_batch_stride, int v_batch_stride,
    int o_batch_stride, bool varlen_padded_input) {
  FA_BEGIN
  FA_CHECK(cu_seqlens_q && cu_seqlens_k, "cu_seqlens_q and cu_seqlens_k must be given");
  Params p = base_params(batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded, num_heads,
                         num_heads_k, head_size, head_size_rounded, softmax_scale, is_causal, is_bf16);
  // mask rows are local to each sequence, its row stride is max_seqlen_k (flash_fwd_kernel.h:1104-1107 with the varlen
  // params.seqlen_k = max_seqlen_k, csrc/capi/flash_attn.cu:730-733)
  set_attn_mask(p, attn_mask, (const int64_t *)mask_dims, softmax_unscale);
  p.p_dropout = p_dropout; p.philox_seed = seed; p.philox_offset = offset;
  if (return_softmax) {
    FA_CHECK(softmax_ptr != nullptr, "return_softmax needs softmax_ptr");
    p.s_dmask = softmax_ptr;
  }
  p.q = tensor(q, q_batch_stride, q_row_stride, q_head_stride);
  p.k = tensor(k, k_batch_stride, k_row_stride, k_head_stride);
  p.v = tensor(v, v_batch_stride, v_row_stride, v_head_stride);
  p.o = tensor(out, o_batch_stride, o_row_stride, o_head_stride);
  p.softmax_lse = (float *)softmax_lse_ptr;
  p.cu_seqlens_q = cu_seqlens_q; p.cu_seqlens_k = cu_seqlens_k;
  p.varlen_padded_input = varlen_padded_input ? 1 : 0;
  resolve_varlen_totals(p, stream);
  if (rng_state) {
    uint64_t st[2] = {seed, offset};
    FA_CUDA(cudaMemcpyAsync(rng_state, st, sizeof(st), cudaMemcpyHostToDevice, stream));
  }
  (void)softmax_ptr;
  fa::run_fwd(p, stream);
  FA_END
}

bool flash_attn_bwd(
    const void *dout, const void *q, const void *k, const void *v, const void *out, const void *softmax_d,
    const void *softmax_lse, void *rng_state, void *dq, void *dk, void *dv, void *dq_accum, int batch_size,
    int seqlen_q, int seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k,
    int head_size, int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale, bool is_causal,
    bool is_bf16, int num_splits, cudaStream_t stream, uint64_t seed, uint64_t offset, const void *attn_mask,
    const int64_t *mask_dims, const void *flashmask_downstart_ptr, const int64_t *flashmask_dims,
    const void *flashmask_upend_ptr, const void *flashmask_downend_ptr, const void *flashmask_upstart_ptr,
    const void *flashmask_maxmin_ptr, int q_row_stride, int k_row_stride, int v_row_stride, int q_head_stride,
    int k_head_stride, int v_head_stride, int o_row_stride, int o_head_stride, int q_batch_stride, int k_batch_stride,
    int v_batch_stride, int o_batch_stride, int dq_row_stride, int dk_row_stride, int dv_row_stride,
    int dq_head_stride, int dk_head_stride, int dv_head_stride, int do_row_stride, int do_head_stride,
    int dq_batch_stride, int dk_batch_stride, int dv_batch_stride, int do_batch_stride) {
  FA_BEGIN
  Params p = base_params(batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded, num_heads, num_heads_k,
                         head_size, head_size_rounded, softmax_scale, is_causal, is_bf16);
  set_attn_mask(p, attn_mask, mask_dims, softmax_unscale);
  p.p_dropout = p_dropout; p.philox_seed = seed; p.philox_offset = offset;
  p.q = tensor(q, q_batch_stride, q_row_stride, q_head_stride);
  p.k = tensor(k, k_batch_stride, k_row_stride, k_head_stride);
  p.v = tensor(v, v_batch_stride, v_row_stride, v_head_stride);
  p.o = tensor(out, o_batch_stride, o_row_stride, o_head_stride);
  p.dout = tensor(dout, do_batch_stride, do_row_stride, do_head_stride);
  p.dq = tensor(dq, dq_batch_stride, dq_row_stride, dq_head_stride);
  p.dk = tensor(dk, dk_batch_stride, dk_row_stride, dk_head_stride);
  p.dv = tensor(dv, dv_batch_stride, dv_row_stride, dv_head_stride);
  p.softmax_lse = (float *)const_cast<void *>(softmax_lse);
  p.softmax_d = (float *)const_cast<void *>(softmax_d);
  p.dq_accum = (float *)dq_accum;
  p.dkv_expanded = 1;                       // csrc/flash_attn/src/flash_bwd_kernel.h:1077-1080
  p.deterministic = (num_splits == 1);      // csrc/flash_attn/src/flash_bwd_kernel.h:3203-3208
  set_flashmask(p, flashmask_downstart_ptr, flashmask_dims, flashmask_upend_ptr, flashmask_downend_ptr,
                flashmask_upstart_ptr, flashmask_maxmin_ptr);
  (void)rng_state;
  fa::run_bwd(p, stream);
  FA_END
}

bool flash_attn_varlen_bwd(
    const void *dout, const void *q, const void *k, const void *v, const void *out, const void *softmax_d,
    const void *softmax_lse, const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state, void *dq,
    void *dk, void *dv, void *dq_accum, int batch_size, int max_seqlen_q, int max_seqlen_k, int seqlen_q_rounded,
    int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size, int head_size_rounded, float p_dropout,
    float softmax_scale, float softmax_unscale, bool is_causal, bool is_bf16, int num_splits, cudaStream_t stream,
    uint64_t seed, uint64_t offset, const void *attn_mask, const int64_t *mask_dims, int q_row_stride,
    int k_row_stride, int v_row_stride, int q_head_stride, int k_head_stride, int v_head_stride, int o_row_stride,
    int o_head_stride, int q_batch_stride, int k_batch_stride, int v_batch_stride, int o_batch_stride,
    int dq_row_stride, int dk_row_stride, int dv_row_stride, int dq_head_stride, int dk_head_stride,
    int dv_head_stride, int do_row_stride, int do_head_stride, int dq_batch_stride, int dk_batch_stride,
    int dv_batch_stride, int do_batch_stride, bool varlen_padded_input) {
  FA_BEGIN
  FA_CHECK(cu_seqlens_q && cu_seqlens_k, "cu_seqlens_q and cu_seqlens_k must be given");
  Params p = base_params(batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded, num_heads,
                         num_heads_k, head_size, head_size_rounded, softmax_scale, is_causal, is_bf16);
  // mask rows are local to each sequence, its row stride is max_seqlen_k (flash_fwd_kernel.h:1104-1107 with the varlen
  // params.seqlen_k = max_seqlen_k, csrc/capi/flash_attn.cu:730-733)
  set_attn_mask(p, attn_mask, (const int64_t *)mask_dims, softmax_unscale);
  p.p_dropout = p_dropout; p.philox_seed = seed; p.philox_offset = offset;
  p.q = tensor(q, q_batch_stride, q_row_stride, q_head_stride);
  p.k = tensor(k, k_batch_stride, k_row_stride, k_head_stride);
  p.v = tensor(v, v_batch_stride, v_row_stride, v_head_stride);
  p.o = tensor(out, o_batch_stride, o_row_stride, o_head_stride);
  p.dout = tensor(dout, do_batch_stride, do_row_stride, do_head_stride);
  p.dq = tensor(dq, dq_batch_stride, dq_row_stride, dq_head_stride);
  p.dk = tensor(dk, dk_batch_stride, dk_row_stride, dk_head_stride);
  p.dv = tensor(dv, dv_batch_stride, dv_row_stride, dv_head_stride);
  p.softmax_lse = (float *)const_cast<void *>(softmax_lse);
  p.softmax_d = (float *)const_cast<void *>(softmax_d);
  p.dq_accum = (float *)dq_accum;
  p.cu_seqlens_q = cu_seqlens_q; p.cu_seqlens_k = cu_seqlens_k;
  p.varlen_padded_input = varlen_padded_input ? 1 : 0;
  resolve_varlen_totals(p, stream);
  p.dq_accum_batch_major = 1;               // the caller's dq_accum is (b, h, seqlen_q_rounded, d_rounded): flash_api.cpp:836-838
  p.dkv_expanded = 1;
  p.deterministic = (num_splits == 1);
  (void)rng_state; (void)seed; (void)offset;
  fa::run_bwd(p, stream);
  FA_END
}

bool calc_reduced_attn_scores(
    const void *q, const void *k, const void *softmax_lse, void *reduced_scores, void *softmax_ptr, int batch_size,
    int seqlen_q, int seqlen_k, int num_heads, int num_heads_k, int head_size, float softmax_scale,
    bool return_softmax, bool is_bf16, int num_splits, cudaStream_t stream, int q_row_stride, int k_row_stride,
    int o_row_stride, int q_head_stride, int k_head_stride, int o_head_stride, int q_batch_stride,
    int k_batch_stride, int o_batch_stride) {
  FA_BEGIN
  FA_CHECK(q && k && softmax_lse && reduced_scores, "calc_reduced_attn_scores: null tensor");
  FA_CHECK(!return_softmax || softmax_ptr != nullptr, "calc_reduced_attn_scores: return_softmax needs softmax_ptr");
  FA_CHECK(head_size % 8 == 0 && head_size <= 256 && num_heads % num_heads_k == 0, "unsupported head configuration");
  fa::reduced_scores(q, k, (const float *)softmax_lse, (float *)reduced_scores,
                     return_softmax ? (float *)softmax_ptr : nullptr, batch_size, seqlen_q, seqlen_k, num_heads,
                     num_heads_k, head_size, softmax_scale, is_bf16, q_batch_stride, q_row_stride, q_head_stride,
                     k_batch_stride, k_row_stride, k_head_stride, stream);
  (void)num_splits; (void)o_row_stride; (void)o_head_stride; (void)o_batch_stride;
  FA_END
}

// ---- legacy FA1 symbols: packed varlen tensors (total, h, d), no GQA, no causal, additive mask and bias, dbias ----
// csrc/flash_attn_with_bias_and_mask/flash_attn_with_bias_mask.cu:268-397 (fwd), 400-541 (bwd).  softmax_lse and
// dsoftmax are (b, h, round_up(max_seqlen_q, 16)) fp32 as there.  The reference strides mask / bias rows by each
// sequence's own length, which describes a dense tensor only when all lengths equal the maximum; here the row stride
// is mask_dims[3] / bias_dims[3] = max_seqlen_k, identical for uniform batches and well defined for ragged ones.
// These run on the generic kernels (the bias path reads and writes O(s^2) tensors; it is HBM-bound by construction).
bool flash_attn_fwd_with_bias_and_mask(
    const void *q, const void *k, const void *v, void *out, const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k,
    int total_q, int total_k, int batch_size, int num_heads, int head_size, int max_seqlen_q_, int max_seqlen_k_,
    float p_dropout, float softmax_scale, bool zero_tensors, bool is_bf16, int num_splits, void *softmax_lse_ptr,
    void *workspace_ptr, uint64_t *workspace_size, cudaStream_t stream, uint64_t seed, uint64_t offset,
    const void *attn_mask, const void *attn_bias, const int64_t *mask_dims, const int64_t *bias_dims) {
  FA_BEGIN
  if (out == nullptr) {                     // workspace query (flash_attn_with_bias_mask.cu:321-330): none needed here
    FA_CHECK(workspace_size != nullptr, "workspace_size must be given for the workspace query");
    *workspace_size = 0;
    return true;
  }
  Params p = fa1_params(batch_size, num_heads, head_size, max_seqlen_q_, max_seqlen_k_, total_q, total_k, p_dropout,
                        softmax_scale, is_bf16, seed, offset, cu_seqlens_q, cu_seqlens_k, attn_mask, attn_bias,
                        mask_dims, bias_dims);
  p.q = tensor(q, 0, num_heads * head_size, head_size);
  p.k = tensor(k, 0, num_heads * head_size, head_size);
  p.v = tensor(v, 0, num_heads * head_size, head_size);
  p.o = tensor(out, 0, num_heads * head_size, head_size);
  FA_CHECK(softmax_lse_ptr != nullptr, "softmax_lse_ptr must be given");
  p.softmax_lse = (float *)softmax_lse_ptr;
  if (zero_tensors) {                       // flash_attn_with_bias_mask.cu:353-356
    FA_CUDA(cudaMemsetAsync(out, 0, (size_t)total_q * num_heads * head_size * 2, stream));
    fa::fill_f32((float *)softmax_lse_ptr, -INFINITY, (int64_t)batch_size * num_heads * p.seqlen_q, stream);
  }
  (void)workspace_ptr; (void)num_splits;
  fa::run_fwd(p, stream);
  FA_END
}

bool flash_attn_bwd_with_bias_and_mask(
    const void *q, const void *k, const void *v, void *dq, void *dk, void *dv, const void *out, const void *dout,
    const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, int total_q, int total_k, int batch_size, int num_heads,
    int head_size, int max_seqlen_q_, int max_seqlen_k_, float p_dropout, float softmax_scale, bool zero_tensors,
    bool is_bf16, int num_splits, const void *softmax_lse_ptr, void *dsoftmax_ptr, void *dbias_ptr, void *workspace_ptr,
    uint64_t *workspace_size, cudaStream_t stream, uint64_t seed, uint64_t offset, const void *attn_mask,
    const void *attn_bias, const int64_t *mask_dims, const int64_t *bias_dims) {
  FA_BEGIN
  if (out == nullptr) {                     // workspace query (flash_attn_with_bias_mask.cu:456-468)
    FA_CHECK(workspace_size != nullptr, "workspace_size must be given for the workspace query");
    *workspace_size = 0;
    return true;
  }
  Params p = fa1_params(batch_size, num_heads, head_size, max_seqlen_q_, max_seqlen_k_, total_q, total_k, p_dropout,
                        softmax_scale, is_bf16, seed, offset, cu_seqlens_q, cu_seqlens_k, attn_mask, attn_bias,
                        mask_dims, bias_dims);
  FA_CHECK(dq && dk && dv && dout && softmax_lse_ptr && dsoftmax_ptr, "backward tensors must be given");
  const int rs = num_heads * head_size;
  p.q = tensor(q, 0, rs, head_size); p.k = tensor(k, 0, rs, head_size); p.v = tensor(v, 0, rs, head_size);
  p.o = tensor(out, 0, rs, head_size); p.dout = tensor(dout, 0, rs, head_size);
  p.dq = tensor(dq, 0, rs, head_size); p.dk = tensor(dk, 0, rs, head_size); p.dv = tensor(dv, 0, rs, head_size);
  p.softmax_lse = (float *)const_cast<void *>(softmax_lse_ptr);
  p.softmax_d = (float *)dsoftmax_ptr;
  p.dkv_expanded = 1;
  if (attn_bias) {                          // flash_attn_with_bias_mask.cu:471-476: dbias is (b, h, sq, sk), zero-filled
    FA_CHECK(dbias_ptr != nullptr, "dbias_ptr must be given with attn_bias");
    FA_CUDA(cudaMemsetAsync(dbias_ptr, 0, (size_t)batch_size * num_heads * max_seqlen_q_ * max_seqlen_k_ * 2, stream));
    p.ds_out = dbias_ptr;
  }
  if (zero_tensors) {                       // flash_attn_with_bias_mask.cu:493-498
    FA_CUDA(cudaMemsetAsync(dq, 0, (size_t)total_q * rs * 2, stream));
    FA_CUDA(cudaMemsetAsync(dk, 0, (size_t)total_k * rs * 2, stream));
    FA_CUDA(cudaMemsetAsync(dv, 0, (size_t)total_k * rs * 2, stream));
    FA_CUDA(cudaMemsetAsync(dsoftmax_ptr, 0, (size_t)batch_size * num_heads * p.seqlen_q * 4, stream));
  }
  (void)workspace_ptr; (void)num_splits;
  fa::run_bwd(p, stream);
  FA_END
}

}  // extern "C"
