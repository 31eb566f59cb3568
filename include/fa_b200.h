/*
 * fa_b200.h -- the C ABI of libflashattn_b200.so, part 2: the descriptor-style entry points.
 *
 * The flat FA2 symbols in fa_capi.h cannot express a sliding window, bottom-right causal
 * alignment, FlashMask vectors on the varlen path, a KV cache or split-KV partials.  In the
 * reference those live behind the setter-style handles of csrc/flashmask_v2/flash_api.h:14-266 and
 * csrc/flash_attn_v3/flash_api.h (one getter+setter per field of csrc/flashmask_v2/flash.h:13-231).
 * Here the same fields are one plain C struct (pointers and sizes only) that the caller fills and
 * passes by pointer -- the cgo / ctypes / dlsym friendly form of the same contract.
 *
 * Error convention: identical to fa_capi.h (bool return + flash_attn_error()).
 */
#ifndef FA_B200_H_
#define FA_B200_H_

#include <stdint.h>
#include <stdbool.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FA_B200_ABI_VERSION 6

/* One strided (batch, row, head, d) view; strides in ELEMENTS, innermost stride 1.
 * For packed varlen tensors batch_stride is ignored (rows are indexed through cu_seqlens). */
typedef struct fa_b200_tensor {
  void   *ptr;
  int64_t batch_stride;
  int64_t row_stride;
  int64_t head_stride;
} fa_b200_tensor;

/* Mirrors Flash_fwd_params / Flash_bwd_params (reference: csrc/flash_attn/src/flash.h:30-168 and
 * csrc/flashmask_v2/flash.h:13-231), with 64-bit strides throughout. */
typedef struct fa_b200_params {
  int32_t abi_version;            /* = FA_B200_ABI_VERSION */
  int32_t is_bf16;                /* 1 = bf16, 0 = fp16 */

  /* ---- tensors ---- */
  fa_b200_tensor q, k, v, o;      /* o: forward output / backward input */
  fa_b200_tensor dout, dq, dk, dv;/* backward only; dk/dv have h (expanded) heads iff dkv_expanded */
  float *softmax_lse;             /* (b, h, seqlen_q) fp32; varlen: (b, h, max_seqlen_q) */
  float *softmax_d;               /* bwd workspace (b, h, seqlen_q_rounded) fp32 */
  float *dq_accum;                /* bwd workspace (b, h, seqlen_q_rounded, d_rounded) fp32 */
  float *dkv_accum;               /* reserved (the reference's GQA dk/dv fp32 workspace): unused, GQA groups are reduced in TMEM */

  /* ---- sizes ---- */
  int32_t b, seqlen_q, seqlen_k;  /* varlen: max_seqlen_q / max_seqlen_k */
  int32_t h, h_k, d;
  int32_t seqlen_q_rounded, seqlen_k_rounded, d_rounded;
  int64_t total_q, total_k;       /* varlen: number of packed rows if known, else -1 */

  /* ---- varlen / cache lengths ---- */
  const int32_t *cu_seqlens_q;    /* (b+1) or NULL */
  const int32_t *cu_seqlens_k;    /* (b+1) or NULL */
  const int32_t *seqused_k;       /* (b) valid keys per batch element (kv cache), or NULL */
  int32_t varlen_padded_input;    /* cu_seqlens give lengths, rows addressed by batch_stride */

  /* ---- softmax ---- */
  float   softmax_scale;
  int32_t lse_empty_is_neg_inf;   /* 0: +inf for rows with no visible key (FA2 surface); 1: -inf (v2/v3) */

  /* ---- masks ---- */
  int32_t is_causal;
  int32_t causal_bottom_right;    /* 0: top-left (flash_attn_*func, FA2 capi); 1: bottom-right (v2/v3, kvcache) */
  int32_t window_left, window_right; /* -1 = unbounded; bottom-right aligned (csrc/flashmask_v2/mask.h:101-118) */
  const int32_t *lt_start, *lt_end, *ut_start, *ut_end; /* FlashMask vectors (b, h_mask, seqlen_k) int32 or NULL */
  int32_t h_mask;                 /* mask heads; q-head hq uses mask head hq / (h / h_mask) */
  int32_t *flashmask_maxmin;      /* workspace 8 * b * h_mask * ceil(seqlen_k/64 rounded to 4) int32, or NULL */

  /* ---- behaviour ---- */
  int32_t dkv_expanded;           /* bwd: dk/dv are written per q-head (FA2 capi contract) */
  int32_t deterministic;          /* bwd: serialise dQ accumulation */
  int32_t force_generic;          /* testing: take the generic SIMT kernels instead of the tcgen05 ones */

  /* ---- split-KV (decode) ---- */
  int32_t num_splits;             /* 0/1 = no split */
  float  *oaccum;                 /* (num_splits, b, seqlen_q, h, d) fp32 */
  float  *lseaccum;               /* (num_splits, b, h, seqlen_q) fp32 */

  /* ---- dropout (ABI v2; reference: csrc/flash_attn/src/flash.h:88-94, softmax.h:381-447) ---- */
  float    p_dropout;             /* probability of DROPPING an attention weight; 0 = off */
  uint64_t philox_seed, philox_offset; /* keep iff philox_byte(seed, offset, b, h, row, col) <= floor(255 * (1 - p)) */
  void    *s_dmask;               /* optional (b, h, seqlen_q_rounded, seqlen_k_rounded) in the element type: softmax
                                     probabilities, NEGATED where the weight was dropped (return_softmax), or NULL */

  /* ---- dense additive mask (reference: csrc/flash_attn/src/softmax.h:360-379, flash_fwd_kernel.h:1104-1107) ---- */
  const void *attn_mask;          /* (b, mask_head_mod, mask_seq_mod, seqlen_k) in the element type, or NULL:
                                     logits = scale * (q.k + mask[b, h % head_mod, row % seq_mod, col] * unscale_softmax) */
  int32_t mask_head_mod, mask_seq_mod;
  float   unscale_softmax;

  /* ---- KV-cache addressing, softcap, ALiBi (ABI v3; reference: csrc/flash_attn_v3/flash.h:96-140, seqlen.h:56-91,
   *      paged_kv.h, mainloop_fwd_sm90_tma_gmma_ws.hpp:537-556,639,1064-1067).  Forward only, except softcap/alibi. ---- */
  const int32_t *page_table;      /* (b, page_table_batch_stride) int32 or NULL.  When set, k/v are
                                     (num_pages, page_size, h_k, d): batch_stride is the PAGE stride, and key t of batch
                                     element i is row t % page_size of page page_table[i][t / page_size];
                                     seqlen_k = pages per sequence * page_size */
  int64_t page_table_batch_stride;
  int32_t page_size, num_pages;
  const int32_t *kv_batch_idx;    /* (b) cache row used by batch element i, or NULL (= i); not with page_table */
  int32_t b_k;                    /* batch size of the cache when kv_batch_idx is set (0 = b) */
  const int32_t *leftpad_k;       /* (b) index of the first valid key; seqused_k counts from index 0, so the valid keys
                                     are [leftpad_k[i], seqused_k[i]) (seqlen.h:73-80) */
  float   softcap;                /* > 0: logits = softcap * tanh(scale * q.k / softcap)  (before masks and ALiBi) */
  const float *alibi_slopes;      /* (h) or (b, h) fp32 or NULL: logits -= slope * |col - (row + seqlen_k - seqlen_q)| */
  int64_t alibi_batch_stride;     /* 0 for (h) */

  /* ---- dense additive bias with its gradient (ABI v4; the legacy flash_attn_*_with_bias_and_mask symbols:
   *      csrc/flash_attn_with_bias_and_mask/src/fmha/gmem_tile.h:735-760,880-925, softmax.h:480-498) ---- */
  const void *attn_bias;          /* (bias_batch_mod, h, bias_rows, seqlen_k) in the element type, or NULL:
                                     logits += bias[b % bias_batch_mod, h, row, col] * unscale_softmax */
  int32_t bias_batch_mod;         /* 1 (shared by the batch) or b */
  int32_t bias_rows;              /* row count of attn_bias / ds_out (the unrounded max query length) */
  void   *ds_out;                 /* backward, optional: (b, h, bias_rows, seqlen_k) in the element type, receives
                                     dS = P * (dP - dPsum) for every visible (row, col); the caller zero-fills it and
                                     reduces it over the batch when the bias is shared.  Generic kernels only. */
  int32_t dq_accum_batch_major;   /* packed varlen backward: 1 = dq_accum is sized (b, h, seqlen_q_rounded, d_rounded) as
                                     the flat varlen symbol's callers allocate it (csrc/flash_attn/flash_api.cpp:836-838);
                                     0 = the compact h * floor((total_q + 128 * b) / 128) tiles */
  const int32_t *seqused_q;       /* (b) int32 or NULL: only the first seqused_q[i] query rows of batch element i are
                                     attended (csrc/flashmask_v2/seqlen.h:17-53); rows beyond are left untouched in out,
                                     lse, dq.  Bottom-right alignment then uses seqused_q[i] as the query length. */
  /* ---- per-batch split counts (ABI v5; reference: num_splits_dynamic_ptr, csrc/flash_attn_v3/flash.h:96-125,
   *      filled by prepare_varlen_num_blocks, flash_prepare_scheduler.cu:15-124) ---- */
  const int32_t *num_splits_dynamic; /* (b) int32 or NULL: batch element i is cut into num_splits_dynamic[i] <= num_splits
                                        key ranges; the partial results of the remaining splits are written as empty
                                        (lse = -inf) and the combine skips them.  fa_b200_prepare_num_splits fills it. */
  /* ---- FP8 forward (ABI v6; reference: is_e4m3 + *_descale_ptr, csrc/flashmask_v2/flash.h:55-63,144 and
   *      flash_api.cu:107-131) ---- */
  int32_t is_e4m3;                /* 1 = q, k, v hold e4m3 bytes (strides in elements = bytes); out is bf16; forward only */
  const float *q_descale;         /* per (batch, kv head) fp32 factors or NULL (= 1): logits are scaled by            */
  const float *k_descale;         /*   q_descale * k_descale (mainloop_fwd_sm90_tma_gmma_ws.hpp:1612-1614), out by    */
  const float *v_descale;         /*   v_descale (:1868-1869)                                                         */
  int64_t q_descale_batch_stride, q_descale_head_stride;
  int64_t k_descale_batch_stride, k_descale_head_stride;
  int64_t v_descale_batch_stride, v_descale_head_stride;
} fa_b200_params;

/* forward: O, LSE.  Equivalent of {flashmaskv2,fa3}_run_mha_fwd (csrc/flashmask_v2/flash_api.h). */
bool fa_b200_fwd(const fa_b200_params *p, cudaStream_t stream);

/* backward: dQ, dK, dV from dO, O, LSE.  Equivalent of flashmaskv2_run_mha_bwd. */
bool fa_b200_bwd(const fa_b200_params *p, cudaStream_t stream);

/* split-KV decode forward: writes oaccum/lseaccum when num_splits > 1 (then call fa_b200_fwd_combine),
 * or o/softmax_lse directly when num_splits <= 1. */
bool fa_b200_fwd_splitkv(const fa_b200_params *p, cudaStream_t stream);

/* LSE-combine of split partials.  Equivalent of fa3_run_mha_fwd_combine
 * (csrc/flash_attn_v3/flash_fwd_combine_kernel.h:194-478). */
bool fa_b200_fwd_combine(const fa_b200_params *p, cudaStream_t stream);

/* Split-KV decode + combine in ONE call that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  Same params as fa_b200_fwd_splitkv, with oaccum and lseaccum NULL: the split partials live
 * in a static __device__ array of FA_B200_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it until
 * the process exits); writes o and softmax_lse.  A call needs num_splits * b * seqlen_q * h * (d + 1) * 4 bytes of it and
 * is rejected (false + flash_attn_error()) when that exceeds the array.  cu_seqlens_*, FlashMask, num_splits_dynamic and
 * is_e4m3 are rejected; seqused_k, paged / indexed / left-padded caches, causal / window masks are taken as usual.
 * Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device share the array: calls that overlap on the
 * GPU (different streams or threads) race -- keep them on one stream.  Nothing in the array needs clearing between
 * calls (every partial a call reads it has written first). */
#define FA_B200_NOALLOC_SCRATCH_BYTES (16u << 20)
bool fa_b200_fwd_splitkv_noalloc(const fa_b200_params *p, cudaStream_t stream);

/* SMs the CALLING THREAD's next persistent forward launches leave free (default 0).  The forward kernels run one CTA per SM
 * over a statically cut tile list; when a collective's kernels hold some SMs while such a launch starts (ring attention:
 * NCCL's send / recv of the next K/V block), the CTAs that cannot become resident start late and the whole launch ends
 * late by as much.  Reserving those SMs costs n / 148 of the forward's throughput and removes the delay (measured on 8
 * GPUs, 16 384 tokens per rank: ring step 70.2 -> 67.3 ms with 8 reserved).  No counterpart in the reference, whose
 * forward grids are not persistent.  FlashMask launches (dynamic tile scheduler) ignore it. */
void fa_b200_set_forward_sm_reserve(int num_sms);

/* Split heuristic.  Equivalent of fa3_get_num_splits (csrc/flashmask_v2/heuristics.h:22-59). */
int fa_b200_num_splits(int batch, int num_heads_k, int seqlen_q_packed, int seqlen_k, int head_size, int num_sm);

/* Split count for the TMA-fed decode kernel of this library (decode.cu): two CTAs are resident per SM, a CTA streams
 * 128-key blocks; picks the split count that minimises waves x (blocks per split + fixed per-CTA cost).  The reference
 * heuristic above is kept for callers that want its exact answer; this one is what flash_attn_with_kvcache uses. */
int fa_b200_decode_num_splits(int batch, int num_heads_k, int seqlen_k, int num_sm);

/* Per-batch split counts for a batch of unequal lengths (cu_seqlens_k / seqused_k / leftpad_k): restates
 * prepare_varlen_num_blocks (csrc/flash_attn_v3/flash_prepare_scheduler.cu:15-124) for this library's tiles --
 * blocks_per_sm = ceil(1.1 * heads_k * sum_b m_blocks(b) * n_blocks(b) / num_sm),
 * out[b] = clamp(ceil(n_blocks(b) / blocks_per_sm), 1, p->num_splits), with 16 packed query rows per m block and 128 keys
 * per n block.  One launch of one CTA; batch <= 1024. */
bool fa_b200_prepare_num_splits(const fa_b200_params *p, int32_t *num_splits_dynamic_out, int num_sm, cudaStream_t stream);

/* Append new K/V rows into the cache at position seqused_k[b] (flash_attn_with_kvcache's k=/v= arguments;
 * reference: csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:1542 store_kv_new). */
bool fa_b200_append_kv(const void *k_new, const void *v_new, void *k_cache, void *v_cache,
                       const int32_t *cache_seqlens, int batch, int seqlen_new, int seqlen_cache,
                       int num_heads_k, int head_size, int is_bf16,
                       int64_t new_batch_stride, int64_t new_row_stride, int64_t new_head_stride,
                       int64_t cache_batch_stride, int64_t cache_row_stride, int64_t cache_head_stride,
                       cudaStream_t stream);

/* Same, for every cache layout fa_b200_params can describe: uses p->k / p->v (cache), p->page_table, p->kv_batch_idx
 * and p->seqused_k (the lengths BEFORE the append; NULL = 0); new rows go to positions seqused_k[i] .. + seqlen_new.
 * k_new / v_new: (b, seqlen_new, h_k, d). */
bool fa_b200_append_kv_paged(const fa_b200_params *p, const void *k_new, const void *v_new, int seqlen_new,
                             int64_t new_batch_stride, int64_t new_row_stride, int64_t new_head_stride,
                             cudaStream_t stream);

/* Rotary embedding, out-of-place (reference: csrc/flash_attn_v3/rotary.h; flash_attn_with_kvcache's rotary_cos /
 * rotary_sin): x, out (b, s, h, d); cos, sin (seqlen_ro, rotary_dim / 2) in the element type.  Row r of batch element
 * i is rotated at position seqlen_offsets[i] + (fixed_position ? 0 : r) (seqlen_offsets NULL = 0).  interleaved pairs
 * (2j, 2j+1), otherwise (j, j + rotary_dim/2); columns >= rotary_dim are copied.  fp32 arithmetic, one rounding. */
bool fa_b200_rotary(const void *x, void *out, const void *cos, const void *sin, const int32_t *seqlen_offsets,
                    int batch, int seqlen, int num_heads, int head_size, int rotary_dim, int seqlen_ro,
                    int interleaved, int fixed_position, int is_bf16,
                    int64_t x_batch_stride, int64_t x_row_stride, int64_t x_head_stride,
                    int64_t out_batch_stride, int64_t out_row_stride, int64_t out_head_stride, cudaStream_t stream);

/* FlashMask block min/max scan (reference: csrc/flashmask_v2/flash_mask.hpp:162-245; order
 * [lts_max, lts_min, ute_max, ute_min, lte_max, lte_min, uts_max, uts_min]). Exposed for tests. */
bool fa_b200_flashmask_prepare(const fa_b200_params *p, int block_n, cudaStream_t stream);

/* ring-attention step merge: (o, lse) <- merge((o, lse), (o_step, lse_step)) elementwise by LSE.
 * o (rows, h, d) bf16/fp16 or fp32 per o_is_fp32; lse / lse_step layouts are free: index = row * row_stride + head * head_stride
 * (the running lse is kept (rows, h); a kernel's own lse is (h, rows)).  +inf and -inf both mean "no contribution". */
bool fa_b200_lse_merge(void *o, float *lse, const void *o_step, const float *lse_step,
                       int64_t rows, int heads, int head_size, int is_bf16, int o_is_fp32,
                       int64_t lse_row_stride, int64_t lse_head_stride,
                       int64_t lse_step_row_stride, int64_t lse_step_head_stride, cudaStream_t stream);

/* The same merge over a batch in ONE launch: o / o_step rows of batch element i start at i * batch stride (elements;
 * rows inside a batch element are contiguous (rows, h, d)), lse / lse_step index = i * batch stride + row * row stride +
 * head * head stride.  Lets ring attention merge into the second half of its running (b, s, h, d) output in place. */
bool fa_b200_lse_merge_batched(void *o, float *lse, const void *o_step, const float *lse_step, int batch, int64_t rows,
                               int heads, int head_size, int is_bf16, int o_is_fp32, int64_t o_batch_stride,
                               int64_t o_step_batch_stride, int64_t lse_batch_stride, int64_t lse_row_stride,
                               int64_t lse_head_stride, int64_t lse_step_batch_stride, int64_t lse_step_row_stride,
                               int64_t lse_step_head_stride, cudaStream_t stream);

/* Per-kernel device timing for bench.py's roofline (CUDA events recorded on the launch stream around the
 * dominant kernels).  slot 0 = forward kernel, slot 1 = backward main kernel.  Process-wide; not for concurrent use. */
void fa_b200_profile_enable(int on);
float fa_b200_profile_last_ms(int slot);

/* Introspection for tests/bench. */
int fa_b200_abi_version(void);
/* which kernel family served the last call on this thread: 0 none, 1 generic SIMT, 2 tcgen05, 3 TMA-fed decode */
int fa_b200_last_kernel_kind(void);
/* number of kernels launched by this library (all threads) since the last reset */
int64_t fa_b200_launch_count(int reset);
/* Why the last fa_b200_fwd / fa_b200_bwd / flash_attn_* call of this thread left the tcgen05 kernels for the generic
 * SIMT family ("" if it did not, or if the caller asked for it with force_generic).  The same text is printed to
 * stderr once per distinct reason unless the environment variable FA_B200_QUIET is set: a drop-in must not get two
 * orders of magnitude slower without saying so.  fa_b200_fallback_count: such calls, process-wide. */
const char *fa_b200_last_fallback_reason(void);
int64_t fa_b200_fallback_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* FA_B200_H_ */
