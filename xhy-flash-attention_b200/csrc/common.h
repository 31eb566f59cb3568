// common.h -- shared host/device definitions for libflashattn_b200.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdexcept>
#include <string>

#include "../../include/fa_b200.h"

namespace fa {

using Params = fa_b200_params;

struct Error : public std::runtime_error {
  explicit Error(const std::string &m) : std::runtime_error(m) {}
};

#define FA_CHECK(cond, msg)                                                              \
  do {                                                                                   \
    if (!(cond)) throw ::fa::Error(std::string(msg) + " [" #cond "] at " __FILE__ ":" +  \
                                   std::to_string(__LINE__));                            \
  } while (0)

#define FA_CUDA(call)                                                                    \
  do {                                                                                   \
    cudaError_t e__ = (call);                                                            \
    if (e__ != cudaSuccess)                                                              \
      throw ::fa::Error(std::string("CUDA error: ") + cudaGetErrorString(e__) + " in " + \
                        #call + " at " __FILE__ ":" + std::to_string(__LINE__));         \
  } while (0)

// per-thread bookkeeping (api.cu)
void note_launch(int kind);  // kind: 1 generic SIMT, 2 tcgen05
int device_sm_count();
// optional per-kernel timing (fa_b200_profile_*): slot 0 = forward kernel, slot 1 = backward main kernel
void profile_mark(int slot, bool begin, cudaStream_t s);
int device_cc_major();
// CTAs of a persistent forward launch: every SM, minus what the calling thread reserved with fa_b200_set_forward_sm_reserve
int forward_grid_sms();

// dispatchers (api.cu)
void run_fwd(const Params &p, cudaStream_t s);
void run_bwd(const Params &p, cudaStream_t s);
void run_splitkv(const Params &p, cudaStream_t s);
void run_splitkv_noalloc(const Params &p, cudaStream_t s);   // noalloc.cu: partials in a static __device__ array

// ---- kernel families -------------------------------------------------------------------------
// generic SIMT kernels: any head_size % 8 == 0 up to 256, every mask kind, both dtypes.
void generic_fwd(const Params &p, cudaStream_t s);
void generic_bwd(const Params &p, cudaStream_t s);
void generic_splitkv_fwd(const Params &p, cudaStream_t s);
void reduced_scores(const void *q, const void *k, const float *lse, float *out, float *probs, int b, int sq, int sk, int h,
                    int h_k, int d, float scale, int is_bf16, int64_t qbs, int64_t qrs, int64_t qhs, int64_t kbs,
                    int64_t krs, int64_t khs, cudaStream_t s);
// tcgen05 / TMA / TMEM kernels (sm_100a): return false when the shape is outside their envelope.
bool sm100_fwd_supported(const Params &p, std::string *why);
void sm100_fwd(const Params &p, cudaStream_t s);
bool sm100_fwd2_eligible(const Params &p);     // split-row forward (sm100_fwd2.cu): plain calls, head dims up to 256
void sm100_fwd2(const Params &p, cudaStream_t s);
void sm100_fwd2_e4m3(const Params &p, cudaStream_t s);   // FP8 forward: e4m3 q/k/v, bf16 out (same kernel, byte operands)
bool sm100_probs_eligible(const Params &p);    // return_softmax pass (sm100_probs.cu): S_dmask from Q, K and the LSE
void sm100_probs(const Params &p, cudaStream_t s);
bool sm100_bwd_supported(const Params &p, std::string *why);
void sm100_bwd(const Params &p, cudaStream_t s);
bool sm100_bwd_wide_eligible(const Params &p);   // head dims 129..256 (sm100_bwd_wide.cu): plain calls only
void sm100_bwd_wide(const Params &p, cudaStream_t s);
int64_t bwd_pre_dpsum(const Params &p, cudaStream_t s);        // shared pre / post passes of the backward (sm100_bwd.cu)
void bwd_post_convert_dq(const Params &p, cudaStream_t s);
// bandwidth kernels
void splitkv_fwd(const Params &p, cudaStream_t s);
void combine_fwd(const Params &p, cudaStream_t s);
void flashmask_prepare(const Params &p, int block_n, cudaStream_t s, bool with_l2 = false);
void append_kv(const void *k_new, const void *v_new, void *k_cache, void *v_cache, const int32_t *cache_seqlens,
               int batch, int seqlen_new, int seqlen_cache, int h_k, int d, int is_bf16,
               int64_t nbs, int64_t nrs, int64_t nhs, int64_t cbs, int64_t crs, int64_t chs, cudaStream_t s);
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device setting: remember it per (kernel instantiation, device)
// so that one process driving several GPUs (Paddle's usual mode) sets it on each of them.  `done` is the caller's
// function-local static array, one flag per device ordinal.
template <typename Kern> inline void ensure_dynamic_smem(Kern kern, int bytes, std::atomic<bool> (&done)[64]) {
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && done[dev].load(std::memory_order_acquire)) return;
  FA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (dev >= 0 && dev < 64) done[dev].store(true, std::memory_order_release);
}

void fill_f32(float *dst, float value, int64_t n, cudaStream_t s);
void prepare_num_splits(const Params &p, int32_t *out, int num_sm, cudaStream_t s);
void append_kv_paged(const Params &p, const void *k_new, const void *v_new, int seqlen_new, int64_t nbs, int64_t nrs,
                     int64_t nhs, cudaStream_t s);
void rotary(const void *x, void *out, const void *cosp, const void *sinp, const int32_t *offsets, int batch, int seqlen,
            int heads, int d, int rdim, int seqlen_ro, int interleaved, int fixed, int is_bf16, int64_t xbs, int64_t xrs,
            int64_t xhs, int64_t obs, int64_t ors, int64_t ohs, cudaStream_t s);
void lse_merge(void *o, float *lse, const void *o_step, const float *lse_step, int batch, int64_t rows, int heads, int d,
               int is_bf16, int o_is_fp32, int64_t lse_row_stride, int64_t lse_head_stride, int64_t step_row_stride,
               int64_t step_head_stride, int64_t o_batch_stride, int64_t o_step_batch_stride, int64_t lse_batch_stride,
               int64_t lse_step_batch_stride, cudaStream_t s);

// ---- device helpers shared by every kernel ------------------------------------------------------
struct SeqInfo {
  int64_t q_off, k_off;   // element offsets of row 0 (before head offset) for q-like / k-like tensors: use with strides
  int q_row0, k_row0;     // first packed row (varlen) or 0
  int sq, sk;             // actual lengths of this batch element
};

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
// The flashmask_maxmin workspace is sized by the caller for 64-key blocks (8 * b * h_mask * round_up(ceil(sk/64), 4)
// int32, the reference's contract).  The tcgen05 kernels scan with 128-key blocks, which leaves room behind the eight
// first-level planes for a second level over groups of 32 blocks whenever this holds.
__host__ __device__ inline bool flashmask_has_l2(int seqlen_k, int block_n) {
  const int nbp = (ceil_div(seqlen_k, block_n) + 3) & ~3;
  const int nbp64 = (ceil_div(seqlen_k, 64) + 3) & ~3;
  return block_n >= 128 && nbp + ceil_div(nbp, 32) <= nbp64;
}

#ifdef __CUDACC__
// Philox-4x32 with 7 rounds (Salmon et al., SC'11: the smallest round count that passes BigCrush, and the count the
// reference runs, csrc/flash_attn/src/philox.cuh:42-47): key = seed, counter = (offset, subsequence).  One call yields
// 16 random bytes.  Round 1 of this library ran the 10 rounds of cuRAND; three rounds fewer is 30 % of the generator's
// cost in the dropout instantiations of the forward and backward kernels.
__device__ __forceinline__ uint4 philox4x32_7(uint64_t seed, uint64_t subsequence, uint64_t offset) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint4 c = make_uint4((uint32_t)offset, (uint32_t)(offset >> 32), (uint32_t)subsequence, (uint32_t)(subsequence >> 32));
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c;
}
// Dropout decision for attention weight (b, h, row, col): 16 consecutive columns share one Philox call; the weight is
// KEPT iff its byte <= floor(255 * p_keep) -- the reference's 8-bit threshold (flash.h:88-94, softmax.h:392-436).
struct DropCtx {
  uint64_t seed, offset, row_base;   // row_base = ((b * h + head) * seqlen_q + row) * ceil(seqlen_k / 16)
  uint32_t threshold;
};
__device__ __forceinline__ bool dropout_keep(const DropCtx &d, int col) {
  const uint4 r = philox4x32_7(d.seed, d.row_base + (uint32_t)(col >> 4), d.offset);
  const uint32_t w = ((col >> 2) & 3) == 0 ? r.x : ((col >> 2) & 3) == 1 ? r.y : ((col >> 2) & 3) == 2 ? r.z : r.w;
  return ((w >> ((col & 3) * 8)) & 0xffu) <= d.threshold;
}
__device__ __forceinline__ DropCtx make_drop_ctx(const Params &p, int bi, int head, int row) {
  DropCtx d;
  d.seed = p.philox_seed; d.offset = p.philox_offset;
  d.row_base = (((uint64_t)bi * p.h + head) * (uint64_t)p.seqlen_q + row) * (uint64_t)((p.seqlen_k + 15) / 16);
  d.threshold = (uint32_t)floorf((1.f - p.p_dropout) * 255.f);
  return d;
}

// Restates csrc/flash_attn/src/block_info.h:12-37 (sum_s = cu_seqlens[b], actual = cu[b+1]-cu[b];
// varlen_padded_input keeps the batch stride) plus seqused_k of csrc/flashmask_v2/seqlen.h:56-91.
__device__ __forceinline__ void seq_bounds(const Params &p, int bi, int &q_row0, int &sq, int &k_row0, int &sk,
                                           bool &q_batched, bool &k_batched) {
  if (p.cu_seqlens_q) {
    int a = p.cu_seqlens_q[bi];
    sq = p.cu_seqlens_q[bi + 1] - a;
    q_row0 = p.varlen_padded_input ? 0 : a;
    q_batched = p.varlen_padded_input != 0;
  } else {
    sq = p.seqlen_q; q_row0 = 0; q_batched = true;
  }
  if (p.cu_seqlens_k) {
    int a = p.cu_seqlens_k[bi];
    sk = p.cu_seqlens_k[bi + 1] - a;
    k_row0 = p.varlen_padded_input ? 0 : a;
    k_batched = p.varlen_padded_input != 0;
  } else {
    sk = p.seqlen_k; k_row0 = 0; k_batched = true;
  }
  if (p.seqused_k) sk = p.seqused_k[bi];
  if (p.seqused_q) sq = min(sq, p.seqused_q[bi]);
}

// Element-level mask rule. TRUE = masked.
//   seqlen tail + top-left causal : csrc/flash_attn/src/softmax.h:143-178 (col >= min(sk, row+1))
//   bottom-right causal / local   : csrc/flashmask_v2/mask.h:85-118
//   flashmask vectors             : csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:1401-1461
struct MaskCtx {
  int sq, sk, shift;      // shift = sk - sq
  int causal, bottom_right, wl, wr;
  const int32_t *lts, *lte, *uts, *ute;  // already offset to (batch, mask head), indexed by column
};
__device__ __forceinline__ bool is_masked(const MaskCtx &m, int row, int col) {
  if (col >= m.sk) return true;
  if (m.causal && col > row + (m.bottom_right ? m.shift : 0)) return true;
  if (m.wl >= 0 && col < row + m.shift - m.wl) return true;
  if (m.wr >= 0 && col > row + m.shift + m.wr) return true;
  if (m.lts) {
    int s = m.lts[col];
    int e = m.lte ? m.lte[col] : 0x7fffffff;
    if (row >= s && row < e) return true;
    if (m.ute) {
      int us = m.uts ? m.uts[col] : 0;
      if (row >= us && row < m.ute[col]) return true;
    }
  }
  return false;
}
// bit i (0 <= i < 32) is set iff position base + i lies OUTSIDE [lo, hi): the masked positions of a 32-wide chunk
// when the visible set is an interval (seqlen tail, causal, sliding window).
__device__ __forceinline__ uint32_t outside_bits(int lo, int hi, int base) {
  lo = min(max(lo - base, 0), 32);
  hi = min(max(hi - base, 0), 32);
  if (hi <= lo) return 0xffffffffu;
  const uint32_t below_hi = hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
  const uint32_t below_lo = (1u << lo) - 1u;  // lo < 32 here
  return ~(below_hi & ~below_lo);
}
// visible key columns of a query row / visible query rows of a key column, as half-open intervals
__device__ __forceinline__ void visible_cols(const MaskCtx &m, int row, int &lo, int &hi) {
  lo = 0; hi = m.sk;
  if (m.causal) hi = min(hi, row + (m.bottom_right ? m.shift : 0) + 1);
  if (m.wr >= 0) hi = min(hi, row + m.shift + m.wr + 1);
  if (m.wl >= 0) lo = max(lo, row + m.shift - m.wl);
}
__device__ __forceinline__ void visible_rows(const MaskCtx &m, int col, int &lo, int &hi) {
  lo = 0; hi = 0x7fffffff;
  if (col >= m.sk) { hi = 0; return; }
  if (m.causal) lo = max(lo, col - (m.bottom_right ? m.shift : 0));
  if (m.wr >= 0) lo = max(lo, col - m.shift - m.wr);
  if (m.wl >= 0) hi = col - m.shift + m.wl + 1;
}
__device__ __forceinline__ MaskCtx make_mask_ctx(const Params &p, int bi, int head, int sq, int sk) {
  MaskCtx m;
  m.sq = sq; m.sk = sk; m.shift = sk - sq;
  m.causal = p.is_causal; m.bottom_right = p.causal_bottom_right;
  m.wl = p.window_left; m.wr = p.window_right;
  m.lts = m.lte = m.uts = m.ute = nullptr;
  if (p.lt_start) {
    int64_t off = ((int64_t)bi * p.h_mask + head / (p.h / p.h_mask)) * p.seqlen_k;
    m.lts = p.lt_start + off;
    m.lte = p.lt_end ? p.lt_end + off : nullptr;
    m.uts = p.ut_start ? p.ut_start + off : nullptr;
    m.ute = p.ut_end ? p.ut_end + off : nullptr;
  }
  return m;
}
#endif

}  // namespace fa
