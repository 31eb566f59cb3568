// generic.cu -- the generic SIMT kernel family of libflashattn_b200.
//
// One warp per query row (forward, dQ) or per key column (dK/dV).  These kernels take every shape the
// boundary admits (head_size % 8 == 0 up to 256, any strides, fp16/bf16, every mask kind, varlen, split-KV)
// and are the kernels behind shapes the tcgen05 family (sm100_fwd.cu / sm100_bwd.cu) does not cover.
// They are CUDA kernels, not a host fallback: nothing in this library computes attention on the CPU.
//
// Semantics restated from the reference:
//   online softmax + LSE           csrc/flash_attn/src/flash_fwd_kernel.h:71-100,505-582
//   exp2 with folded scale         csrc/flash_attn/src/softmax.h:76-82
//   P = exp2(S*c - LSE*log2e), dS = P o (dP - dPsum)   csrc/flash_attn/src/flash_bwd_kernel.h:847,897-906
//   dPsum = rowsum(dO o O)         csrc/flash_attn/src/flash_bwd_kernel.h:112-188
//   split ranges                   csrc/flashmask_v2/block.h:36-46
#include <math_constants.h>

#include "common.h"

namespace fa {
namespace {

constexpr int kWarps = 4;
constexpr int kMaxD = 256;
constexpr int kMaxC = kMaxD / 32;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <typename T> __device__ __forceinline__ float to_f(T x);
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_f<__half>(__half x) { return __half2float(x); }
template <typename T> __device__ __forceinline__ T from_f(float x);
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }
template <> __device__ __forceinline__ __half from_f<__half>(float x) { return __float2half_rn(x); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// dot(smem fp32 vector, global row of T) over d elements (d % 8 == 0, row 16-byte aligned or not: scalar-safe).
template <typename T>
__device__ __forceinline__ float dot_row(const float *__restrict__ a, const T *__restrict__ row, int d) {
  float acc = 0.f;
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const uint4 *r4 = reinterpret_cast<const uint4 *>(row);
    for (int c = 0; c < d / 8; ++c) {
      uint4 u = r4[c];
      const T *e = reinterpret_cast<const T *>(&u);
#pragma unroll
      for (int t = 0; t < 8; ++t) acc = fmaf(a[c * 8 + t], to_f(e[t]), acc);
    }
  } else {
    for (int c = 0; c < d; ++c) acc = fmaf(a[c], to_f(row[c]), acc);
  }
  return acc;
}

// dense additive mask term, already multiplied by unscale_softmax (0 when there is no mask):
// csrc/flash_attn/src/softmax.h:360-379 adds it to the unscaled score, row offset as flash_fwd_kernel.h:1104-1107
template <typename T>
__device__ __forceinline__ float mask_term(const Params &p, int bi, int hq, int row, int col) {
  float t = 0.f;
  if (p.attn_mask != nullptr) {
    const int64_t off = ((((int64_t)bi * p.mask_head_mod + hq % p.mask_head_mod) * p.mask_seq_mod + row % p.mask_seq_mod) *
                         (int64_t)p.seqlen_k) + col;
    t = to_f(((const T *)p.attn_mask)[off]);
  }
  if (p.attn_bias != nullptr) {   // (b % bias_batch_mod, h, row, col): fmha/gmem_tile.h:747-752 of the FA1 tree
    const int64_t off = ((((int64_t)(bi % p.bias_batch_mod) * p.h + hq) * p.bias_rows + row) * (int64_t)p.seqlen_k) + col;
    t += to_f(((const T *)p.attn_bias)[off]);
  }
  return t * p.unscale_softmax;
}

struct Seq {
  int q_row0, k_row0, sq, sk;
  int64_t q_boff, k_boff;  // batch offsets in units of batch_stride multipliers (0 or bi)
};

__device__ __forceinline__ Seq get_seq(const Params &p, int bi) {
  Seq s;
  bool qb, kb;
  seq_bounds(p, bi, s.q_row0, s.sq, s.k_row0, s.sk, qb, kb);
  s.q_boff = qb ? bi : 0;
  s.k_boff = kb ? (p.kv_batch_idx ? p.kv_batch_idx[bi] : bi) : 0;
  if (p.leftpad_k) {                       // valid keys are [leftpad, seqused) (csrc/flash_attn_v3/seqlen.h:73-80)
    const int lp = p.leftpad_k[bi];
    s.k_row0 += lp;
    s.sk = max(s.sk - lp, 0);
  }
  return s;
}

__device__ __forceinline__ int64_t elem_off(const fa_b200_tensor &t, int64_t boff, int64_t row, int head) {
  return boff * t.batch_stride + row * t.row_stride + (int64_t)head * t.head_stride;
}

// K/V row of key `col` of batch element bi: contiguous cache or paged (csrc/flash_attn_v3/paged_kv.h)
template <typename T>
__device__ __forceinline__ const T *kv_row(const Params &p, const fa_b200_tensor &t, const Seq &s, int bi, int hk, int col) {
  const int64_t r = (int64_t)s.k_row0 + col;
  if (p.page_table) {
    const int page = p.page_table[(int64_t)bi * p.page_table_batch_stride + r / p.page_size];
    return (const T *)t.ptr + (int64_t)page * t.batch_stride + (r % p.page_size) * t.row_stride + (int64_t)hk * t.head_stride;
  }
  return (const T *)t.ptr + s.k_boff * t.batch_stride + r * t.row_stride + (int64_t)hk * t.head_stride;
}

// logit in the log2 domain from the raw dot product (+ dense-mask term): scale, softcap (tanh, before masks and ALiBi:
// csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:537-556,1064-1067), ALiBi.  *dfac = d logit / d (scale * dot).
struct LogitCtx {
  float scale, scale_log2, softcap, slope;
  int shift;
  bool plain;
};
__device__ __forceinline__ LogitCtx make_logit_ctx(const Params &p, int bi, int hq, int shift) {
  LogitCtx c;
  c.scale = p.softmax_scale; c.scale_log2 = p.softmax_scale * kLog2e; c.softcap = p.softcap;
  c.slope = p.alibi_slopes ? p.alibi_slopes[(int64_t)bi * p.alibi_batch_stride + hq] : 0.f;
  c.shift = shift;
  c.plain = !(p.softcap > 0.f) && p.alibi_slopes == nullptr;
  return c;
}
__device__ __forceinline__ float logit2(const LogitCtx &c, float dot, int row, int col, float *dfac = nullptr) {
  if (c.plain) { if (dfac) *dfac = 1.f; return dot * c.scale_log2; }
  float x = dot * c.scale, f = 1.f;
  if (c.softcap > 0.f) { const float t = tanhf(x / c.softcap); x = c.softcap * t; f = 1.f - t * t; }
  x -= c.slope * fabsf((float)(col - (row + c.shift)));
  if (dfac) *dfac = f;
  return x * kLog2e;
}

// column range [lo, hi) that causal / window / split leave open for a query row
__device__ __forceinline__ void col_range(const Params &p, const MaskCtx &m, int row, int split, int &lo, int &hi) {
  lo = 0; hi = m.sk;
  if (m.causal) hi = min(hi, row + (m.bottom_right ? m.shift : 0) + 1);
  if (m.wr >= 0) hi = min(hi, row + m.shift + m.wr + 1);
  if (m.wl >= 0) lo = max(lo, row + m.shift - m.wl);
  if (p.num_splits > 1) {
    int nblk = ceil_div(m.sk, 128);
    int per = ceil_div(nblk, p.num_splits) * 128;
    lo = max(lo, split * per);
    hi = min(hi, (split + 1) * per);
  }
}

// ------------------------------------------------------------------------------------------------
// forward (also the split-KV partial forward when Split)
// ------------------------------------------------------------------------------------------------
template <typename T, bool Split>
__global__ void __launch_bounds__(kWarps * 32) generic_fwd_kernel(const Params p) {
  __shared__ float q_s[kWarps][kMaxD];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int hq = blockIdx.y;
  const int bi = Split ? blockIdx.z / p.num_splits : blockIdx.z;
  const int split = Split ? blockIdx.z % p.num_splits : 0;
  const Seq sq_ = get_seq(p, bi);
  const int row = blockIdx.x * kWarps + warp;
  if (row >= sq_.sq) return;
  const int d = p.d;
  const int hk = hq / (p.h / p.h_k);
  const MaskCtx m = make_mask_ctx(p, bi, hq, sq_.sq, sq_.sk);

  const T *q = (const T *)p.q.ptr + elem_off(p.q, sq_.q_boff, sq_.q_row0 + row, hq);
  for (int e = lane; e < d; e += 32) q_s[warp][e] = to_f(q[e]);
  __syncwarp();

  const LogitCtx lc = make_logit_ctx(p, bi, hq, sq_.sk - sq_.sq);
  int lo, hi;
  col_range(p, m, row, split, lo, hi);

  float mx = -CUDART_INF_F, l = 0.f;
  float o[kMaxC];
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) o[c] = 0.f;
  const bool drop = p.p_dropout > 0.f;
  const DropCtx dc = make_drop_ctx(p, bi, hq, row);
  const float rp = drop ? 1.f / (1.f - p.p_dropout) : 1.f;

  for (int j0 = lo; j0 < hi; j0 += 32) {
    const int col = j0 + lane;
    float s = -CUDART_INF_F;
    if (col < hi && !is_masked(m, row, col))
      s = logit2(lc, dot_row(q_s[warp], kv_row<T>(p, p.k, sq_, bi, hk, col), d) + mask_term<T>(p, bi, hq, row, col), row, col);
    const float mnew = fmaxf(mx, warp_max(s));
    if (mnew == -CUDART_INF_F) continue;  // nothing visible so far (warp-uniform)
    const float corr = exp2f(mx - mnew);  // mx = -inf -> 0
    float pj = exp2f(s - mnew);           // s = -inf -> 0
    l = l * corr + warp_sum(pj);          // the row sum is taken BEFORE dropout (flash_fwd_kernel.h:399-409)
    if (drop && col < hi && !dropout_keep(dc, col)) pj = 0.f;
    mx = mnew;
#pragma unroll
    for (int c = 0; c < kMaxC; ++c) o[c] *= corr;
    const int nj = min(32, hi - j0);
    for (int jj = 0; jj < nj; ++jj) {
      const float pv = __shfl_sync(0xffffffffu, pj, jj);
      if (pv == 0.f) continue;  // warp-uniform
      const T *vrow = kv_row<T>(p, p.v, sq_, bi, hk, j0 + jj);
#pragma unroll
      for (int c = 0; c < kMaxC; ++c) {
        const int e = lane + 32 * c;
        if (e < d) o[c] = fmaf(pv, to_f(vrow[e]), o[c]);
      }
    }
  }

  const bool empty = !(l > 0.f);
  const float inv = empty ? 0.f : rp / l;
  if (!Split && p.s_dmask != nullptr) {
    // return_softmax: the normalised probabilities, negated where dropped (plain row-major (b, h, sq_r, sk_r); the
    // reference stores them in its MMA fragment order and de-swizzles in the test, tests/test_flash_attn.py:218-264)
    T *sd = (T *)p.s_dmask + (((int64_t)bi * p.h + hq) * p.seqlen_q_rounded + row) * (int64_t)p.seqlen_k_rounded;
    const float lse2 = mx + log2f(l);
    for (int j0 = 0; j0 < p.seqlen_k_rounded; j0 += 32) {
      const int col = j0 + lane;
      float pr = 0.f;
      if (!empty && col >= lo && col < hi && !is_masked(m, row, col)) {
        pr = exp2f(logit2(lc, dot_row(q_s[warp], kv_row<T>(p, p.k, sq_, bi, hk, col), d) + mask_term<T>(p, bi, hq, row, col),
                          row, col) - lse2);
        if (drop && !dropout_keep(dc, col)) pr = -pr;
      }
      if (col < p.seqlen_k_rounded) sd[col] = from_f<T>(pr);
    }
  }
  if (Split) {
    // partials: oaccum (splits, b, sq, h, d) fp32; lseaccum (splits, b, h, sq) fp32, -inf for an empty split
    float *oa = p.oaccum + ((((int64_t)split * p.b + bi) * p.seqlen_q + row) * p.h + hq) * (int64_t)d;
#pragma unroll
    for (int c = 0; c < kMaxC; ++c) {
      const int e = lane + 32 * c;
      if (e < d) oa[e] = o[c] * inv;
    }
    if (lane == 0)
      p.lseaccum[(((int64_t)split * p.b + bi) * p.h + hq) * p.seqlen_q + row] =
          empty ? -CUDART_INF_F : (mx + log2f(l)) * kLn2;
  } else {
    T *out = (T *)p.o.ptr + elem_off(p.o, sq_.q_boff, sq_.q_row0 + row, hq);
#pragma unroll
    for (int c = 0; c < kMaxC; ++c) {
      const int e = lane + 32 * c;
      if (e < d) out[e] = from_f<T>(o[c] * inv);
    }
    if (lane == 0)
      p.softmax_lse[((int64_t)bi * p.h + hq) * p.seqlen_q + row] =
          empty ? (p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F) : (mx + log2f(l)) * kLn2;
  }
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
// dPsum[b, h, row] = sum_d dO * O   (layout (b, h, seqlen_q_rounded) fp32)
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) generic_dpsum_kernel(const Params p) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int hq = blockIdx.y, bi = blockIdx.z;
  const Seq s = get_seq(p, bi);
  const int row = blockIdx.x * kWarps + warp;
  if (row >= p.seqlen_q_rounded) return;
  float acc = 0.f;
  if (row < s.sq) {
    const T *o = (const T *)p.o.ptr + elem_off(p.o, s.q_boff, s.q_row0 + row, hq);
    const T *g = (const T *)p.dout.ptr + elem_off(p.dout, s.q_boff, s.q_row0 + row, hq);
    for (int e = lane; e < p.d; e += 32) acc = fmaf(to_f(o[e]), to_f(g[e]), acc);
    acc = warp_sum(acc);
  }
  if (lane == 0) p.softmax_d[((int64_t)bi * p.h + hq) * p.seqlen_q_rounded + row] = acc;
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32) generic_dq_kernel(const Params p) {
  __shared__ float q_s[kWarps][kMaxD];
  __shared__ float g_s[kWarps][kMaxD];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int hq = blockIdx.y, bi = blockIdx.z;
  const Seq s_ = get_seq(p, bi);
  const int row = blockIdx.x * kWarps + warp;
  if (row >= s_.sq) return;
  const int d = p.d;
  const int hk = hq / (p.h / p.h_k);
  const MaskCtx m = make_mask_ctx(p, bi, hq, s_.sq, s_.sk);
  const T *q = (const T *)p.q.ptr + elem_off(p.q, s_.q_boff, s_.q_row0 + row, hq);
  const T *g = (const T *)p.dout.ptr + elem_off(p.dout, s_.q_boff, s_.q_row0 + row, hq);
  for (int e = lane; e < d; e += 32) { q_s[warp][e] = to_f(q[e]); g_s[warp][e] = to_f(g[e]); }
  __syncwarp();
  const LogitCtx lc = make_logit_ctx(p, bi, hq, s_.sk - s_.sq);
  const float lse = p.softmax_lse[((int64_t)bi * p.h + hq) * p.seqlen_q + row];
  const float dsum = p.softmax_d[((int64_t)bi * p.h + hq) * p.seqlen_q_rounded + row];
  const bool dead = isinf(lse);
  const float lse2 = dead ? 0.f : lse * kLog2e;
  const bool drop = p.p_dropout > 0.f;
  const DropCtx dc = make_drop_ctx(p, bi, hq, row);
  const float rp = drop ? 1.f / (1.f - p.p_dropout) : 1.f;
  int lo, hi;
  col_range(p, m, row, 0, lo, hi);
  float acc[kMaxC];
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) acc[c] = 0.f;
  const T *kbase = (const T *)p.k.ptr + elem_off(p.k, s_.k_boff, s_.k_row0, hk);
  const T *vbase = (const T *)p.v.ptr + elem_off(p.v, s_.k_boff, s_.k_row0, hk);
  for (int j0 = lo; j0 < hi && !dead; j0 += 32) {
    const int col = j0 + lane;
    float ds = 0.f;
    if (col < hi && !is_masked(m, row, col)) {
      float dfac;
      const float sc = logit2(lc, dot_row(q_s[warp], kbase + (int64_t)col * p.k.row_stride, d) + mask_term<T>(p, bi, hq, row, col),
                              row, col, &dfac);
      const float pij = exp2f(sc - lse2);
      float dp = dot_row(g_s[warp], vbase + (int64_t)col * p.v.row_stride, d);
      if (drop) dp = dropout_keep(dc, col) ? dp * rp : 0.f;   // dP = keep / p_keep * (dO . V)
      ds = pij * (dp - dsum) * dfac;
      // gradient of the additive bias, one element per visible score (fmha/gmem_tile.h:880-925 of the FA1 tree)
      if (p.ds_out != nullptr)
        ((T *)p.ds_out)[((((int64_t)bi * p.h + hq) * p.bias_rows + row) * (int64_t)p.seqlen_k) + col] = from_f<T>(ds * p.unscale_softmax);
    }
    const int nj = min(32, hi - j0);
    for (int jj = 0; jj < nj; ++jj) {
      const float dsj = __shfl_sync(0xffffffffu, ds, jj);
      if (dsj == 0.f) continue;
      const T *krow = kbase + (int64_t)(j0 + jj) * p.k.row_stride;
#pragma unroll
      for (int c = 0; c < kMaxC; ++c) {
        const int e = lane + 32 * c;
        if (e < d) acc[c] = fmaf(dsj, to_f(krow[e]), acc[c]);
      }
    }
  }
  T *dq = (T *)p.dq.ptr + elem_off(p.dq, s_.q_boff, s_.q_row0 + row, hq);
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) {
    const int e = lane + 32 * c;
    if (e < d) dq[e] = from_f<T>(acc[c] * p.softmax_scale);
  }
}

// one warp per key column; loops the q-heads of the group (or just one head when dk/dv are expanded)
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) generic_dkv_kernel(const Params p) {
  __shared__ float k_s[kWarps][kMaxD];
  __shared__ float v_s[kWarps][kMaxD];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int ho = blockIdx.y, bi = blockIdx.z;  // ho: output head (q-head if expanded else kv-head)
  const Seq s_ = get_seq(p, bi);
  const int col = blockIdx.x * kWarps + warp;
  if (col >= s_.sk) return;
  const int d = p.d;
  const int g = p.h / p.h_k;
  const int hk = p.dkv_expanded ? ho / g : ho;
  const int hq0 = p.dkv_expanded ? ho : ho * g;
  const int hq1 = p.dkv_expanded ? ho + 1 : ho * g + g;
  const T *krow = (const T *)p.k.ptr + elem_off(p.k, s_.k_boff, s_.k_row0 + col, hk);
  const T *vrow = (const T *)p.v.ptr + elem_off(p.v, s_.k_boff, s_.k_row0 + col, hk);
  for (int e = lane; e < d; e += 32) { k_s[warp][e] = to_f(krow[e]); v_s[warp][e] = to_f(vrow[e]); }
  __syncwarp();
  float dk[kMaxC], dv[kMaxC];
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) { dk[c] = 0.f; dv[c] = 0.f; }
  const bool drop = p.p_dropout > 0.f;
  const float rp = drop ? 1.f / (1.f - p.p_dropout) : 1.f;
  for (int hq = hq0; hq < hq1; ++hq) {
    const MaskCtx m = make_mask_ctx(p, bi, hq, s_.sq, s_.sk);
    const LogitCtx lc = make_logit_ctx(p, bi, hq, s_.sk - s_.sq);
    int rlo = 0, rhi = s_.sq;
    if (m.causal) rlo = max(rlo, col - (m.bottom_right ? m.shift : 0));
    if (m.wr >= 0) rlo = max(rlo, col - m.shift - m.wr);
    if (m.wl >= 0) rhi = min(rhi, col - m.shift + m.wl + 1);
    const T *qbase = (const T *)p.q.ptr + elem_off(p.q, s_.q_boff, s_.q_row0, hq);
    const T *gbase = (const T *)p.dout.ptr + elem_off(p.dout, s_.q_boff, s_.q_row0, hq);
    const float *lse = p.softmax_lse + ((int64_t)bi * p.h + hq) * p.seqlen_q;
    const float *dsum = p.softmax_d + ((int64_t)bi * p.h + hq) * p.seqlen_q_rounded;
    for (int i0 = rlo; i0 < rhi; i0 += 32) {
      const int row = i0 + lane;
      float pij = 0.f, ds = 0.f;
      if (row < rhi && !is_masked(m, row, col)) {
        const float l = lse[row];
        if (!isinf(l)) {
          float dfac;
          const float sc = logit2(lc, dot_row(k_s[warp], qbase + (int64_t)row * p.q.row_stride, d) + mask_term<T>(p, bi, hq, row, col),
                                  row, col, &dfac);
          pij = exp2f(sc - l * kLog2e);
          float dp = dot_row(v_s[warp], gbase + (int64_t)row * p.dout.row_stride, d);
          float pd = pij;                                  // the weight dV sees: keep / p_keep * P
          if (drop) {
            const bool keep = dropout_keep(make_drop_ctx(p, bi, hq, row), col);
            dp = keep ? dp * rp : 0.f;
            pd = keep ? pij * rp : 0.f;
          }
          ds = pij * (dp - dsum[row]) * dfac;
          pij = pd;
        }
      }
      const int ni = min(32, rhi - i0);
      for (int ii = 0; ii < ni; ++ii) {
        const float pj = __shfl_sync(0xffffffffu, pij, ii);
        const float dsj = __shfl_sync(0xffffffffu, ds, ii);
        if (pj == 0.f && dsj == 0.f) continue;
        const T *qrow = qbase + (int64_t)(i0 + ii) * p.q.row_stride;
        const T *grow = gbase + (int64_t)(i0 + ii) * p.dout.row_stride;
#pragma unroll
        for (int c = 0; c < kMaxC; ++c) {
          const int e = lane + 32 * c;
          if (e < d) {
            dv[c] = fmaf(pj, to_f(grow[e]), dv[c]);
            dk[c] = fmaf(dsj, to_f(qrow[e]), dk[c]);
          }
        }
      }
    }
  }
  T *dko = (T *)p.dk.ptr + elem_off(p.dk, s_.k_boff, s_.k_row0 + col, ho);
  T *dvo = (T *)p.dv.ptr + elem_off(p.dv, s_.k_boff, s_.k_row0 + col, ho);
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) {
    const int e = lane + 32 * c;
    if (e < d) { dko[e] = from_f<T>(dk[c] * p.softmax_scale); dvo[e] = from_f<T>(dv[c]); }
  }
}

// reduced[b, h, k] += sum_q exp(scale * q.k - lse[q])   (csrc/flash_attn/src/calc_reduced_attn_scores_kernel.h:93-110;
// the reference ADDS into the caller's buffer with atomicAdd, :111-116, so does this -- one writer per element here).
// probs != NULL (return_softmax): the fp32 probabilities themselves, (b, h, sq, sk) row-major (ibid.:56-90,262-278).
template <typename T>
__global__ void __launch_bounds__(kWarps * 32)
reduced_scores_kernel(const T *q, const T *k, const float *lse, float *out, float *probs, int sq, int sk, int h, int h_k,
                      int d, float scale, int64_t qbs, int64_t qrs, int64_t qhs, int64_t kbs, int64_t krs, int64_t khs) {
  __shared__ float k_s[kWarps][kMaxD];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int hq = blockIdx.y, bi = blockIdx.z;
  const int col = blockIdx.x * kWarps + warp;
  if (col >= sk) return;
  const int hk = hq / (h / h_k);
  const T *krow = k + bi * kbs + (int64_t)col * krs + hk * khs;
  for (int e = lane; e < d; e += 32) k_s[warp][e] = to_f(krow[e]);
  __syncwarp();
  const float *l = lse + ((int64_t)bi * h + hq) * sq;
  float *prow = probs ? probs + ((int64_t)bi * h + hq) * sq * (int64_t)sk + col : nullptr;
  float acc = 0.f;
  for (int i = lane; i < sq; i += 32) {
    const float li = l[i];
    float pv = 0.f;
    if (!isinf(li)) {
      const float s = dot_row(k_s[warp], q + bi * qbs + (int64_t)i * qrs + hq * qhs, d) * scale;
      pv = __expf(s - li);
    }
    acc += pv;
    if (prow) prow[(int64_t)i * sk] = pv;
  }
  acc = warp_sum(acc);
  if (lane == 0) out[((int64_t)bi * h + hq) * sk + col] += acc;
}

}  // namespace

void generic_fwd(const Params &p, cudaStream_t s) {
  dim3 grid(ceil_div(p.seqlen_q, kWarps), p.h, p.b);
  if (p.is_bf16) generic_fwd_kernel<__nv_bfloat16, false><<<grid, kWarps * 32, 0, s>>>(p);
  else generic_fwd_kernel<__half, false><<<grid, kWarps * 32, 0, s>>>(p);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void generic_splitkv_fwd(const Params &p, cudaStream_t s) {
  if (p.num_splits > 1) {
    dim3 grid(ceil_div(p.seqlen_q, kWarps), p.h, p.b * p.num_splits);
    if (p.is_bf16) generic_fwd_kernel<__nv_bfloat16, true><<<grid, kWarps * 32, 0, s>>>(p);
    else generic_fwd_kernel<__half, true><<<grid, kWarps * 32, 0, s>>>(p);
    FA_CUDA(cudaGetLastError());
    note_launch(1);
  } else {
    generic_fwd(p, s);
  }
}

void generic_bwd(const Params &p, cudaStream_t s) {
  dim3 g1(ceil_div(p.seqlen_q_rounded, kWarps), p.h, p.b);
  dim3 g2(ceil_div(p.seqlen_q, kWarps), p.h, p.b);
  dim3 g3(ceil_div(p.seqlen_k, kWarps), p.dkv_expanded ? p.h : p.h_k, p.b);
  if (p.is_bf16) {
    generic_dpsum_kernel<__nv_bfloat16><<<g1, kWarps * 32, 0, s>>>(p);
    generic_dq_kernel<__nv_bfloat16><<<g2, kWarps * 32, 0, s>>>(p);
    generic_dkv_kernel<__nv_bfloat16><<<g3, kWarps * 32, 0, s>>>(p);
  } else {
    generic_dpsum_kernel<__half><<<g1, kWarps * 32, 0, s>>>(p);
    generic_dq_kernel<__half><<<g2, kWarps * 32, 0, s>>>(p);
    generic_dkv_kernel<__half><<<g3, kWarps * 32, 0, s>>>(p);
  }
  FA_CUDA(cudaGetLastError());
  note_launch(1); note_launch(1); note_launch(1);
}

void reduced_scores(const void *q, const void *k, const float *lse, float *out, float *probs, int b, int sq, int sk, int h,
                    int h_k, int d, float scale, int is_bf16, int64_t qbs, int64_t qrs, int64_t qhs, int64_t kbs,
                    int64_t krs, int64_t khs, cudaStream_t s) {
  dim3 grid(ceil_div(sk, kWarps), h, b);
  if (is_bf16)
    reduced_scores_kernel<__nv_bfloat16><<<grid, kWarps * 32, 0, s>>>((const __nv_bfloat16 *)q, (const __nv_bfloat16 *)k,
                                                                     lse, out, probs, sq, sk, h, h_k, d, scale, qbs,
                                                                     qrs, qhs, kbs, krs, khs);
  else
    reduced_scores_kernel<__half><<<grid, kWarps * 32, 0, s>>>((const __half *)q, (const __half *)k, lse, out, probs, sq,
                                                              sk, h, h_k, d, scale, qbs, qrs, qhs, kbs, krs, khs);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

}  // namespace fa
