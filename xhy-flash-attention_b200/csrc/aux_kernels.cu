// aux_kernels.cu -- the HBM-bound helper kernels: split-KV LSE combine, FlashMask block min/max scan,
// KV-cache append (contiguous / paged), rotary embedding, ring-attention LSE merge.
#include <climits>
#include <math_constants.h>

#include "common.h"

namespace fa {
namespace {

template <typename T> __device__ __forceinline__ float to_f(T x);
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_f<__half>(__half x) { return __half2float(x); }
template <> __device__ __forceinline__ float to_f<float>(float x) { return x; }
template <typename T> __device__ __forceinline__ T from_f(float x);
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }
template <> __device__ __forceinline__ __half from_f<__half>(float x) { return __float2half_rn(x); }
template <> __device__ __forceinline__ float from_f<float>(float x) { return x; }

// ------------------------------------------------------------------------------------------------
// LSE combine of split partials.  Restates csrc/flash_attn_v3/flash_fwd_combine_kernel.h:340-376,423-453:
//   lse = logsumexp_s(lse_s); o = sum_s exp(lse_s - lse) * o_s; splits with lse_s = -inf contribute nothing.
// One warp per (batch, row, head); partials oaccum (S, b, sq, h, d) fp32 / lseaccum (S, b, h, sq) fp32.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) combine_kernel(const Params p) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t item = (int64_t)blockIdx.x * 4 + warp;  // over b * sq * h
  const int64_t n_items = (int64_t)p.b * p.seqlen_q * p.h;
  if (item >= n_items) return;
  const int hq = item % p.h;
  const int row = (item / p.h) % p.seqlen_q;
  const int bi = item / ((int64_t)p.h * p.seqlen_q);
  // varlen (cu_seqlens_q): the partials are indexed by (batch, row < max_seqlen_q); the output is packed
  int q_row0, sq_b, k_row0, sk_b;
  bool q_batched, k_batched;
  seq_bounds(p, bi, q_row0, sq_b, k_row0, sk_b, q_batched, k_batched);
  if (row >= sq_b) return;
  const int S = p.num_splits;
  const int64_t lse_split_stride = (int64_t)p.b * p.h * p.seqlen_q;
  const float *lse_in = p.lseaccum + ((int64_t)bi * p.h + hq) * p.seqlen_q + row;
  float mx = -CUDART_INF_F;
  for (int s = lane; s < S; s += 32) mx = fmaxf(mx, lse_in[s * lse_split_stride]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
  if (mx != -CUDART_INF_F)
    for (int s = lane; s < S; s += 32) sum += __expf(lse_in[s * lse_split_stride] - mx);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const bool empty = (mx == -CUDART_INF_F);
  const float lse = empty ? -CUDART_INF_F : mx + __logf(sum);
  if (lane == 0)
    p.softmax_lse[((int64_t)bi * p.h + hq) * p.seqlen_q + row] =
        empty ? (p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F) : lse;
  const int d = p.d;
  const int64_t o_split_stride = (int64_t)p.b * p.seqlen_q * p.h * d;
  const float *o_in = p.oaccum + (((int64_t)bi * p.seqlen_q + row) * p.h + hq) * (int64_t)d;
  T *out = (T *)p.o.ptr + (q_batched ? bi * p.o.batch_stride : 0) + (int64_t)(q_row0 + row) * p.o.row_stride +
           (int64_t)hq * p.o.head_stride;
  for (int e = lane * 4; e < d; e += 128) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < S && !empty; ++s) {
      const float ls = lse_in[s * lse_split_stride];
      if (ls == -CUDART_INF_F) continue;
      const float w = __expf(ls - lse);
      const float4 v = *reinterpret_cast<const float4 *>(o_in + s * o_split_stride + e);
      acc.x = fmaf(w, v.x, acc.x); acc.y = fmaf(w, v.y, acc.y);
      acc.z = fmaf(w, v.z, acc.z); acc.w = fmaf(w, v.w, acc.w);
    }
    out[e] = from_f<T>(acc.x); out[e + 1] = from_f<T>(acc.y);
    out[e + 2] = from_f<T>(acc.z); out[e + 3] = from_f<T>(acc.w);
  }
}

// ------------------------------------------------------------------------------------------------
// FlashMask block min/max scan.  Restates csrc/flashmask_v2/flash_mask.hpp:80-160,162-245:
// eight arrays [b * h_mask][nblock_padded] in the order
//   lts_max, lts_min, ute_max, ute_min, lte_max, lte_min, uts_max, uts_min
// with nblock_padded = round_up(ceil(sk / block_n), 4); min starts at INT_MAX, max at 0 (ragged tails).
// One warp per (mask row, n_block).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) flashmask_scan_kernel(const Params p, int block_n, int nblock, int nblock_pad) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int nb = blockIdx.x * 4 + warp;
  const int bh = blockIdx.y;
  if (nb >= nblock_pad) return;
  const int64_t plane = (int64_t)p.b * p.h_mask * nblock_pad;
  const int32_t *vecs[4] = {p.lt_start, p.ut_end, p.lt_end, p.ut_start};
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    if (vecs[v] == nullptr) continue;
    const int32_t *src = vecs[v] + (int64_t)bh * p.seqlen_k;
    int mx = 0, mn = INT_MAX;
    if (nb < nblock) {
      const int c1 = min(p.seqlen_k, (nb + 1) * block_n);
      for (int c = nb * block_n + lane; c < c1; c += 32) {
        const int x = src[c];
        mx = max(mx, x); mn = min(mn, x);
      }
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    mn = __reduce_min_sync(0xffffffffu, mn);
    if (lane == 0) {
      p.flashmask_maxmin[(2 * v) * plane + (int64_t)bh * nblock_pad + nb] = mx;
      p.flashmask_maxmin[(2 * v + 1) * plane + (int64_t)bh * nblock_pad + nb] = mn;
    }
  }
}

// Second level of the block scan: max-of-max / min-of-min over groups of 32 consecutive blocks, stored behind the eight
// first-level planes (same order).  Lets the forward kernel discard 32 fully masked key blocks (4096 keys) per test
// when a tile's causal range is long but the mask leaves only a narrow band (documents in a packed 256k-token row).
__global__ void __launch_bounds__(128) flashmask_scan_l2_kernel(const Params p, int nblock_pad, int nsup) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int g = blockIdx.x * 4 + warp;
  const int bh = blockIdx.y;
  if (g >= nsup) return;
  const int64_t plane = (int64_t)p.b * p.h_mask * nblock_pad;
  const int64_t plane2 = (int64_t)p.b * p.h_mask * nsup;
  int32_t *l2 = p.flashmask_maxmin + 8 * plane;
  const bool present[4] = {p.lt_start != nullptr, p.ut_end != nullptr, p.lt_end != nullptr, p.ut_start != nullptr};
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    if (!present[v]) continue;
    const int nb = g * 32 + lane;
    int mx = 0, mn = INT_MAX;
    if (nb < nblock_pad) {
      mx = p.flashmask_maxmin[(2 * v) * plane + (int64_t)bh * nblock_pad + nb];
      mn = p.flashmask_maxmin[(2 * v + 1) * plane + (int64_t)bh * nblock_pad + nb];
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    mn = __reduce_min_sync(0xffffffffu, mn);
    if (lane == 0) {
      l2[(2 * v) * plane2 + (int64_t)bh * nsup + g] = mx;
      l2[(2 * v + 1) * plane2 + (int64_t)bh * nsup + g] = mn;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// KV-cache append: cache[b, cache_seqlens[b] + i] = new[b, i]
// (flash_attn_with_kvcache's k=/v=; csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:1542 store_kv_new)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
append_kv_kernel(const uint4 *k_new, const uint4 *v_new, uint4 *k_cache, uint4 *v_cache, const int32_t *cache_seqlens,
                 int batch, int seqlen_new, int seqlen_cache, int h_k, int d8, int64_t nbs, int64_t nrs, int64_t nhs,
                 int64_t cbs, int64_t crs, int64_t chs) {
  const int64_t total = (int64_t)batch * seqlen_new * h_k * d8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = i % d8;
    const int hh = (i / d8) % h_k;
    const int r = (i / ((int64_t)d8 * h_k)) % seqlen_new;
    const int bi = i / ((int64_t)d8 * h_k * seqlen_new);
    const int dst_row = (cache_seqlens ? cache_seqlens[bi] : 0) + r;
    if (dst_row >= seqlen_cache) continue;
    // strides are in elements; a uint4 holds 8 elements
    const int64_t so = (bi * nbs + (int64_t)r * nrs + (int64_t)hh * nhs) / 8 + c;
    const int64_t dof = (bi * cbs + (int64_t)dst_row * crs + (int64_t)hh * chs) / 8 + c;
    k_cache[dof] = k_new[so];
    v_cache[dof] = v_new[so];
  }
}

// Append for every cache layout fa_b200_params can describe: contiguous, kv_batch_idx-indirected, paged
// (csrc/flash_attn_v3/paged_kv.h store_K / store_V).  Row r of batch element bi lands at physical row
// seqused_k[bi] + r; K and V may have different strides, so one launch handles one tensor.
__global__ void __launch_bounds__(256)
append_paged_kernel(const uint4 *src, uint4 *cache, const int32_t *cache_seqlens, const int32_t *page_table,
                    int64_t table_stride, int page_size, const int32_t *kv_batch_idx, int batch, int seqlen_new,
                    int seqlen_cache, int h_k, int d8, int64_t nbs, int64_t nrs, int64_t nhs, int64_t cbs, int64_t crs,
                    int64_t chs) {
  const int64_t total = (int64_t)batch * seqlen_new * h_k * d8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = i % d8;
    const int hh = (i / d8) % h_k;
    const int r = (i / ((int64_t)d8 * h_k)) % seqlen_new;
    const int bi = i / ((int64_t)d8 * h_k * seqlen_new);
    const int dst_row = (cache_seqlens ? cache_seqlens[bi] : 0) + r;
    if (dst_row >= seqlen_cache) continue;
    int64_t slab = kv_batch_idx ? kv_batch_idx[bi] : bi, row = dst_row;
    if (page_table) {
      slab = page_table[(int64_t)bi * table_stride + dst_row / page_size];
      row = dst_row % page_size;
    }
    cache[(slab * cbs + row * crs + (int64_t)hh * chs) / 8 + c] = src[(bi * nbs + (int64_t)r * nrs + (int64_t)hh * nhs) / 8 + c];
  }
}

// ------------------------------------------------------------------------------------------------
// Rotary embedding, out of place (csrc/flash_attn_v3/rotary.h: interleaved pairs (2j, 2j+1), contiguous pairs
// (j, j + rotary_dim/2); x1' = x1 cos - x2 sin, x2' = x1 sin + x2 cos in fp32, one rounding; position of row r of batch
// element bi = seqlen_offsets[bi] + (fixed_position ? 0 : r), as flash_attn_with_kvcache rotates q and k_new).
// One thread per (batch, row, head, pair-or-tail-element).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
rotary_kernel(const T *x, T *out, const T *cosp, const T *sinp, const int32_t *offsets, int batch, int seqlen, int heads,
              int d, int rdim, int seqlen_ro, int interleaved, int fixed, int64_t xbs, int64_t xrs, int64_t xhs,
              int64_t obs, int64_t ors, int64_t ohs) {
  const int half = rdim / 2;
  const int per_head = half + (d - rdim);                     // rotated pairs + copied tail elements
  const int64_t total = (int64_t)batch * seqlen * heads * per_head;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = i % per_head;
    const int hh = (i / per_head) % heads;
    const int r = (i / ((int64_t)per_head * heads)) % seqlen;
    const int bi = i / ((int64_t)per_head * heads * seqlen);
    const T *xi = x + bi * xbs + (int64_t)r * xrs + (int64_t)hh * xhs;
    T *oi = out + bi * obs + (int64_t)r * ors + (int64_t)hh * ohs;
    if (j >= half) { oi[rdim + (j - half)] = xi[rdim + (j - half)]; continue; }
    const int pos = (offsets ? offsets[bi] : 0) + (fixed ? 0 : r);
    const int e1 = interleaved ? 2 * j : j, e2 = interleaved ? 2 * j + 1 : j + half;
    const float x1 = to_f(xi[e1]), x2 = to_f(xi[e2]);
    float c = 1.f, sn = 0.f;                                  // positions past the table are left unrotated
    if (pos < seqlen_ro) { c = to_f(cosp[(int64_t)pos * half + j]); sn = to_f(sinp[(int64_t)pos * half + j]); }
    oi[e1] = from_f<T>(x1 * c - x2 * sn);
    oi[e2] = from_f<T>(x1 * sn + x2 * c);
  }
}

// ------------------------------------------------------------------------------------------------
// ring-attention step merge (same algebra as the combine): treats +inf and -inf LSE as "no contribution".
// One warp per (row, head); o contiguous (rows, heads, d).
// ------------------------------------------------------------------------------------------------
template <typename TO, typename TS>
__global__ void __launch_bounds__(128)
lse_merge_kernel(TO *o, float *lse, const TS *o_step, const float *lse_step, int64_t rows, int heads, int d,
                 int64_t lrs, int64_t lhs, int64_t srs, int64_t shs, int64_t obs, int64_t sbs, int64_t lbs, int64_t slbs) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int64_t item = (int64_t)blockIdx.x * 4 + warp;
  if (item >= rows * heads) return;
  const int bi = blockIdx.y;
  const int hh = item % heads;
  const int64_t row = item / heads;
  const int64_t li = bi * lbs + row * lrs + hh * lhs;
  const float l1 = lse[li], l2 = lse_step[bi * slbs + row * srs + hh * shs];
  const bool e1 = isinf(l1), e2 = isinf(l2);
  if (e2) return;  // nothing to add
  float w1, w2, ln;
  if (e1) { w1 = 0.f; w2 = 1.f; ln = l2; }
  else {
    const float mx = fmaxf(l1, l2);
    const float s = __expf(l1 - mx) + __expf(l2 - mx);
    ln = mx + __logf(s);
    w1 = __expf(l1 - ln); w2 = __expf(l2 - ln);
  }
  TO *po = o + bi * obs + item * d;
  const TS *ps = o_step + bi * sbs + item * d;
  if (d % 4 == 0) {         // 4 elements per lane and step: 16-byte (fp32) / 8-byte (16-bit) accesses
    for (int e = lane * 4; e < d; e += 128) {
      float a[4], b[4];
      if constexpr (sizeof(TO) == 4) { const float4 t = *reinterpret_cast<const float4 *>(po + e); a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w; }
      else { const uint2 t = *reinterpret_cast<const uint2 *>(po + e); const TO *te = reinterpret_cast<const TO *>(&t);
             for (int i = 0; i < 4; ++i) a[i] = to_f(te[i]); }
      { const uint2 t = *reinterpret_cast<const uint2 *>(ps + e); const TS *te = reinterpret_cast<const TS *>(&t);
        for (int i = 0; i < 4; ++i) b[i] = to_f(te[i]); }
      for (int i = 0; i < 4; ++i) a[i] = w1 * a[i] + w2 * b[i];
      if constexpr (sizeof(TO) == 4) *reinterpret_cast<float4 *>(po + e) = make_float4(a[0], a[1], a[2], a[3]);
      else { uint2 t; TO *te = reinterpret_cast<TO *>(&t); for (int i = 0; i < 4; ++i) te[i] = from_f<TO>(a[i]);
             *reinterpret_cast<uint2 *>(po + e) = t; }
    }
  } else {
    for (int e = lane; e < d; e += 32) po[e] = from_f<TO>(w1 * to_f(po[e]) + w2 * to_f(ps[e]));
  }
  __syncwarp();
  if (lane == 0) lse[li] = ln;
}

// Per-batch split counts: restates prepare_varlen_num_blocks_kernel (csrc/flash_attn_v3/flash_prepare_scheduler.cu:15-106)
// with this library's tile sizes (16 packed query rows per m block -- the decode kernel's row tile -- and 128 keys per
// n block).  One CTA; thread i handles batch element i.
__global__ void __launch_bounds__(1024) prepare_num_splits_kernel(const Params p, int32_t *out, int num_sm) {
  __shared__ int total_blocks;
  if (threadIdx.x == 0) total_blocks = 0;
  __syncthreads();
  const int bi = threadIdx.x;
  int n_blocks = 0, m_blocks = 0;
  if (bi < p.b) {
    int q_row0, sq, k_row0, sk;
    bool qb, kb;
    seq_bounds(p, bi, q_row0, sq, k_row0, sk, qb, kb);
    const int leftpad = p.leftpad_k ? p.leftpad_k[bi] : 0;
    sk = max(min(sk, p.seqlen_k) - leftpad, 0);
    m_blocks = ceil_div(sq * (p.h / p.h_k), 16);
    n_blocks = ceil_div(sk, 128);
    atomicAdd(&total_blocks, m_blocks * n_blocks);
  }
  __syncthreads();
  const int blocks_per_sm = max((int)ceilf((float)total_blocks * 1.1f * (float)p.h_k / (float)num_sm), 1);   // 10% margin
  if (bi < p.b) out[bi] = max(min(ceil_div(n_blocks, blocks_per_sm), p.num_splits), 1);
}

__global__ void __launch_bounds__(256) fill_f32_kernel(float *dst, float value, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) dst[i] = value;
}

}  // namespace

void combine_fwd(const Params &p, cudaStream_t s) {
  FA_CHECK(p.d % 4 == 0, "combine: head_size must be a multiple of 4");
  const int64_t n_items = (int64_t)p.b * p.seqlen_q * p.h;
  dim3 grid((unsigned)((n_items + 3) / 4));
  if (p.is_bf16) combine_kernel<__nv_bfloat16><<<grid, 128, 0, s>>>(p);
  else combine_kernel<__half><<<grid, 128, 0, s>>>(p);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void flashmask_prepare(const Params &p, int block_n, cudaStream_t s, bool with_l2) {
  FA_CHECK(block_n > 0, "flashmask_prepare: block_n must be positive");
  const int nblock = ceil_div(p.seqlen_k, block_n);
  const int nblock_pad = (nblock + 3) & ~3;
  dim3 grid(ceil_div(nblock_pad, 4), p.b * p.h_mask);
  flashmask_scan_kernel<<<grid, 128, 0, s>>>(p, block_n, nblock, nblock_pad);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
  if (with_l2 && flashmask_has_l2(p.seqlen_k, block_n)) {
    const int nsup = ceil_div(nblock_pad, 32);
    dim3 grid2(ceil_div(nsup, 4), p.b * p.h_mask);
    flashmask_scan_l2_kernel<<<grid2, 128, 0, s>>>(p, nblock_pad, nsup);
    FA_CUDA(cudaGetLastError());
    note_launch(1);
  }
}

void append_kv(const void *k_new, const void *v_new, void *k_cache, void *v_cache, const int32_t *cache_seqlens,
               int batch, int seqlen_new, int seqlen_cache, int h_k, int d, int is_bf16, int64_t nbs, int64_t nrs,
               int64_t nhs, int64_t cbs, int64_t crs, int64_t chs, cudaStream_t s) {
  (void)is_bf16;
  FA_CHECK(nbs % 8 == 0 && nrs % 8 == 0 && nhs % 8 == 0 && cbs % 8 == 0 && crs % 8 == 0 && chs % 8 == 0,
           "append_kv: strides must be multiples of 8 elements");
  FA_CHECK(((uintptr_t)k_new | (uintptr_t)v_new | (uintptr_t)k_cache | (uintptr_t)v_cache) % 16 == 0,
           "append_kv: tensors must be 16-byte aligned");
  const int64_t total = (int64_t)batch * seqlen_new * h_k * (d / 8);
  int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
  if (blocks < 1) blocks = 1;
  append_kv_kernel<<<blocks, 256, 0, s>>>((const uint4 *)k_new, (const uint4 *)v_new, (uint4 *)k_cache, (uint4 *)v_cache,
                                         cache_seqlens, batch, seqlen_new, seqlen_cache, h_k, d / 8, nbs, nrs, nhs, cbs,
                                         crs, chs);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void append_kv_paged(const Params &p, const void *k_new, const void *v_new, int seqlen_new, int64_t nbs, int64_t nrs,
                     int64_t nhs, cudaStream_t s) {
  FA_CHECK(nbs % 8 == 0 && nrs % 8 == 0 && nhs % 8 == 0, "append_kv: strides must be multiples of 8 elements");
  FA_CHECK(p.d % 8 == 0, "append_kv: head_size must be a multiple of 8");
  FA_CHECK(!(p.page_table && p.kv_batch_idx), "append_kv: page_table and kv_batch_idx are exclusive");
  FA_CHECK(!p.page_table || p.page_size > 0, "append_kv: page_size");
  const int64_t total = (int64_t)p.b * seqlen_new * p.h_k * (p.d / 8);
  int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
  if (blocks < 1) blocks = 1;
  const fa_b200_tensor *ts[2] = {&p.k, &p.v};
  const void *src[2] = {k_new, v_new};
  for (int i = 0; i < 2; ++i) {
    const fa_b200_tensor &t = *ts[i];
    FA_CHECK(t.batch_stride % 8 == 0 && t.row_stride % 8 == 0 && t.head_stride % 8 == 0,
             "append_kv: cache strides must be multiples of 8 elements");
    FA_CHECK(((uintptr_t)t.ptr | (uintptr_t)src[i]) % 16 == 0, "append_kv: tensors must be 16-byte aligned");
    append_paged_kernel<<<blocks, 256, 0, s>>>((const uint4 *)src[i], (uint4 *)t.ptr, p.seqused_k, p.page_table,
                                               p.page_table_batch_stride, p.page_size, p.kv_batch_idx, p.b, seqlen_new,
                                               p.seqlen_k, p.h_k, p.d / 8, nbs, nrs, nhs, t.batch_stride, t.row_stride,
                                               t.head_stride);
    FA_CUDA(cudaGetLastError());
    note_launch(1);
  }
}

void rotary(const void *x, void *out, const void *cosp, const void *sinp, const int32_t *offsets, int batch, int seqlen,
            int heads, int d, int rdim, int seqlen_ro, int interleaved, int fixed, int is_bf16, int64_t xbs, int64_t xrs,
            int64_t xhs, int64_t obs, int64_t ors, int64_t ohs, cudaStream_t s) {
  FA_CHECK(rdim >= 0 && rdim <= d && rdim % 2 == 0, "rotary: rotary_dim must be even and <= head_size");
  const int64_t total = (int64_t)batch * seqlen * heads * (rdim / 2 + d - rdim);
  int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
  if (blocks < 1) blocks = 1;
  if (is_bf16)
    rotary_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>((const __nv_bfloat16 *)x, (__nv_bfloat16 *)out,
                                                        (const __nv_bfloat16 *)cosp, (const __nv_bfloat16 *)sinp, offsets,
                                                        batch, seqlen, heads, d, rdim, seqlen_ro, interleaved, fixed, xbs,
                                                        xrs, xhs, obs, ors, ohs);
  else
    rotary_kernel<__half><<<blocks, 256, 0, s>>>((const __half *)x, (__half *)out, (const __half *)cosp,
                                                 (const __half *)sinp, offsets, batch, seqlen, heads, d, rdim, seqlen_ro,
                                                 interleaved, fixed, xbs, xrs, xhs, obs, ors, ohs);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void lse_merge(void *o, float *lse, const void *o_step, const float *lse_step, int batch, int64_t rows, int heads, int d,
               int is_bf16, int o_is_fp32, int64_t lrs, int64_t lhs, int64_t srs, int64_t shs, int64_t obs, int64_t sbs,
               int64_t lbs, int64_t slbs, cudaStream_t s) {
  if (batch <= 0 || rows <= 0) return;
  dim3 grid((unsigned)((rows * heads + 3) / 4), (unsigned)batch);
#define FA_MERGE(TO, TS) lse_merge_kernel<TO, TS><<<grid, 128, 0, s>>>((TO *)o, lse, (const TS *)o_step, lse_step, rows, \
                                                                      heads, d, lrs, lhs, srs, shs, obs, sbs, lbs, slbs)
  if (o_is_fp32) { if (is_bf16) FA_MERGE(float, __nv_bfloat16); else FA_MERGE(float, __half); }
  else { if (is_bf16) FA_MERGE(__nv_bfloat16, __nv_bfloat16); else FA_MERGE(__half, __half); }
#undef FA_MERGE
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void prepare_num_splits(const Params &p, int32_t *out, int num_sm, cudaStream_t s) {
  FA_CHECK(p.b <= 1024, "prepare_num_splits: batch must be <= 1024");
  FA_CHECK(out != nullptr && p.num_splits >= 1 && p.h_k > 0, "prepare_num_splits: output and num_splits must be set");
  prepare_num_splits_kernel<<<1, 1024, 0, s>>>(p, out, num_sm > 0 ? num_sm : device_sm_count());
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

void fill_f32(float *dst, float value, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  const int64_t blocks = (n + 255) / 256;
  fill_f32_kernel<<<(unsigned)(blocks < 1184 ? blocks : 1184), 256, 0, s>>>(dst, value, n);
  FA_CUDA(cudaGetLastError());
  note_launch(1);
}

}  // namespace fa
