// decode.cu -- the split-KV decode kernel (flash_attn_with_kvcache): an HBM-bandwidth kernel.
//
// What it computes is the reference's split forward (csrc/flash_attn_v3/flash.h:96-125 num_splits / oaccum_ptr /
// softmax_lseaccum_ptr; split ranges csrc/flashmask_v2/block.h:36-46; bottom-right causal / local masks
// csrc/flashmask_v2/mask.h:85-118; pack-GQA csrc/flash_attn_v3/pack_gqa.h): per (batch, kv-head, split) the partial
// O_s = softmax-partial(Q K_s^T) V_s in fp32 plus LSE_s, merged afterwards by the LSE-combine kernel.
//
// Decode has one (or a few) query rows per sequence, so every K/V byte is used once: the roofline is HBM.  Structure:
//   * grid (split, kv-head, batch); the seqlen_q * (h / h_k) query rows that share this kv-head are packed into the
//     16 rows of one MMA tile (pack-GQA), so each K/V byte is read from HBM exactly once for the whole group;
//   * warp 4 = TMA producer: K and V tiles of 64 keys stream through a 3-stage shared-memory ring (128-byte swizzle),
//     96 KB per CTA so that two CTAs are resident per SM (~190 KB of loads in flight per SM);
//   * warps 0-3 = consumers: warp w owns keys [16w, 16w+16) of every tile -- S = Q K^T with mma.sync m16n8k16
//     (Q fragments live in registers), online softmax with exp2 and the folded scale, quad-shuffle row max / row sum,
//     O += P V with P fed straight from the S accumulator registers; the four partial (m, l, O) are merged through
//     shared memory at the end.  The tensor pipe is nowhere near a limit here (16 query rows), so the legacy
//     warp-level MMA is the right tool: tcgen05 would buy nothing for a kernel that waits on HBM.
#include <math_constants.h>
#include <type_traits>

#include "common.h"
#include "sm100.cuh"

namespace fa {
namespace {

using namespace sm100;

constexpr int kDT = 64;        // keys per pipeline stage
constexpr int kDStages = 3;
constexpr int kDWarps = 4;     // consumer warps
constexpr int kDThreads = (kDWarps + 1) * 32;
constexpr int kDRows = 16;     // packed query rows (seqlen_q * group size) one CTA takes
constexpr int kDMaxRows = 128; // packed rows per (batch, kv-head) served by this kernel, as ceil(rows / 16) row tiles
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct DecTma {
  CUtensorMap k, v;
};

template <int HD> struct DecSmem {
  static constexpr int kTile = kDT * HD * 2;                 // one K or V tile
  static constexpr int kK = 0;
  static constexpr int kV = kDStages * kTile;
  static constexpr int kBar = 2 * kDStages * kTile;
  static constexpr int kTotal = kBar + 2 * kDStages * 8;
  static constexpr int kDynBytes = kTotal + 1024;            // slack for manual 1024-byte alignment
  // merge scratch (after the main loop, over the K ring): [warp][16][HD] O + [warp][16] m + [warp][16] l
  static constexpr int kMergeBytes = kDWarps * kDRows * (HD + 2) * 4;
  static_assert(kMergeBytes <= 2 * kDStages * kTile, "merge scratch must fit in the ring");
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2d(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool kBf16>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (kBf16)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}

// byte offset of the 16-byte chunk `chunk` (0 .. HD/8) of key row `key` inside a TMA-written, 128B-swizzled tile that
// is stored as HD/64 boxes of [64 keys][128 bytes]
__device__ __forceinline__ uint32_t tile_off(int key, int chunk) {
  return (uint32_t)((chunk >> 3) * (kDT * 128) + key * 128 + (((chunk & 7) ^ (key & 7)) << 4));
}

// visible key interval of one packed query row (split range, seqused_k, bottom-right causal, window)
__device__ __forceinline__ void row_interval(const Params &p, int qi, int sq, int sk, int split_lo, int split_hi, int &lo,
                                             int &hi) {
  const int shift = sk - sq;
  lo = split_lo; hi = min(split_hi, sk);
  if (p.is_causal) hi = min(hi, qi + (p.causal_bottom_right ? shift : 0) + 1);
  if (p.window_right >= 0) hi = min(hi, qi + shift + p.window_right + 1);
  if (p.window_left >= 0) lo = max(lo, qi + shift - p.window_left);
}

template <int HD, bool kBf16>
__global__ void __launch_bounds__(kDThreads, 2)
decode_kernel(const __grid_constant__ Params p, const __grid_constant__ DecTma tm) {
  using L = DecSmem<HD>;
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  // grid.x = splits x row tiles, row tile fastest: the CTAs that stream the same K/V range for different groups of 16
  // packed rows are launched next to each other, so all but the first find it in L2
  const int n_rt = (p.seqlen_q * (p.h / p.h_k) + kDRows - 1) / kDRows;
  const int split = blockIdx.x / n_rt, rt = blockIdx.x % n_rt, hk = blockIdx.y, bi = blockIdx.z;
  const int g = p.h / p.h_k;
  // lengths and first rows of this batch element (csrc/flashmask_v2/seqlen.h:56-91): cu_seqlens_q / cu_seqlens_k give
  // packed tensors (no batch dimension), seqused_k a cache length
  int q_row0, sq, k_row0, sk_all;
  bool q_batched, k_batched;
  seq_bounds(p, bi, q_row0, sq, k_row0, sk_all, q_batched, k_batched);
  const int G = sq * g;                                       // packed rows of this (batch, kv-head); this CTA takes
  const int row0 = rt * kDRows;                               // rows [row0, row0 + 16)
  // valid keys are the physical cache rows [leftpad, leftpad + sk) (csrc/flash_attn_v3/seqlen.h:73-80); everything
  // below the mask intervals works on PHYSICAL rows so that 64-key tiles never straddle a page
  const int leftpad = p.leftpad_k ? p.leftpad_k[bi] : 0;
  const int sk = max(min(sk_all, p.seqlen_k) - leftpad, 0);
  // splits of this batch element: fixed, or per batch (num_splits_dynamic, flash_prepare_scheduler.cu:15-124)
  const int nsplit = p.num_splits > 1 ? (p.num_splits_dynamic ? max(min(p.num_splits_dynamic[bi], p.num_splits), 1) : p.num_splits) : 1;

  auto full_bar = [&](int s) { return sbase + L::kBar + s * 8; };
  auto empty_bar = [&](int s) { return sbase + L::kBar + (kDStages + s) * 8; };
  if (threadIdx.x == 0) {
    for (int s = 0; s < kDStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), kDWarps); }
    fence_mbar_init();
  }
  if (warp == kDWarps && lane == 0) { tma_prefetch_desc(&tm.k); tma_prefetch_desc(&tm.v); }
  __syncthreads();

  // key range of this split (same partition as csrc/flashmask_v2/block.h:36-46: whole 128-key blocks per split)
  int split_lo = 0, split_hi = sk;
  if (p.num_splits > 1) {
    const int per = ceil_div(ceil_div(sk, 128), nsplit) * 128;
    split_lo = split < nsplit ? split * per : sk;            // splits beyond this batch element's count are empty
    split_hi = split < nsplit ? min(sk, (split + 1) * per) : sk;
  }
  int cta_lo = 0x7fffffff, cta_hi = 0;
  for (int qi = row0 / g; qi < sq && qi * g < row0 + kDRows; ++qi) {   // the query rows this row tile touches
    int lo, hi;
    row_interval(p, qi, sq, sk, split_lo, split_hi, lo, hi);
    if (hi > lo) { cta_lo = min(cta_lo, lo + leftpad); cta_hi = max(cta_hi, hi + leftpad); }
  }
  const int t_lo = cta_hi > cta_lo ? cta_lo / kDT : 0;
  const int t_hi = cta_hi > cta_lo ? ceil_div(cta_hi, kDT) : 0;
  const int n_tiles = t_hi - t_lo;

  constexpr int kHalves = HD / 64;
  if (warp == kDWarps) {
    // ======================= TMA producer =======================
    if (lane == 0) {
      const int32_t *table = p.page_table ? p.page_table + (int64_t)bi * p.page_table_batch_stride : nullptr;
      const int bkv = p.kv_batch_idx ? p.kv_batch_idx[bi] : bi;
      // paged cache: the page index of tile it+1 is fetched while waiting for the slot of tile it
      int page_next = (table && n_tiles > 0) ? table[(t_lo * kDT) / p.page_size] : 0;
      for (int it = 0; it < n_tiles; ++it) {
        const int st = it % kDStages;
        int row = k_row0 + (t_lo + it) * kDT, slab = k_batched ? bkv : 0;
        if (table) {
          slab = page_next;
          row -= (row / p.page_size) * p.page_size;
          if (it + 1 < n_tiles) page_next = table[((t_lo + it + 1) * kDT) / p.page_size];
        }
        mbar_wait(empty_bar(st), ((it / kDStages) & 1) ^ 1);
        mbar_expect_tx(full_bar(st), 2 * L::kTile);
#pragma unroll
        for (int hf = 0; hf < kHalves; ++hf) {
          tma_load_4d(sbase + L::kK + st * L::kTile + hf * (kDT * 128), &tm.k, full_bar(st), hf * 64, hk, row, slab);
          tma_load_4d(sbase + L::kV + st * L::kTile + hf * (kDT * 128), &tm.v, full_bar(st), hf * 64, hk, row, slab);
        }
      }
    }
    // falls through to the final __syncthreads-free exit: the producer warp takes no part in the merge
    return;
  }

  // ======================= consumers =======================
  const int quad = lane % 4, rA = row0 + lane / 4, rB = rA + 8, lA_ = lane / 4, lB_ = lA_ + 8;   // packed rows / tile-local rows
  // Q fragments (A operand, 16 packed rows x HD): packed row r -> query index r / g, head hk * g + r % g
  uint32_t qa[HD / 16][4];
  {
    const T *qb = (const T *)p.q.ptr + (q_batched ? (int64_t)bi * p.q.batch_stride : 0) + (int64_t)q_row0 * p.q.row_stride;
    const T *qrowA = rA < G ? qb + (int64_t)(rA / g) * p.q.row_stride + (int64_t)(hk * g + rA % g) * p.q.head_stride : nullptr;
    const T *qrowB = rB < G ? qb + (int64_t)(rB / g) * p.q.row_stride + (int64_t)(hk * g + rB % g) * p.q.head_stride : nullptr;
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
      const int c = ks * 16 + 2 * quad;
      qa[ks][0] = qrowA ? *reinterpret_cast<const uint32_t *>(qrowA + c) : 0u;
      qa[ks][1] = qrowB ? *reinterpret_cast<const uint32_t *>(qrowB + c) : 0u;
      qa[ks][2] = qrowA ? *reinterpret_cast<const uint32_t *>(qrowA + c + 8) : 0u;
      qa[ks][3] = qrowB ? *reinterpret_cast<const uint32_t *>(qrowB + c + 8) : 0u;
    }
  }
  int loA = 0, hiA = 0, loB = 0, hiB = 0;
  if (rA < G) { row_interval(p, rA / g, sq, sk, split_lo, split_hi, loA, hiA); loA += leftpad; hiA += leftpad; }
  if (rB < G) { row_interval(p, rB / g, sq, sk, split_lo, split_hi, loB, hiB); loB += leftpad; hiB += leftpad; }
  const int kv_end = leftpad + sk;

  // softcap: logits = softcap * tanh(scale * s / softcap) (csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:537-556);
  // ALiBi: logits -= slope * |key - (query + sk - sq)|, per packed row
  const bool has_cap = p.softcap > 0.f, has_alibi = p.alibi_slopes != nullptr;
  const float scale_log2 = has_cap ? p.softcap * kLog2e : p.softmax_scale * kLog2e;
  const float pre_cap = has_cap ? p.softmax_scale / p.softcap : 0.f;
  float slopeA = 0.f, slopeB = 0.f;
  int diagA = 0, diagB = 0;                                    // physical key on the (bottom-right) diagonal of the row
  if (has_alibi) {
    const float *sl = p.alibi_slopes + (int64_t)bi * p.alibi_batch_stride + hk * g;
    if (rA < G) { slopeA = sl[rA % g] * kLog2e; diagA = rA / g + sk - sq + leftpad; }
    if (rB < G) { slopeB = sl[rB % g] * kLog2e; diagB = rB / g + sk - sq + leftpad; }
  }
  float mA = -CUDART_INF_F, mB = -CUDART_INF_F, lA = 0.f, lB = 0.f;
  float o[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }

  // ldmatrix source rows of this lane: K (non-transposed: matrices = [keys 0-7 | d lo], [keys 0-7 | d hi],
  // [keys 8-15 | d lo], [keys 8-15 | d hi]) and V (transposed: [keys 0-7 | d lo], [keys 8-15 | d lo], [keys 0-7 | d hi], ...)
  const int k_key = warp * 16 + (lane / 16) * 8 + (lane % 8), k_sub = (lane / 8) % 2;
  const int v_key = warp * 16 + ((lane / 8) % 2) * 8 + (lane % 8), v_sub = lane / 16;

  for (int it = 0; it < n_tiles; ++it) {
    const int st = it % kDStages;
    const uint32_t kb = sbase + L::kK + st * L::kTile, vb = sbase + L::kV + st * L::kTile;
    const int c0 = (t_lo + it) * kDT;
    mbar_wait(full_bar(st), (it / kDStages) & 1);
    bool touched = false;
    if (c0 + kDT > kv_end || c0 < leftpad) {
      // cache rows at or beyond the used length hold arbitrary bits (possibly NaN): zero this warp's V rows so that
      // P = 0 times garbage cannot poison O (the reference zero-fills the V tail: flash_fwd_kernel.h:346-349)
      for (int i = lane; i < 16 * (HD / 8); i += 32) {
        const int key = warp * 16 + i / (HD / 8), chunk = i % (HD / 8);
        if (c0 + key >= kv_end || c0 + key < leftpad) sts128(vb + tile_off(key, chunk), 0u, 0u, 0u, 0u);
      }
      __syncwarp();
      touched = true;
    }
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
      uint32_t r0, r1, r2, r3;
      ldsm_x4(kb + tile_off(k_key, ks * 2 + k_sub), r0, r1, r2, r3);
      mma16816<kBf16>(s[0], qa[ks], r0, r1);
      mma16816<kBf16>(s[1], qa[ks], r2, r3);
    }
    if (has_cap) {
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[j][e] = tanh_fast(s[j][e] * pre_cap);
    }
    // mask + online softmax (log2 domain)
    float mxA = -CUDART_INF_F, mxB = -CUDART_INF_F;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int key = c0 + warp * 16 + j * 8 + 2 * quad + e;
        float a = s[j][e] * scale_log2, b2 = s[j][2 + e] * scale_log2;
        if (has_alibi) {
          a -= slopeA * fabsf((float)(key - diagA));
          b2 -= slopeB * fabsf((float)(key - diagB));
        }
        s[j][e] = (key >= loA && key < hiA) ? a : -CUDART_INF_F;
        s[j][2 + e] = (key >= loB && key < hiB) ? b2 : -CUDART_INF_F;
        mxA = fmaxf(mxA, s[j][e]);
        mxB = fmaxf(mxB, s[j][2 + e]);
      }
    }
    mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 1)); mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 2));
    mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 1)); mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 2));
    const float nA = fmaxf(mA, mxA), nB = fmaxf(mB, mxB);
    const float subA = nA == -CUDART_INF_F ? 0.f : nA, subB = nB == -CUDART_INF_F ? 0.f : nB;
    const float corrA = ex2d(mA - subA), corrB = ex2d(mB - subB);
    mA = nA; mB = nB;
    float sumA = 0.f, sumB = 0.f;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        s[j][e] = ex2d(s[j][e] - subA); sumA += s[j][e];
        s[j][2 + e] = ex2d(s[j][2 + e] - subB); sumB += s[j][2 + e];
      }
    }
    lA = lA * corrA + sumA; lB = lB * corrB + sumB;
    uint32_t pa[4];
    pa[0] = pack2<kBf16>(s[0][0], s[0][1]); pa[1] = pack2<kBf16>(s[0][2], s[0][3]);
    pa[2] = pack2<kBf16>(s[1][0], s[1][1]); pa[3] = pack2<kBf16>(s[1][2], s[1][3]);
#pragma unroll
    for (int dp = 0; dp < HD / 16; ++dp) {
      uint32_t r0, r1, r2, r3;
      ldsm_x4_t(vb + tile_off(v_key, dp * 2 + v_sub), r0, r1, r2, r3);
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        o[2 * dp][e] *= corrA; o[2 * dp][2 + e] *= corrB;
        o[2 * dp + 1][e] *= corrA; o[2 * dp + 1][2 + e] *= corrB;
      }
      mma16816<kBf16>(o[2 * dp], pa, r0, r1);
      mma16816<kBf16>(o[2 * dp + 1], pa, r2, r3);
    }
    if (touched) fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar(st));
  }

  // ---------------- merge the four warps' partial (m, l, O) and write the result ----------------
  lA += __shfl_xor_sync(0xffffffffu, lA, 1); lA += __shfl_xor_sync(0xffffffffu, lA, 2);
  lB += __shfl_xor_sync(0xffffffffu, lB, 1); lB += __shfl_xor_sync(0xffffffffu, lB, 2);
  named_bar_sync(1, kDWarps * 32);   // every consumer is done with the ring: it becomes the merge scratch
  float *so = (float *)smem;                                   // [warp][16][HD]
  float *sm = so + kDWarps * kDRows * HD;                      // [warp][16]
  float *sl = sm + kDWarps * kDRows;                           // [warp][16]
#pragma unroll
  for (int nt = 0; nt < HD / 8; ++nt) {
    *reinterpret_cast<float2 *>(so + (warp * kDRows + lA_) * HD + nt * 8 + 2 * quad) = make_float2(o[nt][0], o[nt][1]);
    *reinterpret_cast<float2 *>(so + (warp * kDRows + lB_) * HD + nt * 8 + 2 * quad) = make_float2(o[nt][2], o[nt][3]);
  }
  if (quad == 0) {
    sm[warp * kDRows + lA_] = mA; sl[warp * kDRows + lA_] = lA;
    sm[warp * kDRows + lB_] = mB; sl[warp * kDRows + lB_] = lB;
  }
  named_bar_sync(1, kDWarps * 32);
  {
    const int row = threadIdx.x / 8, cg = threadIdx.x % 8;    // 16 rows x 8 column groups of HD / 8 floats
    const int prow = row0 + row;                              // packed row
    if (prow < G) {
      float M = -CUDART_INF_F;
#pragma unroll
      for (int w = 0; w < kDWarps; ++w) M = fmaxf(M, sm[w * kDRows + row]);
      float wgt[kDWarps], Lsum = 0.f;
#pragma unroll
      for (int w = 0; w < kDWarps; ++w) {
        wgt[w] = (M == -CUDART_INF_F) ? 0.f : ex2d(sm[w * kDRows + row] - M);
        Lsum += wgt[w] * sl[w * kDRows + row];
      }
      const bool empty = !(Lsum > 0.f);
      const float inv = empty ? 0.f : 1.f / Lsum;
      const int qi = prow / g, hq = hk * g + prow % g;
      const float lse = empty ? -CUDART_INF_F : (M + log2f(Lsum)) * kLn2;
      constexpr int kPer = HD / 8;
      float acc[kPer];
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        float a = 0.f;
#pragma unroll
        for (int w = 0; w < kDWarps; ++w) a = fmaf(wgt[w], so[(w * kDRows + row) * HD + cg * kPer + i], a);
        acc[i] = a * inv;
      }
      if (p.num_splits > 1) {
        float *oa = p.oaccum + ((((int64_t)split * p.b + bi) * p.seqlen_q + qi) * p.h + hq) * (int64_t)HD + cg * kPer;
#pragma unroll
        for (int i = 0; i < kPer; i += 4) *reinterpret_cast<float4 *>(oa + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
        if (cg == 0) p.lseaccum[(((int64_t)split * p.b + bi) * p.h + hq) * p.seqlen_q + qi] = lse;
      } else {
        T *out = (T *)p.o.ptr + (q_batched ? (int64_t)bi * p.o.batch_stride : 0) + (int64_t)(q_row0 + qi) * p.o.row_stride +
                 (int64_t)hq * p.o.head_stride + cg * kPer;
#pragma unroll
        for (int i = 0; i < kPer; i += 8) {
          uint4 v;
          v.x = pack2<kBf16>(acc[i], acc[i + 1]); v.y = pack2<kBf16>(acc[i + 2], acc[i + 3]);
          v.z = pack2<kBf16>(acc[i + 4], acc[i + 5]); v.w = pack2<kBf16>(acc[i + 6], acc[i + 7]);
          *reinterpret_cast<uint4 *>(out + i) = v;
        }
        if (cg == 0)
          p.softmax_lse[((int64_t)bi * p.h + hq) * p.seqlen_q + qi] =
              empty ? (p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F) : lse;
      }
    }
  }
}

TmaDesc4 kv_desc(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * 2; x.strides[1] = (uint64_t)t.row_stride * 2;
  x.strides[2] = (uint64_t)t.batch_stride * 2;
  x.box[0] = 64; x.box[1] = 1; x.box[2] = kDT; x.box[3] = 1;
  // a dimension of extent 1 never advances: give it a legal, monotonic stride
  for (int i = 1; i < 4; ++i)
    if (x.dims[i] == 1) x.strides[i - 1] = (i == 1 ? x.dims[0] * 2 : x.strides[i - 2] * x.dims[i - 1] + 15) / 16 * 16;
  return x;
}

bool kv_ok(const fa_b200_tensor &t, int heads, int batch) {
  if (((uintptr_t)t.ptr & 15) != 0) return false;
  if (t.row_stride % 8 != 0 || t.row_stride <= 0) return false;
  if (heads > 1 && (t.head_stride % 8 != 0 || t.head_stride <= 0)) return false;
  if (batch > 1 && (t.batch_stride % 8 != 0 || t.batch_stride <= 0)) return false;
  return true;
}

template <int HD, bool kBf16> void launch_decode(const Params &p, cudaStream_t s) {
  using L = DecSmem<HD>;
  DecTma tm;
  // contiguous cache: (d, h_k, seqlen_k, cache batch); paged: (d, h_k, page_size, num_pages)
  const bool packed_k = p.cu_seqlens_k && !p.varlen_padded_input;
  const int64_t rows = p.page_table ? p.page_size : (packed_k ? p.total_k : p.seqlen_k);
  const int slabs = p.page_table ? p.num_pages : packed_k ? 1 : (p.kv_batch_idx && p.b_k > 0 ? p.b_k : p.b);
  tm.k = make_tma_4d(kv_desc(p.k, p.d, p.h_k, rows, slabs));
  tm.v = make_tma_4d(kv_desc(p.v, p.d, p.h_k, rows, slabs));
  auto kern = decode_kernel<HD, kBf16>;
  static std::atomic<bool> smem_set[64];
  ensure_dynamic_smem(kern, L::kDynBytes, smem_set);
  const int n_rt = ceil_div(p.seqlen_q * (p.h / p.h_k), kDRows);
  dim3 grid((p.num_splits > 1 ? p.num_splits : 1) * n_rt, p.h_k, p.b);
  profile_mark(0, true, s);
  kern<<<grid, kDThreads, L::kDynBytes, s>>>(p, tm);
  profile_mark(0, false, s);
  FA_CUDA(cudaGetLastError());
  note_launch(3);
}

bool decode_supported(const Params &p) {
  if (p.force_generic) return false;
  if (p.d != 64 && p.d != 128) return false;
  if (p.lt_start || p.seqused_q) return false;
  const bool packed_k = p.cu_seqlens_k && !p.varlen_padded_input;
  if (packed_k && (p.total_k < 0 || p.page_table || p.kv_batch_idx)) return false;   // packed K/V: total rows must be known
  if (p.seqlen_q * (p.h / p.h_k) > kDMaxRows) return false;   // beyond that the tcgen05 forward is the better tool
  if (p.page_table && (p.page_size <= 0 || p.page_size % kDT != 0 || p.num_pages <= 0)) return false;
  const int slabs = p.page_table ? p.num_pages : packed_k ? 1 : (p.kv_batch_idx && p.b_k > 0 ? p.b_k : p.b);
  if (!kv_ok(p.k, p.h_k, slabs) || !kv_ok(p.v, p.h_k, slabs)) return false;
  if (((uintptr_t)p.q.ptr & 3) != 0 || p.q.row_stride % 2 != 0 || p.q.head_stride % 2 != 0 || p.q.batch_stride % 2 != 0)
    return false;
  if (p.num_splits > 1) return ((uintptr_t)p.oaccum & 15) == 0;
  return ((uintptr_t)p.o.ptr & 15) == 0 && p.o.row_stride % 8 == 0 && p.o.head_stride % 8 == 0 && p.o.batch_stride % 8 == 0;
}

}  // namespace

// kind 3 = the TMA-fed decode kernel, 2 = tcgen05 forward (many query rows, no split), 1 = generic SIMT
void splitkv_fwd(const Params &p, cudaStream_t s) {
  if (!decode_supported(p)) {
    // many query rows per kv-head (chunked prefill over a cache): the tcgen05 forward takes seqused_k and bottom-right
    // masks itself; only the cache addressing modes / softcap / ALiBi and real splits need the generic kernels
    const bool special = p.alibi_slopes || p.page_table || p.kv_batch_idx || p.leftpad_k ||
                         p.attn_mask || p.s_dmask || p.p_dropout > 0.f;
    std::string why;
    if (p.num_splits <= 1 && !p.force_generic && !special && sm100_fwd_supported(p, &why)) { sm100_fwd(p, s); return; }
    generic_splitkv_fwd(p, s);
    return;
  }
  if (p.d == 64) {
    if (p.is_bf16) launch_decode<64, true>(p, s); else launch_decode<64, false>(p, s);
  } else {
    if (p.is_bf16) launch_decode<128, true>(p, s); else launch_decode<128, false>(p, s);
  }
}

}  // namespace fa
