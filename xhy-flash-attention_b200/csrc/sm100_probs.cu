// sm100_probs.cu -- return_softmax (S_dmask) on the tensor cores.
//
// The reference's forward kernel stores the attention weights it has in registers when Return_softmax is set
// (csrc/flash_attn/src/flash_fwd_kernel.h:419,487; buffer allocated at csrc/flash_attn/flash_api.cpp:276-279), and half
// of its test grid recovers the dropout mask from that buffer (tests/test_flash_attn.py:345,218-264).  Here the
// forward proper (out, LSE) runs on the tcgen05 kernels exactly as it does without return_softmax, and this kernel
// then writes   S_dmask[b, h, row, col] = +/- exp(scale * q_row . k_col - LSE[row])   (0 where masked, negative where the
// Philox byte of (b, h, row, col) dropped the weight) in plain row-major (b, h, seqlen_q_rounded, seqlen_k_rounded):
// one CTA per (128-row query tile, head, batch element) walks the key blocks, S = Q K^T by tcgen05.mma into a
// double-buffered TMEM accumulator, four epilogue warps (one query row per thread) turn scores into probabilities and
// store them.  The work is the O(sq * sk) store: 2 bytes per weight, HBM-write bound.
//
//   warps 0-3  epilogue (TMEM lanes 32 * warp ..)      warp 4  TMA producer (Q once, K ring)      warp 5  MMA issuer
#include "common.h"
#include "sm100.cuh"

#include <algorithm>
#include <cmath>

namespace fa {
namespace {

using namespace sm100;

constexpr int kBM = 128, kBN = 128;
constexpr float kLog2e = 1.4426950408889634f;

struct ProbsTma { CUtensorMap q, k; };

template <int HD> struct PL {
  static constexpr int kTileBytes = 128 * HD * 2;
  // two CTAs per SM (2 x 256 TMEM columns, <= 113 KB of shared memory each): the epilogue of one covers the S = Q K^T of the other
  static constexpr int kStages = HD <= 64 ? 3 : 2;   // (head dims above 128: 192 KB, one CTA per SM as before)
  static constexpr int kQOff = 0;
  static constexpr int kKOff = kTileBytes;
  static constexpr int kBarOff = kKOff + kStages * kTileBytes;
  static constexpr int kNumBars = 1 + 2 * kStages + 4;
  static constexpr int kTmemSlotOff = kBarOff + kNumBars * 8;
  static constexpr int kDynBytes = kTmemSlotOff + 16 + 1024;   // + alignment slack
  static constexpr int kThreads = 192;
};

template <int HD, bool kBf16, bool kDrop>
__global__ void __launch_bounds__(PL<HD>::kThreads, 2)
probs_kernel(const __grid_constant__ Params p, const __grid_constant__ ProbsTma tm_param, const ProbsTma *__restrict__ tm_dev) {
  using L = PL<HD>;
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  const ProbsTma &tm = tm_dev ? *tm_dev : tm_param;
  constexpr int kStages = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  auto bar = [&](int i) { return sbase + L::kBarOff + i * 8; };
  constexpr int B_QFULL = 0, B_KFULL = 1, B_KEMPTY = 1 + kStages, B_SFULL = 1 + 2 * kStages, B_SEMPTY = B_SFULL + 2;
  uint32_t *tmem_slot = (uint32_t *)(smem + L::kTmemSlotOff);

  if (threadIdx.x == 0) {
    mbar_init(bar(B_QFULL), 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(bar(B_KFULL + s), 1); mbar_init(bar(B_KEMPTY + s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(bar(B_SFULL + s), 1); mbar_init(bar(B_SEMPTY + s), kBM); }
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc<256>(smem_u32(tmem_slot));
  if (warp == 4 && lane == 0) {
    if (tm_dev) { tma_acquire_desc(&tm.q); tma_acquire_desc(&tm.k); }
    tma_prefetch_desc(&tm.q); tma_prefetch_desc(&tm.k);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  const int bi = blockIdx.z, hq = blockIdx.y, m0 = blockIdx.x * kBM;
  const int hk = hq / (p.h / p.h_k);
  int q_row0, sq, k_row0, sk;
  bool q_batched, k_batched;
  seq_bounds(p, bi, q_row0, sq, k_row0, sk, q_batched, k_batched);
  const MaskCtx m = make_mask_ctx(p, bi, hq, sq, sk);
  // key blocks that causal / window / the sequence length leave open for the rows of this tile: S is only computed there
  int lo_blk = 0, hi_blk = 0;
  if (m0 < sq) {
    int lo_a, hi_a, lo_b, hi_b;
    visible_cols(m, m0, lo_a, hi_a);
    visible_cols(m, min(m0 + kBM, sq) - 1, lo_b, hi_b);
    const int lo_c = max(min(lo_a, lo_b), 0), hi_c = min(max(hi_a, hi_b), sk);
    if (hi_c > lo_c) { lo_blk = lo_c / kBN; hi_blk = ceil_div(hi_c, kBN); }
  }
  const int nvis = hi_blk - lo_blk;
  const int nb_total = p.seqlen_k_rounded / kBN;
  constexpr int kBoxes = HD / 64;
  constexpr uint32_t kBoxBytes = kBM * 128;

  if (warp == 4) {
    // ======================= TMA producer =======================
    if (lane == 0 && nvis > 0) {
      mbar_expect_tx(bar(B_QFULL), L::kTileBytes);
#pragma unroll
      for (int bx = 0; bx < kBoxes; ++bx)
        tma_load_4d(sbase + L::kQOff + bx * kBoxBytes, &tm.q, bar(B_QFULL), bx * 64, hq, q_row0 + m0, q_batched ? bi : 0);
      for (int n = 0; n < nvis; ++n) {
        const uint32_t st = n % kStages, ph = (n / kStages) & 1;
        mbar_wait(bar(B_KEMPTY + st), ph ^ 1);
        mbar_expect_tx(bar(B_KFULL + st), L::kTileBytes);
#pragma unroll
        for (int bx = 0; bx < kBoxes; ++bx)
          tma_load_4d(sbase + L::kKOff + st * L::kTileBytes + bx * kBoxBytes, &tm.k, bar(B_KFULL + st), bx * 64, hk,
                      k_row0 + (lo_blk + n) * kBN, k_batched ? bi : 0);
      }
    }
  } else if (warp == 5) {
    // ======================= MMA issuer =======================
    constexpr uint32_t idesc_s = make_idesc(kBf16, kBM, kBN, false, false);
    if (nvis > 0) mbar_wait(bar(B_QFULL), 0);
    for (int n = 0; n < nvis; ++n) {
      const uint32_t st = n % kStages, slot = n & 1;
      mbar_wait(bar(B_KFULL + st), (n / kStages) & 1);
      mbar_wait(bar(B_SEMPTY + slot), ((n >> 1) & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = sbase + L::kQOff, ka = sbase + L::kKOff + st * L::kTileBytes;
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) {
          const uint32_t off = (k / 4) * kBoxBytes + (k % 4) * 32;
          umma_ss(tmem + slot * kBN, make_sdesc_sw128(qa + off, 16, 1024), make_sdesc_sw128(ka + off, 16, 1024), idesc_s, k > 0);
        }
        umma_commit(bar(B_SFULL + slot));
        umma_commit(bar(B_KEMPTY + st));
      }
      __syncwarp();
    }
  } else {
    // ======================= epilogue: one query row per thread =======================
    const int row = m0 + warp * 32 + lane;
    const bool valid = row < sq;
    float lse2 = 0.f;
    bool dead = !valid;
    if (valid) {
      const float lse = p.softmax_lse[((int64_t)bi * p.h + hq) * p.seqlen_q + row];
      dead = isinf(lse) || isnan(lse);     // fully masked row: every weight is 0
      lse2 = lse * kLog2e;
    }
    const float scale_log2 = p.softmax_scale * kLog2e;
    T *sd = (T *)p.s_dmask + (((int64_t)bi * p.h + hq) * p.seqlen_q_rounded + (valid ? row : 0)) * (int64_t)p.seqlen_k_rounded;
    DropCtx dc;
    if constexpr (kDrop) dc = make_drop_ctx(p, bi, hq, valid ? row : 0);
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    const bool elementwise = m.lts != nullptr;   // FlashMask vectors: per-element rule; otherwise an interval of columns
    int vlo = 0, vhi = 0;
    if (valid && !dead) visible_cols(m, row, vlo, vhi);
    for (int jb = 0; jb < nb_total; ++jb) {
      const bool vis = jb >= lo_blk && jb < hi_blk;
      const int n = jb - lo_blk, slot = n & 1;
      if (vis) {
        mbar_wait(bar(B_SFULL + slot), (n >> 1) & 1);
        tc_fence_after();
      }
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t s[32];
        if (vis) {
          tmem_ld32(tmem + lane_base + slot * kBN + c * 32, s);
          tmem_wait_ld();
        }
        const int col0 = jb * kBN + c * 32;
        uint32_t packed[16];
        const bool any = vis && valid && !dead && col0 < vhi && col0 + 32 > vlo;
        if (!any) {
#pragma unroll
          for (int i = 0; i < 16; ++i) packed[i] = 0u;
        } else {
          float pr[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int col = col0 + i;
            bool open = col >= vlo && col < vhi;
            if (elementwise && open) open = !is_masked(m, row, col);
            pr[i] = open ? exp2f(fmaf(__uint_as_float(s[i]), scale_log2, -lse2)) : 0.f;
          }
          if constexpr (kDrop) {
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              const uint4 r = philox4x32_7(dc.seed, dc.row_base + (uint32_t)((col0 >> 4) + g), dc.offset);
              const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (((w[i >> 2] >> ((i & 3) * 8)) & 0xffu) > dc.threshold) pr[g * 16 + i] = -pr[g * 16 + i];
            }
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) packed[i] = pack2<kBf16>(pr[2 * i], pr[2 * i + 1]);
        }
        if (valid) {
          uint4 *dst = reinterpret_cast<uint4 *>(sd + col0);
#pragma unroll
          for (int i = 0; i < 4; ++i) dst[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
        }
      }
      if (vis) {
        tc_fence_before();
        mbar_arrive(bar(B_SEMPTY + slot));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc<256>(tmem);
}

TmaDesc4 probs_desc(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * 2; x.strides[1] = (uint64_t)t.row_stride * 2;
  x.strides[2] = (uint64_t)t.batch_stride * 2;
  x.box[0] = 64; x.box[1] = 1; x.box[2] = 128; x.box[3] = 1;
  // a dimension of extent 1 never advances: give it a legal, monotonic stride
  for (int i = 1; i < 4; ++i)
    if (x.dims[i] == 1) x.strides[i - 1] = (i == 1 ? x.dims[0] * 2 : x.strides[i - 2] * x.dims[i - 1] + 15) / 16 * 16;
  return x;
}

template <int HD, bool kBf16, bool kDrop> void launch_probs(const Params &p, cudaStream_t s) {
  using L = PL<HD>;
  const bool packed_q = p.cu_seqlens_q && !p.varlen_padded_input;
  const bool packed_k = p.cu_seqlens_k && !p.varlen_padded_input;
  const bool dev_rows = (packed_q && p.total_q < 0) || (packed_k && p.total_k < 0);
  const int64_t rows_q = packed_q ? (p.total_q >= 0 ? p.total_q : (int64_t)p.b * p.seqlen_q) : p.seqlen_q;
  const int64_t rows_k = packed_k ? (p.total_k >= 0 ? p.total_k : (int64_t)p.b * p.seqlen_k) : p.seqlen_k;
  ProbsTma tm;
  // columns at or beyond p.d are zero-filled by the TMA unit: they add nothing to the dot products
  tm.q = make_tma_4d(probs_desc(p.q, p.d, p.h, rows_q, packed_q ? 1 : p.b));
  tm.k = make_tma_4d(probs_desc(p.k, p.d, p.h_k, rows_k, packed_k ? 1 : p.b));
  ProbsTma *tm_dev = nullptr;
  if (dev_rows) {
    const int k_like[2] = {0, 1};
    tm_dev = reinterpret_cast<ProbsTma *>(patch_tma_rows(&tm.q, k_like, 2, p.cu_seqlens_q, p.cu_seqlens_k, p.b, s));
  }
  auto kern = probs_kernel<HD, kBf16, kDrop>;
  static std::atomic<bool> smem_set[64];
  ensure_dynamic_smem(kern, L::kDynBytes, smem_set);
  const dim3 grid(ceil_div(p.seqlen_q, kBM), p.h, p.b);
  kern<<<grid, L::kThreads, L::kDynBytes, s>>>(p, tm, tm_dev);
  const cudaError_t launch_err = cudaGetLastError();
  if (tm_dev) cudaFreeAsync(tm_dev, s);
  FA_CUDA(launch_err);
  note_launch(2);
}

template <int HD> void probs_hd(const Params &p, cudaStream_t s) {
  const bool drop = p.p_dropout > 0.f;
  if (p.is_bf16) { if (drop) launch_probs<HD, true, true>(p, s); else launch_probs<HD, true, false>(p, s); }
  else { if (drop) launch_probs<HD, false, true>(p, s); else launch_probs<HD, false, false>(p, s); }
}

}  // namespace

// The logit of this pass is scale * q . k: calls whose logits carry more (dense additive mask, softcap, ALiBi, the FA1
// bias) or whose keys are not plain rows (paged / indexed / left-padded caches) keep the element-wise generic pass.
bool sm100_probs_eligible(const Params &p) {
  return p.s_dmask != nullptr && p.attn_mask == nullptr && p.softcap == 0.f && p.alibi_slopes == nullptr &&
         p.attn_bias == nullptr && p.page_table == nullptr && p.kv_batch_idx == nullptr && p.leftpad_k == nullptr &&
         p.d % 8 == 0 && p.d <= 256 && p.seqlen_k_rounded % kBN == 0 && p.seqlen_k_rounded >= p.seqlen_k &&
         p.seqlen_q_rounded >= p.seqlen_q && p.num_splits <= 1;
}

void sm100_probs(const Params &p, cudaStream_t s) {
  if (p.d <= 64) probs_hd<64>(p, s);
  else if (p.d <= 128) probs_hd<128>(p, s);
  else probs_hd<256>(p, s);
}

}  // namespace fa
