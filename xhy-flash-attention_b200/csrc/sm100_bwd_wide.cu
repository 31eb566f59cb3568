// sm100_bwd_wide.cu -- the tcgen05 / TMEM / TMA backward for head dims 129..256 (sm_100a).
//
// Same mathematics as the reference's compute_dq_dk_dv_1colblock (csrc/flash_attn/src/flash_bwd_kernel.h:427-1113);
// the reference serves these head dims with 64 x 64 / 64 x 80 tiles and, in its v2 kernels, by slicing the dQ/dK/dV
// GEMMs over the head dim (Slice_dQKV_Mma, csrc/flashmask_v2/flash_bwd_launch_template.h:371-472).  On Blackwell the
// limits are TMEM (512 columns: dK and dV of a 128-key block at d = 256 would take all of them) and shared memory
// (four 128 x 256 operand tiles would take 256 KB).  Hence:
//
//   * the OUTPUT head dim is cut into two slices of 128 columns; a CTA owns (128-key block, kv-head, batch, slice) and
//     produces dV[:, slice], dK[:, slice] (accumulated in TMEM over the whole GQA group, no atomics) and its part of
//     dQ[:, slice] (fp32 reduce-add into the tiled dq_accum, as for d <= 128);
//   * S^T = K Q^T and dP^T = V dO^T contract over the FULL head dim (256 columns, zero-filled by TMA beyond d) and are
//     recomputed by both slice CTAs (14 GEMM units instead of 10);
//   * query blocks are 64 rows (K 64 KB + V 64 KB + Q 32 KB + dO 32 KB + dS^T 16 KB + dQ staging 16 KB = 224 KB);
//   * dQ is formed TRANSPOSED, dQ^T[slice] = K[:, slice]^T dS^T (M = 128 head-dim columns on the TMEM lanes, N = 64
//     queries) -- an MMA with M = 64 queries would run at half rate -- and the four reducer warps, one per 32-column
//     chunk of dq_accum, transpose it back through shared memory.
//
//   warp 12   : MMA issuer   S^T = K Q^T, dP^T = V dO^T (SS); dV += P^T dO_s, dK += dS^T Q_s (TS); dQ^T = K_s^T dS^T (SS)
//   warp 13   : producer     TMA loads of K, V (once), Q and dO (one stage each); LSE*log2e and dPsum rows into smem
//   warps 0-7 : compute      two warpgroups, each owns 32 of the 64 query columns of S^T / dP^T
//   warps 8-11: dQ reducers  warp w owns dq_accum chunk slice*4 + w
// TMEM columns: S^T/P^T [0,64)  dP^T/dS^T [64,128)  dQ^T [128,192)  dV [256,384)  dK [384,512).
//
// Features: causal (both alignments), sliding window, varlen, seqused, GQA/MQA, per-q-head dK/dV (dkv_expanded),
// bf16/fp16, dropout (kDrop).  FlashMask / dense mask / softcap / deterministic with d > 128 stay on the generic kernels.
#include <math_constants.h>
#include <type_traits>

#include "common.h"
#include "sm100.cuh"

namespace fa {
namespace {

using namespace sm100;

constexpr int kBK = 128;           // keys per block (TMEM lanes)
constexpr int kBQ = 64;            // query rows per block
constexpr int kHDC = 256;          // contraction width of S^T / dP^T
constexpr int kSL = 128;           // output slice width
constexpr int kWThreads = 512;
constexpr float kLog2e = 1.4426950408889634f;

struct WTma {
  CUtensorMap q, k, v, dout;
};

struct WSmem {
  static constexpr int kKBox = kBK * 128;                 // one K/V box: 128 rows x 64 columns (16 KB)
  static constexpr int kQBox = kBQ * 128;                 // one Q/dO box: 64 rows x 64 columns (8 KB)
  static constexpr int kK = 0;
  static constexpr int kV = kK + 4 * kKBox;
  static constexpr int kQ = kV + 4 * kKBox;
  static constexpr int kDO = kQ + 4 * kQBox;
  static constexpr int kDS = kDO + 4 * kQBox;             // dS^T: 128 keys x 64 queries (128 bytes per row)
  static constexpr int kStg = kDS + kBK * 128;            // dQ staging: 4 warps x 2 buffers x (16 rows x 32 floats)
  static constexpr int kVec = kStg + 4 * 2 * 16 * 128;    // [lse2 | dpsum][64] floats
  static constexpr int kBar = kVec + 2 * kBQ * 4;
  static constexpr int kNumBars = 16;
  static constexpr int kTmemSlot = kBar + kNumBars * 8;
  static constexpr int kTotal = kTmemSlot + 16;
  static_assert(kTotal <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");
};

enum WBar {
  WB_KV_FULL = 0, WB_Q_FULL, WB_Q_EMPTY, WB_DO_FULL, WB_DO_EMPTY, WB_S_FULL, WB_P_FULL, WB_DP_FULL, WB_DS_FULL,
  WB_DQ_FULL, WB_DQ_EMPTY, WB_DKV_FULL
};

__device__ __forceinline__ float ex2w(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct WTile {
  int bi, hk, ho, n0, slice;
  int hq0, hq1;
  int sq, sk, q_row0, k_row0, q_b, k_b;
  int m_lo, m_hi;            // 64-row query blocks
  int acc_tile0, acc_tiles_per_head;
};

__device__ __forceinline__ bool wide_decode(const Params &p, WTile &t) {
  t.bi = blockIdx.z;
  t.ho = blockIdx.y;
  t.slice = blockIdx.x & 1;                 // sibling slices next to each other: they read the same tiles (L2)
  const int kb = blockIdx.x >> 1;
  const int g = p.h / p.h_k;
  if (p.dkv_expanded) { t.hk = t.ho / g; t.hq0 = t.ho; t.hq1 = t.ho + 1; }
  else { t.hk = t.ho; t.hq0 = t.ho * g; t.hq1 = t.ho * g + g; }
  bool qb, kbt;
  seq_bounds(p, t.bi, t.q_row0, t.sq, t.k_row0, t.sk, qb, kbt);
  t.q_b = qb ? t.bi : 0;
  t.k_b = kbt ? t.bi : 0;
  t.n0 = kb * kBK;
  if (t.n0 >= t.sk) return false;
  const int shift = t.sk - t.sq;
  const int shiftc = p.causal_bottom_right ? shift : 0;
  int rlo = 0, rhi = t.sq;
  const int c_last = min(t.n0 + kBK, t.sk) - 1;
  if (p.is_causal) rlo = max(rlo, t.n0 - shiftc);
  if (p.window_right >= 0) rlo = max(rlo, t.n0 - shift - p.window_right);
  if (p.window_left >= 0) rhi = min(rhi, c_last - shift + p.window_left + 1);
  t.m_lo = rlo / kBQ;
  t.m_hi = (max(rhi, 0) + kBQ - 1) / kBQ;
  if (t.m_hi < t.m_lo) t.m_hi = t.m_lo;
  if (p.cu_seqlens_q && !p.varlen_padded_input && !p.dq_accum_batch_major) {
    t.acc_tile0 = (p.cu_seqlens_q[t.bi] + 128 * t.bi) / 128;
    t.acc_tiles_per_head = (int)((p.total_q + (int64_t)128 * p.b) / 128);
  } else {
    t.acc_tiles_per_head = p.seqlen_q_rounded / 128;
    t.acc_tile0 = t.bi * p.h * t.acc_tiles_per_head;
  }
  return true;
}

// dq_accum tile (128 rows) that holds 64-row query block m of head hq
__device__ __forceinline__ int64_t wide_acc_tile(const Params &p, const WTile &t, int hq, int m) {
  if (p.cu_seqlens_q && !p.varlen_padded_input) return (int64_t)hq * t.acc_tiles_per_head + t.acc_tile0 + (m >> 1);
  return (int64_t)t.acc_tile0 + (int64_t)hq * t.acc_tiles_per_head + (m >> 1);
}

template <bool kBf16, bool kDrop>
__global__ void __launch_bounds__(kWThreads, 1)
bwd_wide_kernel(const __grid_constant__ Params p, const __grid_constant__ WTma tm_param, const WTma *__restrict__ tm_dev) {
  using L = WSmem;
  const WTma &tm = tm_dev ? *tm_dev : tm_param;   // device-patched maps (flat varlen symbols) or the launch's own
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();  // 128-byte swizzled tiles need a 1024-byte aligned base
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  auto bar = [&](int i) { return sbase + L::kBar + i * 8; };
  uint32_t *tmem_slot = (uint32_t *)(smem + L::kTmemSlot);
  float *vec = (float *)(smem + L::kVec);  // [0: lse*log2e, 1: dpsum][64]

  WTile t;
  if (!wide_decode(p, t)) return;  // uniform for the whole CTA
  const int nm = t.m_hi - t.m_lo;
  const int n_it = (t.hq1 - t.hq0) * nm;
  const int n_it_u = __shfl_sync(0xffffffffu, n_it, 0);

  if (threadIdx.x == 0) {
    mbar_init(bar(WB_KV_FULL), 1);
    mbar_init(bar(WB_Q_FULL), 1); mbar_init(bar(WB_Q_EMPTY), 1);
    mbar_init(bar(WB_DO_FULL), 1); mbar_init(bar(WB_DO_EMPTY), 1);
    mbar_init(bar(WB_S_FULL), 1); mbar_init(bar(WB_P_FULL), 256);
    mbar_init(bar(WB_DP_FULL), 1); mbar_init(bar(WB_DS_FULL), 256);
    mbar_init(bar(WB_DQ_FULL), 1); mbar_init(bar(WB_DQ_EMPTY), 128);
    mbar_init(bar(WB_DKV_FULL), 1);
    fence_mbar_init();
  }
  if (warp == 12) tmem_alloc<512>(smem_u32(tmem_slot));
  if (warp == 13 && lane == 0) {
    if (tm_dev) { tma_acquire_desc(&tm.q); tma_acquire_desc(&tm.k); tma_acquire_desc(&tm.v); tma_acquire_desc(&tm.dout); }
    tma_prefetch_desc(&tm.q); tma_prefetch_desc(&tm.k); tma_prefetch_desc(&tm.v); tma_prefetch_desc(&tm.dout);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  constexpr uint32_t TS_ = 0, TDP = 64, TDQ = 128, TDV = 256, TDK = 384;  // TMEM column map
  const float scale_log2 = p.softmax_scale * kLog2e;

  if (warp >= 12) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 13) {
      // ======================= producer =======================
      if (lane == 0 && n_it > 0) {
        mbar_expect_tx(bar(WB_KV_FULL), 8 * L::kKBox);
#pragma unroll
        for (int bx = 0; bx < 4; ++bx) {
          tma_load_4d(sbase + L::kK + bx * L::kKBox, &tm.k, bar(WB_KV_FULL), bx * 64, t.hk, t.k_row0 + t.n0, t.k_b);
          tma_load_4d(sbase + L::kV + bx * L::kKBox, &tm.v, bar(WB_KV_FULL), bx * 64, t.hk, t.k_row0 + t.n0, t.k_b);
        }
      }
      for (int it = 0; it < n_it; ++it) {
        const int hq = t.hq0 + it / nm, m = t.m_lo + it % nm;
        mbar_wait(bar(WB_Q_EMPTY), (it & 1) ^ 1);
        const float *lse = p.softmax_lse + ((int64_t)t.bi * p.h + hq) * p.seqlen_q;
        const float *dps = p.softmax_d + ((int64_t)t.bi * p.h + hq) * p.seqlen_q_rounded;
#pragma unroll
        for (int i = lane; i < kBQ; i += 32) {
          const int row = m * kBQ + i;
          float l = CUDART_INF_F, dsum = 0.f;
          if (row < t.sq) {
            l = lse[row];
            l = isinf(l) ? CUDART_INF_F : l * kLog2e;
            dsum = dps[row];
          }
          vec[i] = l;
          vec[kBQ + i] = dsum;
        }
        __syncwarp();
        if (lane == 0) {
          mbar_expect_tx(bar(WB_Q_FULL), 4 * L::kQBox);
#pragma unroll
          for (int bx = 0; bx < 4; ++bx)
            tma_load_4d(sbase + L::kQ + bx * L::kQBox, &tm.q, bar(WB_Q_FULL), bx * 64, hq, t.q_row0 + m * kBQ, t.q_b);
          mbar_wait(bar(WB_DO_EMPTY), (it & 1) ^ 1);
          mbar_expect_tx(bar(WB_DO_FULL), 4 * L::kQBox);
#pragma unroll
          for (int bx = 0; bx < 4; ++bx)
            tma_load_4d(sbase + L::kDO + bx * L::kQBox, &tm.dout, bar(WB_DO_FULL), bx * 64, hq, t.q_row0 + m * kBQ, t.q_b);
        }
        __syncwarp();
      }
    } else if (warp == 12 && n_it_u > 0) {
      // ======================= MMA issuer =======================
      const uint32_t tmb = (uint32_t)__shfl_sync(0xffffffffu, (int)tmem, 0);
      const int slice = __shfl_sync(0xffffffffu, t.slice, 0);
      constexpr uint32_t idesc_st = make_idesc(kBf16, kBK, kBQ, false, false);   // S^T, dP^T: K-major x K-major, N = 64
      constexpr uint32_t idesc_kv = make_idesc(kBf16, kBK, kSL, false, true);    // dV, dK: A in TMEM, B MN-major
      constexpr uint32_t idesc_dq = make_idesc(kBf16, kSL, kBQ, true, true);     // dQ^T: A (K slice) MN-major, B (dS^T) MN-major
      const uint32_t sK = sbase + L::kK, sV = sbase + L::kV, sQ = sbase + L::kQ, sDO = sbase + L::kDO, sDS = sbase + L::kDS;
      // D[128 keys][64 q] = A[128 keys][256] . B[64 q][256]^T, both K-major, 16 K16 steps over four 64-column boxes
      auto contract = [&](uint32_t d_col, uint32_t a_base, uint32_t b_base, uint32_t done_bar) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kHDC / 16; ++k) {
            const uint32_t ao = (k / 4) * L::kKBox + (k % 4) * 32, bo = (k / 4) * L::kQBox + (k % 4) * 32;
            umma_ss(tmb + d_col, make_sdesc_sw128(a_base + ao, 16, 1024), make_sdesc_sw128(b_base + bo, 16, 1024), idesc_st,
                    k > 0);
          }
          umma_commit(done_bar);
        }
        __syncwarp();
      };
      // D[128 keys][128 slice cols] (+)= A[128 keys][64 q] (TMEM, packed 16-bit) . B[64 q][slice] (MN-major boxes 2s, 2s+1)
      auto outer = [&](uint32_t d_col, uint32_t a_col, uint32_t b_base, bool accumulate, uint32_t done_bar) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBQ / 16; ++k) {
            const uint32_t a = tmb + a_col + (k < 2 ? k * 8 : 32 + (k - 2) * 8);   // each compute warpgroup packs into its own half
            umma_ts(tmb + d_col, a, make_sdesc_sw128(b_base + 2 * slice * L::kQBox + k * 2048, L::kQBox, 1024), idesc_kv,
                    (accumulate || k > 0) ? 1u : 0u);
          }
          umma_commit(done_bar);
        }
        __syncwarp();
      };
      mbar_wait(bar(WB_KV_FULL), 0);
      if (p.seqused_k != nullptr && t.n0 + kBK > t.sk) {
        // KV cache: rows at or beyond the used length hold arbitrary bits (possibly NaN): zero the tail rows of both tiles
        const int first = max(t.sk - t.n0, 0);
        for (int i = first * 8 + lane; i < kBK * 8; i += 32)
#pragma unroll
          for (int bx = 0; bx < 4; ++bx) {
            sts128(sK + bx * L::kKBox + i * 16, 0u, 0u, 0u, 0u);
            sts128(sV + bx * L::kKBox + i * 16, 0u, 0u, 0u, 0u);
          }
        fence_proxy_async_smem();
        __syncwarp();
      }
      mbar_wait(bar(WB_Q_FULL), 0);
      tc_fence_after();
      contract(TS_, sK, sQ, bar(WB_S_FULL));
      for (int it = 0; it < n_it_u; ++it) {
        mbar_wait(bar(WB_DO_FULL), it & 1);
        tc_fence_after();
        contract(TDP, sV, sDO, bar(WB_DP_FULL));       // in order behind dK(it-1), which read dS^T from this region
        mbar_wait(bar(WB_P_FULL), it & 1);
        tc_fence_after();
        outer(TDV, TS_, sDO, it > 0, bar(WB_DO_EMPTY));
        mbar_wait(bar(WB_DS_FULL), it & 1);
        tc_fence_after();
        outer(TDK, TDP, sQ, it > 0, bar(WB_Q_EMPTY));
        if (it > 0) mbar_wait(bar(WB_DQ_EMPTY), (it - 1) & 1);   // the reducers have read dQ^T(it-1) out of TMEM
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)
            umma_ss(tmb + TDQ, make_sdesc_sw128(sK + 2 * slice * L::kKBox + k * 2048, L::kKBox, 1024),
                    make_sdesc_sw128(sDS + k * 2048, 1024, 1024), idesc_dq, k > 0);
          umma_commit(bar(WB_DQ_FULL));
        }
        __syncwarp();
        if (it + 1 < n_it_u) {
          mbar_wait(bar(WB_Q_FULL), (it + 1) & 1);
          tc_fence_after();
          contract(TS_, sK, sQ, bar(WB_S_FULL));       // in order behind dV(it), which read P^T from this region
        }
      }
      if (elect_one()) umma_commit(bar(WB_DKV_FULL));
      __syncwarp();
    }
  } else if (warp >= 8) {
    // ======================= dQ reducers =======================
    // dQ^T sits in TMEM with the slice's head-dim columns on the lanes: warp w holds columns 32w..32w+31 of the slice =
    // one 32-float chunk of dq_accum, 64 query rows deep.  Each lane writes its column into the staging rows (a row of
    // 32 floats per query, the accumulator's swizzle applied), 16 rows at a time, and lane 0 issues the bulk reduce-add.
    const int w = warp - 8;                                  // (keeps the 128 registers of the launch)
    const uint32_t lane_base = (uint32_t)(w * 32) << 16;
    const int chunk = t.slice * 4 + w;                       // 32-float chunk of the head dim
    const bool chunk_ok = chunk < p.d_rounded / 32;
    uint8_t *stg = smem + L::kStg + w * (2 * 16 * 128);
    uint32_t gcount = 0;                                     // bulk groups committed by this warp
    for (int it = 0; it < n_it; ++it) {
      const int hq = t.hq0 + it / nm, m = t.m_lo + it % nm;
      float *gtile = p.dq_accum + (wide_acc_tile(p, t, hq, m) * (int64_t)(p.d_rounded / 32) + chunk) * (128 * 32) +
                     (m & 1) * (kBQ * 32);
      mbar_wait(bar(WB_DQ_FULL), it & 1);
      tc_fence_after();
      uint32_t v[2][32];
      tmem_ld32(tmem + lane_base + TDQ, v[0]);
      tmem_ld32(tmem + lane_base + TDQ + 32, v[1]);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(bar(WB_DQ_EMPTY));
      if (chunk_ok) {
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {               // 16 query rows per pass, two staging buffers
          uint8_t *buf = stg + (gcount & 1) * (16 * 128);
          if (lane == 0 && gcount >= 2) bulk_wait_read<1>();  // the group that last used this buffer has been read out
          __syncwarp();
#pragma unroll
          for (int r = 0; r < 16; ++r) {
            const int q = pass * 16 + r;                     // row inside the 64-row block; its swizzle phase is q % 8
            *reinterpret_cast<uint32_t *>(buf + r * 128 + (((lane >> 2) ^ (q & 7)) * 16) + (lane & 3) * 4) = v[q >> 5][q & 31];
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            bulk_reduce_add_f32(gtile + pass * (16 * 32), smem_u32(buf), 16 * 128);
            bulk_commit();
          }
          ++gcount;
        }
      }
    }
    if (lane == 0) bulk_wait<0>();
  } else {
    // ======================= compute warpgroups =======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 136;");
    const int wg = warp / 4;
    const int r = (warp % 4) * 32 + lane;          // key row inside the block == TMEM lane
    const int col = t.n0 + r;                      // key index
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const uint32_t tS = tmem + lane_base + TS_ + wg * 32, tP = tS;
    const uint32_t tDP = tmem + lane_base + TDP + wg * 32, tDS = tDP;
    const uint32_t ds_row_a = sbase + L::kDS + r * 128;
    const int shift = t.sk - t.sq;
    const int shiftc = p.causal_bottom_right ? shift : 0;
    const float rp_drop = kDrop ? 1.f / (1.f - p.p_dropout) : 1.f;
    const uint32_t lse2_a = sbase + L::kVec + (wg * 32) * 4;
    const uint32_t dpsum_a = sbase + L::kVec + (kBQ + wg * 32) * 4;
    for (int it = 0; it < n_it; ++it) {
      const int hq = t.hq0 + it / nm, m = t.m_lo + it % nm;
      const int m0 = m * kBQ;
      const MaskCtx mc = make_mask_ctx(p, t.bi, hq, t.sq, t.sk);
      const bool full = (t.n0 + kBK <= t.sk) && (!mc.causal || t.n0 + kBK - 1 <= m0 + shiftc) &&
                        (mc.wr < 0 || t.n0 + kBK - 1 <= m0 + shift + mc.wr) &&
                        (mc.wl < 0 || t.n0 >= m0 + kBQ - 1 + shift - mc.wl);
      uint32_t mbits = 0u;
      if (!full) {   // seqlen tail / causal / window: the visible rows of a key column form an interval
        int vlo, vhi;
        visible_rows(mc, col, vlo, vhi);
        mbits = outside_bits(vlo, vhi, m0 + wg * 32);
      }
      // Dropout (kDrop): keep bits of this thread's 32 (query, key) pairs -- the scheme of sm100_bwd.cu: a lane runs the
      // generator for one query row of a 16-row group and the 16 key columns of its half warp, turns the 16 bytes into a
      // 16-bit keep mask, and the half warp transposes the 16 x 16 bit matrix with four butterfly shuffles.
      uint32_t kb = 0xffffffffu;
      if constexpr (kDrop) {
        kb = 0u;
        const int l16 = lane & 15;
        const uint32_t thr4 = (uint32_t)floorf((1.f - p.p_dropout) * 255.f) * 0x01010101u;
        const uint64_t ncg = (uint64_t)((p.seqlen_k + 15) / 16);
        const uint64_t cg = (uint64_t)((t.n0 + (warp % 4) * 32 + (lane & 16)) >> 4);
        const uint64_t head_base = ((uint64_t)t.bi * p.h + hq) * (uint64_t)p.seqlen_q;
        auto keep4 = [&](uint32_t w) { return ((__vcmpleu4(w, thr4) & 0x08040201u) * 0x01010101u) >> 24; };
#pragma unroll
        for (int ig = 0; ig < 2; ++ig) {
          const uint64_t rrow = (uint64_t)(m0 + wg * 32 + ig * 16 + l16);
          const uint4 rnd = philox4x32_7(p.philox_seed, (head_base + rrow) * ncg + cg, p.philox_offset);
          uint32_t x = keep4(rnd.x) | (keep4(rnd.y) << 4) | (keep4(rnd.z) << 8) | (keep4(rnd.w) << 12);
#pragma unroll
          for (int s_ = 8; s_ >= 1; s_ >>= 1) {
            const uint32_t lo = s_ == 8 ? 0x00ffu : s_ == 4 ? 0x0f0fu : s_ == 2 ? 0x3333u : 0x5555u;
            const uint32_t y = __shfl_xor_sync(0xffffffffu, x, s_);
            x = (l16 & s_) ? (((y >> s_) & lo) | (x & (lo << s_))) : ((x & lo) | ((y & lo) << s_));
          }
          kb |= x << (ig * 16);
        }
      }
      mbar_wait(bar(WB_Q_FULL), it & 1);     // LSE / dPsum rows of this query block are in smem
      mbar_wait(bar(WB_S_FULL), it & 1);
      tc_fence_after();
      float pr[32];
      {
        uint32_t s[32];
        tmem_ld32(tS, s);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 l4 = lds128(lse2_a + i * 4);
          pr[i + 0] = ex2w(fmaf(__uint_as_float(s[i + 0]), scale_log2, -l4.x));
          pr[i + 1] = ex2w(fmaf(__uint_as_float(s[i + 1]), scale_log2, -l4.y));
          pr[i + 2] = ex2w(fmaf(__uint_as_float(s[i + 2]), scale_log2, -l4.z));
          pr[i + 3] = ex2w(fmaf(__uint_as_float(s[i + 3]), scale_log2, -l4.w));
        }
        if (!full) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if ((mbits >> i) & 1u) pr[i] = 0.f;
        }
        uint32_t pk[16];
        if constexpr (kDrop) {   // dV sees the dropped weights (1 / p_keep is applied to dV in the epilogue)
#pragma unroll
          for (int i = 0; i < 16; ++i)
            pk[i] = pack2<kBf16>(((kb >> (2 * i)) & 1u) ? pr[2 * i] : 0.f, ((kb >> (2 * i + 1)) & 1u) ? pr[2 * i + 1] : 0.f);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = pack2<kBf16>(pr[2 * i], pr[2 * i + 1]);
        }
        tmem_st16(tP, pk);
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(bar(WB_P_FULL));

      mbar_wait(bar(WB_DP_FULL), it & 1);
      if (it > 0) mbar_wait(bar(WB_DQ_FULL), (it - 1) & 1);  // dQ^T(it-1) no longer reads the dS^T smem tile
      tc_fence_after();
      {
        uint32_t dp[32];
        tmem_ld32(tDP, dp);
        tmem_wait_ld();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 d4 = lds128(dpsum_a + i * 4);
          float g0 = __uint_as_float(dp[i + 0]), g1 = __uint_as_float(dp[i + 1]), g2 = __uint_as_float(dp[i + 2]),
                g3 = __uint_as_float(dp[i + 3]);
          if constexpr (kDrop) {   // dP = keep / p_keep * (dO . V)
            g0 = ((kb >> (i + 0)) & 1u) ? g0 * rp_drop : 0.f;
            g1 = ((kb >> (i + 1)) & 1u) ? g1 * rp_drop : 0.f;
            g2 = ((kb >> (i + 2)) & 1u) ? g2 * rp_drop : 0.f;
            g3 = ((kb >> (i + 3)) & 1u) ? g3 * rp_drop : 0.f;
          }
          const float e0 = pr[i + 0] * (g0 - d4.x);
          const float e1 = pr[i + 1] * (g1 - d4.y);
          const float e2 = pr[i + 2] * (g2 - d4.z);
          const float e3 = pr[i + 3] * (g3 - d4.w);
          pk[i / 2] = pack2<kBf16>(e0, e1);
          pk[i / 2 + 1] = pack2<kBf16>(e2, e3);
        }
        tmem_st16(tDS, pk);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          sts128(ds_row_a + (((wg * 4 + j) ^ (r & 7)) * 16), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
      }
      tmem_wait_st();
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar(WB_DS_FULL));
    }
    // ---------------- epilogue: dK (x scale), dV: this slice's 128 columns, 64 per warpgroup ----------------
    const bool row_ok = col < t.sk;
    T *dk_row = (T *)p.dk.ptr + (int64_t)t.k_b * p.dk.batch_stride + (int64_t)(t.k_row0 + col) * p.dk.row_stride +
                (int64_t)t.ho * p.dk.head_stride;
    T *dv_row = (T *)p.dv.ptr + (int64_t)t.k_b * p.dv.batch_stride + (int64_t)(t.k_row0 + col) * p.dv.row_stride +
                (int64_t)t.ho * p.dv.head_stride;
    if (n_it > 0) {
      mbar_wait(bar(WB_DKV_FULL), 0);
      tc_fence_after();
    }
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      T *dst = which == 0 ? dv_row : dk_row;
      const float mul = which == 0 ? rp_drop : p.softmax_scale;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        const int c0 = wg * 64 + c * 32;             // column inside the slice
        if (n_it > 0) {
          tmem_ld32(tmem + lane_base + (which == 0 ? TDV : TDK) + c0, v);
          tmem_wait_ld();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0;
        }
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int dcol = t.slice * kSL + c0 + j * 8;
            if (dcol < p.d) {
              uint4 o;
              o.x = pack2<kBf16>(__uint_as_float(v[j * 8 + 0]) * mul, __uint_as_float(v[j * 8 + 1]) * mul);
              o.y = pack2<kBf16>(__uint_as_float(v[j * 8 + 2]) * mul, __uint_as_float(v[j * 8 + 3]) * mul);
              o.z = pack2<kBf16>(__uint_as_float(v[j * 8 + 4]) * mul, __uint_as_float(v[j * 8 + 5]) * mul);
              o.w = pack2<kBf16>(__uint_as_float(v[j * 8 + 6]) * mul, __uint_as_float(v[j * 8 + 7]) * mul);
              *reinterpret_cast<uint4 *>(dst + dcol) = o;
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 12) tmem_dealloc<512>(tmem);
}

void sanitize_w(TmaDesc4 &d) {
  for (int i = 1; i < 4; ++i)
    if (d.dims[i] == 1) d.strides[i - 1] = (i == 1 ? d.dims[0] * 2 : d.strides[i - 2] * d.dims[i - 1] + 15) / 16 * 16;
}
TmaDesc4 wdesc(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch, int box_rows) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * 2; x.strides[1] = (uint64_t)t.row_stride * 2;
  x.strides[2] = (uint64_t)t.batch_stride * 2;
  x.box[0] = 64; x.box[1] = 1; x.box[2] = (uint32_t)box_rows; x.box[3] = 1;
  sanitize_w(x);
  return x;
}

template <bool kBf16, bool kDrop> void launch_wide(const Params &p, cudaStream_t s) {
  using L = WSmem;
  const bool packed = p.cu_seqlens_q && !p.varlen_padded_input;
  const bool dev_rows = packed && (p.total_q < 0 || p.total_k < 0);
  const int64_t rows_q = packed ? (p.total_q >= 0 ? p.total_q : (int64_t)p.b * p.seqlen_q) : p.seqlen_q;
  const int64_t rows_k = packed ? (p.total_k >= 0 ? p.total_k : (int64_t)p.b * p.seqlen_k) : p.seqlen_k;
  WTma tm;   // member order q, k, v, dout
  // columns at or beyond p.d are zero-filled by the TMA unit: the contraction always runs over 256
  tm.q = make_tma_4d(wdesc(p.q, p.d, p.h, rows_q, packed ? 1 : p.b, kBQ));
  tm.dout = make_tma_4d(wdesc(p.dout, p.d, p.h, rows_q, packed ? 1 : p.b, kBQ));
  tm.k = make_tma_4d(wdesc(p.k, p.d, p.h_k, rows_k, packed ? 1 : p.b, kBK));
  tm.v = make_tma_4d(wdesc(p.v, p.d, p.h_k, rows_k, packed ? 1 : p.b, kBK));
  WTma *tm_dev = nullptr;
  if (dev_rows) {
    const int k_like[4] = {0, 1, 1, 0};
    tm_dev = reinterpret_cast<WTma *>(patch_tma_rows(&tm.q, k_like, 4, p.cu_seqlens_q, p.cu_seqlens_k, p.b, s));
  }
  bwd_pre_dpsum(p, s);
  auto kern = bwd_wide_kernel<kBf16, kDrop>;
  static std::atomic<bool> smem_set[64];
  ensure_dynamic_smem(kern, L::kTotal, smem_set);
  dim3 grid(2 * ceil_div(p.seqlen_k, kBK), p.dkv_expanded ? p.h : p.h_k, p.b);
  profile_mark(1, true, s);
  kern<<<grid, kWThreads, L::kTotal, s>>>(p, tm, tm_dev);
  profile_mark(1, false, s);
  const cudaError_t launch_err = cudaGetLastError();
  if (tm_dev) cudaFreeAsync(tm_dev, s);
  FA_CUDA(launch_err);
  note_launch(2);
  bwd_post_convert_dq(p, s);
}

}  // namespace

bool sm100_bwd_wide_eligible(const Params &p) {
  return p.d > 128 && p.d <= 256 && p.lt_start == nullptr && p.attn_mask == nullptr && p.softcap == 0.f &&
         !p.deterministic;
}

void sm100_bwd_wide(const Params &p, cudaStream_t s) {
  const bool drop = p.p_dropout > 0.f;
  if (p.is_bf16) { if (drop) launch_wide<true, true>(p, s); else launch_wide<true, false>(p, s); }
  else { if (drop) launch_wide<false, true>(p, s); else launch_wide<false, false>(p, s); }
}

}  // namespace fa
