// sm100_bwd.cu -- the tcgen05 / TMEM / TMA backward kernels (sm_100a).
//
// Same mathematics as the reference's compute_dq_dk_dv_1colblock (csrc/flash_attn/src/flash_bwd_kernel.h:427-1113):
//   S = Q K^T, P = exp2(S * scale_log2 - LSE * log2e), dP = dO V^T, dS = P o (dP - dPsum),
//   dV += P^T dO, dK += dS^T Q (x scale), dQ += dS K (x scale, fp32 accumulation across key blocks),
// preceded by dPsum = rowsum(dO o O) (:112-188) and followed by the fp32 -> 16-bit dQ conversion (:232-310).
//
// Blackwell structure.  One CTA owns one 128-key block of one (batch, kv-head) and loops over the query blocks of
// every q-head of the group, so dK/dV accumulate in TMEM with no atomics and GQA needs no expanded buffers.
// Everything is computed TRANSPOSED (keys on TMEM lanes) so that P^T and dS^T can feed the dV / dK MMAs straight
// from TMEM (an A operand in TMEM cannot be transposed):
//   warp 12   : MMA issuer   S^T = K Q^T, dP^T = V dO^T (SS), dV += P^T dO, dK += dS^T Q (TS), dQ = dS K (SS, dS^T in smem)
//   warp 13   : producer     TMA loads of K, V (once) and Q (2 stages), dO; LSE*log2e and dPsum rows into smem
//   warps 0-7 : compute      two warpgroups, each owns 64 of the 128 query columns of S^T / dP^T
//   warps 8-11: dQ reducers  TMEM -> registers -> smem -> cp.reduce.async.bulk (fp32 add) into the dq_accum workspace
// TMEM columns: S^T/P^T [0,128)  dV [128,128+HD)  dP^T/dS^T/dQ [256,384)  dK [384,384+HD).
//
// dq_accum layout (a private workspace): tiles of 128 rows; per tile HD/32 chunks; a chunk is 128 rows x 32 floats
// (16 KiB, the unit of one bulk reduce-add) whose 16-byte groups are XOR-swizzled with (row % 8).
#include <math_constants.h>
#include <type_traits>

#include "common.h"
#include "sm100.cuh"

#ifdef FA_TRACE
// Timeline probe (variant builds only, tools/trace_bwd.py): CTA (0,0,0) records clock64() at the hand-over points of its
// first iterations -- role 0/1 = compute warpgroup 0/1 (first warp), role 2 = MMA issuer, role 3 = dQ reducer warp 8.
__device__ unsigned long long fa_trace_bwd_buf[4][96][8];
#define FB_TR(role, it, ev) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (it) < 96) \
  fa_trace_bwd_buf[role][it][ev] = clock64(); } while (0)
extern "C" int fa_b200_trace_bwd_dump(unsigned long long *host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, fa_trace_bwd_buf, sizeof(fa_trace_bwd_buf));
}
#define FB_TR_DEP(role, it, ev, val) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (it) < 96) { \
  unsigned long long t__; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t__) : "r"(val)); fa_trace_bwd_buf[role][it][ev] = t__; } } while (0)
#else
#define FB_TR(role, it, ev) do { } while (0)
#define FB_TR_DEP(role, it, ev, val) do { } while (0)
#endif

namespace fa {
namespace {

using namespace sm100;

constexpr int kB = 128;            // query rows per block == keys per block
constexpr int kBwdThreads = 512;   // 16 warps (the last two idle; they only pad the register-budget warpgroup)
constexpr float kLog2e = 1.4426950408889634f;

struct BwdTma {
  CUtensorMap q, k, v, dout;
};

template <int HD> struct BwdSmem {
  static constexpr int kTile = kB * HD * 2;              // one 128 x HD 16-bit tile
  static constexpr int kK = 0;
  static constexpr int kV = kK + kTile;
  static constexpr int kQ = kV + kTile;                  // 2 stages
  static constexpr int kDO = kQ + 2 * kTile;
  static constexpr int kDS = kDO + kTile;                // dS^T: 2 halves x (128 keys x 128 bytes)
  static constexpr int kDQ = kDS + kB * kB * 2;          // dQ staging: 2 chunks of 128 rows x 32 floats
  static constexpr int kVec = kDQ + 2 * kB * 32 * 4;     // [2 stages][lse2 | dpsum][128] floats
  static constexpr int kBar = kVec + 2 * 2 * kB * 4;
  static constexpr int kNumBars = 16;
  static constexpr int kTmemSlot = kBar + kNumBars * 8;
  static constexpr int kTotal = kTmemSlot + 16;
  static constexpr int kDynBytes = kTotal;               // the dynamic smem base is 1024-byte aligned (checked at run time)
  static_assert(kTotal <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");
};

enum BwdBar {
  BB_KV_FULL = 0, BB_Q_FULL = 1 /*2*/, BB_Q_EMPTY = 3 /*2*/, BB_DO_FULL = 5, BB_DO_EMPTY = 6, BB_S_FULL = 7, BB_P_FULL = 8,
  BB_DP_FULL = 9, BB_DS_FULL = 10, BB_DQ_FULL = 11, BB_DQ_EMPTY = 12, BB_DKV_FULL = 13
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2f(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int N> __device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// What one CTA works on: key block n0 of (batch bi, kv-head hk); q-heads [hq0, hq1); per head query blocks [m_lo, m_hi).
struct BwdTile {
  int bi, hk, ho, n0;
  int hq0, hq1;
  int sq, sk, q_row0, k_row0, q_b, k_b;
  int m_lo, m_hi;
  int acc_tile0;     // first dq_accum tile of this batch element (before head offset)
  int acc_tiles_per_head;
};

__device__ __forceinline__ bool bwd_decode(const Params &p, BwdTile &t) {
  // grid: x = key block (descending work with ascending index for causal), y = output head, z = batch
  t.bi = blockIdx.z;
  t.ho = blockIdx.y;
  const int g = p.h / p.h_k;
  if (p.dkv_expanded) { t.hk = t.ho / g; t.hq0 = t.ho; t.hq1 = t.ho + 1; }
  else { t.hk = t.ho; t.hq0 = t.ho * g; t.hq1 = t.ho * g + g; }
  bool qb, kb;
  seq_bounds(p, t.bi, t.q_row0, t.sq, t.k_row0, t.sk, qb, kb);
  t.q_b = qb ? t.bi : 0;
  t.k_b = kb ? t.bi : 0;
  t.n0 = blockIdx.x * kB;
  if (t.n0 >= t.sk) return false;
  const int shift = t.sk - t.sq;
  const int shiftc = p.causal_bottom_right ? shift : 0;
  // rows that can see any key of [n0, n0 + 128): causal: row >= col - shiftc; window_right: row >= col - shift - wr;
  // window_left: row <= col - shift + wl
  int rlo = 0, rhi = t.sq;
  const int c_last = min(t.n0 + kB, t.sk) - 1;
  if (p.is_causal) rlo = max(rlo, t.n0 - shiftc);
  if (p.window_right >= 0) rlo = max(rlo, t.n0 - shift - p.window_right);
  if (p.window_left >= 0) rhi = min(rhi, c_last - shift + p.window_left + 1);
  t.m_lo = rlo / kB;
  t.m_hi = (max(rhi, 0) + kB - 1) / kB;
  if (t.m_hi < t.m_lo) t.m_hi = t.m_lo;
  if (p.cu_seqlens_q && !p.varlen_padded_input && !p.dq_accum_batch_major) {
    t.acc_tile0 = (p.cu_seqlens_q[t.bi] + kB * t.bi) / kB;
    t.acc_tiles_per_head = (int)((p.total_q + (int64_t)kB * p.b) / kB);
  } else {
    t.acc_tiles_per_head = p.seqlen_q_rounded / kB;
    t.acc_tile0 = t.bi * p.h * t.acc_tiles_per_head;  // batch-major: (b, h, tile)
  }
  return true;
}

// index of the dq_accum tile for (head hq, query block m)
__device__ __forceinline__ int64_t acc_tile_index(const Params &p, const BwdTile &t, int hq, int m) {
  if (p.cu_seqlens_q && !p.varlen_padded_input) return (int64_t)hq * t.acc_tiles_per_head + t.acc_tile0 + m;
  return (int64_t)t.acc_tile0 + (int64_t)hq * t.acc_tiles_per_head + m;
}

// Deterministic mode (fa_b200_params.deterministic; the flat C ABI's num_splits == 1, csrc/flash_attn/src/
// flash_bwd_kernel.h:3203-3208; the v3 handle's dq_semaphore, mainloop_bwd_sm90_tma_gmma_ws.hpp:751-890): the fp32 adds
// into a dq_accum tile happen in key-block order.  One int per tile counts the key blocks that have added so far; the
// CTA of key block j waits for the count to reach (j - first key block that contributes to this query block).
// CTAs are dispatched in blockIdx.x order, so the CTA being waited for is always resident or finished.
__device__ __forceinline__ bool key_block_touches(const Params &p, const BwdTile &t, int j, int m) {
  const int shift = t.sk - t.sq, shiftc = p.causal_bottom_right ? shift : 0;
  const int n0 = j * kB;
  if (n0 >= t.sk) return false;
  int rlo = 0, rhi = t.sq;
  const int c_last = min(n0 + kB, t.sk) - 1;
  if (p.is_causal) rlo = max(rlo, n0 - shiftc);
  if (p.window_right >= 0) rlo = max(rlo, n0 - shift - p.window_right);
  if (p.window_left >= 0) rhi = min(rhi, c_last - shift + p.window_left + 1);
  const int m_lo = rlo / kB, m_hi = (max(rhi, 0) + kB - 1) / kB;
  return m >= m_lo && m < m_hi;
}
__device__ __forceinline__ int first_key_block(const Params &p, const BwdTile &t, int m) {
  if (p.window_left < 0) return 0;           // without a left window every key block up to the causal bound contributes
  int j = blockIdx.x;
  while (j > 0 && key_block_touches(p, t, j - 1, m)) --j;
  return j;
}
__device__ __forceinline__ int ld_acquire(const int *ptr) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(ptr) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add1(int *ptr) {
  asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(ptr) : "memory");
}

// The (q-head, query block) pairs one CTA visits, in the order every role walks them.  Per q-head the query blocks
// are [m_lo, m_hi) minus the blocks that FlashMask covers completely for ALL keys of this CTA's key block: rows
// [lts_max, lte_min) and [uts_max, ute_min) (per-block min/max of the mask vectors, flashmask_scan_kernel) -- at most
// three surviving ranges, the same cut the reference makes (mainloop_bwd_sm90_tma_gmma_ws.hpp:1249-1299).
template <bool kFM> struct MCursor {
  int hq, hq1, ri, m, end, s0, e0, s1, e1, s2, e2;   // scalars: no dynamically indexed arrays (local memory)
  int v[8];   // lts_max, lts_min, ute_max, ute_min, lte_max, lte_min, uts_max, uts_min of (mask head, key block)
  bool has_ute;

  __device__ __forceinline__ void load_head(const Params &p, const BwdTile &t, int h) {
    s0 = t.m_lo; e0 = t.m_hi; s1 = e1 = s2 = e2 = 0;
    if constexpr (kFM) {
      const int nbp = (ceil_div(p.seqlen_k, kB) + 3) & ~3;
      const int64_t plane = (int64_t)p.b * p.h_mask * nbp;
      const int32_t *mm = p.flashmask_maxmin + ((int64_t)t.bi * p.h_mask + h / (p.h / p.h_mask)) * nbp + t.n0 / kB;
      has_ute = p.ut_end != nullptr;
      v[0] = mm[0]; v[1] = mm[plane];
      v[2] = has_ute ? mm[2 * plane] : 0; v[3] = has_ute ? mm[3 * plane] : 0;
      v[4] = p.lt_end ? mm[4 * plane] : 0x7fffffff; v[5] = p.lt_end ? mm[5 * plane] : 0x7fffffff;
      v[6] = p.ut_start ? mm[6 * plane] : 0; v[7] = p.ut_start ? mm[7 * plane] : 0;
      // fully masked block intervals [a0, a1) (lower-triangle band) and [b0, b1) (upper-triangle band), clipped
      int a0 = min(max((int)(((int64_t)v[0] + kB - 1) / kB), t.m_lo), t.m_hi);
      int a1 = v[5] >= t.sq ? t.m_hi : min(max(v[5] / kB, t.m_lo), t.m_hi);
      int b0 = t.m_hi, b1 = t.m_hi;
      if (has_ute) {
        b0 = min(max((v[6] + kB - 1) / kB, t.m_lo), t.m_hi);
        b1 = v[3] >= t.sq ? t.m_hi : min(max(v[3] / kB, t.m_lo), t.m_hi);
      }
      if (a1 < a0) a1 = a0;
      if (b1 < b0) b1 = b0;
      if (b0 < a0) { int x = a0; a0 = b0; b0 = x; x = a1; a1 = b1; b1 = x; }   // order by start
      s0 = t.m_lo; e0 = a0;
      s1 = a1; e1 = max(b0, a1);
      s2 = max(b1, a1); e2 = t.m_hi;
    }
  }
  __device__ __forceinline__ int count() const {
    return max(e0 - s0, 0) + max(e1 - s1, 0) + max(e2 - s2, 0);
  }
  __device__ __forceinline__ void init(const BwdTile &t) { hq = t.hq0 - 1; hq1 = t.hq1; ri = 2; m = 0; end = 0; }
  __device__ __forceinline__ bool next(const Params &p, const BwdTile &t) {
    ++m;
    while (true) {
      if (m < end) return true;
      ++ri;
      if (ri == 1) { m = s1; end = e1; }
      else if (ri == 2) { m = s2; end = e2; }
      else {
        if (++hq >= hq1) return false;
        load_head(p, t, hq);
        ri = 0; m = s0; end = e0;
      }
    }
  }
  // Position of iteration `it` (called with it = 0, 1, 2, ... in order).  Without FlashMask this is the closed form the
  // counted loops always used (keeps their code generation); with FlashMask the cursor walks the surviving ranges.
  __device__ __forceinline__ void step(const Params &p, const BwdTile &t, int it) {
    if constexpr (!kFM) {
      const int nm = t.m_hi - t.m_lo;
      hq = t.hq0 + it / nm; m = t.m_lo + it % nm;
    } else {
      next(p, t);
    }
  }
  // may FlashMask mask single elements of query block m (for this key block)?
  __device__ __forceinline__ bool partial(const BwdTile &t) const {
    if constexpr (!kFM) return false;
    const int r0 = m * kB, r1 = min(r0 + kB, t.sq);
    const bool clear = (r1 <= v[1] || r0 >= v[4]) && (!has_ute || r1 <= v[7] || r0 >= v[2]);
    return !clear;
  }
};

template <int HD, bool kBf16, bool kFM, bool kDet, bool kDrop, bool kMask, bool kCap>
__global__ void __launch_bounds__(kBwdThreads, 1)
bwd_kernel(const __grid_constant__ Params p, const __grid_constant__ BwdTma tm_param, const BwdTma *__restrict__ tm_dev,
           int *__restrict__ dq_sem) {
  using L = BwdSmem<HD>;
  const BwdTma &tm = tm_dev ? *tm_dev : tm_param;   // device-patched maps (flat varlen symbols) or the launch's own
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = smem_raw;
  const uint32_t sbase = smem_u32(smem);
  if ((sbase & 1023u) != 0) __trap();  // 128-byte swizzled tiles need a 1024-byte aligned base
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  auto bar = [&](int i) { return sbase + L::kBar + i * 8; };
  uint32_t *tmem_slot = (uint32_t *)(smem + L::kTmemSlot);
  float *vec = (float *)(smem + L::kVec);  // [stage][0: lse*log2e, 1: dpsum][128]

  BwdTile t;
  const bool live = bwd_decode(p, t);
  if (!live) return;  // uniform for the whole CTA
  int n_it = (t.hq1 - t.hq0) * (t.m_hi - t.m_lo);
  if constexpr (kFM) {
    MCursor<kFM> c;
    n_it = 0;
    for (int h = t.hq0; h < t.hq1; ++h) { c.load_head(p, t, h); n_it += c.count(); }
  }
  const int n_it_u = __shfl_sync(0xffffffffu, n_it, 0);  // same value, provably warp-uniform for the issuer warp

  if (threadIdx.x == 0) {
    mbar_init(bar(BB_KV_FULL), 1);
    for (int s = 0; s < 2; ++s) { mbar_init(bar(BB_Q_FULL + s), 1); mbar_init(bar(BB_Q_EMPTY + s), 1); }
    mbar_init(bar(BB_DO_FULL), 1); mbar_init(bar(BB_DO_EMPTY), 1);
    mbar_init(bar(BB_S_FULL), 1); mbar_init(bar(BB_P_FULL), 256);
    mbar_init(bar(BB_DP_FULL), 1); mbar_init(bar(BB_DS_FULL), 256);
    mbar_init(bar(BB_DQ_FULL), 1); mbar_init(bar(BB_DQ_EMPTY), 128);
    mbar_init(bar(BB_DKV_FULL), 1);
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

  constexpr int kHalves = HD / 64;
  constexpr uint32_t kHalfBytes = kB * 128;
  constexpr uint32_t TS_ = 0, TDV = 128, TDP = 256, TDK = 384;  // TMEM column map
  const float scale_log2 = p.softmax_scale * kLog2e;

  if (warp >= 12) {
    setmaxnreg_dec<56>();
    if (warp == 13) {
      // ======================= producer =======================
      if (lane == 0 && n_it > 0) {
        mbar_expect_tx(bar(BB_KV_FULL), 2 * L::kTile);
#pragma unroll
        for (int hf = 0; hf < kHalves; ++hf) {
          tma_load_4d(sbase + L::kK + hf * kHalfBytes, &tm.k, bar(BB_KV_FULL), hf * 64, t.hk, t.k_row0 + t.n0, t.k_b);
          tma_load_4d(sbase + L::kV + hf * kHalfBytes, &tm.v, bar(BB_KV_FULL), hf * 64, t.hk, t.k_row0 + t.n0, t.k_b);
        }
      }
      MCursor<kFM> cur;
      cur.init(t);
      for (int it = 0; it < n_it; ++it) {
        cur.step(p, t, it);
        const int st = it & 1;
        const int hq = cur.hq, m = cur.m;
        mbar_wait(bar(BB_Q_EMPTY + st), ((it >> 1) & 1) ^ 1);
        const float *lse = p.softmax_lse + ((int64_t)t.bi * p.h + hq) * p.seqlen_q;
        const float *dps = p.softmax_d + ((int64_t)t.bi * p.h + hq) * p.seqlen_q_rounded;
#pragma unroll
        for (int i = lane; i < kB; i += 32) {
          const int row = m * kB + i;
          float l = CUDART_INF_F, dsum = 0.f;
          if (row < t.sq) {
            l = lse[row];
            l = isinf(l) ? CUDART_INF_F : l * kLog2e;
            dsum = dps[row];
          }
          vec[(st * 2 + 0) * kB + i] = l;
          vec[(st * 2 + 1) * kB + i] = dsum;
        }
        __syncwarp();
        if (lane == 0) {
          mbar_expect_tx(bar(BB_Q_FULL + st), L::kTile);
#pragma unroll
          for (int hf = 0; hf < kHalves; ++hf)
            tma_load_4d(sbase + L::kQ + st * L::kTile + hf * kHalfBytes, &tm.q, bar(BB_Q_FULL + st), hf * 64, hq,
                        t.q_row0 + m * kB, t.q_b);
          mbar_wait(bar(BB_DO_EMPTY), (it & 1) ^ 1);
          mbar_expect_tx(bar(BB_DO_FULL), L::kTile);
#pragma unroll
          for (int hf = 0; hf < kHalves; ++hf)
            tma_load_4d(sbase + L::kDO + hf * kHalfBytes, &tm.dout, bar(BB_DO_FULL), hf * 64, hq, t.q_row0 + m * kB,
                        t.q_b);
        }
        __syncwarp();
      }
    } else if (warp == 12 && n_it_u > 0) {
      // ======================= MMA issuer =======================
      // Whole warp, warp-uniform control flow (descriptors stay in uniform registers); one elected lane issues.
      const uint32_t tm = (uint32_t)__shfl_sync(0xffffffffu, (int)tmem, 0);
      constexpr uint32_t idesc_st = make_idesc(kBf16, kB, kB, false, false);   // S^T, dP^T: K-major x K-major
      constexpr uint32_t idesc_kv = make_idesc(kBf16, kB, HD, false, true);    // dV, dK: A in TMEM, B MN-major
      constexpr uint32_t idesc_dq = make_idesc(kBf16, kB, HD, true, true);     // dQ: A (dS^T) MN-major, B (K) MN-major
      const uint32_t sK = sbase + L::kK, sV = sbase + L::kV, sDO = sbase + L::kDO, sDS = sbase + L::kDS;
      auto kmajor_pair = [&](uint32_t d_col, uint32_t a_base, uint32_t b_base, uint32_t done_bar) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t off = (k / 4) * kHalfBytes + (k % 4) * 32;
            umma_ss(tm + d_col, make_sdesc_sw128(a_base + off, 16, 1024), make_sdesc_sw128(b_base + off, 16, 1024),
                    idesc_st, k > 0);
          }
          umma_commit(done_bar);
        }
        __syncwarp();
      };
      auto tmem_a_mn_b = [&](uint32_t d_col, uint32_t a_col, uint32_t b_base, bool accumulate, uint32_t done_bar) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kB / 16; ++k) {
            const uint32_t a = tm + a_col + (k < 4 ? k * 8 : 64 + (k - 4) * 8);  // each compute warpgroup packs into its own half
            umma_ts(tm + d_col, a, make_sdesc_sw128(b_base + k * 2048, kHalfBytes, 1024), idesc_kv,
                    (accumulate || k > 0) ? 1u : 0u);
          }
          umma_commit(done_bar);
        }
        __syncwarp();
      };
      mbar_wait(bar(BB_KV_FULL), 0);
      if (p.seqused_k != nullptr && t.n0 + kB > t.sk) {
        // KV cache: rows at or beyond the used length hold arbitrary bits (possibly NaN).  Their P is masked to zero, but
        // 0 * NaN would still reach dS (through dP = dO.V) and dQ (through dS.K): zero the tail rows of both tiles once.
        // A row is 128 bytes in each 64-column box, whatever the swizzle (same as the forward's V tail).
        const int first = max(t.sk - t.n0, 0);
        for (int i = first * 8 + lane; i < kB * 8; i += 32)
#pragma unroll
          for (int hf = 0; hf < HD / 64; ++hf) {
            sts128(sK + hf * kHalfBytes + i * 16, 0u, 0u, 0u, 0u);
            sts128(sV + hf * kHalfBytes + i * 16, 0u, 0u, 0u, 0u);
          }
        fence_proxy_async_smem();
        __syncwarp();
      }
      mbar_wait(bar(BB_Q_FULL + 0), 0);
      tc_fence_after();
      kmajor_pair(TS_, sK, sbase + L::kQ, bar(BB_S_FULL));
      for (int it = 0; it < n_it_u; ++it) {
        const int st = it & 1;
        const uint32_t sQ = sbase + L::kQ + st * L::kTile;
        FB_TR(2, it, 0);
        mbar_wait(bar(BB_DO_FULL), it & 1);
        if (it > 0) mbar_wait(bar(BB_DQ_EMPTY), (it - 1) & 1);   // dQ(it-1) has left TMEM: the dP region is free
        FB_TR(2, it, 1);
        tc_fence_after();
        kmajor_pair(TDP, sV, sDO, bar(BB_DP_FULL));
        mbar_wait(bar(BB_P_FULL), it & 1);
        FB_TR(2, it, 2);
        tc_fence_after();
        tmem_a_mn_b(TDV, TS_, sDO, it > 0, bar(BB_DO_EMPTY));
        // S(it+1) goes here, BEFORE dK(it) / dQ(it): its P phase (the longest compute phase) then runs under those two MMAs.
        // Measured alternative (round 2): S(it+1) behind dQ(it), to fill the bubble while the reducers drain dQ(it) out of
        // the dP/dS/dQ region -- 16 % slower on C3 (9.73 -> 11.32 ms), 26 % at d64: P(it+1) arrives too late for dV(it+1).
        if (it + 1 < n_it_u) {
          mbar_wait(bar(BB_Q_FULL + (st ^ 1)), ((it + 1) >> 1) & 1);
          tc_fence_after();
          kmajor_pair(TS_, sK, sbase + L::kQ + (st ^ 1) * L::kTile, bar(BB_S_FULL));
        }
        FB_TR(2, it, 3);
        mbar_wait(bar(BB_DS_FULL), it & 1);
        FB_TR(2, it, 4);
        tc_fence_after();
        // dK reads dS^T (packed 16-bit) from the dP/dS/dQ TMEM region and dQ overwrites that region: dK first on the
        // in-order pipe.  Tried in round 2 (timeline probe: the cycle dS ready -> dK -> dQ -> reducers drain dQ -> dP of the
        // next block -> dS phase is the critical one): dQ first with dK reading dS^T from SHARED memory instead, so that dK
        // runs while the reducers drain.  Correct, but 7.5 % SLOWER on C3 (1180 -> 1092 TFLOPS): an SS MMA of this shape is
        // shared-memory-bandwidth bound (tools/tmem_bench.cu: 675 cycles per 128^3 against 544 for TS), and the block's
        // MMA stream -- S, dP, dQ in SS mode, dV, dK in TS mode: 3 x 675 + 2 x 544 = 3113 cycles -- is already 84 % of the
        // measured 3700-cycle period.  What would help is cta_group::2 (a CTA pair shares the B operand), not reordering.
        tmem_a_mn_b(TDK, TDP, sQ, it > 0, bar(BB_Q_EMPTY + st));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)
            umma_ss(tm + TDP, make_sdesc_sw128(sDS + k * 2048, kHalfBytes, 1024),
                    make_sdesc_sw128(sK + k * 2048, kHalfBytes, 1024), idesc_dq, k > 0);
          umma_commit(bar(BB_DQ_FULL));
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(bar(BB_DKV_FULL));
      __syncwarp();
    }
  } else if (warp >= 8) {
    // ======================= dQ reducers =======================
    // All of dQ (128 fp32 per row) is pulled out of TMEM first so that the dP/dS/dQ region is handed back to the MMA
    // warp as early as possible; the smem staging + bulk reduce-add then runs off the critical path.
    setmaxnreg_inc<kDrop ? 168 : 184>();   // the dropout instantiation moves 16 registers per thread to the compute warpgroups
    const int r = (warp % 4) * 32 + lane;
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const int tid = threadIdx.x - 256;
    const int n_chunks = p.d_rounded / 32 < HD / 32 ? p.d_rounded / 32 : HD / 32;
    uint8_t *stg = smem + L::kDQ;
    uint32_t g = 0;  // bulk groups committed so far; group g uses staging buffer g & 1, last used by group g - 2
    MCursor<kFM> cur;
    cur.init(t);
    for (int it = 0; it < n_it; ++it) {
      cur.step(p, t, it);
      const int hq = cur.hq, m = cur.m;
      const int64_t tile_idx = acc_tile_index(p, t, hq, m);
      float *gtile = p.dq_accum + tile_idx * (int64_t)(p.d_rounded / 32) * (kB * 32);
      if (warp == 8) FB_TR(3, it, 0);
      mbar_wait(bar(BB_DQ_FULL), it & 1);
      if (warp == 8) FB_TR(3, it, 1);
      tc_fence_after();
      uint32_t v[HD / 32][32];
#pragma unroll
      for (int c = 0; c < HD / 32; ++c)
        if (c < n_chunks) tmem_ld32(tmem + lane_base + TDP + c * 32, v[c]);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(bar(BB_DQ_EMPTY));
      if (warp == 8) FB_TR(3, it, 2);
      // Measured alternative (round 2): red.global.add.v4.f32 straight from the registers into the same tile, to spare the
      // shared-memory traffic of the staging (64 KB in + 64 KB out per query block on top of ~380 KB of MMA operand reads,
      // TMA writes and dS^T) -- 83 % SLOWER on C3: 16-byte reductions scattered over 32 rows per instruction are no match
      // for the bulk unit's 16 KB reduce-adds.
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        if (c < n_chunks) {
          uint8_t *buf = stg + (g & 1) * (kB * 32 * 4);
          if (tid == 0 && g >= 2) bulk_wait_read<1>();   // the group that last used this buffer has been read out
          named_bar_sync(3, 128);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4 *>(buf + r * 128 + ((j ^ (r & 7)) * 16)) =
                make_uint4(v[c][4 * j], v[c][4 * j + 1], v[c][4 * j + 2], v[c][4 * j + 3]);
          fence_proxy_async_smem();
          named_bar_sync(4, 128);
          if (tid == 0) {
            if (kDet && c == 0) {                // deterministic: wait for this tile's turn
              const int turn = (int)blockIdx.x - first_key_block(p, t, m);
              const long long t0 = clock64();
              while (ld_acquire(dq_sem + tile_idx) != turn) {
                __nanosleep(200);
                if (clock64() - t0 > 20000000000ll) __trap();
              }
            }
            bulk_reduce_add_f32(gtile + c * (kB * 32), smem_u32(buf), kB * 32 * 4);
            bulk_commit();
          }
          ++g;
        }
      }
      if (kDet && tid == 0) {                    // all of this key block's adds have landed: pass the tile on
        bulk_wait<0>();
        red_release_add1(dq_sem + tile_idx);
      }
    }
    if (tid == 0) bulk_wait<0>();
  } else {
    // ======================= compute warpgroups =======================
    setmaxnreg_inc<kDrop ? 144 : 136>();
    const int wg = warp / 4;
    const int r = (warp % 4) * 32 + lane;          // key row inside the block == TMEM lane
    const int col = t.n0 + r;                      // key index
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const uint32_t tS = tmem + lane_base + TS_ + wg * 64, tP = tmem + lane_base + TS_ + wg * 64;
    const uint32_t tDP = tmem + lane_base + TDP + wg * 64, tDS = tmem + lane_base + TDP + wg * 64;
    const uint32_t ds_row_a = sbase + L::kDS + wg * kHalfBytes + r * 128;
    const int shift = t.sk - t.sq;
    const int shiftc = p.causal_bottom_right ? shift : 0;
    const float rp_drop = kDrop ? 1.f / (1.f - p.p_dropout) : 1.f;
    const float pre_cap = kCap ? p.softmax_scale / p.softcap : 0.f, cap_log2 = kCap ? p.softcap * kLog2e : 0.f;
    const float inv_cap_log2 = kCap ? 1.f / (p.softcap * kLog2e) : 0.f;
    MCursor<kFM> cur;
    cur.init(t);
    for (int it = 0; it < n_it; ++it) {
      cur.step(p, t, it);
      const int st = it & 1;
      const int hq = cur.hq, m = cur.m;
      const int m0 = m * kB;
      const bool fm_partial = cur.partial(t);
      const MaskCtx mc = make_mask_ctx(p, t.bi, hq, t.sq, t.sk);
      const bool full = (t.n0 + kB <= t.sk) && (!mc.causal || t.n0 + kB - 1 <= m0 + shiftc) &&
                        (mc.wr < 0 || t.n0 + kB - 1 <= m0 + shift + mc.wr) &&
                        (mc.wl < 0 || t.n0 >= m0 + kB - 1 + shift - mc.wl) &&
                        (kFM ? !fm_partial : (mc.lts == nullptr));
      const uint32_t lse2_a = sbase + L::kVec + ((st * 2 + 0) * kB + wg * 64) * 4;
      const uint32_t dpsum_a = sbase + L::kVec + ((st * 2 + 1) * kB + wg * 64) * 4;
      // element mask of this thread's 64 (query, key) pairs: built in a rolled loop so that the generic mask rule is
      // emitted once (boundary / flashmask blocks only); the hot loops below stay small enough for the instruction cache
      uint32_t mbits[2] = {0u, 0u};
      if (!full) {
        if (mc.lts == nullptr || kFM) {  // seqlen tail / causal / window: the visible rows of a key column form an interval
          int vlo, vhi;
          visible_rows(mc, col, vlo, vhi);
          mbits[0] = outside_bits(vlo, vhi, m0 + wg * 64);
          mbits[1] = outside_bits(vlo, vhi, m0 + wg * 64 + 32);
          if (kFM && fm_partial && col < t.sk) {
            // FlashMask element rule for this thread's key column: the masked rows are the two intervals
            // [lts, lte) and [uts, ute), so the bit masks follow arithmetically
            const int lts_ = mc.lts[col], lte_ = mc.lte ? mc.lte[col] : 0x7fffffff;
            mbits[0] |= ~outside_bits(lts_, lte_, m0 + wg * 64);
            mbits[1] |= ~outside_bits(lts_, lte_, m0 + wg * 64 + 32);
            if (mc.ute) {
              const int uts_ = mc.uts ? mc.uts[col] : 0, ute_ = mc.ute[col];
              mbits[0] |= ~outside_bits(uts_, ute_, m0 + wg * 64);
              mbits[1] |= ~outside_bits(uts_, ute_, m0 + wg * 64 + 32);
            }
          }
        } else {
#pragma unroll 1
          for (int i = 0; i < 64; ++i)
            if (is_masked(mc, m0 + wg * 64 + i, col)) mbits[i >> 5] |= 1u << (i & 31);
        }
      }
      // Dropout (kDrop): keep bits of this thread's 64 (query, key) pairs.  The mask is dropout_keep() of common.h: one
      // Philox call covers 16 consecutive KEY columns of one query row, i.e. the 16 lanes of a half warp want one byte
      // each of the same call.  Each lane therefore runs the generator for one row of a 16-row group, turns its 16 bytes
      // into a 16-bit keep mask (bit j = column j of its row), and the half warp TRANSPOSES the 16 x 16 bit matrix with
      // four butterfly exchanges (rows on lanes -> columns on lanes): lane j ends up with the 16 rows of its own column.
      // (Round 1 exchanged the bytes themselves in 16 shuffle steps: 2.4 x the instructions.)  Done before the wait for S,
      // under the MMA that is in flight.
      uint32_t kb[2] = {0xffffffffu, 0xffffffffu};
      if constexpr (kDrop) {
        kb[0] = kb[1] = 0u;
        const int l16 = lane & 15;
        const uint32_t thr4 = (uint32_t)floorf((1.f - p.p_dropout) * 255.f) * 0x01010101u;
        const uint64_t ncg = (uint64_t)((p.seqlen_k + 15) / 16);
        const uint64_t cg = (uint64_t)((t.n0 + (warp % 4) * 32 + (lane & 16)) >> 4);
        const uint64_t head_base = ((uint64_t)t.bi * p.h + hq) * (uint64_t)p.seqlen_q;
        auto keep4 = [&](uint32_t w) {   // 4 bytes -> 4 bits: bit b = (byte b <= threshold)
          return ((__vcmpleu4(w, thr4) & 0x08040201u) * 0x01010101u) >> 24;
        };
#pragma unroll
        for (int ig = 0; ig < 4; ++ig) {
          const uint64_t rrow = (uint64_t)(m0 + wg * 64 + ig * 16 + l16);
          const uint4 rnd = philox4x32_7(p.philox_seed, (head_base + rrow) * ncg + cg, p.philox_offset);
          uint32_t x = keep4(rnd.x) | (keep4(rnd.y) << 4) | (keep4(rnd.z) << 8) | (keep4(rnd.w) << 12);
#pragma unroll
          for (int s_ = 8; s_ >= 1; s_ >>= 1) {
            const uint32_t lo = s_ == 8 ? 0x00ffu : s_ == 4 ? 0x0f0fu : s_ == 2 ? 0x3333u : 0x5555u;
            const uint32_t y = __shfl_xor_sync(0xffffffffu, x, s_);
            x = (l16 & s_) ? (((y >> s_) & lo) | (x & (lo << s_))) : ((x & lo) | ((y & lo) << s_));
          }
          kb[ig >> 1] |= x << ((ig & 1) * 16);   // bit (row in the 64-row half) of this thread's key column
        }
      }
      if (warp % 4 == 0) FB_TR(wg, it, 0);
      mbar_wait(bar(BB_Q_FULL + st), (it >> 1) & 1);   // LSE / dPsum rows of this query block are in smem
      mbar_wait(bar(BB_S_FULL), it & 1);
      if (warp % 4 == 0) FB_TR(wg, it, 1);
      tc_fence_after();
      float pr[64];
#pragma unroll
      for (int c2 = 0; c2 < 2; ++c2) {
        uint32_t s[32];
        tmem_ld32(tS + c2 * 32, s);
        tmem_wait_ld();
        if constexpr (kMask) {
          // dense additive mask: this thread's key column, 32 query rows (lanes are consecutive columns, so every load
          // of the warp is one contiguous run of 32 elements)
          const T *mcol = (const T *)p.attn_mask +
                          ((int64_t)t.bi * p.mask_head_mod + hq % p.mask_head_mod) * p.mask_seq_mod * (int64_t)p.seqlen_k + col;
          const bool col_ok = col < p.seqlen_k;
          // mask_seq_mod is 1 (one row for all queries) or the query length: no per-element integer division
          const int64_t rstride = p.mask_seq_mod == 1 ? 0 : (int64_t)p.seqlen_k;
          const int row_first = m0 + wg * 64 + c2 * 32;
          uint16_t raw[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            int rowi = row_first + i;
            if (rowi >= p.mask_seq_mod) rowi = p.mask_seq_mod == 1 ? 0 : rowi % p.mask_seq_mod;
            raw[i] = col_ok ? *reinterpret_cast<const uint16_t *>(mcol + (int64_t)rowi * rstride) : (uint16_t)0;
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            T mv;
            *reinterpret_cast<uint16_t *>(&mv) = raw[i];
            s[i] = __float_as_uint(fmaf(to_float(mv), p.unscale_softmax, __uint_as_float(s[i])));
          }
        }
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 l4 = lds128(lse2_a + (c2 * 32 + i) * 4);
          if constexpr (kCap) {   // P = exp2(softcap * log2e * tanh(scale * s / softcap) - LSE * log2e)
            pr[c2 * 32 + i + 0] = ex2f(fmaf(tanh_fast(__uint_as_float(s[i + 0]) * pre_cap), cap_log2, -l4.x));
            pr[c2 * 32 + i + 1] = ex2f(fmaf(tanh_fast(__uint_as_float(s[i + 1]) * pre_cap), cap_log2, -l4.y));
            pr[c2 * 32 + i + 2] = ex2f(fmaf(tanh_fast(__uint_as_float(s[i + 2]) * pre_cap), cap_log2, -l4.z));
            pr[c2 * 32 + i + 3] = ex2f(fmaf(tanh_fast(__uint_as_float(s[i + 3]) * pre_cap), cap_log2, -l4.w));
          } else {
            pr[c2 * 32 + i + 0] = ex2f(fmaf(__uint_as_float(s[i + 0]), scale_log2, -l4.x));
            pr[c2 * 32 + i + 1] = ex2f(fmaf(__uint_as_float(s[i + 1]), scale_log2, -l4.y));
            pr[c2 * 32 + i + 2] = ex2f(fmaf(__uint_as_float(s[i + 2]), scale_log2, -l4.z));
            pr[c2 * 32 + i + 3] = ex2f(fmaf(__uint_as_float(s[i + 3]), scale_log2, -l4.w));
          }
        }
        if (!full) {
          const uint32_t mb = mbits[c2];
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if ((mb >> i) & 1u) pr[c2 * 32 + i] = 0.f;
        }
        uint32_t pk[16];
        if constexpr (kDrop) {   // dV sees the dropped weights (1 / p_keep is applied to dV in the epilogue)
#pragma unroll
          for (int i = 0; i < 16; ++i)
            pk[i] = pack2<kBf16>(((kb[c2] >> (2 * i)) & 1u) ? pr[c2 * 32 + 2 * i] : 0.f,
                                 ((kb[c2] >> (2 * i + 1)) & 1u) ? pr[c2 * 32 + 2 * i + 1] : 0.f);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = pack2<kBf16>(pr[c2 * 32 + 2 * i], pr[c2 * 32 + 2 * i + 1]);
        }
        if (warp % 4 == 0 && c2 == 1) FB_TR_DEP(wg, it, 5, pk[15]);
        tmem_st16(tP + c2 * 16, pk);
      }
      tmem_wait_st();
      tc_fence_before();
      mbar_arrive(bar(BB_P_FULL));
      if (warp % 4 == 0) FB_TR(wg, it, 2);

      mbar_wait(bar(BB_DP_FULL), it & 1);
      if (it > 0) mbar_wait(bar(BB_DQ_FULL), (it - 1) & 1);  // dQ(it-1) no longer reads the dS^T smem tile
      if (warp % 4 == 0) FB_TR(wg, it, 3);
      tc_fence_after();
#pragma unroll
      for (int c2 = 0; c2 < 2; ++c2) {
        uint32_t dp[32];
        tmem_ld32(tDP + c2 * 32, dp);
        tmem_wait_ld();
        if (warp % 4 == 0 && c2 == 1) FB_TR_DEP(wg, it, 6, dp[31]);
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 d4 = lds128(dpsum_a + (c2 * 32 + i) * 4);
          float g0 = __uint_as_float(dp[i + 0]), g1 = __uint_as_float(dp[i + 1]), g2 = __uint_as_float(dp[i + 2]),
                g3 = __uint_as_float(dp[i + 3]);
          if constexpr (kDrop) {   // dP = keep / p_keep * (dO . V)
            g0 = ((kb[c2] >> (i + 0)) & 1u) ? g0 * rp_drop : 0.f;
            g1 = ((kb[c2] >> (i + 1)) & 1u) ? g1 * rp_drop : 0.f;
            g2 = ((kb[c2] >> (i + 2)) & 1u) ? g2 * rp_drop : 0.f;
            g3 = ((kb[c2] >> (i + 3)) & 1u) ? g3 * rp_drop : 0.f;
          }
          float e0 = pr[c2 * 32 + i + 0] * (g0 - d4.x);
          float e1 = pr[c2 * 32 + i + 1] * (g1 - d4.y);
          float e2 = pr[c2 * 32 + i + 2] * (g2 - d4.z);
          float e3 = pr[c2 * 32 + i + 3] * (g3 - d4.w);
          if constexpr (kCap) {
            // d logit / d (scale * s) = 1 - tanh^2; tanh is recovered from P itself, tanh = (log2 P + LSE * log2e) /
            // (softcap * log2e), clamped to [-1, 1] (P = 0 gives -inf), instead of holding 64 more registers
            const float4 l4 = lds128(lse2_a + (c2 * 32 + i) * 4);
            const float t0 = fminf(fmaxf((lg2f(pr[c2 * 32 + i + 0]) + l4.x) * inv_cap_log2, -1.f), 1.f);
            const float t1 = fminf(fmaxf((lg2f(pr[c2 * 32 + i + 1]) + l4.y) * inv_cap_log2, -1.f), 1.f);
            const float t2 = fminf(fmaxf((lg2f(pr[c2 * 32 + i + 2]) + l4.z) * inv_cap_log2, -1.f), 1.f);
            const float t3 = fminf(fmaxf((lg2f(pr[c2 * 32 + i + 3]) + l4.w) * inv_cap_log2, -1.f), 1.f);
            e0 *= fmaf(-t0, t0, 1.f); e1 *= fmaf(-t1, t1, 1.f); e2 *= fmaf(-t2, t2, 1.f); e3 *= fmaf(-t3, t3, 1.f);
          }
          pk[i / 2] = pack2<kBf16>(e0, e1);
          pk[i / 2 + 1] = pack2<kBf16>(e2, e3);
        }
        tmem_st16(tDS + c2 * 16, pk);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          sts128(ds_row_a + (((c2 * 4 + j) ^ (r & 7)) * 16), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
      }
      tmem_wait_st();
      if (warp % 4 == 0) FB_TR(wg, it, 7);
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar(BB_DS_FULL));
      if (warp % 4 == 0) FB_TR(wg, it, 4);
    }
    // ---------------- epilogue: dK (x scale), dV ----------------
    const bool row_ok = col < t.sk;
    T *dk_row = (T *)p.dk.ptr + (int64_t)t.k_b * p.dk.batch_stride + (int64_t)(t.k_row0 + col) * p.dk.row_stride +
                (int64_t)t.ho * p.dk.head_stride;
    T *dv_row = (T *)p.dv.ptr + (int64_t)t.k_b * p.dv.batch_stride + (int64_t)(t.k_row0 + col) * p.dv.row_stride +
                (int64_t)t.ho * p.dv.head_stride;
    if (n_it > 0) {
      mbar_wait(bar(BB_DKV_FULL), 0);
      tc_fence_after();
    }
    constexpr int kColsPerWg = HD / 2;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      T *dst = which == 0 ? dv_row : dk_row;
      const float mul = which == 0 ? rp_drop : p.softmax_scale;
#pragma unroll
      for (int c = 0; c < kColsPerWg / 32; ++c) {
        uint32_t v[32];
        const int c0 = wg * kColsPerWg + c * 32;
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
            if (c0 + j * 8 < p.d) {
              uint4 o;
              o.x = pack2<kBf16>(__uint_as_float(v[j * 8 + 0]) * mul, __uint_as_float(v[j * 8 + 1]) * mul);
              o.y = pack2<kBf16>(__uint_as_float(v[j * 8 + 2]) * mul, __uint_as_float(v[j * 8 + 3]) * mul);
              o.z = pack2<kBf16>(__uint_as_float(v[j * 8 + 4]) * mul, __uint_as_float(v[j * 8 + 5]) * mul);
              o.w = pack2<kBf16>(__uint_as_float(v[j * 8 + 6]) * mul, __uint_as_float(v[j * 8 + 7]) * mul);
              *reinterpret_cast<uint4 *>(dst + c0 + j * 8) = o;
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

// ------------------------------------------------------------------------------------------------
// pre: dPsum[b, h, row] = sum_d dO * O  (fp32, (b, h, seqlen_q_rounded), 0 for padded rows).  One warp per row.
// (csrc/flash_attn/src/flash_bwd_kernel.h:112-188)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) bwd_dpsum_kernel(const Params p, int64_t acc_vec4) {
  // (a) zero the fp32 dQ accumulator (grid-stride, 16-byte stores): fused here so that the write stream overlaps the
  //     O / dO read stream of (b) instead of costing a separate memset pass
  {
    const int64_t nthreads = (int64_t)gridDim.x * gridDim.y * gridDim.z * 256;
    const int64_t tid = (((int64_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 256 + threadIdx.x;
    float4 *acc = reinterpret_cast<float4 *>(p.dq_accum);
#ifndef FA_AB_MEMSET
    for (int64_t i = tid; i < acc_vec4; i += nthreads) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#endif
  }
  // (b) dPsum: 16 lanes per row (8 elements each cover d <= 128), two rows per warp, 16 rows per block
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int sub = lane / 16, l16 = lane % 16;
  const int hq = blockIdx.y, bi = blockIdx.z;
  const int row = blockIdx.x * 16 + warp * 2 + sub;
  int q_row0, sq, k_row0, sk;
  bool qb, kb;
  seq_bounds(p, bi, q_row0, sq, k_row0, sk, qb, kb);
  float acc = 0.f;
  if (row < sq) {
    const T *o = (const T *)p.o.ptr + (qb ? bi : 0) * p.o.batch_stride + (int64_t)(q_row0 + row) * p.o.row_stride +
                 (int64_t)hq * p.o.head_stride;
    const T *g = (const T *)p.dout.ptr + (qb ? bi : 0) * p.dout.batch_stride +
                 (int64_t)(q_row0 + row) * p.dout.row_stride + (int64_t)hq * p.dout.head_stride;
    for (int e = l16 * 8; e < p.d; e += 128) {
      const uint4 a = *reinterpret_cast<const uint4 *>(o + e);
      const uint4 b = *reinterpret_cast<const uint4 *>(g + e);
      const T *ae = reinterpret_cast<const T *>(&a);
      const T *be = reinterpret_cast<const T *>(&b);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc = fmaf((float)ae[i], (float)be[i], acc);
    }
  }
#pragma unroll
  for (int o2 = 8; o2 > 0; o2 >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o2);
  if (l16 == 0 && row < p.seqlen_q_rounded) p.softmax_d[((int64_t)bi * p.h + hq) * p.seqlen_q_rounded + row] = acc;
}

// ------------------------------------------------------------------------------------------------
// post: dQ = T(scale * dq_accum), un-tiling the accumulator layout.  One thread per 8 output elements.
// (csrc/flash_attn/src/flash_bwd_kernel.h:232-310)
// ------------------------------------------------------------------------------------------------
template <typename T, bool kBf16>
__global__ void __launch_bounds__(256) bwd_convert_dq_kernel(const Params p) {
  const int hq = blockIdx.y, bi = blockIdx.z;
  int q_row0, sq, k_row0, sk;
  bool qb, kb;
  seq_bounds(p, bi, q_row0, sq, k_row0, sk, qb, kb);
  const int d8 = p.d / 8;
  const int idx = blockIdx.x * 256 + threadIdx.x;
  const int row = idx / d8, c8 = idx % d8;
  if (row >= sq) return;
  int64_t tile;
  if (p.cu_seqlens_q && !p.varlen_padded_input && !p.dq_accum_batch_major) {
    const int tiles_per_head = (int)((p.total_q + (int64_t)kB * p.b) / kB);
    tile = (int64_t)hq * tiles_per_head + (p.cu_seqlens_q[bi] + kB * bi) / kB + row / kB;
  } else {
    const int tiles_per_head = p.seqlen_q_rounded / kB;
    tile = ((int64_t)bi * p.h + hq) * tiles_per_head + row / kB;
  }
  const int r = row % kB, col = c8 * 8;
  const float *chunk = p.dq_accum + (tile * (p.d_rounded / 32) + col / 32) * (int64_t)(kB * 32) + r * 32;
  const int g0 = (col % 32) / 4;  // 16-byte group inside the 32-float row; two consecutive groups
  const float4 a = *reinterpret_cast<const float4 *>(chunk + ((g0 ^ (r & 7)) * 4));
  const float4 b = *reinterpret_cast<const float4 *>(chunk + (((g0 + 1) ^ (r & 7)) * 4));
  const float s = p.softmax_scale;
  uint4 o;
  o.x = pack2<kBf16>(a.x * s, a.y * s); o.y = pack2<kBf16>(a.z * s, a.w * s);
  o.z = pack2<kBf16>(b.x * s, b.y * s); o.w = pack2<kBf16>(b.z * s, b.w * s);
  T *dq = (T *)p.dq.ptr + (qb ? bi : 0) * p.dq.batch_stride + (int64_t)(q_row0 + row) * p.dq.row_stride +
          (int64_t)hq * p.dq.head_stride;
  *reinterpret_cast<uint4 *>(dq + col) = o;
}

void sanitize_b(TmaDesc4 &d) {
  // a dimension of extent 1 never advances: give it a legal, monotonic stride
  for (int i = 1; i < 4; ++i)
    if (d.dims[i] == 1) d.strides[i - 1] = (i == 1 ? d.dims[0] * 2 : d.strides[i - 2] * d.dims[i - 1] + 15) / 16 * 16;
}
TmaDesc4 bdesc(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * 2; x.strides[1] = (uint64_t)t.row_stride * 2;
  x.strides[2] = (uint64_t)t.batch_stride * 2;
  x.box[0] = 64; x.box[1] = 1; x.box[2] = kB; x.box[3] = 1;
  sanitize_b(x);
  return x;
}
bool tma_ok_b(const fa_b200_tensor &t, int heads, int batch) {
  if (((uintptr_t)t.ptr & 15) != 0) return false;
  if (t.row_stride % 8 != 0 || t.row_stride <= 0) return false;
  if (heads > 1 && (t.head_stride % 8 != 0 || t.head_stride <= 0)) return false;
  if (batch > 1 && (t.batch_stride % 8 != 0 || t.batch_stride <= 0)) return false;
  return true;
}
bool out_ok(const fa_b200_tensor &t) {
  return ((uintptr_t)t.ptr & 15) == 0 && t.row_stride % 8 == 0 && t.head_stride % 8 == 0 && t.batch_stride % 8 == 0;
}

template <int HD, bool kBf16, bool kFM, bool kDet = false, bool kDrop = false, bool kMask = false, bool kCap = false> void launch_bwd(const Params &p, cudaStream_t s) {
  using L = BwdSmem<HD>;
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  const bool packed = p.cu_seqlens_q && !p.varlen_padded_input;
  // packed tensors without host-known totals (the flat varlen symbols): the row extents are patched on the device
  const bool dev_rows = packed && (p.total_q < 0 || p.total_k < 0);
  const int64_t rows_q = packed ? (p.total_q >= 0 ? p.total_q : (int64_t)p.b * p.seqlen_q) : p.seqlen_q;
  const int64_t rows_k = packed ? (p.total_k >= 0 ? p.total_k : (int64_t)p.b * p.seqlen_k) : p.seqlen_k;
  BwdTma tm;   // member order q, k, v, dout
  tm.q = make_tma_4d(bdesc(p.q, p.d, p.h, rows_q, packed ? 1 : p.b));
  tm.dout = make_tma_4d(bdesc(p.dout, p.d, p.h, rows_q, packed ? 1 : p.b));
  tm.k = make_tma_4d(bdesc(p.k, p.d, p.h_k, rows_k, packed ? 1 : p.b));
  tm.v = make_tma_4d(bdesc(p.v, p.d, p.h_k, rows_k, packed ? 1 : p.b));
  BwdTma *tm_dev = nullptr;
  if (dev_rows) {
    const int k_like[4] = {0, 1, 1, 0};
    tm_dev = reinterpret_cast<BwdTma *>(patch_tma_rows(&tm.q, k_like, 4, p.cu_seqlens_q, p.cu_seqlens_k, p.b, s));
  }
  // 1. dPsum + zeroing of the fp32 dQ accumulator (one kernel)
  const int64_t acc_tiles = bwd_pre_dpsum(p, s);
  // 3. main kernel
  {
    if (kFM) flashmask_prepare(p, kB, s);   // per-128-key-block min/max of the mask vectors
    auto kern = bwd_kernel<HD, kBf16, kFM, kDet, kDrop, kMask, kCap>;
    static std::atomic<bool> smem_set[64];
    ensure_dynamic_smem(kern, L::kDynBytes, smem_set);
    dim3 grid(ceil_div(p.seqlen_k, kB), p.dkv_expanded ? p.h : p.h_k, p.b);
    int *dq_sem = nullptr;
    if (kDet) {                                 // one turn counter per dq_accum tile, stream-ordered scratch
      FA_CUDA(cudaMallocAsync(&dq_sem, sizeof(int) * (size_t)acc_tiles, s));
      FA_CUDA(cudaMemsetAsync(dq_sem, 0, sizeof(int) * (size_t)acc_tiles, s));
    }
    profile_mark(1, true, s);
    kern<<<grid, kBwdThreads, L::kDynBytes, s>>>(p, tm, tm_dev, dq_sem);
    profile_mark(1, false, s);
    const cudaError_t launch_err = cudaGetLastError();
    if (tm_dev) cudaFreeAsync(tm_dev, s);
    if (dq_sem) FA_CUDA(cudaFreeAsync(dq_sem, s));
    FA_CUDA(launch_err);
    note_launch(2);
  }
  // 4. dQ conversion
  bwd_post_convert_dq(p, s);
}

template <typename T> void pre_dpsum_t(const Params &p, int64_t acc_vec4, cudaStream_t s) {
  dim3 grid(ceil_div(p.seqlen_q_rounded, 16), p.h, p.b);
  bwd_dpsum_kernel<T><<<grid, 256, 0, s>>>(p, acc_vec4);
}
template <typename T, bool kBf16> void post_convert_t(const Params &p, cudaStream_t s) {
  dim3 grid(ceil_div(p.seqlen_q * (p.d / 8), 256), p.h, p.b);
  bwd_convert_dq_kernel<T, kBf16><<<grid, 256, 0, s>>>(p);
}

}  // namespace

// dPsum = rowsum(dO o O) and the zeroing of the tiled fp32 dQ accumulator, one launch; returns the number of 128-row
// accumulator tiles.  Shared by the d <= 128 kernel above and the sliced kernel of sm100_bwd_wide.cu.
int64_t bwd_pre_dpsum(const Params &p, cudaStream_t s) {
  const bool packed = p.cu_seqlens_q && !p.varlen_padded_input;
  int64_t tiles;
  if (packed && !p.dq_accum_batch_major) tiles = (int64_t)p.h * ((p.total_q + (int64_t)kB * p.b) / kB);
  else tiles = (int64_t)p.b * p.h * (p.seqlen_q_rounded / kB);
  const int64_t acc_vec4 = tiles * (p.d_rounded / 32) * (kB * 32) / 4;
#ifdef FA_AB_MEMSET
  FA_CUDA(cudaMemsetAsync(p.dq_accum, 0, (size_t)acc_vec4 * 16, s));
#endif
  if (p.is_bf16) pre_dpsum_t<__nv_bfloat16>(p, acc_vec4, s); else pre_dpsum_t<__half>(p, acc_vec4, s);
  FA_CUDA(cudaGetLastError());
  note_launch(2);
  return tiles;
}

// dQ = T(scale * dq_accum), un-tiling the accumulator
void bwd_post_convert_dq(const Params &p, cudaStream_t s) {
  if (p.is_bf16) post_convert_t<__nv_bfloat16, true>(p, s); else post_convert_t<__half, false>(p, s);
  FA_CUDA(cudaGetLastError());
  note_launch(2);
}

bool sm100_bwd_supported(const Params &p, std::string *why) {
  auto no = [&](const char *w) { if (why) *why = w; return false; };
  if (p.d > 128 && !sm100_bwd_wide_eligible(p))
    return no("head_size > 128 together with FlashMask / attn_mask / softcap / deterministic");
  if (!p.dq_accum) return no("no dq_accum workspace");
  if (p.d_rounded != round_up(p.d, 32) || p.seqlen_q_rounded % kB != 0 || p.seqlen_q_rounded < p.seqlen_q)
    return no("rounded sizes are not the documented ones");
  if (p.deterministic && p.lt_start) return no("deterministic + FlashMask (key blocks skip query blocks irregularly)");
  if (p.deterministic && p.p_dropout > 0.f) return no("deterministic + dropout");
  if (p.attn_mask && (p.lt_start || p.p_dropout > 0.f || p.deterministic)) return no("dense attn_mask with FlashMask / dropout / deterministic");
  if (p.softcap > 0.f && (p.lt_start || p.p_dropout > 0.f || p.deterministic || p.attn_mask)) return no("softcap with FlashMask / dropout / deterministic / attn_mask");
  const bool packed = p.cu_seqlens_q && !p.varlen_padded_input;
  if ((p.cu_seqlens_q != nullptr) != (p.cu_seqlens_k != nullptr)) return no("need both cu_seqlens");
  if (packed && (p.total_q < 0 || p.total_k < 0) && !p.dq_accum_batch_major)
    return no("varlen totals unknown and dq_accum not batch-major");   // the tiled accumulator is sized by total_q
  const int nb = packed ? 1 : p.b;
  if (!tma_ok_b(p.q, p.h, nb) || !tma_ok_b(p.dout, p.h, nb) || !tma_ok_b(p.k, p.h_k, nb) || !tma_ok_b(p.v, p.h_k, nb))
    return no("strides/alignment not TMA-compatible");
  if (!out_ok(p.dq) || !out_ok(p.dk) || !out_ok(p.dv) || !out_ok(p.o)) return no("outputs not 16-byte aligned");
  if (((uintptr_t)p.dq_accum & 127) != 0) return no("dq_accum not 128-byte aligned");
  return true;
}

void sm100_bwd(const Params &p, cudaStream_t s) {
  if (p.d > 128) { sm100_bwd_wide(p, s); return; }   // sliced kernel (sm100_bwd_wide.cu)
  const bool fm = p.lt_start != nullptr && p.flashmask_maxmin != nullptr;   // block skipping needs the workspace
  if (p.deterministic && !fm) {   // ordered dQ accumulation: its own instantiation, the default kernels carry none of it
    if (p.d <= 64) { if (p.is_bf16) launch_bwd<64, true, false, true>(p, s); else launch_bwd<64, false, false, true>(p, s); }
    else { if (p.is_bf16) launch_bwd<128, true, false, true>(p, s); else launch_bwd<128, false, false, true>(p, s); }
    return;
  }
  if (p.attn_mask != nullptr) {   // dense additive mask (no FlashMask / dropout / deterministic with it)
    if (p.d <= 64) { if (p.is_bf16) launch_bwd<64, true, false, false, false, true>(p, s); else launch_bwd<64, false, false, false, false, true>(p, s); }
    else { if (p.is_bf16) launch_bwd<128, true, false, false, false, true>(p, s); else launch_bwd<128, false, false, false, false, true>(p, s); }
    return;
  }
  if (p.softcap > 0.f) {          // softcap (no FlashMask / dropout / deterministic / dense mask with it)
    if (p.d <= 64) { if (p.is_bf16) launch_bwd<64, true, false, false, false, false, true>(p, s); else launch_bwd<64, false, false, false, false, false, true>(p, s); }
    else { if (p.is_bf16) launch_bwd<128, true, false, false, false, false, true>(p, s); else launch_bwd<128, false, false, false, false, false, true>(p, s); }
    return;
  }
  if (p.p_dropout > 0.f) {        // dropout: its own instantiations as well
    if (p.d <= 64) {
      if (p.is_bf16) { if (fm) launch_bwd<64, true, true, false, true>(p, s); else launch_bwd<64, true, false, false, true>(p, s); }
      else { if (fm) launch_bwd<64, false, true, false, true>(p, s); else launch_bwd<64, false, false, false, true>(p, s); }
    } else {
      if (p.is_bf16) { if (fm) launch_bwd<128, true, true, false, true>(p, s); else launch_bwd<128, true, false, false, true>(p, s); }
      else { if (fm) launch_bwd<128, false, true, false, true>(p, s); else launch_bwd<128, false, false, false, true>(p, s); }
    }
    return;
  }
  if (p.d <= 64) {
    if (p.is_bf16) { if (fm) launch_bwd<64, true, true>(p, s); else launch_bwd<64, true, false>(p, s); }
    else { if (fm) launch_bwd<64, false, true>(p, s); else launch_bwd<64, false, false>(p, s); }
  } else {
    if (p.is_bf16) { if (fm) launch_bwd<128, true, true>(p, s); else launch_bwd<128, true, false>(p, s); }
    else { if (fm) launch_bwd<128, false, true>(p, s); else launch_bwd<128, false, false>(p, s); }
  }
}

}  // namespace fa
