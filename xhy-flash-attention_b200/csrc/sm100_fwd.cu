// sm100_fwd.cu -- the tcgen05 / TMEM / TMA forward kernel (sm_100a).
//
// Computes what the reference's compute_attn_1rowblock computes (csrc/flash_attn/src/flash_fwd_kernel.h:121-583:
// O = softmax(scale * Q K^T + mask) V and LSE), restructured for Blackwell:
//
//   * persistent CTAs, one per SM; a work tile is 256 query rows (two 128-row Q tiles) of one (batch, q-head);
//   * warp 9  : TMA producer -- Q tiles once per work tile, K/V tiles through a 4-stage shared-memory ring;
//   * warp 8  : MMA issuer   -- S_t = Q_t K_j^T (tcgen05.mma, SS) into TMEM, O_t += P_t V_j (tcgen05.mma, TS: P read from TMEM);
//   * warps 0-3 / 4-7 : softmax warpgroups for Q tile 0 / 1 -- one thread per query row (TMEM lane == row, so row max
//     and row sum need no shuffles), exp2 with the folded scale, lazy (thresholded) O rescale, P written back over S in
//     TMEM as packed 16-bit, epilogue O -> smem -> coalesced global stores, LSE.
//   The two Q tiles ping-pong on the tensor core: while one warpgroup runs softmax the MMA warp works for the other.
//
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,256+HD) O1 [256+HD, 256+2HD).
#include <cstdlib>
#include <math_constants.h>
#include <type_traits>

#include "common.h"
#include "sm100.cuh"

#ifdef FA_TRACE
// Timeline probe (variant builds only, tools/trace_fwd.py): CTA 0 records clock64() at the hand-over points of its first
// key blocks -- role 0/1 = softmax warpgroup of Q tile 0/1 (first warp), role 2 = MMA issuer.
__device__ unsigned long long fa_trace_buf[3][96][8];
#define FA_TR(role, blk, ev) do { if (blockIdx.x == 0 && lane == 0 && (blk) < 96u) fa_trace_buf[role][blk][ev] = clock64(); } while (0)
extern "C" int fa_b200_trace_dump(unsigned long long *host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, fa_trace_buf, sizeof(fa_trace_buf));
}
// the clock is read only after `val` is available (the asm takes it as an input): times a data arrival, not an issue
#define FA_TR_DEP(role, blk, ev, val) do { if (blockIdx.x == 0 && lane == 0 && (blk) < 96u) { unsigned long long t__; \
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t__) : "r"(val)); fa_trace_buf[role][blk][ev] = t__; } } while (0)
#else
#define FA_TR(role, blk, ev) do { } while (0)
#define FA_TR_DEP(role, blk, ev, val) do { } while (0)
#endif

namespace fa {
namespace {

using namespace sm100;

constexpr int kBM = 128;          // rows per Q tile
constexpr int kBN = 128;          // keys per K/V tile
constexpr int kTilesPerCta = 2;   // Q tiles per work tile
constexpr int kStages = 4;        // K/V ring depth
constexpr int kThreads = 384;     // 2 softmax warpgroups + {MMA warp, TMA warp, 2 idle warps} (register-budget warpgroup)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.0f;  // log2 domain: P stays below 2^8

struct FwdTma {
  CUtensorMap q, k, v;
};

template <int HD> struct FwdSmem {
  static constexpr int kTileBytes = kBM * HD * 2;               // one Q or K or V tile
  static constexpr int kQOff = 0;
  static constexpr int kKVOff = kTilesPerCta * kTileBytes;
  static constexpr int kBarOff = kKVOff + kStages * kTileBytes;
  static constexpr int kNumBars = 2 + 2 + kStages * 2 + 2 + 2 + 2 + 2;
  static constexpr int kTmemSlotOff = kBarOff + kNumBars * 8;
  // FlashMask: record ring + its barriers, and per softmax warpgroup the staged mask vectors of one key block
  static constexpr int kFmRecOff = (kTmemSlotOff + 16 + 31) / 32 * 32;
  static constexpr int kFmBarOff = kFmRecOff + 16 * 32;
  static constexpr int kFmVecOff = kFmBarOff + 2 * 16 * 8;
  // dynamic tile scheduler: ring of tile indices + full/empty barriers
  static constexpr int kSchedOff = kFmVecOff + 2 * 4 * kBN * 4;
  static constexpr int kTotal = kSchedOff + 4 * 4 + 2 * 4 * 8;   // room for up to 4 entries
  static constexpr int kDynBytes = kTotal + 1024;               // slack for manual 1024-byte alignment
};

struct TileInfo {
  int bi, hq, hk, m0;
  int sq, sk, q_row0, k_row0, q_b, k_b;
  bool active[2];
  int n_lo[2], n_hi[2];
  int lo, hi;  // union block range
};

__device__ __forceinline__ bool decode_tile(const Params &p, int tile, int num_m, TileInfo &ti) {
  const int g = p.h / p.h_k;
  const int per_seg = g * num_m;
  const int seg = tile / per_seg, pos = tile % per_seg;
  ti.bi = seg / p.h_k;
  ti.hk = seg % p.h_k;
  ti.hq = ti.hk * g + pos % g;
  ti.m0 = (num_m - 1 - pos / g) * (kBM * kTilesPerCta);
  bool qb, kb;
  seq_bounds(p, ti.bi, ti.q_row0, ti.sq, ti.k_row0, ti.sk, qb, kb);
  ti.q_b = qb ? ti.bi : 0;
  ti.k_b = kb ? ti.bi : 0;
  if (ti.m0 >= ti.sq) return false;
  const int shift = ti.sk - ti.sq;
  const int shiftc = p.causal_bottom_right ? shift : 0;
  ti.lo = 0x7fffffff; ti.hi = 0;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int r0 = ti.m0 + t * kBM;
    ti.active[t] = r0 < ti.sq;
    const int r1 = min(r0 + kBM, ti.sq);
    int hi_col = ti.sk, lo_col = 0;
    if (p.is_causal) hi_col = min(hi_col, r1 - 1 + shiftc + 1);
    if (p.window_right >= 0) hi_col = min(hi_col, r1 - 1 + shift + p.window_right + 1);
    if (p.window_left >= 0) lo_col = max(lo_col, r0 + shift - p.window_left);
    int nlo = lo_col / kBN, nhi = (max(hi_col, 0) + kBN - 1) / kBN;
    if (!ti.active[t] || nhi <= nlo) { nlo = 0; nhi = 0; }
    ti.n_lo[t] = nlo; ti.n_hi[t] = nhi;
    if (nhi > nlo) { ti.lo = min(ti.lo, nlo); ti.hi = max(ti.hi, nhi); }
  }
  if (ti.hi == 0) ti.lo = 0;
  return true;
}

// Broadcast the (data-dependent, e.g. cu_seqlens-derived) tile description from lane 0 so that the compiler can keep
// loop bounds, stage indices and MMA descriptors in uniform registers in the single-issuer warps.
__device__ __forceinline__ int bcast0(int v) { return __shfl_sync(0xffffffffu, v, 0); }
__device__ __forceinline__ void uniformize(TileInfo &ti) {
  ti.hq = bcast0(ti.hq); ti.hk = bcast0(ti.hk); ti.m0 = bcast0(ti.m0);
  ti.q_row0 = bcast0(ti.q_row0); ti.k_row0 = bcast0(ti.k_row0); ti.q_b = bcast0(ti.q_b); ti.k_b = bcast0(ti.k_b);
  ti.lo = bcast0(ti.lo); ti.hi = bcast0(ti.hi);
#pragma unroll
  for (int t = 0; t < 2; ++t) { ti.n_lo[t] = bcast0(ti.n_lo[t]); ti.n_hi[t] = bcast0(ti.n_hi[t]); }
}

// The ordered list of key blocks a work tile visits, as seen identically by every role (producer, MMA issuer, both
// softmax warpgroups).  Without FlashMask it is the union range [lo, hi) of decode_tile.  With FlashMask (kFM) blocks
// that the column-sparse mask covers completely for a Q tile are dropped for that tile, and blocks dropped for both Q
// tiles are never loaded: the classification restates csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:835-853 on
// the per-block min/max of the four vectors (flash_mask.hpp:162-245, computed by flashmask_scan_kernel with block 128).
//
// The classification reads global memory, so it is done AHEAD of the consumers by a dedicated scanner warp (warp 10),
// 32 blocks per step (one per lane, results balloted), and published as records in a shared-memory ring guarded by
// full/empty mbarriers -- the reference uses generator warps the same way (flash_fwd_kernel_sm90.h:435).  A record is
// {first block of the 32-group, active bits for Q tile 0 / 1, partial bits for Q tile 0 / 1}; base < 0 ends a tile.
// Every consumer warp (producer, issuer, 8 softmax warps) reads every record.
constexpr int kFmRing = 16;        // records in flight
constexpr int kFmConsumers = 10;   // warps that read each record
struct Blk {
  int j;
  bool a[2];   // block takes part for Q tile t
  bool pm[2];  // FlashMask may mask single elements of the block for Q tile t (needs the element rule)
};
struct FmRing {
  uint32_t rec, full, empty;   // shared-memory addresses: records (8 x u32 each), full barriers, empty barriers
};
template <bool kFM> struct BlockSeq {
  int j, hi, nlo[2], nhi[2];
  // FlashMask consumer state
  FmRing ring;
  uint32_t rec_it, act[2], par[2];
  int base, lane;
  bool done;

  __device__ __forceinline__ void setup(const FmRing &r, int lane_) { ring = r; rec_it = 0; lane = lane_; }
  __device__ __forceinline__ void init(const Params &p, const TileInfo &ti) {
    j = ti.lo; hi = ti.hi;
#pragma unroll
    for (int t = 0; t < 2; ++t) { nlo[t] = ti.n_lo[t]; nhi[t] = ti.n_hi[t]; }
    if constexpr (kFM) { act[0] = act[1] = par[0] = par[1] = 0u; done = false; base = 0; }
  }
  __device__ __forceinline__ bool next(Blk &b) {
    if constexpr (!kFM) {
      if (j >= hi) return false;
      b.j = j;
#pragma unroll
      for (int t = 0; t < 2; ++t) { b.a[t] = j >= nlo[t] && j < nhi[t]; b.pm[t] = false; }
      ++j;
      return true;
    } else {
      while ((act[0] | act[1]) == 0u) {
        if (done) return false;
        const uint32_t slot = rec_it % kFmRing;
        mbar_wait(ring.full + slot * 8, (rec_it / kFmRing) & 1);
        uint32_t r0, r1, r2, r3, r4;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(ring.rec + slot * 32));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r4) : "r"(ring.rec + slot * 32 + 16));
        __syncwarp();
        if (lane == 0) mbar_arrive(ring.empty + slot * 8);
        ++rec_it;
        base = (int)r0;
        if (base < 0) { done = true; return false; }
        act[0] = r1; act[1] = r2; par[0] = r3; par[1] = r4;
      }
      const int bit = __ffs(act[0] | act[1]) - 1;
      b.j = base + bit;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        b.a[t] = (act[t] >> bit) & 1u; b.pm[t] = (par[t] >> bit) & 1u;
        act[t] &= ~(1u << bit); par[t] &= ~(1u << bit);
      }
      return true;
    }
  }
  // a consumer that leaves a tile early must still drain its records (not used today: every role walks to the end)
};

// Scanner (one warp): classifies the key blocks of every work tile of this CTA and publishes the records.
// Two levels: groups of 32 blocks are first tested against the group-wide min/max (flashmask_scan_l2_kernel) so that a
// long causal range with a narrow visible band (documents in a 256k-token row) costs one test per 4096 keys.
__device__ __forceinline__ void fm_scan_tile(const Params &p, const TileInfo &ti, const FmRing &ring, uint32_t &rec_it, int lane) {
  const int nbp = (ceil_div(p.seqlen_k, kBN) + 3) & ~3;
  const int64_t plane = (int64_t)p.b * p.h_mask * nbp;
  const int64_t row = (int64_t)ti.bi * p.h_mask + ti.hq / (p.h / p.h_mask);
  const int32_t *mm = p.flashmask_maxmin + row * nbp;
  const bool has_lte = p.lt_end != nullptr, has_ute = p.ut_end != nullptr, has_uts = p.ut_start != nullptr;
  const bool has_l2 = flashmask_has_l2(p.seqlen_k, kBN);
  const int nsup = ceil_div(nbp, 32);
  const int64_t plane2 = (int64_t)p.b * p.h_mask * nsup;
  const int32_t *mm2 = p.flashmask_maxmin + 8 * plane + row * nsup;
  int r0[2], r1[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) { r0[t] = ti.m0 + t * kBM; r1[t] = min(r0[t] + kBM, ti.sq); }
  auto publish = [&](int base, uint32_t a0, uint32_t a1, uint32_t p0, uint32_t p1) {
    const uint32_t slot = rec_it % kFmRing;
    mbar_wait(ring.empty + slot * 8, ((rec_it / kFmRing) & 1) ^ 1);
    if (lane == 0) {
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(ring.rec + slot * 32), "r"((uint32_t)base), "r"(a0), "r"(a1), "r"(p0) : "memory");
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(ring.rec + slot * 32 + 16), "r"(p1) : "memory");
      mbar_arrive(ring.full + slot * 8);
    }
    __syncwarp();
    ++rec_it;
  };
  for (int wbase = ti.lo / 32; wbase * 32 < ti.hi; wbase += 32) {   // window of 32 groups
    const int g = wbase + lane;
    bool cand = g * 32 < ti.hi;
    if (cand && has_l2) {
      const int lts_max = mm2[g];
      const int lte_min = has_lte ? mm2[5 * plane2 + g] : 0x7fffffff;
      const int ute_min = has_ute ? mm2[3 * plane2 + g] : 0;
      const int uts_max = has_uts ? mm2[6 * plane2 + g] : 0;
      bool all_full = true;
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const bool in = ti.n_hi[t] > ti.n_lo[t];
        const bool fully = (r0[t] >= lts_max && r1[t] <= lte_min) || (has_ute && r0[t] >= uts_max && r1[t] <= ute_min);
        all_full = all_full && (!in || fully);
      }
      cand = !all_full;
    }
    uint32_t sup = __ballot_sync(0xffffffffu, cand);
    while (sup) {
      const int gbit = __ffs(sup) - 1;
      sup &= sup - 1;
      const int base = (wbase + gbit) * 32;
      const int jb = base + lane;
      bool a_[2] = {false, false}, p_[2] = {false, false};
      if (jb >= ti.lo && jb < ti.hi) {
        const int lts_max = mm[jb], lts_min = mm[plane + jb];
        const int lte_max = has_lte ? mm[4 * plane + jb] : 0x7fffffff, lte_min = has_lte ? mm[5 * plane + jb] : 0x7fffffff;
        const int ute_max = has_ute ? mm[2 * plane + jb] : 0, ute_min = has_ute ? mm[3 * plane + jb] : 0;
        const int uts_max = has_uts ? mm[6 * plane + jb] : 0, uts_min = has_uts ? mm[7 * plane + jb] : 0;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const bool in = jb >= ti.n_lo[t] && jb < ti.n_hi[t];
          const bool fully = (r0[t] >= lts_max && r1[t] <= lte_min) || (has_ute && r0[t] >= uts_max && r1[t] <= ute_min);
          const bool clear = (r1[t] <= lts_min || r0[t] >= lte_max) && (!has_ute || r1[t] <= uts_min || r0[t] >= ute_max);
          a_[t] = in && !fully;
          p_[t] = a_[t] && !clear;
        }
      }
      const uint32_t a0 = __ballot_sync(0xffffffffu, a_[0]), a1 = __ballot_sync(0xffffffffu, a_[1]);
      const uint32_t p0 = __ballot_sync(0xffffffffu, p_[0]), p1 = __ballot_sync(0xffffffffu, p_[1]);
      if ((a0 | a1) != 0u) publish(base, a0, a1, p0, p1);
    }
  }
  publish(-1, 0u, 0u, 0u, 0u);   // end of tile
}

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// ---- packed fp32x2 arithmetic (FFMA2 / FADD2: two elements per issue slot) ----
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk2(uint64_t v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// 2^x for a pair, on the FMA pipe instead of the MUFU: x = j + f with j = round(x), f in [-0.5, 0.5];
// 2^f by a degree-3 polynomial (max relative error 7.7e-5, well below the 16-bit rounding of P), 2^j by adding j to the
// exponent field.  The forward softmax is MUFU-bound on B200 (16 ex2/clk/SM against 8192 tensor flop/clk/SM at d128:
// the exponentials take as long as the MMAs), so a quarter of the exponentials are moved here.
__device__ __forceinline__ void exp2_poly2(uint64_t x2, float &p0, float &p1) {
  float x0, x1;
  unpk2(x2, x0, x1);
  x0 = fmaxf(x0, -125.f); x1 = fmaxf(x1, -125.f);   // keeps the result a normal number (and maps -inf to ~0)
  const uint64_t xc = pk2(x0, x1);
  const uint64_t magic = pk2(12582912.f, 12582912.f);           // 1.5 * 2^23: low mantissa bits = round(x)
  const uint64_t jm = fadd2(xc, magic);
  const uint64_t jf = fadd2(jm, pk2(-12582912.f, -12582912.f)); // float(j)
  const uint64_t f = ffma2(jf, pk2(-1.f, -1.f), xc);            // x - j
  uint64_t q = ffma2(pk2(0.05508868396282196f, 0.05508868396282196f), f, pk2(0.24260404706001282f, 0.24260404706001282f));
  q = ffma2(q, f, pk2(0.6932762265205383f, 0.6932762265205383f));
  q = ffma2(q, f, pk2(0.9999289512634277f, 0.9999289512634277f));
  float q0, q1, j0, j1;
  unpk2(q, q0, q1);
  unpk2(jm, j0, j1);
  p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(j0) << 23));
  p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(j1) << 23));
}

// Tile walk of one CTA.  Two policies:
//   static  (dense, varlen): tile = blockIdx.x + i * gridDim.x.  Dense causal tiles of one head repeat with a period that
//           the stride steps through evenly, so the static walk is balanced and costs nothing (measured: any ring in
//           front of the roles costs ~3 % on BASELINE config 3).
//   dynamic (FlashMask): the reference's DynamicPersistentTileScheduler (csrc/flashmask_v2/tile_scheduler.hpp:618-765,
//           next tile by atomicAdd on a global counter).  With a column-sparse mask the work of a tile depends on the
//           documents it touches, and a static stride left SMs idle (document mask over 256k tokens: 18.5 -> 13.7 ms).
//           Warp 11 claims tiles -- the first is blockIdx.x, then gridDim.x + atomicAdd(counter) -- skips empty slots
//           and publishes the indices through a 4-entry shared-memory ring that every other role reads.  It claims
//           eagerly (one tile in hand beyond the ring): FlashMask tiles can be a few microseconds long, and a claim
//           made only when a slot frees (tried: 2 entries, lazy) put the atomic's latency back on the producer's
//           path (16.0 ms instead of 13.6).  -1 ends the stream.
constexpr int kTileRing = 4;
template <bool kDyn> struct TileWalk {
  uint32_t val, full, empty, it;
  int lane, cur, num_tiles;
  __device__ __forceinline__ void setup(uint32_t base, int lane_, int num_tiles_) {
    val = base; full = base + kTileRing * 4; empty = full + kTileRing * 8; it = 0; lane = lane_;
    cur = blockIdx.x; num_tiles = num_tiles_;
  }
  __device__ __forceinline__ int next() {
    if constexpr (!kDyn) {
      const int t = cur;
      cur += gridDim.x;
      return t < num_tiles ? t : -1;
    } else {
      const uint32_t slot = it % kTileRing;
      mbar_wait(full + slot * 8, (it / kTileRing) & 1);
      int tile;
      asm volatile("ld.shared.s32 %0, [%1];" : "=r"(tile) : "r"(val + slot * 4));
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + slot * 8);
      ++it;
      return tile;
    }
  }
};

template <int HD, bool kBf16, bool kFM, bool kDrop, bool kMask, bool kCap>
__global__ void __launch_bounds__(kThreads, 1)
fwd_kernel(const __grid_constant__ Params p, const __grid_constant__ FwdTma tm_param, const FwdTma *__restrict__ tm_dev,
           int num_m, int num_tiles, int *__restrict__ tile_counter) {
  using L = FwdSmem<HD>;
  const FwdTma &tm = tm_dev ? *tm_dev : tm_param;   // device-patched maps (flat varlen symbols) or the launch's own
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  // barrier addresses
  auto bar = [&](int i) { return sbase + L::kBarOff + i * 8; };
  const int B_QFULL = 0, B_QEMPTY = 2, B_KVFULL = 4, B_KVEMPTY = 4 + kStages, B_SFULL = 4 + 2 * kStages,
            B_PFULL = B_SFULL + 2, B_OFULL = B_PFULL + 2, B_PHALF = B_OFULL + 2;
  uint32_t *tmem_slot = (uint32_t *)(smem + L::kTmemSlotOff);
  const FmRing fm_ring = {sbase + L::kFmRecOff, sbase + L::kFmBarOff, sbase + L::kFmBarOff + kFmRing * 8};

  if (threadIdx.x == 0) {
    for (int t = 0; t < 2; ++t) {
      mbar_init(bar(B_QFULL + t), 1);
      mbar_init(bar(B_QEMPTY + t), kBM);
      mbar_init(bar(B_SFULL + t), 1);
      mbar_init(bar(B_PFULL + t), kBM);
      mbar_init(bar(B_PHALF + t), kBM);
      mbar_init(bar(B_OFULL + t), 1);
    }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar(B_KVFULL + s), 1);
      mbar_init(bar(B_KVEMPTY + s), 1);
    }
    if constexpr (kFM) {
      for (int r = 0; r < kFmRing; ++r) {
        mbar_init(sbase + L::kFmBarOff + r * 8, 1);
        mbar_init(sbase + L::kFmBarOff + (kFmRing + r) * 8, kFmConsumers);
      }
    }
    if constexpr (kFM) {
      for (int r = 0; r < kTileRing; ++r) {
        mbar_init(sbase + L::kSchedOff + kTileRing * 4 + r * 8, 1);
        mbar_init(sbase + L::kSchedOff + kTileRing * 4 + (kTileRing + r) * 8, 11);   // consumer warps 0-10
      }
    }
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc<512>(smem_u32(tmem_slot));
  if (warp == 9 && lane == 0) {
    if (tm_dev) { tma_acquire_desc(&tm.q); tma_acquire_desc(&tm.k); tma_acquire_desc(&tm.v); }
    tma_prefetch_desc(&tm.q); tma_prefetch_desc(&tm.k); tma_prefetch_desc(&tm.v);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_raw = *tmem_slot;
  const uint32_t tmem = tmem_raw;

  constexpr int kHalves = HD / 64;            // 64-element (128-byte) boxes per row
  constexpr uint32_t kHalfBytes = kBM * 128;  // one box: 128 rows x 128 bytes
  const uint32_t scale_dummy = 0; (void)scale_dummy;

  // register budget: 168 per thread at launch (384 threads); the issuer/producer warpgroup gives most of its registers
  // to the two softmax warpgroups, which hold a full 128-column S row per thread
  if (warp >= 8) {
   asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
   if (warp == 9) {
    // ======================= TMA producer =======================
    // the whole warp walks the block sequence (FlashMask classification ballots); lane 0 waits and issues
    {
      uint32_t kv_it = 0, qcnt[2] = {0, 0};
      BlockSeq<kFM> seq;
      seq.setup(fm_ring, lane);
      TileWalk<kFM> tw;
      tw.setup(sbase + L::kSchedOff, lane, num_tiles);
      for (int tile = tw.next(); tile >= 0; tile = tw.next()) {
        TileInfo ti;
        if (!decode_tile(p, tile, num_m, ti)) continue;
        seq.init(p, ti);
        bool q_loaded[2] = {false, false};
        Blk blk;
        while (seq.next(blk)) {
          if (lane == 0) {
            // K/V first: at a tile boundary the Q buffers are still owned by the previous tile's epilogue, the K/V ring
            // is not -- the first key block of the next tile is on its way while that epilogue runs
#pragma unroll
            for (int kv = 0; kv < 2; ++kv) {
              const uint32_t st = kv_it % kStages, ph = (kv_it / kStages) & 1;
              mbar_wait(bar(B_KVEMPTY + st), ph ^ 1);
              mbar_expect_tx(bar(B_KVFULL + st), L::kTileBytes);
#pragma unroll
              for (int hf = 0; hf < kHalves; ++hf)
                tma_load_4d(sbase + L::kKVOff + st * L::kTileBytes + hf * kHalfBytes, kv == 0 ? &tm.k : &tm.v,
                            bar(B_KVFULL + st), hf * 64, ti.hk, ti.k_row0 + blk.j * kBN, ti.k_b);
              ++kv_it;
            }
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              if (!blk.a[t] || q_loaded[t]) continue;
              mbar_wait(bar(B_QEMPTY + t), (qcnt[t] & 1) ^ 1);
              mbar_expect_tx(bar(B_QFULL + t), L::kTileBytes);
#pragma unroll
              for (int hf = 0; hf < kHalves; ++hf)
                tma_load_4d(sbase + L::kQOff + t * L::kTileBytes + hf * kHalfBytes, &tm.q, bar(B_QFULL + t), hf * 64, ti.hq,
                            ti.q_row0 + ti.m0 + t * kBM, ti.q_b);
              ++qcnt[t];
            }
          }
#pragma unroll
          for (int t = 0; t < 2; ++t) q_loaded[t] |= blk.a[t];
          if constexpr (kFM) __syncwarp();
        }
      }
    }
  } else if (warp == 8) {
    // ======================= MMA issuer =======================
    // The whole warp runs this (warp-uniform control flow keeps descriptors in uniform registers); one elected lane
    // issues the tcgen05 instructions.
    {
      const uint32_t tmem = bcast0((int)tmem_raw);
      constexpr uint32_t idesc_s = make_idesc(kBf16, kBM, kBN, false, false);
      constexpr uint32_t idesc_o = make_idesc(kBf16, kBM, HD, false, true);
      uint32_t kv_it = 0, qcnt[2] = {0, 0}, pcnt[2] = {0, 0};
      auto commit = [&](uint32_t b) {
        if (elect_one()) umma_commit(b);
        __syncwarp();
      };
      auto issue_s = [&](int t, uint32_t kstage) {
        const uint32_t qa = sbase + L::kQOff + t * L::kTileBytes;
        const uint32_t ka = sbase + L::kKVOff + kstage * L::kTileBytes;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) {
            const uint32_t off = (k / 4) * kHalfBytes + (k % 4) * 32;
            umma_ss(tmem + t * kBN, make_sdesc_sw128(qa + off, 16, 1024), make_sdesc_sw128(ka + off, 16, 1024), idesc_s,
                    k > 0);
          }
          umma_commit(bar(B_SFULL + t));
        }
        __syncwarp();
      };
      // O_t += P_t V in two halves of 64 keys: the first half is issued as soon as the softmax warpgroup has stored the
      // first 64 columns of P (B_PHALF), so it runs under the second half of the exponentials; only the second half is
      // left on the S -> softmax -> PV -> S dependency chain of the tile.
      auto issue_pv_half = [&](int t, uint32_t vstage, int half, bool accumulate) {
        const uint32_t va = sbase + L::kKVOff + vstage * L::kTileBytes;
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBN / 32; ++kk) {
            const int k = half * (kBN / 32) + kk;
            umma_ts(tmem + 2 * kBN + t * HD, tmem + t * kBN + k * 8, make_sdesc_sw128(va + k * 2048, kHalfBytes, 1024),
                    idesc_o, (accumulate || k > 0) ? 1u : 0u);
          }
          if (half == 1) umma_commit(bar(B_OFULL + t));
        }
        __syncwarp();
      };
      BlockSeq<kFM> seq;
      seq.setup(fm_ring, lane);
      TileWalk<kFM> tw;
      tw.setup(sbase + L::kSchedOff, lane, num_tiles);
      for (int tile = tw.next(); tile >= 0; tile = tw.next()) {
        TileInfo ti;
        if (!bcast0(decode_tile(p, tile, num_m, ti) ? 1 : 0)) continue;
        uniformize(ti);
        seq.init(p, ti);
        Blk cur, nxt;
        if (!seq.next(cur)) continue;
        bool q_ready[2] = {false, false}, first[2] = {true, true};
        auto need_q = [&](int t) {
          if (!q_ready[t]) { mbar_wait(bar(B_QFULL + t), qcnt[t] & 1); ++qcnt[t]; q_ready[t] = true; }
        };
        // first K tile
        uint32_t kst = kv_it % kStages;
        mbar_wait(bar(B_KVFULL + kst), (kv_it / kStages) & 1);
        ++kv_it;
#pragma unroll
        for (int t = 0; t < 2; ++t)
          if (cur.a[t]) need_q(t);
        tc_fence_after();
#pragma unroll
        for (int t = 0; t < 2; ++t)
          if (cur.a[t]) issue_s(t, kst);
        commit(bar(B_KVEMPTY + kst));
        bool have = true;
        while (have) {
          const bool has_next = seq.next(nxt);
          const uint32_t vst = kv_it % kStages;
          mbar_wait(bar(B_KVFULL + vst), (kv_it / kStages) & 1);
          ++kv_it;
          if (p.seqused_k != nullptr && cur.j * kBN + kBN > ti.sk) {
            // KV cache: rows at or beyond the used length hold arbitrary bits (possibly NaN) and P = 0 times NaN would
            // poison O -- zero the tail rows of this V tile (the reference zero-fills the V tail,
            // flash_fwd_kernel.h:346-349).  A row is 128 bytes in each 64-column box, whatever the swizzle.
            const int first = max(ti.sk - cur.j * kBN, 0);
            const uint32_t vt = sbase + L::kKVOff + vst * L::kTileBytes;
            for (int i = first * 8 + lane; i < kBN * 8; i += 32)
#pragma unroll
              for (int hf = 0; hf < kHalves; ++hf) sts128(vt + hf * kHalfBytes + i * 16, 0u, 0u, 0u, 0u);
            fence_proxy_async_smem();
            __syncwarp();
          }
          uint32_t nkst = 0;
          bool next_waited = false;
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (cur.a[t]) {
              FA_TR(2, pcnt[t], t * 4 + 0);
              mbar_wait(bar(B_PHALF + t), pcnt[t] & 1);
              FA_TR(2, pcnt[t], t * 4 + 1);
              tc_fence_after();
              issue_pv_half(t, vst, 0, !first[t]);
              mbar_wait(bar(B_PFULL + t), pcnt[t] & 1);
              FA_TR(2, pcnt[t], t * 4 + 2);
              ++pcnt[t];
              tc_fence_after();
              issue_pv_half(t, vst, 1, !first[t]);
              first[t] = false;
            }
            if (has_next && nxt.a[t]) {
              if (!next_waited) {
                nkst = kv_it % kStages;
                mbar_wait(bar(B_KVFULL + nkst), (kv_it / kStages) & 1);
                next_waited = true;
              }
              need_q(t);
              tc_fence_after();
              issue_s(t, nkst);
              FA_TR(2, pcnt[t] - 1, t * 4 + 3);
            }
          }
          commit(bar(B_KVEMPTY + vst));
          if (has_next) {   // a block of the sequence is active for at least one Q tile, so next_waited holds
            ++kv_it;
            commit(bar(B_KVEMPTY + nkst));
          }
          cur = nxt;
          have = has_next;
        }
      }
    }
   } else if (warp == 10) {
    // ======================= FlashMask scanner =======================
    if constexpr (kFM) {
      uint32_t rec_it = 0;
      TileWalk<kFM> tw;
      tw.setup(sbase + L::kSchedOff, lane, num_tiles);
      for (int tile = tw.next(); tile >= 0; tile = tw.next()) {
        TileInfo ti;
        if (!decode_tile(p, tile, num_m, ti)) continue;
        fm_scan_tile(p, ti, fm_ring, rec_it, lane);
      }
    }
   } else if (warp == 11) {
    // ======================= tile scheduler (FlashMask only) =======================
    if constexpr (kFM) {
      const uint32_t val = sbase + L::kSchedOff, full = val + kTileRing * 4, empty = full + kTileRing * 8;
      auto claim = [&]() {
        int t = 0;
        if (lane == 0) t = (int)gridDim.x + atomicAdd(tile_counter, 1);
        return bcast0(t);
      };
      int tile = blockIdx.x;
      for (uint32_t it = 0;; ++it) {
        while (tile < num_tiles) {             // skip slots that hold no rows (m0 beyond the sequence)
          TileInfo ti;
          if (bcast0(decode_tile(p, tile, num_m, ti) ? 1 : 0)) break;
          tile = claim();
        }
        const int pub = tile < num_tiles ? tile : -1;
        const uint32_t slot = it % kTileRing;
        mbar_wait_relaxed(empty + slot * 8, ((it / kTileRing) & 1) ^ 1);
        if (lane == 0) {
          asm volatile("st.shared.s32 [%0], %1;" ::"r"(val + slot * 4), "r"(pub) : "memory");
          mbar_arrive(full + slot * 8);
        }
        __syncwarp();
        if (pub < 0) break;
        tile = claim();
      }
    }
   }
  } else {
    // ======================= softmax warpgroups =======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    const int t = warp / 4;                    // Q tile of this warpgroup
    const int r_in_tile = threadIdx.x - t * kBM;
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const uint32_t tS = tmem + lane_base + t * kBN;
    const uint32_t tO = tmem + lane_base + 2 * kBN + t * HD;
    // softcap (kCap instantiation): logits = softcap * tanh(scale * s / softcap), applied to S right after the load and
    // before any mask (csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:537-556,1064-1067); from there on the values
    // are tanh outputs and the "scale" of the softmax is softcap
    const float scale_nat = kCap ? p.softcap : p.softmax_scale;
    const float scale_log2 = scale_nat * kLog2e;
    const float pre_cap = kCap ? p.softmax_scale / p.softcap : 0.f;
    constexpr bool kPipe = !kCap && !kMask && !kDrop;       // fast path of the block loop: plain and FlashMask instantiations
    uint32_t scnt = 0, ocnt = 0;
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    BlockSeq<kFM> seq;
    seq.setup(fm_ring, lane);
    TileWalk<kFM> tw;
    tw.setup(sbase + L::kSchedOff, lane, num_tiles);
    for (int tile = tw.next(); tile >= 0; tile = tw.next()) {
      TileInfo ti;
      if (!decode_tile(p, tile, num_m, ti)) continue;
      if (!ti.active[t]) {
        if constexpr (kFM) {   // the records of this tile must still be consumed
          seq.init(p, ti);
          Blk skip;
          while (seq.next(skip)) {}
        }
        continue;
      }
      const int row = ti.m0 + t * kBM + r_in_tile;
      // dropout (kDrop instantiation): the mask of common.h's dropout_keep -- one Philox-4x32-7 call per 16
      // consecutive columns of a row, a weight is kept iff its byte <= floor(255 * p_keep).  A thread owns a row, so
      // every 32-column chunk of P costs it two calls.  The row sum is taken BEFORE dropout and 1 / p_keep is folded
      // into the final normalisation (flash_fwd_kernel.h:399-409, softmax.h:381-447).
      DropCtx dctx;
      if constexpr (kDrop) dctx = make_drop_ctx(p, ti.bi, ti.hq, row);
      T *optr = (T *)p.o.ptr + (int64_t)ti.q_b * p.o.batch_stride + (int64_t)(ti.q_row0) * p.o.row_stride +
                (int64_t)ti.hq * p.o.head_stride;
      float *lse_ptr = p.softmax_lse + ((int64_t)ti.bi * p.h + ti.hq) * p.seqlen_q;
      const MaskCtx mc = make_mask_ctx(p, ti.bi, ti.hq, ti.sq, ti.sk);
      const int shiftc = mc.bottom_right ? mc.shift : 0;
      const int r0 = ti.m0 + t * kBM;
      float m_ref = -CUDART_INF_F, l = 0.f;
      // dense additive mask (kMask instantiation; the flat C ABI's attn_mask, csrc/flash_attn/src/softmax.h:360-379):
      // logits = scale * (q.k + mask[b, h % head_mod, row % seq_mod, col] * unscale_softmax); a thread adds its row
      const T *mask_row = nullptr;
      if constexpr (kMask)
        mask_row = (const T *)p.attn_mask + (((int64_t)ti.bi * p.mask_head_mod + ti.hq % p.mask_head_mod) * p.mask_seq_mod +
                                             row % p.mask_seq_mod) * (int64_t)p.seqlen_k;
      seq.init(p, ti);
      Blk blk;
      uint32_t nb = 0;   // blocks of this Q tile processed so far
      while (seq.next(blk)) {
        if (!blk.a[t]) continue;
        const int j = blk.j;
        if (warp % 4 == 0) FA_TR(t, scnt, 0);
        mbar_wait(bar(B_SFULL + t), scnt & 1);
        if (warp % 4 == 0) FA_TR(t, scnt, 1);
        ++scnt;
        tc_fence_after();
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32(tS + c * 32, s[c]);
        tmem_wait_ld();
        if (warp % 4 == 0) FA_TR_DEP(t, scnt - 1, 2, s[3][31]);   // the last loaded column has arrived
        const int c0 = j * kBN;
        if constexpr (kCap) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              float y;
              asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(__uint_as_float(s[c][i]) * pre_cap));
              s[c][i] = __float_as_uint(y);
            }
        }
        if constexpr (kMask) {
          // (fetching the 128 mask values before the wait for S was tried: 64 more live registers, spills, slower)
          const bool vec_ok = (p.seqlen_k % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.attn_mask) & 15) == 0) &&
                              c0 + kBN <= p.seqlen_k;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              const int col = c0 + c * 32 + i;
              if (vec_ok) {
                const uint4 mv = *reinterpret_cast<const uint4 *>(mask_row + col);
                const T *me = reinterpret_cast<const T *>(&mv);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  s[c][i + e] = __float_as_uint(fmaf(to_float(me[e]), p.unscale_softmax, __uint_as_float(s[c][i + e])));
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if (col + e < p.seqlen_k)
                    s[c][i + e] = __float_as_uint(fmaf(to_float(mask_row[col + e]), p.unscale_softmax, __uint_as_float(s[c][i + e])));
              }
            }
          }
        }
        const bool full = (c0 + kBN <= ti.sk) && (!mc.causal || c0 + kBN - 1 <= r0 + shiftc) &&
                          (mc.wr < 0 || c0 + kBN - 1 <= r0 + mc.shift + mc.wr) &&
                          (mc.wl < 0 || c0 >= r0 + kBM - 1 + mc.shift - mc.wl) &&
                          (kFM ? !blk.pm[t] : (mc.lts == nullptr));
        if (!full) {
          // boundary / flashmask blocks: the generic mask rule runs in a rolled loop (emitted once, keeps the hot loop
          // inside the instruction cache); the unrolled pass below only applies the resulting bit mask
          uint32_t mb[4];
          if (mc.lts == nullptr || kFM) {  // seqlen tail / causal / window: the visible columns of a row form an interval
            int vlo, vhi;
            visible_cols(mc, row, vlo, vhi);
#pragma unroll
            for (int c = 0; c < 4; ++c) mb[c] = outside_bits(vlo, vhi, c0 + c * 32);
            if (kFM && blk.pm[t]) {
              // FlashMask element rule for a partially masked block: the warpgroup stages the block's 128 entries of
              // each vector in shared memory (one per thread), then every thread tests its row against all columns
              const uint32_t vec = sbase + L::kFmVecOff + t * (4 * kBN * 4);
              named_bar_sync(3 + t, kBM);   // the previous partial block has been read out
              {
                const int col = c0 + r_in_tile;
                const bool ok = col < ti.sk;
                const int lts_ = ok ? mc.lts[col] : 0x7fffffff;
                const int lte_ = ok ? (mc.lte ? mc.lte[col] : 0x7fffffff) : 0x7fffffff;
                const int ute_ = ok && mc.ute ? mc.ute[col] : 0;
                const int uts_ = ok && mc.ute && mc.uts ? mc.uts[col] : 0;
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(vec + (0 * kBN + r_in_tile) * 4), "r"(lts_) : "memory");
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(vec + (1 * kBN + r_in_tile) * 4), "r"(lte_) : "memory");
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(vec + (2 * kBN + r_in_tile) * 4), "r"(uts_) : "memory");
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(vec + (3 * kBN + r_in_tile) * 4), "r"(ute_) : "memory");
              }
              named_bar_sync(3 + t, kBM);
              const bool two_bands = mc.ute != nullptr;
              uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = 0u;
#pragma unroll 1
              for (int c = 0; c < 4; ++c) {
                uint32_t word = 0u;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                  const float4 a = lds128(vec + (0 * kBN + c * 32 + i) * 4), b4 = lds128(vec + (1 * kBN + c * 32 + i) * 4);
                  const int as[4] = {__float_as_int(a.x), __float_as_int(a.y), __float_as_int(a.z), __float_as_int(a.w)};
                  const int bs[4] = {__float_as_int(b4.x), __float_as_int(b4.y), __float_as_int(b4.z), __float_as_int(b4.w)};
#pragma unroll
                  for (int e = 0; e < 4; ++e) word |= (row >= as[e] && row < bs[e]) ? (1u << (i + e)) : 0u;
                }
                if (two_bands) {
#pragma unroll
                  for (int i = 0; i < 32; i += 4) {
                    const float4 a = lds128(vec + (2 * kBN + c * 32 + i) * 4), b4 = lds128(vec + (3 * kBN + c * 32 + i) * 4);
                    const int as[4] = {__float_as_int(a.x), __float_as_int(a.y), __float_as_int(a.z), __float_as_int(a.w)};
                    const int bs[4] = {__float_as_int(b4.x), __float_as_int(b4.y), __float_as_int(b4.z), __float_as_int(b4.w)};
#pragma unroll
                    for (int e = 0; e < 4; ++e) word |= (row >= as[e] && row < bs[e]) ? (1u << (i + e)) : 0u;
                  }
                }
                w0 = c == 0 ? word : w0; w1 = c == 1 ? word : w1; w2 = c == 2 ? word : w2; w3 = c == 3 ? word : w3;
              }
              mb[0] |= w0; mb[1] |= w1; mb[2] |= w2; mb[3] |= w3;
            }
          } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) mb[c] = 0u;
#pragma unroll 1
            for (int i = 0; i < kBN; ++i)
              if (is_masked(mc, row, c0 + i)) mb[i >> 5] |= 1u << (i & 31);
          }
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if ((mb[c] >> i) & 1u) s[c][i] = __float_as_uint(-CUDART_INF_F);
        }
        // ---- fast path: exponentials against the EXISTING reference maximum, row maximum computed alongside ----
        // Timeline probe (FA_TRACE build, C3): of the 2150 cycles a softmax warp spends per key block, ~750 went into the
        // row-maximum pass, during which MUFU idles; the exponentials that follow are MUFU-bound.  The lazy-rescale rule
        // only needs "block maximum <= m_ref + 2^8 (log2 domain)", so for every block but a tile's first the
        // exponentials of the first 64 columns start at once against m_ref and the 64 FMNMX3 of the maximum are
        // interleaved with them (independent chains that fill MUFU's issue gaps).  The check is made before the first
        // half of P is announced; if some row of the warp fails it, the block is redone on the full path below (S is
        // still in registers; only P columns that nobody has been told about were written).
        bool done = false;
        if constexpr (kPipe) {
          if (!__any_sync(0xffffffffu, nb == 0 || m_ref == -CUDART_INF_F)) {
            const float msc_f = m_ref * scale_log2;
            const uint64_t sc2 = pk2(scale_log2, scale_log2), nm2 = pk2(-msc_f, -msc_f);
            uint64_t suma = pk2(0.f, 0.f), sumb = pk2(0.f, 0.f);
            float mxa = -CUDART_INF_F, mxb = -CUDART_INF_F;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const uint64_t x2 = ffma2(pk2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, nm2);
                float p0, p1;
                if (i % 4 == 3) {
                  exp2_poly2(x2, p0, p1);
                } else {
                  float x0, x1;
                  unpk2(x2, x0, x1);
                  p0 = ex2(x0); p1 = ex2(x1);
                }
                // the row maximum over all four chunks rides along: two 3-input maxima per pair of exponentials
                mxa = fmax3(mxa, __uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1]));
                mxb = fmax3(mxb, __uint_as_float(s[c + 2][2 * i]), __uint_as_float(s[c + 2][2 * i + 1]));
                if (i & 1) sumb = fadd2(sumb, pk2(p0, p1)); else suma = fadd2(suma, pk2(p0, p1));
                pk[i] = pack2<kBf16>(p0, p1);
              }
              tmem_st16(tS + c * 16, pk);
            }
            const float mxf = fmaxf(mxa, mxb);
            if (!__any_sync(0xffffffffu, (mxf - m_ref) * scale_log2 > kRescaleThreshold)) {
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(bar(B_PHALF + t));
              if (warp % 4 == 0) FA_TR(t, scnt - 1, 4);
#pragma unroll
              for (int c = 2; c < 4; ++c) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                  const uint64_t x2 = ffma2(pk2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, nm2);
                  float p0, p1;
                  if (i % 4 == 3) {
                    exp2_poly2(x2, p0, p1);
                  } else {
                    float x0, x1;
                    unpk2(x2, x0, x1);
                    p0 = ex2(x0); p1 = ex2(x1);
                  }
                  if (i & 1) sumb = fadd2(sumb, pk2(p0, p1)); else suma = fadd2(suma, pk2(p0, p1));
                  pk[i] = pack2<kBf16>(p0, p1);
                }
                tmem_st16(tS + c * 16, pk);
              }
              float a0, a1;
              unpk2(fadd2(suma, sumb), a0, a1);
              l += a0 + a1;
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(bar(B_PFULL + t));
              if (warp % 4 == 0) FA_TR(t, scnt - 1, 5);
              done = true;
            }
          }
        }
        if (!done) {
          float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F, mx2 = -CUDART_INF_F, mx3 = -CUDART_INF_F;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            mx0 = fmaxf(mx0, __uint_as_float(s[0][i]));
            mx1 = fmaxf(mx1, __uint_as_float(s[1][i]));
            mx2 = fmaxf(mx2, __uint_as_float(s[2][i]));
            mx3 = fmaxf(mx3, __uint_as_float(s[3][i]));
          }
          const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
          if (nb == 0) {
            m_ref = mx;
          } else {
            const bool need = (mx > m_ref) && ((mx - m_ref) * scale_log2 > kRescaleThreshold || m_ref == -CUDART_INF_F);
            if (__any_sync(0xffffffffu, need)) {
              // the previous PV of this tile must have landed in O before it is rescaled
              mbar_wait(bar(B_OFULL + t), (ocnt + nb - 1) & 1);
              tc_fence_after();
              const float m_new = fmaxf(m_ref, mx);
              const float corr = (m_new == -CUDART_INF_F) ? 1.f : ex2((m_ref - m_new) * scale_log2);
              l *= corr;
              m_ref = m_new;
#pragma unroll
              for (int c = 0; c < HD / 32; ++c) {
                uint32_t o[32];
                tmem_ld32(tO + c * 32, o);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
                tmem_st32(tO + c * 32, o);
              }
            }
          }
          const float msc = (m_ref == -CUDART_INF_F) ? 0.f : m_ref * scale_log2;
          if (warp % 4 == 0) FA_TR_DEP(t, scnt - 1, 3, __float_as_uint(msc));   // the row maximum is known
          if (warp % 4 == 0) FA_TR_DEP(t, scnt - 1, 6, __float_as_uint(mx));
          const uint64_t sc2 = pk2(scale_log2, scale_log2), nm2 = pk2(-msc, -msc);
          uint64_t suma = pk2(0.f, 0.f), sumb = pk2(0.f, 0.f);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t pk[16];
            uint4 rnd = make_uint4(0, 0, 0, 0);
            (void)rnd;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const uint64_t x2 = ffma2(pk2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, nm2);
              float p0, p1;
              if (i % 4 == 3) {          // every fourth pair: polynomial on the FMA pipe
                exp2_poly2(x2, p0, p1);
              } else {
                float x0, x1;
                unpk2(x2, x0, x1);
                p0 = ex2(x0); p1 = ex2(x1);
              }
              if (i & 1) sumb = fadd2(sumb, pk2(p0, p1)); else suma = fadd2(suma, pk2(p0, p1));
              pk[i] = pack2<kBf16>(p0, p1);
              if constexpr (kDrop) {
                // one byte-wise compare per 4 columns (0xff where the weight is kept), widened to the two 16-bit lanes
                // of a packed pair by a byte permute: the mask is applied to the PACKED weights, 7 instructions per 4
                // columns instead of a shift / mask / compare / select per column
                if (i % 8 == 0) {
                  rnd = philox4x32_7(dctx.seed, dctx.row_base + (uint32_t)((blk.j * kBN + c * 32 + 2 * i) >> 4), dctx.offset);
                  const uint32_t thr4 = dctx.threshold * 0x01010101u;
                  rnd = make_uint4(__vcmpleu4(rnd.x, thr4), __vcmpleu4(rnd.y, thr4), __vcmpleu4(rnd.z, thr4), __vcmpleu4(rnd.w, thr4));
                }
                const int e = (2 * i) & 15;                       // position of the pair inside its 16-column group
                const uint32_t w = e < 4 ? rnd.x : e < 8 ? rnd.y : e < 12 ? rnd.z : rnd.w;
                pk[i] &= __byte_perm(w, 0u, (e & 3) == 0 ? 0x1100u : 0x3322u);
              }
            }
            tmem_st16(tS + c * 16, pk);
            if (c == 1) {   // the first 64 columns of P are in TMEM: let the issuer start the first half of PV
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(bar(B_PHALF + t));
              if (warp % 4 == 0) FA_TR(t, scnt - 1, 4);
            }
          }
          {
            float a0, a1;
            unpk2(fadd2(suma, sumb), a0, a1);
            l += a0 + a1;
          }
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar(B_PFULL + t));
          if (warp % 4 == 0) FA_TR(t, scnt - 1, 5);
        }
        ++nb;
      }
      if (nb == 0) {
        // no visible key block for this Q tile: out = 0, lse = +-inf (no barrier was touched)
        if (row < ti.sq) {
          T *orow = optr + (int64_t)row * p.o.row_stride;
          for (int e = 0; e < p.d; e += 8) *reinterpret_cast<uint4 *>(orow + e) = make_uint4(0, 0, 0, 0);
          lse_ptr[row] = p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F;
        }
        continue;
      }
      // ---------------- epilogue ----------------
      mbar_wait(bar(B_OFULL + t), (ocnt + nb - 1) & 1);
      ocnt += nb;
      tc_fence_after();
      // A row with at least one visible key has l >= ~1 (its maximum contributes 2^0).  Masked elements that went
      // through the polynomial exp2 contribute 2^-125 each instead of 0, so "no visible key" is l below a tiny bound.
      const bool nonempty = l > 1e-30f;
      const float inv = nonempty ? (kDrop ? 1.f / ((1.f - p.p_dropout) * l) : 1.f / l) : 0.f;
      uint8_t *stage = smem + L::kQOff + t * L::kTileBytes;  // Q_t is dead: all S MMAs of this tile have completed
      constexpr int kChunks = HD / 8;                        // 16-byte chunks per row
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t o[32];
        tmem_ld32(tO + c * 32, o);
        tmem_wait_ld();
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint4 v;
          v.x = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 0]) * inv, __uint_as_float(o[q4 * 8 + 1]) * inv);
          v.y = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 2]) * inv, __uint_as_float(o[q4 * 8 + 3]) * inv);
          v.z = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 4]) * inv, __uint_as_float(o[q4 * 8 + 5]) * inv);
          v.w = pack2<kBf16>(__uint_as_float(o[q4 * 8 + 6]) * inv, __uint_as_float(o[q4 * 8 + 7]) * inv);
          const int chunk = c * 4 + q4;
          *reinterpret_cast<uint4 *>(stage + r_in_tile * (HD * 2) + ((chunk ^ (r_in_tile & (kChunks - 1))) * 16)) = v;
        }
      }
      tc_fence_before();
      if (row < ti.sq)
        lse_ptr[row] = nonempty ? (m_ref * scale_nat + __logf(l))
                                 : (p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F);
      named_bar_sync(1 + t, kBM);
      {
        const int chunk = r_in_tile % kChunks;
        const int rsub = r_in_tile / kChunks;
        constexpr int kRowsPerPass = kBM / kChunks;
#pragma unroll 4
        for (int pass = 0; pass < kBM / kRowsPerPass; ++pass) {
          const int rr = pass * kRowsPerPass + rsub;
          const int grow = ti.m0 + t * kBM + rr;
          if (grow < ti.sq && chunk * 8 < p.d) {
            const uint4 v = *reinterpret_cast<const uint4 *>(stage + rr * (HD * 2) + ((chunk ^ (rr & (kChunks - 1))) * 16));
            *reinterpret_cast<uint4 *>(optr + (int64_t)grow * p.o.row_stride + chunk * 8) = v;
          }
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(bar(B_QEMPTY + t));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tmem);
}

void sanitize(TmaDesc4 &d) {
  // a dimension of extent 1 never advances: give it any legal (16-byte multiple) stride
  // a dimension of extent 1 never advances: give it a legal, monotonic stride
  for (int i = 1; i < 4; ++i)
    if (d.dims[i] == 1) d.strides[i - 1] = (i == 1 ? d.dims[0] * 2 : d.strides[i - 2] * d.dims[i - 1] + 15) / 16 * 16;
}

TmaDesc4 desc_for(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch, int box_rows) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * 2; x.strides[1] = (uint64_t)t.row_stride * 2;
  x.strides[2] = (uint64_t)t.batch_stride * 2;
  x.box[0] = 64; x.box[1] = 1; x.box[2] = (uint32_t)box_rows; x.box[3] = 1;
  sanitize(x);
  return x;
}

bool tma_ok(const fa_b200_tensor &t, int heads, int batch) {
  if (((uintptr_t)t.ptr & 15) != 0) return false;
  if (t.row_stride % 8 != 0 || t.row_stride <= 0) return false;
  if (heads > 1 && (t.head_stride % 8 != 0 || t.head_stride <= 0)) return false;
  if (batch > 1 && (t.batch_stride % 8 != 0 || t.batch_stride <= 0)) return false;
  return true;
}

template <int HD, bool kBf16, bool kFM, bool kDrop, bool kMask = false, bool kCap = false> void launch_fwd(const Params &p, cudaStream_t s) {
  using L = FwdSmem<HD>;
  const bool packed_q = p.cu_seqlens_q && !p.varlen_padded_input;
  const bool packed_k = p.cu_seqlens_k && !p.varlen_padded_input;
  // packed tensors without host-known totals (the flat varlen symbols): the row extents are patched on the device
  const bool dev_rows = (packed_q && p.total_q < 0) || (packed_k && p.total_k < 0);
  const int64_t rows_q = packed_q ? (p.total_q >= 0 ? p.total_q : (int64_t)p.b * p.seqlen_q) : p.seqlen_q;
  const int64_t rows_k = packed_k ? (p.total_k >= 0 ? p.total_k : (int64_t)p.b * p.seqlen_k) : p.seqlen_k;
  FwdTma tm;
  tm.q = make_tma_4d(desc_for(p.q, p.d, p.h, rows_q, packed_q ? 1 : p.b, kBM));
  tm.k = make_tma_4d(desc_for(p.k, p.d, p.h_k, rows_k, packed_k ? 1 : p.b, kBN));
  tm.v = make_tma_4d(desc_for(p.v, p.d, p.h_k, rows_k, packed_k ? 1 : p.b, kBN));
  FwdTma *tm_dev = nullptr;
  if (dev_rows) {
    const int k_like[3] = {0, 1, 1};
    tm_dev = reinterpret_cast<FwdTma *>(patch_tma_rows(&tm.q, k_like, 3, p.cu_seqlens_q, p.cu_seqlens_k, p.b, s));
  }
  const int num_m = ceil_div(p.seqlen_q, kBM * kTilesPerCta);
  const int num_tiles = p.b * p.h * num_m;
  if (kFM) flashmask_prepare(p, kBN, s, true);   // per-128-key-block min/max of the mask vectors
  auto kern = fwd_kernel<HD, kBf16, kFM, kDrop, kMask, kCap>;
  static std::atomic<bool> smem_set[64];
  ensure_dynamic_smem(kern, L::kDynBytes, smem_set);
  const int grid = std::min(num_tiles, kFM ? device_sm_count() : forward_grid_sms());
  int *counter = nullptr;                          // stream-ordered scratch: the dynamic scheduler's tile counter
  if (kFM) {
    FA_CUDA(cudaMallocAsync(&counter, sizeof(int), s));
    FA_CUDA(cudaMemsetAsync(counter, 0, sizeof(int), s));
  }
  profile_mark(0, true, s);
  kern<<<grid, kThreads, L::kDynBytes, s>>>(p, tm, tm_dev, num_m, num_tiles, counter);
  profile_mark(0, false, s);
  const cudaError_t launch_err = cudaGetLastError();
  if (tm_dev) cudaFreeAsync(tm_dev, s);
  if (counter) FA_CUDA(cudaFreeAsync(counter, s));
  FA_CUDA(launch_err);
  note_launch(2);
}

}  // namespace

bool sm100_fwd_supported(const Params &p, std::string *why) {
  auto no = [&](const char *w) { if (why) *why = w; return false; };
  if (p.is_e4m3) {
    // FP8 forward (sm100_fwd2.cu, byte operands): strides are in bytes, TMA wants multiples of 16
    if (!sm100_fwd2_eligible(p)) return no("FP8 together with FlashMask / dropout / attn_mask / softcap");
    if (p.num_splits > 1) return no("FP8 split-KV");
    if (p.d % 16 != 0) return no("FP8 needs head_size % 16 == 0");
    if ((bool)p.cu_seqlens_q != (bool)p.cu_seqlens_k) return no("cu_seqlens_q and cu_seqlens_k go together");
    const bool pk = p.cu_seqlens_q && !p.varlen_padded_input;
    auto ok8 = [&](const fa_b200_tensor &t, int heads, int batch) {
      return ((uintptr_t)t.ptr & 15) == 0 && t.row_stride > 0 && t.row_stride % 16 == 0 &&
             (heads <= 1 || (t.head_stride > 0 && t.head_stride % 16 == 0)) &&
             (batch <= 1 || (t.batch_stride > 0 && t.batch_stride % 16 == 0));
    };
    if (!ok8(p.q, p.h, pk ? 1 : p.b) || !ok8(p.k, p.h_k, pk ? 1 : p.b) || !ok8(p.v, p.h_k, pk ? 1 : p.b))
      return no("FP8 strides/alignment not TMA-compatible");
    if (((uintptr_t)p.o.ptr & 15) != 0 || p.o.row_stride % 8 != 0 || p.o.head_stride % 8 != 0 || p.o.batch_stride % 8 != 0)
      return no("output not 16-byte aligned");
    return true;
  }
  if (p.d > 128 && !sm100_fwd2_eligible(p)) return no("head_size > 128 together with FlashMask / attn_mask / softcap");
  if (p.num_splits > 1) return no("split-KV");
  if (p.attn_mask && (p.lt_start || p.p_dropout > 0.f)) return no("dense attn_mask together with FlashMask / dropout");
  if (p.softcap > 0.f && (p.lt_start || p.p_dropout > 0.f || p.attn_mask)) return no("softcap together with FlashMask / dropout / attn_mask");
  const bool packed = p.cu_seqlens_q && !p.varlen_padded_input;
  if (p.cu_seqlens_q && !p.cu_seqlens_k) return no("cu_seqlens_q without cu_seqlens_k");
  if (p.cu_seqlens_k && !p.cu_seqlens_q) return no("cu_seqlens_k without cu_seqlens_q");
  if (!tma_ok(p.q, p.h, packed ? 1 : p.b) || !tma_ok(p.k, p.h_k, packed ? 1 : p.b) ||
      !tma_ok(p.v, p.h_k, packed ? 1 : p.b))
    return no("strides/alignment not TMA-compatible");
  if (((uintptr_t)p.o.ptr & 15) != 0 || p.o.row_stride % 8 != 0 || p.o.head_stride % 8 != 0 || p.o.batch_stride % 8 != 0)
    return no("output not 16-byte aligned");
  return true;
}

namespace {

template <int HD, bool kBf16> void dispatch_fwd(const Params &p, cudaStream_t s) {
  const bool fm = p.lt_start != nullptr && p.flashmask_maxmin != nullptr;   // block skipping needs the workspace
  const bool drop = p.p_dropout > 0.f;
  if (p.attn_mask != nullptr) { launch_fwd<HD, kBf16, false, false, true>(p, s); return; }   // dense mask: no FlashMask / dropout
  if (p.softcap > 0.f) { launch_fwd<HD, kBf16, false, false, false, true>(p, s); return; }       // softcap: likewise
  if (fm) { if (drop) launch_fwd<HD, kBf16, true, true>(p, s); else launch_fwd<HD, kBf16, true, false>(p, s); }
  else { if (drop) launch_fwd<HD, kBf16, false, true>(p, s); else launch_fwd<HD, kBf16, false, false>(p, s); }
}

}  // namespace

void sm100_fwd(const Params &p, cudaStream_t s) {
  // every head dim above 128 takes the split-row kernel of sm100_fwd2.cu.  For head dims <= 128 the one-thread-per-row
  // kernel below is faster (measured on C3: 1123 vs 1028 TFLOPS with two Q tiles per CTA, 1081 with one);
  // FA_B200_FWD2=1 selects the split-row kernel there too (A/B measurements)
  static const bool use_fwd2 = [] { const char *e = getenv("FA_B200_FWD2"); return e && e[0] == '1'; }();
  if (p.is_e4m3) { sm100_fwd2_e4m3(p, s); return; }
  if (sm100_fwd2_eligible(p) && (p.d > 128 || use_fwd2)) { sm100_fwd2(p, s); return; }
  if (p.d <= 64) { if (p.is_bf16) dispatch_fwd<64, true>(p, s); else dispatch_fwd<64, false>(p, s); }
  else { if (p.is_bf16) dispatch_fwd<128, true>(p, s); else dispatch_fwd<128, false>(p, s); }
}

}  // namespace fa
