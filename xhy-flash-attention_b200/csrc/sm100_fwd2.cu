// sm100_fwd2.cu -- the tcgen05 forward for the plain instantiation (no FlashMask / dropout / dense mask / softcap) and
// for head dims above 128: the softmax of a Q tile is SPLIT OVER TWO WARPGROUPS, half a row of S per thread.
//
// Same mathematics as the reference's compute_attn_1rowblock (csrc/flash_attn/src/flash_fwd_kernel.h:121-583); the
// structure the reference gets from its two ping-ponged consumer warpgroups + intra-warpgroup overlap
// (csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:1734-1795, flash_fwd_kernel_sm90.h:474-542) is obtained here as:
//
//   head dim <= 128 : two 128-row Q tiles per CTA ping-pong on the tensor core (S0/S1, O0/O1 in TMEM); FOUR softmax
//                     warpgroups (16 warps), two per Q tile.  One thread per (row, half): 64 columns of S.  The round-1
//                     kernel (sm100_fwd.cu) ran one thread per full row: its softmax warps were latency bound (issue slots
//                     45 % busy, 1 warp per scheduler per tile).  Twice the warps hide each other's dependency stalls.
//   head dim  > 128 : one Q tile per CTA (O alone takes up to 256 TMEM columns); the two S buffers hold CONSECUTIVE key
//                     blocks of the same tile, so S(j+1) = Q K(j+1)^T runs under the softmax of block j and P(j) V(j)
//                     under the softmax of block j+1; two softmax warpgroups (half a row each).
//
// The two threads of a row agree on the reference maximum through shared memory (one named barrier per key block);
// everything else of round 1 is kept: exp2 with the folded scale, a quarter of the exponentials on the FMA pipe, lazy
// (thresholded) O rescale, P written back over S in TMEM as packed 16-bit, P V issued in two halves of 64 keys -- here
// each half is produced by its own warpgroup.  Row maxima use the 3-input max (FMNMX3).
//
// TMEM columns: S slot 0 [0,128)  S slot 1 [128,256)  O0 [256,256+HD)  O1 [256+HD,256+2HD) (second O only for HD <= 128).
#include <cstdlib>
#include <math_constants.h>
#include <type_traits>

#include "common.h"
#include "sm100.cuh"

namespace fa {
namespace {

using namespace sm100;

constexpr int kBM = 128;          // rows per Q tile
constexpr int kBN = 128;          // keys per K/V tile
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.0f;  // log2 domain: P stays below 2^8

struct Fwd2Tma {
  CUtensorMap q, k, v;
};

// EB = bytes per q/k/v element: 2 (bf16 / fp16) or 1 (e4m3, the FP8 forward of csrc/flashmask_v2/flash_api.cu:107-131;
// out stays 16-bit).  Shared-memory tiles are 128 rows of HD * EB bytes in 128-byte swizzled boxes either way.
template <int HD, int QT, int EB = 2> struct F2 {
  static constexpr int kQT = QT;                                         // Q tiles per CTA (2 needs 256 + 2 HD <= 512 TMEM columns)
  static constexpr int kStages = HD <= 128 ? (QT == 2 ? 4 : 5) : (HD <= 192 ? 3 : 2);    // K/V ring depth (tiles of 128 x HD)
  static constexpr int kSmWarps = 8 * kQT;                               // softmax warps: two warpgroups per Q tile
  static constexpr int kThreads = (kSmWarps + 4) * 32;                   // + {issuer, producer, 2 idle}
  // setmaxnreg moves registers inside the CTA's OWN allocation (threads x registers at launch), not the SM's file:
  // 640 threads launch with 96 -> 61440 = 512 x 104 + 128 x 64; 384 threads launch with 168 -> 64512 >= 256 x 200 + 128 x 96
  static constexpr int kSmRegs = QT == 2 ? 104 : 200;                    // register budget of a softmax thread
  static constexpr int kAuxRegs = QT == 2 ? 64 : 96;                     // ... of an issuer / producer / idle thread
  static_assert(kSmWarps * 32 * kSmRegs + 128 * kAuxRegs <= kThreads * (QT == 2 ? 96 : 168), "register pool of the CTA");
  static constexpr int kTileBytes = kBM * HD * EB;
  static constexpr int kOTileBytes = kBM * HD * 2;                       // one staged output tile (16-bit)
  static constexpr int kQOff = 0;
  static constexpr int kKVOff = kQT * kTileBytes;
  // output staging: 16-bit inputs reuse the (dead) Q tile; an e4m3 Q tile is half the size of its output tile
  static constexpr int kOStageOff = EB == 2 ? kQOff : kKVOff + kStages * kTileBytes;
  static constexpr int kXchOff = kKVOff + kStages * kTileBytes + (EB == 2 ? 0 : kQT * kOTileBytes);  // row-max exchange: [2 parities][kQT][2 halves][128]
  static constexpr int kSumOff = kXchOff + 2 * kQT * 2 * kBM * 4;        // row-sum exchange: [kQT][2 halves][128]
  static constexpr int kBarOff = kSumOff + kQT * 2 * kBM * 4;
  static constexpr int kNumBars = 4 + 2 * kStages + 8;
  static constexpr int kTmemSlotOff = kBarOff + kNumBars * 8;
  static constexpr int kTotal = kTmemSlotOff + 16;
  static constexpr int kDynBytes = kTotal + 1024;                        // slack for manual 1024-byte alignment
  static_assert(kDynBytes <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");
  static_assert(256 + kQT * HD <= 512, "TMEM columns");
};

struct Tile2 {
  int bi, hq, hk, m0;
  int sq, sk, q_row0, k_row0, q_b, k_b;
  bool active[2];
  int n_lo[2], n_hi[2];
  int lo, hi;  // union block range
};

template <int kQT> __device__ __forceinline__ bool decode2(const Params &p, int tile, int num_m, Tile2 &ti) {
  const int g = p.h / p.h_k;
  const int per_seg = g * num_m;
  const int seg = tile / per_seg, pos = tile % per_seg;
  ti.bi = seg / p.h_k;
  ti.hk = seg % p.h_k;
  ti.hq = ti.hk * g + pos % g;
  ti.m0 = (num_m - 1 - pos / g) * (kBM * kQT);      // longest (causal) tiles first inside a (batch, kv-head) segment
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
    ti.active[t] = t < kQT && r0 < ti.sq;
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

__device__ __forceinline__ int bc0(int v) { return __shfl_sync(0xffffffffu, v, 0); }
__device__ __forceinline__ void uniformize2(Tile2 &ti) {
  ti.hq = bc0(ti.hq); ti.hk = bc0(ti.hk); ti.m0 = bc0(ti.m0);
  ti.q_row0 = bc0(ti.q_row0); ti.k_row0 = bc0(ti.k_row0); ti.q_b = bc0(ti.q_b); ti.k_b = bc0(ti.k_b);
  ti.lo = bc0(ti.lo); ti.hi = bc0(ti.hi); ti.sk = bc0(ti.sk);
#pragma unroll
  for (int t = 0; t < 2; ++t) { ti.n_lo[t] = bc0(ti.n_lo[t]); ti.n_hi[t] = bc0(ti.n_hi[t]); }
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
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
// 2^x on the FMA pipe (degree-3 polynomial after a Cody-Waite split; max relative error 7.7e-5, below the 16-bit
// rounding of P): see sm100_fwd.cu.  MUFU issues 16 ex2/clk/SM -- as long as the tile's MMAs at d128 -- so a quarter of
// the exponentials go here.  Scalar FFMA with literal coefficients: the immediate form issues at twice the rate of the
// three-register form and, unlike the packed f32x2 version, keeps no coefficient pairs in registers (the softmax thread
// has 112).
__device__ __forceinline__ float exp2_poly(float x) {
  x = fmaxf(x, -125.f);                       // keeps the result a normal number (and maps -inf to ~0)
  const float jm = x + 12582912.f;            // 1.5 * 2^23: the low mantissa bits are round(x)
  const float f = x - (jm - 12582912.f);      // in [-0.5, 0.5]
  float q = fmaf(0.05508868396282196f, f, 0.24260404706001282f);
  q = fmaf(q, f, 0.6932762265205383f);
  q = fmaf(q, f, 0.9999289512634277f);
  return __int_as_float(__float_as_int(q) + (__float_as_int(jm) << 23));
}

template <int HD, bool kBf16, int QT, int EB, bool kDrop>
__global__ void __launch_bounds__(F2<HD, QT, EB>::kThreads, 1)
fwd2_kernel(const __grid_constant__ Params p, const __grid_constant__ Fwd2Tma tm_param, const Fwd2Tma *__restrict__ tm_dev,
            int num_m, int num_tiles) {
  using L = F2<HD, QT, EB>;
  constexpr bool kFp8 = EB == 1;
  const Fwd2Tma &tm = tm_dev ? *tm_dev : tm_param;   // device-patched maps (flat varlen symbols) or the launch's own
  constexpr int kQT = L::kQT, kStages = L::kStages, kSmWarps = L::kSmWarps;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

  auto bar = [&](int i) { return sbase + L::kBarOff + i * 8; };
  constexpr int B_QFULL = 0, B_QEMPTY = 2, B_KVFULL = 4, B_KVEMPTY = 4 + kStages, B_SFULL = 4 + 2 * kStages,
                B_P0 = B_SFULL + 2, B_P1 = B_P0 + 2, B_OFULL = B_P1 + 2;
  uint32_t *tmem_slot = (uint32_t *)(smem + L::kTmemSlotOff);

  if (threadIdx.x == 0) {
    for (int t = 0; t < 2; ++t) {
      mbar_init(bar(B_QFULL + t), 1);
      mbar_init(bar(B_QEMPTY + t), 2 * kBM);
      mbar_init(bar(B_SFULL + t), 1);
      mbar_init(bar(B_P0 + t), kBM);
      mbar_init(bar(B_P1 + t), kBM);
      mbar_init(bar(B_OFULL + t), 1);
    }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar(B_KVFULL + s), 1);
      mbar_init(bar(B_KVEMPTY + s), 1);
    }
    fence_mbar_init();
  }
#ifdef FA_POISON
  // debug variant: fill the operand tiles with NaN patterns so that any byte TMA does not write shows up in the output
  for (int i = threadIdx.x; i < L::kXchOff / 16; i += L::kThreads) sts128(sbase + i * 16, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
  fence_proxy_async_smem();
#endif
  if (warp == kSmWarps) tmem_alloc<512>(smem_u32(tmem_slot));
  if (warp == kSmWarps + 1 && lane == 0) {
    if (tm_dev) { tma_acquire_desc(&tm.q); tma_acquire_desc(&tm.k); tma_acquire_desc(&tm.v); }
    tma_prefetch_desc(&tm.q); tma_prefetch_desc(&tm.k); tma_prefetch_desc(&tm.v);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_raw = *tmem_slot;

  constexpr int kBoxes = HD * EB / 128;        // 128-byte boxes per row
  constexpr int kBoxElems = 128 / EB;          // elements per box row
  constexpr uint32_t kBoxBytes = kBM * 128;    // one box: 128 rows x 128 bytes

  if (warp >= kSmWarps) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(L::kAuxRegs));
    if (warp == kSmWarps + 1) {
      // ======================= TMA producer =======================
      // K/V tiles go through ONE ring in the order the issuer consumes them:
      //   two Q tiles : K0 V0 K1 V1 ...          one Q tile : K0 K1 V0 K2 V1 ... (S runs one block ahead of P V)
      if (lane == 0) {
        uint32_t kv_it = 0, qcnt[2] = {0, 0};
        auto load_kv = [&](const CUtensorMap *m, const Tile2 &ti, int j) {
          const uint32_t st = kv_it % kStages, ph = (kv_it / kStages) & 1;
          mbar_wait(bar(B_KVEMPTY + st), ph ^ 1);
          mbar_expect_tx(bar(B_KVFULL + st), L::kTileBytes);
#pragma unroll
          for (int bx = 0; bx < kBoxes; ++bx)
            tma_load_4d(sbase + L::kKVOff + st * L::kTileBytes + bx * kBoxBytes, m, bar(B_KVFULL + st), bx * kBoxElems, ti.hk,
                        ti.k_row0 + j * kBN, ti.k_b);
          ++kv_it;
        };
        auto load_q = [&](const Tile2 &ti, int t) {
          mbar_wait(bar(B_QEMPTY + t), (qcnt[t] & 1) ^ 1);
          mbar_expect_tx(bar(B_QFULL + t), L::kTileBytes);
#pragma unroll
          for (int bx = 0; bx < kBoxes; ++bx)
            tma_load_4d(sbase + L::kQOff + t * L::kTileBytes + bx * kBoxBytes, &tm.q, bar(B_QFULL + t), bx * kBoxElems, ti.hq,
                        ti.q_row0 + ti.m0 + t * kBM, ti.q_b);
          ++qcnt[t];
        };
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
          Tile2 ti;
          if (!decode2<kQT>(p, tile, num_m, ti)) continue;
          if constexpr (kQT == 2) {
            bool q_loaded[2] = {false, false};
            for (int j = ti.lo; j < ti.hi; ++j) {
              // K/V first: at a tile boundary the Q buffers still belong to the previous tile's epilogue, the ring does not
              load_kv(&tm.k, ti, j);
              load_kv(&tm.v, ti, j);
#pragma unroll
              for (int t = 0; t < 2; ++t) {
                const bool a = j >= ti.n_lo[t] && j < ti.n_hi[t];
                if (a && !q_loaded[t]) { load_q(ti, t); q_loaded[t] = true; }
              }
            }
          } else {
            if (ti.hi <= ti.lo) continue;
            load_kv(&tm.k, ti, ti.lo);
            load_q(ti, 0);
            if (ti.lo + 1 < ti.hi) load_kv(&tm.k, ti, ti.lo + 1);
            for (int j = ti.lo; j < ti.hi; ++j) {
              load_kv(&tm.v, ti, j);
              if (j + 2 < ti.hi) load_kv(&tm.k, ti, j + 2);
            }
          }
        }
      }
    } else if (warp == kSmWarps) {
      // ======================= MMA issuer =======================
      // The whole warp runs the control flow (descriptors stay in uniform registers); one elected lane issues.
      const uint32_t tmem = bc0((int)tmem_raw);
      constexpr uint32_t idesc_s = kFp8 ? make_idesc_e4m3(kBM, kBN, false) : make_idesc(kBf16, kBM, kBN, false, false);
      constexpr uint32_t idesc_o = kFp8 ? make_idesc_e4m3(kBM, HD, true) : make_idesc(kBf16, kBM, HD, false, true);
      uint32_t kv_it = 0, qcnt[2] = {0, 0}, pcnt[2] = {0, 0};
      auto commit = [&](uint32_t b) {
        if (elect_one()) umma_commit(b);
        __syncwarp();
      };
      auto wait_kv = [&]() {        // next ring item is in shared memory; returns its stage
        const uint32_t st = kv_it % kStages;
        mbar_wait(bar(B_KVFULL + st), (kv_it / kStages) & 1);
        ++kv_it;
        return st;
      };
      // S[slot] = Q_t K^T
      auto issue_s = [&](int t, int slot, uint32_t kstage) {
        const uint32_t qa = sbase + L::kQOff + t * L::kTileBytes;
        const uint32_t ka = sbase + L::kKVOff + kstage * L::kTileBytes;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < HD * EB / 32; ++k) {       // 32 bytes of the head dim per instruction (K16 x 16-bit, K32 x 8-bit)
            const uint32_t off = (k / 4) * kBoxBytes + (k % 4) * 32;
            if constexpr (kFp8)
              umma_ss_f8(tmem + slot * kBN, make_sdesc_sw128(qa + off, 16, 1024), make_sdesc_sw128(ka + off, 16, 1024), idesc_s,
                         k > 0);
            else
              umma_ss(tmem + slot * kBN, make_sdesc_sw128(qa + off, 16, 1024), make_sdesc_sw128(ka + off, 16, 1024), idesc_s,
                      k > 0);
          }
          umma_commit(bar(B_SFULL + slot));
        }
        __syncwarp();
      };
      // O_t += P[slot][:, half] V[half]: 64 keys = 4 K16 steps; P is read from TMEM (packed 16-bit, 8 columns per step)
      auto issue_pv_half = [&](int t, int slot, uint32_t vstage, int half, bool accumulate) {
        const uint32_t va = sbase + L::kKVOff + vstage * L::kTileBytes;
        if (elect_one()) {
          constexpr int kKeys = 32 / EB;                  // keys per instruction: P is packed 2 (16-bit) or 4 (e4m3) per column
#pragma unroll
          for (int kk = 0; kk < kBN / 2 / kKeys; ++kk) {
            const int k = half * (kBN / 2 / kKeys) + kk;
            if constexpr (kFp8)
              umma_ts_f8(tmem + 2 * kBN + t * HD, tmem + slot * kBN + k * 8, make_sdesc_sw128(va + k * (kKeys * 128), kBoxBytes, 1024),
                         idesc_o, (accumulate || k > 0) ? 1u : 0u);
            else
              umma_ts(tmem + 2 * kBN + t * HD, tmem + slot * kBN + k * 8, make_sdesc_sw128(va + k * (kKeys * 128), kBoxBytes, 1024),
                      idesc_o, (accumulate || k > 0) ? 1u : 0u);
          }
          // completion of this P V.  One Q tile: signalled per S SLOT, because there S runs two key blocks ahead of P V and
          // a single barrier could be two phases ahead of a waiter (a parity wait cannot tell phase n from phase n + 2)
          if (half == 1) umma_commit(bar(B_OFULL + (kQT == 2 ? t : slot)));
        }
        __syncwarp();
      };
      auto zero_v_tail = [&](const Tile2 &ti, int j, uint32_t vst) {
        if (p.seqused_k != nullptr && j * kBN + kBN > ti.sk) {
          // KV cache: rows at or beyond the used length hold arbitrary bits (possibly NaN) and P = 0 times NaN would
          // poison O (the reference zero-fills the V tail, flash_fwd_kernel.h:346-349)
          const int first = max(ti.sk - j * kBN, 0);
          const uint32_t vt = sbase + L::kKVOff + vst * L::kTileBytes;
          for (int i = first * 8 + lane; i < kBN * 8; i += 32)
#pragma unroll
            for (int bx = 0; bx < kBoxes; ++bx) sts128(vt + bx * kBoxBytes + i * 16, 0u, 0u, 0u, 0u);
          fence_proxy_async_smem();
          __syncwarp();
        }
      };
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        Tile2 ti;
        if (!bc0(decode2<kQT>(p, tile, num_m, ti) ? 1 : 0)) continue;
        uniformize2(ti);
        if (ti.hi <= ti.lo) continue;
        if constexpr (kQT == 2) {
          bool q_ready[2] = {false, false}, first[2] = {true, true};
          auto need_q = [&](int t) {
            if (!q_ready[t]) { mbar_wait(bar(B_QFULL + t), qcnt[t] & 1); ++qcnt[t]; q_ready[t] = true; }
          };
          auto act = [&](int t, int j) { return j >= ti.n_lo[t] && j < ti.n_hi[t]; };
          uint32_t kst = wait_kv();
#pragma unroll
          for (int t = 0; t < 2; ++t)
            if (act(t, ti.lo)) need_q(t);
          tc_fence_after();
#pragma unroll
          for (int t = 0; t < 2; ++t)
            if (act(t, ti.lo)) issue_s(t, t, kst);
          commit(bar(B_KVEMPTY + kst));
          for (int j = ti.lo; j < ti.hi; ++j) {
            const bool has_next = j + 1 < ti.hi;
            const uint32_t vst = wait_kv();
            zero_v_tail(ti, j, vst);
            uint32_t nkst = 0;
            bool next_waited = false;
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              if (act(t, j)) {
                mbar_wait(bar(B_P0 + t), pcnt[t] & 1);
                tc_fence_after();
                issue_pv_half(t, t, vst, 0, !first[t]);
                mbar_wait(bar(B_P1 + t), pcnt[t] & 1);
                ++pcnt[t];
                tc_fence_after();
                issue_pv_half(t, t, vst, 1, !first[t]);
                first[t] = false;
              }
              if (has_next && act(t, j + 1)) {
                if (!next_waited) { nkst = wait_kv(); next_waited = true; }
                need_q(t);
                tc_fence_after();
                issue_s(t, t, nkst);
              }
            }
            commit(bar(B_KVEMPTY + vst));
            if (has_next) {
              if (!next_waited) nkst = wait_kv();     // block j+1 is in the union range: active for at least one tile
              commit(bar(B_KVEMPTY + nkst));
            }
          }
        } else {
          // one Q tile, S double-buffered over consecutive key blocks; g counts this CTA's blocks across tiles:
          // block g uses S slot g & 1, its barriers are in their (g >> 1)-th phase
          mbar_wait(bar(B_QFULL), qcnt[0] & 1);
          ++qcnt[0];
          const int nblk = ti.hi - ti.lo;
          uint32_t &g = pcnt[0];
          {
            const uint32_t kst = wait_kv();
            tc_fence_after();
            issue_s(0, g & 1, kst);
            commit(bar(B_KVEMPTY + kst));
          }
          if (nblk > 1) {
            const uint32_t kst = wait_kv();
            tc_fence_after();
            issue_s(0, (g + 1) & 1, kst);
            commit(bar(B_KVEMPTY + kst));
          }
          for (int i = 0; i < nblk; ++i) {
            const uint32_t slot = g & 1, par = (g >> 1) & 1;
            const uint32_t vst = wait_kv();
            zero_v_tail(ti, ti.lo + i, vst);
            mbar_wait(bar(B_P0 + slot), par);
            tc_fence_after();
            issue_pv_half(0, slot, vst, 0, i > 0);
            mbar_wait(bar(B_P1 + slot), par);
            tc_fence_after();
            issue_pv_half(0, slot, vst, 1, i > 0);
            commit(bar(B_KVEMPTY + vst));
            if (i + 2 < nblk) {
              const uint32_t kst = wait_kv();
              tc_fence_after();
              issue_s(0, slot, kst);       // in order behind P V(i): the slot's P has been consumed
              commit(bar(B_KVEMPTY + kst));
            }
            ++g;
          }
        }
      }
    }
  } else {
    // ======================= softmax warpgroups =======================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(L::kSmRegs));
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    const int wg = warp / 4;
    const int t = kQT == 2 ? wg / 2 : 0;        // Q tile of this warpgroup
    const int hf = wg % 2;                      // which 64 columns of S / which half of the O columns
    const int r_in_tile = (warp % 4) * 32 + lane;
    const int tid2 = hf * kBM + r_in_tile;      // index among the 256 threads of the tile
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const uint32_t tmem = tmem_raw;
    constexpr int kHalfCols = HD / 2;           // O columns per thread
    const uint32_t tO = tmem + lane_base + 2 * kBN + t * HD + hf * kHalfCols;
    float scale_nat = p.softmax_scale;
    float scale_log2 = scale_nat * kLog2e;
    float out_scale = 1.f;
    float *xch = (float *)(smem + L::kXchOff);  // [parity][t][hf][row]
    float *xsum = (float *)(smem + L::kSumOff); // [t][hf][row]
    uint32_t g = 0;       // key blocks processed by this thread so far (all tiles)
    uint32_t ocnt = 0;    // P V commits on this tile's O before the current work tile
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      Tile2 ti;
      if (!decode2<kQT>(p, tile, num_m, ti)) continue;
      if (!ti.active[t]) continue;
      const int row = ti.m0 + t * kBM + r_in_tile;
      T *optr = (T *)p.o.ptr + (int64_t)ti.q_b * p.o.batch_stride + (int64_t)(ti.q_row0) * p.o.row_stride +
                (int64_t)ti.hq * p.o.head_stride;
      float *lse_ptr = p.softmax_lse + ((int64_t)ti.bi * p.h + ti.hq) * p.seqlen_q;
      const MaskCtx mc = make_mask_ctx(p, ti.bi, ti.hq, ti.sq, ti.sk);
      const int shiftc = mc.bottom_right ? mc.shift : 0;
      const int r0 = ti.m0 + t * kBM;
      if constexpr (kFp8) {
        // per (batch, kv head) dequantisation factors: logits by q_descale * k_descale, out by v_descale
        // (csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:1612-1614,1868-1869)
        const float qd = p.q_descale ? p.q_descale[ti.bi * p.q_descale_batch_stride + ti.hk * p.q_descale_head_stride] : 1.f;
        const float kd = p.k_descale ? p.k_descale[ti.bi * p.k_descale_batch_stride + ti.hk * p.k_descale_head_stride] : 1.f;
        out_scale = p.v_descale ? p.v_descale[ti.bi * p.v_descale_batch_stride + ti.hk * p.v_descale_head_stride] : 1.f;
        scale_nat = p.softmax_scale * qd * kd;
        scale_log2 = scale_nat * kLog2e;
      }
      float m_ref = -CUDART_INF_F, l = 0.f;
      // dropout (kDrop): the mask of common.h's dropout_keep -- one Philox call per 16 consecutive columns of a row; this
      // thread's 64 columns of a key block are four calls.  Row sum before dropout, 1 / p_keep folded into the
      // normalisation (flash_fwd_kernel.h:399-409, softmax.h:381-447), as in sm100_fwd.cu.
      DropCtx dctx;
      if constexpr (kDrop) { dctx = make_drop_ctx(p, ti.bi, ti.hq, row); out_scale = 1.f / (1.f - p.p_dropout); }
      const int nb_total = ti.n_hi[t] - ti.n_lo[t];
      for (int nb = 0; nb < nb_total; ++nb, ++g) {
        const int j = ti.n_lo[t] + nb;
        const uint32_t slot = kQT == 2 ? (uint32_t)t : (g & 1u);
        const uint32_t par = kQT == 2 ? (g & 1u) : ((g >> 1) & 1u);
        const uint32_t tS = tmem + lane_base + slot * kBN + hf * 64;
        const uint32_t tP = tmem + lane_base + slot * kBN + hf * (kFp8 ? 16 : 32);   // this thread's 64 weights: 32 / 16 columns
        mbar_wait(bar(B_SFULL + slot), par);
        tc_fence_after();
        uint32_t s[2][32];
        tmem_ld32(tS, s[0]);
        tmem_ld32(tS + 32, s[1]);
        tmem_wait_ld();
        const int c0 = j * kBN + hf * 64;        // first key column of this thread
        const bool full = (j * kBN + kBN <= ti.sk) && (!mc.causal || j * kBN + kBN - 1 <= r0 + shiftc) &&
                          (mc.wr < 0 || j * kBN + kBN - 1 <= r0 + mc.shift + mc.wr) &&
                          (mc.wl < 0 || j * kBN >= r0 + kBM - 1 + mc.shift - mc.wl);
        if (!full) {
          // boundary blocks: seqlen tail / causal / window -- the visible columns of a row form an interval
          int vlo, vhi;
          visible_cols(mc, row, vlo, vhi);
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const uint32_t mb = outside_bits(vlo, vhi, c0 + c * 32);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if ((mb >> i) & 1u) s[c][i] = __float_as_uint(-CUDART_INF_F);
          }
        }
        float mx0 = __uint_as_float(s[0][0]), mx1 = __uint_as_float(s[1][0]);
#pragma unroll
        for (int i = 1; i + 1 < 32; i += 2) {
          mx0 = fmax3(mx0, __uint_as_float(s[0][i]), __uint_as_float(s[0][i + 1]));
          mx1 = fmax3(mx1, __uint_as_float(s[1][i]), __uint_as_float(s[1][i + 1]));
        }
        float mx = fmax3(mx0, mx1, fmaxf(__uint_as_float(s[0][31]), __uint_as_float(s[1][31])));
        // the two threads of a row agree on the block maximum (double-buffered slot: a thread can be a whole block ahead
        // of its partner's read only after the next barrier)
        {
          float *slotp = xch + ((g & 1u) * kQT + t) * (2 * kBM);
          slotp[hf * kBM + r_in_tile] = mx;
          named_bar_sync(1 + t, 2 * kBM);
          mx = fmaxf(mx, slotp[(hf ^ 1) * kBM + r_in_tile]);
        }
        if (nb == 0) {
          m_ref = mx;
        } else {
          const bool need = (mx > m_ref) && ((mx - m_ref) * scale_log2 > kRescaleThreshold || m_ref == -CUDART_INF_F);
          if (__any_sync(0xffffffffu, need)) {
            // the previous P V of this tile must have landed in O before it is rescaled
            if constexpr (kQT == 2) mbar_wait(bar(B_OFULL + t), (ocnt + nb - 1) & 1);
            else mbar_wait(bar(B_OFULL + ((g - 1) & 1u)), ((g - 1) >> 1) & 1u);   // block g - 1: slot (g-1)&1, its (g-1)>>1-th use
            tc_fence_after();
            const float m_new = fmaxf(m_ref, mx);
            const float corr = (m_new == -CUDART_INF_F) ? 1.f : ex2((m_ref - m_new) * scale_log2);
            l *= corr;
            m_ref = m_new;
            // rare path (the running maximum grew by more than 2^8): rolled, 16 columns at a time, so that it does not
            // push the 64 live S values of the hot path out of the register file
#pragma unroll 1
            for (int c = 0; c < kHalfCols / 16; ++c) {
              uint32_t o[16];
              tmem_ld16(tO + c * 16, o);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
              tmem_st16(tO + c * 16, o);
            }
          }
        }
        const float msc = (m_ref == -CUDART_INF_F) ? 0.f : m_ref * scale_log2;
        const uint64_t sc2 = pk2(scale_log2, scale_log2), nm2 = pk2(-msc, -msc);
        uint64_t suma = pk2(0.f, 0.f), sumb = pk2(0.f, 0.f);
        uint32_t pk8[16];      // e4m3: the 64 weights of this thread, four per column
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t pk[16];
          uint4 km = make_uint4(0, 0, 0, 0);
          (void)km;
          float pprev0 = 0.f, pprev1 = 0.f;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint64_t x2 = ffma2(pk2(__uint_as_float(s[c][2 * i]), __uint_as_float(s[c][2 * i + 1])), sc2, nm2);
            float p0, p1;
            float x0, x1;
            unpk2(x2, x0, x1);
            if (i % 4 == 3) {          // every fourth pair: polynomial on the FMA pipe
              p0 = exp2_poly(x0); p1 = exp2_poly(x1);
            } else {
              p0 = ex2(x0); p1 = ex2(x1);
            }
            if (i & 1) sumb = fadd2(sumb, pk2(p0, p1)); else suma = fadd2(suma, pk2(p0, p1));
            if constexpr (kFp8) {
              if (i & 1) pk8[c * 8 + i / 2] = pack4_e4m3(pprev0, pprev1, p0, p1);
              else { pprev0 = p0; pprev1 = p1; }
            } else {
              pk[i] = pack2<kBf16>(p0, p1);
              if constexpr (kDrop) {   // keep mask on the packed pair: byte-wise compare, widened by a byte permute
                if (i % 8 == 0) {
                  const uint4 r4 = philox4x32_7(dctx.seed, dctx.row_base + (uint32_t)((c0 + c * 32 + 2 * i) >> 4), dctx.offset);
                  const uint32_t thr4 = dctx.threshold * 0x01010101u;
                  km = make_uint4(__vcmpleu4(r4.x, thr4), __vcmpleu4(r4.y, thr4), __vcmpleu4(r4.z, thr4), __vcmpleu4(r4.w, thr4));
                }
                const int e = (2 * i) & 15;
                const uint32_t w = e < 4 ? km.x : e < 8 ? km.y : e < 12 ? km.z : km.w;
                pk[i] &= __byte_perm(w, 0u, (e & 3) == 0 ? 0x1100u : 0x3322u);
              }
            }
          }
          if constexpr (!kFp8) tmem_st16(tP + c * 16, pk);
        }
        if constexpr (kFp8) tmem_st16(tP, pk8);
        {
          float a0, a1;
          unpk2(fadd2(suma, sumb), a0, a1);
          l += a0 + a1;
        }
        tmem_wait_st();
        tc_fence_before();
        mbar_arrive(bar((hf == 0 ? B_P0 : B_P1) + slot));
      }
      if (nb_total == 0) {
        // no visible key block for this Q tile: out = 0, lse = +-inf (no barrier was touched)
        if (hf == 0 && row < ti.sq) {
          T *orow = optr + (int64_t)row * p.o.row_stride;
          for (int e = 0; e < p.d; e += 8) *reinterpret_cast<uint4 *>(orow + e) = make_uint4(0, 0, 0, 0);
          lse_ptr[row] = p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F;
        }
        continue;
      }
      // ---------------- epilogue ----------------
      xsum[(t * 2 + hf) * kBM + r_in_tile] = l;
      if constexpr (kQT == 2) {
        mbar_wait(bar(B_OFULL + t), (ocnt + nb_total - 1) & 1);
      } else {   // g already counts this tile's blocks: the last two P V are blocks g - 2 and g - 1 (P V complete in order)
        if (nb_total > 1) mbar_wait(bar(B_OFULL + ((g - 2) & 1u)), ((g - 2) >> 1) & 1u);
        mbar_wait(bar(B_OFULL + ((g - 1) & 1u)), ((g - 1) >> 1) & 1u);
      }
      ocnt += nb_total;
      tc_fence_after();
      named_bar_sync(1 + t, 2 * kBM);
      l += xsum[(t * 2 + (hf ^ 1)) * kBM + r_in_tile];
      // A row with at least one visible key has l >= ~1 (its maximum contributes 2^0).  Masked elements that went
      // through the polynomial exp2 contribute 2^-125 each instead of 0, so "no visible key" is l below a tiny bound.
      const bool nonempty = l > 1e-30f;
      const float inv = nonempty ? out_scale / l : 0.f;
      uint8_t *stage = smem + L::kOStageOff + t * L::kOTileBytes;  // 16-bit inputs: Q_t, dead once this tile's S MMAs have completed
      constexpr int kChunks = HD / 8;                        // 16-byte chunks per row
#pragma unroll
      for (int c = 0; c < kHalfCols / 32; ++c) {
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
          const int chunk = hf * (kHalfCols / 8) + c * 4 + q4;
          *reinterpret_cast<uint4 *>(stage + r_in_tile * (HD * 2) + ((chunk ^ (r_in_tile & 7)) * 16)) = v;
        }
      }
      tc_fence_before();
      if (hf == 0 && row < ti.sq)
        lse_ptr[row] = nonempty ? (m_ref * scale_nat + __logf(l))
                                 : (p.lse_empty_is_neg_inf ? -CUDART_INF_F : CUDART_INF_F);
      named_bar_sync(1 + t, 2 * kBM);
#pragma unroll 4
      for (int i = tid2; i < kBM * kChunks; i += 2 * kBM) {
        const int rr = i / kChunks, chunk = i % kChunks;
        const int grow = ti.m0 + t * kBM + rr;
        if (grow < ti.sq && chunk * 8 < p.d) {
          const uint4 v = *reinterpret_cast<const uint4 *>(stage + rr * (HD * 2) + ((chunk ^ (rr & 7)) * 16));
          *reinterpret_cast<uint4 *>(optr + (int64_t)grow * p.o.row_stride + chunk * 8) = v;
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(bar(B_QEMPTY + t));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kSmWarps) tmem_dealloc<512>(tmem_raw);
}

void sanitize2(TmaDesc4 &d) {
  // a dimension of extent 1 never advances: give it a legal, monotonic stride
  for (int i = 1; i < 4; ++i)
    if (d.dims[i] == 1) d.strides[i - 1] = (i == 1 ? d.dims[0] * d.elem_bytes : d.strides[i - 2] * d.dims[i - 1] + 15) / 16 * 16;
}

TmaDesc4 desc2(const fa_b200_tensor &t, int d, int heads, int64_t rows, int batch, int box_rows, int eb = 2) {
  TmaDesc4 x;
  x.ptr = t.ptr;
  x.elem_bytes = eb;
  x.dims[0] = (uint64_t)d; x.dims[1] = (uint64_t)heads; x.dims[2] = (uint64_t)rows; x.dims[3] = (uint64_t)batch;
  x.strides[0] = (uint64_t)t.head_stride * eb; x.strides[1] = (uint64_t)t.row_stride * eb;
  x.strides[2] = (uint64_t)t.batch_stride * eb;
  x.box[0] = 128 / eb; x.box[1] = 1; x.box[2] = (uint32_t)box_rows; x.box[3] = 1;
  sanitize2(x);
  return x;
}

template <int HD, bool kBf16, int QT = (HD <= 128 ? 2 : 1), int EB = 2, bool kDrop = false> void launch_fwd2(const Params &p, cudaStream_t s) {
  using L = F2<HD, QT, EB>;
  const bool packed_q = p.cu_seqlens_q && !p.varlen_padded_input;
  const bool packed_k = p.cu_seqlens_k && !p.varlen_padded_input;
  // packed tensors without host-known totals (the flat varlen symbols): the row extents are patched on the device
  const bool dev_rows = (packed_q && p.total_q < 0) || (packed_k && p.total_k < 0);
  const int64_t rows_q = packed_q ? (p.total_q >= 0 ? p.total_q : (int64_t)p.b * p.seqlen_q) : p.seqlen_q;
  const int64_t rows_k = packed_k ? (p.total_k >= 0 ? p.total_k : (int64_t)p.b * p.seqlen_k) : p.seqlen_k;
  Fwd2Tma tm;
  // columns at or beyond p.d (head dims that are not a multiple of 64) are zero-filled by the TMA unit
  tm.q = make_tma_4d(desc2(p.q, p.d, p.h, rows_q, packed_q ? 1 : p.b, kBM, EB));
  tm.k = make_tma_4d(desc2(p.k, p.d, p.h_k, rows_k, packed_k ? 1 : p.b, kBN, EB));
  tm.v = make_tma_4d(desc2(p.v, p.d, p.h_k, rows_k, packed_k ? 1 : p.b, kBN, EB));
  Fwd2Tma *tm_dev = nullptr;
  if (dev_rows) {
    const int k_like[3] = {0, 1, 1};
    tm_dev = reinterpret_cast<Fwd2Tma *>(patch_tma_rows(&tm.q, k_like, 3, p.cu_seqlens_q, p.cu_seqlens_k, p.b, s));
  }
  const int num_m = ceil_div(p.seqlen_q, kBM * L::kQT);
  const int num_tiles = p.b * p.h * num_m;
  auto kern = fwd2_kernel<HD, kBf16, QT, EB, kDrop>;
  static std::atomic<bool> smem_set[64];
  ensure_dynamic_smem(kern, L::kDynBytes, smem_set);
  const int grid = std::min(num_tiles, forward_grid_sms());
  profile_mark(0, true, s);
  kern<<<grid, L::kThreads, L::kDynBytes, s>>>(p, tm, tm_dev, num_m, num_tiles);
  profile_mark(0, false, s);
  const cudaError_t launch_err = cudaGetLastError();
  if (tm_dev) cudaFreeAsync(tm_dev, s);
  FA_CUDA(launch_err);
  note_launch(2);
}

}  // namespace

// The split-row forward serves calls without FlashMask / dropout / dense mask / softcap (those keep the one-thread-per-row
// kernel of sm100_fwd.cu, head dims <= 128 only).
bool sm100_fwd2_eligible(const Params &p) {
  // dropout: instantiated for head dims above 128 only (below, sm100_fwd.cu has its own kDrop instantiations)
  return p.lt_start == nullptr && (p.p_dropout == 0.f || (p.d > 128 && !p.is_e4m3)) && p.attn_mask == nullptr && p.softcap == 0.f;
}

// FP8 (e4m3) forward: same kernel with byte operands, tcgen05.mma kind::f8f6f4 (twice the 16-bit rate), P rounded to
// e4m3 for the P V product as the reference's sm90 FP8 path does, output in bf16.
void sm100_fwd2_e4m3(const Params &p, cudaStream_t s) {
  if (p.d <= 128) launch_fwd2<128, true, 2, 1>(p, s);
  else launch_fwd2<256, true, 1, 1>(p, s);
}

void sm100_fwd2(const Params &p, cudaStream_t s) {
  // experiment switch: FA_B200_FWD2_QT=1 runs head dims <= 128 with ONE Q tile per CTA (S double-buffered over key blocks)
  static const bool one_tile = [] { const char *e = getenv("FA_B200_FWD2_QT"); return e && e[0] == '1'; }();
  if (one_tile && p.d <= 128) {
    if (p.d <= 64) { if (p.is_bf16) launch_fwd2<64, true, 1>(p, s); else launch_fwd2<64, false, 1>(p, s); }
    else { if (p.is_bf16) launch_fwd2<128, true, 1>(p, s); else launch_fwd2<128, false, 1>(p, s); }
    return;
  }
  if (p.d <= 64) { if (p.is_bf16) launch_fwd2<64, true>(p, s); else launch_fwd2<64, false>(p, s); }
  else if (p.d <= 128) { if (p.is_bf16) launch_fwd2<128, true>(p, s); else launch_fwd2<128, false>(p, s); }
  else if (p.p_dropout > 0.f) {   // head dims above 128 with dropout
    if (p.d <= 192) { if (p.is_bf16) launch_fwd2<192, true, 1, 2, true>(p, s); else launch_fwd2<192, false, 1, 2, true>(p, s); }
    else { if (p.is_bf16) launch_fwd2<256, true, 1, 2, true>(p, s); else launch_fwd2<256, false, 1, 2, true>(p, s); }
  }
  else if (p.d <= 192) { if (p.is_bf16) launch_fwd2<192, true>(p, s); else launch_fwd2<192, false>(p, s); }
  else { if (p.is_bf16) launch_fwd2<256, true>(p, s); else launch_fwd2<256, false>(p, s); }
}

}  // namespace fa
