// pair_bench.cu -- does a CTA pair (tcgen05.mma.cta_group::2, M = 256 over two SMs) lift the shared-memory limit of the SS
// form?  tmem_bench.cu measured a cta_group::1 SS product of 128^3 at ~675 cycles (64 KB of operands through a 128 B/clk
// port) against ~544 for TS.  With a pair each SM reads its own 128 rows of A but only HALF of B (64 of the 128 N rows).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I xhy-flash-attention_b200/csrc -I include tools/pair_bench.cu -o tools/_bin/pair_bench
// Also checks the numbers: A[m][k] = m % 7 + 1, B[n][k] = n % 5 + 1 (n global over the pair) -> D[m][n] = 128 a b.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "sm100.cuh"

using namespace fa::sm100;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void umma_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_ts_pair(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}
// arrive on the barrier at the same shared-memory offset in every CTA of the mask when the MMAs issued so far are done
__device__ __forceinline__ void umma_commit_pair(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"(mask) : "memory");
}

// write element (row, col) of a K-major 128-byte-swizzled tile of boxes [rows x 64 elements]
__device__ __forceinline__ void put(uint8_t *tile, int rows, int row, int col, float v) {
  const int box = col / 64, c = col % 64;
  const int chunk = (c / 8) ^ (row & 7);
  *reinterpret_cast<__nv_bfloat16 *>(tile + box * rows * 128 + row * 128 + chunk * 16 + (c % 8) * 2) = __float2bfloat16(v);
}

template <int mode> __global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(160, 1) bench(int iters, long long *out, int *bad, int pattern) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t rank = cluster_ctarank();
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long mbar[4];   // batch it commits to mbar[it % 4]: one outstanding phase per barrier
  uint8_t *A = smem, *B = smem + 32768;          // A: 128 x 128 (32 KB); B: 128 x 128 for one CTA, 64 x 128 per CTA of a pair
  constexpr bool pair = mode >= 4;
  const int brows = pair ? 64 : 128;
  for (int i = threadIdx.x; i < 128 * 128; i += blockDim.x) {
    const int r = i / 128, c = i % 128;
    put(A, 128, r, c, (float)((r + 128 * (pair ? rank : 0)) % 7 + 1));
  }
  for (int i = threadIdx.x; i < brows * 128; i += blockDim.x) {
    const int r = i / 128, c = i % 128;
    put(B, brows, r, c, (float)((r + (pair ? 64 * rank : 0)) % 5 + 1));
  }
  // operand data patterns (does the tensor pipe slow down with the data it chews on?): 1 = one constant, 2 = zeros, 3 = random bits
  if (pattern) {
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) {
      uint32_t x = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
      x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
      const uint32_t rnd = (x & 0x3fff3fffu) | 0x30003000u;          // finite bf16 pairs of mixed sign-free magnitude
      ((uint32_t *)smem)[i] = pattern == 1 ? 0x3c003c00u : pattern == 2 ? 0u : (rnd ^ ((x << 3) & 0x80008000u));
    }
  }
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&mbar[i]), 1); fence_mbar_init(); }
  fence_proxy_async_smem();
  if (warp == 4) {
    if (pair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      tmem_alloc<512>(smem_u32(&tmem_slot));
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  long long t = 0;
  if (warp == 4 && (!pair || rank == 0)) {
    constexpr uint32_t idesc = make_idesc(true, pair ? 256 : 128, 128, false, false);
    const uint32_t bbox = brows * 128;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (it >= 4) mbar_wait(smem_u32(&mbar[it % 4]), ((it - 4) / 4) & 1);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t da = make_sdesc_sw128(sbase + (k / 4) * 16384 + (k % 4) * 32, 16, 1024);
          const uint64_t db = make_sdesc_sw128(sbase + 32768 + (k / 4) * bbox + (k % 4) * 32, 16, 1024);
          const uint32_t d = tmem + (it & 1) * 128;
          if constexpr (mode == 0) umma_ss(d, da, db, idesc, k > 0);
          else if constexpr (mode == 1) umma_ts(d, tmem + 256 + k * 8, db, idesc, k > 0);
          else if constexpr (mode == 4) umma_ss_pair(d, da, db, idesc, k > 0);
          else umma_ts_pair(d, tmem + 256 + k * 8, db, idesc, k > 0);
        }
        if (pair) umma_commit_pair(smem_u32(&mbar[it % 4]), 1);
        else umma_commit(smem_u32(&mbar[it % 4]));
      }
      __syncwarp();
    }
    for (int it = iters - 4; it < iters; ++it) mbar_wait(smem_u32(&mbar[it % 4]), (it / 4) & 1);
    t = clock64() - t0;
    if (lane == 0 && blockIdx.x == 0) out[mode] = t;
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  // check D (SS modes only): last product went to slot (iters - 1) & 1
  if (warp < 4 && (mode == 0 || mode == 4) && pattern == 0) {
    uint32_t v[32];
    const uint32_t base = tmem + ((uint32_t)(warp * 32) << 16) + ((iters - 1) & 1) * 128;
    int nbad = 0;
    for (int c = 0; c < 4; ++c) {
      tmem_ld32(base + c * 32, v);
      tmem_wait_ld();
      const int m = warp * 32 + lane + (pair ? 128 * rank : 0);
      for (int i = 0; i < 32; ++i) {
        const int n = c * 32 + i;
        const float want = 128.f * (m % 7 + 1) * (n % 5 + 1);
        if (__uint_as_float(v[i]) != want) ++nbad;
      }
    }
    if (nbad) atomicAdd(bad + mode, nbad);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 4) {
    if (pair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    else tmem_dealloc<512>(tmem);
  }
}

// Mixed use: TMEM allocated with cta_group::2; the leader alternates pair MMAs (M 256, N 128, into columns [0,128)) with
// its own cta_group::1 MMAs (M 128, N 64, into columns [128,192)) while the peer CTA issues cta_group::1 MMAs of its own.
// Is that legal on the hardware (results checked) and what does it cost?
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(160, 1) mix(int iters, long long *out, int *bad) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t rank = cluster_ctarank();
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long mbar[4], mbar1[4];
  uint8_t *A = smem, *B = smem + 32768;          // A: 128 x 128; B: 64 x 128 (this CTA's half of the pair's N = 128)
  for (int i = threadIdx.x; i < 128 * 128; i += blockDim.x) put(A, 128, i / 128, i % 128, (float)((i / 128 + 128 * rank) % 7 + 1));
  for (int i = threadIdx.x; i < 64 * 128; i += blockDim.x) put(B, 64, i / 128, i % 128, (float)((i / 128 + 64 * rank) % 5 + 1));
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) { mbar_init(smem_u32(&mbar[i]), 1); mbar_init(smem_u32(&mbar1[i]), 1); } fence_mbar_init(); }
  fence_proxy_async_smem();
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before(); __syncthreads(); cluster_sync_all(); tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 4) {
    constexpr uint32_t idesc2 = make_idesc(true, 256, 128, false, false), idesc1 = make_idesc(true, 128, 64, false, false);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (it >= 4) { if (rank == 0) mbar_wait(smem_u32(&mbar[it % 4]), ((it - 4) / 4) & 1); mbar_wait(smem_u32(&mbar1[it % 4]), ((it - 4) / 4) & 1); }
      if (elect_one()) {
        if (rank == 0) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ss_pair(tmem, make_sdesc_sw128(sbase + (k / 4) * 16384 + (k % 4) * 32, 16, 1024),
                         make_sdesc_sw128(sbase + 32768 + (k / 4) * 8192 + (k % 4) * 32, 16, 1024), idesc2, k > 0);
          umma_commit_pair(smem_u32(&mbar[it % 4]), 1);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_ss(tmem + 128, make_sdesc_sw128(sbase + (k / 4) * 16384 + (k % 4) * 32, 16, 1024),
                  make_sdesc_sw128(sbase + 32768 + (k / 4) * 8192 + (k % 4) * 32, 16, 1024), idesc1, k > 0);
        umma_commit(smem_u32(&mbar1[it % 4]));
      }
      __syncwarp();
    }
    for (int it = iters - 4; it < iters; ++it) { if (rank == 0) mbar_wait(smem_u32(&mbar[it % 4]), (it / 4) & 1); mbar_wait(smem_u32(&mbar1[it % 4]), (it / 4) & 1); }
    if (lane == 0 && blockIdx.x < 2) out[8 + rank] = clock64() - t0;
  }
  tc_fence_before(); __syncthreads(); cluster_sync_all(); tc_fence_after();
  if (warp < 4) {
    uint32_t v[32];
    const uint32_t base = tmem + ((uint32_t)(warp * 32) << 16);
    int nbad = 0;
    const int m = warp * 32 + lane + 128 * rank;
    for (int c = 0; c < 6; ++c) {                 // columns [0,128): pair product; [128,192): this CTA's own product
      tmem_ld32(base + c * 32, v);
      tmem_wait_ld();
      for (int i = 0; i < 32; ++i) {
        const int n = c < 4 ? c * 32 + i : (c - 4) * 32 + i + 64 * rank;
        if (__uint_as_float(v[i]) != 128.f * (m % 7 + 1) * (n % 5 + 1)) ++nbad;
      }
    }
    if (nbad) atomicAdd(bad + 8, nbad);
  }
  tc_fence_before(); __syncthreads(); cluster_sync_all();
  if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

int main() {
  long long *out;
  int *bad;
  cudaMalloc(&out, 64 * 8);
  cudaMalloc(&bad, 64 * 4);
  cudaMemset(out, 0, 64 * 8);
  cudaMemset(bad, 0, 64 * 4);
  auto launch = [&](int mode, int iters, int pattern) {
    auto go = [&](auto kern) {
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 2048);
      kern<<<148, 160, 65536 + 1024>>>(iters, out, bad, pattern);
    };
    if (mode == 0) go(bench<0>); else if (mode == 1) go(bench<1>); else if (mode == 4) go(bench<4>); else go(bench<5>);
  };
  const int iters = 2000;
  const char *names[8] = {"cta_group::1 SS", "cta_group::1 TS", "", "", "cta_group::2 SS (M256, half of B per SM)", "cta_group::2 TS", "", ""};
  const char *pat[4] = {"small integers", "one constant", "zeros", "random bits"};
  for (int pattern : {0, 3})
  for (int mode : {0, 1, 4, 5}) {
    for (int rep = 0; rep < 2; ++rep) launch(mode, iters, pattern);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mode %d error: %s\n", mode, cudaGetErrorString(e)); return 1; }
    long long h[8];
    int hb[8];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    cudaMemcpy(hb, bad, sizeof(hb), cudaMemcpyDeviceToHost);
    printf("%-16s %-45s %.0f cycles per 128x128x128 per SM   wrong elements: %d\n", pat[pattern], names[mode], (double)h[mode] / iters, hb[mode]);
  }
  cudaFuncSetAttribute(mix, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 2048);
  for (int rep = 0; rep < 2; ++rep) mix<<<148, 160, 65536 + 1024>>>(iters, out, bad);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("mixed cta_group error: %s\n", cudaGetErrorString(e)); return 1; }
  long long h[16];
  int hb[16];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  cudaMemcpy(hb, bad, sizeof(hb), cudaMemcpyDeviceToHost);
  printf("mixed: pair SS 128^3 + own cta_group::1 SS (M128 N64 K128) per iteration: leader %.0f cycles, peer (cta_group::1 only) %.0f cycles; wrong elements: %d\n",
         (double)h[8] / iters, (double)h[9] / iters, hb[8]);
  return 0;
}
