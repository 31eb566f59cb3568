// tmem_bench.cu -- microbenchmark behind DESIGN.md's forward-kernel model: how fast is tcgen05.ld (TMEM -> registers),
// how fast is tcgen05.mma (SS, M128 N128 K16, bf16), and do they slow each other down when they run at the same time?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I xhy-flash-attention_b200/csrc -I include tools/tmem_bench.cu -o gpurun_out/tmem_bench
// One CTA per SM.  Warps 0-3 (or 0-7) load a 128 x 128 fp32 tile from TMEM per iteration (4 x tcgen05.ld.32x32b.x32 per
// thread, the forward softmax's access); warp 8 issues batches of 8 MMAs (one 128x128x128 product) back to back.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "sm100.cuh"

using namespace fa::sm100;

__global__ void __launch_bounds__(320, 1) bench(int mode, int iters, int ld_warps, long long *out, int depth) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long mbar;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0x3c003c00u;   // finite operands
  if (threadIdx.x == 0) { mbar_init(smem_u32(&mbar), 1); fence_mbar_init(); }
  fence_proxy_async_smem();
  if (warp == 8) tmem_alloc<512>(smem_u32(&tmem_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  long long t_ld = 0, t_mma = 0;
  const bool do_ld = mode & 1, do_mma = (mode & 2) != 0 || mode == 8;
  if (warp < ld_warps && do_ld) {
    const uint32_t lane_base = (uint32_t)((warp % 4) * 32) << 16;
    const uint32_t col0 = (warp / 4) * 128;           // warps 4-7 read a second tile
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      uint32_t s[4][32];
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32(tmem + lane_base + col0 + c * 32, s[c]);
      tmem_wait_ld();
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; i += 8) acc ^= s[c][i];
    }
    t_ld = clock64() - t0;
    if (acc == 0x12345678u) out[1000] = acc;
  }
  if (warp == 8 && mode == 8) {
    // the forward kernel's MMA stream with nothing to wait for: per iteration, for each of the two Q tiles,
    // O_t += P_t V (TS, P read from S slot t, two halves of 4 K16 steps) then S_t = Q_t K^T (SS into slot t)
    constexpr uint32_t idesc_s = make_idesc(true, 128, 128, false, false);
    constexpr uint32_t idesc_o = make_idesc(true, 128, 128, false, true);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
#pragma unroll
        for (int t = 0; t < 2; ++t) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_ts(tmem + 256 + t * 128, tmem + t * 128 + k * 8, make_sdesc_sw128(sbase + k * 2048, 16384, 1024), idesc_o, 1u);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t off = (k / 4) * 16384 + (k % 4) * 32;
            umma_ss(tmem + t * 128, make_sdesc_sw128(sbase + 32768 + off, 16, 1024), make_sdesc_sw128(sbase + off, 16, 1024), idesc_s, k > 0);
          }
        }
        umma_commit(smem_u32(&mbar));
      }
      __syncwarp();
      if (it >= depth) mbar_wait(smem_u32(&mbar), (it - depth) & 1);
    }
    mbar_wait(smem_u32(&mbar), (iters - 1) & 1);
    t_mma = clock64() - t0;
  }
  if (warp == 8 && do_mma && mode != 8) {
    constexpr uint32_t idesc = make_idesc(true, 128, 128, false, false);
    constexpr uint32_t idesc_ts = make_idesc(true, 128, 128, false, true);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t off = (k / 4) * 16384 + (k % 4) * 32;
          if (mode & 4)   // TS: A (packed 16-bit, 8 columns per K16 step) from TMEM columns [448, 512), B MN-major from smem
            umma_ts(tmem + 256 + (it & 1) * 128, tmem + 192 + k * 8, make_sdesc_sw128(sbase + k * 2048, 16384, 1024), idesc_ts, k > 0);
          else
            umma_ss(tmem + 256 + (it & 1) * 128, make_sdesc_sw128(sbase + off, 16, 1024), make_sdesc_sw128(sbase + 32768 + off, 16, 1024),
                    idesc, k > 0);
        }
        umma_commit(smem_u32(&mbar));
      }
      __syncwarp();
      if (it >= depth) mbar_wait(smem_u32(&mbar), (it - depth) & 1);   // depth + 1 batches in flight at most
    }
    mbar_wait(smem_u32(&mbar), (iters - 1) & 1);
    t_mma = clock64() - t0;
  }
  if (lane == 0 && blockIdx.x == 0) {
    if (warp < ld_warps && do_ld) out[mode * 16 + warp] = t_ld;
    if (warp == 8 && do_mma) out[mode * 16 + 8] = t_mma;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc<512>(tmem);
}

int main() {
  long long *out;
  cudaMalloc(&out, 4096 * 8);
  cudaMemset(out, 0, 4096 * 8);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 1024);
  const int iters = 2000;
  for (int depth : {1, 3, 7}) {
    const int ldw = 4;
    printf("---- MMA batches in flight: %d\n", depth + 1);
    for (int mode : {1, 2, 3, 8}) {
      if (mode == 1 && depth > 1) continue;
      for (int rep = 0; rep < 2; ++rep) bench<<<148, 320, 65536, 0>>>(mode, iters, ldw, out, depth);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
      long long h[160];
      cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
      printf("ld_warps %d mode %d (%s):", ldw, mode, mode == 1 ? "ld only" : mode == 2 ? "SS mma only" : mode == 3 ? "ld + SS mma" : mode == 6 ? "TS mma only" : mode == 7 ? "ld + TS mma" : "forward stream");
      if (mode & 1) {
        double c = (double)h[mode * 16 + 0] / iters;
        printf("  ld: %.0f cyc per 16 KB per warp (%.1f B/clk/warp, %.1f B/clk/SM with %d warps)", c, 16384.0 / c, ldw * 16384.0 / c, ldw);
      }
      if (mode == 8) printf("  forward MMA stream (PV0 S0 PV1 S1): %.0f cyc per iteration (4 x 128^3; ideal 2048)", (double)h[mode * 16 + 8] / iters);
      else if (mode & 2) printf("  mma: %.0f cyc per 128x128x128 (ideal 512)", (double)h[mode * 16 + 8] / iters);
      printf("\n");
    }
  }
  return 0;
}
