// pipe_bench.cu -- issue rate (cycles per warp instruction, 8 independent chains) and latency (1 dependent chain) of the
// instructions the softmax loops are made of, on one SM sub-partition: 1 warp per scheduler (128 threads) and 2 (256).
#include <cstdio>
#include <cuda_runtime.h>
#define N 2048
template <int OP, int CH> __global__ void k(float *out, long long *t, float a0) {
  float x[8]; unsigned long long y[8];
  for (int i = 0; i < 8; ++i) { x[i] = a0 + i + threadIdx.x; y[i] = ((unsigned long long)__float_as_uint(x[i]) << 32) | __float_as_uint(a0); }
  float b = a0 * 0.5f, c = a0 * 0.25f;
  unsigned long long b2 = y[1], c2 = y[2];
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < N; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (OP == 0) asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(b));
      if (OP == 1) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(b), "f"(c));
      if (OP == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if (OP == 3) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(y[i]) : "l"(b2), "l"(c2));
      if (OP == 4) { unsigned r_; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r_) : "f"(x[i]), "f"(b)); x[i] = __uint_as_float(r_); }
      if (OP == 5) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(y[i]) : "l"(b2));
      if (OP == 6) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(b), "f"(c));
    }
  }
  const long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += x[i] + (float)(y[i] & 0xff);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) t[0] = t1 - t0;
}
template <int OP> void run(const char *name, float *out, long long *t) {
  printf("%-10s", name);
  for (int threads : {128, 256}) {
    long long h;
    k<OP, 8><<<148, threads>>>(out, t, 1.5f); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
    printf("  %d warp/sched: %.2f cyc/instr (8 chains)", threads / 128, (double)h / (N * 4 * 8));
    k<OP, 1><<<148, threads>>>(out, t, 1.5f); cudaDeviceSynchronize(); cudaMemcpy(&h, t, 8, cudaMemcpyDeviceToHost);
    printf(", %.1f cyc dependent", (double)h / (N * 4));
  }
  printf("\n");
}
int main() {
  float *out; long long *t; cudaMalloc(&out, 148 * 256 * 4); cudaMalloc(&t, 8);
  run<0>("FMNMX", out, t); run<1>("FMNMX3", out, t); run<2>("MUFU.EX2", out, t); run<3>("FFMA2", out, t);
  run<4>("F2FP.BF16", out, t); run<5>("FADD2", out, t); run<6>("FFMA", out, t);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
