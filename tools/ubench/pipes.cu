// Developer microbenchmark: per-SM issue rates of the instructions the modulator leans on
// (FFMA, FFMA2 = fma.rn.f32x2, IMAD.WIDE, IMAD.HI+IMAD, LOP3, MUFU.*, I2FP).  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipes tools/ubench/pipes.cu && /tmp/pipes
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITERS = 2048, UNROLL = 8;

template <int OP>
__global__ void __launch_bounds__(256) k(float* out, uint32_t seed) {
  float f[UNROLL];
  uint32_t u[UNROLL];
  unsigned long long w[UNROLL];
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) {
    f[i] = 1.0f + 1e-3f * (threadIdx.x + i);
    u[i] = seed + threadIdx.x * 977u + i;
    w[i] = ((unsigned long long)__float_as_uint(f[i]) << 32) | __float_as_uint(f[i]);
  }
  const float a = 0.999f, b = 1e-4f;
  const unsigned long long aa = ((unsigned long long)__float_as_uint(a) << 32) | __float_as_uint(a);
  const unsigned long long bb = ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(b);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) {
      if (OP == 0) f[i] = fmaf(f[i], a, b);
      if (OP == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(w[i]) : "l"(aa), "l"(bb));
      if (OP == 2) { unsigned long long t = (unsigned long long)u[i] * 0xD2511F53u; u[i] = (uint32_t)(t >> 32) ^ (uint32_t)t; }
      if (OP == 3) u[i] = __umulhi(u[i], 0xD2511F53u) + 1u;
      if (OP == 4) u[i] = (u[i] ^ (u[(i + 1) % UNROLL] & seed)) ;
      if (OP == 5) asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      if (OP == 6) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      if (OP == 7) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      if (OP == 8) { f[i] = (float)(int)u[i]; u[i] = __float_as_uint(f[i]) + 1u; }
      if (OP == 9) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
      if (OP == 10) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) s += f[i] + (float)u[i] + __uint_as_float((uint32_t)w[i]) + __uint_as_float((uint32_t)(w[i] >> 32));
  if (s == 123.456f) out[0] = s;
}

template <int OP>
void run(const char* name, float* d, int sms, double clk_hz, int ops_per_inst) {
  const int blocks = sms * 4;
  k<OP><<<blocks, 256>>>(d, 1);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<OP><<<blocks, 256>>>(d, 2);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double warp_insts = (double)blocks * 8 * ITERS * UNROLL * ops_per_inst;
  const double per_clk_sm = warp_insts / (ms * 1e-3 * clk_hz) / sms;
  printf("%-28s %8.3f ms  %6.3f warp-inst/clk/SM  (%.1f lanes/clk/SM)\n", name, ms, per_clk_sm, per_clk_sm * 32);
}

int main() {
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double clk = khz * 1e3;
  printf("%s, %d SMs, clock %.0f MHz (max; rates assume it)\n", pr.name, pr.multiProcessorCount, clk / 1e6);
  float* d; cudaMalloc(&d, 4);
  const int s = pr.multiProcessorCount;
  run<0>("FFMA", d, s, clk, 1);
  run<1>("FFMA2 (fma.rn.f32x2)", d, s, clk, 1);
  run<2>("IMAD.WIDE + LOP3", d, s, clk, 2);
  run<3>("IMAD.HI + IADD", d, s, clk, 2);
  run<4>("LOP3", d, s, clk, 1);
  run<5>("FMUL.RZ + MUFU.SIN", d, s, clk, 1);
  run<6>("MUFU.LG2", d, s, clk, 1);
  run<7>("MUFU.SQRT (sqrt.approx)", d, s, clk, 1);
  run<8>("I2FP + IADD", d, s, clk, 2);
  run<9>("MUFU.EX2", d, s, clk, 1);
  run<10>("MUFU.RSQ", d, s, clk, 1);
  return 0;
}
