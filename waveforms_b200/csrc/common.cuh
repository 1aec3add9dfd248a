// common.cuh -- shared host/device helpers of libwfb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdarg.h>
#include "../../include/wfb200.h"

namespace wf {

// ---- per-thread error text -------------------------------------------------
char* err_buf();
int fail(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* where);

#define WF_REQUIRE(cond, ...)                                  \
  do {                                                         \
    if (!(cond)) return ::wf::fail(WF_ERR_ARG, __VA_ARGS__);   \
  } while (0)

#define WF_LAUNCH_CHECK(where)                                 \
  do {                                                         \
    cudaError_t e__ = cudaGetLastError();                      \
    if (e__ != cudaSuccess) return ::wf::cuda_fail(e__, where);\
  } while (0)

int require_device();  // WF_OK or WF_ERR_DEVICE (cached per device)
int sm_count();

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- Philox4x32-10 (Salmon et al., SC'11), counter-based ---------------------
struct u32x4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

__host__ __device__ __forceinline__ u32x4 philox4x32_10(u32x4 c, uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mulhi32(M0, c.x), lo0 = M0 * c.x;
    uint32_t hi1 = mulhi32(M1, c.z), lo1 = M1 * c.z;
    u32x4 n;
    n.x = hi1 ^ c.y ^ k0;
    n.y = lo1;
    n.z = hi0 ^ c.w ^ k1;
    n.w = lo0;
    c = n;
    k0 += W0;
    k1 += W1;
  }
  return c;
}

// Same function with the ten round keys (k + r W) precomputed by the host: they are then
// constant-bank operands of the XORs instead of per-call integer adds.
struct PhiloxKeys { uint32_t k[20]; };
static inline PhiloxKeys philox_round_keys(uint64_t seed) {
  PhiloxKeys rk;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; ++r) { rk.k[2 * r] = k0; rk.k[2 * r + 1] = k1; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
  return rk;
}
__host__ __device__ __forceinline__ u32x4 philox4x32_10_rk(u32x4 c, const PhiloxKeys& rk) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mulhi32(M0, c.x), lo0 = M0 * c.x;
    uint32_t hi1 = mulhi32(M1, c.z), lo1 = M1 * c.z;
    u32x4 n;
    n.x = hi1 ^ c.y ^ rk.k[2 * r];
    n.y = lo1;
    n.z = hi0 ^ c.w ^ rk.k[2 * r + 1];
    n.w = lo0;
    c = n;
  }
  return c;
}

// Noise layout shared by every noise kernel: results depend only on
// (seed, point, global frame, sample) -- never on batching or sharding.
// One Philox call feeds FOUR complex samples (one 32-bit word each).  Sample pair p
// (samples 2p, 2p+1) of a frame uses
//   call(p) = 4 (p >> 3) + (p & 3),   words (x, y) if (p & 4) == 0 else (z, w),
// i.e. pairs p and p ^ 4 (same offset inside two neighbouring sps-8 symbol intervals)
// share a call, so a thread that walks consecutive intervals amortises one call over two.
__host__ __device__ __forceinline__ uint64_t noise_call_of_pair(uint64_t pair) { return ((pair >> 3) << 2) | (pair & 3u); }
__host__ __device__ __forceinline__ u32x4 noise_counter(uint64_t call, uint64_t frame, uint32_t point) {
  u32x4 c;
  c.x = (uint32_t)call;
  c.y = (uint32_t)frame;
  c.z = (uint32_t)(frame >> 32) ^ ((uint32_t)(call >> 32) << 16);
  c.w = point;  // < 2^31; bit 31 marks the tail-refinement call (noise_tail_log2)
  return c;
}

#ifdef __CUDACC__
__device__ __forceinline__ float fast_sqrt(float x) {
#ifdef WF_HOSTEMU
  return sqrtf(x);
#else
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}

__device__ __forceinline__ float fast_log2(float x) {  // MUFU.LG2 without the denormal fix-up (x >= 2^-33 here)
#ifdef WF_HOSTEMU
  return log2f(x);
#else
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#endif
}

// Box-Muller on ONE 32-bit word -> one complex sigma*(N(0,1)+jN(0,1)) sample.
//   low 16 bits m:  x = (m + 0.5) 2^-16 in (0, 1).  m == 0 (probability 2^-16) refines the
//                   tail with a fresh 32-bit draw t: x = (t + 0.5) 2^-48, so the radius
//                   distribution is exact down to 2^-48 (8.1 sigma) -- never clipped at 4.7.
//   high 16 bits a: angle = (a + 0.5) 2 pi 2^-16 - pi, 65536 directions symmetric about 0.
// sigma rides inside the radius: r = sqrt(coef (log2(m + 0.5) - 16)), coef = -2 ln2 sigma^2, so
// the caller adds the result to the signal with no further multiply.  Both 16-bit fields become
// floats by OR-ing them under the exponent of 2^23 (one LOP3 / PRMT) and an FADD / FFMA on the
// FMA pipe: no I2F on the XU pipe, which already carries LG2 / SQRT / SIN / COS.
struct NoiseCoef { float coef, c16; };
__device__ __forceinline__ NoiseCoef noise_coef(float sigma) {
  NoiseCoef nc;
  nc.coef = -1.3862943611198906f * sigma * sigma;
  nc.c16 = -16.0f * nc.coef;
  return nc;
}

// log2(t + 0.5) - 32 of the refined 32-bit uniform t for word `word` of the call whose counter is c (rare path)
static __device__ __noinline__ float noise_tail_lgt(uint32_t cx, uint32_t cy, uint32_t cz, uint32_t cw, uint32_t k0, uint32_t k1,
                                                    int word) {
  u32x4 c;
  c.x = cx; c.y = cy; c.z = cz; c.w = cw | 0x80000000u;
  const u32x4 r = philox4x32_10(c, k0, k1);
  const uint32_t t = word == 0 ? r.x : (word == 1 ? r.y : (word == 2 ? r.z : r.w));
  return fast_log2((float)t + 0.5f) - 32.0f;
}

// log2(m + 0.5) of the low 16 bits
__device__ __forceinline__ float noise_lgt(uint32_t w) {
  return fast_log2(__uint_as_float((w & 0xFFFFu) | 0x4B000000u) - 8388607.5f);
}
// lgt = log2 of the radius uniform + 16 (noise_lgt, or the tail refinement)
__device__ __forceinline__ float2 box_muller16(uint32_t w, float lgt, NoiseCoef nc) {
  const float r = fast_sqrt(fmaf(lgt, nc.coef, nc.c16));  // sigma sqrt(-2 ln x)
  // (2^23 + a) * 2pi/65536 - (2^23 + 32767.5) * 2pi/65536
  const float ang = fmaf(__uint_as_float((w >> 16) | 0x4B000000u), 9.58738019107841e-05f, -807.3892822265625f);
  float s, c;
  __sincosf(ang, &s, &c);
  return make_float2(r * c, r * s);
}
// the two complex samples of a pair: words (wa, wb) = words (2 half, 2 half + 1) of call counter c
__device__ __forceinline__ void noise_pair(uint32_t wa, uint32_t wb, int half, const u32x4& c, uint32_t k0, uint32_t k1,
                                           NoiseCoef nc, float2& g0, float2& g1) {
  float la = noise_lgt(wa), lb = noise_lgt(wb);
  if (((wa & 0xFFFFu) == 0u) | ((wb & 0xFFFFu) == 0u)) {  // p = 2^-15 per pair
    if ((wa & 0xFFFFu) == 0u) la = noise_tail_lgt(c.x, c.y, c.z, c.w, k0, k1, 2 * half);
    if ((wb & 0xFFFFu) == 0u) lb = noise_tail_lgt(c.x, c.y, c.z, c.w, k0, k1, 2 * half + 1);
  }
  g0 = box_muller16(wa, la, nc);
  g1 = box_muller16(wb, lb, nc);
}
// c + sum of the four signed-byte products of a and b (IDP.4A)
__device__ __forceinline__ int dp4a_s8(uint32_t a, uint32_t b, int c) {
#ifdef WF_HOSTEMU
  for (int i = 0; i < 4; ++i) c += (int)(int8_t)(a >> (8 * i)) * (int)(int8_t)(b >> (8 * i));
  return c;
#else
  return __dp4a((int)a, (int)b, c);
#endif
}

// packed FP32 FMA (Blackwell FFMA2): one issue slot for two independent fused multiply-adds
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
#ifdef WF_HOSTEMU
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#else
  unsigned long long ra, rb, rc, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
#endif
}

// ---- cp.async helpers (LDGSTS): global -> shared without a register round trip ----
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gmem_src) {
#ifdef WF_HOSTEMU
  memcpy(smem_dst, gmem_src, BYTES);
#else
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(d), "l"(gmem_src), "n"(BYTES) : "memory");
#endif
}
// 16-byte copy that bypasses L1 (streaming data read once per block)
__device__ __forceinline__ void cp_async_cg16(void* smem_dst, const void* gmem_src) {
#ifdef WF_HOSTEMU
  memcpy(smem_dst, gmem_src, 16);
#else
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
#endif
}
__device__ __forceinline__ void cp_async_commit() {
#ifndef WF_HOSTEMU
  asm volatile("cp.async.commit_group;" ::: "memory");
#endif
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
#ifndef WF_HOSTEMU
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
#endif
}

#endif

}  // namespace wf
