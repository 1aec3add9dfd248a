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

// Counter layout shared by every noise kernel: results depend only on
// (seed, point, global frame, sample pair) -- never on batching or sharding.
__host__ __device__ __forceinline__ u32x4 noise_counter(uint64_t pair, uint64_t frame, uint32_t point) {
  u32x4 c;
  c.x = (uint32_t)pair;
  c.y = (uint32_t)frame;
  c.z = (uint32_t)(frame >> 32) ^ ((uint32_t)(pair >> 32) << 16);
  c.w = point;
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

// Box-Muller on two 32-bit uniforms -> one complex sigma*(N(0,1)+jN(0,1)) sample.
// `coef` = -2 ln2 sigma^2 (see noise_coef): sigma rides inside the radius, so the caller
// adds the result to the signal with no further multiply.  Uniforms come from I2FP
// (ALU pipe, not the XU pipe that already carries LG2/SQRT/SIN/COS):
//   x = (u1 + 0.5) 2^-32 in (0, 1]   -> radii reach sqrt(2 ln 2^33) = 6.76 sigma,
//   angle = int32(u2) 2 pi 2^-32 in [-pi, pi].
__device__ __forceinline__ float noise_coef(float sigma) { return -1.3862943611198906f * sigma * sigma; }

__device__ __forceinline__ float2 box_muller(uint32_t u1, uint32_t u2, float coef) {
  const float x = fmaf((float)u1, 2.3283064365386963e-10f, 1.1641532182693481e-10f);
  const float r = fast_sqrt(coef * fast_log2(x));  // sigma sqrt(-2 ln x)
  const float ang = (float)(int32_t)u2 * 1.4629180792671596e-09f;  // 2 pi 2^-32
  float s, c;
  __sincosf(ang, &s, &c);
  return make_float2(r * c, r * s);
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
