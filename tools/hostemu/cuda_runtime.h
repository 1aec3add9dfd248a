// DEVELOPMENT TOOL ONLY -- a minimal "CUDA on host threads" shim used to run the
// kernel SOURCES of libwfb200 on the CPU for pre-flight debugging in a container
// without a GPU.  It is never built into, loaded by, or reachable from the product
// (waveforms_b200/); tools/hostemu/build_emu.py compiles a separate throw-away
// library under /tmp.  One OS thread per CUDA thread, blocks run one at a time.
#pragma once
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#define WF_HOSTEMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __grid_constant__
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define __restrict__

struct uint3e { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct uint3 { uint32_t x, y, z; };
static inline uint3 make_uint3(uint32_t a, uint32_t b, uint32_t c) { return {a, b, c}; }
struct uint4 { uint32_t x, y, z, w; };
struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
static inline uint4 make_uint4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) { return {a, b, c, d}; }
static inline float2 make_float2(float a, float b) { return {a, b}; }
static inline float4 make_float4(float a, float b, float c, float d) { return {a, b, c, d}; }
static inline double2 make_double2(double a, double b) { return {a, b}; }

typedef int cudaError_t;
typedef void* cudaStream_t;
enum { cudaSuccess = 0 };
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
struct cudaDeviceProp { int major, minor, multiProcessorCount; };
static inline cudaError_t cudaGetLastError() { return 0; }
// "device" memory is host memory in the emulator, so a __device__ symbol's address is its host address
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return 0; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return 0; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) { p->major = 10; p->minor = 0; p->multiProcessorCount = 4; return 0; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }

namespace emu {

struct Barrier {
  std::mutex mu;
  std::condition_variable cv;
  int expected = 0, waiting = 0;
  uint64_t gen = 0;
  void reset(int n) { expected = n; waiting = 0; }
  void wait() {
    std::unique_lock<std::mutex> lk(mu);
    uint64_t g = gen;
    if (++waiting >= expected) { waiting = 0; ++gen; cv.notify_all(); return; }
    cv.wait(lk, [&] { return gen != g; });
  }
  void drop() {
    std::unique_lock<std::mutex> lk(mu);
    --expected;
    if (expected > 0 && waiting >= expected) { waiting = 0; ++gen; cv.notify_all(); }
  }
};

struct WarpCtx {
  Barrier bar;
  uint64_t slot[32];
  int pred[32];
};

struct BlockCtx {
  Barrier bar;
  std::vector<WarpCtx> warps;
  std::vector<unsigned char> dyn;
};

extern thread_local uint3e t_threadIdx, t_blockIdx;
extern thread_local dim3 t_blockDim, t_gridDim;
extern thread_local BlockCtx* t_block;
extern thread_local WarpCtx* t_warp;
extern thread_local int t_lane;

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body);

template <typename T>
static inline T shfl_from(T v, int src) {
  static_assert(sizeof(T) <= 8, "shuffle of wide type");
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  t_warp->slot[t_lane] = raw;
  t_warp->bar.wait();
  uint64_t got = t_warp->slot[src & 31];
  t_warp->bar.wait();
  T out;
  std::memcpy(&out, &got, sizeof(T));
  return out;
}

}  // namespace emu

#define threadIdx (::emu::t_threadIdx)
#define blockIdx (::emu::t_blockIdx)
#define blockDim (::emu::t_blockDim)
#define gridDim (::emu::t_gridDim)

static inline void __syncthreads() { ::emu::t_block->bar.wait(); }
template <typename T> static inline T __shfl_sync(unsigned, T v, int src) { return ::emu::shfl_from(v, src); }
template <typename T> static inline T __shfl_up_sync(unsigned, T v, int d) {
  int src = ::emu::t_lane - d;
  T got = ::emu::shfl_from(v, src < 0 ? ::emu::t_lane : src);
  return src < 0 ? v : got;
}
template <typename T> static inline T __shfl_xor_sync(unsigned, T v, int m) { return ::emu::shfl_from(v, ::emu::t_lane ^ m); }
static inline unsigned __ballot_sync(unsigned, int pred) {
  ::emu::t_warp->pred[::emu::t_lane] = pred ? 1 : 0;
  ::emu::t_warp->bar.wait();
  unsigned m = 0;
  for (int i = 0; i < 32; ++i) m |= (unsigned)::emu::t_warp->pred[i] << i;
  ::emu::t_warp->bar.wait();
  return m;
}
static inline int __all_sync(unsigned mk, int pred) { return __ballot_sync(mk, pred) == 0xFFFFFFFFu; }
static inline int __any_sync(unsigned mk, int pred) { return __ballot_sync(mk, pred) != 0u; }

static inline float fmaxf_emu(float a, float b) { return a > b ? a : b; }
static inline unsigned __brev(unsigned v) { unsigned r = 0; for (int i = 0; i < 32; ++i) r |= ((v >> i) & 1u) << (31 - i); return r; }
static inline double cospi(double a) { return cos(M_PI * a); }
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __ffs(unsigned v) { return __builtin_ffs((int)v); }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
static inline float __uint_as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline void emu_sincosf(float a, float* s, float* c) { *s = sinf(a); *c = cosf(a); }
static inline float emu_log2f(float x) { return log2f(x); }
#define __sincosf emu_sincosf
#define __log2f emu_log2f
static inline void sincospif(float a, float* s, float* c) { *s = (float)sin(M_PI * (double)a); *c = (float)cos(M_PI * (double)a); }
static inline void sincospi(double a, double* s, double* c) { *s = sin(M_PI * a); *c = cos(M_PI * a); }
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) {
  return __atomic_fetch_add(p, v, __ATOMIC_RELAXED);
}
template <typename T> static inline T min(T a, T b) { return b < a ? b : a; }
template <typename T> static inline T max(T a, T b) { return a < b ? b : a; }
using std::fma;
using std::fmin;
using std::rint;
using std::floor;
static inline void __syncwarp(unsigned = 0xFFFFFFFFu) { ::emu::t_warp->bar.wait(); }
template <typename T> static inline T __ldg(const T* p) { return *p; }
