// noalloc.cu -- allocation-free SOQPSK detection for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
// reference.
//
// wf_soqpsk_detect_static == wf_mf_bank(rx -> z, complex64, combine given) followed by
// wf_viterbi_soqpsk(z -> decisions / counters) with the intermediate z (the detector's 3-vector per call,
// complex64[B][n_calls][3]) held in the file-scope device array below instead of a caller buffer.
// The array is part of the loaded module: every GPU that loads libwfb200 holds kStaticScratchBytes of
// HBM for it until the process exits.  All calls share it; two calls that overlap on the GPU (different
// streams, no dependency) race on it.
#include "common.cuh"
#include <stdlib.h>
#include <mutex>

namespace wf {

constexpr size_t kStaticScratchBytes = 32u << 20;  // 32 MiB = 1 398 101 detector calls of 3 complex64
constexpr int64_t kStaticScratchCalls = (int64_t)(kStaticScratchBytes / (3 * sizeof(float2)));

// In HOST code this name is a stand-in handle, not a device address: never pass it to a kernel.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// Device code may name the array directly.  The bank overwrites every element the detector reads in the
// same call, so nothing has to start at zero for correctness; this kernel poisons the part a call is about
// to use with quiet NaNs when WF_STATIC_SCRATCH_POISON is set in the environment, so that a stale read
// would show up as wrong decisions in the tests instead of silently reusing the previous call's values.
__global__ void static_scratch_poison_kernel(size_t n_words) {
  uint32_t* w = reinterpret_cast<uint32_t*>(g_static_scratch);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x)
    w[i] = 0x7FC00000u;
}

static std::mutex g_scratch_mu;
static void* g_scratch_ptr[64];  // device address of g_static_scratch, per device ordinal

// Device address of the array on the CURRENT device (looked up once per device, with it current).
static int static_scratch_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= 64) return fail(WF_ERR_DEVICE, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
    if (!ptr) return fail(WF_ERR_CUDA, "cudaGetSymbolAddress returned NULL for the static scratch array");
    g_scratch_ptr[dev] = ptr;
  }
  *out = g_scratch_ptr[dev];
  return WF_OK;
}

}  // namespace wf

extern "C" int64_t wf_static_scratch_max_calls(void) { return wf::kStaticScratchCalls; }

extern "C" int wf_soqpsk_detect_static(const void* d_rx, int64_t B, int64_t T, int64_t ld_rx,
                                       const double* taps_re, const double* taps_im, int n_filt, int n_taps,
                                       int center, const double* combine_re, const double* combine_im,
                                       int sps, int64_t first_instant, int64_t n_calls,
                                       int diff, int length, uint8_t* d_dec_bits, int8_t* d_dec_syms,
                                       const uint32_t* d_tx_packed, int64_t ld_words,
                                       const int8_t* d_tx_syms, int64_t S,
                                       const int32_t* d_point, int64_t* d_counters, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_calls >= 0, "negative shape");
  WF_REQUIRE(combine_re || n_filt == 3, "the detector needs 3 outputs per call: give combine_re[3][n_filt] or a 3-filter bank");
  // overflow-safe B * n_calls <= kStaticScratchCalls
  if (B > 0 && n_calls > 0 && (n_calls > kStaticScratchCalls || B > kStaticScratchCalls / n_calls))
    return fail(WF_ERR_ARG,
                "wf_soqpsk_detect_static: B * n_calls = %lld * %lld detector calls need more than the %zu-byte "
                "static scratch (limit %lld calls per launch); split the batch or use wf_mf_bank + wf_viterbi_soqpsk",
                (long long)B, (long long)n_calls, kStaticScratchBytes, (long long)kStaticScratchCalls);
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || n_calls == 0) return WF_OK;
  void* z = nullptr;
  if ((rc = static_scratch_address(&z))) return rc;
  static const bool poison = getenv("WF_STATIC_SCRATCH_POISON") != nullptr;
  if (poison) {
    const size_t n_words = (size_t)B * (size_t)n_calls * 3 * 2;
    static_scratch_poison_kernel<<<sm_count() * 4, 256, 0, as_stream(stream)>>>(n_words);
    WF_LAUNCH_CHECK("static_scratch_poison_kernel");
  }
  rc = wf_mf_bank(d_rx, WF_F32, B, T, ld_rx, taps_re, taps_im, n_filt, n_taps, center, combine_re, combine_im, 3,
                  sps, first_instant, n_calls, z, n_calls, stream);
  if (rc) return rc;
  return wf_viterbi_soqpsk(z, WF_F32, B, n_calls, n_calls, diff, length, d_dec_bits, d_dec_syms, d_tx_packed,
                           ld_words, d_tx_syms, S, d_point, d_counters, nullptr, stream);
}
