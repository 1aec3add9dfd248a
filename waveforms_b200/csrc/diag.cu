// diag.cu -- SURVEY 8(f)3/4: signal diagnostics as CUDA reductions / gathers over the baseband the
// modulator produces.  The reference only PLOTS these (waveforms/viz/*.py, matplotlib -- figure output is
// unavailable on this box); here they produce the ARRAYS the figures are drawn from.
//
//   welch_psd_kernel      viz/psd.py:16-54 = matplotlib Axes.psd(x, NFFT, Fs, scale_by_freq=False), i.e.
//                         matplotlib.mlab.psd (third-party, pinned 3.7.5 / 3.9.4 / 3.10.1 in the reference's
//                         uv.lock, absent from /root/reference): segments of NFFT samples, no overlap, no
//                         detrend, symmetric Hann window, |FFT|^2 averaged over segments, divided by
//                         (sum window)^2, two-sided and fft-shifted for complex input.
//   eye_traces_kernel     viz/eye.py:39-56: trace i = signal[i*span : i*span + span + 1], span = sps * modulo.
//   phase_tree_kernel     viz/tree.py:64-71: per chunk of `span` samples np.unwrap(np.angle(.)) minus an offset
//                         (the chunk's first value unless `off` is given).
//   timing_metric_kernel  a replacement for the hard-coded timing offsets of examples/soqpsk_detection.py:
//                         183-192: |mean (-1)^i y[ph + sps i]^2| of a matched-filter output per sampling phase ph.
#include "common.cuh"
#include <math.h>

namespace wf {

constexpr int kPsdThreads = 256;
constexpr int kPsdMaxLog = 12;  // NFFT <= 4096: 32 KB of shared memory for the in-place transform

// One block per frame.  Each segment is windowed into shared memory in bit-reversed order, transformed in
// place by log2(N) radix-2 stages (twiddles from a shared table), and |X_k|^2 is accumulated in registers:
// thread t owns bins t, t + 256, ...  No atomics: the result is deterministic.
__global__ void __launch_bounds__(kPsdThreads)
welch_psd_kernel(const float2* __restrict__ sig, int64_t ld, int T, int logn, float amp2, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int N = 1 << logn;
  float2* x = reinterpret_cast<float2*>(smem_raw);   // [N]
  float2* tw = x + N;                                 // [N / 2]: exp(-2 pi j k / N)
  float* win = reinterpret_cast<float*>(tw + N / 2);  // [N]: np.hanning(N)
  const int tid = threadIdx.x;
  const float2* row = sig + (int64_t)blockIdx.x * ld;
  double wsum_local = 0.0;
  for (int k = tid; k < N; k += kPsdThreads) {
    // symmetric Hann, 0.5 - 0.5 cos(2 pi n / (N - 1)), evaluated in double like numpy
    const double w = N > 1 ? 0.5 - 0.5 * cospi(2.0 * (double)k / (double)(N - 1)) : 1.0;
    win[k] = (float)w;
    wsum_local += w;
    if (k < N / 2) {
      double s, c;
      sincospi(-2.0 * (double)k / (double)N, &s, &c);
      tw[k] = make_float2((float)c, (float)s);
    }
  }
  // block sum of the window (double)
  __shared__ double red[kPsdThreads / 32];
  for (int o = 16; o > 0; o >>= 1) wsum_local += __shfl_xor_sync(0xFFFFFFFFu, wsum_local, o);
  if ((tid & 31) == 0) red[tid >> 5] = wsum_local;
  __syncthreads();
  double wsum = 0.0;
  for (int i = 0; i < kPsdThreads / 32; ++i) wsum += red[i];

  const int n_seg = T / N;
  float acc[(1 << kPsdMaxLog) / kPsdThreads];
#pragma unroll
  for (int j = 0; j < (1 << kPsdMaxLog) / kPsdThreads; ++j) acc[j] = 0.f;
  for (int sg = 0; sg < n_seg; ++sg) {
    const float2* seg = row + (int64_t)sg * N;
    for (int k = tid; k < N; k += kPsdThreads) {
      const float2 v = seg[k];
      const float w = win[k];
      const int r = (int)(__brev((unsigned)k) >> (32 - logn));
      x[r] = make_float2(v.x * w, v.y * w);
    }
    __syncthreads();
    for (int s = 0; s < logn; ++s) {
      const int half = 1 << s;
      for (int b = tid; b < N / 2; b += kPsdThreads) {
        const int j = b & (half - 1);
        const int i0 = ((b >> s) << (s + 1)) | j, i1 = i0 + half;
        const float2 w = tw[j << (logn - 1 - s)];
        const float2 a = x[i0], c = x[i1];
        const float2 t = make_float2(c.x * w.x - c.y * w.y, c.x * w.y + c.y * w.x);
        x[i0] = make_float2(a.x + t.x, a.y + t.y);
        x[i1] = make_float2(a.x - t.x, a.y - t.y);
      }
      __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < (1 << kPsdMaxLog) / kPsdThreads; ++j) {
      const int k = tid + j * kPsdThreads;
      if (k < N) { const float2 v = x[k]; acc[j] = fmaf(v.x, v.x, fmaf(v.y, v.y, acc[j])); }
    }
    __syncthreads();
  }
  // mean over segments, / (sum window)^2, x amp2 (the caller's signal scaling squared), fft-shifted
  const float scale = n_seg > 0 ? (float)((double)amp2 / ((double)n_seg * wsum * wsum)) : 0.f;
  float* orow = out + (int64_t)blockIdx.x * N;
#pragma unroll
  for (int j = 0; j < (1 << kPsdMaxLog) / kPsdThreads; ++j) {
    const int k = tid + j * kPsdThreads;
    if (k < N) orow[(k + N / 2) & (N - 1)] = acc[j] * scale;
  }
}

// trace i of frame b = sig[b][i*span .. i*span + span] (span + 1 samples: neighbouring traces share one)
__global__ void __launch_bounds__(256)
eye_traces_kernel(const float2* __restrict__ sig, int64_t B, int64_t ld, int n_tr, int span, float2* __restrict__ out) {
  const int64_t per = (int64_t)n_tr * (span + 1);
  const int64_t total = B * per;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = e / per, r = e - b * per;
    const int tr = (int)(r / (span + 1)), k = (int)(r - (int64_t)tr * (span + 1));
    out[e] = sig[b * ld + (int64_t)tr * span + k];
  }
}


// One warp per (frame, chunk): out[k] = unwrap(angle(sig[chunk*span + k])) - offset.  np.unwrap adds to every
// sample the accumulated correction -2 pi floor((d + pi) / 2 pi) of the differences d before it (a difference of
// exactly +pi... -pi maps to -pi unless d > 0): the correction is an integer number of turns, so it is
// accumulated exactly as an int by a warp scan and applied once.
constexpr int kTreeWarps = 4;
__global__ void __launch_bounds__(kTreeWarps * 32)
phase_tree_kernel(const float2* __restrict__ sig, int64_t B, int64_t ld, int n_chunks, int span, int use_off, float off,
                  float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * kTreeWarps + (threadIdx.x >> 5);
  if (item >= B * n_chunks) return;
  const int64_t b = item / n_chunks;
  const int ch = (int)(item - b * n_chunks);
  const float2* src = sig + b * ld + (int64_t)ch * span;
  float* dst = out + item * (int64_t)span;
  const float two_pi = 6.283185307179586f, pi = 3.14159274101257324f;
  float prev_last = 0.f, first = 0.f;
  int turns_carry = 0;
  for (int base = 0; base < span; base += 32) {
    const int k = base + lane;
    float p = 0.f;
    if (k < span) { const float2 v = src[k]; p = atan2f(v.y, v.x); }
    float pm1 = __shfl_up_sync(0xFFFFFFFFu, p, 1);
    if (lane == 0) pm1 = prev_last;
    int t = 0;
    if (k > 0 && k < span) {
      const float d = p - pm1;
      const float fl = floorf((d + pi) / two_pi);
      float ddmod = (d + pi) - fl * two_pi - pi;
      t = -(int)fl;
      if (ddmod == -pi && d > 0.f) t += 1;
      if (fabsf(d) < pi) t = 0;
    }
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int u = __shfl_up_sync(0xFFFFFFFFu, t, s);
      if (lane >= s) t += u;
    }
    t += turns_carry;
    if (base == 0) first = __shfl_sync(0xFFFFFFFFu, p, 0);
    if (k < span) dst[k] = (p + two_pi * (float)t) - (use_off ? off : first);
    const int last = min(31, span - 1 - base);
    turns_carry = __shfl_sync(0xFFFFFFFFu, t, last);
    prev_last = __shfl_sync(0xFFFFFFFFu, p, last);
  }
}

// Sampling-phase metric for offset-QPSK-like CPM (SOQPSK), one block per frame:
//   m[ph] = | (1 / cnt) sum_i (-1)^i y[ph + sps i]^2 |,   cnt = number of samples at phase ph.
// The envelope |y|^2 of a constant-envelope signal carries no timing information; y^2 does: at the right
// sampling phase the matched-filter output alternates between the I axis (y^2 = +1) and the Q axis (y^2 = -1)
// from bit to bit, so the sign-alternated sum adds coherently there and the magnitude makes it independent of
// the carrier phase.  Deterministic (fixed summation tree).
__global__ void __launch_bounds__(256)
timing_metric_kernel(const float2* __restrict__ y, int64_t ld, int T, int sps, float* __restrict__ out) {
  __shared__ float red_re[256], red_im[256];
  const int tid = threadIdx.x;
  const int G = 256 / sps;              // sample groups walked in parallel
  const int phase = tid % sps, g = tid / sps;
  const float2* row = y + (int64_t)blockIdx.x * ld;
  float re = 0.f, im = 0.f;
  if (g < G)
    for (int i = g; phase + (int64_t)sps * i < T; i += G) {
      const float2 v = row[phase + sps * i];
      const float sq_re = fmaf(v.x, v.x, -v.y * v.y), sq_im = 2.f * v.x * v.y;
      if (i & 1) { re -= sq_re; im -= sq_im; } else { re += sq_re; im += sq_im; }
    }
  red_re[tid] = re;
  red_im[tid] = im;
  __syncthreads();
  if (tid < sps) {
    float sr = 0.f, si = 0.f;
    for (int k = 0; k < G; ++k) { sr += red_re[tid + k * sps]; si += red_im[tid + k * sps]; }
    const int cnt = (T - tid + sps - 1) / sps;
    out[(int64_t)blockIdx.x * sps + tid] = cnt > 0 ? sqrtf(sr * sr + si * si) / (float)cnt : 0.f;
  }
}

}  // namespace wf

extern "C" {

int wf_welch_psd(const void* d_sig, int64_t B, int64_t T, int64_t ld, int nfft, double amplitude, float* d_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld >= T, "bad shape");
  int logn = 0;
  while ((1 << logn) < nfft) ++logn;
  WF_REQUIRE(nfft >= 2 && (1 << logn) == nfft && logn <= kPsdMaxLog, "nfft must be a power of two in 2..%d", 1 << kPsdMaxLog);
  WF_REQUIRE(T < 0x7FFFFFF0, "frame too long");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  WF_REQUIRE(d_sig && d_out, "NULL device pointer");
  const size_t sm = (size_t)nfft * 8 + (size_t)nfft / 2 * 8 + (size_t)nfft * 4;
  if (sm > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(welch_psd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(welch_psd_kernel)");
  }
  welch_psd_kernel<<<(unsigned)B, kPsdThreads, sm, as_stream(stream)>>>(reinterpret_cast<const float2*>(d_sig), ld, (int)T, logn,
                                                                         (float)(amplitude * amplitude), d_out);
  WF_LAUNCH_CHECK("welch_psd_kernel");
  return WF_OK;
}

int wf_eye_traces(const void* d_sig, int64_t B, int64_t T, int64_t ld, int span, void* d_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld >= T && span >= 1, "bad shape");
  int rc;
  if ((rc = require_device())) return rc;
  const int64_t n_tr = T > 0 ? (T - 1) / span : 0;
  if (B == 0 || n_tr == 0) return WF_OK;
  WF_REQUIRE(d_sig && d_out, "NULL device pointer");
  WF_REQUIRE(n_tr < 0x7FFFFFF0, "too many traces");
  const int64_t total = B * n_tr * (span + 1);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  eye_traces_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float2*>(d_sig), B, ld, (int)n_tr, span,
                                                                      reinterpret_cast<float2*>(d_out));
  WF_LAUNCH_CHECK("eye_traces_kernel");
  return WF_OK;
}

int wf_phase_tree(const void* d_sig, int64_t B, int64_t T, int64_t ld, int span, int use_off, double off, float* d_out,
                  void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld >= T && span >= 1, "bad shape");
  int rc;
  if ((rc = require_device())) return rc;
  const int64_t n_chunks = T / span;
  if (B == 0 || n_chunks == 0) return WF_OK;
  WF_REQUIRE(d_sig && d_out, "NULL device pointer");
  WF_REQUIRE(n_chunks < 0x7FFFFFF0 && B * n_chunks / kTreeWarps < 0x7FFFFFF0, "too many chunks");
  const unsigned grid = (unsigned)((B * n_chunks + kTreeWarps - 1) / kTreeWarps);
  phase_tree_kernel<<<grid, kTreeWarps * 32, 0, as_stream(stream)>>>(reinterpret_cast<const float2*>(d_sig), B, ld, (int)n_chunks, span,
                                                                      use_off, (float)off, d_out);
  WF_LAUNCH_CHECK("phase_tree_kernel");
  return WF_OK;
}

int wf_timing_metric(const void* d_y, int64_t B, int64_t T, int64_t ld, int sps, float* d_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld >= T && T < 0x7FFFFFF0, "bad shape");
  WF_REQUIRE(sps >= 1 && sps <= 256, "sps=%d outside 1..256", sps);
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  WF_REQUIRE(d_y && d_out, "NULL device pointer");
  timing_metric_kernel<<<(unsigned)B, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float2*>(d_y), ld, (int)T, sps, d_out);
  WF_LAUNCH_CHECK("timing_metric_kernel");
  return WF_OK;
}

}  // extern "C"
