// mfbank.cu -- A8/A9 decimated matched-filter bank.
//
// The reference computes np.convolve(rx, taps, "same") at EVERY sample for each
// filter and the driver loop then uses one column in sps
// (examples/soqpsk_detection.py:134-173, :189-196).  Here only the used columns
// are computed: for detector call i at sample n_i = first_instant + i*sps
//     y_f[i] = sum_t rx[n_i + center - t] * taps[f][t]        (rx = 0 outside [0,T))
//     z_r[i] = sum_f combine[r][f] * y_f[i]
// One block = one frame x (R*128) consecutive calls.  The rx tile is staged in
// shared memory TRANSPOSED as [n mod sps][q mod R][q / R] (q = n / sps) so that the
// 32 lanes of a warp -- which own consecutive groups of R calls -- read
// consecutive 8-byte words: conflict free.  Each thread keeps R calls in registers
// so every staged sample is read once per R*(taps/sps) uses; taps are kernel
// parameters, i.e. constant-bank operands of the FFMAs in the unrolled loop.
// Results are staged back through shared memory and stored coalesced.
#include "common.cuh"
#include <string.h>

namespace wf {

constexpr int kMfThreads = 128;
constexpr int kMfMaxTapEntries = 576;  // n_filt * n_taps (x 2 for the duplicated real taps of the tiled kernel)
constexpr int kMfMaxComb = 96;         // n_out * n_filt

// row stride (complex elements) of the transposed tile; 2*qs*R == 4 (mod 32) words keeps the
// transposing stores spread over the banks for sps = 8
__host__ __device__ constexpr int mf_qs(int sps, int nt, int R) {
  const int sym = R * kMfThreads;
  int qs = (((sym - 1) * sps + nt - 1) / sps) / R + 2;
  for (int guard = 0; guard < 16 && (qs * R) % 16 != 2; ++guard) ++qs;
  return qs;
}

template <typename T> struct Cplx;
template <> struct Cplx<float> { using type = float2; };
template <> struct Cplx<double> { using type = double2; };

template <typename T>
struct MfParams {
  const void* rx;
  void* out;
  int64_t B, Tn, ld_rx, ld_out, first_instant, n_calls;
  int n_filt, n_taps, center, n_out, sps, cplx_taps, has_comb;
  int qs;  // row stride of the transposed tile, in complex elements
  T tap_re[kMfMaxTapEntries];
  T tap_im[kMfMaxTapEntries];
  T comb_re[kMfMaxComb];
  T comb_im[kMfMaxComb];
};

// NT/NF/SPS > 0: compile-time shape (fully unrolled, constant-bank taps).
// NT == 0: run-time shape (API-parity path for arbitrary banks).
template <typename T, int SPS, int NF, int NT, bool CPLX, int R>
__global__ void __launch_bounds__(kMfThreads)
mf_bank_kernel(const __grid_constant__ MfParams<T> p) {
  using C = typename Cplx<T>::type;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  C* tile = reinterpret_cast<C*>(smem_raw);

  constexpr bool CT = NT > 0;
  const int sps = (CT || SPS > 0) ? SPS : p.sps;  // SPS > 0 with NT == 0: divisions by the sample rate become shifts / multiplies
  const int nt = CT ? NT : p.n_taps;
  const int nf = CT ? NF : p.n_filt;
  const int n_out = p.n_out;
  const int qs = CT ? mf_qs(SPS > 0 ? SPS : 1, NT > 0 ? NT : 1, R) : p.qs;  // compile-time stride: LDS with immediate offsets
  const int tid = threadIdx.x;
  constexpr int SYM = R * kMfThreads;

  const int64_t n_tiles = (p.n_calls + SYM - 1) / SYM;
  const int64_t frame = (int64_t)blockIdx.x / n_tiles;  // consecutive blocks share a frame (halo hits L2)
  const int64_t i0 = ((int64_t)blockIdx.x - frame * n_tiles) * SYM;
  const int n_sym = (int)min((int64_t)SYM, p.n_calls - i0);
  const C* rx = reinterpret_cast<const C*>(p.rx) + frame * p.ld_rx;
  const int64_t n_lo = p.first_instant + i0 * sps + p.center - (nt - 1);
  const int n_samp = (SYM - 1) * sps + nt;

  // ---- stage the rx tile, transposed ----------------------------------------------------
  // interior float tiles: 128-bit global loads, four in flight per thread, each split into two
  // 64-bit shared stores (samples n, n+1 live in adjacent rows of the transposed tile)
  bool staged = false;
  if constexpr (sizeof(T) == 4) {
    const bool fast = ((n_lo & 1) == 0) && ((p.ld_rx & 1) == 0) && ((sps & 1) == 0) && n_lo >= 0 &&
                      n_lo + n_samp <= p.Tn && ((reinterpret_cast<uintptr_t>(p.rx) & 15) == 0);
    if (fast) {
      const float4* src = reinterpret_cast<const float4*>(rx + n_lo);
      const int npair = n_samp >> 1;
      for (int base = 0; base < npair; base += 4 * kMfThreads) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int pi = base + k * kMfThreads + tid;
          if (pi < npair) v[k] = __ldg(src + pi);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int pi = base + k * kMfThreads + tid;
          if (pi < npair) {
            const int mm = 2 * pi, r = mm % sps, q = mm / sps;
            C* d = &tile[(r * R + (q % R)) * qs + q / R];
            d[0] = make_float2(v[k].x, v[k].y);
            d[R * qs] = make_float2(v[k].z, v[k].w);
          }
        }
      }
      if ((n_samp & 1) && tid == 0) {
        const int mm = n_samp - 1, r = mm % sps, q = mm / sps;
        tile[(r * R + (q % R)) * qs + q / R] = rx[n_lo + mm];
      }
      staged = true;
    }
  }
  if (!staged) {  // edges of the frame, float64, odd alignments: cp.async per sample, zeros outside [0, T)
#pragma unroll 4
    for (int m = tid; m < n_samp; m += kMfThreads) {
      const int64_t n = n_lo + m;
      const int r = m % sps, q = m / sps;
      C* d = &tile[(r * R + (q % R)) * qs + q / R];
      if (n >= 0 && n < p.Tn) {
        cp_async<sizeof(C)>(d, rx + n);
      } else {
        C zero;
        zero.x = 0; zero.y = 0;
        *d = zero;
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
  }
  __syncthreads();

  // ---- R calls per thread ----------------------------------------------------------
  // run-time tap count: NF > 0 still fixes the filter count at compile time (accumulators in registers, real or
  // complex taps chosen by CPLX); NF == 0 is the any-shape case: up to 32 filters (multi-h bank), accumulators in
  // local memory
  constexpr bool CF = !CT && NF > 0;
  constexpr int NFM = (CT || CF) ? NF : 32;
  T yr[R][NFM], yi[R][NFM];
#pragma unroll
  for (int e = 0; e < R; ++e)
#pragma unroll
    for (int f = 0; f < NFM; ++f) { yr[e][f] = 0; yi[e][f] = 0; }

  if constexpr (CT) {
    constexpr int NV = (R - 1) * SPS + NT;  // distinct samples this thread touches
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int r = v % SPS, c = v / SPS;
      const C x = tile[(r * R + (c % R)) * qs + tid + c / R];
#pragma unroll
      for (int e = 0; e < R; ++e) {
        const int u = v - e * SPS;
        if (u >= 0 && u < NT) {
          const int tt = NT - 1 - u;
#pragma unroll
          for (int f = 0; f < NF; ++f) {
            const T tr = p.tap_re[f * NT + tt];
            if (CPLX) {
              const T ti = p.tap_im[f * NT + tt];
              yr[e][f] = fma(x.x, tr, yr[e][f]); yr[e][f] = fma(-x.y, ti, yr[e][f]);
              yi[e][f] = fma(x.x, ti, yi[e][f]); yi[e][f] = fma(x.y, tr, yi[e][f]);
            } else {
              yr[e][f] = fma(x.x, tr, yr[e][f]);
              yi[e][f] = fma(x.y, tr, yi[e][f]);
            }
          }
        }
      }
    }
  } else {
    const int nv = (R - 1) * sps + nt;
#pragma unroll 8
    for (int v = 0; v < nv; ++v) {
      const int r = v % sps, c = v / sps;
      const C x = tile[(r * R + (c % R)) * qs + tid + c / R];
#pragma unroll
      for (int e = 0; e < R; ++e) {
        const int u = v - e * sps;
        if (u >= 0 && u < nt) {
          const int tt = nt - 1 - u;
          if constexpr (CF) {
#pragma unroll
            for (int f = 0; f < NFM; ++f) {
              const T tr = p.tap_re[f * nt + tt];
              if (CPLX) {
                const T ti = p.tap_im[f * nt + tt];
                yr[e][f] = fma(x.x, tr, yr[e][f]); yr[e][f] = fma(-x.y, ti, yr[e][f]);
                yi[e][f] = fma(x.x, ti, yi[e][f]); yi[e][f] = fma(x.y, tr, yi[e][f]);
              } else {
                yr[e][f] = fma(x.x, tr, yr[e][f]);
                yi[e][f] = fma(x.y, tr, yi[e][f]);
              }
            }
          } else {
            for (int f = 0; f < nf; ++f) {
              const T tr = p.tap_re[f * nt + tt];
              const T ti = p.cplx_taps ? p.tap_im[f * nt + tt] : (T)0;
              yr[e][f] = fma(x.x, tr, yr[e][f]); yr[e][f] = fma(-x.y, ti, yr[e][f]);
              yi[e][f] = fma(x.x, ti, yi[e][f]); yi[e][f] = fma(x.y, tr, yi[e][f]);
            }
          }
        }
      }
    }
  }
  __syncthreads();  // every thread is done reading the rx tile; reuse it for the output

  // ---- combine + stage the output tile [SYM][n_out] ----------------------------------
  // rows are `po` elements apart, po odd: a thread writes one row, so with an even n_out (32 for the
  // multi-h bank: 256-byte thread stride) every lane of a store would land in the same bank pair
  const int po = n_out | 1;
#pragma unroll
  for (int e = 0; e < R; ++e) {
    const int s = R * tid + e;
    if (p.has_comb) {
      for (int r = 0; r < n_out; ++r) {
        T zr = 0, zi = 0;
#pragma unroll
        for (int f = 0; f < NFM; ++f) {
          if (f < nf) {
            const T cr = p.comb_re[r * nf + f], ci = p.comb_im[r * nf + f];
            zr = fma(cr, yr[e][f], zr); zr = fma(-ci, yi[e][f], zr);
            zi = fma(cr, yi[e][f], zi); zi = fma(ci, yr[e][f], zi);
          }
        }
        C z; z.x = zr; z.y = zi;
        tile[s * po + r] = z;
      }
    } else {
#pragma unroll
      for (int f = 0; f < NFM; ++f) {
        if (f < nf) { C z; z.x = yr[e][f]; z.y = yi[e][f]; tile[s * po + f] = z; }
      }
    }
  }
  __syncthreads();
  C* out = reinterpret_cast<C*>(p.out) + (frame * p.ld_out + i0) * n_out;
  const int n_el = n_sym * n_out;
  int row = tid / n_out, col = tid - row * n_out;
  const int dr = kMfThreads / n_out, dc = kMfThreads - dr * n_out;
  for (int k = tid; k < n_el; k += kMfThreads) {
    out[k] = tile[row * po + col];
    row += dr; col += dc;
    if (col >= n_out) { col -= n_out; ++row; }
  }
}

// ---------------------------------------------------------------------------
// complex64 register-tiled path (compile-time shapes, even sps, 16-byte aligned rows).
// Each thread owns R = 4 consecutive calls: R*SPS + NT - SPS staged samples feed
// R*NT*NF*(2|4) FFMAs, so one 128-bit shared load (two samples) is amortised over up to
// 2*R*(NT/SPS) FFMA groups.  The rx tile keeps its NATURAL sample order -- it lands in
// shared memory with 16-byte cp.async.cg straight from HBM, no register round trip, no
// transposing stores -- and every thread's R*SPS-sample segment is followed by one 16-byte
// pad: thread stride = (R*SPS/2 + 1) 16-byte units, an odd number, so the eight lanes of a
// quarter-warp LDS.128 hit eight different bank groups (conflict free).  All shared
// addresses are thread base + immediate.  Output: R*n_out consecutive values per thread
// staged through the (reused) tile and stored with coalesced 128-bit words.
// ---------------------------------------------------------------------------
template <int SPS, int NT, int R>
struct MfTile {
  static constexpr int RS = R * SPS;                       // samples per thread segment
  static constexpr int SEG = RS + 2;                       // + one 16-byte pad
  static constexpr int SYM = R * kMfThreads;               // calls per block
  static constexpr int NSAMP = (SYM - 1) * SPS + NT;       // samples per block tile
  static constexpr int NV = (R - 1) * SPS + NT;            // samples one thread touches
  static constexpr int SLOTS = NSAMP + 2 * ((NSAMP + RS - 1) / RS) + 2;
  __host__ __device__ static constexpr int pos(int m) { return m + 2 * (m / RS); }
};

template <int SPS, int NF, int NT, bool CPLX, int R>
__global__ void __launch_bounds__(kMfThreads)
mf_bank_f32_kernel(const __grid_constant__ MfParams<float> p) {
  using TL = MfTile<SPS, NT, R>;
  static_assert((TL::RS % 4) == 0 && (SPS % 2) == 0, "segment stride must be an odd number of 16-byte units");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tile = reinterpret_cast<float2*>(smem_raw);
  const int tid = threadIdx.x;
  const int n_out = p.n_out;

  const int n_tiles = (int)((p.n_calls + TL::SYM - 1) / TL::SYM);
  const int frame = (int)(blockIdx.x / (unsigned)n_tiles);  // consecutive blocks share a frame (halo hits L2)
  const int i0 = ((int)blockIdx.x - frame * n_tiles) * TL::SYM;
  const int n_sym = min(TL::SYM, (int)p.n_calls - i0);
  const float2* rx = reinterpret_cast<const float2*>(p.rx) + (int64_t)frame * p.ld_rx;
  const int n_lo = (int)p.first_instant + i0 * SPS + p.center - (NT - 1);  // even (checked by the launcher)
  const int Tn = (int)p.Tn;

  // ---- stage: pairs of samples, 16 bytes each; zeros outside [0, T) -------------------
  // pair pi = tid + 128 k sits at tile offset pos(2 tid) + k (256 + 16): source and
  // destination both advance by compile-time strides, so the unrolled loop is LDGSTS +
  // immediates.  Interior tiles skip the bounds checks altogether.
  {
    constexpr int NPAIR = (TL::NSAMP + 1) / 2;
    constexpr int NK = (NPAIR + kMfThreads - 1) / kMfThreads;
    constexpr int DSTEP = 2 * kMfThreads + 2 * (2 * kMfThreads / TL::RS);  // pos() advance per k
    constexpr bool kConstStep = (2 * kMfThreads) % TL::RS == 0;  // pair stride = whole segments (sps 8)
    const float2* src = rx + n_lo + 2 * tid;
    float2* dst = tile + TL::pos(2 * tid);
    const bool interior = n_lo >= 0 && n_lo + TL::NSAMP + 1 <= Tn;
    if (interior) {
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        const int m = 2 * (tid + k * kMfThreads);
        if ((k + 1) * kMfThreads <= NPAIR || m < TL::NSAMP)
          cp_async_cg16(kConstStep ? dst + k * DSTEP : tile + TL::pos(m), src + k * 2 * kMfThreads);
      }
    } else {
      const int n0 = n_lo + 2 * tid;
#pragma unroll
      for (int k = 0; k < NK; ++k) {
        const int m = 2 * (tid + k * kMfThreads), n = n0 + k * 2 * kMfThreads;
        if ((k + 1) * kMfThreads <= NPAIR || m < TL::NSAMP) {
          float2* d = kConstStep ? dst + k * DSTEP : tile + TL::pos(m);
          if ((unsigned)n < (unsigned)(Tn - 1)) {
            cp_async_cg16(d, src + k * 2 * kMfThreads);
          } else {
#pragma unroll
            for (int s = 0; s < 2; ++s) {
              if (n + s >= 0 && n + s < Tn) cp_async<8>(d + s, rx + n + s);
              else d[s] = make_float2(0.f, 0.f);
            }
          }
        }
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
  }
  __syncthreads();

  // ---- R calls per thread --------------------------------------------------------------
  // y[e][f] = (re, im) of filter f at call e.  Taps arrive DUPLICATED, (t, t) per entry, so one
  // packed FFMA2 does (re, im) += (x_re, x_im) * (t, t); complex taps add (-x_im, x_re) * (ti, ti).
  const float2* tap2_re = reinterpret_cast<const float2*>(p.tap_re);
  const float2* tap2_im = reinterpret_cast<const float2*>(p.tap_im);
  float2 y[R][NF];
#pragma unroll
  for (int e = 0; e < R; ++e)
#pragma unroll
    for (int f = 0; f < NF; ++f) y[e][f] = make_float2(0.f, 0.f);

  const float2* seg = tile + tid * TL::SEG;
#pragma unroll
  for (int v = 0; v < TL::NV; v += 2) {
    float4 xx;
    if (v + 1 < TL::NV) {
      xx = *reinterpret_cast<const float4*>(seg + TL::pos(v));
    } else {
      const float2 x1 = seg[TL::pos(v)];
      xx = make_float4(x1.x, x1.y, 0.f, 0.f);
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (v + s < TL::NV) {
        const float2 x = s ? make_float2(xx.z, xx.w) : make_float2(xx.x, xx.y);
        const float2 xj = make_float2(-x.y, x.x);  // j x
#pragma unroll
        for (int e = 0; e < R; ++e) {
          const int u = v + s - e * SPS;
          if (u >= 0 && u < NT) {
            const int tt = NT - 1 - u;
#pragma unroll
            for (int f = 0; f < NF; ++f) {
              y[e][f] = ffma2(x, tap2_re[f * NT + tt], y[e][f]);
              if (CPLX) y[e][f] = ffma2(xj, tap2_im[f * NT + tt], y[e][f]);
            }
          }
        }
      }
    }
  }
  __syncthreads();  // every thread is done reading the rx tile; reuse it for the output

  // ---- combine + stage the output tile [SYM][n_out] -------------------------------------
#pragma unroll
  for (int e = 0; e < R; ++e) {
    const int s = R * tid + e;
    if (p.has_comb) {
      for (int r = 0; r < n_out; ++r) {
        float zr = 0.f, zi = 0.f;
#pragma unroll
        for (int f = 0; f < NF; ++f) {
          const float cr = p.comb_re[r * NF + f], ci = p.comb_im[r * NF + f];
          zr = fmaf(cr, y[e][f].x, zr); zr = fmaf(-ci, y[e][f].y, zr);
          zi = fmaf(cr, y[e][f].y, zi); zi = fmaf(ci, y[e][f].x, zi);
        }
        tile[s * n_out + r] = make_float2(zr, zi);
      }
    } else {
#pragma unroll
      for (int f = 0; f < NF; ++f) tile[s * NF + f] = y[e][f];
    }
  }
  __syncthreads();
  float2* out = reinterpret_cast<float2*>(p.out) + ((int64_t)frame * p.ld_out + i0) * n_out;
  const int n_el = n_sym * n_out;
  if (((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
    const int n4 = n_el >> 1;
    float4* o4 = reinterpret_cast<float4*>(out);
    const float4* t4 = reinterpret_cast<const float4*>(tile);
    for (int k = tid; k < n4; k += kMfThreads) o4[k] = t4[k];
    if ((n_el & 1) && tid == 0) out[n_el - 1] = tile[n_el - 1];
  } else {
    for (int k = tid; k < n_el; k += kMfThreads) out[k] = tile[k];
  }
}

template <typename T>
static size_t mf_smem_bytes(int sps, int nt, int R, int n_out, int& qs) {
  const int SYM = R * kMfThreads;
  qs = mf_qs(sps, nt, R);
  size_t in_bytes = (size_t)sps * R * qs * 2 * sizeof(T);
  size_t out_bytes = (size_t)SYM * (n_out | 1) * 2 * sizeof(T);  // odd row pitch, see the kernel
  return in_bytes > out_bytes ? in_bytes : out_bytes;
}

template <typename T>
static int launch_mf(const void* d_rx, int64_t B, int64_t Tn, int64_t ld_rx, const double* taps_re,
                     const double* taps_im, int n_filt, int n_taps, int center, const double* comb_re,
                     const double* comb_im, int n_out, int sps, int64_t first_instant, int64_t n_calls,
                     void* d_out, int64_t ld_out, cudaStream_t stream) {
  MfParams<T> p;
  memset(&p, 0, sizeof(p));
  p.rx = d_rx; p.out = d_out; p.B = B; p.Tn = Tn; p.ld_rx = ld_rx; p.ld_out = ld_out;
  p.first_instant = first_instant; p.n_calls = n_calls;
  p.n_filt = n_filt; p.n_taps = n_taps; p.center = center; p.n_out = n_out; p.sps = sps;
  p.cplx_taps = taps_im != nullptr; p.has_comb = comb_re != nullptr;
  for (int i = 0; i < n_filt * n_taps; ++i) {
    p.tap_re[i] = (T)taps_re[i];
    p.tap_im[i] = taps_im ? (T)taps_im[i] : (T)0;
  }
  if (comb_re)
    for (int i = 0; i < n_out * n_filt; ++i) {
      p.comb_re[i] = (T)comb_re[i];
      p.comb_im[i] = comb_im ? (T)comb_im[i] : (T)0;
    }

#define WF_MF_CASE(SPS_, NF_, NT_, CPLX_, R_)                                                         \
  do {                                                                                                \
    int qs;                                                                                           \
    size_t sm = mf_smem_bytes<T>(sps, n_taps, R_, n_out, qs);                                         \
    p.qs = qs;                                                                                        \
    auto kfn = mf_bank_kernel<T, SPS_, NF_, NT_, CPLX_, R_>;                                          \
    if (sm > 48 * 1024) {                                                                             \
      cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);\
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(mf_bank_kernel)");              \
    }                                                                                                 \
    const int SYM = R_ * kMfThreads;                                                                  \
    dim3 grid((unsigned)(((n_calls + SYM - 1) / SYM) * B));                                           \
    kfn<<<grid, kMfThreads, sm, stream>>>(p);                                                         \
  } while (0)

  const bool real_taps = taps_im == nullptr;
  if constexpr (sizeof(T) == 4) {
    // register-tiled complex64 path: even tile origin, even row pitch, 16-byte aligned rows
    const bool tiled_ok = (((first_instant + center - (n_taps - 1)) & 1) == 0) && ((ld_rx & 1) == 0) &&
                          ((reinterpret_cast<uintptr_t>(d_rx) & 15) == 0) && Tn < 0x7FFFFFF0 && n_calls < 0x7FFFFFF0 &&
                          first_instant < 0x7FFFFFF0 && (!comb_re || n_out <= n_filt + 2);
#define WF_MFT_CASE(SPS_, NF_, NT_, CPLX_, R_)                                                        \
  do {                                                                                                \
    using TL = MfTile<SPS_, NT_, R_>;                                                                 \
    static_assert(2 * NF_ * NT_ <= kMfMaxTapEntries, "duplicated taps do not fit");                   \
    for (int i = 0; i < NF_ * NT_; ++i) { /* (t, t) pairs: operands of the packed FFMA2 */            \
      p.tap_re[2 * i] = p.tap_re[2 * i + 1] = (T)taps_re[i];                                          \
      p.tap_im[2 * i] = p.tap_im[2 * i + 1] = taps_im ? (T)taps_im[i] : (T)0;                         \
    }                                                                                                 \
    size_t sm = (size_t)TL::SLOTS * sizeof(float2);                                                   \
    const size_t so = (size_t)TL::SYM * n_out * sizeof(float2);                                       \
    if (so > sm) sm = so;                                                                             \
    auto kfn = mf_bank_f32_kernel<SPS_, NF_, NT_, CPLX_, R_>;                                         \
    if (sm > 48 * 1024) {                                                                             \
      cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);\
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(mf_bank_f32_kernel)");          \
    }                                                                                                 \
    dim3 grid((unsigned)(((n_calls + TL::SYM - 1) / TL::SYM) * B));                                   \
    kfn<<<grid, kMfThreads, sm, stream>>>(p);                                                         \
    WF_LAUNCH_CHECK("mf_bank_f32_kernel");                                                            \
    return WF_OK;                                                                                     \
  } while (0)
    if (tiled_ok) {
      if (sps == 8 && n_filt == 2 && n_taps == 73 && real_taps) WF_MFT_CASE(8, 2, 73, false, 4);       // PAM bank, TG / A
      // (SOQPSK-B's 137-tap PAM bank takes the run-time-shape kernel below: a compiled 137-tap instance of this
      // fully unrolled kernel costs ptxas 12 minutes, too much for a waveform outside the benchmark configs)
      else if (sps == 8 && n_filt == 3 && n_taps == 9 && !real_taps) WF_MFT_CASE(8, 3, 9, true, 4);   // PT bank
      else if (sps == 8 && n_filt == 2 && n_taps == 17 && real_taps) WF_MFT_CASE(8, 2, 17, false, 4);  // PAM bank, SOQPSK-MIL
      else if (sps == 10 && n_filt == 2 && n_taps == 21 && real_taps) WF_MFT_CASE(10, 2, 21, false, 4);
      else if (sps == 10 && n_filt == 2 && n_taps == 91 && real_taps) WF_MFT_CASE(10, 2, 91, false, 4);
      else if (sps == 10 && n_filt == 3 && n_taps == 11 && !real_taps) WF_MFT_CASE(10, 3, 11, true, 4);
      else if (sps == 8 && n_filt == 4 && n_taps == 8 && !real_taps && !comb_re) WF_MFT_CASE(8, 4, 8, true, 4);  // PCM/FM, Lp = 2
    }
    // Odd tile origin (IRIG PCM/FM: first instant -1): one zero tap in front of every filter moves the origin
    // one sample back to an even index -- same sums, 9/8 of the FFMAs, but the 16-byte cp.async staging applies.
    const bool odd_ok = (((first_instant + center - n_taps) & 1) == 0) && ((ld_rx & 1) == 0) &&
                        ((reinterpret_cast<uintptr_t>(d_rx) & 15) == 0) && Tn < 0x7FFFFFF0 && n_calls < 0x7FFFFFF0 &&
                        first_instant < 0x7FFFFFF0;
    if (odd_ok && sps == 8 && n_filt == 4 && n_taps == 8 && !real_taps && !comb_re) {
      double pad_re[4 * 9], pad_im[4 * 9];
      for (int f = 0; f < 4; ++f)
        for (int t = 0; t < 9; ++t) {  // tap index NT-1 multiplies the OLDEST sample of the window
          pad_re[f * 9 + t] = t < 8 ? taps_re[f * 8 + t] : 0.0;
          pad_im[f * 9 + t] = t < 8 ? taps_im[f * 8 + t] : 0.0;
        }
      {
        const double* taps_re = pad_re;  // the macro reads these names
        const double* taps_im = pad_im;
        p.n_taps = 9;
        WF_MFT_CASE(8, 4, 9, true, 4);
      }
    }
#undef WF_MFT_CASE
  }
  if (sps == 8 && n_filt == 2 && n_taps == 73 && real_taps) WF_MF_CASE(8, 2, 73, false, 2);       // PAM bank, TG
  else if (sps == 8 && n_filt == 3 && n_taps == 9 && !real_taps) WF_MF_CASE(8, 3, 9, true, 2);   // PT bank
  else if (sps == 10 && n_filt == 2 && n_taps == 91 && real_taps) WF_MF_CASE(10, 2, 91, false, 2);
  else if (sps == 10 && n_filt == 3 && n_taps == 11 && !real_taps) WF_MF_CASE(10, 3, 11, true, 2);
  else if (sps == 8 && n_filt == 4 && n_taps == 8 && !real_taps && !comb_re) WF_MF_CASE(8, 4, 8, true, 2);    // PCM/FM, Lp = 2
  else if (sps == 8 && n_filt == 32 && n_taps == 8 && !real_taps && !comb_re) WF_MF_CASE(8, 32, 8, true, 1);  // multi-h, Lp = 2
  // any tap count, filter count fixed: every other PAM bank (SOQPSK-B: 137 taps at sps 8, 171 at sps 10) and PT bank
  else if (sps == 8 && n_filt == 2 && real_taps) WF_MF_CASE(8, 2, 0, false, 2);
  else if (sps == 10 && n_filt == 2 && real_taps) WF_MF_CASE(10, 2, 0, false, 2);
  else if (n_filt == 2 && real_taps) WF_MF_CASE(0, 2, 0, false, 2);
  else if (n_filt == 3 && !real_taps) WF_MF_CASE(0, 3, 0, true, 2);
  else WF_MF_CASE(0, 0, 0, true, 1);
#undef WF_MF_CASE
  WF_LAUNCH_CHECK("mf_bank_kernel");
  return WF_OK;
}

}  // namespace wf

extern "C" {

int wf_mf_bank(const void* d_rx, int dtype, int64_t B, int64_t T, int64_t ld_rx, const double* taps_re,
               const double* taps_im, int n_filt, int n_taps, int center, const double* combine_re,
               const double* combine_im, int n_out, int sps, int64_t first_instant, int64_t n_calls,
               void* d_out, int64_t ld_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && n_calls >= 0, "negative shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  WF_REQUIRE(taps_re && n_filt >= 1 && n_filt <= 32 && n_taps >= 1, "bad filter bank shape (1..32 filters)");
  WF_REQUIRE(n_filt * n_taps <= kMfMaxTapEntries, "filter bank has %d taps in total (max %d)", n_filt * n_taps, kMfMaxTapEntries);
  WF_REQUIRE(sps >= 1 && sps <= 64, "sps=%d outside 1..64", sps);
  WF_REQUIRE(ld_rx >= T && ld_out >= n_calls, "leading dimension too small");
  if (combine_re) WF_REQUIRE(n_out >= 1 && n_out * n_filt <= kMfMaxComb, "combine matrix too large");
  else WF_REQUIRE(n_out == n_filt, "n_out must equal n_filt without a combine matrix");
  WF_REQUIRE(B * ((n_calls + 127) / 128) <= 0x7FFFFFFF, "batch too large for one launch; split the call");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || n_calls == 0) return WF_OK;
  WF_REQUIRE(d_rx && d_out, "NULL device pointer");
  if (dtype == WF_F32)
    return launch_mf<float>(d_rx, B, T, ld_rx, taps_re, taps_im, n_filt, n_taps, center, combine_re, combine_im, n_out,
                            sps, first_instant, n_calls, d_out, ld_out, as_stream(stream));
  return launch_mf<double>(d_rx, B, T, ld_rx, taps_re, taps_im, n_filt, n_taps, center, combine_re, combine_im, n_out,
                           sps, first_instant, n_calls, d_out, ld_out, as_stream(stream));
}

}  // extern "C"
