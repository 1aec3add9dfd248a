// modulate.cu -- A6 CPM modulator (cpm/modulate.py:57-101) with optional fused
// A7 AWGN (noise.py:8-32), plus the stand-alone AWGN kernels.
//
// The reference up-samples h*alpha, convolves with the frequency pulse ("same"),
// accumulates sample by sample and takes exp(j phase).  Here the accumulation is
// closed-form: with Q the cumulative pulse, off=(K-1)//2,
//   revs[n] = sum_k w_k ( Q[clip(n+off+1-(k+1)sps, 0, K)] - Q[clip(off-(k+1)sps, 0, K)] ),  w_k = alpha_k h_k.
// Symbols whose pulse is complete contribute w_k Q[K]: that part is an exact
// 64-bit fixed-point prefix sum in turns (wraps mod 1 turn, order independent,
// no drift over long frames); only the J = ceil((K-1)/sps) newest symbols need
// Q-table terms per sample.  Three kernels share this arithmetic:
//   cpm_modulate_kernel<T>            any shape, float / double: one block per frame, 1024-symbol chunks (kChunkSyms),
//                                     threads own sample PAIRS (one 128-bit store each, 512 contiguous bytes per warp).
//   cpm_modulate_f32s8_kernel         complex64 at sps 8 (J = 8, 3, 1 or 16): whole-frame scan segments of up to 5120
//                                     symbols, 8-interval runs per thread, FFMA2 phase sums, one Philox call per two pairs.
//   cpm_modulate_bank_warp_kernel     the DEFAULT SOQPSK link (wf_cpm_modulate_bank, tune >= 1): one WARP per frame,
//                                     64-interval chunks, modulator + AWGN phase and PAM-bank phase by the same lanes
//                                     on a private shared-memory tile, no block barrier.  tune = 0 keeps
//                                     the one-block-per-frame fused instance of cpm_modulate_f32s8_kernel (FUSE = true).
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <type_traits>

#ifndef WF_MOD_FFMA2
#define WF_MOD_FFMA2 1
#endif

namespace wf {

constexpr int kModThreads = 256;
constexpr int kChunkSyms = 1024;
constexpr int kRing = 2048;       // power of two >= chunk + look-back + look-ahead
constexpr int kQMax = 768;        // padded Q table entries
constexpr int kScanPer = 2;       // symbols per thread per scan step (512 per step)
constexpr int kMaxH = 8;

template <typename T> struct ModTraits;
template <> struct ModTraits<float> {
  using Fix = uint32_t;  // top 32 bits of the 64-bit turn fraction
  using Vec2 = float2;
  __device__ static __forceinline__ Fix narrow(uint64_t f) { return (uint32_t)(f >> 32); }
  __device__ static __forceinline__ float turns(Fix f) { return (float)(int32_t)f * 2.3283064365386963e-10f; }
  __device__ static __forceinline__ void unit(float t, float& c, float& s) {
    t -= (t + 12582912.f) - 12582912.f;  // t - rint(t), |t| < 2^22
    __sincosf(t * 6.283185307179586f, &s, &c);
  }
};
template <> struct ModTraits<double> {
  using Fix = uint64_t;
  using Vec2 = double2;
  __device__ static __forceinline__ Fix narrow(uint64_t f) { return f; }
  __device__ static __forceinline__ double turns(Fix f) { return (double)(int64_t)f * 5.421010862427522e-20; }
  __device__ static __forceinline__ void unit(double t, double& c, double& s) {
    t -= rint(t);
    sincospi(2.0 * t, &s, &c);
  }
};

template <typename T>
struct ModParams {
  const int8_t* syms;
  void* out;
  const float* sigma;
  const int32_t* point;
  const uint64_t* frame_ids;  // global frame id per row (NULL: first_frame + row)
  int64_t B, S, Tn, ld_out;
  uint64_t seed, first_frame;
  int n_h, K, sps, off, J;
  T phase0_turns;
  T inv_sps;
  T hw[kMaxH];
  uint64_t hfix[kMaxH];
  T q[kQMax];  // Q[0..K], padded with Q[K]
  PhiloxKeys rk;  // round keys of `seed`
};

// SPS > 0: compile-time samples/symbol; JT > 0: compile-time J with the thread's
// Q coefficients hoisted into registers (needs (2*threads) % SPS == 0, off even).
template <typename T, int SPS, int JT, bool NOISE>
__global__ void __launch_bounds__(kModThreads)
cpm_modulate_kernel(const __grid_constant__ ModParams<T> p) {
  constexpr int SPSD = SPS > 0 ? SPS : 1;  // dead-code divisor when SPS is a run-time value
  using Tr = ModTraits<T>;
  using Fix = typename Tr::Fix;
  using Vec2 = typename Tr::Vec2;
  __shared__ T w_s[kRing + 16];  // slots [kRing, kRing+16) mirror [0, 16): a J-wide window never wraps
  __shared__ Fix f_s[kRing];
  __shared__ uint64_t warp_tot[kModThreads / 32];
  __shared__ T c0_s;

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t frame = blockIdx.x;
  const int sps = SPS > 0 ? SPS : p.sps;
  const int J = JT > 0 ? JT : p.J;
  const int off = p.off;
  const int64_t S = p.S, Tn = p.Tn;
  const int8_t* sy = p.syms + frame * S;
  Vec2* out = reinterpret_cast<Vec2*>(p.out) + frame * p.ld_out;
  const bool vec_ok = ((p.ld_out & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.out) & (2 * sizeof(Vec2) - 1)) == 0);

  // zero the ring (covers symbols k < 0 and F[m <= 0])
  for (int i = tid; i < kRing + 16; i += kModThreads) { w_s[i] = (T)0; if (i < kRing) f_s[i] = 0; }
  if (tid == 0) {
    // frame-start truncation of "same" convolution: sum_k w_k Q[clip(off-(k+1)sps,0,K)]
    T c0 = 0;
    for (int64_t k = 0; k < S && (k + 1) * sps < off; ++k) {
      int idx = off - (int)(k + 1) * sps;
      if (idx > p.K) idx = p.K;
      c0 += (T)sy[k] * p.hw[k % p.n_h] * p.q[idx];
    }
    c0_s = c0;
  }

  NoiseCoef ncoef = noise_coef(0.f);
  uint32_t pt = 0;
  uint64_t gframe = 0;
  if (NOISE) {
    ncoef = noise_coef(p.sigma[frame]);
    pt = p.point ? (uint32_t)p.point[frame] : 0u;
    gframe = p.frame_ids ? p.frame_ids[frame] : p.first_frame + (uint64_t)frame;
  }

  // hoisted Q coefficients (fast path): rho of this thread's even sample is fixed
  T qa[JT > 0 ? JT : 1], qb[JT > 0 ? JT : 1];
  if (JT > 0) {
    const int rho = (2 * tid + off) % SPSD;  // even sample; odd sample is rho + 1 (same symbol since rho even < SPS-1)
#pragma unroll
    for (int j = 0; j < JT; ++j) { qa[j] = p.q[rho + 1 + j * SPSD]; qb[j] = p.q[rho + 2 + j * SPSD]; }
  }

  uint64_t carry = 0;           // F[scanned]
  int64_t scanned = 0;          // symbols [0, scanned) are in the ring
  const int ahead = off / sps + 2;
  __syncthreads();
  const T c0 = c0_s;

  for (int64_t cs = 0; cs * sps < Tn; cs += kChunkSyms) {
    // ---- (a) extend the scan to symbol cs + chunk + ahead ----------------------
    const int64_t want = cs + kChunkSyms + ahead;
    while (scanned < want) {  // uniform across the block
      // each thread takes kScanPer consecutive symbols; the ring then holds
      // [cs - J, cs + 1536): never more than kRing entries (J + ahead <= 512)
      const int64_t k0 = scanned + kScanPer * (int64_t)tid;
      uint64_t inc[kScanPer];
      T wv[kScanPer];
      uint64_t run = 0;
#pragma unroll
      for (int i = 0; i < kScanPer; ++i) {
        const int64_t k = k0 + i;
        int a = (k < S) ? (int)sy[k] : 0;
        int hi = (int)(k % p.n_h);
        wv[i] = (T)a * p.hw[hi];
        run += (uint64_t)((int64_t)a) * p.hfix[hi];
        inc[i] = run;
      }
      // block-wide exclusive scan of `run`
      uint64_t incl = run;
#pragma unroll
      for (int s = 1; s < 32; s <<= 1) {
        uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, s);
        if (lane >= s) incl += t;
      }
      if (lane == 31) warp_tot[wid] = incl;
      __syncthreads();
      uint64_t base = carry;
      for (int i = 0; i < wid; ++i) base += warp_tot[i];
      uint64_t total = 0;
      for (int i = 0; i < kModThreads / 32; ++i) total += warp_tot[i];
      base += incl - run;
#pragma unroll
      for (int i = 0; i < kScanPer; ++i) {
        const int64_t k = k0 + i;
        const int slot = (int)(k & (kRing - 1));
        w_s[slot] = wv[i];
        if (slot < 16) w_s[slot + kRing] = wv[i];
        f_s[(k + 1) & (kRing - 1)] = Tr::narrow(base + inc[i]);  // F[k+1]
      }
      carry += total;
      scanned += kScanPer * kModThreads;
      __syncthreads();
    }

    // ---- (b) samples of symbols-intervals [cs, cs + chunk) --------------------
    const int n_base = (int)cs * sps;  // Tn < 2^31 (checked on the host): 32-bit sample indices
    const int npairs = kChunkSyms * sps / 2;
    const int Ti = (int)Tn;
    for (int q = tid; q < npairs; q += kModThreads) {
      const int n = n_base + 2 * q;
      if (n >= Ti) break;
      T ph[2];
      if (JT > 0) {
        const int khi = (int)((n + off) / SPSD) - 1;
        const T* wwin = &w_s[(khi - (JT - 1)) & (kRing - 1)];  // symbols khi-JT+1 .. khi, contiguous (mirrored tail)
        T acc0 = 0, acc1 = 0;
#pragma unroll
        for (int j = 0; j < JT; ++j) {
          const T w = wwin[JT - 1 - j];
          acc0 = fma(w, qa[j], acc0);
          acc1 = fma(w, qb[j], acc1);
        }
        const T old = Tr::turns(f_s[(khi - JT + 1) & (kRing - 1)]) + p.phase0_turns;
        ph[0] = fma(acc0 - c0, p.inv_sps, old);
        ph[1] = fma(acc1 - c0, p.inv_sps, old);
      } else {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int a = n + e + off;
          const int khi = a / sps - 1;
          const int rho = a - (khi + 1) * sps;
          T acc = 0;
          for (int j = 0; j < J; ++j) acc = fma(w_s[(khi - j) & (kRing - 1)], p.q[rho + 1 + j * sps], acc);
          const T old = Tr::turns(f_s[(khi - J + 1) & (kRing - 1)]) + p.phase0_turns;
          ph[e] = fma(acc - c0, p.inv_sps, old);
        }
      }
      T c[2], s[2];
      Tr::unit(ph[0], c[0], s[0]);
      Tr::unit(ph[1], c[1], s[1]);
      if (NOISE) {
        const uint32_t pr = (uint32_t)(n >> 1);
        const int half = (int)((pr >> 2) & 1u);
        const u32x4 ctr = noise_counter(noise_call_of_pair(pr), gframe, pt);
        const uint32_t k0 = (uint32_t)p.seed, k1 = (uint32_t)(p.seed >> 32);
        const u32x4 r = philox4x32_10(ctr, k0, k1);
        float2 g0, g1;
        noise_pair(half ? r.z : r.x, half ? r.w : r.y, half, ctr, k0, k1, ncoef, g0, g1);
        c[0] += (T)g0.x; s[0] += (T)g0.y;
        c[1] += (T)g1.x; s[1] += (T)g1.y;
      }
      if (vec_ok && n + 1 < Ti) {
        if (sizeof(T) == 4) {
          *reinterpret_cast<float4*>(out + n) = make_float4((float)c[0], (float)s[0], (float)c[1], (float)s[1]);
        } else {
          double2* o = reinterpret_cast<double2*>(out + n);
          o[0] = make_double2((double)c[0], (double)s[0]);
          o[1] = make_double2((double)c[1], (double)s[1]);
        }
      } else {
        Vec2 v0; v0.x = c[0]; v0.y = s[0];
        out[n] = v0;
        if (n + 1 < Ti) { Vec2 v1; v1.x = c[1]; v1.y = s[1]; out[n + 1] = v1; }
      }
    }
    // next chunk's scan starts with a __syncthreads-protected phase; ring slots it
    // overwrites are older than anything chunk `cs` reads (kRing >= 2*chunk).
  }
}

// ---------------------------------------------------------------------------
// complex64 fast path for sps = 8 and an even "same"-mode offset (SOQPSK-TG/A/B/MIL, multi-h).
// Same maths as above; what differs is the schedule.
//  * The block first turns a whole SEGMENT of the frame (kFSeg symbol intervals: the whole
//    frame when S <= ~5100) into w_k (float) and the fixed-point prefix sum F (top 32 bits)
//    in shared memory -- five scan steps back to back -- and then produces the samples with
//    NO further barrier, so the eight warps of a block drift freely through the long
//    MUFU/Philox part instead of meeting at a barrier every 1024 symbols.
//  * Four lanes own the four sample PAIRS of one symbol interval and walk RUN consecutive
//    intervals: the J-symbol window slides through registers (one LDS per new symbol), every
//    shared / global address is base + immediate, the thread's 2J Q coefficients (pre-scaled
//    to radians) stay in registers, and a warp store covers eight 64-byte runs of full sectors.
//  * One Philox call feeds the same pair offset of TWO neighbouring intervals (common.cuh),
//    i.e. two loop iterations of a thread.  Lanes whose pair index would start on the odd
//    half of a call begin one interval later (`e`), so every thread's calls are whole.
//  * The frame-start constant and the initial phase ride inside the prefix sum, so the angle
//    is fixed(F) * 2pi/2^32 in [-pi, pi) plus the J-tap partial: no range reduction.
// ---------------------------------------------------------------------------
constexpr int kFSeg = 5120;                   // symbol intervals per segment (multiple of 1024)
constexpr int kFLB = 24;                      // look-back slots in front of a segment (>= J + 2; J = 16 for SOQPSK-B)
constexpr int kFSlots = kFLB + kFSeg + 8;     // + spill of the lanes that start one interval late
constexpr size_t kFSmemBytes = (size_t)kFSlots * 8;

// ---------------------------------------------------------------------------
// Fused modulate -> AWGN -> PAM bank (wf_cpm_modulate_bank): the fast kernel below with FUSE = true
// never writes the noisy baseband to HBM.  A chunk of 512 symbol intervals (4096 samples) lands in a
// shared-memory tile laid out exactly like mf_bank_f32_kernel's (mfbank.cu: natural sample order, one
// 16-byte pad after every 32 samples, so the bank's per-thread 128-bit loads are conflict free), one
// barrier, warps 0-3 then run the bank's R = 4 calls-per-thread FFMA2 loop over the tile -- SAME
// operation order as mf_bank_f32_kernel, so the filter outputs are bit-identical to the two-kernel
// path -- and store 16 bytes per call (two raw PAM filter outputs); warps 4-7 carry the newest LB
// samples to the front of the tile for the next chunk.  Chunk j' (intervals 512 j' ..) completes calls
// [512 j' - CB, 512 (j'+1) - CB); the tile origin is the first sample of call 512 j' - CB's window.
// HBM per bit: 1 B in + 16 B out instead of 65 + 80.
// ---------------------------------------------------------------------------
constexpr int kFzNT = 73, kFzNF = 2, kFzR = 4;   // PAM bank of SOQPSK-TG at sps 8 (mfbank.cu WF_MFT_CASE(8, 2, 73, false, 4))
constexpr int kFzThreads = 128;                   // threads that run the bank phase: 4 x 128 = 512 calls per chunk
constexpr int kFzChunk = kFzR * kFzThreads;       // calls per chunk == intervals per chunk
constexpr int kFzLBMax = 96;                      // look-back samples kept from the previous chunk
constexpr int kFzSlots = 4464;                    // pos(8 * kFzChunk + kFzLBMax) rounded up
constexpr int kFzSeg = 2048;                      // scan segment of the fused instance (symbol intervals)
constexpr size_t kFzSmemBytes = (size_t)(kFLB + kFzSeg + 8) * 8 + (size_t)kFzSlots * 8;
__host__ __device__ constexpr int fz_pos(int m) { return m + 2 * (m >> 5); }

struct FuseParams {
  float2* z;            // [B][ld_z][2] raw outputs of the two real-tap filters
  int64_t ld_z;
  int n_calls, CB, LB;
  alignas(16) float tap2[2 * kFzNF * kFzNT];  // (t, t) pairs, filter-major: 64-bit uniform operands of the packed FFMA2
};
struct FuseNone { int unused; };

// R = 4 consecutive calls of thread t, BOTH filters (warps 0-3 only); mirrors mf_bank_f32_kernel<8, 2, 73, false, 4>.
static __device__ __forceinline__ void fz_bank_phase2(const FuseParams& fz, const float2* tile, int t, int cbase, float2* zrow) {
  const int c0 = cbase + kFzR * t;
  if (c0 + kFzR - 1 < 0 || c0 >= fz.n_calls) return;
  constexpr int NV = (kFzR - 1) * 8 + kFzNT;
  const float2* tap2 = reinterpret_cast<const float2*>(fz.tap2);
  float2 y[kFzR][kFzNF];
#pragma unroll
  for (int e = 0; e < kFzR; ++e)
#pragma unroll
    for (int f = 0; f < kFzNF; ++f) y[e][f] = make_float2(0.f, 0.f);
  const float2* seg = tile + t * (kFzR * 8 + 2);
#pragma unroll
  for (int v = 0; v < NV; v += 2) {
    float4 xx;
    if (v + 1 < NV) {
      xx = *reinterpret_cast<const float4*>(seg + fz_pos(v));
    } else {
      const float2 x1 = seg[fz_pos(v)];
      xx = make_float4(x1.x, x1.y, 0.f, 0.f);
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (v + s < NV) {
        const float2 x = s ? make_float2(xx.z, xx.w) : make_float2(xx.x, xx.y);
#pragma unroll
        for (int e = 0; e < kFzR; ++e) {
          const int u = v + s - e * 8;
          if (u >= 0 && u < kFzNT) {
#pragma unroll
            for (int f = 0; f < kFzNF; ++f) y[e][f] = ffma2(x, tap2[f * kFzNT + kFzNT - 1 - u], y[e][f]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < kFzR; ++e) {
    const int c = c0 + e;
    if ((unsigned)c < (unsigned)fz.n_calls)
      *reinterpret_cast<float4*>(zrow + 2 * (int64_t)c) = make_float4(y[e][0].x, y[e][0].y, y[e][1].x, y[e][1].y);
  }
}

// Rare path (p = 2^-15 per pair): one pair again, this time with the tail refinement of the
// radius uniform.  Kept out of line so the unrolled main loop stays branch free.
// Sample a: window wa (symbols A-JT .. A-1 of ITS interval), fa = F[A - JT], Q index qa0 + 8 j;
// sample b likewise (the two differ only for an odd "same" offset, see ODD in the kernel).
template <int JT>
static __device__ __noinline__ float4 mod_f32s8_slow_pair(const ModParams<float>& p, const float* wa, const float* wb, uint32_t fa,
                                                          uint32_t fb, int qa0, int qb0, u32x4 ctr, int half, NoiseCoef ncoef) {
  float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
  for (int j = 0; j < JT; ++j) {
    acc0 = fmaf(wa[JT - 1 - j], p.q[qa0 + 8 * j] * 0.78539816339744831f, acc0);
    acc1 = fmaf(wb[JT - 1 - j], p.q[qb0 + 8 * j] * 0.78539816339744831f, acc1);
  }
  const float olda = (float)(int32_t)fa * 1.4629180792671596e-09f, oldb = (float)(int32_t)fb * 1.4629180792671596e-09f;
  float c0f, s0f, c1f, s1f;
  __sincosf(acc0 + olda, &s0f, &c0f);
  __sincosf(acc1 + oldb, &s1f, &c1f);
  const uint32_t k0 = (uint32_t)p.seed, k1 = (uint32_t)(p.seed >> 32);
  const u32x4 r = philox4x32_10(ctr, k0, k1);
  float2 g0, g1;
  noise_pair(half ? r.z : r.x, half ? r.w : r.y, half, ctr, k0, k1, ncoef, g0, g1);
  return make_float4(c0f + g0.x, s0f + g0.y, c1f + g1.x, s1f + g1.y);
}

// ODD: the "same"-mode offset off = (K-1)/2 is odd (PCM/FM, K = 24).  Sample n = a - off is then
// even for ODD a, so a thread's pair is a = 8 A + 2 sub - 1, 8 A + 2 sub; for sub = 0 the first of
// the two is the LAST sample (rho = 7) of interval A - 1: one symbol further back in the window and
// in F.  The pair then runs on two scalar FFMA chains instead of one packed FFMA2 chain.
template <int JT, bool NOISE, int RUN, int MINB, bool ODD, bool FUSE = false, int FSEG = kFSeg>
__global__ void __launch_bounds__(kModThreads, MINB)
cpm_modulate_f32s8_kernel(const __grid_constant__ ModParams<float> p,
                          const __grid_constant__ typename std::conditional<FUSE, FuseParams, FuseNone>::type fz) {
  constexpr int kFChunk = RUN * (kModThreads / 4);
  constexpr int kFSeg = FSEG;                    // shadows the namespace constant: segment length of THIS instance
  constexpr int kFSlots = kFLB + FSEG + 8;
  constexpr size_t kFSmemBytes = (size_t)kFSlots * 8;
  static_assert(!FUSE || (!ODD && kFChunk == kFzChunk), "the fused bank wants 512-interval chunks and an even offset");
  constexpr int kScanStep = 4 * kModThreads;
  static_assert(kFSeg % kFChunk == 0 && kFSeg % kScanStep == 0, "segment = whole chunks and whole scan steps");
  static_assert(JT + 2 <= kFLB && (RUN % 2) == 0, "look-back too short / odd run");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* w_s = reinterpret_cast<float*>(smem_raw);          // slot(k) = k - k_lo + kFLB holds w_k
  uint32_t* f_s = reinterpret_cast<uint32_t*>(w_s + kFSlots);  // slot(k) holds F[k + 1]
  __shared__ uint64_t warp_tot[kModThreads / 32];
  __shared__ uint64_t base_s;

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t frame = blockIdx.x;
  float2* tile = reinterpret_cast<float2*>(smem_raw + kFSmemBytes);  // FUSE only
  if constexpr (FUSE) {  // samples before the frame are zero; published by the scan's barriers
    if (tid < kFzLBMax / 2 + 8) *reinterpret_cast<float4*>(tile + fz_pos(2 * tid)) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int off = p.off;
  const int S = (int)p.S, Ti = (int)p.Tn;
  const int8_t* sy = p.syms + frame * (int64_t)S;
  float4* out = reinterpret_cast<float4*>(reinterpret_cast<float2*>(p.out) + frame * p.ld_out);  // indexed by pair
  const bool sym_vec = ((S & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.syms) & 3) == 0);

  if (tid == 0) {
    // per-frame constant folded into the prefix sum: initial phase minus the frame-start
    // truncation of "same" convolution, sum_k w_k Q[clip(off-(k+1)8, 0, K)] / 8  (turns).
    // At most off/8 terms; float is ample (a constant phase error of 1e-7 turns).  No barrier
    // here: the value is published by the first barrier of the scan, so the other warps are
    // already fetching symbols while this lane works.
    float c0 = 0.f;
    for (int k = 0; k < S && (k + 1) * 8 < off; ++k) {
      int idx = off - (k + 1) * 8;
      if (idx > p.K) idx = p.K;
      c0 = fmaf((float)sy[k] * p.hw[k % p.n_h], p.q[idx], c0);
    }
    float t = fmaf(c0, -0.125f, p.phase0_turns);
    t -= rintf(t);                                                 // [-1/2, 1/2]
    base_s = (uint64_t)(int64_t)llrintf(t * 4611686018427387904.0f) << 2;  // t 2^64 without overflowing
  }
  uint64_t carry = 0;  // F[scanned] (+ the constant), valid after the first scan barrier

  NoiseCoef ncoef = noise_coef(0.f);
  u32x4 ctr;
  ctr.x = 0; ctr.y = 0; ctr.z = 0; ctr.w = 0;
  if (NOISE) {
    ncoef = noise_coef(p.sigma[frame]);
    const uint64_t gframe = p.frame_ids ? p.frame_ids[frame] : p.first_frame + (uint64_t)frame;
    ctr.y = (uint32_t)gframe; ctr.z = (uint32_t)(gframe >> 32);
    ctr.w = p.point ? (uint32_t)p.point[frame] : 0u;
  }

  // a = n + off = 8 A + rho: this thread owns rho = 2 sub, 2 sub + 1 of RUN consecutive A.
  // Its pair index at interval A is 4 A + sub - off/2; e makes (pair >> 2) even at the first
  // interval of every run, so iterations (2m, 2m+1) share Philox call c0 + 4m.
  const int sub = tid & 3, grp = tid >> 2;
  const int rho = 2 * sub - (ODD ? 1 : 0);          // a of the pair's first sample (-1: rho 7 of the interval before)
  const bool sha = ODD && sub == 0;
  const int qa0 = sha ? 8 : rho + 1, qb0 = rho + 2;  // Q index of the newest symbol's tap for samples a, b
  const int e = ((sub - ((off + 1) >> 1)) >> 2) & 1;
  const int A_last = (Ti - 1 + off) >> 3;
  const bool one_h = (p.n_h == 1);

  for (int k_lo = 0; k_lo <= A_last; k_lo += kFSeg) {
    if (k_lo > 0) {  // carry the newest kFLB slots of the previous segment to the front
      __syncthreads();
      float wv = 0.f; uint32_t fv = 0;
      if (tid < kFLB) { wv = w_s[kFSeg + tid]; fv = f_s[kFSeg + tid]; }
      __syncthreads();
      if (tid < kFLB) { w_s[tid] = wv; f_s[tid] = fv; }
    }
    // ---- (a) one-pass scan: w_k and F[k+1] for k in [k_lo, k_lo + kFSeg) -------------------
    // Thread t owns kPer = 20 consecutive symbols (five packed words): local total -> one block
    // scan of the 256 totals (two barriers per segment) -> second local pass from the thread's
    // base writes w and F with 128-bit stores (thread stride 20 words: conflict free).
    constexpr int kPer = kFSeg / kModThreads;
    static_assert(kPer % 4 == 0, "whole packed words per thread");
    const int k0 = k_lo + kPer * tid;
    uint32_t sv[kPer / 4];
#pragma unroll
    for (int j = 0; j < kPer / 4; ++j) {
      const int kk = k0 + 4 * j;
      sv[j] = 0u;
      if (sym_vec && kk + 3 < S) {
        sv[j] = *reinterpret_cast<const uint32_t*>(sy + kk);
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (kk + i < S) sv[j] |= (uint32_t)(uint8_t)sy[kk + i] << (8 * i);
      }
    }
    const int hi0 = one_h ? 0 : k0 % p.n_h;
    // thread total: byte sums by IDP.4A (even / odd positions apart for two-h patterns), one or two
    // 64-bit multiplies -- deliberately NOT the per-symbol products of the second pass
    uint64_t run = 0;
    if (p.n_h <= 2) {
      int se = 0, so = 0;
#pragma unroll
      for (int j = 0; j < kPer / 4; ++j) { se = dp4a_s8(sv[j], 0x00010001u, se); so = dp4a_s8(sv[j], 0x01000100u, so); }
      run = one_h ? (uint64_t)((int64_t)(se + so)) * p.hfix[0]
                  : (uint64_t)((int64_t)se) * p.hfix[hi0] + (uint64_t)((int64_t)so) * p.hfix[hi0 ^ 1];
    } else {
      int hi = hi0;
#pragma unroll 1
      for (int i = 0; i < kPer; ++i) {
        const int a = (int8_t)((sv[0] >> (8 * (i & 3))) & 0xFF);
        run += (uint64_t)((int64_t)a) * p.hfix[hi];
        if (++hi == p.n_h) hi = 0;
        if ((i & 3) == 3) {  // rotate the packed words so the loop indexes only sv[0]
          const uint32_t t0 = sv[0];
#pragma unroll
          for (int j = 0; j + 1 < kPer / 4; ++j) sv[j] = sv[j + 1];
          sv[kPer / 4 - 1] = t0;
        }
      }
    }
    uint64_t incl = run;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();  // also publishes base_s (first segment) / the look-back copy (later ones)
    if (k_lo == 0) {
      carry = base_s;
      if (tid < kFLB) { w_s[tid] = 0.f; f_s[tid] = (uint32_t)(carry >> 32); }  // symbols k < 0, F[k <= 0]
    }
    uint64_t base = carry + (incl - run), total = 0;
#pragma unroll
    for (int i = 0; i < kModThreads / 32; ++i) {
      const uint64_t t = warp_tot[i];
      if (i < wid) base += t;
      total += t;
    }
    {
      int hi = hi0;
#pragma unroll
      for (int j = 0; j < kPer / 4; ++j) {
        float wv[4];
        uint32_t fv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int a = (int8_t)((sv[j] >> (8 * i)) & 0xFF);
          wv[i] = (float)a * p.hw[hi];
          base += (uint64_t)((int64_t)a) * p.hfix[hi];
          fv[i] = (uint32_t)(base >> 32);  // F[k + 1]
          if (!one_h && ++hi == p.n_h) hi = 0;
        }
        const int slot = kFLB + kPer * tid + 4 * j;  // 16-byte aligned
        *reinterpret_cast<float4*>(w_s + slot) = make_float4(wv[0], wv[1], wv[2], wv[3]);
        *reinterpret_cast<uint4*>(f_s + slot) = make_uint4(fv[0], fv[1], fv[2], fv[3]);
      }
    }
    carry += total;
    __syncthreads();  // the segment is published; warp_tot is free for the next one

    // ---- (b) RUN consecutive intervals per thread, barrier free ---------------------------
    float qa[JT], qb[JT];  // (re)built per segment so that they are not live across the scan
#pragma unroll
    for (int j = 0; j < JT; ++j) {
      qa[j] = p.q[qa0 + 8 * j] * 0.78539816339744831f;  // Q * 2 pi / sps  (radians per unit w)
      qb[j] = p.q[qb0 + 8 * j] * 0.78539816339744831f;
    }
    for (int cs = 0; cs < kFSeg && k_lo + cs <= A_last; cs += kFChunk) {
      const int r0 = cs + grp * RUN + e;             // first interval of this run, relative to k_lo
      const bool live = k_lo + r0 <= A_last;
      if (!FUSE && !live) break;                     // this run (and every later one of the thread) is past the frame
      int mb = 0;                                    // FUSE: tile sample index of this thread's first pair
      if constexpr (FUSE) mb = fz.LB + 8 * (grp * RUN + e) + 2 * sub;
      if (live) {
      const float* wp = w_s + (r0 - JT + kFLB);      // symbols A0-JT .. A0+RUN-2
      const uint32_t* fp = f_s + (r0 - JT - 1 + kFLB);  // F[A0-JT ..]
      float w[JT + RUN - 1];   // sample b's window (and sample a's unless it sits one interval back)
      float wa[ODD ? JT + RUN - 1 : 1];
      uint32_t fa[ODD ? RUN : 1];
#pragma unroll
      for (int j = 0; j < JT + RUN - 1; ++j) w[j] = wp[j];
      if (ODD) {
        const float* wq = sha ? wp - 1 : wp;
        const uint32_t* fq = sha ? fp - 1 : fp;
#pragma unroll
        for (int j = 0; j < JT + RUN - 1; ++j) wa[j] = wq[j];
#pragma unroll
        for (int j = 0; j < RUN; ++j) fa[j] = fq[j];
      }
      const int n0 = 8 * (k_lo + r0) + rho - off;    // first sample of the first pair (< 0 before the frame starts)
      float4* o = out + (n0 >> 1);
#ifndef WF_HOSTEMU
      asm volatile("" : "+l"(o));  // keep the row pointer live: every store is then pointer + immediate
#endif
      const int P0 = n0 >> 1;                         // pair index; (P0 >> 2) is even by construction
      const uint32_t c0 = (uint32_t)(((P0 >> 3) << 2) | (P0 & 3));
      u32x4 r;
      r.x = r.y = r.z = r.w = 0;
      uint32_t m_min = 0xFFFFu;
#pragma unroll
      for (int i = 0; i < RUN; ++i) {
        float acc0 = 0.f, acc1 = 0.f;
        if (ODD) {
#pragma unroll
          for (int j = 0; j < JT; ++j) {
            acc0 = fmaf(wa[i + JT - 1 - j], qa[j], acc0);
            acc1 = fmaf(w[i + JT - 1 - j], qb[j], acc1);
          }
        } else {
#if WF_MOD_FFMA2
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < JT; ++j) {
          const float wj = w[i + JT - 1 - j];  // symbol A-1-j
          acc = ffma2(make_float2(wj, wj), make_float2(qa[j], qb[j]), acc);
        }
        acc0 = acc.x; acc1 = acc.y;
#else
#pragma unroll
        for (int j = 0; j < JT; ++j) {
          const float wj = w[i + JT - 1 - j];  // symbol A-1-j
          acc0 = fmaf(wj, qa[j], acc0);
          acc1 = fmaf(wj, qb[j], acc1);
        }
#endif
        }
        const float old = (float)(int32_t)fp[i] * 1.4629180792671596e-09f;  // F[A-JT] in radians, [-pi, pi)
        const float olda = ODD ? (float)(int32_t)fa[i] * 1.4629180792671596e-09f : old;
        float c0f, s0f, c1f, s1f;
        __sincosf(acc0 + olda, &s0f, &c0f);
        __sincosf(acc1 + old, &s1f, &c1f);
        if (NOISE) {
          if ((i & 1) == 0) {
            ctr.x = c0 + 2u * (uint32_t)i;
            r = philox4x32_10_rk(ctr, p.rk);
          }
          const uint32_t wa = (i & 1) ? r.z : r.x, wb = (i & 1) ? r.w : r.y;
          m_min = min(m_min, min(wa & 0xFFFFu, wb & 0xFFFFu));
          const float2 g0 = box_muller16(wa, noise_lgt(wa), ncoef), g1 = box_muller16(wb, noise_lgt(wb), ncoef);
          c0f += g0.x; s0f += g0.y; c1f += g1.x; s1f += g1.y;
        }
        if constexpr (FUSE) {
          const bool in = (unsigned)(n0 + 8 * i) < (unsigned)Ti;
          const int m = mb + 8 * i;
          *reinterpret_cast<float4*>(tile + m + ((m >> 5) << 1)) =
              in ? make_float4(c0f, s0f, c1f, s1f) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
          if ((unsigned)(n0 + 8 * i) < (unsigned)Ti) o[4 * i] = make_float4(c0f, s0f, c1f, s1f);
        }
      }
      if (NOISE && m_min == 0u) {  // some radius uniform of this run needs its tail refined: redo those pairs
#pragma unroll 1
        for (int i = 0; i < RUN; ++i) {
          ctr.x = c0 + 4u * (uint32_t)(i >> 1);
          const u32x4 rr = philox4x32_10_rk(ctr, p.rk);
          const uint32_t wa = (i & 1) ? rr.z : rr.x, wb = (i & 1) ? rr.w : rr.y;
          if (((wa & 0xFFFFu) == 0u || (wb & 0xFFFFu) == 0u) && (unsigned)(n0 + 8 * i) < (unsigned)Ti) {
            const float4 v4 = mod_f32s8_slow_pair<JT>(p, (sha ? wp - 1 : wp) + i, wp + i, (sha ? fp - 1 : fp)[i], fp[i], qa0,
                                                      qb0, ctr, i & 1, ncoef);
            if constexpr (FUSE) {
              const int m = mb + 8 * i;
              *reinterpret_cast<float4*>(tile + m + ((m >> 5) << 1)) = v4;
            } else {
              o[4 * i] = v4;
            }
          }
        }
      }
      } else if constexpr (FUSE) {  // run past the end of the frame: its samples are zeros for the bank
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
          const int m = mb + 8 * i;
          *reinterpret_cast<float4*>(tile + m + ((m >> 5) << 1)) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      if constexpr (FUSE) {
        __syncthreads();  // the chunk's 4096 samples (and the carried look-back) are in the tile
        const int q = tid - kFzThreads;
        const bool carry_lane = q >= 0 && 2 * q < fz.LB;
        float4 keep = make_float4(0.f, 0.f, 0.f, 0.f);
        if (carry_lane) keep = *reinterpret_cast<const float4*>(tile + fz_pos(8 * kFzChunk + 2 * q));
        if (tid < kFzThreads) fz_bank_phase2(fz, tile, tid, (k_lo + cs) - fz.CB, fz.z + frame * fz.ld_z * kFzNF);
        __syncthreads();  // bank reads done: the front of the tile may take the next chunk's look-back
        if (carry_lane) *reinterpret_cast<float4*>(tile + fz_pos(2 * q)) = keep;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Warp-autonomous fused modulate -> AWGN -> PAM bank (wf_cpm_modulate_bank, tune >= 1).
// Same arithmetic per sample and per (call, filter) as the block kernel above -- so z is still bit-identical
// to wf_cpm_modulate + wf_mf_bank -- but ONE WARP owns a frame and there is no block barrier anywhere:
//   * the frame is walked in chunks of CH = 32 R symbol intervals; per chunk the warp (a) scans the chunk's
//     symbols into w_k / F (a 5-step 64-bit shuffle scan; exact fixed point, so chunking changes nothing),
//     (b) produces the chunk's 8 CH samples into its private shared-memory tile with the 4-lanes-per-interval
//     register-window loop of the fast modulator (R / 2 passes of 8 intervals per lane), (c) runs the bank on
//     R consecutive calls per lane (all 32 lanes; R 8 + 2-sample segments, 16-byte pad: LDS.128 conflict free)
//     and stores 16 R bytes per lane straight to HBM, (d) carries the newest LB samples to the tile front.
//   * only __syncwarp orders the phases, so the warps of an SM drift apart and the schedulers always see a mix
//     of MUFU / integer work (produce) and FFMA2 streams (bank); nobody idles at a barrier and no warp waits
//     for another frame's scan.
//   * w_k / F live in a padded ring (one pad word per 8 slots): the 8 interval groups of a warp read their
//     windows from 8 different banks.
// Shared memory per warp: 6.3 KB (R = 2) or 10.9 KB (R = 4).
// ---------------------------------------------------------------------------
constexpr int kFwWarps = 4;
template <int R> struct FwGeom {
  static constexpr int CH = 32 * R;                                   // intervals (= calls) per chunk
  static constexpr int RS = 8 * R;                                    // samples per lane segment
  __host__ __device__ static constexpr int pos(int m) { return m + 2 * (int)((unsigned)m / (unsigned)RS); }
  static constexpr int SLOTS = kFLB + CH + 8;                         // w / F ring slots
  static constexpr int RINGW = SLOTS + SLOTS / 8 + 2;                 // + one pad word per 8 slots
  static constexpr int RINGW4 = (RINGW + 3) & ~3;
  static constexpr int TILE = ((kFzLBMax + 8 * CH) + 2 * ((kFzLBMax + 8 * CH) / RS) + 4 + 1) & ~1;  // float2 slots
  static constexpr size_t WARP_BYTES = (size_t)TILE * 8 + (size_t)RINGW4 * 8;
};
__device__ __forceinline__ int fw_pad(int s) { return s + (s >> 3); }

// NT1: taps of the second filter that can be non-zero (rho_1 is 65 taps zero-padded to 73: a zero tap adds
// exactly nothing, fma(x, 0, y) == y, so skipping it keeps the outputs bit-identical)
template <int R, int NT1>
static __device__ __forceinline__ void fw_bank_phase(const FuseParams& fz, const float2* tile, int t, int cbase, float2* zrow) {
  using G = FwGeom<R>;
  const int c0 = cbase + R * t;
  if (c0 + R - 1 < 0 || c0 >= fz.n_calls) return;
  constexpr int NV = (R - 1) * 8 + kFzNT;
  const float2* tap2 = reinterpret_cast<const float2*>(fz.tap2);
  float2 y[R][kFzNF];
#pragma unroll
  for (int e = 0; e < R; ++e)
#pragma unroll
    for (int f = 0; f < kFzNF; ++f) y[e][f] = make_float2(0.f, 0.f);
  const float2* seg = tile + t * (G::RS + 2);
#pragma unroll
  for (int v = 0; v < NV; v += 2) {
    float4 xx;
    if (v + 1 < NV) {
      xx = *reinterpret_cast<const float4*>(seg + G::pos(v));
    } else {
      const float2 x1 = seg[G::pos(v)];
      xx = make_float4(x1.x, x1.y, 0.f, 0.f);
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (v + s < NV) {
        const float2 x = s ? make_float2(xx.z, xx.w) : make_float2(xx.x, xx.y);
#pragma unroll
        for (int e = 0; e < R; ++e) {
          const int u = v + s - e * 8;
          if (u >= 0 && u < kFzNT) {
#pragma unroll
            for (int f = 0; f < kFzNF; ++f)
              if (f == 0 || u >= kFzNT - NT1) y[e][f] = ffma2(x, tap2[f * kFzNT + kFzNT - 1 - u], y[e][f]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < R; ++e) {
    const int c = c0 + e;
    if ((unsigned)c < (unsigned)fz.n_calls)
      *reinterpret_cast<float4*>(zrow + 2 * (int64_t)c) = make_float4(y[e][0].x, y[e][0].y, y[e][1].x, y[e][1].y);
  }
}

template <bool NOISE, int R, int MINB, int NT1>
__global__ void __launch_bounds__(kFwWarps * 32, MINB)
cpm_modulate_bank_warp_kernel(const __grid_constant__ ModParams<float> p, const __grid_constant__ FuseParams fz, int n_chunks) {
  using G = FwGeom<R>;
  constexpr int JT = 8, RUN = 8;
  constexpr int PASSES = G::CH / (8 * RUN);   // 8 interval groups per warp x RUN intervals per pass
  constexpr int PER = G::CH / 32;             // symbols per lane per chunk scan
  static_assert((kFLB - JT) % 8 == 0 && PASSES >= 1 && PER >= 2 && PER <= 4, "window bases must stay multiples of 8");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t frame = (int64_t)blockIdx.x * kFwWarps + wid;
  if (frame >= p.B) return;  // whole warp; only __syncwarp below
  unsigned char* wbase = smem_raw + (size_t)wid * G::WARP_BYTES;
  float2* tile = reinterpret_cast<float2*>(wbase);
  float* w_s = reinterpret_cast<float*>(wbase + (size_t)G::TILE * 8);     // padded slot(k) = fw_pad(k - K0 + kFLB) holds w_k
  uint32_t* f_s = reinterpret_cast<uint32_t*>(w_s + G::RINGW4);            // same slot holds F[k + 1]

  const int off = p.off;
  const int S = (int)p.S, Ti = (int)p.Tn;
  const int8_t* sy = p.syms + frame * (int64_t)S;
  float2* zrow = fz.z + frame * fz.ld_z * kFzNF;
  const bool one_h = (p.n_h == 1);

  // samples before the frame (tile look-back of chunk 0) are zero
  for (int m = 2 * lane; m < kFzLBMax + 16; m += 64) *reinterpret_cast<float4*>(tile + G::pos(m)) = make_float4(0.f, 0.f, 0.f, 0.f);

  // per-frame constant folded into the prefix sum (see the block kernel): computed by lane 0, broadcast
  uint64_t carry = 0;
  if (lane == 0) {
    float c0 = 0.f;
    for (int k = 0; k < S && (k + 1) * 8 < off; ++k) {
      int idx = off - (k + 1) * 8;
      if (idx > p.K) idx = p.K;
      c0 = fmaf((float)sy[k] * p.hw[k % p.n_h], p.q[idx], c0);
    }
    float t = fmaf(c0, -0.125f, p.phase0_turns);
    t -= rintf(t);
    carry = (uint64_t)(int64_t)llrintf(t * 4611686018427387904.0f) << 2;
  }
  carry = __shfl_sync(0xFFFFFFFFu, carry, 0);
  if (lane < kFLB) { w_s[fw_pad(lane)] = 0.f; f_s[fw_pad(lane)] = (uint32_t)(carry >> 32); }  // symbols k < 0, F[k <= 0]

  NoiseCoef ncoef = noise_coef(0.f);
  u32x4 ctr;
  ctr.x = 0; ctr.y = 0; ctr.z = 0; ctr.w = 0;
  if (NOISE) {
    ncoef = noise_coef(p.sigma[frame]);
    const uint64_t gframe = p.frame_ids ? p.frame_ids[frame] : p.first_frame + (uint64_t)frame;
    ctr.y = (uint32_t)gframe; ctr.z = (uint32_t)(gframe >> 32);
    ctr.w = p.point ? (uint32_t)p.point[frame] : 0u;
  }
  const int sub = lane & 3, grp = lane >> 2;
  const int rho = 2 * sub;
  const int qa0 = rho + 1, qb0 = rho + 2;
  const int A_last = (Ti - 1 + off) >> 3;
  float qa[JT], qb[JT];
#pragma unroll
  for (int j = 0; j < JT; ++j) {
    qa[j] = p.q[qa0 + 8 * j] * 0.78539816339744831f;  // Q * 2 pi / sps  (radians per unit w)
    qb[j] = p.q[qb0 + 8 * j] * 0.78539816339744831f;
  }

  // symbols of the chunk this lane scans: PER consecutive bytes, fetched one chunk ahead
  auto fetch = [&](int K0) {
    uint32_t v = 0u;
    const int k0 = K0 + PER * lane;
#pragma unroll
    for (int i = 0; i < PER; ++i)
      if (k0 + i < S) v |= (uint32_t)(uint8_t)sy[k0 + i] << (8 * i);
    return v;
  };
  uint32_t sv_next = fetch(0);

  for (int jc = 0; jc < n_chunks; ++jc) {
    const int K0 = jc * G::CH;
    // ---- (a) scan: w_k and F[k + 1] for k in [K0, K0 + CH) ---------------------------------------
    const uint32_t sv = sv_next;
    sv_next = fetch(K0 + G::CH);
    if (jc > 0) {  // the newest kFLB slots of the previous chunk become the look-back (its readers are done: (e) ended with a __syncwarp)
      float wv = 0.f; uint32_t fv = 0;
      if (lane < kFLB) { wv = w_s[fw_pad(G::CH + lane)]; fv = f_s[fw_pad(G::CH + lane)]; }
      __syncwarp();
      if (lane < kFLB) { w_s[fw_pad(lane)] = wv; f_s[fw_pad(lane)] = fv; }
    }
    {
      const int k0 = K0 + PER * lane;
      int hi = one_h ? 0 : k0 % p.n_h;
      uint64_t run = 0;
      int hh = hi;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const int a = (int8_t)((sv >> (8 * i)) & 0xFF);
        run += (uint64_t)((int64_t)a) * p.hfix[hh];
        if (!one_h && ++hh == p.n_h) hh = 0;
      }
      uint64_t incl = run;
#pragma unroll
      for (int s = 1; s < 32; s <<= 1) {
        const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, s);
        if (lane >= s) incl += t;
      }
      uint64_t base = carry + (incl - run);
      carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const int a = (int8_t)((sv >> (8 * i)) & 0xFF);
        base += (uint64_t)((int64_t)a) * p.hfix[hi];
        const int slot = fw_pad(kFLB + PER * lane + i);
        w_s[slot] = (float)a * p.hw[hi];
        f_s[slot] = (uint32_t)(base >> 32);  // F[k + 1]
        if (!one_h && ++hi == p.n_h) hi = 0;
      }
    }
    __syncwarp();
    // ---- (b) produce the chunk's samples into the tile -----------------------------------------
#pragma unroll 1
    for (int ps = 0; ps < PASSES; ++ps) {
      const int r0 = ps * 8 * RUN + grp * RUN;       // first interval of this lane's run, relative to K0
      const bool live = K0 + r0 <= A_last;
      const int mb = fz.LB + 8 * r0 + 2 * sub;       // tile sample index of this lane's first pair
      if (live) {
        const int wb = fw_pad(r0 - JT + kFLB);       // a multiple of 9: symbols A0-JT .. sit at wb + j + (j >> 3)
        const int fb = r0 - JT - 1 + kFLB;           // == 7 (mod 8): F[A0-JT + i] at fw_pad(fb + i)
        float w[JT + RUN - 1];
#pragma unroll
        for (int j = 0; j < JT + RUN - 1; ++j) w[j] = w_s[wb + j + (j >> 3)];
        uint32_t fpv[RUN];
#pragma unroll
        for (int i = 0; i < RUN; ++i) fpv[i] = f_s[fw_pad(fb + i)];
        const int n0 = 8 * (K0 + r0) + rho - off;    // first sample of the first pair (< 0 before the frame starts)
        const int P0 = n0 >> 1;
        const uint32_t c0 = (uint32_t)(((P0 >> 3) << 2) | (P0 & 3));
        u32x4 r;
        r.x = r.y = r.z = r.w = 0;
        uint32_t m_min = 0xFFFFu;
#pragma unroll
        for (int i = 0; i < RUN; ++i) {
          float2 acc = make_float2(0.f, 0.f);
#pragma unroll
          for (int j = 0; j < JT; ++j) {
            const float wj = w[i + JT - 1 - j];  // symbol A-1-j
            acc = ffma2(make_float2(wj, wj), make_float2(qa[j], qb[j]), acc);
          }
          const float old = (float)(int32_t)fpv[i] * 1.4629180792671596e-09f;  // F[A-JT] in radians, [-pi, pi)
          float c0f, s0f, c1f, s1f;
          __sincosf(acc.x + old, &s0f, &c0f);
          __sincosf(acc.y + old, &s1f, &c1f);
          if (NOISE) {
            if ((i & 1) == 0) {
              ctr.x = c0 + 2u * (uint32_t)i;
              r = philox4x32_10_rk(ctr, p.rk);
            }
            const uint32_t wa = (i & 1) ? r.z : r.x, wbw = (i & 1) ? r.w : r.y;
            m_min = min(m_min, min(wa & 0xFFFFu, wbw & 0xFFFFu));
            const float2 g0 = box_muller16(wa, noise_lgt(wa), ncoef), g1 = box_muller16(wbw, noise_lgt(wbw), ncoef);
            c0f += g0.x; s0f += g0.y; c1f += g1.x; s1f += g1.y;
          }
          const bool in = (unsigned)(n0 + 8 * i) < (unsigned)Ti;
          const int m = mb + 8 * i;
          *reinterpret_cast<float4*>(tile + G::pos(m)) = in ? make_float4(c0f, s0f, c1f, s1f) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (NOISE && m_min == 0u) {  // some radius uniform of this run needs its tail refined: redo those pairs
#pragma unroll 1
          for (int i = 0; i < RUN; ++i) {
            ctr.x = c0 + 4u * (uint32_t)(i >> 1);
            const u32x4 rr = philox4x32_10_rk(ctr, p.rk);
            const uint32_t wa = (i & 1) ? rr.z : rr.x, wbw = (i & 1) ? rr.w : rr.y;
            if (((wa & 0xFFFFu) == 0u || (wbw & 0xFFFFu) == 0u) && (unsigned)(n0 + 8 * i) < (unsigned)Ti) {
              float wl[JT];
#pragma unroll
              for (int j = 0; j < JT; ++j) wl[j] = w_s[fw_pad(r0 - JT + kFLB + i + j)];
              const uint32_t fv = f_s[fw_pad(fb + i)];
              const float4 v4 = mod_f32s8_slow_pair<JT>(p, wl, wl, fv, fv, qa0, qb0, ctr, i & 1, ncoef);
              *reinterpret_cast<float4*>(tile + G::pos(mb + 8 * i)) = v4;
            }
          }
        }
      } else {  // run past the end of the frame: its samples are zeros for the bank
#pragma unroll
        for (int i = 0; i < RUN; ++i) *reinterpret_cast<float4*>(tile + G::pos(mb + 8 * i)) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    __syncwarp();  // the chunk's samples (and the carried look-back) are in the tile
    // ---- (c) bank: R consecutive calls per lane, straight to HBM ---------------------------------
    fw_bank_phase<R, NT1>(fz, tile, lane, K0 - fz.CB, zrow);
    // ---- (d) carry the newest LB samples to the front of the tile -------------------------------
    float4 keep[(kFzLBMax / 2 + 31) / 32];
#pragma unroll
    for (int q = 0; q < (kFzLBMax / 2 + 31) / 32; ++q) {
      const int u = lane + 32 * q;
      keep[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (2 * u < fz.LB) keep[q] = *reinterpret_cast<const float4*>(tile + G::pos(8 * G::CH + 2 * u));
    }
    __syncwarp();  // bank reads done: the front of the tile may take the next chunk's look-back
#pragma unroll
    for (int q = 0; q < (kFzLBMax / 2 + 31) / 32; ++q) {
      const int u = lane + 32 * q;
      if (2 * u < fz.LB) *reinterpret_cast<float4*>(tile + G::pos(2 * u)) = keep[q];
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// Stand-alone AWGN: in-place add / fill.  One thread per Philox call (= the two sample
// pairs p, p + 4 that share it, see common.cuh), grid-stride.
// ---------------------------------------------------------------------------
template <typename T, bool ADD>
__global__ void __launch_bounds__(256)
awgn_kernel(T* __restrict__ sig, int64_t B, int64_t Tn, int64_t ld, const float* __restrict__ sigma,
            const int32_t* __restrict__ point, uint64_t seed, uint64_t first_frame, const uint64_t* __restrict__ frame_ids) {
  const int64_t pairs_per_frame = (Tn + 1) >> 1;
  const int64_t calls_per_frame = ((pairs_per_frame + 7) >> 3) << 2;
  const int64_t total = B * calls_per_frame;
  const bool vec_ok = ((ld & 1) == 0) && ((reinterpret_cast<uintptr_t>(sig) & (4 * sizeof(T) - 1)) == 0);
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / calls_per_frame;
    const int64_t call = i - b * calls_per_frame;
    const NoiseCoef nc = noise_coef(sigma ? sigma[b] : 1.0f);
    const uint32_t pt = point ? (uint32_t)point[b] : 0u;
    const u32x4 ctr = noise_counter((uint64_t)call, frame_ids ? frame_ids[b] : first_frame + (uint64_t)b, pt);
    const u32x4 r = philox4x32_10(ctr, k0, k1);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int64_t pr = ((call >> 2) << 3) + (call & 3) + 4 * half;
      if (pr >= pairs_per_frame) continue;
      const int64_t n = pr << 1;
      float2 g0, g1;
      noise_pair(half ? r.z : r.x, half ? r.w : r.y, half, ctr, k0, k1, nc, g0, g1);
      T* row = sig + 2 * (b * ld + n);
      if (vec_ok && n + 1 < Tn && sizeof(T) == 4) {
        float4* q = reinterpret_cast<float4*>(row);
        float4 v = ADD ? *q : make_float4(0.f, 0.f, 0.f, 0.f);
        v.x += g0.x; v.y += g0.y; v.z += g1.x; v.w += g1.y;
        *q = v;
      } else {
        row[0] = (ADD ? row[0] : (T)0) + (T)g0.x;
        row[1] = (ADD ? row[1] : (T)0) + (T)g0.y;
        if (n + 1 < Tn) {
          row[2] = (ADD ? row[2] : (T)0) + (T)g1.x;
          row[3] = (ADD ? row[3] : (T)0) + (T)g1.y;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// frequency_modulate (cpm/modulate.py:28-54) for arbitrary per-sample frequency
// pulses: phase[n] = ((sum_{m<=n} f[m]) mod sps) * 2pi/sps + initial_phase, and
// phase_modulate (:12-25): exp(j * sensitivity * phase).  One block per frame,
// float64 block scan in chunks of 1024 samples.
// ---------------------------------------------------------------------------
template <typename TO>
__global__ void __launch_bounds__(256)
frequency_modulate_kernel(const double* __restrict__ f, int64_t Tn, int64_t ld_in, int sps, double initial_phase,
                          TO* __restrict__ out, int64_t ld_out) {
  __shared__ double warp_tot[8];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const double* row = f + (int64_t)blockIdx.x * ld_in;
  TO* orow = out + 2 * (int64_t)blockIdx.x * ld_out;
  double carry = 0.0;
  const double k = 6.283185307179586476925286766559 / (double)sps;
  for (int64_t base = 0; base < Tn; base += 1024) {
    double v[4], run = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t n = base + 4 * tid + i;
      run += (n < Tn) ? row[n] : 0.0;
      v[i] = run;
    }
    double incl = run;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      double t = __shfl_up_sync(0xFFFFFFFFu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    double before = carry, total = 0.0;
    for (int i = 0; i < 8; ++i) { if (i < wid) before += warp_tot[i]; total += warp_tot[i]; }
    before += incl - run;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t n = base + 4 * tid + i;
      if (n < Tn) {
        double revs = before + v[i];
        revs -= floor(revs / sps) * sps;  // Python's % : result in [0, sps)
        double sn, cs;
        sincos(revs * k + initial_phase, &sn, &cs);
        orow[2 * n] = (TO)cs;
        orow[2 * n + 1] = (TO)sn;
      }
    }
    carry += total;
    carry -= floor(carry / sps) * sps;
    __syncthreads();
  }
}

template <typename TO>
__global__ void __launch_bounds__(256)
phase_modulate_kernel(const double* __restrict__ ph, int64_t n, double sensitivity, TO* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double sn, cs;
    sincos(sensitivity * ph[i], &sn, &cs);
    out[2 * i] = (TO)cs;
    out[2 * i + 1] = (TO)sn;
  }
}

template <typename T>
static int fill_mod_params(ModParams<T>& p, const int8_t* d_syms, int64_t B, int64_t S, const double* mod_index, int n_h,
                           const double* pulse, int K, int sps, double initial_phase, double derotate,
                           void* d_out, int64_t ld_out, const float* d_sigma, const int32_t* d_point,
                           uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids) {
  memset(&p, 0, sizeof(p));
  p.syms = d_syms; p.out = d_out; p.sigma = d_sigma; p.point = d_point; p.frame_ids = d_frame_ids;
  p.B = B; p.S = S; p.Tn = (S + 1) * sps; p.ld_out = ld_out;
  p.seed = seed; p.first_frame = first_frame; p.rk = philox_round_keys(seed);
  p.n_h = n_h; p.K = K; p.sps = sps; p.off = (K - 1) / 2;
  p.J = (K - 1 + sps - 1) / sps;
  if (p.J < 1) p.J = 1;
  WF_REQUIRE(p.J * sps + 1 <= kQMax, "pulse too long: K=%d at sps=%d needs %d table entries (max %d)", K, sps, p.J * sps + 1, kQMax);
  WF_REQUIRE(p.J + p.off / sps + 2 <= 512, "pulse look-back too long for the symbol ring");
  // cumulative pulse in long double, then narrowed once
  long double acc = 0;
  p.q[0] = (T)0;
  for (int i = 0; i < K; ++i) { acc += (long double)pulse[i]; p.q[i + 1] = (T)acc; }
  const long double qK = acc;
  for (int i = K + 1; i < kQMax; ++i) p.q[i] = (T)qK;
  for (int i = 0; i < n_h; ++i) {
    p.hw[i] = (T)mod_index[i];
    long double turns = (long double)mod_index[i] * qK / (long double)sps;  // per unit alpha
    turns -= floorl(turns);                                                  // [0,1)
    long double scaled = turns * 18446744073709551616.0L;                    // 2^64
    uint64_t fx = (scaled >= 18446744073709551615.0L) ? 0ull : (uint64_t)(scaled + 0.5L);
    p.hfix[i] = fx;
  }
  const double two_pi = 6.283185307179586476925286766559;
  double ph0 = (initial_phase + derotate) / two_pi;
  p.phase0_turns = (T)(ph0 - floor(ph0 + 0.5));
  p.inv_sps = (T)(1.0 / (double)sps);
  return WF_OK;
}

template <typename T>
static int launch_modulate(const int8_t* d_syms, int64_t B, int64_t S, const double* mod_index, int n_h,
                           const double* pulse, int K, int sps, double initial_phase, double derotate,
                           void* d_out, int64_t ld_out, const float* d_sigma, const int32_t* d_point,
                           uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, cudaStream_t stream) {
  ModParams<T> p;
  int rcp = fill_mod_params<T>(p, d_syms, B, S, mod_index, n_h, pulse, K, sps, initial_phase, derotate, d_out, ld_out,
                               d_sigma, d_point, seed, first_frame, d_frame_ids);
  if (rcp) return rcp;
  const bool noise = d_sigma != nullptr;
  const dim3 grid((unsigned)B), block(kModThreads);
  const bool fast8 = (sps == 8) && (p.off % 2 == 0);
  const bool fast8_odd = (sps == 8) && (p.off % 2 == 1) && p.J == 3;  // PCM/FM (K = 24)
#define WF_MOD_LAUNCH(SPS_, JT_)                                                            \
  do {                                                                                      \
    if (noise) cpm_modulate_kernel<T, SPS_, JT_, true><<<grid, block, 0, stream>>>(p);     \
    else cpm_modulate_kernel<T, SPS_, JT_, false><<<grid, block, 0, stream>>>(p);          \
  } while (0)
  if (sizeof(T) == 4 && ((fast8 && (p.J == 16 || p.J == 8 || p.J == 3 || p.J == 1)) || fast8_odd) && (ld_out & 1) == 0 &&
      (reinterpret_cast<uintptr_t>(d_out) & 15) == 0 && p.Tn < 0x7FFFF000) {
    const ModParams<float>& pf = reinterpret_cast<const ModParams<float>&>(p);  // T == float on this branch
#define WF_MODF_LAUNCH(JT_, RUN_, MINB_, ODD_)                                                                   \
  do {                                                                                                          \
    if (noise) cpm_modulate_f32s8_kernel<JT_, true, RUN_, MINB_, ODD_><<<grid, block, kFSmemBytes, stream>>>(pf, FuseNone{0});  \
    else cpm_modulate_f32s8_kernel<JT_, false, RUN_, MINB_, ODD_><<<grid, block, kFSmemBytes, stream>>>(pf, FuseNone{0});       \
  } while (0)
    // RUN = 8 intervals per thread at 4 resident blocks/SM (64 registers) measured fastest on B200
    // among RUN in {4, 8, 16} x blocks/SM in {3, 4, 5} (profiles/r1f_kbench_mod_variants.log)
    if (fast8_odd) WF_MODF_LAUNCH(3, 8, 4, true);
    else if (p.J == 16) WF_MODF_LAUNCH(16, 8, 2, false);  // SOQPSK-B (129-tap pulse): 32 Q coefficients per thread, 2 blocks/SM
    else if (p.J == 8) WF_MODF_LAUNCH(8, 8, 4, false);
    else if (p.J == 3) WF_MODF_LAUNCH(3, 8, 4, false);
    else WF_MODF_LAUNCH(1, 8, 4, false);
#undef WF_MODF_LAUNCH
    WF_LAUNCH_CHECK("cpm_modulate_f32s8_kernel");
    return WF_OK;
  }
  if (fast8 && p.J == 8) WF_MOD_LAUNCH(8, 8);
  else if (fast8 && p.J == 3) WF_MOD_LAUNCH(8, 3);
  else if (fast8 && p.J == 1) WF_MOD_LAUNCH(8, 1);
  else if (sps == 8) WF_MOD_LAUNCH(8, 0);
  else if (sps == 10) WF_MOD_LAUNCH(10, 0);
  else WF_MOD_LAUNCH(0, 0);
#undef WF_MOD_LAUNCH
  WF_LAUNCH_CHECK("cpm_modulate_kernel");
  return WF_OK;
}

// Geometry of the fused kernel for a bank described the way wf_mf_bank takes it.  Returns false when
// the shape is not the one the fused kernel is built for (callers then run the two-kernel path).
static bool fuse_geometry(int K, int sps, int n_filt, int n_taps, int center, int64_t first_instant, int64_t S,
                          int64_t n_calls, int& CB, int& LB) {
  if (sps != 8 || n_filt != kFzNF || n_taps != kFzNT || K < 2) return false;
  const int off = (K - 1) / 2, J = (K - 1 + sps - 1) / sps;
  if (J != 8 || (off % 8) != 0) return false;            // JT = 8 instance; off % 8 == 0 keeps every lane's e at 0
  if (first_instant < 0 || first_instant >= 4096 || center < 0 || center >= kFzNT) return false;
  if (S < 1 || (S + 1) * 8 >= 0x7FFFF000LL || n_calls < 0 || n_calls >= 0x7FFFF000LL) return false;
  const int x = (int)first_instant + center + off - 7;    // window of call c ends at sample 8 c + x + 7 - off
  CB = x > 0 ? (x + 7) / 8 : 0;
  LB = kFzNT - 1 + 8 * CB - off - (int)first_instant - center;
  if (LB < 0 || LB > kFzLBMax || (LB & 1)) return false;
  const int64_t A_last = ((S + 1) * 8 - 1 + off) >> 3;
  const int64_t n_chunks = A_last / kFzChunk + 1;
  return n_calls <= n_chunks * kFzChunk - CB;
}

static int launch_modulate_bank(const int8_t* d_syms, int64_t B, int64_t S, const double* mod_index, int n_h,
                                const double* pulse, int K, int sps, double initial_phase, double derotate,
                                const float* d_sigma, const int32_t* d_point, uint64_t seed, uint64_t first_frame,
                                const uint64_t* d_frame_ids,
                                const double* taps_re, int n_filt, int n_taps, int center, int64_t first_instant,
                                int64_t n_calls, void* d_z, int64_t ld_z, int tune, cudaStream_t stream) {
  int CB = 0, LB = 0;
  if (!fuse_geometry(K, sps, n_filt, n_taps, center, first_instant, S, n_calls, CB, LB)) {
    return fail(WF_ERR_UNSUPPORTED, "wf_cpm_modulate_bank: shape not covered by the fused kernel (see wf_cpm_modulate_bank_supported)");
  }
  ModParams<float> p;
  int rc = fill_mod_params<float>(p, d_syms, B, S, mod_index, n_h, pulse, K, sps, initial_phase, derotate, nullptr,
                                  (S + 1) * sps, d_sigma, d_point, seed, first_frame, d_frame_ids);
  if (rc) return rc;
  FuseParams fz;
  memset(&fz, 0, sizeof(fz));
  fz.z = reinterpret_cast<float2*>(d_z); fz.ld_z = ld_z; fz.n_calls = (int)n_calls; fz.CB = CB; fz.LB = LB;
  for (int i = 0; i < kFzNF * kFzNT; ++i) fz.tap2[2 * i] = fz.tap2[2 * i + 1] = (float)taps_re[i];
  if (tune >= 1) {  // warp-autonomous kernel: R = 2 calls per lane, 64-interval chunks, 8 blocks of 4 warps per SM
    // (R = 4 / 128-interval chunks / 5 blocks per SM measured the same: 7.00 against 6.94 ms per 65536 frames, r2c)
    const int R = 2, CH = 32 * R;
    bool pad8 = true;  // the second filter ends in 8 exact zeros (rho_1 padded to rho_0's length)?
    for (int t = kFzNT - 8; t < kFzNT; ++t) pad8 = pad8 && (float)taps_re[kFzNT + t] == 0.0f;
    const int64_t A_last = ((S + 1) * 8 - 1 + p.off) >> 3;
    int64_t nch = A_last / CH + 1;
    const int64_t need = (n_calls + CB + CH - 1) / CH;
    if (need > nch) nch = need;
    const dim3 gridw((unsigned)((B + kFwWarps - 1) / kFwWarps)), blockw(kFwWarps * 32);
#define WF_FW_LAUNCH(NOISE_, R_, MINB_, NT1_)                                                                     \
  do {                                                                                                            \
    auto kfn = cpm_modulate_bank_warp_kernel<NOISE_, R_, MINB_, NT1_>;                                             \
    const size_t sm = FwGeom<R_>::WARP_BYTES * kFwWarps;                                                          \
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);              \
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(cpm_modulate_bank_warp_kernel)");             \
    kfn<<<gridw, blockw, sm, stream>>>(p, fz, (int)nch);                                                           \
  } while (0)
    if (pad8) { if (d_sigma) WF_FW_LAUNCH(true, 2, 8, kFzNT - 8); else WF_FW_LAUNCH(false, 2, 8, kFzNT - 8); }
    else { if (d_sigma) WF_FW_LAUNCH(true, 2, 8, kFzNT); else WF_FW_LAUNCH(false, 2, 8, kFzNT); }
#undef WF_FW_LAUNCH
    WF_LAUNCH_CHECK("cpm_modulate_bank_warp_kernel");
    return WF_OK;
  }
  const dim3 grid((unsigned)B), block(kModThreads);
  // 2048-interval scan segments keep the scan arrays at 16.6 KB (8 B per symbol interval): with the 35.7 KB sample
  // tile that is 52 KB per block, four blocks per SM at 64 registers -- the occupancy the produce phase wants.
  // Measured alternatives (profiles/r1p_fused_variants.log): 5120-interval segments at 2 blocks/SM 8.81 ms, 1024 at 3
  // 8.36, 1024 at 4 8.28, this one 8.11 ms per 65536 frames; the bank phase split by filter over all eight warps
  // 8.34; eight modulator warps + four dedicated bank warps on two tiles (named barriers, 2 blocks/SM) 8.36.
#define WF_FZ_LAUNCH(NOISE_)                                                                                \
  do {                                                                                                      \
    auto kfn = cpm_modulate_f32s8_kernel<8, NOISE_, 8, 4, false, true, kFzSeg>;                             \
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFzSmemBytes); \
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(cpm_modulate_f32s8_kernel fused)");    \
    kfn<<<grid, block, kFzSmemBytes, stream>>>(p, fz);                                                     \
  } while (0)
  if (d_sigma) WF_FZ_LAUNCH(true);
  else WF_FZ_LAUNCH(false);
#undef WF_FZ_LAUNCH
  WF_LAUNCH_CHECK("cpm_modulate_f32s8_kernel (fused bank)");
  return WF_OK;
}

}  // namespace wf

extern "C" {

int wf_cpm_modulate_bank_supported(int K, int sps, int n_filt, int n_taps, int center, int64_t first_instant, int64_t S,
                                   int64_t n_calls) {
  int CB, LB;
  return wf::fuse_geometry(K, sps, n_filt, n_taps, center, first_instant, S, n_calls, CB, LB) ? 1 : 0;
}

int wf_cpm_modulate_bank(const int8_t* d_syms, int64_t B, int64_t S, const double* mod_index, int n_h,
                         const double* pulse, int K, int sps, double initial_phase, double derotate,
                         const float* d_sigma, const int32_t* d_point, uint64_t seed, uint64_t first_frame,
                         const uint64_t* d_frame_ids,
                         const double* taps_re, int n_filt, int n_taps, int center, int64_t first_instant,
                         int64_t n_calls, void* d_z, int64_t ld_z, int tune, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && S >= 0 && n_calls >= 0, "negative shape");
  WF_REQUIRE(K >= 1 && pulse, "empty pulse filter");
  WF_REQUIRE(n_h >= 1 && n_h <= kMaxH && mod_index, "n_h=%d outside 1..%d", n_h, kMaxH);
  WF_REQUIRE(taps_re && n_filt >= 1 && n_taps >= 1, "bad filter bank shape");
  WF_REQUIRE(ld_z >= n_calls, "ld_z smaller than n_calls");
  WF_REQUIRE(B <= 0x7FFFFFFF, "batch too large for one launch");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || n_calls == 0) return WF_OK;
  WF_REQUIRE(d_syms && d_z, "NULL device pointer");
  WF_REQUIRE((reinterpret_cast<uintptr_t>(d_z) & 15) == 0, "d_z must be 16-byte aligned");
  return launch_modulate_bank(d_syms, B, S, mod_index, n_h, pulse, K, sps, initial_phase, derotate, d_sigma, d_point, seed,
                              first_frame, d_frame_ids, taps_re, n_filt, n_taps, center, first_instant, n_calls, d_z, ld_z,
                              tune, as_stream(stream));
}

int wf_cpm_modulate(const int8_t* d_syms, int64_t B, int64_t S, const double* mod_index, int n_h,
                    const double* pulse, int K, int sps, double initial_phase, double derotate,
                    int dtype, void* d_out, int64_t ld_out, const float* d_sigma, const int32_t* d_point,
                    uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && S >= 0, "negative batch or symbol count");
  WF_REQUIRE(sps >= 1 && sps <= 64, "sps=%d outside 1..64", sps);
  WF_REQUIRE(K >= 1 && pulse, "empty pulse filter");
  WF_REQUIRE(n_h >= 1 && n_h <= kMaxH && mod_index, "n_h=%d outside 1..%d", n_h, kMaxH);
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  WF_REQUIRE(ld_out >= (S + 1) * sps, "ld_out smaller than (S+1)*sps");
  WF_REQUIRE((S + 1) * sps < 0x7FFFFFFF, "frame longer than 2^31 samples");
  WF_REQUIRE(B <= 0x7FFFFFFF, "batch too large for one launch");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  WF_REQUIRE(d_syms || S == 0, "d_syms is NULL");
  WF_REQUIRE(d_out, "d_out is NULL");
  if (dtype == WF_F32)
    return launch_modulate<float>(d_syms, B, S, mod_index, n_h, pulse, K, sps, initial_phase, derotate, d_out, ld_out,
                                  d_sigma, d_point, seed, first_frame, d_frame_ids, as_stream(stream));
  return launch_modulate<double>(d_syms, B, S, mod_index, n_h, pulse, K, sps, initial_phase, derotate, d_out, ld_out,
                                 d_sigma, d_point, seed, first_frame, d_frame_ids, as_stream(stream));
}

static int awgn_common(void* d_sig, int dtype, int64_t B, int64_t T, int64_t ld, const float* d_sigma,
                       const int32_t* d_point, uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream,
                       bool add) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld >= T, "bad AWGN shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || T == 0) return WF_OK;
  WF_REQUIRE(d_sig, "signal pointer is NULL");
  const int64_t total = B * ((((T + 1) / 2 + 7) / 8) * 4);  // one thread per Philox call
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = as_stream(stream);
  if (dtype == WF_F32) {
    if (add) awgn_kernel<float, true><<<(unsigned)blocks, 256, 0, st>>>((float*)d_sig, B, T, ld, d_sigma, d_point, seed, first_frame, d_frame_ids);
    else awgn_kernel<float, false><<<(unsigned)blocks, 256, 0, st>>>((float*)d_sig, B, T, ld, d_sigma, d_point, seed, first_frame, d_frame_ids);
  } else {
    if (add) awgn_kernel<double, true><<<(unsigned)blocks, 256, 0, st>>>((double*)d_sig, B, T, ld, d_sigma, d_point, seed, first_frame, d_frame_ids);
    else awgn_kernel<double, false><<<(unsigned)blocks, 256, 0, st>>>((double*)d_sig, B, T, ld, d_sigma, d_point, seed, first_frame, d_frame_ids);
  }
  WF_LAUNCH_CHECK("awgn_kernel");
  return WF_OK;
}

int wf_awgn_add(void* d_sig, int dtype, int64_t B, int64_t T, int64_t ld, const float* d_sigma,
                const int32_t* d_point, uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream) {
  WF_REQUIRE(d_sigma, "d_sigma is NULL");
  return awgn_common(d_sig, dtype, B, T, ld, d_sigma, d_point, seed, first_frame, d_frame_ids, stream, true);
}

int wf_awgn_fill(float* d_out, int64_t B, int64_t T, const float* d_sigma, const int32_t* d_point,
                 uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream) {
  return awgn_common(d_out, WF_F32, B, T, T, d_sigma, d_point, seed, first_frame, d_frame_ids, stream, false);
}

int wf_frequency_modulate(const double* d_freq, int64_t B, int64_t T, int64_t ld_in, int sps, double initial_phase,
                          int dtype, void* d_out, int64_t ld_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && T >= 0 && ld_in >= T && ld_out >= T, "bad frequency_modulate shape");
  WF_REQUIRE(sps >= 1, "sps must be positive");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || T == 0) return WF_OK;
  WF_REQUIRE(d_freq && d_out, "NULL device pointer");
  if (dtype == WF_F32)
    frequency_modulate_kernel<float><<<(unsigned)B, 256, 0, as_stream(stream)>>>(d_freq, T, ld_in, sps, initial_phase, (float*)d_out, ld_out);
  else
    frequency_modulate_kernel<double><<<(unsigned)B, 256, 0, as_stream(stream)>>>(d_freq, T, ld_in, sps, initial_phase, (double*)d_out, ld_out);
  WF_LAUNCH_CHECK("frequency_modulate_kernel");
  return WF_OK;
}

int wf_phase_modulate(const double* d_phase, int64_t n, double sensitivity, int dtype, void* d_out, void* stream) {
  using namespace wf;
  WF_REQUIRE(n >= 0, "negative size");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  int rc;
  if ((rc = require_device())) return rc;
  if (n == 0) return WF_OK;
  WF_REQUIRE(d_phase && d_out, "NULL device pointer");
  int64_t blocks = (n + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (dtype == WF_F32)
    phase_modulate_kernel<float><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(d_phase, n, sensitivity, (float*)d_out);
  else
    phase_modulate_kernel<double><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(d_phase, n, sensitivity, (double*)d_out);
  WF_LAUNCH_CHECK("phase_modulate_kernel");
  return WF_OK;
}

}  // extern "C"
