// cpmdet.cu -- A11/A12: pulse-truncated coherent CPM trellis detection
// (PCM/FM, ARTM multi-h CPM, any M-ary single/multi-h CPM).  The reference has NO detector
// for these waveforms (waveforms.viterbi exports only SOQPSKTrellisDetector); the design
// generalises its pulse-truncation bank + Re{e^{-j theta} z} metric
// (examples/soqpsk_detection.py:134-156, viterbi/algorithm.py:57-63) and is specified by the
// float64 statement oracle/cpm_generic.py (parity unpinned).
//
// State s = v * M^(Lp-1) + hist with v the phase state (theta = pi v / p, v in Z_2p) and hist the
// Lp-1 newest symbols.  Decision rule, all kernels (oracle/cpm_generic.py:detect): BLOCK RELEASE with
// survivor memory 2 D -- whenever (t + 1) is a multiple of D and t + 1 >= 2 D the survivor of the best
// state at step t (largest metric, lowest index on ties) releases the symbols of steps
// [t - 2D + 1, t - D]; the best final survivor releases the rest.
//
// Three kernels:
//   viterbi_cpm_kernel         any (M, Lp, h list, p), float / double, bank outputs z as input: one warp
//                              per frame, lane-per-state, metrics and 2D-symbol REGISTER-EXCHANGE
//                              survivor words in shared memory (the parity instantiation).
//   cpm_detect_m2l2_kernel     binary, memory 2, single h (IRIG PCM/FM): one THREAD per frame, reads the
//                              RECEIVED SAMPLES, computes the step's 4 correlations itself (the bank is
//                              fused in), add-compare-select on registers, one decision BIT per state per
//                              step packed into a 64-bit word in shared memory, traceback every D steps.
//   cpm_detect_lanev_kernel    M = 4, memory 2, 32 phase states, two h (ARTM CPM): one WARP per frame, lane
//                              = phase state, bank fused in (16 correlations spread over the lanes),
//                              metrics exchanged by shuffles, 2 decision bits per state per step in shared
//                              memory, warp-uniform traceback every D steps.
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <type_traits>

namespace wf {

constexpr int kCpmWarps = 4;
constexpr int kCpmMaxStates = 256;
constexpr int kCpmMaxCols = 128;  // n_h * M^Lp bank outputs per step
constexpr int kCpmMaxH = 8;

struct CpmParams {
  const void* z;    // bank outputs (viterbi_cpm_kernel)
  const void* rx;   // received samples complex64[B][ld_rx] (fused kernels)
  int8_t* dec_syms;
  const int8_t* tx_syms;
  const int32_t* point;
  unsigned long long* counters;
  int64_t B, n_steps, ld_z, S, skip_head, skip_tail, Tn, ld_rx;
  int M, Lp, p, n_h, depth, bits, n_filt, n_hist, n_states, n_cols;
  int start_parity;  // parity class of the phase-state index v the detector starts from (oracle/cpm_generic.py:start_parity); -1 = every v
  int delta[kCpmMaxH][16];  // (K_h * alphabet[x]) mod 2p
};

struct CpmRot { float c[64], s[64]; };  // cos / sin(pi v / p), v in Z_2p (float32 of the double values)

template <typename T>
__global__ void __launch_bounds__(kCpmWarps * 32)
viterbi_cpm_kernel(const __grid_constant__ CpmParams p) {
  using C = typename std::conditional<sizeof(T) == 4, float2, double2>::type;
  __shared__ T pm_s[kCpmWarps][2][kCpmMaxStates];
  __shared__ unsigned long long hist_s[kCpmWarps][2][kCpmMaxStates];
  __shared__ C z_s[kCpmWarps][kCpmMaxCols];
  __shared__ T rot_c[64], rot_s[64];

  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int p2 = 2 * p.p;
  for (int i = threadIdx.x; i < p2; i += blockDim.x) {
    double sn, cs;
    sincospi((double)i / (double)p.p, &sn, &cs);
    rot_c[i] = (T)cs;
    rot_s[i] = (T)sn;
  }
  __syncthreads();
  const int64_t frame = (int64_t)blockIdx.x * kCpmWarps + wid;
  if (frame >= p.B) return;  // whole warp; no block barrier after this point

  const int M = p.M, Lp = p.Lp, NS = p.n_states, NH = p.n_hist, bits = p.bits, D = p.depth;
  const int NHM = NH >= M ? NH / M : 1;
  // survivor words hold the newest 2 D symbols (2 D bits <= 64, checked by the launcher)
  const unsigned long long hmask = (2 * D * bits >= 64) ? ~0ull : ((1ull << (2 * D * bits)) - 1ull);
  const unsigned sym_mask = (1u << bits) - 1u;
  const C* zrow = reinterpret_cast<const C*>(p.z) + frame * p.ld_z * p.n_cols;
  const int8_t* tx = p.tx_syms ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = p.dec_syms ? p.dec_syms + frame * p.n_steps : nullptr;
  const int64_t n_steps = p.n_steps;
  const int64_t cmp_lo = p.skip_head, cmp_hi = min(p.S, n_steps) - p.skip_tail;

  T* pm[2] = {pm_s[wid][0], pm_s[wid][1]};
  unsigned long long* hs[2] = {hist_s[wid][0], hist_s[wid][1]};
  for (int s = lane; s < NS; s += 32) {
    const bool dead = p.start_parity >= 0 && (((s / NH) & 1) != p.start_parity);
    pm[0][s] = dead ? (sizeof(T) == 4 ? (T)-3.0e38f : (T)-1e300) : (T)0;  // absorbs every finite increment: stays dead
    hs[0][s] = 0ull;
  }
  __syncwarp();

  unsigned long long sym_err = 0, bit_err = 0, compared = 0;
  auto emit = [&](int64_t idx, unsigned v) {  // lane 0 only
    const int a = 2 * (int)v - (M - 1);
    if (out) out[idx] = (int8_t)a;
    if (tx && idx >= cmp_lo && idx < cmp_hi) {
      const unsigned t = (unsigned)(((int)tx[idx] + (M - 1)) >> 1);
      sym_err += (t != v) ? 1u : 0u;
      bit_err += __popc((t ^ v) & sym_mask);
      compared += (unsigned)bits;  // column 2 counts compared BITS for every waveform
    }
  };

  int cur = 0;
  int64_t released = 0;
  for (int64_t i = 0; i < n_steps; ++i) {
    const int par = (int)(i % p.n_h);
    int lpar = (int)((i - Lp + 1) % p.n_h);
    if (lpar < 0) lpar += p.n_h;
    const C* zi = zrow + i * p.n_cols + par * p.n_filt;
    for (int f = lane; f < p.n_filt; f += 32) z_s[wid][f] = zi[f];
    __syncwarp();
    const T* pmo = pm[cur];
    const unsigned long long* ho = hs[cur];
    T* pmn = pm[cur ^ 1];
    unsigned long long* hn = hs[cur ^ 1];
    T lane_max = (T)-1e300;
    if (sizeof(T) == 4) lane_max = (T)-3.0e38f;
    for (int s = lane; s < NS; s += 32) {
      const int vnew = s / NH, hnew = s - vnew * NH;
      T best = lane_max;  // overwritten by x = 0 below
      int arg = 0, argx = 0;
      for (int x = 0; x < M; ++x) {
        int vpred = vnew - p.delta[lpar][x];
        if (vpred < 0) vpred += p2;
        const int pred = (Lp == 1) ? vpred : vpred * NH + (hnew % NHM) * M + x;
        const int f = (Lp == 1) ? x : hnew * M + x;
        const C zz = z_s[wid][f];
        // pm + Re{e^{-j pi v/p} z} = pm + (cos * zr + sin * zi), evaluated in this order
        const T cand = pmo[pred] + (rot_c[vpred] * zz.x + rot_s[vpred] * zz.y);
        if (x == 0 || cand > best) { best = cand; arg = pred; argx = x; }
      }
      const unsigned newest = (Lp == 1) ? (unsigned)argx : (unsigned)(hnew / NHM);
      hn[s] = ((ho[arg] << bits) | (unsigned long long)newest) & hmask;
      pmn[s] = best;
      lane_max = (s == lane) ? best : (best > lane_max ? best : lane_max);
    }
    // warp max (for normalisation) and argmax with the LOWEST state index on ties
    T wmax = lane_max;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const T t = __shfl_xor_sync(0xFFFFFFFFu, wmax, o);
      wmax = t > wmax ? t : wmax;
    }
    __syncwarp();
    int best_s = 0x7FFFFFFF;
    for (int s = lane; s < NS; s += 32) {
      const T v = pmn[s];
      if (v == wmax && s < best_s) best_s = s;
      pmn[s] = v - wmax;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const int t = __shfl_xor_sync(0xFFFFFFFFu, best_s, o);
      best_s = t < best_s ? t : best_s;
    }
    __syncwarp();
    cur ^= 1;
    // block release: the oldest D symbols of the best survivor's 2 D-symbol word
    if ((i + 1) % D == 0 && i + 1 >= 2 * D) {
      if (lane == 0) {
        const unsigned long long h = hs[cur][best_s];
        for (int j = 0; j < D; ++j) emit(i - 2 * D + 1 + j, (unsigned)((h >> ((2 * D - 1 - j) * bits)) & sym_mask));
      }
      released = i + 1 - D;
    }
    if (i == n_steps - 1 && lane == 0) {  // the best final survivor releases what is still held (< 2 D symbols)
      const unsigned long long h = hs[cur][best_s];
      for (int64_t idx = released; idx < n_steps; ++idx) emit(idx, (unsigned)((h >> ((n_steps - 1 - idx) * bits)) & sym_mask));
    }
  }
  if (lane == 0 && p.counters) {
    const int64_t pt = p.point ? p.point[frame] : 0;
    atomicAdd(&p.counters[3 * pt + 0], bit_err);
    atomicAdd(&p.counters[3 * pt + 1], sym_err);
    atomicAdd(&p.counters[3 * pt + 2], compared);
  }
}

// ---- small device helpers -------------------------------------------------------------------------
__device__ __forceinline__ uint32_t funnel_l1(uint32_t hi, uint32_t lo) {  // (hi << 1) | (lo >> 31)
#ifdef WF_HOSTEMU
  return (hi << 1) | (lo >> 31);
#else
  return __funnelshift_l(lo, hi, 1);
#endif
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
#ifdef WF_HOSTEMU
  return make_float2(a.x + b.x, a.y + b.y);
#else
  unsigned long long ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
#endif
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
#ifdef WF_HOSTEMU
  return make_float2(a.x * b.x, a.y * b.y);
#else
  unsigned long long ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
#endif
}
// warp-wide max / min in one instruction (CREDUX; sm_100a has the f32 form)
__device__ __forceinline__ float warp_max_f32(float v) {
#ifdef WF_HOSTEMU
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xFFFFFFFFu, v, o);
    v = t > v ? t : v;
  }
  return v;
#else
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
#endif
}
__device__ __forceinline__ int warp_min_s32(int v) {
#ifdef WF_HOSTEMU
  for (int o = 16; o > 0; o >>= 1) {
    const int t = __shfl_xor_sync(0xFFFFFFFFu, v, o);
    v = t < v ? t : v;
  }
  return v;
#else
  return __reduce_min_sync(0xFFFFFFFFu, v);
#endif
}

// ---------------------------------------------------------------------------
// Binary, memory-2, single-h shape (IRIG PCM/FM: M = 2, Lp = 2, h = K/p, first instant -1), straight from
// the received samples.  One THREAD per frame.
//  * Bank fused in: the step's window is the 8 samples rx[8i-1 .. 8i+6]; the four hypothesis filters are two
//    complex-conjugate pairs (hypothesis (-a0,-a1) is the conjugate of (a0,a1)), so four real sums per pair,
//    (A, C) += (x_re, x_im) (c_re, c_re) and (E, B) += (x_re, x_im) (c_im, c_im) -- two packed FFMA2 per
//    sample and pair, taps as constant-bank operands -- give z = (A - B, E + C) and its partner (A + B, C - E).
//  * The 2p = 20 rotations fold to 10 products per filter (e^{-j pi (v+10)/10} = -e^{-j pi v/10}); the 2 P2
//    path metrics live in registers, predecessors / filters / rotations are compile-time indices.
//  * Survivors: ONE decision bit per state per step (sign of c0 - c1, funnel-shifted into a word) -- 40 bits
//    stored as one 64-bit word in a shared-memory ring of RING steps, laid out [step][thread] (conflict
//    free); every D steps the thread traces its best state back 2 D steps and releases D symbols, which are
//    compared with the transmitted symbols 4 at a time.
//  * Samples reach the thread through a warp-private cp.async double buffer: CS steps (16 CS samples) of the
//    warp's 32 rows per chunk, 16-byte units, rows an odd number of units apart (LDS.128 conflict free); the
//    unit that straddles two steps is carried in registers.
// Same decisions as oracle/cpm_generic.py up to float32 near-ties (tested with the float64 margins).
// ---------------------------------------------------------------------------
constexpr int kBfThreads = 64;  // two autonomous warps
constexpr int kBfCS = 1;        // steps per staged chunk: 4 CS units of 16 bytes per row (CS = 2 measured 6.0 ms at 12 warps / SM)
constexpr int kBfPitch = 4 * kBfCS + 1;
constexpr int kBfBufs = 3;      // staged chunks in flight: chunk c + 2 is requested while chunk c is consumed (5 KB each: with the
                                // 16 KB decision ring, seven blocks still fit an SM, which one wave of 65 536 frames needs)

struct CpmPairTaps {  // [pair][sample u][0: (c_re, c_re), 1: (c_im, c_im)] of the conj(template) of filters 0 and 1
  alignas(16) float t[2][8][2][2];
};

template <int P2, int D0, int D1, int RING, int NORM>
__global__ void __launch_bounds__(kBfThreads)
cpm_detect_m2l2_kernel(const __grid_constant__ CpmParams p, const __grid_constant__ CpmRot rot,
                       const __grid_constant__ CpmPairTaps tp) {
  constexpr int NS = 2 * P2, HP = P2 / 2;
  static_assert(P2 <= 32 && (P2 % 4) == 0 && (D0 & 1) && (D1 & 1) && (RING & (RING - 1)) == 0,
                "both phase steps odd: the P2 states of one parity class of v are live at a time, one 32-bit decision word per step");
  __align__(16) __shared__ float4 tile[kBfBufs][kBfThreads * kBfPitch];
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint32_t* tb = reinterpret_cast<uint32_t*>(smem_raw);  // [RING][kBfThreads]: one decision bit per LIVE state (P2 of them) per step

  const int tid = threadIdx.x, lane = tid & 31, wrow0 = tid & ~31;
  const int64_t wframe0 = (int64_t)blockIdx.x * kBfThreads + wrow0;
  if (wframe0 >= p.B) return;  // whole warp; only warp-level synchronisation below
  const int64_t frame = wframe0 + lane;
  const bool active = frame < p.B;
  const int n_steps = (int)p.n_steps, S = (int)p.S, D = p.depth;
  const int8_t* tx = (active && p.tx_syms) ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = (active && p.dec_syms) ? p.dec_syms + frame * p.n_steps : nullptr;
  const int cmp_lo = (int)p.skip_head, cmp_hi = min(S, n_steps) - (int)p.skip_tail;
  const bool tx_vec = tx != nullptr && ((reinterpret_cast<uintptr_t>(tx) & 3) == 0);
  // staging pattern: unit e = lane + 32 r (r < CS) of the warp's 32 rows x 4 CS units -> row e / (4 CS), unit e % (4 CS)
  constexpr int UPR = 4 * kBfCS;              // units per row per chunk
  constexpr int RPS = 32 / UPR;               // rows covered by one 32-lane pass
  const int64_t row_units = p.ld_rx / 2;      // ld_rx is even (checked by the launcher)
  const int64_t last_row = p.B - 1 - wframe0; // rows beyond it are clamped to it (valid memory, never used)
  const float4* src0 = reinterpret_cast<const float4*>(p.rx) + wframe0 * row_units + (lane % UPR);
  const int dst0 = (wrow0 + lane / UPR) * kBfPitch + (lane % UPR);
  auto stage = [&](int buf, int chunk) {
#pragma unroll
    for (int r = 0; r < UPR; ++r) {
      int64_t row = lane / UPR + RPS * r;
      if (row > last_row) row = last_row;
      cp_async_cg16(&tile[buf][dst0 + RPS * r * kBfPitch], src0 + row * row_units + (int64_t)chunk * UPR);
    }
    cp_async_commit();
  };

  float pm[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) pm[s] = 0.f;  // step 0 reads the start class only; the other half is written before it is read
  const int start_par = p.start_parity;      // 0 or 1 here (odd phase steps)
  unsigned sym_err = 0, compared = 0;
  const float2* tap2 = reinterpret_cast<const float2*>(tp.t);

  // release the symbols of steps [lo, hi) found by tracing state `s` back from step t
  auto trace_release = [&](int t, int s, int lo, int hi) {
    unsigned long long bitsw = 0ull;  // bit (tau - lo) = decided symbol index of step tau
    for (int tau = t; tau >= lo; --tau) {
      const uint32_t w = tb[(tau & (RING - 1)) * kBfThreads + tid];
      const int j = ((s >> 2) << 1) | (s & 1);                 // rank of s = 4 m + 2 Q + h among the live states of its step
      const unsigned b = (w >> (P2 - 1 - j)) & 1u;             // which predecessor won
      if (tau < hi) bitsw |= (unsigned long long)(s & 1) << (tau - lo);
      int vp = (s >> 1) - (b ? D1 : D0);
      vp += vp < 0 ? P2 : 0;
      s = 2 * vp + (int)b;
    }
    const int n = hi - lo;
    if (out)
      for (int j = 0; j < n; ++j) out[lo + j] = (int8_t)(2 * (int)((bitsw >> j) & 1ull) - 1);
    if (tx) {
      for (int j0 = 0; j0 < n; j0 += 4) {
        const int k = lo + j0;
        unsigned txb = 0, msk = 0;
        if (tx_vec && (k & 3) == 0 && j0 + 4 <= n && k + 4 <= S) {
          const uint32_t wv = ~*reinterpret_cast<const uint32_t*>(tx + k);  // +1 -> sign bit clear -> 1 after the inversion
          txb = ((wv >> 7) & 1u) | ((wv >> 14) & 2u) | ((wv >> 21) & 4u) | ((wv >> 28) & 8u);
#pragma unroll
          for (int e = 0; e < 4; ++e) msk |= (k + e >= cmp_lo && k + e < cmp_hi) ? (1u << e) : 0u;
        } else {
          for (int e = 0; e < 4 && j0 + e < n; ++e) {
            if (k + e >= cmp_lo && k + e < cmp_hi) {
              msk |= 1u << e;
              txb |= (unsigned)(((int)tx[k + e] + 1) >> 1) << e;
            }
          }
        }
        sym_err += __popc(((unsigned)(bitsw >> j0) ^ txb) & msk);
        compared += __popc(msk);
      }
    }
  };

  const int n_chunks = (n_steps + kBfCS - 1) / kBfCS;
  // kBfBufs - 1 chunks ahead: with one thread per frame a batch is a single wave of ~3.5 warps per scheduler, too
  // few to hide a DRAM round trip behind one step of arithmetic (long-scoreboard stalls 0.57 cycles per issue with
  // one chunk ahead); every iteration commits exactly one group (empty past the end) so the wait count is constant
#pragma unroll
  for (int k = 0; k < kBfBufs - 1; ++k) {
    if (k < n_chunks) stage(k, k);
    else cp_async_commit();
  }
  float2 carry = make_float2(0.f, 0.f);  // sample 8 i - 1 (zero before the frame)
  int released = 0;
  int next_rel = 2 * D - 1;  // the next step that ends a release block
  int buf = 0, buf_next = kBfBufs - 1;  // c % kBfBufs and (c + kBfBufs - 1) % kBfBufs
  for (int c = 0; c < n_chunks; ++c) {
    if (c + kBfBufs - 1 < n_chunks) stage(buf_next, c + kBfBufs - 1);
    else cp_async_commit();
    cp_async_wait<kBfBufs - 1>();
    __syncwarp();
    const float4* row = &tile[buf][tid * kBfPitch];
    buf_next = buf;
    buf = buf + 1 == kBfBufs ? 0 : buf + 1;
#pragma unroll
    for (int u = 0; u < kBfCS; ++u) {
      const int i = c * kBfCS + u;
      if (i < n_steps) {
        // ---- the step's window and its four correlations --------------------------------------
        float2 w[8];
        w[0] = carry;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 v = row[4 * u + q];
          if (q < 3) { w[2 * q + 1] = make_float2(v.x, v.y); w[2 * q + 2] = make_float2(v.z, v.w); }
          else { w[7] = make_float2(v.x, v.y); carry = make_float2(v.z, v.w); }
        }
        float zx[4], zy[4];
#pragma unroll
        for (int pr = 0; pr < 2; ++pr) {
          float2 ac = make_float2(0.f, 0.f), eb = make_float2(0.f, 0.f);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            ac = ffma2(w[k], tap2[(pr * 8 + k) * 2 + 0], ac);
            eb = ffma2(w[k], tap2[(pr * 8 + k) * 2 + 1], eb);
          }
          zx[pr] = ac.x - eb.y; zy[pr] = eb.x + ac.y;          // filter pr
          zx[3 - pr] = ac.x + eb.y; zy[3 - pr] = ac.y - eb.x;  // its conjugate partner
        }
        // ---- the step proper, for the parity class Q of the phase states it ENDS in -----------------
        // Both phase steps are odd, so every branch flips the parity of v: the states of class Q are written
        // from the states of class 1 - Q, IN PLACE (the two halves of pm[] alternate as source and destination),
        // and only the P2 / 2 rotations Re{e^{-j pi k/p} z_f} of the predecessor class are formed.
        auto step = [&](auto qc) {
          constexpr int Q = decltype(qc)::value;
          float r[HP][4];
#pragma unroll
          for (int k = 0; k < HP; ++k)
            if ((k & 1) != Q) {  // k = vp mod HP has the parity of vp (HP is even): predecessor class 1 - Q
#pragma unroll
              for (int f = 0; f < 4; ++f)
                r[k][f] = k == 0 ? zx[f] : ((2 * k == HP) ? zy[f] : rot.c[k] * zx[f] + rot.s[k] * zy[f]);
            }
          uint32_t w = 0;
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            const int vnew = s >> 1, hnew = s & 1;
            if ((vnew & 1) == Q) {
              const int vp0 = (vnew - D0 + 2 * P2) % P2, vp1 = (vnew - D1 + 2 * P2) % P2;
              const int f0 = hnew * 2, f1 = hnew * 2 + 1;
              const float c0 = vp0 < HP ? pm[2 * vp0] + r[vp0 % HP][f0] : pm[2 * vp0] - r[vp0 % HP][f0];
              const float c1 = vp1 < HP ? pm[2 * vp1 + 1] + r[vp1 % HP][f1] : pm[2 * vp1 + 1] - r[vp1 % HP][f1];
              const float d = c0 - c1;  // negative <=> c1 > c0: strict, so the lower x keeps ties
              pm[s] = fmaxf(c0, c1);
              w = funnel_l1(w, __float_as_uint(d));  // live states in increasing s: rank j ends at bit P2 - 1 - j
            }
          }
          tb[(i & (RING - 1)) * kBfThreads + tid] = w;
          const bool blk = i == next_rel;
          const bool rel = blk || i == n_steps - 1;
          if (rel || NORM == 1 || (i & (NORM - 1)) == NORM - 1) {
            float wmax = -3.0e38f;  // only the steps that normalise or release need the maximum
#pragma unroll
            for (int s = 0; s < NS; ++s)
              if (((s >> 1) & 1) == Q) wmax = fmaxf(wmax, pm[s]);
            if (rel) {
              int best_s = -1;
#pragma unroll
              for (int s = NS - 1; s >= 0; --s)
                if (((s >> 1) & 1) == Q) best_s = (pm[s] == wmax || best_s < 0) ? s : best_s;  // lowest live index wins
              if (blk) {
                trace_release(i, best_s, released, i + 1 - D);
                released = i + 1 - D;
                next_rel += D;
              }
              if (i == n_steps - 1) trace_release(i, best_s, released, n_steps);
            }
#pragma unroll
            for (int s = 0; s < NS; ++s)
              if (((s >> 1) & 1) == Q) pm[s] -= wmax;
          }
        };
        // class after step i: the start class flipped i + 1 times
        if ((start_par + i + 1) & 1) step(std::integral_constant<int, 1>{});
        else step(std::integral_constant<int, 0>{});
      }
    }
    __syncwarp();  // all lanes are done with this buffer before chunk c + kBfBufs is staged into it
  }
  if (p.counters) {
    // binary alphabet: bit errors = symbol errors; one atomic triple per warp when the points agree
    const uint32_t full = 0xFFFFFFFFu;
    const int32_t pt = (active && p.point) ? p.point[frame] : 0;
    const int32_t pt0 = __shfl_sync(full, pt, 0);
    unsigned long long se = active ? sym_err : 0u, nc = active ? compared : 0u;
    if (__all_sync(full, !active || pt == pt0)) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        se += __shfl_xor_sync(full, se, o);
        nc += __shfl_xor_sync(full, nc, o);
      }
      if (lane == 0) {
        atomicAdd(&p.counters[3 * (int64_t)pt0 + 0], se);
        atomicAdd(&p.counters[3 * (int64_t)pt0 + 1], se);
        atomicAdd(&p.counters[3 * (int64_t)pt0 + 2], nc);
      }
    } else if (active) {
      atomicAdd(&p.counters[3 * (int64_t)pt + 0], se);
      atomicAdd(&p.counters[3 * (int64_t)pt + 1], se);
      atomicAdd(&p.counters[3 * (int64_t)pt + 2], nc);
    }
  }
}

// ---------------------------------------------------------------------------
// M = 4, memory 2, 2p = 32 phase states, two modulation indices (ARTM CPM: h = {4, 5}/16 -> 128 states),
// straight from the received samples.  One WARP per frame, lane = phase state v, the M history states of a
// lane in REGISTERS (for Lp = 2 the predecessor of (v, h_new) through oldest symbol x is (v - delta[x], x)
// for every h_new: M metric shuffles per step with per-lane source lanes).
//  * Bank fused in: only the step's OWN modulation-index parity is computed (16 of the 32 filters), and
//    only 8 of those -- hypothesis 15 - f is the complex conjugate of f.  Lane = (pair q, which, half):
//    4 samples x one packed FFMA2 each, (A, C) += x (c_re, c_re) for which = 0, (E, B) += x (c_im, c_im)
//    for which = 1; two shuffles add the halves, two more swap (A, C) <-> (E, B), and the lane forms
//    z_q = (A - B, E + C) or z_{15-q} = (A + B, C - E).  The 16 outputs go to shared memory as
//    (re_x, re_x+1, im_x, im_x+1) quads, so the rotation + add of TWO candidates is one FMUL2 + FFMA2 + FADD2.
//  * Samples: a 256-sample warp-private ring filled by cp.async, 64 samples (one 16-byte unit per lane)
//    per chunk, two chunks ahead.
//  * Survivors: two decision bits per state -- one byte per lane per step -- in a shared-memory ring of RING
//    steps; every D steps the whole warp walks the best state's path back 2 D steps (warp-uniform: broadcast
//    byte load, two integer ops per step) and releases D symbols, lane j comparing / storing symbol j.
// ---------------------------------------------------------------------------
constexpr int kLfWarps = 4;
constexpr int kLfRingSamples = 256;

struct CpmLaneTaps {  // conj(template) of filters 0..7 for both parities: [parity][filter][u][re, im]
  float t[2][8][8][2];
};

// Walk the path of state (v, hn) back from step t to step lo and release steps [lo, hi): lane j handles symbol
// lo + j (+ 32).  Warp-uniform.  Per step: one broadcast byte load, the winning predecessor symbol x from two
// bits of it, the phase step of (parity, x) by PRMT from a packed register, and a 2-bit shift-in of h_new.
// Returns this lane's (symbol errors, bit errors, compared bits).
template <int RING>
static __device__ __noinline__ uint3 lanev_trace_release(const uint8_t* tb, uint32_t dl0, uint32_t dl1, int t, int v, int hn, int lo,
                                                         int hi, int lane, int8_t* out, const int8_t* tx, int cmp_lo, int cmp_hi) {
  constexpr int M = 4, BITS = 2;
  unsigned long long w0 = 0ull, w1 = 0ull;  // 2 bits per released symbol, symbol lo in the lowest bits of w0
  int hn2 = 2 * hn;
  auto back = [&](int tau) {
    // the byte of lane 2 m + (hn >> 1), m = (v - class) / 2, class = live parity after step tau = (tau & 3) >> 1
    const int li = ((v - ((tau & 3) >> 1)) & 30) | (hn2 >> 2);
    const unsigned b = tb[((tau & (RING - 1)) << 5) + li];
    const unsigned x = (b >> (hn2 & 2)) & 3u;
    const uint32_t dl = (tau & 1) ? dl0 : dl1;  // transition parity (tau - Lp + 1) % 2 = (tau + 1) & 1
#ifdef WF_HOSTEMU
    const uint32_t dv = (dl >> (8 * x)) & 0xFFu;
#else
    const uint32_t dv = __byte_perm(dl, 0u, 0x4440u | x);
#endif
    v = (v - (int)dv) & 31;
    hn2 = (int)(2u * x);
  };
  int tau = t;
  for (; tau >= hi; --tau) back(tau);                  // merge part: nothing released
  if (hi - lo <= 32) {
    for (; tau >= lo; --tau) { w0 = (w0 << 2) | (unsigned long long)(hn2 >> 1); back(tau); }
  } else {
    for (; tau >= lo; --tau) { w1 = (w1 << 2) | (w0 >> 62); w0 = (w0 << 2) | (unsigned long long)(hn2 >> 1); back(tau); }
  }
  uint3 r = make_uint3(0u, 0u, 0u);
  const int n = hi - lo;
  for (int j = lane; j < n; j += 32) {
    const unsigned dec = (unsigned)((j < 32 ? w0 >> (2 * j) : w1 >> (2 * (j - 32))) & 3ull);
    const int k = lo + j;
    if (out) out[k] = (int8_t)(2 * (int)dec - (M - 1));
    if (tx && k >= cmp_lo && k < cmp_hi) {
      const unsigned tt = (unsigned)(((int)tx[k] + (M - 1)) >> 1);
      r.x += (tt != dec) ? 1u : 0u;
      r.y += __popc((tt ^ dec) & (unsigned)(M - 1));
      r.z += BITS;  // column 2 counts compared BITS for every waveform
    }
  }
  return r;
}

template <int RING, int NORM>
__global__ void __launch_bounds__(kLfWarps * 32)
cpm_detect_lanev_kernel(const __grid_constant__ CpmParams p, const __grid_constant__ CpmRot rot,
                        const __grid_constant__ CpmLaneTaps tp) {
  constexpr int M = 4, NHH = 2;
  __align__(16) __shared__ float2 ring_s[kLfWarps][kLfRingSamples];
  __align__(16) __shared__ float4 zq_s[kLfWarps][2][2][8];    // [buffer][step of the pair][(re_x, re_x+1, im_x, im_x+1) at 2 h_new + x / 2]
  __shared__ uint8_t tb_s[kLfWarps][RING * 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t frame = (int64_t)blockIdx.x * kLfWarps + wid;
  if (frame >= p.B) return;  // whole warp; only warp-level synchronisation below

  const int D = p.depth;
  const int n_steps = (int)p.n_steps, S = (int)p.S, Tn = (int)p.Tn;
  const int8_t* tx = p.tx_syms ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = p.dec_syms ? p.dec_syms + frame * p.n_steps : nullptr;
  const int cmp_lo = (int)p.skip_head, cmp_hi = min(S, n_steps) - (int)p.skip_tail;
  const float2* rxrow = reinterpret_cast<const float2*>(p.rx) + frame * p.ld_rx;
  float2* ring = ring_s[wid];
  uint8_t* tb = tb_s[wid];

  // Reachable states only.  h = {even, odd} / p: the parity class of the phase-state index v flips on the
  // odd-K transitions, i.e. after step i the live class is LA(i) = (i & 3) >> 1 (start class 1), 16 phase states x 4
  // histories = 64 of the 128 states.  Lane = 2 m + hh owns v = 2 m + LA and the two histories h_new = 2 hh, 2 hh + 1.
  // The predecessor of (v, h_new) through oldest symbol x is (v - delta[lp][x], x), held by lane
  //   (lane & ~1) + (LA - LB) - delta[lp][x] + (x >> 1)   (mod 32; LB = class before the step)
  // in its register pm[x & 1]: four shuffles per step, sources = one lane-constant plus uniform integers.
  // The rotation e^{-j pi v_pred / p} splits into a lane part e^{-j pi 2 m / p} (cm, sm below) and a uniform part
  // e^{-j pi (LA - delta[lp][x]) / p} that the BANK lanes fold into z before they hand it over.
  const int base_lane = lane & ~1, hh = lane & 1;
  const float cm = rot.c[base_lane], sm = rot.s[base_lane];     // cos / sin(pi 2 m / p)
  float2 cm2 = make_float2(cm, cm), sm2 = make_float2(sm, sm);
  int nd[NHH][M];                                               // -delta[lp][x] + (x >> 1), uniform
#pragma unroll
  for (int h = 0; h < NHH; ++h)
#pragma unroll
    for (int x = 0; x < M; ++x) nd[h][x] = (x >> 1) - p.delta[h][x];
  // phase steps of both parities packed one per byte: the traceback picks byte x with one PRMT
  const uint32_t dl0 = (uint32_t)p.delta[0][0] | ((uint32_t)p.delta[0][1] << 8) | ((uint32_t)p.delta[0][2] << 16) | ((uint32_t)p.delta[0][3] << 24);
  const uint32_t dl1 = (uint32_t)p.delta[1][0] | ((uint32_t)p.delta[1][1] << 8) | ((uint32_t)p.delta[1][2] << 16) | ((uint32_t)p.delta[1][3] << 24);
#ifndef WF_HOSTEMU
  asm volatile("" : "+f"(cm2.x), "+f"(cm2.y), "+f"(sm2.x), "+f"(sm2.y));  // keep them in registers (see CpmRot)
#endif
  // bank role of this lane: it serves the step of ITS parity (steps come in pairs (even, odd)), filter pair q,
  // and accumulates either (A, C) (which = 0, taps c_re) or (E, B) (which = 1, taps c_im) over all 8 samples
  const int bstep = lane >> 4, bq = (lane >> 1) & 7, bwhich = lane & 1;
  float2 tapd[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float t = tp.t[bstep][bq][k][bwhich];
    tapd[k] = make_float2(t, t);
  }
#ifndef WF_HOSTEMU
#pragma unroll
  for (int k = 0; k < 8; ++k) asm volatile("" : "+f"(tapd[k].x), "+f"(tapd[k].y));
#endif
  const int bf = bwhich ? 15 - bq : bq;              // the filter this lane finishes
  const int zword = 32 * bstep + 4 * (bf >> 1) + (bf & 1);  // its re slot inside a buffer (floats); im is two words further on
  // uniform part of the rotation for the filter this lane finishes (oldest symbol x = bf & 3) at its steps
  // u = 2 up + bstep: transition parity lp = (bstep + 1) & 1, class after the step LA = up & 1
  float zrc[2], zrs[2];
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    const int n = (a - p.delta[(bstep + 1) & 1][bf & 3]) & 31;
    zrc[a] = rot.c[n];
    zrs[a] = -rot.s[n];  // e^{-j pi n / p} = cos - j sin
  }
#ifndef WF_HOSTEMU
  asm volatile("" : "+f"(zrc[0]), "+f"(zrc[1]), "+f"(zrs[0]), "+f"(zrs[1]));
#endif

  float pm[2] = {0.f, 0.f};  // histories 2 hh, 2 hh + 1 of this lane's live phase state (every live state starts equal)
  unsigned sym_err = 0, bit_err = 0, compared = 0;  // per lane, reduced at the end

  // chunk c = samples [64 c - 8, 64 c + 56) -> ring slots (64 c + j) & 255; lane l moves unit l
  auto stage = [&](int c) {
    const int n = 64 * c - 8 + 2 * lane;
    float2* d = ring + ((64 * c + 2 * lane) & (kLfRingSamples - 1));
    if (n >= 0 && n + 1 < Tn) {
      cp_async_cg16(d, rxrow + n);
    } else {
      d[0] = (n >= 0 && n < Tn) ? rxrow[n] : make_float2(0.f, 0.f);
      d[1] = make_float2(0.f, 0.f);
    }
    cp_async_commit();
  };

  // traceback + release live out of line (lanev_trace_release): the unrolled step loop stays small enough for the
  // instruction cache (inlined eight times it cost 0.9 stall cycles per issue in "no instruction")
  auto trace_release = [&](int t, int v, int hn, int lo, int hi) {
    const uint3 r = lanev_trace_release<RING>(tb, dl0, dl1, t, v, hn, lo, hi, lane, out, tx, cmp_lo, cmp_hi);
    sym_err += r.x; bit_err += r.y; compared += r.z;
  };

  stage(0);
  stage(1);
  int released = 0;
  int next_rel = 2 * D - 1;  // the next step that ends a release block
  for (int i0 = 0; i0 < n_steps; i0 += 8) {
    // group g = i0 / 8 reads chunks g and g + 1; chunk g + 2 replaces chunk g - 2 (last read by step i0 - 1)
    __syncwarp();
    stage((i0 >> 3) + 2);
    cp_async_wait<1>();
    __syncwarp();
#pragma unroll
    for (int up = 0; up < 4; ++up) {
      if (i0 + 2 * up >= n_steps) break;
      // ---- bank of the step pair (i0 + 2 up, + 1): 8 samples, one FFMA2 each, two chains ------------------
      {
        const int s0 = 8 * (i0 + 2 * up + bstep) + 7;  // ring slot of sample 8 i - 1  (slot(n) = (n + 8) & 255)
        const float2 x0 = ring[s0 & (kLfRingSamples - 1)];
        const float4 x12 = *reinterpret_cast<const float4*>(ring + ((s0 + 1) & (kLfRingSamples - 1)));
        const float4 x34 = *reinterpret_cast<const float4*>(ring + ((s0 + 3) & (kLfRingSamples - 1)));
        const float4 x56 = *reinterpret_cast<const float4*>(ring + ((s0 + 5) & (kLfRingSamples - 1)));
        const float2 x7 = ring[(s0 + 7) & (kLfRingSamples - 1)];
        float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
        a0 = ffma2(x0, tapd[0], a0);
        a1 = ffma2(make_float2(x12.x, x12.y), tapd[1], a1);
        a0 = ffma2(make_float2(x12.z, x12.w), tapd[2], a0);
        a1 = ffma2(make_float2(x34.x, x34.y), tapd[3], a1);
        a0 = ffma2(make_float2(x34.z, x34.w), tapd[4], a0);
        a1 = ffma2(make_float2(x56.x, x56.y), tapd[5], a1);
        a0 = ffma2(make_float2(x56.z, x56.w), tapd[6], a0);
        a1 = ffma2(x7, tapd[7], a1);
        const float2 acc = fadd2(a0, a1);
        const float ox = __shfl_xor_sync(0xFFFFFFFFu, acc.x, 1), oy = __shfl_xor_sync(0xFFFFFFFFu, acc.y, 1);
        // which = 0: own (A, C), other (E, B) -> z_q = (A - B, E + C);  which = 1: own (E, B), other (A, C) -> z_{15-q} = (A + B, C - E)
        const float zr0 = bwhich ? ox + acc.y : acc.x - oy;
        const float zi0 = bwhich ? oy - acc.x : ox + acc.y;
        const float zr = fmaf(-zi0, zrs[up & 1], zr0 * zrc[up & 1]);   // z e^{-j pi (LA - delta) / p}
        const float zi = fmaf(zi0, zrc[up & 1], zr0 * zrs[up & 1]);
        float* zs = reinterpret_cast<float*>(zq_s[wid][up & 1]);  // double buffered: the readers of this buffer are two pairs back
        zs[zword] = zr;
        zs[zword + 2] = zi;
        __syncwarp();
      }
#pragma unroll
      for (int us = 0; us < 2; ++us) {
        const int u = 2 * up + us;
        const int i = i0 + u;
        if (i >= n_steps) break;
        const int lp = (us + 1) & 1;      // transition parity (i - 1) % 2  (i0 is a multiple of 8: i % 2 = us)
        const float4* zq = zq_s[wid][up & 1][us];
        // ---- add-compare-select over the live class --------------------------------------------------
        const int LA = (u & 3) >> 1, LB = ((u + 3) & 3) >> 1;  // class after / before step i (i0 is a multiple of 8; folded by the unroll)
        float2 pmx2[M / 2];
#pragma unroll
        for (int j = 0; j < M / 2; ++j) {
          pmx2[j].x = __shfl_sync(0xFFFFFFFFu, pm[0], base_lane + nd[lp][2 * j] + (LA - LB));
          pmx2[j].y = __shfl_sync(0xFFFFFFFFu, pm[1], base_lane + nd[lp][2 * j + 1] + (LA - LB));
        }
        unsigned dec = 0;
        float lmax = -3.0e38f;
#pragma unroll
        for (int k = 0; k < 2; ++k) {  // h_new = 2 hh + k
          float2 c2[M / 2];
#pragma unroll
          for (int j = 0; j < M / 2; ++j) {
            const float4 q = zq[4 * hh + 2 * k + j];
            float2 t2 = fmul2(cm2, make_float2(q.x, q.y));
            t2 = ffma2(sm2, make_float2(q.z, q.w), t2);
            c2[j] = fadd2(pmx2[j], t2);
          }
          // strict '>' tournament: the lowest x keeps ties
          const bool b01 = c2[0].y > c2[0].x, b23 = c2[1].y > c2[1].x;
          const float m01 = fmaxf(c2[0].x, c2[0].y), m23 = fmaxf(c2[1].x, c2[1].y);
          const bool bf2 = m23 > m01;
          const float best = fmaxf(m01, m23);
          const unsigned idx = bf2 ? (2u | (b23 ? 1u : 0u)) : (b01 ? 1u : 0u);
          dec |= idx << (2 * k);
          pm[k] = best;
          lmax = fmaxf(lmax, best);
        }
        tb[((i & (RING - 1)) << 5) + lane] = (uint8_t)dec;
        const bool blk = i == next_rel;
        const bool rel = blk || i == n_steps - 1;
        if (rel || NORM == 1 || (u % NORM) == NORM - 1) {
          const float wmax = warp_max_f32(lmax);
          if (rel) {
            int best_s = 0x7FFFFFFF;  // lowest state index v M + h with the maximum metric
#pragma unroll
            for (int k = 1; k >= 0; --k)
              if (pm[k] == wmax) best_s = (base_lane + LA) * M + 2 * hh + k;
            const int wbest = warp_min_s32(best_s);
            __syncwarp();  // this step's decision bytes are visible to the whole warp
            if (blk) {
              trace_release(i, wbest >> 2, wbest & 3, released, i + 1 - D);
              released = i + 1 - D;
              next_rel += D;
            }
            if (i == n_steps - 1) trace_release(i, wbest >> 2, wbest & 3, released, n_steps);
          }
          pm[0] -= wmax;
          pm[1] -= wmax;
        }
      }
    }
  }
  // per-lane counts -> lane 0
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sym_err += __shfl_xor_sync(0xFFFFFFFFu, sym_err, o);
    bit_err += __shfl_xor_sync(0xFFFFFFFFu, bit_err, o);
    compared += __shfl_xor_sync(0xFFFFFFFFu, compared, o);
  }
  if (lane == 0 && p.counters) {
    const int64_t pt = p.point ? p.point[frame] : 0;
    atomicAdd(&p.counters[3 * pt + 0], (unsigned long long)bit_err);
    atomicAdd(&p.counters[3 * pt + 1], (unsigned long long)sym_err);
    atomicAdd(&p.counters[3 * pt + 2], (unsigned long long)compared);
  }
}

// ---- host side ------------------------------------------------------------------------------------
static int fill_cpm_params(CpmParams& q, int M, int Lp, int p, const int32_t* h_num, int n_h, int depth) {
  WF_REQUIRE(M >= 2 && M <= 16 && (M & (M - 1)) == 0, "alphabet size M must be a power of two in 2..16");
  WF_REQUIRE(Lp >= 1 && Lp <= 4 && p >= 1 && 2 * p <= 64, "need 1 <= Lp <= 4 and p <= 32");
  WF_REQUIRE(n_h >= 1 && n_h <= kCpmMaxH && h_num, "n_h outside 1..8");
  memset(&q, 0, sizeof(q));
  q.M = M; q.Lp = Lp; q.p = p; q.n_h = n_h;
  q.bits = 0;
  while ((1 << q.bits) < M) ++q.bits;
  q.n_filt = 1;
  for (int i = 0; i < Lp; ++i) q.n_filt *= M;
  q.n_hist = q.n_filt / M;
  q.n_states = 2 * p * q.n_hist;
  q.n_cols = n_h * q.n_filt;
  WF_REQUIRE(q.n_states <= kCpmMaxStates, "trellis has %d states (max %d)", q.n_states, kCpmMaxStates);
  WF_REQUIRE(q.n_cols <= kCpmMaxCols, "bank has %d outputs per step (max %d)", q.n_cols, kCpmMaxCols);
  int dflt = 32 / q.bits;  // release block D; survivor memory 2 D symbols
  if (dflt > 16) dflt = 16;
  q.depth = depth > 0 ? depth : dflt;
  for (int h = 0; h < n_h; ++h)
    for (int x = 0; x < M; ++x) {
      int d = (h_num[h] * (2 * x - (M - 1))) % (2 * p);
      if (d < 0) d += 2 * p;
      q.delta[h][x] = d;
    }
  // Every branch of a step moves v by K a with a odd, i.e. flips the parity of v iff K is odd: the trellis is
  // bipartite and only one parity class of phase states is reachable at a time.  The frame starts with no symbol in
  // the air (v = 0 once the Lp - 1 phantom symbols have left the window); walking those steps backwards gives the
  // class at the start.  States of the other class start at -BIG and never recover (oracle/cpm_generic.py).
  bool any_odd = false;
  for (int h = 0; h < n_h; ++h) any_odd |= (h_num[h] & 1) != 0;
  q.start_parity = -1;
  if (any_odd) {
    int par = 0;
    for (int i = Lp - 2; i >= 0; --i) {
      int idx = (i - Lp + 1) % n_h;
      if (idx < 0) idx += n_h;
      par = (par + (h_num[idx] & 1)) & 1;  // minus = plus, mod 2
    }
    q.start_parity = par;
  }
  return WF_OK;
}

static void fill_rot(CpmRot& rot, int p) {
  memset(&rot, 0, sizeof(rot));
  for (int v = 0; v < 2 * p; ++v) {
    rot.c[v] = (float)cos(M_PI * (double)v / (double)p);
    rot.s[v] = (float)sin(M_PI * (double)v / (double)p);
  }
}

// which fused kernel (0 = none, 1 = PCM/FM thread-per-frame, 2 = ARTM lane-per-phase-state) covers the shape.
// taps: [n_h * M^Lp][sps] in wf_mf_bank's convolution form with center = sps - 1 (tap t multiplies rx[c_i + sps-1 - t]).
static int fused_kind(int dtype, int64_t Tn, int64_t ld_rx, const double* tre, const double* tim, int sps, int64_t first_instant,
                      int64_t n_steps, const CpmParams& q) {
  if (dtype != WF_F32 || sps != 8 || first_instant != -1 || !tre || !tim || q.Lp != 2) return 0;
  if ((ld_rx & 1) || n_steps < 1 || n_steps >= 0x0FFFFFF0 || Tn >= 0x7FFFFFF0) return 0;
  const int NF = q.n_filt;
  // hypothesis NF-1-f must be the complex conjugate of f, exactly, in float
  for (int par = 0; par < q.n_h; ++par)
    for (int f = 0; f < NF / 2; ++f)
      for (int t = 0; t < 8; ++t) {
        const int a = (par * NF + f) * 8 + t, b = (par * NF + NF - 1 - f) * 8 + t;
        if ((float)tre[a] != (float)tre[b] || (float)tim[a] != -(float)tim[b]) return 0;
      }
  if (q.M == 2 && q.n_h == 1 && q.p == 10 && q.delta[0][0] == 13 && q.delta[0][1] == 7 && q.depth <= 32 &&
      8 * n_steps <= Tn && 8 * ((n_steps + kBfCS - 1) / kBfCS) * kBfCS <= ld_rx)  // whole chunks are staged
    return 1;
  // the lane layout is built on the class pattern of h = {even, odd} / 16 (ARTM: 4/16, 5/16): start class 1, flips on even steps
  if (q.M == 4 && q.n_h == 2 && q.p == 16 && q.depth <= 32 && (q.delta[0][0] & 1) == 0 && (q.delta[1][0] & 1) == 1 &&
      q.start_parity == 1)
    return 2;
  return 0;
}

}  // namespace wf

extern "C" int wf_viterbi_cpm(const void* d_z, int dtype, int64_t B, int64_t n_steps, int64_t ld_z, int M, int Lp,
                              int p, const int32_t* h_num, int n_h, int depth, int8_t* d_dec_syms,
                              const int8_t* d_tx_syms, int64_t S, int64_t skip_head, int64_t skip_tail,
                              const int32_t* d_point, int64_t* d_counters, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_steps >= 0 && ld_z >= n_steps, "bad shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  CpmParams q;
  int rc;
  if ((rc = fill_cpm_params(q, M, Lp, p, h_num, n_h, depth))) return rc;
  WF_REQUIRE(q.depth >= 1 && 2 * q.depth * q.bits <= 64, "survivor memory 2 x depth x bits/symbol exceeds 64 bits");
  WF_REQUIRE(skip_head >= 0 && skip_tail >= 0, "negative skip");
  WF_REQUIRE(!d_tx_syms || d_counters, "error counting needs d_counters");
  if ((rc = require_device())) return rc;
  if (B == 0 || n_steps == 0) return WF_OK;
  WF_REQUIRE(d_z, "d_z is NULL");
  q.z = d_z; q.dec_syms = d_dec_syms; q.tx_syms = d_tx_syms; q.point = d_point;
  q.counters = reinterpret_cast<unsigned long long*>(d_counters);
  q.B = B; q.n_steps = n_steps; q.ld_z = ld_z; q.S = S; q.skip_head = skip_head; q.skip_tail = skip_tail;
  const unsigned grid = (unsigned)((B + kCpmWarps - 1) / kCpmWarps);
  if (dtype == WF_F32) viterbi_cpm_kernel<float><<<grid, kCpmWarps * 32, 0, as_stream(stream)>>>(q);
  else viterbi_cpm_kernel<double><<<grid, kCpmWarps * 32, 0, as_stream(stream)>>>(q);
  WF_LAUNCH_CHECK("viterbi_cpm_kernel");
  return WF_OK;
}

extern "C" int wf_cpm_detect_supported(int dtype, int64_t T, int64_t ld_rx, const double* taps_re, const double* taps_im,
                                       int sps, int64_t first_instant, int64_t n_steps, int M, int Lp, int p,
                                       const int32_t* h_num, int n_h, int depth) {
  using namespace wf;
  CpmParams q;
  if (fill_cpm_params(q, M, Lp, p, h_num, n_h, depth)) return 0;
  return fused_kind(dtype, T, ld_rx, taps_re, taps_im, sps, first_instant, n_steps, q) != 0;
}

extern "C" int wf_cpm_detect(const void* d_rx, int dtype, int64_t B, int64_t T, int64_t ld_rx, const double* taps_re,
                             const double* taps_im, int sps, int64_t first_instant, int64_t n_steps, int M, int Lp,
                             int p, const int32_t* h_num, int n_h, int depth, int8_t* d_dec_syms,
                             const int8_t* d_tx_syms, int64_t S, int64_t skip_head, int64_t skip_tail,
                             const int32_t* d_point, int64_t* d_counters, int tune, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_steps >= 0 && T >= 0 && ld_rx >= T, "bad shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  CpmParams q;
  int rc;
  if ((rc = fill_cpm_params(q, M, Lp, p, h_num, n_h, depth))) return rc;
  WF_REQUIRE(skip_head >= 0 && skip_tail >= 0, "negative skip");
  WF_REQUIRE(!d_tx_syms || d_counters, "error counting needs d_counters");
  const int kind = fused_kind(dtype, T, ld_rx, taps_re, taps_im, sps, first_instant, n_steps, q);
  if (!kind)
    return fail(WF_ERR_UNSUPPORTED, "wf_cpm_detect: shape not covered by a fused kernel (see wf_cpm_detect_supported); "
                                    "run wf_mf_bank + wf_viterbi_cpm");
  if ((rc = require_device())) return rc;
  if (B == 0 || n_steps == 0) return WF_OK;
  WF_REQUIRE(d_rx, "d_rx is NULL");
  WF_REQUIRE((reinterpret_cast<uintptr_t>(d_rx) & 15) == 0, "d_rx must be 16-byte aligned");
  q.rx = d_rx; q.dec_syms = d_dec_syms; q.tx_syms = d_tx_syms; q.point = d_point;
  q.counters = reinterpret_cast<unsigned long long*>(d_counters);
  q.B = B; q.n_steps = n_steps; q.S = S; q.skip_head = skip_head; q.skip_tail = skip_tail; q.Tn = T; q.ld_rx = ld_rx;
  CpmRot rot;
  fill_rot(rot, p);
  cudaStream_t st = as_stream(stream);
  const int norm_every = tune > 0 ? tune : 1;  // path-metric normalisation every `tune` steps (1, 2, 4 or 8)
  if (kind == 1) {
    CpmPairTaps tp;
    for (int pr = 0; pr < 2; ++pr)
      for (int u = 0; u < 8; ++u) {  // ct[f][u] = taps[f][7 - u]
        const float cr = (float)taps_re[pr * 8 + 7 - u], ci = (float)taps_im[pr * 8 + 7 - u];
        tp.t[pr][u][0][0] = tp.t[pr][u][0][1] = cr;
        tp.t[pr][u][1][0] = tp.t[pr][u][1][1] = ci;
      }
    const unsigned g = (unsigned)((B + kBfThreads - 1) / kBfThreads);
#define WF_BF_LAUNCH(RING_, NORM_)                                                                            \
  do {                                                                                                        \
    const size_t sm = (size_t)RING_ * kBfThreads * sizeof(uint32_t);                                          \
    auto kfn = cpm_detect_m2l2_kernel<20, 13, 7, RING_, NORM_>;                                               \
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);          \
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(cpm_detect_m2l2_kernel)");                \
    kfn<<<g, kBfThreads, sm, st>>>(q, rot, tp);                                                               \
  } while (0)
    if (q.depth <= 16) {
      if (norm_every >= 4) WF_BF_LAUNCH(32, 4);
      else if (norm_every == 2) WF_BF_LAUNCH(32, 2);
      else WF_BF_LAUNCH(32, 1);
    } else {
      if (norm_every >= 4) WF_BF_LAUNCH(64, 4);
      else if (norm_every == 2) WF_BF_LAUNCH(64, 2);
      else WF_BF_LAUNCH(64, 1);
    }
#undef WF_BF_LAUNCH
    WF_LAUNCH_CHECK("cpm_detect_m2l2_kernel");
    return WF_OK;
  }
  CpmLaneTaps tp;
  for (int par = 0; par < 2; ++par)
    for (int f = 0; f < 8; ++f)
      for (int u = 0; u < 8; ++u) {
        tp.t[par][f][u][0] = (float)taps_re[(par * 16 + f) * 8 + 7 - u];
        tp.t[par][f][u][1] = (float)taps_im[(par * 16 + f) * 8 + 7 - u];
      }
  const unsigned g = (unsigned)((B + kLfWarps - 1) / kLfWarps);
  if (q.depth <= 16) {
    if (norm_every >= 4) cpm_detect_lanev_kernel<32, 4><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
    else if (norm_every == 2) cpm_detect_lanev_kernel<32, 2><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
    else cpm_detect_lanev_kernel<32, 1><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
  } else {
    if (norm_every >= 4) cpm_detect_lanev_kernel<64, 4><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
    else if (norm_every == 2) cpm_detect_lanev_kernel<64, 2><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
    else cpm_detect_lanev_kernel<64, 1><<<g, kLfWarps * 32, 0, st>>>(q, rot, tp);
  }
  WF_LAUNCH_CHECK("cpm_detect_lanev_kernel");
  return WF_OK;
}
