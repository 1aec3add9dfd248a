// cpmdet.cu -- A11/A12: generic pulse-truncated coherent CPM trellis detector
// (PCM/FM, ARTM multi-h CPM, any M-ary single/multi-h CPM).  The reference has NO detector
// for these waveforms (waveforms.viterbi exports only SOQPSKTrellisDetector); the design
// generalises its pulse-truncation bank + Re{e^{-j theta} z} metric
// (examples/soqpsk_detection.py:134-156, viterbi/algorithm.py:57-63) and is specified by the
// float64 statement oracle/cpm_generic.py (parity unpinned).
//
// One WARP per frame, lane-per-state: state s = v * M^(Lp-1) + hist with v the phase state
// (theta = pi v / p, v in Z_2p) and hist the Lp-1 newest symbols; lane l owns states l, l+32, ...
// Path metrics and REGISTER-EXCHANGE survivor words (the last `depth` symbols of each survivor,
// 64 bits) ping-pong in shared memory, so there is no traceback: each step emits the oldest symbol
// of the best survivor.  Per step the warp stages the bank outputs of the step, every lane does M
// add-compare-selects per owned state, then two warp reductions (max for normalisation, argmax for
// the output) run on shuffles.
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
  const void* z;
  int8_t* dec_syms;
  const int8_t* tx_syms;
  const int32_t* point;
  unsigned long long* counters;
  int64_t B, n_steps, ld_z, S, skip_head, skip_tail;
  int M, Lp, p, n_h, depth, bits, n_filt, n_hist, n_states, n_cols;
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
  const unsigned long long hmask = (D * bits >= 64) ? ~0ull : ((1ull << (D * bits)) - 1ull);
  const unsigned sym_mask = (1u << bits) - 1u;
  const C* zrow = reinterpret_cast<const C*>(p.z) + frame * p.ld_z * p.n_cols;
  const int8_t* tx = p.tx_syms ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = p.dec_syms ? p.dec_syms + frame * p.n_steps : nullptr;
  const int64_t n_steps = p.n_steps;
  const int64_t cmp_lo = p.skip_head, cmp_hi = min(p.S, n_steps) - p.skip_tail;

  T* pm[2] = {pm_s[wid][0], pm_s[wid][1]};
  unsigned long long* hs[2] = {hist_s[wid][0], hist_s[wid][1]};
  for (int s = lane; s < NS; s += 32) { pm[0][s] = (T)0; hs[0][s] = 0ull; }
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
    if (lane == 0 && i >= D - 1) emit(i - D + 1, (unsigned)((hs[cur][best_s] >> ((D - 1) * bits)) & sym_mask));
    if (i == n_steps - 1 && lane == 0) {  // flush the tail from the best final survivor
      const unsigned long long h = hs[cur][best_s];
      for (int d = 0; d < D - 1 && d < n_steps; ++d) emit(n_steps - 1 - d, (unsigned)((h >> (d * bits)) & sym_mask));
    }
  }
  if (lane == 0 && p.counters) {
    const int64_t pt = p.point ? p.point[frame] : 0;
    atomicAdd(&p.counters[3 * pt + 0], bit_err);
    atomicAdd(&p.counters[3 * pt + 1], sym_err);
    atomicAdd(&p.counters[3 * pt + 2], compared);
  }
}

// ---------------------------------------------------------------------------
// Binary, memory-2, single-h shape (IRIG PCM/FM: M = 2, Lp = 2, h = K/p): the same recursion with
// one THREAD per frame.  The 2 * P2 path metrics live in registers (old and new generation), every
// predecessor index, filter index and rotation is a compile-time constant (rotations are
// constant-bank operands), and only the 64-bit survivor words ping-pong through shared memory,
// laid out [state][thread] so a warp's LDS.64 / STS.64 are conflict free.  Per frame-step this is
// ~25 instructions instead of the ~400 of the lane-per-state kernel (which spends a whole warp, two
// shuffle reductions and run-time index arithmetic on one frame).  Semantics are identical: strict
// '>' keeps x = 0 on ties, the best state is the LOWEST index with the maximum metric, metrics
// are normalised by the maximum, the oldest survivor symbol of the best state is emitted.
// ---------------------------------------------------------------------------
constexpr int kM2Threads = 64;

// SV: survivor word, uint32_t when depth <= 32 (the default: half the shared memory -- 7 blocks per SM, so the
// 1024 blocks of a 65 536-frame batch are ONE wave instead of 1.4 -- and half the select / shift work)
template <int P2, int D0, int D1, typename SV>
__global__ void __launch_bounds__(kM2Threads, sizeof(SV) == 4 ? 7 : 5)
viterbi_cpm_m2l2_kernel(const __grid_constant__ CpmParams p, const __grid_constant__ CpmRot rot) {
  constexpr int NS = 2 * P2;
  __shared__ SV sv_s[2][NS][kM2Threads];
  const int tid = threadIdx.x;
  const int64_t frame = (int64_t)blockIdx.x * kM2Threads + tid;
  if (frame >= p.B) return;  // no barriers in this kernel: every thread is autonomous

  const int D = p.depth;
  const SV hmask = (D >= (int)(8 * sizeof(SV))) ? (SV)~(SV)0 : (SV)(((SV)1 << D) - (SV)1);
  const float4* zrow = reinterpret_cast<const float4*>(p.z) + frame * p.ld_z * 2;  // 4 complex64 = 2 float4 per step
  const int8_t* tx = p.tx_syms ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = p.dec_syms ? p.dec_syms + frame * p.n_steps : nullptr;
  const int64_t n_steps = p.n_steps;
  const int64_t cmp_lo = p.skip_head, cmp_hi = min(p.S, n_steps) - p.skip_tail;

  float pm[NS], pn[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) { pm[s] = 0.f; sv_s[0][s][tid] = (SV)0; }

  unsigned long long sym_err = 0, compared = 0;
  auto emit = [&](int64_t idx, unsigned v) {
    if (out) out[idx] = (int8_t)(2 * (int)v - 1);
    if (tx && idx >= cmp_lo && idx < cmp_hi) {
      const unsigned t = (unsigned)(((int)tx[idx] + 1) >> 1);
      sym_err += (t != v) ? 1u : 0u;
      ++compared;
    }
  };

  int cur = 0;
  float4 za = n_steps > 0 ? zrow[0] : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 zb = n_steps > 0 ? zrow[1] : make_float4(0.f, 0.f, 0.f, 0.f);
  for (int64_t i = 0; i < n_steps; ++i) {
    const float zx[4] = {za.x, za.z, zb.x, zb.z}, zy[4] = {za.y, za.w, zb.y, zb.w};
    if (i + 1 < n_steps) { za = zrow[2 * (i + 1)]; zb = zrow[2 * (i + 1) + 1]; }  // next step's sector, a step ahead
    const SV(*ho)[kM2Threads] = sv_s[cur];
    SV(*hn)[kM2Threads] = sv_s[cur ^ 1];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      constexpr int dummy = 0; (void)dummy;
      const int vnew = s >> 1, hnew = s & 1;
      const int vp0 = (vnew - D0 + 2 * P2) % P2, vp1 = (vnew - D1 + 2 * P2) % P2;
      const int pred0 = vp0 * 2, pred1 = vp1 * 2 + 1;
      const int f0 = hnew * 2, f1 = hnew * 2 + 1;
      // pm + Re{e^{-j pi v/p} z} = pm + (cos * zr + sin * zi), evaluated in this order (as in the generic kernel)
      const float c0 = pm[pred0] + (rot.c[vp0] * zx[f0] + rot.s[vp0] * zy[f0]);
      const float c1 = pm[pred1] + (rot.c[vp1] * zx[f1] + rot.s[vp1] * zy[f1]);
      const bool win1 = c1 > c0;
      pn[s] = win1 ? c1 : c0;
      const SV h = win1 ? ho[pred1][tid] : ho[pred0][tid];
      hn[s][tid] = (SV)(((SV)(h << 1) | (SV)hnew) & hmask);
    }
    float wmax = pn[0];
#pragma unroll
    for (int s = 1; s < NS; ++s) wmax = pn[s] > wmax ? pn[s] : wmax;
    int best_s = NS - 1;
#pragma unroll
    for (int s = NS - 2; s >= 0; --s) best_s = (pn[s] == wmax) ? s : best_s;  // lowest index wins
    if (pn[NS - 1] != wmax && best_s == NS - 1) best_s = 0;                    // unreachable (wmax is one of them)
#pragma unroll
    for (int s = 0; s < NS; ++s) pm[s] = pn[s] - wmax;
    cur ^= 1;
    if (i >= D - 1) emit(i - D + 1, (unsigned)((sv_s[cur][best_s][tid] >> (D - 1)) & (SV)1));
    if (i == n_steps - 1) {  // flush the tail from the best final survivor
      const SV h = sv_s[cur][best_s][tid];
      for (int d = 0; d < D - 1 && d < n_steps; ++d) emit(n_steps - 1 - d, (unsigned)((h >> d) & (SV)1));
    }
  }
  if (p.counters) {
    const int64_t pt = p.point ? p.point[frame] : 0;
    atomicAdd(&p.counters[3 * pt + 0], sym_err);  // binary alphabet: bit errors = symbol errors
    atomicAdd(&p.counters[3 * pt + 1], sym_err);
    atomicAdd(&p.counters[3 * pt + 2], compared);
  }
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
// Memory-2 trellises with 2p = 32 phase states (ARTM multi-h: M = 4, h = {4, 5}/16 -> 128 states):
// one WARP per frame, lane = phase state v, the M history states of a lane in REGISTERS.
// For Lp = 2 the predecessor of (v, h_new) through oldest symbol x is (v - delta[x], x): the same
// lane and register for every h_new, so a step needs only M metric shuffles and M survivor
// shuffles (lane-varying source = v - delta[x]); rotations and source lanes are per-lane registers
// for each modulation-index parity, the step's M^2 bank outputs come from a 4-deep cp.async ring
// read by broadcast LDS.128.  No shared-memory metric/survivor ping-pong, no run-time index
// arithmetic, transmitted symbols fetched 32 at a time.  Same semantics as viterbi_cpm_kernel.
// ---------------------------------------------------------------------------
constexpr int kLvWarps = 4;
constexpr int kLvRing = 4;

template <int M, int NHH>
__global__ void __launch_bounds__(kLvWarps * 32)
viterbi_cpm_lanev_kernel(const __grid_constant__ CpmParams p, const __grid_constant__ CpmRot rot) {
  constexpr int NF = M * M;                 // bank outputs per step and parity
  constexpr int BITS = M == 2 ? 1 : (M == 4 ? 2 : (M == 8 ? 3 : 4));
  __align__(16) __shared__ float2 z_s[kLvWarps][kLvRing][NF];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t frame = (int64_t)blockIdx.x * kLvWarps + wid;
  if (frame >= p.B) return;  // whole warp; only warp-level synchronisation below

  const int D = p.depth;
  const unsigned long long hmask = (D * BITS >= 64) ? ~0ull : ((1ull << (D * BITS)) - 1ull);
  const int n_cols = p.n_cols;
  const int8_t* tx = p.tx_syms ? p.tx_syms + frame * p.S : nullptr;
  int8_t* out = p.dec_syms ? p.dec_syms + frame * p.n_steps : nullptr;
  const int n_steps = (int)p.n_steps, S = (int)p.S;  // < 2^31 (checked by the launcher): 32-bit index arithmetic
  const int cmp_lo = (int)p.skip_head, cmp_hi = min(S, n_steps) - (int)p.skip_tail;
  const int sh_old = (D - 1) * BITS;  // position of the oldest symbol of a survivor word

  // per-lane transition tables for each modulation-index parity
  int src[NHH][M];
  float rc[NHH][M], rs[NHH][M];
#pragma unroll
  for (int h = 0; h < NHH; ++h)
#pragma unroll
    for (int x = 0; x < M; ++x) {
      src[h][x] = (lane - p.delta[h][x]) & 31;
      rc[h][x] = rot.c[src[h][x]];
      rs[h][x] = rot.s[src[h][x]];
#ifndef WF_HOSTEMU
      // pin the values in registers: left alone, the compiler re-reads rot.c[src] in the loop, a
      // lane-divergent constant-bank access that serialises 32 ways (13 stall cycles per issue in ncu)
      asm volatile("" : "+r"(src[h][x]), "+f"(rc[h][x]), "+f"(rs[h][x]));
#endif
    }

  float pm[M];
  unsigned long long sv[M];
#pragma unroll
  for (int h = 0; h < M; ++h) { pm[h] = 0.f; sv[h] = 0ull; }

  // Decisions come out in ascending order, one per step: lane (idx mod 32) keeps decision idx, and once
  // a group of 32 is complete every lane stores / compares ITS symbol (one coalesced 32-byte store and
  // one 32-byte load per 32 steps; no per-step divergent branch, byte store or shuffle).
  unsigned sym_err = 0, bit_err = 0, compared = 0;  // per lane, reduced at the end
  unsigned dec_reg = 0;
  auto flush_group = [&](int last_idx) {  // decisions (last_idx & ~31) .. last_idx are in the lanes' dec_reg
    const int k = (last_idx & ~31) + lane;
    if (k <= last_idx) {
      if (out) out[k] = (int8_t)(2 * (int)dec_reg - (M - 1));
      if (tx && k >= cmp_lo && k < cmp_hi) {
        const unsigned t = (unsigned)(((int)tx[k] + (M - 1)) >> 1);
        sym_err += (t != dec_reg) ? 1u : 0u;
        bit_err += __popc((t ^ dec_reg) & (unsigned)(M - 1));
        compared += BITS;  // column 2 counts compared BITS for every waveform
      }
    }
  };
  auto emit = [&](int idx, unsigned v) {  // executed by every lane with the same (idx, v)
    dec_reg = (lane == (idx & 31)) ? v : dec_reg;
    if ((idx & 31) == 31 || idx == n_steps - 1) flush_group(idx);
  };

  // bank outputs of step j: NF complex64 at column (j mod NHH) NF of row j; 16 bytes per lane, NF/2 lanes
  const float2* zst = reinterpret_cast<const float2*>(p.z) + frame * p.ld_z * n_cols + 2 * lane;
  float2* zdst = &z_s[wid][0][2 * lane];
  int st_j = 0;  // next step to stage
  auto stage = [&](int par) {
    if (st_j < n_steps && lane < NF / 2) cp_async_cg16(zdst + (st_j & (kLvRing - 1)) * NF, zst + par * NF);
    cp_async_commit();
    zst += n_cols;
    ++st_j;
  };
  static_assert((kLvRing & (kLvRing - 1)) == 0 && kLvRing % NHH == 0, "ring slots: power of two, multiple of the h period");
#pragma unroll
  for (int k = 0; k < kLvRing - 1; ++k) stage(k % NHH);

  for (int i0 = 0; i0 < n_steps; i0 += NHH) {
#pragma unroll
    for (int hp = 0; hp < NHH; ++hp) {
      const int i = i0 + hp;
      if (i >= n_steps) break;
      // parity of the filters is i % NHH = hp (i0 is a multiple of NHH); of the transition (i - Lp + 1) % NHH
      constexpr int dummy = 0; (void)dummy;
      const int lp = (hp + NHH - 1) % NHH;
      __syncwarp();              // every lane has finished reading the slot that is staged next
      stage((hp + kLvRing - 1) % NHH);
      cp_async_wait<kLvRing - 1>();
      __syncwarp();              // the step's bank outputs are visible to all lanes
      const float4* zq = reinterpret_cast<const float4*>(z_s[wid][i & (kLvRing - 1)]);
      float pmx[M];
      unsigned long long svx[M];
#pragma unroll
      for (int x = 0; x < M; ++x) {
        pmx[x] = __shfl_sync(0xFFFFFFFFu, pm[x], src[lp][x]);
        svx[x] = __shfl_sync(0xFFFFFFFFu, sv[x], src[lp][x]);
      }
      float lmax = -3.0e38f;
#pragma unroll
      for (int hn = 0; hn < M; ++hn) {
        float best = 0.f;
        unsigned long long bsv = 0ull;
#pragma unroll
        for (int x = 0; x < M; x += 2) {
          const float4 q = zq[(hn * M + x) / 2];  // two complex outputs: f = hn M + x, f + 1
          // pm + Re{e^{-j pi v/p} z} = pm + (cos * zr + sin * zi), evaluated in this order
          const float ca = pmx[x] + (rc[lp][x] * q.x + rs[lp][x] * q.y);
          const float cb = pmx[x + 1] + (rc[lp][x + 1] * q.z + rs[lp][x + 1] * q.w);
          if (x == 0 || ca > best) { best = ca; bsv = svx[x]; }
          if (cb > best) { best = cb; bsv = svx[x + 1]; }
        }
        pm[hn] = best;
        sv[hn] = ((bsv << BITS) | (unsigned long long)hn) & hmask;
        lmax = best > lmax ? best : lmax;
      }
      const float wmax = warp_max_f32(lmax);
      int best_s = 0x7FFFFFFF;  // lowest state index v M + h with the maximum metric
      unsigned long long my_sv = 0ull;
#pragma unroll
      for (int hn = M - 1; hn >= 0; --hn) {
        if (pm[hn] == wmax) { best_s = lane * M + hn; my_sv = sv[hn]; }
        pm[hn] -= wmax;
      }
      const int wbest = warp_min_s32(best_s);
      if (i >= D - 1 || i == n_steps - 1) {
        const unsigned long long h = __shfl_sync(0xFFFFFFFFu, my_sv, wbest / M);  // the best lane's lowest-h survivor
        if (i >= D - 1) emit(i - D + 1, (unsigned)(h >> sh_old) & (unsigned)(M - 1));
        if (i == n_steps - 1)  // flush the tail from the best final survivor
          for (int d = D - 2 < n_steps - 1 ? D - 2 : n_steps - 1; d >= 0; --d)
            emit(n_steps - 1 - d, (unsigned)(h >> (d * BITS)) & (unsigned)(M - 1));
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

}  // namespace wf

extern "C" int wf_viterbi_cpm(const void* d_z, int dtype, int64_t B, int64_t n_steps, int64_t ld_z, int M, int Lp,
                              int p, const int32_t* h_num, int n_h, int depth, int8_t* d_dec_syms,
                              const int8_t* d_tx_syms, int64_t S, int64_t skip_head, int64_t skip_tail,
                              const int32_t* d_point, int64_t* d_counters, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_steps >= 0 && ld_z >= n_steps, "bad shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  WF_REQUIRE(M >= 2 && M <= 16 && (M & (M - 1)) == 0, "alphabet size M must be a power of two in 2..16");
  WF_REQUIRE(Lp >= 1 && Lp <= 4 && p >= 1 && 2 * p <= 64, "need 1 <= Lp <= 4 and p <= 32");
  WF_REQUIRE(n_h >= 1 && n_h <= kCpmMaxH && h_num, "n_h outside 1..8");
  CpmParams q;
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
  q.depth = depth > 0 ? depth : (64 / q.bits < 32 ? 64 / q.bits : 32);  // default: 32 symbols
  WF_REQUIRE(q.depth >= 1 && q.depth * q.bits <= 64, "survivor depth x bits/symbol exceeds 64");
  WF_REQUIRE(skip_head >= 0 && skip_tail >= 0, "negative skip");
  WF_REQUIRE(!d_tx_syms || d_counters, "error counting needs d_counters");
  for (int h = 0; h < n_h; ++h)
    for (int x = 0; x < M; ++x) {
      int d = (h_num[h] * (2 * x - (M - 1))) % (2 * p);
      if (d < 0) d += 2 * p;
      q.delta[h][x] = d;
    }
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || n_steps == 0) return WF_OK;
  WF_REQUIRE(d_z, "d_z is NULL");
  q.z = d_z; q.dec_syms = d_dec_syms; q.tx_syms = d_tx_syms; q.point = d_point;
  q.counters = reinterpret_cast<unsigned long long*>(d_counters);
  q.B = B; q.n_steps = n_steps; q.ld_z = ld_z; q.S = S; q.skip_head = skip_head; q.skip_tail = skip_tail;
  // IRIG PCM/FM shape (binary, memory 2, h = 7/10, 4-complex rows on 16-byte boundaries): thread-per-frame kernel
  if (dtype == WF_F32 && M == 2 && Lp == 2 && n_h == 1 && p == 10 && q.delta[0][0] == 13 && q.delta[0][1] == 7 &&
      (reinterpret_cast<uintptr_t>(d_z) & 15) == 0) {
    CpmRot rot;
    for (int v = 0; v < 2 * p; ++v) {
      rot.c[v] = (float)cos(M_PI * (double)v / (double)p);
      rot.s[v] = (float)sin(M_PI * (double)v / (double)p);
    }
    const unsigned g = (unsigned)((B + kM2Threads - 1) / kM2Threads);
    if (q.depth <= 32) viterbi_cpm_m2l2_kernel<20, 13, 7, uint32_t><<<g, kM2Threads, 0, as_stream(stream)>>>(q, rot);
    else viterbi_cpm_m2l2_kernel<20, 13, 7, unsigned long long><<<g, kM2Threads, 0, as_stream(stream)>>>(q, rot);
    WF_LAUNCH_CHECK("viterbi_cpm_m2l2_kernel");
    return WF_OK;
  }
  // memory-2 trellis with 32 phase states (ARTM multi-h): lane-per-phase-state register kernel
  if (dtype == WF_F32 && M == 4 && Lp == 2 && p == 16 && n_h == 2 && q.depth * q.bits <= 64 && n_steps < 0x7FFFFF00 && S < 0x7FFFFF00 &&
      (reinterpret_cast<uintptr_t>(d_z) & 15) == 0) {
    CpmRot rot;
    for (int v = 0; v < 2 * p; ++v) {
      rot.c[v] = (float)cos(M_PI * (double)v / (double)p);
      rot.s[v] = (float)sin(M_PI * (double)v / (double)p);
    }
    const unsigned g = (unsigned)((B + kLvWarps - 1) / kLvWarps);
    viterbi_cpm_lanev_kernel<4, 2><<<g, kLvWarps * 32, 0, as_stream(stream)>>>(q, rot);
    WF_LAUNCH_CHECK("viterbi_cpm_lanev_kernel");
    return WF_OK;
  }
  const unsigned grid = (unsigned)((B + kCpmWarps - 1) / kCpmWarps);
  if (dtype == WF_F32) viterbi_cpm_kernel<float><<<grid, kCpmWarps * 32, 0, as_stream(stream)>>>(q);
  else viterbi_cpm_kernel<double><<<grid, kCpmWarps * 32, 0, as_stream(stream)>>>(q);
  WF_LAUNCH_CHECK("viterbi_cpm_kernel");
  return WF_OK;
}
