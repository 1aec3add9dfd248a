// viterbi.cu -- A10 SOQPSK trellis detector + A13 error count.
//
// Restates SOQPSKTrellisDetector.iteration (viterbi/algorithm.py:44-101) as driven
// by examples/soqpsk_detection.py:177-209: one detector call per symbol instant,
// element [0] of each call kept, first `length` outputs dropped, decisions
// compared with the transmitted bits/symbols.
//
// One THREAD per frame: the 4 path metrics, the previous increments and the
// decision word live in registers; add-compare-select is straight-line FADD /
// FSETP / FSEL with the trellis folded in as immediates (the 4-state x 2-column
// tables are re-derived from the state convention state = 2*I + Q, not stored).
// The [frame][call][3] matched-filter rows are staged through shared memory in
// chunks of 8 calls so global reads are 192-byte contiguous runs while each
// thread walks its own row.  Order of float operations follows the reference:
// normalise (M - min M) BEFORE adding increments (:65), strict '<' keeps the first
// minimum in branch-list order (:82), argmin takes the lowest state on ties (:92).
#include "common.cuh"
#include <string.h>
#include <type_traits>

namespace wf {

constexpr int kVitThreads = 128;
template <typename T> struct VitChunk { static constexpr int n = sizeof(T) == 4 ? 8 : 4; };  // calls staged per pass
constexpr int kVitMaxLen = 16;

// ---- trellis algebra (cpm/trellis/model.py:205-258, re-derived) ---------------
// column parity c: 0 = even/I rail updates, 1 = odd/Q rail.  state = 2*I + Q.
// Branch list order inside a column: b = 2*start + new_bit.
__host__ __device__ constexpr int tr_end(int c, int s, int nb) { return c == 0 ? ((nb << 1) | (s & 1)) : ((s & 2) | nb); }
__host__ __device__ constexpr int tr_out(int c, int s, int nb) {
  return c == 0 ? -2 * (2 * (s & 1) - 1) * (nb - (s >> 1)) : 2 * (2 * (s >> 1) - 1) * (nb - (s & 1));
}
__host__ __device__ constexpr int tr_inp(int c, int s, int nb, bool diff) {
  return diff ? (nb ^ (c == 0 ? (s >> 1) : (s & 1))) : nb;
}
// k-th (k = 0, 1) branch in list order that ends in state st: its start state and new bit
__host__ __device__ constexpr int tr_pred(int c, int st, int k) { return c == 0 ? ((k << 1) | (st & 1)) : ((st & 2) | k); }
__host__ __device__ constexpr int tr_newbit(int c, int st) { return c == 0 ? (st >> 1) : (st & 1); }

// Re(state_exp_term[s] * z[row]) with state_exp_term = [+j, -1, +1, -j] (algorithm.py:30,57-63)
template <typename T>
__device__ __forceinline__ T branch_inc(int s, const T* zr, const T* zi, int row) {
  return s == 0 ? -zi[row] : (s == 1 ? -zr[row] : (s == 2 ? zr[row] : zi[row]));
}

template <typename T>
struct VitParams {
  const void* z;
  uint8_t* dec_bits;
  int8_t* dec_syms;
  const uint32_t* tx_packed;
  const int8_t* tx_syms;
  const int32_t* point;
  unsigned long long* counters;
  void* state;  // T[B][1 + 4 + 8*length]: call parity, carried metrics, increment history (streaming)
  int64_t B, n_calls, ld_z, ld_words, S;
  int length;
  // n_in = 2: `z` holds the bank's two raw filter outputs y_f per call and the detector forms
  // z_r = sum_f comb[r][f] y_f itself (same fused operations as the bank's combine stage)
  int n_in;
  T comb_re[6], comb_im[6];
};

// one ACS stage over column parity C: prev[4] -> next[4], sel bit per state
template <typename T, int C>
__device__ __forceinline__ void acs_stage(const T prev[4], const T* zr, const T* zi, T next[4], uint32_t& sel) {
  sel = 0;
#pragma unroll
  for (int st = 0; st < 4; ++st) {
    const int s1 = tr_pred(C, st, 0), s2 = tr_pred(C, st, 1);
    const int nb = tr_newbit(C, st);
    const int r1 = (tr_out(C, s1, nb) + 2) / 2, r2 = (tr_out(C, s2, nb) + 2) / 2;
    const T m1 = prev[s1] + branch_inc<T>(s1, zr, zi, r1);
    const T m2 = prev[s2] + branch_inc<T>(s2, zr, zi, r2);
    const bool second = m2 < m1;
    next[st] = second ? m2 : m1;
    sel |= (second ? 1u : 0u) << st;
  }
}

// length = 2 detector call with compile-time column parity PAR = i % 2.
// Returns the decision of this call: bit in bit 0, symbol code (0,1,2 -> -2,0,+2) in bits 1-2.
template <typename T, int PAR, bool DIFF>
__device__ __forceinline__ uint32_t detector_call_l2(T M[4], T zpr[3], T zpi[3], const T zr[3], const T zi[3]) {
  // metrics[:, -1] <- metrics[:, 0] - min (algorithm.py:65)
  const T mn = fmin(fmin(M[0], M[1]), fmin(M[2], M[3]));
  T Mn[4] = {M[0] - mn, M[1] - mn, M[2] - mn, M[3] - mn};
  T M0[4], M1[4];
  uint32_t selA, selB;
  acs_stage<T, 1 - PAR>(Mn, zpr, zpi, M0, selA);  // stage j=0: column (i-1)%2, previous increments
  acs_stage<T, PAR>(M0, zr, zi, M1, selB);        // stage j=1: column i%2, current increments
  // np.argmin: lowest index on ties
  int s1 = 0;
  T best = M1[0];
  if (M1[1] < best) { best = M1[1]; s1 = 1; }
  if (M1[2] < best) { best = M1[2]; s1 = 2; }
  if (M1[3] < best) { best = M1[3]; s1 = 3; }
  // traceback: s0 = start of the surviving stage-B branch into s1
  const int kB = (selB >> s1) & 1;
  const int s0 = PAR == 0 ? ((kB << 1) | (s1 & 1)) : ((s1 & 2) | kB);
  // decision = surviving stage-A branch (column 1-PAR) into s0
  const int kA = (selA >> s0) & 1;
  constexpr int CA = 1 - PAR;
  const int pa = CA == 0 ? ((kA << 1) | (s0 & 1)) : ((s0 & 2) | kA);
  const int nb = CA == 0 ? (s0 >> 1) : (s0 & 1);
  const int old = CA == 0 ? (pa >> 1) : (pa & 1);
  const int other = CA == 0 ? (pa & 1) : (pa >> 1);
  const int bit = DIFF ? (nb ^ old) : nb;
  const int sgn = CA == 0 ? -1 : 1;
  const int sym = 2 * sgn * (2 * other - 1) * (nb - old);
#pragma unroll
  for (int s = 0; s < 4; ++s) M[s] = M0[s];
#pragma unroll
  for (int r = 0; r < 3; ++r) { zpr[r] = zr[r]; zpi[r] = zi[r]; }
  return (uint32_t)bit | ((uint32_t)((sym + 2) >> 1) << 1);
}

// error-count bookkeeping shared by both kernels
struct ErrCount {
  uint32_t decw = 0;
  unsigned long long bit_err = 0, sym_err = 0;
  __device__ __forceinline__ void add(int64_t j, int64_t n_cmp, uint32_t bit, int sym,
                                      const uint32_t* txw, const int8_t* txs) {
    if (j < 0 || j >= n_cmp) return;
    decw |= bit << (j & 31);
    if (txs) sym_err += (sym != (int)txs[j]) ? 1u : 0u;
    if ((j & 31) == 31 || j == n_cmp - 1) {
      if (txw) {
        const int nv = (int)(j & 31) + 1;
        const uint32_t mask = nv == 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u);
        bit_err += __popc((decw ^ txw[j >> 5]) & mask);
      }
      decw = 0;
    }
  }
};

__device__ __forceinline__ void flush_counters(const ErrCount& ec, int64_t n_cmp, bool active, int32_t pt,
                                               unsigned long long* counters) {
  // one atomic triple per warp when the whole warp reports to the same Eb/N0 point
  const uint32_t full = 0xFFFFFFFFu;
  const int32_t pt0 = __shfl_sync(full, pt, 0);
  const bool uniform = __all_sync(full, (!active) || pt == pt0);
  unsigned long long be = active ? ec.bit_err : 0ull, se = active ? ec.sym_err : 0ull,
                     nc = active ? (unsigned long long)n_cmp : 0ull;
  if (uniform && __any_sync(full, active)) {
    // lane 0 may be inactive only if the whole tail is; pick the first active lane's point
    const uint32_t act = __ballot_sync(full, active);
    const int leader = __ffs(act) - 1;
    const int32_t ptl = __shfl_sync(full, pt, leader);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      be += __shfl_xor_sync(full, be, s);
      se += __shfl_xor_sync(full, se, s);
      nc += __shfl_xor_sync(full, nc, s);
    }
    if ((int)(threadIdx.x & 31) == leader) {
      atomicAdd(&counters[3 * (int64_t)ptl + 0], be);
      atomicAdd(&counters[3 * (int64_t)ptl + 1], se);
      atomicAdd(&counters[3 * (int64_t)ptl + 2], nc);
    }
  } else if (active) {
    atomicAdd(&counters[3 * (int64_t)pt + 0], be);
    atomicAdd(&counters[3 * (int64_t)pt + 1], se);
    atomicAdd(&counters[3 * (int64_t)pt + 2], nc);
  }
}

// stage one chunk of calls of every frame of the block: global [frame][call][3] -> smem rows
template <typename C, int kVitChunk>
__device__ __forceinline__ void stage_chunk(const C* __restrict__ z, int64_t frame0, int64_t B, int64_t ld_z,
                                            int64_t c0, int64_t n_calls, C* tile, int pitch) {
  constexpr int ROW = kVitChunk * 3;
  for (int e = threadIdx.x; e < kVitThreads * ROW; e += kVitThreads) {
    const int f = e / ROW, k = e - f * ROW;
    const int64_t fr = frame0 + f;
    const int64_t call = c0 + k / 3;
    if (fr < B && call < n_calls) tile[f * pitch + k] = z[(fr * ld_z + c0) * 3 + k];
  }
}

constexpr int kL2Threads = 64;  // two warps; each warp owns its 32 frames and its own staging buffers

// length-2 detector, the throughput path.  Every WARP is autonomous: it stages the next
// chunk of its 32 frames' rows with cp.async (coalesced 192-byte runs, double buffered)
// while it runs add-compare-select on the current chunk; there is no block barrier.
// FULL = decisions and/or symbol errors requested (API path); otherwise only packed bit
// errors are counted: decisions are collected 32 calls at a time and compared with the
// transmitted bits by funnel-shift (the 2-call decision delay) + XOR + popcount.
template <typename T, bool DIFF, bool FULL, int NIN>
__global__ void __launch_bounds__(kL2Threads)
viterbi_soqpsk_l2_kernel(const __grid_constant__ VitParams<T> p) {
  // Everything is staged and read in 16-byte UNITS (two complex64 or one complex128): a chunk of
  // CH calls is 12 units per frame for both types, rows sit 13 units apart (odd: the quarter-warp
  // phases of LDS.128 are conflict free), one cp.async moves 16 bytes (half the LDGSTS count of an
  // 8-byte copy -- the MIO queue, not HBM, was the limiter) and a float call pair is 3 LDS.128.
  // NIN = 3: rows hold the three combined outputs z_r (24 bytes per call); NIN = 2: the two raw
  // filter outputs (16 bytes per call) and the 3x2 combine runs here, off the metric recursion.
  constexpr int CH = VitChunk<T>::n;       // calls per chunk (divides 32)
  constexpr int ROWU = CH * NIN * 2 * (int)sizeof(T) / 16;  // units per frame per chunk: 12 or 8
  constexpr int PITCHU = ROWU + 1;
  constexpr int NPAT = ROWU == 12 ? 3 : 1;            // unit e = lane + 32 r: the (row, column) pattern repeats
  constexpr int ROWS_PER_STEP = 32 * NPAT / ROWU;     // every NPAT r's, this many rows further on
  constexpr int STEPS = 32 / ROWS_PER_STEP;
  static_assert(ROWU == 12 || ROWU == 8, "chunk must be 12 or 8 units");
  constexpr int NBUF = 3;  // chunk c + 2 is requested while chunk c is consumed (one wave of ~3.5 warps per scheduler
                           // cannot hide a DRAM round trip behind one chunk of arithmetic)
  __align__(16) __shared__ uint4 tile[NBUF][kL2Threads * PITCHU];

  const int lane = threadIdx.x & 31, wrow0 = threadIdx.x & ~31;
  const int64_t wframe0 = (int64_t)blockIdx.x * kL2Threads + wrow0;
  if (wframe0 >= p.B) return;  // whole warp
  const int64_t frame = wframe0 + lane;
  const bool active = frame < p.B;
  const int n_calls = (int)p.n_calls;
  const int n_cmp = max(0, (int)min(p.S, (int64_t)n_calls - 2));
  const uint32_t* txw = (active && p.tx_packed) ? p.tx_packed + frame * p.ld_words : nullptr;
  const int8_t* txs = (FULL && active && p.tx_syms) ? p.tx_syms + frame * p.S : nullptr;
  uint8_t* ob = (FULL && active && p.dec_bits) ? p.dec_bits + frame * p.n_calls : nullptr;
  int8_t* os = (FULL && active && p.dec_syms) ? p.dec_syms + frame * p.n_calls : nullptr;

  // staging pattern of this lane: three (row, unit) pairs, each repeated STEPS times ROWS_PER_STEP rows apart
  const int64_t row_units = p.ld_z * (int64_t)(NIN * 2 * sizeof(T)) / 16;  // row bytes / 16 (a whole number here)
  const uint4* src[NPAT];
  int dst[NPAT], ucall[NPAT];
#pragma unroll
  for (int r = 0; r < NPAT; ++r) {
    const int e = lane + 32 * r, f = e / ROWU, k = e - f * ROWU;
    src[r] = reinterpret_cast<const uint4*>(p.z) + (wframe0 + f) * row_units + k;
    dst[r] = (wrow0 + f) * PITCHU + k;
    ucall[r] = sizeof(T) == 4 ? (2 * k) / NIN : k / NIN;  // call of the unit's first element
  }
  const int64_t src_step = (int64_t)ROWS_PER_STEP * row_units;
  const int64_t last_row = p.B - 1 - wframe0;  // rows beyond it are clamped to it (valid memory, never used)

  auto stage = [&](int buf, int c0) {
    const bool whole = c0 + CH <= n_calls;
#pragma unroll
    for (int r = 0; r < NPAT; ++r) {
      const int f0 = (lane + 32 * r) / ROWU;
#pragma unroll
      for (int m = 0; m < STEPS; ++m) {
        const int64_t rowi = f0 + m * ROWS_PER_STEP;
        const int64_t back = rowi > last_row ? (rowi - last_row) * row_units : 0;
        if (whole || c0 + ucall[r] < n_calls)
          cp_async_cg16(&tile[buf][dst[r] + m * ROWS_PER_STEP * PITCHU], src[r] + m * src_step - back + (int64_t)(c0 / CH) * ROWU);
      }
    }
    cp_async_commit();
  };

  T M[4] = {0, 0, 0, 0};
  T zpr[3] = {0, 0, 0}, zpi[3] = {0, 0, 0};
  uint32_t decw = 0, w_prev = 0;
  unsigned long long bit_err = 0, sym_err = 0;
  const int n_words = (n_cmp + 31) >> 5;
  uint32_t w_next = (txw != nullptr && n_words > 0) ? txw[0] : 0u;  // loaded one group ahead of its use

  // compare the 32 decisions of calls [g, g+32) with transmitted bits [g-2, g+30)
  auto count_group = [&](int g) {
    if (txw == nullptr) return;
    const int lo = max(g, 2), hi = min(g + 32, n_cmp + 2);  // valid call range
    if (hi > lo) {
      const int wi = g >> 5;
      const uint32_t w_cur = w_next;
      w_next = (wi + 1 < n_words) ? txw[wi + 1] : 0u;
      const uint32_t tx = (w_cur << 2) | (w_prev >> 30);
      uint32_t mask = 0xFFFFFFFFu;
      if (lo > g) mask &= 0xFFFFFFFFu << (lo - g);
      if (hi < g + 32) mask &= 0xFFFFFFFFu >> (g + 32 - hi);
      bit_err += __popc((decw ^ tx) & mask);
      w_prev = w_cur;
    }
  };

  const int n_chunks = (n_calls + CH - 1) / CH;
  // every iteration commits exactly one group (empty past the end), so the wait count is a constant
#pragma unroll
  for (int k = 0; k < NBUF - 1; ++k) {
    if (k < n_chunks) stage(k, k * CH);
    else cp_async_commit();
  }
  int buf = 0, buf_next = NBUF - 1;  // c % NBUF and (c + NBUF - 1) % NBUF
  for (int c = 0; c < n_chunks; ++c) {
    const int c0 = c * CH;
    if (c + NBUF - 1 < n_chunks) stage(buf_next, c0 + (NBUF - 1) * CH);
    else cp_async_commit();
    cp_async_wait<NBUF - 1>();
    __syncwarp();
    const uint4* row = &tile[buf][threadIdx.x * PITCHU];
    buf_next = buf;
    buf = buf + 1 == NBUF ? 0 : buf + 1;
    auto decide = [&](int i, const T* zr, const T* zi, int par) {
      const uint32_t d = par ? detector_call_l2<T, 1, DIFF>(M, zpr, zpi, zr, zi)
                             : detector_call_l2<T, 0, DIFF>(M, zpr, zpi, zr, zi);
      decw |= (d & 1u) << (i & 31);
      if (FULL) {
        const int sym = 2 * (int)(d >> 1) - 2;
        if (ob) ob[i] = (uint8_t)(d & 1u);
        if (os) os[i] = (int8_t)sym;
        if (txs && i >= 2 && i - 2 < n_cmp) sym_err += (sym != (int)txs[i - 2]) ? 1u : 0u;
      }
    };
    const bool whole = c0 + CH <= n_calls;
    if (NIN == 2) {
#pragma unroll
      for (int u = 0; u < CH; ++u) {
        if (whole || c0 + u < n_calls) {
          T yr[2], yi[2];
          if (sizeof(T) == 4) {
            const float4 q = *reinterpret_cast<const float4*>(row + u);
            yr[0] = (T)q.x; yi[0] = (T)q.y; yr[1] = (T)q.z; yi[1] = (T)q.w;
          } else {
            const double2 a = *reinterpret_cast<const double2*>(row + 2 * u), b = *reinterpret_cast<const double2*>(row + 2 * u + 1);
            yr[0] = (T)a.x; yi[0] = (T)a.y; yr[1] = (T)b.x; yi[1] = (T)b.y;
          }
          T zr[3], zi[3];
#pragma unroll
          for (int r = 0; r < 3; ++r) {  // operation for operation the bank's combine stage (mfbank.cu)
            T ar = 0, ai = 0;
#pragma unroll
            for (int f = 0; f < 2; ++f) {
              const T cr = p.comb_re[r * 2 + f], ci = p.comb_im[r * 2 + f];
              ar = fma(cr, yr[f], ar); ar = fma(-ci, yi[f], ar);
              ai = fma(cr, yi[f], ai); ai = fma(ci, yr[f], ai);
            }
            zr[r] = ar; zi[r] = ai;
          }
          decide(c0 + u, zr, zi, u & 1);
        }
      }
    } else if (sizeof(T) == 4) {
#pragma unroll
      for (int u = 0; u < CH; u += 2) {  // a call pair is 3 units: (z0 z1) (z2 z0') (z1' z2')
        if (whole || c0 + u < n_calls) {
          const float4 q0 = *reinterpret_cast<const float4*>(row + 3 * (u / 2));
          const float4 q1 = *reinterpret_cast<const float4*>(row + 3 * (u / 2) + 1);
          const float4 q2 = *reinterpret_cast<const float4*>(row + 3 * (u / 2) + 2);
          const T zr0[3] = {(T)q0.x, (T)q0.z, (T)q1.x}, zi0[3] = {(T)q0.y, (T)q0.w, (T)q1.y};
          const T zr1[3] = {(T)q1.z, (T)q2.x, (T)q2.z}, zi1[3] = {(T)q1.w, (T)q2.y, (T)q2.w};
          decide(c0 + u, zr0, zi0, 0);
          if (whole || c0 + u + 1 < n_calls) decide(c0 + u + 1, zr1, zi1, 1);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < CH; ++u) {
        if (whole || c0 + u < n_calls) {
          T zr[3], zi[3];
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const double2 v = *reinterpret_cast<const double2*>(row + 3 * u + r);
            zr[r] = (T)v.x; zi[r] = (T)v.y;
          }
          decide(c0 + u, zr, zi, u & 1);
        }
      }
    }
    if (((c0 + CH) & 31) == 0) { count_group(c0 + CH - 32); decw = 0; }
    __syncwarp();  // all lanes are done with this buffer before chunk c + NBUF is staged into it
  }
  if ((n_chunks * CH) & 31) count_group((n_chunks * CH) & ~31);  // group left open by the last chunk

  if (p.counters) {
    ErrCount ec;
    ec.bit_err = bit_err;
    ec.sym_err = sym_err;
    flush_counters(ec, n_cmp, active, (active && p.point) ? p.point[frame] : 0, p.counters);
  }
}

// ---------------------------------------------------------------------------
// Any window length (API parity, algorithm.py:44-101 verbatim in structure):
// keeps `length` columns of 8 branch increments BY BRANCH INDEX, so odd lengths
// reproduce the reference's column-parity mismatch as well.
// ---------------------------------------------------------------------------
template <typename T, bool DIFF>
__global__ void __launch_bounds__(kVitThreads)
viterbi_soqpsk_any_kernel(const __grid_constant__ VitParams<T> p) {
  using C = typename std::conditional<sizeof(T) == 4, float2, double2>::type;
  constexpr int kVitChunk = VitChunk<T>::n;
  constexpr int PITCH = kVitChunk * 3 + 1;
  __shared__ C tile[kVitThreads * PITCH];

  const int L = p.length;
  const int64_t frame0 = (int64_t)blockIdx.x * kVitThreads;
  const int64_t frame = frame0 + threadIdx.x;
  const bool active = frame < p.B;
  const C* z = reinterpret_cast<const C*>(p.z);
  const int64_t n_calls = p.n_calls;
  const int64_t n_cmp = max((int64_t)0, min(p.S, n_calls - L));
  const uint32_t* txw = (active && p.tx_packed) ? p.tx_packed + frame * p.ld_words : nullptr;
  const int8_t* txs = (active && p.tx_syms) ? p.tx_syms + frame * p.S : nullptr;
  uint8_t* ob = (active && p.dec_bits) ? p.dec_bits + frame * n_calls : nullptr;
  int8_t* os = (active && p.dec_syms) ? p.dec_syms + frame * n_calls : nullptr;

  T inc[kVitMaxLen][8];
  for (int j = 0; j < kVitMaxLen; ++j)
    for (int b = 0; b < 8; ++b) inc[j][b] = 0;
  T carry[4] = {0, 0, 0, 0};
  int64_t ibase = 0;
  T* st_io = (active && p.state) ? reinterpret_cast<T*>(p.state) + frame * (5 + 8 * L) : nullptr;
  if (st_io) {
    ibase = (int64_t)st_io[0];
    for (int s = 0; s < 4; ++s) carry[s] = st_io[1 + s];
    for (int j = 0; j < L; ++j)
      for (int b = 0; b < 8; ++b) inc[j][b] = st_io[5 + 8 * j + b];
  }
  ErrCount ec;
  const C* row = tile + threadIdx.x * PITCH;

  for (int64_t c0 = 0; c0 < n_calls; c0 += kVitChunk) {
    __syncthreads();
    stage_chunk<C, kVitChunk>(z, frame0, p.B, p.ld_z, c0, n_calls, tile, PITCH);
    __syncthreads();
    if (!active) continue;
    for (int u = 0; u < kVitChunk && c0 + u < n_calls; ++u) {
      const int64_t il = c0 + u;       // call index inside this launch
      const int64_t i = il + ibase;    // detector's running call counter (parity is what matters)
      T zr[3], zi[3];
      for (int r = 0; r < 3; ++r) { const C v = row[u * 3 + r]; zr[r] = v.x; zi[r] = v.y; }
      // roll the history left, append this call's increments (column i % 2)
      for (int j = 0; j + 1 < L; ++j)
        for (int b = 0; b < 8; ++b) inc[j][b] = inc[j + 1][b];
      const int cnow = (int)(i & 1);
      for (int b = 0; b < 8; ++b) {
        const int s = b >> 1, nb = b & 1;
        inc[L - 1][b] = branch_inc<T>(s, zr, zi, (tr_out(cnow, s, nb) + 2) / 2);
      }
      const T mn = fmin(fmin(carry[0], carry[1]), fmin(carry[2], carry[3]));
      T prev[4] = {carry[0] - mn, carry[1] - mn, carry[2] - mn, carry[3] - mn};
      uint8_t src[kVitMaxLen];  // per stage: 2 bits per state = start state of the survivor
      for (int j = 0; j < L; ++j) {
        const int c = (int)((i + j + 1) & 1);  // (i + j - 1) mod 2
        T met[4];
        uint32_t sw = 0;
        for (int st = 0; st < 4; ++st) {
          const int s1 = tr_pred(c, st, 0), s2 = tr_pred(c, st, 1), nb = tr_newbit(c, st);
          const T m1 = prev[s1] + inc[j][2 * s1 + nb];
          const T m2 = prev[s2] + inc[j][2 * s2 + nb];
          const bool second = m2 < m1;
          met[st] = second ? m2 : m1;
          sw |= (uint32_t)(second ? s2 : s1) << (2 * st);
        }
        src[j] = (uint8_t)sw;
        if (j == 0)
          for (int st = 0; st < 4; ++st) carry[st] = met[st];
        for (int st = 0; st < 4; ++st) prev[st] = met[st];
      }
      // traceback from argmin of the last stage (lowest index on ties) down to stage 0
      int st = 0;
      {
        T best = prev[0];
        for (int s = 1; s < 4; ++s)
          if (prev[s] < best) { best = prev[s]; st = s; }
      }
      int pstart = 0;
      for (int j = L - 1; j >= 0; --j) {
        pstart = (int)((src[j] >> (2 * st)) & 3u);
        if (j > 0) st = pstart;
      }
      const int c0p = (int)((i + 1) & 1);  // column of stage 0
      const int nb = tr_newbit(c0p, st);
      const int bit = tr_inp(c0p, pstart, nb, DIFF);
      const int sym = tr_out(c0p, pstart, nb);
      if (ob) ob[il] = (uint8_t)bit;
      if (os) os[il] = (int8_t)sym;
      ec.add(il - L, n_cmp, (uint32_t)bit, sym, txw, txs);
    }
  }
  if (st_io) {
    st_io[0] = (T)((ibase + n_calls) & 1);
    for (int s = 0; s < 4; ++s) st_io[1 + s] = carry[s];
    for (int j = 0; j < L; ++j)
      for (int b = 0; b < 8; ++b) st_io[5 + 8 * j + b] = inc[j][b];
  }
  if (p.counters) flush_counters(ec, n_cmp, active, (active && p.point) ? p.point[frame] : 0, p.counters);
}

template <typename T>
static int launch_viterbi(const VitParams<T>& p, int diff, cudaStream_t stream) {
  const unsigned grid = (unsigned)((p.B + kVitThreads - 1) / kVitThreads);
  // the length-2 kernel stages 16-byte units: rows must start on 16-byte boundaries
  const bool rows16 = (reinterpret_cast<uintptr_t>(p.z) & 15) == 0 && (sizeof(T) == 8 || p.n_in == 2 || (p.ld_z & 1) == 0);
  if (p.length == 2 && !p.state && rows16) {
    const unsigned g2 = (unsigned)((p.B + kL2Threads - 1) / kL2Threads);
    const bool full = p.dec_bits || p.dec_syms || p.tx_syms;
    if (p.n_in == 2) {
      if (diff) viterbi_soqpsk_l2_kernel<T, true, false, 2><<<g2, kL2Threads, 0, stream>>>(p);
      else viterbi_soqpsk_l2_kernel<T, false, false, 2><<<g2, kL2Threads, 0, stream>>>(p);
    }
    else if (diff && full) viterbi_soqpsk_l2_kernel<T, true, true, 3><<<g2, kL2Threads, 0, stream>>>(p);
    else if (diff) viterbi_soqpsk_l2_kernel<T, true, false, 3><<<g2, kL2Threads, 0, stream>>>(p);
    else if (full) viterbi_soqpsk_l2_kernel<T, false, true, 3><<<g2, kL2Threads, 0, stream>>>(p);
    else viterbi_soqpsk_l2_kernel<T, false, false, 3><<<g2, kL2Threads, 0, stream>>>(p);
  } else if (p.n_in == 2) {
    return fail(WF_ERR_ARG, "combined-input detection needs length 2, no streaming state and 16-byte aligned rows");
  } else {
    if (diff) viterbi_soqpsk_any_kernel<T, true><<<grid, kVitThreads, 0, stream>>>(p);
    else viterbi_soqpsk_any_kernel<T, false><<<grid, kVitThreads, 0, stream>>>(p);
  }
  WF_LAUNCH_CHECK("viterbi_soqpsk_kernel");
  return WF_OK;
}

}  // namespace wf

extern "C" {

int wf_viterbi_soqpsk(const void* d_z, int dtype, int64_t B, int64_t n_calls, int64_t ld_z, int diff, int length,
                      uint8_t* d_dec_bits, int8_t* d_dec_syms, const uint32_t* d_tx_packed, int64_t ld_words,
                      const int8_t* d_tx_syms, int64_t S, const int32_t* d_point, int64_t* d_counters, void* d_state,
                      void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_calls >= 0, "negative shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  WF_REQUIRE(length >= 1 && length <= kVitMaxLen, "detector length %d outside 1..%d", length, kVitMaxLen);
  WF_REQUIRE(ld_z >= n_calls, "ld_z smaller than n_calls");
  WF_REQUIRE(!d_tx_packed || ld_words >= (S + 31) / 32, "ld_words too small for S");
  WF_REQUIRE(!(d_tx_packed || d_tx_syms) || d_counters, "error counting needs d_counters");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  WF_REQUIRE(d_z || n_calls == 0, "d_z is NULL");
  if (dtype == WF_F32) {
    VitParams<float> p;
    memset(&p, 0, sizeof(p));
    p.z = d_z; p.dec_bits = d_dec_bits; p.dec_syms = d_dec_syms; p.tx_packed = d_tx_packed; p.tx_syms = d_tx_syms;
    p.point = d_point; p.counters = reinterpret_cast<unsigned long long*>(d_counters);
    p.B = B; p.n_calls = n_calls; p.ld_z = ld_z; p.ld_words = ld_words; p.S = S; p.length = length; p.state = d_state;
    p.n_in = 3;
    return launch_viterbi<float>(p, diff, as_stream(stream));
  }
  VitParams<double> p;
  memset(&p, 0, sizeof(p));
  p.z = d_z; p.dec_bits = d_dec_bits; p.dec_syms = d_dec_syms; p.tx_packed = d_tx_packed; p.tx_syms = d_tx_syms;
  p.point = d_point; p.counters = reinterpret_cast<unsigned long long*>(d_counters);
  p.B = B; p.n_calls = n_calls; p.ld_z = ld_z; p.ld_words = ld_words; p.S = S; p.length = length; p.state = d_state;
  p.n_in = 3;
  return launch_viterbi<double>(p, diff, as_stream(stream));
}

// Counter-only detection straight from the bank's RAW filter outputs (n_in = 2, the PAM bank):
// y is [B][ld_y][2] and z_r = sum_f comb[r][f] y_f (r = 0..2) is formed inside the detector with the
// bank's own sequence of fused operations, so the counters equal those of
// wf_mf_bank(.., combine, ..) followed by wf_viterbi_soqpsk(..) bit for bit, while the array between
// the two kernels is 16 instead of 24 bytes per call.  comb_* are HOST float64[3][2].
int wf_viterbi_soqpsk_combined(const void* d_y, int dtype, int64_t B, int64_t n_calls, int64_t ld_y, int n_in,
                               const double* comb_re, const double* comb_im, int diff, const uint32_t* d_tx_packed,
                               int64_t ld_words, int64_t S, const int32_t* d_point, int64_t* d_counters, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && n_calls >= 0, "negative shape");
  WF_REQUIRE(dtype == WF_F32 || dtype == WF_F64, "dtype must be WF_F32 or WF_F64");
  WF_REQUIRE(n_in == 2 && comb_re, "n_in must be 2 with a 3x2 combine matrix");
  WF_REQUIRE(ld_y >= n_calls, "ld_y smaller than n_calls");
  WF_REQUIRE(!d_tx_packed || ld_words >= (S + 31) / 32, "ld_words too small for S");
  WF_REQUIRE(!d_tx_packed || d_counters, "error counting needs d_counters");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  WF_REQUIRE(d_y || n_calls == 0, "d_y is NULL");
  if (dtype == WF_F32) {
    VitParams<float> p;
    memset(&p, 0, sizeof(p));
    p.z = d_y; p.tx_packed = d_tx_packed; p.point = d_point; p.counters = reinterpret_cast<unsigned long long*>(d_counters);
    p.B = B; p.n_calls = n_calls; p.ld_z = ld_y; p.ld_words = ld_words; p.S = S; p.length = 2; p.n_in = 2;
    for (int i = 0; i < 6; ++i) { p.comb_re[i] = (float)comb_re[i]; p.comb_im[i] = comb_im ? (float)comb_im[i] : 0.f; }
    return launch_viterbi<float>(p, diff, as_stream(stream));
  }
  VitParams<double> p;
  memset(&p, 0, sizeof(p));
  p.z = d_y; p.tx_packed = d_tx_packed; p.point = d_point; p.counters = reinterpret_cast<unsigned long long*>(d_counters);
  p.B = B; p.n_calls = n_calls; p.ld_z = ld_y; p.ld_words = ld_words; p.S = S; p.length = 2; p.n_in = 2;
  for (int i = 0; i < 6; ++i) { p.comb_re[i] = comb_re[i]; p.comb_im[i] = comb_im ? comb_im[i] : 0.0; }
  return launch_viterbi<double>(p, diff, as_stream(stream));
}

}  // extern "C"
