// prbs.cu -- A1/A3/A4: Galois-LFSR frames with GF(2) jump-ahead, fused with the
// closed-form bit->symbol encoders.  One warp per frame; lane l owns 32-bit
// words l, l+32, ... of the frame so that every warp store covers 1 KiB of
// contiguous symbols.  Differential encoding is a prefix-XOR on the even/odd
// rails (in-word shifts + a 5-step warp scan); the SOQPSK ternary symbol is a
// 3-tap stencil on the encoded word.  Integer work only: bit-exact.
//
// Reference behaviour restated: waveforms/glfsr/glfsr.py:15-19 (step),
// glfsr/pn.py:93-107 (all-ones start), cpm/trellis/encoder.py:33-46 (FSM walk),
// cpm/trellis/model.py:205-293 (tables), cpm/{soqpsk,pcmfm,multih}/precoder.py.
#include "common.cuh"
#include "lfsr_host.h"
#include <string.h>

namespace wf {

constexpr int kWarpsPerBlock = 4;
constexpr int kWordsPerPass = 32;          // one word per lane per pass
constexpr int kStrideSteps = 32 * 31;      // after emitting 32 bits, skip the other 31 lanes' words

struct PrbsParams {
  uint64_t mask, init_state, period, first_frame;
  const uint64_t* offsets;
  const uint8_t* bits_in;
  int64_t B, N, ld_words;
  int32_t* enc_state;
  uint8_t* bits;
  uint32_t* packed;
  int8_t* syms;
  int degree, nj, encoder;
  uint64_t stride_col[64];  // columns of A^(31*32)
};

template <int MAXW>
struct JumpTable { uint64_t pw[MAXW]; };  // [nj][degree] columns of A^(2^j)

// ---- SWAR helpers: 4 bits -> 4 bytes (bit k -> byte k, value 0/1) ----------
__device__ __forceinline__ uint32_t spread4(uint32_t nib) {
  return ((nib & 0xFu) * 0x00204081u) & 0x01010101u;
}

__device__ __forceinline__ void store_bytes32(int8_t* dst, const uint32_t v[8], int nvalid, bool vec_ok) {
  if (nvalid >= 32 && vec_ok) {
    uint4* q = reinterpret_cast<uint4*>(dst);
    q[0] = make_uint4(v[0], v[1], v[2], v[3]);
    q[1] = make_uint4(v[4], v[5], v[6], v[7]);
  } else {
    for (int i = 0; i < nvalid && i < 32; ++i) dst[i] = (int8_t)((v[i >> 2] >> (8 * (i & 3))) & 0xFF);
  }
}

__device__ __forceinline__ void store_bytes16(int8_t* dst, const uint32_t v[4], int nvalid, bool vec_ok) {
  if (nvalid >= 16 && vec_ok) {
    *reinterpret_cast<uint4*>(dst) = make_uint4(v[0], v[1], v[2], v[3]);
  } else {
    for (int i = 0; i < nvalid && i < 16; ++i) dst[i] = (int8_t)((v[i >> 2] >> (8 * (i & 3))) & 0xFF);
  }
}

template <typename ST>
__device__ __forceinline__ uint32_t lfsr_emit32(ST& s, ST mask) {
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    uint32_t bit = (uint32_t)s & 1u;
    s = (s >> 1) ^ (((ST)0 - (ST)bit) & mask);
    out |= bit << i;
  }
  return out;
}

template <typename ST>
__device__ __forceinline__ ST gf2_apply_cols(const uint64_t* __restrict__ col, int degree, ST v) {
  ST r = 0;
  for (int i = 0; i < degree; ++i) r ^= ((ST)0 - (ST)((v >> i) & 1)) & (ST)col[i];
  return r;
}

__device__ __forceinline__ uint64_t mulmod_u64(uint64_t a, uint64_t b, uint64_t m) {
  return (uint64_t)(((unsigned __int128)a * b) % m);
}

// SRC: 0 = LFSR, 1 = caller bits (uint8).  ST: LFSR state word.
template <int SRC, typename ST, int MAXW>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
encode_frames_kernel(const __grid_constant__ PrbsParams p, const __grid_constant__ JumpTable<MAXW> jt) {
  const int lane = threadIdx.x & 31;
  const int64_t frame = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (frame >= p.B) return;  // whole warp leaves together

  const int64_t N = p.N;
  const int64_t nwords = (N + 31) >> 5;
  const int64_t npass = (nwords + kWordsPerPass - 1) / kWordsPerPass;
  const int enc = p.encoder;
  const bool soqpsk = (enc == WF_ENC_SOQPSK_DIFF || enc == WF_ENC_SOQPSK || enc == WF_ENC_SOQPSK_TERNARY);
  const int64_t S = (enc == WF_ENC_QUATERNARY) ? (N >> 1) : N;

  // ---- LFSR start state of this lane's first word --------------------------
  ST state = 0;
  if (SRC == 0) {
    uint64_t k;
    if (p.offsets) {
      k = p.offsets[frame];
      if (p.period) k %= p.period;
    } else if (p.period) {
      k = mulmod_u64((p.first_frame + (uint64_t)frame) % p.period, (uint64_t)N % p.period, p.period);
    } else {
      k = (p.first_frame + (uint64_t)frame) * (uint64_t)N;
    }
    k += 32ull * (uint64_t)lane;
    if (p.period) k %= p.period;
    state = (ST)p.init_state;
    for (int j = 0; j < p.nj; ++j) {
      ST nxt = gf2_apply_cols<ST>(&jt.pw[j * p.degree], p.degree, state);
      state = ((k >> j) & 1) ? nxt : state;
    }
  }

  // ---- encoder carry-in -----------------------------------------------------
  int32_t i0 = 0, st0 = 0;
  if (p.enc_state) {
    i0 = p.enc_state[2 * frame];
    st0 = p.enc_state[2 * frame + 1];
  }
  const bool i0_odd = (i0 & 1) != 0;
  // d_{-1}, d_{-2} in local indexing (state = 2*even-rail + odd-rail of GLOBAL parity)
  uint32_t dm1_init = i0_odd ? ((st0 >> 1) & 1) : (st0 & 1);
  uint32_t dm2_init = i0_odd ? (st0 & 1) : ((st0 >> 1) & 1);
  uint32_t carry = dm2_init | (dm1_init << 1);            // bit0: even local rail, bit1: odd local rail
  uint32_t d_last = (dm1_init << 31) | (dm2_init << 30);  // "previous word" seen by word 0
  const uint32_t even_global = i0_odd ? 0xAAAAAAAAu : 0x55555555u;

  const bool sym_vec = ((reinterpret_cast<uintptr_t>(p.syms) & 15) == 0) && ((S & 15) == 0);
  const bool bit_vec = ((reinterpret_cast<uintptr_t>(p.bits) & 15) == 0) && ((N & 15) == 0);
  const bool in_vec = ((reinterpret_cast<uintptr_t>(p.bits_in) & 15) == 0) && ((N & 15) == 0);

  uint32_t last_d = 0, last_dm1 = 0;  // kept by the lane that owns the final word

  for (int64_t pass = 0; pass < npass; ++pass) {
    const int64_t w = pass * kWordsPerPass + lane;
    const int64_t bit0 = w << 5;
    int nb = (int)min((int64_t)32, N - bit0);  // valid bits in this word (<= 0: none)
    if (nb < 0) nb = 0;

    // ---- 1. the raw word ---------------------------------------------------
    uint32_t word = 0;
    if (SRC == 0) {
      word = lfsr_emit32<ST>(state, (ST)p.mask);
      if (pass + 1 < npass) state = gf2_apply_cols<ST>(p.stride_col, p.degree, state);
    } else if (nb > 0) {
      const uint8_t* src = p.bits_in + frame * N + bit0;
      if (nb == 32 && in_vec) {
        const uint4* q = reinterpret_cast<const uint4*>(src);
        uint4 a = q[0], b = q[1];
        uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int g = 0; g < 8; ++g) word |= ((((v[g] & 0x01010101u) * 0x00204081u) >> 21) & 0xFu) << (4 * g);
      } else {
        for (int i = 0; i < nb; ++i) word |= (uint32_t)(src[i] & 1) << i;
      }
    }
    if (nb < 32) word &= (nb == 0) ? 0u : (0xFFFFFFFFu >> (32 - nb));

    if (nb > 0) {
      if (p.packed) p.packed[frame * p.ld_words + w] = word;
      if (p.bits) {
        uint32_t v[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) v[g] = spread4(word >> (4 * g));
        store_bytes32(reinterpret_cast<int8_t*>(p.bits + frame * N + bit0), v, nb, bit_vec);
      }
    }

    // ---- 2. symbols ----------------------------------------------------------
    if (soqpsk) {
      uint32_t d = word;
      if (enc == WF_ENC_SOQPSK_DIFF) {
        // in-word prefix XOR on each rail
        d ^= d << 2; d ^= d << 4; d ^= d << 8; d ^= d << 16;
        uint32_t par = (__popc(word & 0x55555555u) & 1) | ((__popc(word & 0xAAAAAAAAu) & 1) << 1);
        uint32_t inc = par;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
          uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, s);
          if (lane >= s) inc ^= t;
        }
        uint32_t excl = inc ^ par ^ carry;                // rail parities of everything before this word
        carry ^= __shfl_sync(0xFFFFFFFFu, inc, 31);       // running carry for the next pass
        d ^= ((excl & 1) ? 0x55555555u : 0u) ^ ((excl & 2) ? 0xAAAAAAAAu : 0u);
        if (nb < 32) d &= (nb == 0) ? 0u : (0xFFFFFFFFu >> (32 - nb));
      }
      uint32_t dprev = __shfl_up_sync(0xFFFFFFFFu, d, 1);
      if (lane == 0) dprev = d_last;
      d_last = __shfl_sync(0xFFFFFFFFu, d, 31);
      const uint32_t dm1 = (d << 1) | (dprev >> 31);
      const uint32_t dm2 = (d << 2) | (dprev >> 30);
      const uint32_t nz = d ^ dm2;                        // d_i != d_{i-2}
      const uint32_t neg = (even_global ^ dm1 ^ d) & nz;  // see DESIGN.md: sign of 2(-1)^(i+1)(2d_{i-1}-1)(d_i-d_{i-2})
      if (nb > 0 && bit0 + nb == N) { last_d = d; last_dm1 = dm1; }
      if (nb > 0 && p.syms) {
        uint32_t v[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint32_t nzb = spread4(nz >> (4 * g));
          uint32_t ngb = spread4(neg >> (4 * g));
          if (enc == WF_ENC_SOQPSK_TERNARY) {
            uint32_t m = nzb & ~ngb;                      // -(alpha)/2: non-negated alpha -> -1
            v[g] = (nzb ^ (m * 0xFFu)) + m;
          } else {
            v[g] = ((nzb << 1) ^ (ngb * 0xFFu)) + ngb;    // 0, +2 or -2 per byte
          }
        }
        store_bytes32(p.syms + frame * S + bit0, v, nb, sym_vec);
      }
    } else if (enc == WF_ENC_BINARY) {
      if (nb > 0 && bit0 + nb == N) last_d = word;
      if (nb > 0 && p.syms) {
        uint32_t v[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint32_t bb = spread4(word >> (4 * g));
          v[g] = bb | ((bb ^ 0x01010101u) * 0xFFu);      // 1 -> +1, 0 -> -1
        }
        store_bytes32(p.syms + frame * S + bit0, v, nb, sym_vec);
      }
    } else {  // WF_ENC_QUATERNARY: bit pairs (b0 first) -> 2(2 b0 + b1) - 3
      if (nb > 0 && bit0 + nb == N) last_d = word;
      if (nb > 0 && p.syms) {
        uint32_t ev = word & 0x55555555u, od = (word >> 1) & 0x55555555u;
        // compress the 16 even-position bits / odd-position bits to 16 contiguous bits
        auto squeeze = [](uint32_t x) {
          x = (x | (x >> 1)) & 0x33333333u;
          x = (x | (x >> 2)) & 0x0F0F0F0Fu;
          x = (x | (x >> 4)) & 0x00FF00FFu;
          x = (x | (x >> 8)) & 0x0000FFFFu;
          return x;
        };
        uint32_t e16 = squeeze(ev), o16 = squeeze(od);
        uint32_t v[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t x = (spread4(e16 >> (4 * g)) << 2) + (spread4(o16 >> (4 * g)) << 1);  // 4 b0 + 2 b1
          v[g] = ((x | 0x80808080u) - 0x03030303u) ^ 0x80808080u;                       // minus 3 per byte
        }
        store_bytes16(p.syms + frame * S + (bit0 >> 1), v, nb >> 1, sym_vec);
      }
    }
  }

  // ---- 3. encoder state out (encoder.py:44-46) --------------------------------
  if (p.enc_state && N > 0) {
    const int64_t wl = nwords - 1;
    if ((int)(wl & 31) == lane) {
      const int nbl = (int)(N - (wl << 5));
      int32_t st, iout;
      if (soqpsk) {
        uint32_t a = (last_d >> (nbl - 1)) & 1;    // d_{N-1}
        uint32_t b = (last_dm1 >> (nbl - 1)) & 1;  // d_{N-2}
        bool last_even_global = (((int64_t)i0 + N - 1) & 1) == 0;
        st = last_even_global ? (int32_t)(2 * a + b) : (int32_t)(2 * b + a);
        iout = i0 + (int32_t)N;
      } else if (enc == WF_ENC_BINARY) {
        st = (last_d >> (nbl - 1)) & 1;
        iout = i0 + (int32_t)N;
      } else {
        uint32_t b1 = (last_d >> (nbl - 1)) & 1, b0 = (nbl >= 2) ? ((last_d >> (nbl - 2)) & 1) : 0;
        st = (int32_t)(2 * b0 + b1);
        iout = i0 + (int32_t)(N >> 1);
      }
      p.enc_state[2 * frame] = iout;
      p.enc_state[2 * frame + 1] = st;
    }
  }
}

// ---------------------------------------------------------------------------
// Arbitrary Trellis, table driven: one frame per thread (API-parity path, not
// the throughput path).
// ---------------------------------------------------------------------------
struct TrellisParams {
  const uint8_t* bits;
  int8_t* syms;
  int32_t* enc_state;
  int64_t B, N;
  int columns, states, card;
  int8_t next_state[1024];
  int8_t out_symbol[1024];
};

__global__ void trellis_encode_kernel(const __grid_constant__ TrellisParams p) {
  int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= p.B) return;
  const int ninp = 1 << p.card;
  const int64_t S = p.N / p.card;
  int32_t i = 0, st = 0;
  if (p.enc_state) { i = p.enc_state[2 * f]; st = p.enc_state[2 * f + 1]; }
  const uint8_t* b = p.bits + f * p.N;
  int8_t* o = p.syms + f * S;
  for (int64_t n = 0; n < S; ++n) {
    int v = 0;
    for (int x = 0; x < p.card; ++x) v = (v << 1) | (b[n * p.card + x] & 1);
    int col = (int)(((i % p.columns) + p.columns) % p.columns);
    int idx = (col * p.states + st) * ninp + v;
    int ns = p.next_state[idx];
    if (ns >= 0) { o[n] = p.out_symbol[idx]; st = ns; } else { o[n] = 0; }
    ++i;
  }
  if (p.enc_state) { p.enc_state[2 * f] = i; p.enc_state[2 * f + 1] = st; }
}

// ---------------------------------------------------------------------------
template <int SRC>
static int launch_encode(PrbsParams& p, cudaStream_t stream) {
  const unsigned grid = (unsigned)((p.B + kWarpsPerBlock - 1) / kWarpsPerBlock);
  const dim3 block(kWarpsPerBlock * 32);
  if (SRC == 1) {
    JumpTable<1> jt;
    jt.pw[0] = 0;
    encode_frames_kernel<1, uint32_t, 1><<<grid, block, 0, stream>>>(p, jt);
  } else {
    Gf2Mat pw[64];
    gf2_step_powers(p.mask, p.degree, pw);
    Gf2Mat st;
    gf2_step_power(p.mask, p.degree, (uint64_t)kStrideSteps, st);
    memcpy(p.stride_col, st.col, sizeof(p.stride_col));
    if (p.nj * p.degree <= 400) {
      JumpTable<400> jt;
      memset(&jt, 0, sizeof(jt));
      for (int j = 0; j < p.nj; ++j) memcpy(&jt.pw[j * p.degree], pw[j].col, sizeof(uint64_t) * p.degree);
      if (p.degree <= 32) encode_frames_kernel<0, uint32_t, 400><<<grid, block, 0, stream>>>(p, jt);
      else encode_frames_kernel<0, uint64_t, 400><<<grid, block, 0, stream>>>(p, jt);
    } else {
      JumpTable<3840> jt;  // 30 KiB of kernel parameters
      memset(&jt, 0, sizeof(jt));
      for (int j = 0; j < p.nj; ++j) memcpy(&jt.pw[j * p.degree], pw[j].col, sizeof(uint64_t) * p.degree);
      if (p.degree <= 32) encode_frames_kernel<0, uint32_t, 3840><<<grid, block, 0, stream>>>(p, jt);
      else encode_frames_kernel<0, uint64_t, 3840><<<grid, block, 0, stream>>>(p, jt);
    }
  }
  WF_LAUNCH_CHECK("encode_frames_kernel");
  return WF_OK;
}

static int check_encoder(int encoder, int64_t N) {
  WF_REQUIRE(encoder >= WF_ENC_SOQPSK_DIFF && encoder <= WF_ENC_SOQPSK_TERNARY, "unknown encoder id %d", encoder);
  WF_REQUIRE(!(encoder == WF_ENC_QUATERNARY && (N & 1)), "Odd length bit array passed into quaternary mapper.");
  return WF_OK;
}

}  // namespace wf

extern "C" {

int wf_prbs_encode(uint64_t mask, int degree, uint64_t init_state, uint64_t period,
                   const uint64_t* d_bit_offsets, uint64_t first_frame,
                   int64_t B, int64_t N, int encoder, int32_t* d_enc_state,
                   uint8_t* d_bits, uint32_t* d_packed, int64_t ld_words,
                   int8_t* d_syms, void* stream) {
  using namespace wf;
  WF_REQUIRE(degree >= 1 && degree <= 64, "LFSR degree %d outside 1..64", degree);
  WF_REQUIRE(B >= 0 && N >= 0, "negative batch or frame size");
  WF_REQUIRE(degree == 64 || (mask >> degree) == 0, "mask has bits above the LFSR degree");
  WF_REQUIRE(degree == 64 || (init_state >> degree) == 0, "state has bits above the LFSR degree");
  int rc = check_encoder(encoder, N);
  if (rc) return rc;
  WF_REQUIRE(!d_packed || ld_words >= (N + 31) / 32, "ld_words too small for N");
  WF_REQUIRE(B <= (int64_t)0x7FFFFFFF * kWarpsPerBlock, "batch too large for one launch");
  if ((rc = require_device())) return rc;
  if (B == 0 || N == 0) return WF_OK;
  PrbsParams p;
  memset(&p, 0, sizeof(p));
  p.mask = mask; p.init_state = init_state; p.period = period; p.first_frame = first_frame;
  p.offsets = d_bit_offsets; p.B = B; p.N = N; p.ld_words = ld_words;
  p.enc_state = d_enc_state; p.bits = d_bits; p.packed = d_packed; p.syms = d_syms;
  p.degree = degree; p.encoder = encoder;
  int nj = 60;
  if (period) { nj = 0; while (nj < 64 && (period >> nj)) ++nj; }
  if (nj > 60) nj = 60;
  p.nj = nj;
  return launch_encode<0>(p, as_stream(stream));
}

int wf_bits_encode(const uint8_t* d_bits_in, int64_t B, int64_t N, int encoder,
                   int32_t* d_enc_state, uint32_t* d_packed, int64_t ld_words,
                   int8_t* d_syms, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && N >= 0, "negative batch or frame size");
  WF_REQUIRE(d_bits_in || B * N == 0, "d_bits_in is NULL");
  int rc = check_encoder(encoder, N);
  if (rc) return rc;
  WF_REQUIRE(!d_packed || ld_words >= (N + 31) / 32, "ld_words too small for N");
  if ((rc = require_device())) return rc;
  if (B == 0 || N == 0) return WF_OK;
  PrbsParams p;
  memset(&p, 0, sizeof(p));
  p.bits_in = d_bits_in; p.B = B; p.N = N; p.ld_words = ld_words;
  p.enc_state = d_enc_state; p.packed = d_packed; p.syms = d_syms;
  p.degree = 1; p.encoder = encoder;
  return launch_encode<1>(p, as_stream(stream));
}

int wf_trellis_encode(const uint8_t* d_bits_in, int64_t B, int64_t N,
                      const int8_t* next_state, const int8_t* out_symbol,
                      int columns, int states, int bits_per_symbol,
                      int32_t* d_enc_state, int8_t* d_syms, void* stream) {
  using namespace wf;
  WF_REQUIRE(columns >= 1 && states >= 1 && bits_per_symbol >= 1, "bad trellis shape");
  WF_REQUIRE((int64_t)columns * states * (1 << bits_per_symbol) <= 1024, "trellis tables exceed 1024 entries");
  WF_REQUIRE(N % bits_per_symbol == 0, "Input length must be a multiple of FSM cardinality.");
  WF_REQUIRE(next_state && out_symbol, "NULL trellis table");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0 || N == 0) return WF_OK;
  TrellisParams p;
  memset(&p, 0, sizeof(p));
  p.bits = d_bits_in; p.syms = d_syms; p.enc_state = d_enc_state; p.B = B; p.N = N;
  p.columns = columns; p.states = states; p.card = bits_per_symbol;
  size_t n = (size_t)columns * states * (1u << bits_per_symbol);
  memcpy(p.next_state, next_state, n);
  memcpy(p.out_symbol, out_symbol, n);
  trellis_encode_kernel<<<(unsigned)((B + 127) / 128), 128, 0, as_stream(stream)>>>(p);
  WF_LAUNCH_CHECK("trellis_encode_kernel");
  return WF_OK;
}

}  // extern "C"
