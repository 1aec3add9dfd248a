// bert.cu -- A13 stand-alone error count on unpacked decisions
// (examples/soqpsk_detection.py:200-209: drop the first `skip` detector outputs,
// compare the next n_cmp with the transmitted bits / symbols, count non-zeros).
// One warp per frame, lanes stride over the compared positions; warp reduce and
// one atomic triple per frame into counters[point] = (bit errors, symbol errors, compared).
#include "common.cuh"

namespace wf {

__global__ void __launch_bounds__(128)
bert_kernel(const uint8_t* __restrict__ dec_bits, const int8_t* __restrict__ dec_syms, int64_t ld_dec,
            const uint8_t* __restrict__ tx_bits, const int8_t* __restrict__ tx_syms, int64_t ld_tx,
            int64_t B, int64_t skip, int64_t n_cmp, const int32_t* __restrict__ point,
            unsigned long long* counters) {
  const int lane = threadIdx.x & 31;
  const int64_t frame = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (frame >= B) return;
  unsigned long long be = 0, se = 0;
  for (int64_t j = lane; j < n_cmp; j += 32) {
    if (dec_bits && tx_bits) be += (dec_bits[frame * ld_dec + skip + j] != tx_bits[frame * ld_tx + j]) ? 1u : 0u;
    if (dec_syms && tx_syms) se += (dec_syms[frame * ld_dec + skip + j] != tx_syms[frame * ld_tx + j]) ? 1u : 0u;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    be += __shfl_xor_sync(0xFFFFFFFFu, be, s);
    se += __shfl_xor_sync(0xFFFFFFFFu, se, s);
  }
  if (lane == 0) {
    const int64_t pt = point ? point[frame] : 0;
    atomicAdd(&counters[3 * pt + 0], be);
    atomicAdd(&counters[3 * pt + 1], se);
    atomicAdd(&counters[3 * pt + 2], (unsigned long long)n_cmp);
  }
}

}  // namespace wf

extern "C" int wf_bert(const uint8_t* d_dec_bits, const int8_t* d_dec_syms, int64_t ld_dec,
                       const uint8_t* d_tx_bits, const int8_t* d_tx_syms, int64_t ld_tx, int64_t B,
                       int64_t skip, int64_t n_cmp, const int32_t* d_point, int64_t* d_counters, void* stream) {
  using namespace wf;
  WF_REQUIRE(B >= 0 && skip >= 0 && n_cmp >= 0, "negative shape");
  WF_REQUIRE(ld_dec >= skip + n_cmp && ld_tx >= n_cmp, "leading dimension too small");
  WF_REQUIRE(d_counters, "d_counters is NULL");
  int rc;
  if ((rc = require_device())) return rc;
  if (B == 0) return WF_OK;
  bert_kernel<<<(unsigned)((B + 3) / 4), 128, 0, as_stream(stream)>>>(
      d_dec_bits, d_dec_syms, ld_dec, d_tx_bits, d_tx_syms, ld_tx, B, skip, n_cmp, d_point,
      reinterpret_cast<unsigned long long*>(d_counters));
  WF_LAUNCH_CHECK("bert_kernel");
  return WF_OK;
}
