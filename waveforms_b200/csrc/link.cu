// link.cu -- wf_link_sim_fused: one C call per batch of the SOQPSK Monte-Carlo link
// (examples/soqpsk_detection.py:45-216 for B frames at once): PN bits + precode (wf_prbs_encode) ->
// modulate + AWGN + PAM bank in one kernel (wf_cpm_modulate_bank) -> 4-state detector + error count fed by
// the two raw bank outputs (wf_viterbi_soqpsk_combined).  Three kernels, one host call; the only arrays that
// reach HBM are the packed bits, the symbols and the 16 B-per-call bank outputs, all carved from ONE
// caller-provided workspace (the library never allocates).  Counters are bit-identical to the three calls.
#include "common.cuh"

namespace wf {

static_assert(sizeof(wf_link_cfg) == 152, "wf_link_cfg layout is part of the ABI (waveforms_b200/_ffi.py:LinkCfg mirrors it)");

struct LinkLayout {
  int64_t n_sym, ld_words, ld_z;
  size_t off_packed, off_syms, off_z, total;
};

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int link_layout(const wf_link_cfg* c, int64_t B, LinkLayout* L) {
  WF_REQUIRE(c, "cfg is NULL");
  WF_REQUIRE(B >= 0 && c->n_bits > 0 && c->n_calls > 0, "bad shape (B %lld, n_bits %lld, n_calls %lld)", (long long)B,
             (long long)c->n_bits, (long long)c->n_calls);
  WF_REQUIRE(c->encoder == WF_ENC_SOQPSK_DIFF || c->encoder == WF_ENC_SOQPSK,
             "wf_link_sim_fused covers the SOQPSK encoders only (encoder %d)", c->encoder);
  WF_REQUIRE(c->n_filt == 2 && c->combine_re && c->combine_im, "the fused link needs the 2-filter PAM bank and its 3x2 combine");
  L->n_sym = c->n_bits;
  L->ld_words = (c->n_bits + 31) / 32;
  L->ld_z = (c->n_calls + 3) & ~(int64_t)3;
  size_t o = 0;
  L->off_packed = o; o = align_up(o + (size_t)B * L->ld_words * sizeof(uint32_t), 256);
  L->off_syms = o;   o = align_up(o + (size_t)B * L->n_sym, 256);
  L->off_z = o;      o = align_up(o + (size_t)B * L->ld_z * 2 * sizeof(float2), 256);
  L->total = o;
  return WF_OK;
}

}  // namespace wf

extern "C" int64_t wf_link_sim_workspace_bytes(const wf_link_cfg* cfg, int64_t B) {
  wf::LinkLayout L;
  int rc = wf::link_layout(cfg, B, &L);
  return rc ? (int64_t)rc : (int64_t)L.total;
}

extern "C" int wf_link_sim_fused(const wf_link_cfg* c, int64_t B, uint64_t first_frame, const uint64_t* d_frame_ids,
                                 const uint64_t* d_bit_offsets, const float* d_sigma, const int32_t* d_point,
                                 void* d_workspace, int64_t workspace_bytes, int64_t* d_counters, void* stream) {
  using namespace wf;
  LinkLayout L;
  int rc;
  if ((rc = link_layout(c, B, &L))) return rc;
  WF_REQUIRE(d_counters, "d_counters is NULL");
  WF_REQUIRE((d_frame_ids == nullptr) == (d_bit_offsets == nullptr), "d_frame_ids and d_bit_offsets go together");
  WF_REQUIRE(d_workspace && workspace_bytes >= (int64_t)L.total, "workspace of %lld bytes given, %zu needed "
             "(wf_link_sim_workspace_bytes)", (long long)workspace_bytes, L.total);
  WF_REQUIRE((reinterpret_cast<uintptr_t>(d_workspace) & 255) == 0, "workspace must be 256-byte aligned");
  if (!wf_cpm_modulate_bank_supported(c->K, c->sps, c->n_filt, c->n_taps, c->center, c->first_instant, L.n_sym, c->n_calls))
    return fail(WF_ERR_UNSUPPORTED, "wf_link_sim_fused: shape not covered by wf_cpm_modulate_bank (K %d, sps %d, taps %d, "
                "first instant %lld); run the separate calls", c->K, c->sps, c->n_taps, (long long)c->first_instant);
  if (B == 0) return WF_OK;
  char* ws = static_cast<char*>(d_workspace);
  uint32_t* packed = reinterpret_cast<uint32_t*>(ws + L.off_packed);
  int8_t* syms = reinterpret_cast<int8_t*>(ws + L.off_syms);
  void* z = ws + L.off_z;
  if ((rc = wf_prbs_encode(c->mask, c->degree, c->init_state, c->period, d_bit_offsets, first_frame, B, c->n_bits,
                           c->encoder, nullptr, nullptr, packed, L.ld_words, syms, stream)))
    return rc;
  if ((rc = wf_cpm_modulate_bank(syms, B, L.n_sym, c->mod_index, c->n_h, c->pulse, c->K, c->sps, c->initial_phase,
                                 c->derotate, d_sigma, d_point, c->seed, first_frame, d_frame_ids, c->taps_re, c->n_filt,
                                 c->n_taps, c->center, c->first_instant, c->n_calls, z, L.ld_z, c->tune, stream)))
    return rc;
  return wf_viterbi_soqpsk_combined(z, WF_F32, B, c->n_calls, L.ld_z, c->n_filt, c->combine_re, c->combine_im, c->diff,
                                    packed, L.ld_words, L.n_sym, d_point, d_counters, stream);
}
