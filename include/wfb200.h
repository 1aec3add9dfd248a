/* wfb200.h -- C ABI of libwfb200.so, the B200 (sm_100a) hot path of the CPM
 * Monte-Carlo link simulator.
 *
 * Boundary rules:
 *   - every bulk pointer named d_* is DEVICE memory owned by the caller; the
 *     library never allocates, frees or keeps it.  Small tables (pulses, taps,
 *     mod indices) are HOST pointers, copied into kernel parameter space.
 *   - every compute entry point is asynchronous on `stream` (a cudaStream_t
 *     passed as void*; NULL = legacy default stream) and re-entrant.
 *   - return value: 0 = ok, negative = wf_status; wf_last_error_string() gives
 *     the text for the calling thread.  There is no CPU fallback: on a machine
 *     without an sm_100 device the compute calls return WF_ERR_DEVICE.
 *
 * Each entry point names the reference interface it replaces (paths relative
 * to the reference repository mcdiarmid/waveforms).
 */
#ifndef WFB200_H
#define WFB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WF_ABI_VERSION 2

#if defined(__GNUC__)
#define WF_API __attribute__((visibility("default")))
#else
#define WF_API
#endif

enum wf_status {
  WF_OK = 0,
  WF_ERR_ARG = -1,     /* bad argument (text says which) */
  WF_ERR_DEVICE = -2,  /* no usable sm_100 CUDA device */
  WF_ERR_CUDA = -3,    /* CUDA runtime error (text = cudaGetErrorString) */
  WF_ERR_UNSUPPORTED = -4
};

/* element type of baseband / matched-filter / metric arrays */
enum wf_dtype { WF_F32 = 0 /* complex64 */, WF_F64 = 1 /* complex128 */ };

/* closed-form bit->symbol encoders (waveforms/cpm/trellis/model.py:179-293,
 * cpm/{soqpsk,pcmfm,multih}/precoder.py) */
enum wf_encoder {
  WF_ENC_SOQPSK_DIFF = 0,    /* TrellisEncoder(SOQPSKTrellis4x2DiffEncoded)            */
  WF_ENC_SOQPSK = 1,         /* TrellisEncoder(SOQPSKTrellis4x2) == SOQPSKTrellis8x1   */
  WF_ENC_BINARY = 2,         /* SimpleTrellis2 == PCMFMSymbolMapper: 2b-1              */
  WF_ENC_QUATERNARY = 3,     /* SimpleTrellis4 == MultiHSymbolMapper: 2(2b0+b1)-3      */
  WF_ENC_SOQPSK_TERNARY = 4  /* SOQPSKPrecoder: {-1,0,+1} = -(SOQPSKTrellis4x2)/2      */
};

/* ---- library / device ---------------------------------------------------- */
WF_API int wf_abi_version(void);
WF_API const char* wf_last_error_string(void);
/* 0 when the current CUDA device is compute capability 10.x, else WF_ERR_DEVICE */
WF_API int wf_device_check(void);

/* ---- A1  PRBS: waveforms/glfsr/glfsr.py:6-19, glfsr/pn.py:6-107 --------- */
/* generate_mask(degree) (pn.py:75-90); returns 0 when 1 < degree < 65 fails
 * (the Python shim raises the reference's KeyError). Host only. */
WF_API uint64_t wf_lfsr_mask(int degree);
/* state after `steps` applications of GLFSR.next_bit (glfsr.py:15-19) by
 * GF(2) jump-ahead. Host only. */
WF_API uint64_t wf_lfsr_jump(uint64_t mask, int degree, uint64_t state, uint64_t steps);

/* B frames of N bits from one Galois LFSR stream plus their symbols.
 * Frame b starts `d_bit_offsets[b]` steps after `init_state` when
 * d_bit_offsets != NULL, else (first_frame + b) * N steps (reduced mod 2^degree-1
 * when the state is non-zero and the polynomial maximal -- caller's choice via
 * `period`; period == 0 means "do not reduce").
 * Outputs (any may be NULL): d_bits uint8[B][N] (0/1, the reference layout),
 * d_packed uint32[B][ld_words] (bit i of a frame = word i/32, bit i%32),
 * d_syms int8[B][S] with S = N (N/2 for WF_ENC_QUATERNARY).
 * d_enc_state int32[B][2] = (i, state) of TrellisEncoder carried in/out
 * (encoder.py:11-12,44-46); NULL = fresh encoder per frame.
 * Replaces PNSequence/GLFSR.next_bit + TrellisEncoder.encode / the mappers. */
WF_API int wf_prbs_encode(uint64_t mask, int degree, uint64_t init_state, uint64_t period,
                   const uint64_t* d_bit_offsets, uint64_t first_frame,
                   int64_t B, int64_t N, int encoder,
                   int32_t* d_enc_state,
                   uint8_t* d_bits, uint32_t* d_packed, int64_t ld_words,
                   int8_t* d_syms, void* stream);

/* Same encoders on caller-supplied bits d_bits_in uint8[B][N] (0/1).
 * Replaces TrellisEncoder.encode (encoder.py:17-48) and the mapper __call__s. */
WF_API int wf_bits_encode(const uint8_t* d_bits_in, int64_t B, int64_t N, int encoder,
                   int32_t* d_enc_state,
                   uint32_t* d_packed, int64_t ld_words, int8_t* d_syms, void* stream);

/* Arbitrary Trellis object, table driven (one frame per thread).
 * next_state / out_symbol are HOST tables [columns][states][1<<bits_per_symbol]
 * built from forward_map (model.py:127-137; last branch wins on duplicates);
 * entries < 0 in next_state mark a missing branch (-> WF_ERR_ARG is NOT raised
 * on device; the symbol written is 0 and the state is kept). */
WF_API int wf_trellis_encode(const uint8_t* d_bits_in, int64_t B, int64_t N,
                      const int8_t* next_state, const int8_t* out_symbol,
                      int columns, int states, int bits_per_symbol,
                      int32_t* d_enc_state, int8_t* d_syms, void* stream);

/* ---- A6/A7  modulator and AWGN: cpm/modulate.py:28-101, noise.py:8-32 ---- */
/* cpm_modulate(symbols, mod_index, pulse_filter, sps) for B frames.
 * d_syms int8[B][S]; mod_index double[n_h] (cycled per symbol, modulate.py:91-92);
 * pulse double[K]; output d_out complex[B][ld_out] (ld_out >= T=(S+1)*sps, in
 * complex elements), dtype selects complex64 / complex128.
 * phase[n] = 2pi/sps * cumsum(convolve(upsampled, pulse, "same"))[n] + initial_phase;
 * the output is exp(j(phase + derotate)) (derotate = -pi/4 reproduces
 * examples/soqpsk_detection.py:85; 0 = plain cpm_modulate).
 * When d_sigma != NULL complex AWGN is added in the same pass (see wf_awgn_add).
 * d_frame_ids uint64[B] (may be NULL): the GLOBAL frame id of every row, which keys its noise
 * stream; NULL means row b is frame first_frame + b.  With it one launch can carry frames of
 * several Eb/N0 points (LinkSimulator.run packs a whole small sweep into one batch). */
WF_API int wf_cpm_modulate(const int8_t* d_syms, int64_t B, int64_t S,
                    const double* mod_index, int n_h,
                    const double* pulse, int K, int sps,
                    double initial_phase, double derotate,
                    int dtype, void* d_out, int64_t ld_out,
                    const float* d_sigma, const int32_t* d_point,
                    uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream);

/* Fused cpm_modulate -> + AWGN -> decimated real-tap bank (examples/soqpsk_detection.py:78-90 followed by
 * the PAM bank of :158-173) in ONE kernel: the noisy baseband stays in shared memory, only the bank's
 * raw filter outputs d_z complex64[B][ld_z][n_filt] reach HBM (same values, bit for bit, as
 * wf_cpm_modulate(.., WF_F32, ..) followed by wf_mf_bank(.., combine = NULL, ..) on the same arguments).
 * Built for the SOQPSK-TG shape only: sps 8, pulse with ceil((K-1)/8) == 8 and ((K-1)/2) % 8 == 0,
 * n_filt == 2 real filters of n_taps == 73; wf_cpm_modulate_bank_supported() says whether a shape is
 * covered (1) or not (0); uncovered shapes return WF_ERR_UNSUPPORTED and the caller runs the two calls.
 * tune selects the kernel (same results, bit for bit): 0 = one BLOCK per frame (512-interval chunks, bank
 * phase on four of eight warps, two block barriers per chunk); >= 1 = one WARP per frame, no block barrier
 * (64-interval chunks, every lane does both phases, 2 bank calls per lane). */
WF_API int wf_cpm_modulate_bank_supported(int K, int sps, int n_filt, int n_taps, int center,
                                          int64_t first_instant, int64_t S, int64_t n_calls);
WF_API int wf_cpm_modulate_bank(const int8_t* d_syms, int64_t B, int64_t S,
                    const double* mod_index, int n_h,
                    const double* pulse, int K, int sps,
                    double initial_phase, double derotate,
                    const float* d_sigma, const int32_t* d_point,
                    uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids,
                    const double* taps_re, int n_filt, int n_taps, int center,
                    int64_t first_instant, int64_t n_calls,
                    void* d_z, int64_t ld_z, int tune, void* stream);

/* frequency_modulate(freq_pulses, sps, initial_phase) (cpm/modulate.py:28-54) for B rows
 * of arbitrary per-sample frequency pulses d_freq double[B][ld_in]:
 * out[n] = exp(j(((sum_{m<=n} f[m]) mod sps) * 2pi/sps + initial_phase)). */
WF_API int wf_frequency_modulate(const double* d_freq, int64_t B, int64_t T, int64_t ld_in, int sps,
                                 double initial_phase, int dtype, void* d_out, int64_t ld_out, void* stream);

/* phase_modulate(phase, sensitivity) (cpm/modulate.py:12-25): exp(j * sensitivity * phase[i]) */
WF_API int wf_phase_modulate(const double* d_phase, int64_t n, double sensitivity, int dtype, void* d_out,
                             void* stream);

/* generate_complex_awgn(sigma, size) added in place: d_sig[b][n] += sigma[b] *
 * (N(0,1) + j N(0,1)), sigma per real component as in noise.py.  The noise of a sample depends
 * only on (seed, point, global frame, sample index) -- never on batching or sharding:
 *   Philox4x32-10, key = (seed lo, seed hi), counter = (call, frame lo, frame hi ^ (call hi << 16), point).
 *   One call yields FOUR 32-bit words = FOUR complex samples.  Sample pair q (samples 2q, 2q+1) uses
 *   call(q) = 4 (q >> 3) + (q & 3) and its words (x, y) when (q & 4) == 0, else (z, w): pairs q and
 *   q ^ 4 (the same offset inside two neighbouring sps-8 symbol intervals) share a call.
 *   Word w -> one complex sample: radius uniform u = ((w & 0xFFFF) + 0.5) 2^-16 (when the 16-bit field
 *   is 0, u = (t + 0.5) 2^-48 with t the matching word of the call whose counter has bit 31 of
 *   `point` set), angle = ((w >> 16) + 0.5) 2 pi 2^-16 - pi,
 *   sample = sigma sqrt(-2 ln u) (cos angle + j sin angle).
 * Global frame of row b: d_frame_ids[b] when given, else first_frame + b.
 * d_sigma float[B]; d_point int32[B] (0 if NULL). */
WF_API int wf_awgn_add(void* d_sig, int dtype, int64_t B, int64_t T, int64_t ld,
                const float* d_sigma, const int32_t* d_point,
                uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids, void* stream);

/* raw N(0,1) pairs, for statistical tests: d_out float2[B][T] */
WF_API int wf_awgn_fill(float* d_out, int64_t B, int64_t T, const float* d_sigma,
                 const int32_t* d_point, uint64_t seed, uint64_t first_frame, const uint64_t* d_frame_ids,
                 void* stream);

/* ---- A8/A9  decimated matched-filter bank:
 *      examples/soqpsk_detection.py:134-173 (np.convolve(rx, taps, "same") then
 *      the driver's 1-in-sps column pick, :189-196) ------------------------- */
/* y_f[i] = sum_t rx[n_i + center - t] * taps[f][t],  n_i = first_instant + i*sps,
 * rx outside [0,T) = 0;  z_r[i] = sum_f combine[r][f] * y_f[i] (combine NULL ->
 * z = y, n_out = n_filt).  taps_im NULL -> real taps.  All tables HOST double.
 * d_rx complex[B][ld_rx]; d_out complex[B][ld_out][n_out] (ld_out >= n_calls).
 * center = (n_taps-1)/2 reproduces "same". */
WF_API int wf_mf_bank(const void* d_rx, int dtype, int64_t B, int64_t T, int64_t ld_rx,
               const double* taps_re, const double* taps_im, int n_filt, int n_taps,
               int center, const double* combine_re, const double* combine_im, int n_out,
               int sps, int64_t first_instant, int64_t n_calls,
               void* d_out, int64_t ld_out, void* stream);

/* ---- A10/A13  SOQPSK trellis detector + error count:
 *      viterbi/algorithm.py:19-101 driven as examples/soqpsk_detection.py:177-209 */
/* d_z complex[B][ld_z][3] = one matched-filter 3-vector (alpha=-2,0,+2) per
 * detector call.  Runs n_calls calls of SOQPSKTrellisDetector(length,
 * differantial_encoding=diff).iteration per frame and keeps element [0] of each.
 * Optional outputs: d_dec_bits uint8[B][n_calls], d_dec_syms int8[B][n_calls].
 * Optional error count (soqpsk_detection.py:200-209): decisions [length:] are
 * compared with d_tx_packed (bits, wf_prbs_encode layout) and d_tx_syms
 * int8[B][S]; n_cmp = min(S, n_calls-length) per frame; adds to
 * d_counters int64[n_points][3] = (bit errors, symbol errors, compared) at row
 * d_point[b] (row 0 if NULL).
 * d_state (NULL = fresh detector per frame): real[B][5 + 8*length] of the z element
 * type = (call parity, metrics[:,0], bi_history) carried in and out, so a stream can
 * be detected in chunks exactly like repeated .iteration() calls (algorithm.py:25-42,100). */
WF_API int wf_viterbi_soqpsk(const void* d_z, int dtype, int64_t B, int64_t n_calls, int64_t ld_z,
                      int diff, int length,
                      uint8_t* d_dec_bits, int8_t* d_dec_syms,
                      const uint32_t* d_tx_packed, int64_t ld_words,
                      const int8_t* d_tx_syms, int64_t S,
                      const int32_t* d_point, int64_t* d_counters, void* d_state, void* stream);

/* Counter-only variant fed by the bank's RAW filter outputs (the PAM bank of
 * examples/soqpsk_detection.py:158-173 has n_in = 2 real filters):
 * d_y complex[B][ld_y][n_in]; the detector forms z_r = sum_f combine[r][f] y_f (r = 0..2,
 * HOST float64[3][n_in], the conj(pseudo-symbol) matrix of :165-173) itself, with the same
 * sequence of fused operations as wf_mf_bank's combine stage, and then runs
 * SOQPSKTrellisDetector(length=2, differantial_encoding=diff) + the error count exactly as
 * wf_viterbi_soqpsk does.  Counters are bit-identical to wf_mf_bank(.., combine, ..) +
 * wf_viterbi_soqpsk(..); the intermediate array is n_in/3 of the size.  Rows must start on
 * 16-byte boundaries. */
WF_API int wf_viterbi_soqpsk_combined(const void* d_y, int dtype, int64_t B, int64_t n_calls, int64_t ld_y,
                               int n_in, const double* combine_re, const double* combine_im, int diff,
                               const uint32_t* d_tx_packed, int64_t ld_words, int64_t S,
                               const int32_t* d_point, int64_t* d_counters, void* stream);

/* stand-alone error count of unpacked decisions (soqpsk_detection.py:200-209):
 * drops the first `skip` decisions of each frame, compares n_cmp of them. */
WF_API int wf_bert(const uint8_t* d_dec_bits, const int8_t* d_dec_syms, int64_t ld_dec,
            const uint8_t* d_tx_bits, const int8_t* d_tx_syms, int64_t ld_tx,
            int64_t B, int64_t skip, int64_t n_cmp,
            const int32_t* d_point, int64_t* d_counters, void* stream);

/* ---- A11/A12  generic pulse-truncated coherent CPM trellis detector (PCM/FM, ARTM multi-h).
 *      NO reference implementation exists (waveforms.viterbi has only the SOQPSK detector);
 *      waveform parameters from examples/irig_comparison.py:50-72, design in DESIGN.md 4.5,
 *      float64 statement in oracle/cpm_generic.py.  Parity unpinned. ---------------------- */
/* d_z complex[B][ld_z][n_h * M^Lp]: per detection step the bank outputs for every symbol
 * hypothesis (newest symbol most significant), one block of M^Lp per step parity i % n_h.
 * h = h_num[k % n_h] / p.  State count 2p * M^(Lp-1) <= 256.  Start: every history equal; of the 2p phase states only
 * the parity class reachable from an empty channel (v = 0 after the Lp - 1 phantom symbols; every step changes the
 * parity of v by K mod 2 for all its branches) -- the other class starts at -inf.  Decision rule (oracle/cpm_generic.py):
 * BLOCK RELEASE with survivor memory 2 D, D = `depth` (0 -> min(16, 32 / log2 M)): whenever t + 1 is a
 * multiple of D and t + 1 >= 2 D the survivor of the best state at step t releases the symbols of
 * steps [t - 2D + 1, t - D]; the best final survivor releases the rest.  This entry point keeps
 * 2 D-symbol register-exchange words (2 D log2 M <= 64).  Outputs: d_dec_syms int8[B][n_steps] (alphabet values,
 * may be NULL) and, when d_tx_syms int8[B][S] is given, counters[point][3] += (bit errors,
 * symbol errors, compared BITS = log2(M) per compared symbol) over symbols
 * [skip_head, min(S, n_steps) - skip_tail): column 0 / column 2 is a bit-error ratio for every
 * waveform, exactly as in wf_viterbi_soqpsk. */
WF_API int wf_viterbi_cpm(const void* d_z, int dtype, int64_t B, int64_t n_steps, int64_t ld_z,
                          int M, int Lp, int p, const int32_t* h_num, int n_h, int depth,
                          int8_t* d_dec_syms, const int8_t* d_tx_syms, int64_t S,
                          int64_t skip_head, int64_t skip_tail,
                          const int32_t* d_point, int64_t* d_counters, void* stream);

/* The same detector straight from the RECEIVED SAMPLES: the hypothesis bank of
 * wf_mf_bank(.., center = sps - 1, ..) is computed inside the detector kernel, so neither the bank
 * launch nor its output array exists (the detector reads 8 sps bytes of rx per step instead of
 * 8 n_h M^Lp bytes of z), and the survivors are decision bits in shared memory traced back every D
 * steps.  d_rx complex64[B][ld_rx]; taps_re / taps_im HOST double[n_h * M^Lp][sps] exactly as wf_mf_bank
 * takes them.  Built for the two IRIG shapes (sps 8, Lp 2, first_instant -1, hypothesis M^Lp-1-f the
 * complex conjugate of f): PCM/FM (M 2, h 7/10, depth <= 32) and ARTM CPM (M 4, h {4,5}/16, depth <= 32);
 * wf_cpm_detect_supported() says whether a shape is covered (1) or not (0); uncovered shapes return
 * WF_ERR_UNSUPPORTED and the caller runs wf_mf_bank + wf_viterbi_cpm.  Same decision rule and counters as
 * wf_viterbi_cpm (float32 arithmetic: decisions may differ from the float64 statement at near-ties only).
 * tune: path metrics are re-normalised every `tune` steps (0 / 1 = every step; 2, 4). */
WF_API int wf_cpm_detect_supported(int dtype, int64_t T, int64_t ld_rx, const double* taps_re, const double* taps_im,
                                   int sps, int64_t first_instant, int64_t n_steps,
                                   int M, int Lp, int p, const int32_t* h_num, int n_h, int depth);
WF_API int wf_cpm_detect(const void* d_rx, int dtype, int64_t B, int64_t T, int64_t ld_rx,
                         const double* taps_re, const double* taps_im, int sps, int64_t first_instant, int64_t n_steps,
                         int M, int Lp, int p, const int32_t* h_num, int n_h, int depth,
                         int8_t* d_dec_syms, const int8_t* d_tx_syms, int64_t S,
                         int64_t skip_head, int64_t skip_tail,
                         const int32_t* d_point, int64_t* d_counters, int tune, void* stream);

/* ---- Whole-chain entry point (SURVEY 8(b) export list): one C call per batch of the SOQPSK Monte-Carlo link,
 *      i.e. examples/soqpsk_detection.py:45-216 for B frames at once.  Equivalent, counters bit for bit, to
 *      wf_prbs_encode -> wf_cpm_modulate_bank -> wf_viterbi_soqpsk_combined on the same arguments; the
 *      packed bits, symbols and raw bank outputs live in ONE caller-provided workspace. ---------------- */
typedef struct wf_link_cfg {
  /* source: PN bits + precode (wf_prbs_encode) */
  uint64_t mask; uint64_t init_state; uint64_t period; int32_t degree; int32_t encoder; /* WF_ENC_SOQPSK_DIFF | WF_ENC_SOQPSK */
  int64_t n_bits;                       /* bits (= symbols) per frame */
  /* modulator + AWGN (wf_cpm_modulate) */
  const double* mod_index; const double* pulse; int32_t n_h; int32_t K; int32_t sps; int32_t tune;
  double initial_phase; double derotate; uint64_t seed;
  /* PAM bank (wf_mf_bank): taps_re HOST double[n_filt][n_taps], n_filt == 2 */
  const double* taps_re; int32_t n_filt; int32_t n_taps; int32_t center; int32_t diff;
  int64_t first_instant; int64_t n_calls;
  /* detector: conj(pseudo-symbol) combine HOST double[3][n_filt] (wf_viterbi_soqpsk_combined) */
  const double* combine_re; const double* combine_im;
} wf_link_cfg;
/* bytes of workspace wf_link_sim_fused needs for B frames (negative wf_status on a bad cfg) */
WF_API int64_t wf_link_sim_workspace_bytes(const wf_link_cfg* cfg, int64_t B);
/* d_sigma float[B], d_point int32[B] (NULL = row 0), d_frame_ids / d_bit_offsets uint64[B] (both or neither;
 * NULL = frames first_frame + b, bits (first_frame + b) * n_bits into the stream), d_workspace 256-byte aligned
 * device memory of at least wf_link_sim_workspace_bytes(cfg, B); adds (bit errors, symbol errors, compared bits)
 * to d_counters int64[n_points][3].  Shapes wf_cpm_modulate_bank does not cover return WF_ERR_UNSUPPORTED. */
WF_API int wf_link_sim_fused(const wf_link_cfg* cfg, int64_t B, uint64_t first_frame, const uint64_t* d_frame_ids,
                             const uint64_t* d_bit_offsets, const float* d_sigma, const int32_t* d_point,
                             void* d_workspace, int64_t workspace_bytes, int64_t* d_counters, void* stream);

/* ---- signal diagnostics as CUDA reductions (the reference only plots these:
 *      waveforms/viz/*.py, matplotlib; the figures are unavailable on this box, the arrays are not) ---- */
/* viz/psd.py:16-54 = matplotlib.mlab.psd(x * amplitude, NFFT = nfft, noverlap = 0, window = hanning,
 * scale_by_freq = False), two-sided and fft-shifted: d_out float[B][nfft] = mean over the T / nfft whole
 * segments of |FFT(hann * x)|^2 / (sum hann)^2, bin nfft/2 = DC.  d_sig complex64[B][ld].  nfft: power of two <= 4096. */
WF_API int wf_welch_psd(const void* d_sig, int64_t B, int64_t T, int64_t ld, int nfft, double amplitude,
                        float* d_out, void* stream);
/* viz/eye.py:39-56: d_out complex64[B][(T-1)/span][span+1], trace i = sig[i*span .. i*span+span] (span = sps * modulo) */
WF_API int wf_eye_traces(const void* d_sig, int64_t B, int64_t T, int64_t ld, int span, void* d_out, void* stream);
/* viz/tree.py:64-71: d_out float[B][T/span][span] = np.unwrap(np.angle(chunk)) - (use_off ? off : chunk's first value) */
WF_API int wf_phase_tree(const void* d_sig, int64_t B, int64_t T, int64_t ld, int span, int use_off, double off,
                         float* d_out, void* stream);
/* replacement for the hard-coded timing offsets of examples/soqpsk_detection.py:183-192: d_out float[B][sps],
 * d_out[b][ph] = | mean_i (-1)^i y[ph + sps i]^2 | of a full-rate matched-filter output d_y complex64[B][ld] of an
 * offset-QPSK-like CPM signal (y^2 alternates in sign between I and Q bits at the right sampling phase; the
 * envelope is flat).  The best sampling phase maximises it (timing_offset = -phase in the script's convention). */
WF_API int wf_timing_metric(const void* d_y, int64_t B, int64_t T, int64_t ld, int sps, float* d_out, void* stream);

/* ---- Allocation-free SOQPSK detection.  NO counterpart in the reference: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 *      serving loops, harnesses that forbid it). ------------------------------------------------------- */
/* wf_soqpsk_detect_static(..) == wf_mf_bank(d_rx, WF_F32, .., combine_re, combine_im, 3, .., z, n_calls, stream)
 * followed by wf_viterbi_soqpsk(z, WF_F32, B, n_calls, n_calls, diff, length, .., NULL, stream): same arguments,
 * same decisions and counters bit for bit, but the intermediate z (complex64[B][n_calls][3]) lives in a
 * file-scope device array inside the library (csrc/noalloc.cu) instead of a caller buffer, and the call
 * performs no cudaMalloc / cudaFree / allocating library routine.
 * LIMIT: B * n_calls <= wf_static_scratch_max_calls() = 1 398 101 (a 32 MiB array; every GPU that loads
 * the library holds it until the process exits); larger launches return WF_ERR_ARG with a message and do
 * nothing -- split the batch, or use the two-call form with your own buffer.
 * All calls share the one array (per GPU) and it keeps the last call's contents: the bank overwrites every
 * element the detector reads, so nothing needs clearing, but two calls that OVERLAP on the same GPU
 * (different streams without a dependency, or two host threads) race on it -- serialise them.
 * d_rx complex64[B][ld_rx]; the other arguments as in wf_mf_bank / wf_viterbi_soqpsk
 * (combine_re[3][n_filt] required unless n_filt == 3, the pulse-truncated bank, whose outputs are used as they are). */
WF_API int64_t wf_static_scratch_max_calls(void);
WF_API int wf_soqpsk_detect_static(const void* d_rx, int64_t B, int64_t T, int64_t ld_rx,
                                   const double* taps_re, const double* taps_im, int n_filt, int n_taps,
                                   int center, const double* combine_re, const double* combine_im,
                                   int sps, int64_t first_instant, int64_t n_calls,
                                   int diff, int length, uint8_t* d_dec_bits, int8_t* d_dec_syms,
                                   const uint32_t* d_tx_packed, int64_t ld_words,
                                   const int8_t* d_tx_syms, int64_t S,
                                   const int32_t* d_point, int64_t* d_counters, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WFB200_H */
