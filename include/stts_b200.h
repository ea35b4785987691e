/* stts_b200 -- C ABI of the B200-native Stylish-TTS synthesis forward path.
 *
 * The reference (Fannovel16/stylish-tts) has no FFI / plugin interface for this path: its
 * boundary is the Python nn.Module API.  This header is therefore the
 * boundary *underneath* the Python module mirrors in stylish_tts_b200/modules; each entry point
 * names the reference call it replaces (paths relative to src/stylish_tts/train/).
 *
 * Conventions
 *   - every function returns 0 on success or a negative stts_status; stts_last_error() gives text
 *   - all `d_` pointers are DEVICE pointers, fp32 unless stated; `stream` is a cudaStream_t
 *   - no hot call allocates, synchronises, or keeps a caller pointer past its return
 *   - "dense" tensors are the reference layout of ONE utterance ([C, L] row-major, i.e. a B = 1
 *     slice of [B, C, L]) concatenated over the utterances of the batch; per-utterance element
 *     offsets follow from the layout set with stts_set_layout (off = C * cumsum(L))
 *   - the batch is a set of independent utterances ("segments"); statistics never cross them
 */
#ifndef STTS_B200_H
#define STTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct stts_ctx stts_ctx;

enum stts_status {
  STTS_OK = 0,
  STTS_ERR_BAD_ARG = -1,
  STTS_ERR_BAD_SHAPE = -2,
  STTS_ERR_MISSING_WEIGHT = -3,
  STTS_ERR_NO_LAYOUT = -4,
  STTS_ERR_CUDA = -5,
  STTS_ERR_UNSUPPORTED = -6,
  STTS_ERR_UNVOICED_MIN = -7 /* voiced frames but no f0 > 20 Hz: the reference raises (generator.py:285) */
};

/* arithmetic of the dense contractions */
enum stts_precision {
  STTS_FP32 = 0,   /* CUDA-core fp32 FMA everywhere (parity reference on device)                      */
  STTS_F16 = 1,    /* 16-bit mode: tcgen05 kind::f16 on IEEE fp16 operand twins, fp32 accumulate, one
                      pass; the forward STFT and the token-rate encoders run the three-pass fp16 split
                      (f16x3, fp32-grade); statistics fp32 / fp64                                      */
  STTS_TF32X3 = 2  /* tcgen05 kind::tf32, error-compensated 3-pass split (fp32-grade) everywhere      */
};

/* ---- lifetime ------------------------------------------------------------------------------ */
int stts_create(stts_ctx** out, int device);
void stts_destroy(stts_ctx* ctx);
const char* stts_last_error(const stts_ctx* ctx);
const char* stts_version(void);

/* ---- weights: state_dict() of the reference modules, key by key -------------------------------
 * key is the reference state_dict key (SURVEY 8b "Checkpoint layout"); h_data is a HOST fp32 buffer, copied;
 * module in {"speech_predictor", "pitch_energy_predictor", "pe_text_encoder", "pe_text_style_encoder",
 * "duration_predictor"}.  Unknown keys are kept but unused.                                       */
int stts_load_tensor(stts_ctx* ctx, const char* module, const char* key, const float* h_data,
                     const int64_t* shape, int ndim);
/* folds both weight-norm dialects, builds the K-major fp32 / tf32 hi-lo / fp16 hi-lo packs and the
 * STFT / iSTFT bases.  Replaces nothing in the reference (it re-parametrises on every forward). */
int stts_finalize_weights(stts_ctx* ctx);
int stts_set_precision(stts_ctx* ctx, int precision);
/* Programmatic dependent launch of the library's kernels (on by default): the next kernel of the
 * stream is scheduled while the current one runs and waits on-device for its completion.  It
 * shortens a single stream of small dependent kernels (one utterance: 3.6 -> 3.3 ms); with
 * several contexts feeding one GPU concurrently the waiting CTAs cost 1-2 %, so turn it off
 * there.  No reference counterpart (launch mechanics).                                          */
int stts_set_pdl(stts_ctx* ctx, int on);
/* Small-batch kernel shapes (on by default): when a batch is so small that the throughput shapes
 * would occupy a fraction of the SMs (one 5 s utterance: 13 row tiles, 4-32 attention CTAs), the
 * GEMMs take 128- instead of 256-column tiles and the attention kernels one query per lane group
 * with 4-16 key sets (one utterance: 2.77 -> 2.56 ms with the split-K LayerNorm GEMM).  The shapes
 * differ in summation order: a small batch and the same utterances inside a large batch agree to
 * rounding (16-bit mode: to its operand rounding), not bit for bit; off = one shape per layer
 * whatever the batch.  No reference counterpart (launch mechanics).                              */
int stts_set_small_batch_shapes(stts_ctx* ctx, int on);

/* ---- packed layout ---------------------------------------------------------------------------
 * len_f[b]  = frames at hop 300 (F), len_f4[b] = frames at hop 75 (normally 4 F).  Either may be
 * NULL.  (Re)allocates and zeroes the context workspace when the layout changes.              */
int stts_set_layout(stts_ctx* ctx, int n_seg, const int* len_f, const int* len_f4);
int64_t stts_workspace_bytes(const stts_ctx* ctx);

/* ---- token-rate encoders ------------------------------------------------------------------------
 * len_t[b] = phoneme tokens of utterance b (pad tokens included, like text_lengths).              */
int stts_set_token_layout(stts_ctx* ctx, int n_seg, const int* len_t);
/* TextEncoder.forward, text_encoder.py:433-462.  which: "speech_predictor" (its text_encoder.*),
 * "pe_text_encoder", "duration_predictor" (its text_encoder.*) -- the module the weights were
 * loaded under.  d_tokens int32 [sum T] (utterances concatenated).  Outputs, token-major dense and
 * either may be NULL: d_mu_tok [sum T, inter_dim] (= mu.transpose(1,2)), d_x_tok [sum T, 128].
 * Always fp32-grade arithmetic (the duration logits decide integer durations): fp32 FMA in STTS_FP32,
 * tf32x3 in STTS_TF32X3, f16x3 in STTS_F16.                                                         */
int stts_text_encoder(stts_ctx* ctx, const char* which, const int32_t* d_tokens, float* d_mu_tok,
                      float* d_x_tok, void* stream);

/* ProsodyEncoder.forward, prosody_encoder.py:63-81.  which: "duration_predictor" or
 * "pitch_energy_predictor" (their prosody_encoder.*).  d_enc_tok token-major dense [sum T, d_model]
 * (= text encoding .transpose(1,2)), d_style [n_seg, 64]; d_out_tok [sum T, d_model + 64] (may be
 * NULL); d_logits_tok [sum T, 16] = DurationPredictor.duration_proj of the result
 * (duration_predictor.py:30-36; "duration_predictor" only, may be NULL).                            */
int stts_prosody_encoder(stts_ctx* ctx, const char* which, const float* d_enc_tok, const float* d_style,
                         float* d_out_tok, float* d_logits_tok, void* stream);

/* TextStyleEncoder.forward, text_style_encoder.py:20-26 (conv k7, BasicConvNeXtBlock x n
 * conv_next.py:18-51, mean over the utterance).  which: "speech_predictor" (its style_encoder.*),
 * "pe_text_style_encoder", "duration_predictor" (its style_encoder.*).  d_enc_tok token-major dense
 * [sum T, inter_dim], or NULL = the mu rows stts_text_encoder of the same chain just produced
 * (speech_predictor <- speech_predictor, pe_text_style_encoder <- pe_text_encoder,
 * duration_predictor <- duration_predictor).  d_style [n_seg, 64].                                   */
int stts_text_style_encoder(stts_ctx* ctx, const char* which, const float* d_enc_tok, float* d_style,
                            void* stream);

/* ---- length regulation ------------------------------------------------------------------------
 * DurationProcessor.prediction_to_duration, utils.py:468-474: logits [n_tokens,16] -> int32.    */
int stts_durations(stts_ctx* ctx, const float* d_logits, int n_tokens, int32_t* d_dur, void* stream);
/* DurationProcessor.duration_to_alignment, utils.py:476-489, as indices: for utterance b the
 * token owning frame f is d_idx[d_cu_frames[b] + f] (= argmax over tokens of the one-hot
 * alignment column).  d_cu_tok [n_seg+1] are token offsets.  *d_status != 0 if cap too small.   */
int stts_frame_index(stts_ctx* ctx, const int32_t* d_dur, const int32_t* d_cu_tok, int n_seg,
                     int32_t* d_cu_frames, int32_t* d_idx, int idx_cap, int32_t* d_status, void* stream);
/* No counterpart in the reference: added for callers that
 * cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same result as stts_frame_index, but the scratch the kernels need (the n_seg + 1 exclusive frame
 * offsets, d_cu_frames above) lives in a file-scope `static __device__` array of 4 352 bytes that
 * every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree and
 * calls no allocating routine.  Limit: 1 <= n_seg <= 1024 (STTS_ERR_BAD_ARG with a message beyond
 * it).  d_idx [idx_cap] and d_status [1] are the caller's; *d_status must be 0 on entry and is
 * set to 1 if idx_cap is too small (same contract as stts_frame_index).  Every call on a device
 * shares the one array: calls that overlap on the GPU (different streams) race; calls on one
 * stream are ordered and never see the previous call's contents.                               */
int stts_frame_index_static(stts_ctx* ctx, const int32_t* d_dur, const int32_t* d_cu_tok, int n_seg,
                            int32_t* d_idx, int idx_cap, int32_t* d_status, void* stream);

/* ---- source / activation variants outside the default model.yml -------
 * SineGenerator.forward, train/models/cfm/cfm_mel_decoder.py:16-101 (incl. _f02sine :58-83 and
 * _f02uv :52-56) on a ragged batch.  d_f0: per-SAMPLE f0 in Hz, the utterances back to back
 * (sum(len) floats; the reference's [B, L, 1]); len [n_seg] HOST lengths in samples;
 * d_rand_ini [n_seg, dim] the uniform initial-phase draws (column 0 is ignored, :69) and d_noise
 * [sum(len), dim] the normal draws the reference makes inside forward (:66, :97), dim =
 * harmonic_num + 1 <= 16; merge_weight [dim] HOST = merge[0].weight (Linear without bias, :47-50).
 * d_out [sum(len)] = tanh(merge(...)); d_phase (optional, may be NULL) [sum(len), dim] = the
 * wrap-corrected cumulative phase in cycles (the argument of :78 before * 2 pi), bit-equal to the
 * reference's: torch's CPU cumsum accumulates in fp64 and rounds each output to fp32.            */
int stts_source_sinegen(stts_ctx* ctx, const float* d_f0, const int* len, int n_seg, const float* d_rand_ini,
                        const float* d_noise, const float* merge_weight, int harmonic_num, float samp_rate,
                        float sine_amp, float noise_std, float voiced_threshold, float* d_out, float* d_phase,
                        void* stream);
/* AdaptiveGeneratorBlock.forward, train/models/ada_norm.py:109-120 (AdaptiveInstance :130-140,
 * Snake1D x + sin^2(a x) / a, dilated convs :14-83; built by UpsampleGenerator generator.py:114-163
 * with k in {3, 7, 11}, dilation (1, 3, 5)).  Weights: stts_load_tensor(ctx, module, key, ...) with
 * the block's own state_dict keys (convs1.i.*, convs2.i.*, adain1.i.fc.*, adain2.i.fc.*,
 * alpha1.i, alpha2.i; both weight-norm dialects), packed on the first call.  d_x / d_out: the
 * utterances' dense [channels, len_b] tensors back to back; len [n_seg] HOST; d_style [n_seg, 64].
 * channels % 64 == 0, kernel_size odd <= 15, padding (k d - d) / 2 <= 32.  The GEMM arithmetic
 * follows stts_set_precision.  Re-plans its private workspace (cudaMalloc) when len changes.    */
int stts_generator_block_snake(stts_ctx* ctx, const char* module, int channels, int kernel_size,
                               const int* dilation /*[3]*/, int n_seg, const int* len, const float* d_x,
                               const float* d_style, float* d_out, void* stream);

/* STFT.transform / STFT.inverse, train/models/stft.py:98-136 / :138-186 -- the conv1d / conv_transpose1d DFT module
 * swapped in for the ONNX export (convert_to_onnx.py:31-36).  Basis matrices: the module's
 * own buffers weight_forward_real / weight_forward_imag / weight_backward_real / weight_backward_imag
 * ([filter_length / 2 + 1, 1, filter_length], stft.py:69-96) loaded under `module` with stts_load_tensor.
 * transform: d_wave = the utterances back to back, len [n_seg] HOST samples; pad_mode 0 = replicate (the module's
 * default), 1 = constant; outputs per utterance a dense [bins, frames_b] block (frames_b = len_b / hop + 1), blocks
 * back to back: magnitude = sqrt(re^2 + im^2 + 1e-14), re / magnitude, im / magnitude.
 * inverse: same block layout in; writes out_len[b] <= (n_frames[b] - 1) * hop samples per utterance (NULL = all; the
 * module's `length` argument).  filter_length even, ceil(win_length / hop_length) <= 33.  fp32-grade arithmetic
 * (CUDA cores in fp32 mode, three-pass tf32 otherwise).                                                              */
int stts_stft_conv_transform(stts_ctx* ctx, const char* module, int filter_length, int hop_length, int win_length,
                             int pad_mode, int n_seg, const int* len, const float* d_wave, float* d_mag, float* d_x,
                             float* d_y, void* stream);
int stts_stft_conv_inverse(stts_ctx* ctx, const char* module, int filter_length, int hop_length, int win_length,
                           int n_seg, const int* n_frames, const int* out_len, const float* d_mag, const float* d_x,
                           const float* d_y, float* d_wave, void* stream);

/* ---- pitch / energy towers --------------------------------------------------------------------
 * PitchEnergyPredictor.forward lines 111-119 (pitch_energy_predictor.py): 2 x 3
 * AdaptiveDecoderBlock(320) + 1x1 projections on the cross-attended frames.
 * d_x: dense [320, F_b] (NULL = take the frames stts_pe_cross left in the workspace);
 * d_style [n_seg, 64]; outputs dense [sum F].                                                    */
int stts_pe_towers(stts_ctx* ctx, const float* d_x, const float* d_style, float* d_pitch,
                   float* d_energy, void* stream);
/* PitchEnergyPredictor.compute_cross, pitch_energy_predictor.py:83-102: tokens gathered to frames,
 * query / key AdaptiveLayerNorm, 8-head cross attention (partial RoPE, inverted +-5 band bias),
 * depthwise k5 + SiLU + 1x1, (base + attention) / sqrt 2.  Needs both the token layout and the
 * hop-300 layout.  d_prosody_tok token-major dense [sum T, 320] or NULL = the rows
 * stts_prosody_encoder("pitch_energy_predictor") just produced; d_idx / d_cu_frames from
 * stts_frame_index.  The result stays in the workspace as the input of stts_pe_towers (call it with
 * d_x = NULL); d_x (may be NULL) additionally receives it dense [320, F_b].                          */
int stts_pe_cross(stts_ctx* ctx, const float* d_prosody_tok, const float* d_style, const int32_t* d_idx,
                  const int32_t* d_cu_frames, float* d_x, void* stream);

/* ---- module-level stages of SpeechPredictor.forward (speech_predictor.py:85-129) --------------
 * Decoder.forward, decoder.py:47-60.  d_asr dense [128, L_b]; d_f0/d_n dense [sum L];
 * d_out dense [512, L_b].                                                                      */
int stts_decoder(stts_ctx* ctx, const float* d_asr, const float* d_f0, const float* d_n,
                 const float* d_style, float* d_out, void* stream);
/* PriorEncoder.forward (flow.py:311-315) + ResidualCouplingBlock.forward(reverse=True)
 * (flow.py:132-151) + post_flow (speech_predictor.py:111).  d_x dense [512, L_b]; d_noise dense
 * [128, L_b] is the randn_like draw; d_mel dense [512, L_b]; optional d_z_prior/d_z_flow dense
 * [128, L_b] (may be NULL).                                                                    */
int stts_prior_flow(stts_ctx* ctx, const float* d_x, const float* d_noise, const float* d_style,
                    float* d_mel, float* d_z_prior, float* d_z_flow, void* stream);
/* generate_pcph (generator.py:247-315): d_pitch4 dense [sum L] (Hz per hop-75 frame), d_noise
 * dense [sum 75 L] raw randn draw, d_init_phase [n_seg] the torch.rand((1,1)) draw per utterance;
 * d_out dense [sum 75 L].                                                                      */
int stts_source_pcph(stts_ctx* ctx, const float* d_pitch4, const float* d_noise,
                     const float* d_init_phase, float* d_out, void* stream);
/* Generator.forward (generator.py:402-438).  d_mel dense [512, L_b]; outputs: d_audio dense
 * [sum 75 L]; optional d_logamp / d_phase dense [1025, L_b + 1] (replicate-padded frame
 * included, like DecoderPrediction.magnitude / .phase).                                         */
int stts_generator(stts_ctx* ctx, const float* d_mel, const float* d_style, const float* d_pitch4,
                   const float* d_noise, const float* d_init_phase, float* d_audio, float* d_logamp,
                   float* d_phase, void* stream);

/* Generator.forward from generator.py:412 on, with the STFT of the source (|X| and atan2 phase,
 * dense [1025, L_b], generator.py:407-410) supplied by the caller.  Test / analysis entry: the
 * phase of near-real bins is rounding noise in the reference (DESIGN.md "ill-conditioned phase"). */
int stts_generator_from_stft(stts_ctx* ctx, const float* d_mel, const float* d_style,
                             const float* d_har_spec, const float* d_har_phase, float* d_audio,
                             float* d_logamp, float* d_phase, void* stream);

/* ---- fused frame-rate path --------------------------------------------------------------------
 * Everything of SpeechPredictor.forward after the token-rate encoders: alignment expansion
 * (:88,:93), x4 linear upsample (:89-90), decoder, prior, flow^-1, post_flow, generator.
 * d_enc_tok: token-major text encoding [sum T, 128]; d_cu_tok [n_seg+1]; d_idx / d_cu_frames from
 * stts_frame_index; d_pitch / d_energy dense [sum F]; d_style [n_seg,64]; noise as above.        */
int stts_speech_frames(stts_ctx* ctx, const float* d_enc_tok, const int32_t* d_cu_tok,
                       const int32_t* d_idx, const int32_t* d_cu_frames, const float* d_pitch,
                       const float* d_energy, const float* d_style, const float* d_noise_prior,
                       const float* d_noise_source, const float* d_init_phase, float* d_audio,
                       float* d_logamp, float* d_phase, void* stream);
/* after a synchronised call: STTS_ERR_UNVOICED_MIN if any utterance hit the reference's error path */
int stts_check_source_status(stts_ctx* ctx, void* stream);

/* 16-bit mode range audit (synchronises): number of elements of the workspace's fp16 operand twins that sit
 * at the saturation value +-65504 after the last run.  The fp16 twins clip where the reference's fp32 (or a
 * true bfloat16 path) would not; a non-zero count means the checkpoint's activations left the range the
 * 16-bit mode was validated for and an fp32-grade mode (STTS_TF32X3) should be used.  Replaces nothing in
 * the reference (it has no reduced-precision inference mode, config/config.yml:12).                  */
int64_t stts_saturation_count(stts_ctx* ctx, void* stream);

/* int16 PCM of n device samples: trunc(x * 32768), saturated -- replaces
 * `np.multiply(outputs[0], 32768).astype(np.int16)` (train/test_onnx.py:79) so that the device -> host
 * copy of a synthesised batch moves 2 bytes per sample.  d_pcm: int16 [n].                      */
int stts_pcm16(stts_ctx* ctx, const float* d_audio, int64_t n, int16_t* d_pcm, void* stream);

/* ---- test / profiling hooks -------------------------------------------------------------------
 * Raw conv-as-GEMM on caller buffers (packed [R, ld] layout of the current hop-75 layout):
 * Y[r, n] = bias[n] + sum_j sum_c X[r + j - pad, c] * W[n, c, j];  d_w is [N, cin, taps] fp32
 * (PyTorch Conv1d layout).  backend: stts_precision.  Used by the kernel parity tests.          */
int stts_test_conv(stts_ctx* ctx, int backend, const float* d_x, int ldx, int cin, int taps, int pad,
                   const float* h_w, const float* h_bias, int n_out, float* d_y, int ldy, void* stream);
/* copy a named workspace buffer (packed rows) to d_dst; returns rows*ld written or negative.   */
int64_t stts_debug_read(stts_ctx* ctx, const char* name, float* d_dst, int64_t cap, int* rows, int* ld,
                        void* stream);
int stts_layout_info(const stts_ctx* ctx, int rate, int* rows, int* seg_start /*[n_seg]*/);
/* per-GEMM CUDA-event timing (bench.py roofline): stts_set_profiling(ctx, 1) starts recording an
 * event pair around every conv-as-GEMM launch on the launching stream; stts_profile_json
 * synchronises the device and returns a JSON array of {name, backend, launches, ms, flops}
 * (algorithmic FLOPs = 2 * valid rows * N * un-padded K) aggregated per layer, and clears.      */
int stts_set_profiling(stts_ctx* ctx, int on);
const char* stts_profile_json(stts_ctx* ctx);
/* number of kernels launched by this context since creation (bench.py "gpu_launches")          */
int64_t stts_launch_count(const stts_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* STTS_B200_H */
