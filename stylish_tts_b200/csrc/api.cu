// extern "C" surface declared in include/stts_b200.h
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/stts_b200.h"
#include "engine.h"

using namespace stts;

struct stts_ctx {
  Context c;
};

namespace {
struct DeviceGuard {
  int prev = 0;
  explicit DeviceGuard(const Context& c) {      // also: this context's programmatic-dependent-launch choice for the call
    const int dev = c.device;
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev); else prev = -1;
    pdl_set(c.pdl);
    small_shapes_set(c.small_shapes);
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }
}  // namespace

extern "C" {

const char* stts_version(void) { return "stts_b200 0.1 (sm_100a)"; }

int stts_create(stts_ctx** out, int device) {
  if (out == nullptr) return STTS_ERR_BAD_ARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return STTS_ERR_CUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return STTS_ERR_CUDA;
  if (prop.major != 10) return STTS_ERR_UNSUPPORTED;  // built for sm_100a only, no fallback
  stts_ctx* ctx = new stts_ctx();
  ctx->c.device = device;
  *out = ctx;
  return STTS_OK;
}

void stts_destroy(stts_ctx* ctx) {
  if (ctx == nullptr) return;
  DeviceGuard g(ctx->c);
  cudaDeviceSynchronize();
  free_variant_work(ctx->c);
  for (void* p : ctx->c.weight_allocs) cudaFree(p);
  if (ctx->c.arena) cudaFree(ctx->c.arena);
  if (ctx->c.tarena) cudaFree(ctx->c.tarena);
  if (ctx->c.h_meta) cudaFreeHost(ctx->c.h_meta);
  if (ctx->c.ev_fork) cudaEventDestroy(ctx->c.ev_fork);
  if (ctx->c.ev_join) cudaEventDestroy(ctx->c.ev_join);
  if (ctx->c.st_aux) cudaStreamDestroy(ctx->c.st_aux);
  delete ctx;
}

const char* stts_last_error(const stts_ctx* ctx) { return ctx ? ctx->c.err.c_str() : "null context"; }

int stts_load_tensor(stts_ctx* ctx, const char* module, const char* key, const float* h_data, const int64_t* shape,
                     int ndim) {
  if (!ctx || !module || !key || !h_data || ndim < 0 || ndim > 8) return STTS_ERR_BAD_ARG;
  HostTensor t;
  int64_t n = 1;
  for (int i = 0; i < ndim; ++i) { t.shape.push_back(shape[i]); n *= shape[i]; }
  t.data.assign(h_data, h_data + n);
  ctx->c.tensors[std::string(module) + "/" + key] = std::move(t);
  ctx->c.finalized = false;
  ctx->c.snake.erase(module);      // a lazily packed AdaptiveGeneratorBlock of this name is stale now
  return STTS_OK;
}

int stts_finalize_weights(stts_ctx* ctx) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  ctx->c.sig.clear();  // style-FC scratch sizes depend on the packs: force a re-plan
  return finalize_weights(ctx->c);
}

int stts_set_precision(stts_ctx* ctx, int precision) {
  if (!ctx || precision < 0 || precision > 2) return STTS_ERR_BAD_ARG;
  ctx->c.precision = precision;
  return STTS_OK;
}

int stts_set_pdl(stts_ctx* ctx, int on) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  ctx->c.pdl = on ? 1 : 0;
  return STTS_OK;
}

int stts_set_small_batch_shapes(stts_ctx* ctx, int on) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  ctx->c.small_shapes = on ? 1 : 0;
  return STTS_OK;
}

int stts_set_token_layout(stts_ctx* ctx, int n_seg, const int* len_t) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return set_token_layout(ctx->c, n_seg, len_t);
}

int stts_text_encoder(stts_ctx* ctx, const char* which, const int32_t* d_tokens, float* d_mu_tok, float* d_x_tok,
                      void* stream) {
  if (!ctx || !which || !d_tokens) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  const std::string w(which);
  const char* nm = w == "speech_predictor" ? "te.sp" : (w == "pe_text_encoder" ? "te.pe"
                   : (w == "duration_predictor" ? "te.dp" : nullptr));
  if (nm == nullptr) return fail(ctx->c, STTS_ERR_BAD_ARG, "stts_text_encoder: unknown module " + w);
  return run_text_encoder(ctx->c, nm, d_tokens, d_mu_tok, d_x_tok, S(stream));
}

int stts_prosody_encoder(stts_ctx* ctx, const char* which, const float* d_enc_tok, const float* d_style,
                         float* d_out_tok, float* d_logits_tok, void* stream) {
  if (!ctx || !which || !d_enc_tok || !d_style) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  const std::string w(which);
  const char* nm = w == "duration_predictor" ? "pr.dp" : (w == "pitch_energy_predictor" ? "pr.pe" : nullptr);
  if (nm == nullptr) return fail(ctx->c, STTS_ERR_BAD_ARG, "stts_prosody_encoder: unknown module " + w);
  return run_prosody_encoder(ctx->c, nm, d_enc_tok, d_style, d_out_tok, d_logits_tok, S(stream));
}

int stts_text_style_encoder(stts_ctx* ctx, const char* which, const float* d_enc_tok, float* d_style, void* stream) {
  if (!ctx || !which || !d_style) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  const std::string w(which);
  const char* nm = w == "speech_predictor" ? "se.sp" : (w == "pe_text_style_encoder" ? "se.pe"
                   : (w == "duration_predictor" ? "se.dp" : nullptr));
  if (nm == nullptr) return fail(ctx->c, STTS_ERR_BAD_ARG, "stts_text_style_encoder: unknown module " + w);
  return run_style_encoder(ctx->c, nm, d_enc_tok, d_style, S(stream));
}

int stts_pe_cross(stts_ctx* ctx, const float* d_prosody_tok, const float* d_style, const int32_t* d_idx,
                  const int32_t* d_cu_frames, float* d_x, void* stream) {
  if (!ctx || !d_style || !d_idx || !d_cu_frames) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_pe_cross(ctx->c, d_prosody_tok, d_style, d_idx, d_cu_frames, d_x, S(stream));
}

int stts_set_layout(stts_ctx* ctx, int n_seg, const int* len_f, const int* len_f4) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return set_layout(ctx->c, n_seg, len_f, len_f4);
}

int64_t stts_workspace_bytes(const stts_ctx* ctx) { return ctx ? (int64_t)ctx->c.arena_used : 0; }

int stts_durations(stts_ctx* ctx, const float* d_logits, int n_tokens, int32_t* d_dur, void* stream) {
  if (!ctx || !d_logits || !d_dur || n_tokens < 0) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  DeviceGuard g(c);
  STTS_CK(durations_from_logits(d_logits, n_tokens, d_dur, S(stream)));
  return STTS_OK;
}

int stts_frame_index(stts_ctx* ctx, const int32_t* d_dur, const int32_t* d_cu_tok, int n_seg, int32_t* d_cu_frames,
                     int32_t* d_idx, int idx_cap, int32_t* d_status, void* stream) {
  if (!ctx || !d_dur || !d_cu_tok || !d_cu_frames || !d_idx || !d_status) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  DeviceGuard g(c);
  STTS_CK(frame_index(d_dur, d_cu_tok, n_seg, d_cu_frames, d_idx, idx_cap, d_status, S(stream)));
  c.launches++;
  return STTS_OK;
}

int stts_frame_index_static(stts_ctx* ctx, const int32_t* d_dur, const int32_t* d_cu_tok, int n_seg, int32_t* d_idx,
                            int idx_cap, int32_t* d_status, void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (!d_dur || !d_cu_tok || !d_idx || !d_status) return fail(c, STTS_ERR_BAD_ARG, "stts_frame_index_static: null pointer");
  if (n_seg <= 0 || n_seg > frame_index_static_max_seg())
    return fail(c, STTS_ERR_BAD_ARG, "stts_frame_index_static: n_seg = " + std::to_string(n_seg) + " is outside 1 .. " +
                                         std::to_string(frame_index_static_max_seg()) +
                                         " (the static scratch array holds n_seg + 1 <= 1025 int32 frame offsets)");
  DeviceGuard g(c);
  STTS_CK(frame_index_static(c.device, d_dur, d_cu_tok, n_seg, d_idx, idx_cap, d_status, S(stream)));
  c.launches++;
  return STTS_OK;
}

int stts_source_sinegen(stts_ctx* ctx, const float* d_f0, const int* len, int n_seg, const float* d_rand_ini,
                        const float* d_noise, const float* merge_weight, int harmonic_num, float samp_rate,
                        float sine_amp, float noise_std, float voiced_threshold, float* d_out, float* d_phase,
                        void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (!d_f0 || !len || !d_rand_ini || !d_noise || !merge_weight || !d_out)
    return fail(c, STTS_ERR_BAD_ARG, "stts_source_sinegen: null pointer");
  DeviceGuard g(c);
  return run_sinegen(c, d_f0, len, n_seg, d_rand_ini, d_noise, merge_weight, harmonic_num, samp_rate, sine_amp, noise_std,
                     voiced_threshold, d_out, d_phase, S(stream));
}

int stts_generator_block_snake(stts_ctx* ctx, const char* module, int channels, int kernel_size, const int* dilation,
                               int n_seg, const int* len, const float* d_x, const float* d_style, float* d_out,
                               void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (!module || !dilation || !len || !d_x || !d_style || !d_out)
    return fail(c, STTS_ERR_BAD_ARG, "stts_generator_block_snake: null pointer");
  DeviceGuard g(c);
  return run_snake_block(c, module, channels, kernel_size, dilation, n_seg, len, d_x, d_style, d_out, S(stream));
}

int stts_stft_conv_transform(stts_ctx* ctx, const char* module, int filter_length, int hop_length, int win_length,
                             int pad_mode, int n_seg, const int* len, const float* d_wave, float* d_mag, float* d_x,
                             float* d_y, void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (!module || !len || !d_wave || !d_mag || !d_x || !d_y) return fail(c, STTS_ERR_BAD_ARG, "stts_stft_conv_transform: null pointer");
  DeviceGuard g(c);
  return run_stft_conv_transform(c, module, filter_length, hop_length, win_length, pad_mode, n_seg, len, d_wave, d_mag, d_x,
                                 d_y, S(stream));
}

int stts_stft_conv_inverse(stts_ctx* ctx, const char* module, int filter_length, int hop_length, int win_length, int n_seg,
                           const int* n_frames, const int* out_len, const float* d_mag, const float* d_x, const float* d_y,
                           float* d_wave, void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (!module || !n_frames || !d_wave || !d_mag || !d_x || !d_y) return fail(c, STTS_ERR_BAD_ARG, "stts_stft_conv_inverse: null pointer");
  DeviceGuard g(c);
  return run_stft_conv_inverse(c, module, filter_length, hop_length, win_length, n_seg, n_frames, out_len, d_mag, d_x, d_y,
                               d_wave, S(stream));
}

int stts_pe_towers(stts_ctx* ctx, const float* d_x, const float* d_style, float* d_pitch, float* d_energy,
                   void* stream) {
  if (!ctx || !d_style || !d_pitch || !d_energy) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_pe_towers(ctx->c, d_x, d_style, d_pitch, d_energy, S(stream));
}

int stts_decoder(stts_ctx* ctx, const float* d_asr, const float* d_f0, const float* d_n, const float* d_style,
                 float* d_out, void* stream) {
  if (!ctx || !d_asr || !d_f0 || !d_n || !d_style || !d_out) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_decoder(ctx->c, d_asr, d_f0, d_n, d_style, d_out, S(stream));
}

int stts_prior_flow(stts_ctx* ctx, const float* d_x, const float* d_noise, const float* d_style, float* d_mel,
                    float* d_z_prior, float* d_z_flow, void* stream) {
  if (!ctx || !d_x || !d_noise || !d_style || !d_mel) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_prior_flow(ctx->c, d_x, d_noise, d_style, d_mel, d_z_prior, d_z_flow, S(stream));
}

int stts_source_pcph(stts_ctx* ctx, const float* d_pitch4, const float* d_noise, const float* d_init_phase,
                     float* d_out, void* stream) {
  if (!ctx || !d_pitch4 || !d_noise || !d_init_phase || !d_out) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_source(ctx->c, d_pitch4, d_noise, d_init_phase, d_out, S(stream));
}

int stts_generator(stts_ctx* ctx, const float* d_mel, const float* d_style, const float* d_pitch4,
                   const float* d_noise, const float* d_init_phase, float* d_audio, float* d_logamp, float* d_phase,
                   void* stream) {
  if (!ctx || !d_mel || !d_style || !d_pitch4 || !d_noise || !d_init_phase || !d_audio) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_generator(ctx->c, d_mel, d_style, d_pitch4, d_noise, d_init_phase, d_audio, d_logamp, d_phase, S(stream));
}

int stts_generator_from_stft(stts_ctx* ctx, const float* d_mel, const float* d_style, const float* d_har_spec,
                             const float* d_har_phase, float* d_audio, float* d_logamp, float* d_phase, void* stream) {
  if (!ctx || !d_mel || !d_style || !d_har_spec || !d_har_phase || !d_audio) return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_generator_from_stft(ctx->c, d_mel, d_style, d_har_spec, d_har_phase, d_audio, d_logamp, d_phase,
                                 S(stream));
}

int stts_speech_frames(stts_ctx* ctx, const float* d_enc_tok, const int32_t* d_cu_tok, const int32_t* d_idx,
                       const int32_t* d_cu_frames, const float* d_pitch, const float* d_energy, const float* d_style,
                       const float* d_noise_prior, const float* d_noise_source, const float* d_init_phase,
                       float* d_audio, float* d_logamp, float* d_phase, void* stream) {
  if (!ctx || !d_enc_tok || !d_cu_tok || !d_idx || !d_cu_frames || !d_pitch || !d_energy || !d_style ||
      !d_noise_prior || !d_noise_source || !d_init_phase || !d_audio)
    return STTS_ERR_BAD_ARG;
  DeviceGuard g(ctx->c);
  return run_speech_frames(ctx->c, d_enc_tok, d_cu_tok, d_idx, d_cu_frames, d_pitch, d_energy, d_style, d_noise_prior,
                           d_noise_source, d_init_phase, d_audio, d_logamp, d_phase, S(stream));
}

int stts_check_source_status(stts_ctx* ctx, void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (c.raw.count("META") == 0 || c.h_meta == nullptr) return fail(c, -4, "no layout set");
  DeviceGuard g(c);
  const int B = c.L4.n_seg;
  cudaError_t e = cudaMemcpyAsync(c.h_meta, c.raw["META"], (size_t)B * sizeof(SourceMeta), cudaMemcpyDeviceToHost,
                                  S(stream));
  if (e == cudaSuccess) e = cudaStreamSynchronize(S(stream));
  if (e != cudaSuccess) return cuda_fail(c, e, "source status read");
  for (int b = 0; b < B; ++b)
    if (c.h_meta[b].error)
      return fail(c, STTS_ERR_UNVOICED_MIN,
                  "utterance " + std::to_string(b) + ": voiced frames but no f0 > 20 Hz (reference raises in torch.min)");
  return STTS_OK;
}

int64_t stts_saturation_count(stts_ctx* ctx, void* stream) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (c.raw.count("STATUS") == 0) return fail(c, -4, "no layout set");
  DeviceGuard g(c);
  unsigned long long* cnt = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c.raw["STATUS"]) + 128);
  cudaError_t e = cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), S(stream));
  if (e != cudaSuccess) return cuda_fail(c, e, "saturation counter reset");
  for (auto* acts : {&c.act, &c.tact})
    for (auto& kv : *acts) {
      const Act& a = kv.second;
      if (a.h != nullptr) STTS_CK(saturation_count(a.h, (size_t)a.rows * a.ld, cnt, S(stream)));
      if (a.g != nullptr) STTS_CK(saturation_count(a.g, (size_t)a.rows * a.ldg, cnt, S(stream)));
    }
  unsigned long long h = 0;
  e = cudaMemcpyAsync(&h, cnt, sizeof(h), cudaMemcpyDeviceToHost, S(stream));
  if (e == cudaSuccess) e = cudaStreamSynchronize(S(stream));
  if (e != cudaSuccess) return cuda_fail(c, e, "saturation counter read");
  return (int64_t)h;
}

int stts_pcm16(stts_ctx* ctx, const float* d_audio, int64_t n, int16_t* d_pcm, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!d_audio || !d_pcm))) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  DeviceGuard g(c);
  STTS_CK(pcm16(d_audio, d_pcm, (long long)n, S(stream)));
  return STTS_OK;
}

int stts_test_conv(stts_ctx* ctx, int backend, const float* d_x, int ldx, int cin, int taps, int pad,
                   const float* h_w, const float* h_bias, int n_out, float* d_y, int ldy, void* stream) {
  if (!ctx || !d_x || !h_w || !d_y || backend < 0 || backend > 2) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  if (c.L4.total() == 0) return fail(c, -4, "no hop-75 layout set");
  DeviceGuard g(c);
  PackedGemm pg = pack_from_conv_weight(c, h_w, h_bias, n_out, cin, taps, pad);
  if (pg.cin_pad[0] > ldx) { free_pack(pg); return fail(c, -2, "ldx must cover cin padded to 64"); }
  Act a;
  a.f = const_cast<float*>(d_x);
  a.ld = ldx;
  a.rows = c.L4.R;
  // 16-bit backend: build a temporary twin of the operand
  twin_t* twin = nullptr;
  if (backend == 1) {
    if (cudaMalloc(&twin, (size_t)c.L4.R * ldx * 2) != cudaSuccess) { free_pack(pg); return fail(c, -5, "twin alloc"); }
    std::vector<float> hx((size_t)c.L4.R * ldx);
    cudaMemcpy(hx.data(), d_x, hx.size() * 4, cudaMemcpyDeviceToHost);
    std::vector<twin_t> hh(hx.size());
    for (size_t i = 0; i < hx.size(); ++i) hh[i] = to_twin(hx[i]);
    cudaMemcpy(twin, hh.data(), hh.size() * 2, cudaMemcpyHostToDevice);
    a.h = twin;
  }
  Epi e;
  std::memset(&e, 0, sizeof(e));
  e.kind = EPI_BIAS;
  e.out = d_y; e.ldo = ldy; e.col0 = 0;
  e.scale = 1.f;
  int r = run_gemm(c, pg, GC_BODY, &a, 0, nullptr, 0, e, c.L4.map(), S(stream), backend);
  cudaError_t ce = cudaStreamSynchronize(S(stream));
  if (twin) cudaFree(twin);
  free_pack(pg);
  tc_map_cache_clear();
  if (r != 0) return r;
  if (ce != cudaSuccess) return cuda_fail(c, ce, "test_conv");
  return STTS_OK;
}

int64_t stts_debug_read(stts_ctx* ctx, const char* name, float* d_dst, int64_t cap, int* rows, int* ld,
                        void* stream) {
  if (!ctx || !name) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  DeviceGuard g(c);
  auto it = c.act.find(name);
  const float* src = nullptr;
  int64_t n = 0;
  if (it != c.act.end()) {
    src = it->second.f;
    n = (int64_t)it->second.rows * it->second.ld;
    if (rows) *rows = it->second.rows;
    if (ld) *ld = it->second.ld;
  } else {
    auto jt = c.raw.find(name);
    if (jt == c.raw.end()) return fail(c, -1, std::string("unknown buffer ") + name);
    src = reinterpret_cast<const float*>(jt->second);
    n = (int64_t)(c.raw_bytes[name] / 4);
    if (rows) *rows = (int)n;
    if (ld) *ld = 1;
  }
  if (d_dst == nullptr) return n;
  if (cap < n) return fail(c, -2, "debug_read: destination too small");
  cudaError_t e = cudaMemcpyAsync(d_dst, src, (size_t)n * 4, cudaMemcpyDeviceToDevice, S(stream));
  if (e != cudaSuccess) return cuda_fail(c, e, "debug_read");
  return n;
}

int stts_layout_info(const stts_ctx* ctx, int rate, int* rows, int* seg_start) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  const Layout& L = (rate == 4) ? ctx->c.L4 : ctx->c.LF;
  if (rows) *rows = L.R;
  if (seg_start)
    for (int b = 0; b < L.n_seg; ++b) seg_start[b] = L.seg_start[b];
  return L.n_seg;
}

int stts_set_profiling(stts_ctx* ctx, int on) {
  if (!ctx) return STTS_ERR_BAD_ARG;
  Context& c = ctx->c;
  DeviceGuard g(c);
  cudaDeviceSynchronize();
  for (auto& r : c.prof) { c.ev_pool.push_back(r.e0); c.ev_pool.push_back(r.e1); }
  c.prof.clear();
  c.prof_on = on != 0;
  return STTS_OK;
}

const char* stts_profile_json(stts_ctx* ctx) {
  if (!ctx) return "[]";
  Context& c = ctx->c;
  DeviceGuard g(c);
  cudaDeviceSynchronize();
  struct Agg { double ms = 0, flops = 0; long n = 0; int backend = 0; };
  std::map<std::string, Agg> agg;
  for (auto& r : c.prof) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    Agg& a = agg[r.name + "#" + std::to_string(r.backend)];
    a.ms += ms; a.flops += r.flops; a.n++; a.backend = r.backend;
    c.ev_pool.push_back(r.e0);
    c.ev_pool.push_back(r.e1);
  }
  c.prof.clear();
  static thread_local std::string out;
  out = "[";
  bool first = true;
  for (auto& kv : agg) {
    char buf[512];
    const std::string nm = kv.first.substr(0, kv.first.find('#'));
    snprintf(buf, sizeof(buf), "%s{\"name\":\"%s\",\"backend\":%d,\"launches\":%ld,\"ms\":%.6f,\"flops\":%.1f}",
             first ? "" : ",", nm.c_str(), kv.second.backend, kv.second.n, kv.second.ms, kv.second.flops);
    out += buf;
    first = false;
  }
  out += "]";
  return out.c_str();
}

int64_t stts_launch_count(const stts_ctx* ctx) { return ctx ? ctx->c.launches : 0; }

}  // extern "C"
