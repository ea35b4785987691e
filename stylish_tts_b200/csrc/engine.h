// Context of the synthesis engine: loaded state_dict tensors, prepacked device weights,
// packed-layout descriptors and the workspace arena.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.cuh"
#include "kernels.h"

namespace stts {

struct HostTensor {
  std::vector<int64_t> shape;
  std::vector<float> data;
  int64_t dim(int i) const { return i < (int)shape.size() ? shape[i] : 1; }
  int64_t numel() const { int64_t n = 1; for (auto s : shape) n *= s; return n; }
};

// One prepacked dense contraction: W[Npad, Ktot] K-major in three formats.
struct PackedGemm {
  int N = 0, Npad = 0, Ktot = 0;
  int n_kseg = 0;
  int taps[2] = {0, 0}, pad[2] = {0, 0}, cin_pad[2] = {0, 0};
  int dil[2] = {1, 1};            // row step between taps (dilated convs)
  float* Wf = nullptr;            // exact fp32                     [Npad, Ktot]
  float* Whl = nullptr;           // tf32 hi rows then lo rows      [2 * Npad, Ktot]
  __half* Wg = nullptr;           // fp16(W * wscale): one-pass 16-bit GEMMs and hi part of f16x3  [Npad, Ktot]
  __half* Wl = nullptr;           // fp16(W * wscale - Wg)          [Npad, Ktot]
  float wscale = 1.f;             // power of two
  float* bias = nullptr;          // [Npad] (zeros when the layer has no bias)
  std::string name;               // layer name (profiling)
  double k_real = 0;              // un-padded K (algorithmic FLOPs = 2 * rows * N * k_real)
};

struct Act {          // an activation buffer: fp32 + optional 16-bit twin, same ld
  float* f = nullptr;
  twin_t* h = nullptr;
  __half* g = nullptr;          // fp16 split twins g = fp16(s x), l = fp16(s x - g): operands of the
  __half* l = nullptr;          //   island GEMMs (f16x3; g alone for the one-pass fp16 GEMM)
  int ldg = 0;                  // row pitch of g / l (elements)
  float gscale = 1.f;           // s: power-of-two pre-scale of the fp16 twins
  int ld = 0;
  int rows = 0;
};

struct Layout {       // host + device description of one frame rate
  int n_seg = 0, R = 0, n_tiles = 0;
  std::vector<int> seg_start, seg_len, tile_seg, tile_nvalid, tile_nvalid_p1, tile_seg_p1, cu;  // cu: dense offsets
  int* d_tile_seg = nullptr; int* d_tile_nvalid = nullptr; int* d_seg_start = nullptr; int* d_seg_len = nullptr;
  int* d_tile_seg_p1 = nullptr; int* d_tile_nvalid_p1 = nullptr; int* d_seg_len_p1 = nullptr;
  int* d_cu = nullptr;            // [n_seg + 1] dense offsets (frames)
  long long* d_off[4] = {nullptr, nullptr, nullptr, nullptr};  // dense element offsets for C = {see engine}
  int* d_cs = nullptr;            // [n_seg] channel stride (= seg_len)
  int* d_cs_p1 = nullptr;         // [n_seg] seg_len + 1
  RowMap map() const { return RowMap{R, n_tiles, n_seg, d_tile_seg, d_tile_nvalid, d_seg_start, d_seg_len}; }
  RowMap map_p1() const { return RowMap{R, n_tiles, n_seg, d_tile_seg_p1, d_tile_nvalid_p1, d_seg_start, d_seg_len_p1}; }
  int total() const { return cu.empty() ? 0 : cu.back(); }
};

struct SnakeCfg { int C = 0, k = 0, dil[3] = {1, 1, 1}, J = 0; };   // a packed AdaptiveGeneratorBlock (prepack.cu)
struct VariantWork;                                                  // workspace of the section-8 f2 operators (variants.cu)

enum GemmClass { GC_BODY = 0, GC_ISLAND = 1 };  // ISLAND: STFT, output heads, iSTFT stay fp32-grade

struct Context {
  int device = 0;
  int precision = 0;
  int small_shapes = -1;                       // small-batch kernel shapes (stts_set_small_batch_shapes): -1 / 1 on, 0 off
  int pdl = -1;                                // programmatic dependent launch: -1 default (on), 0 off, 1 on
  std::string err;
  int64_t launches = 0;
  std::map<std::string, HostTensor> tensors;   // "module/key"
  bool finalized = false;

  // ---- prepacked weights
  std::map<std::string, PackedGemm> gemm;      // by layer name
  std::map<std::string, float*> vec;           // small device vectors / matrices by name
  std::map<std::string, int> vec_len;
  std::vector<void*> weight_allocs;
  // style-FC concatenations: row offset of each norm inside the stage matrix
  std::map<std::string, int> fc_off;
  int J_dec = 0, J_gen = 0, J_flow = 0, J_pe = 0;

  // ---- layouts + workspace
  Layout L4, LF;
  // token-rate layout + its own arena (stts_set_token_layout)
  Layout LT;
  std::vector<int> tsig;
  char* tarena = nullptr; size_t tarena_bytes = 0;
  std::map<std::string, Act> tact;
  std::map<std::string, void*> traw;
  int max_t = 0;
  std::map<int, float2*> rope_tab;   // rd -> (cos, sin) table [kRopePositions][rd / 2] (token_pipeline.cu)
  std::vector<int> sig;                        // layout signature
  char* arena = nullptr; size_t arena_bytes = 0, arena_used = 0;
  std::map<std::string, Act> act;              // named activation buffers
  std::map<std::string, void*> raw;            // other named scratch
  std::map<std::string, size_t> raw_bytes;
  SourceMeta* h_meta = nullptr;                // pinned
  cudaStream_t st_aux = nullptr;               // second branch of fork / join stages (the energy tower), with its events
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // ---- optional per-GEMM CUDA-event profiling (bench.py roofline)
  bool prof_on = false;
  struct ProfRec { std::string name; cudaEvent_t e0, e1; double flops; int backend; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;

  // ---- SURVEY section 8 f2 operators (SineGenerator, Snake AdaptiveGeneratorBlock): own layout + workspace
  std::map<std::string, SnakeCfg> snake;
  VariantWork* vwork = nullptr;

  bool has(const std::string& k) const { return tensors.count(k) != 0; }
};

// prepack.cu
int finalize_weights(Context& c);
PackedGemm pack_from_conv_weight(Context& c, const float* w, const float* bias, int n_out, int cin, int taps, int pad);
void free_pack(PackedGemm& p);
int pack_snake_block(Context& c, const std::string& module, int channels, int k, const int* dil);

// pipeline.cu
void build_layout(Layout& L, int n_seg, const int* len, int guard = 0);   // guard: zero rows after every utterance (0 = STTS_GUARD)
int set_layout(Context& c, int n_seg, const int* len_f, const int* len_f4);
int run_gemm(Context& c, const PackedGemm& pg, int cls, const Act* a0, int a0_col, const Act* a1, int a1_col,
             const Epi& epi, const RowMap& map, cudaStream_t st, int forced_backend = -1,
             const twin_t* seg_weights = nullptr);   // per-utterance fp16 weights [n_seg * Npad, Ktot]
int run_pe_towers(Context& c, const float* d_x, const float* d_style, float* d_pitch, float* d_energy, cudaStream_t st);
int run_decoder_core(Context& c, const float* d_style, cudaStream_t st);
int run_decoder(Context& c, const float* d_asr, const float* d_f0, const float* d_n, const float* d_style, float* d_out,
                cudaStream_t st);
int run_prior_flow_core(Context& c, const Act& x, const float* d_style, cudaStream_t st);
int run_prior_flow(Context& c, const float* d_x, const float* d_noise, const float* d_style, float* d_mel,
                   float* d_z_prior, float* d_z_flow, cudaStream_t st);
int run_source(Context& c, const float* d_pitch4_dense, const float* d_noise, const float* d_init_phase, float* d_out,
               cudaStream_t st);
int run_generator_core(Context& c, const float* d_style, const float* d_noise, const float* d_init_phase,
                       float* d_audio, float* d_logamp, float* d_phase, cudaStream_t st, bool lean = false);
int run_generator_tail(Context& c, const float* d_style, float* d_audio, float* d_logamp, float* d_phase,
                       cudaStream_t st, bool priors_done = false);
int run_generator_from_stft(Context& c, const float* d_mel, const float* d_style, const float* d_har_spec,
                            const float* d_har_phase, float* d_audio, float* d_logamp, float* d_phase,
                            cudaStream_t st);
int run_generator(Context& c, const float* d_mel, const float* d_style, const float* d_pitch4, const float* d_noise,
                  const float* d_init_phase, float* d_audio, float* d_logamp, float* d_phase, cudaStream_t st);
int run_speech_frames(Context& c, const float* d_enc_tok, const int* d_cu_tok, const int* d_idx,
                      const int* d_cu_frames, const float* d_pitch, const float* d_energy, const float* d_style,
                      const float* d_noise_prior, const float* d_noise_source, const float* d_init_phase,
                      float* d_audio, float* d_logamp, float* d_phase, cudaStream_t st);

// token_pipeline.cu / prepack_tokens.cu
int set_token_layout(Context& c, int n_seg, const int* len_t);
int run_text_encoder(Context& c, const std::string& name, const int* d_tokens, float* d_mu_tok, float* d_x_tok,
                     cudaStream_t st);

int run_prosody_encoder(Context& c, const std::string& name, const float* d_enc_tok, const float* d_style,
                        float* d_out_tok, float* d_logits_tok, cudaStream_t st);

int run_style_encoder(Context& c, const std::string& name, const float* d_enc_tok, float* d_style_out,
                      cudaStream_t st);
int run_pe_cross(Context& c, const float* d_prosody_tok, const float* d_style, const int* d_idx, const int* d_cu_frames,
                 float* d_x_out, cudaStream_t st);

// variants.cu
int run_sinegen(Context& c, const float* d_f0, const int* h_len, int n_seg, const float* d_rand_ini, const float* d_noise,
                const float* h_merge_w, int harmonic_num, float samp_rate, float sine_amp, float noise_std,
                float voiced_threshold, float* d_out, float* d_phase, cudaStream_t st);
int run_snake_block(Context& c, const std::string& module, int channels, int k, const int* dil, int n_seg,
                    const int* h_len, const float* d_x, const float* d_style, float* d_out, cudaStream_t st);
int run_stft_conv_transform(Context& c, const std::string& module, int n_fft, int hop, int win, int pad_mode, int n_seg,
                            const int* h_len, const float* d_wave, float* d_mag, float* d_x, float* d_y, cudaStream_t st);
int run_stft_conv_inverse(Context& c, const std::string& module, int n_fft, int hop, int win, int n_seg,
                          const int* h_frames, const int* h_out_len, const float* d_mag, const float* d_x, const float* d_y,
                          float* d_wave, cudaStream_t st);
void free_variant_work(Context& c);

int fail(Context& c, int code, const std::string& msg);
int cuda_fail(Context& c, cudaError_t e, const char* where);

#define STTS_CK(call)                                                  \
  do {                                                                 \
    cudaError_t _e = (call);                                           \
    if (_e != cudaSuccess) return cuda_fail(c, _e, #call);             \
    c.launches++;                                                      \
  } while (0)
#define STTS_RET(call)                \
  do {                                \
    int _r = (call);                  \
    if (_r != 0) return _r;           \
  } while (0)

}  // namespace stts
