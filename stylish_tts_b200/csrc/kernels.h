// Internal launcher declarations (one per kernel family).  Every launcher enqueues on the
// caller's stream, never synchronises and never allocates.
#pragma once
#include "common.cuh"

namespace stts {

// ---- conv-as-GEMM (gemm_simt.cu / gemm_tc.cu)
cudaError_t conv_gemm_simt(const ConvGemm& g, const RowMap& map, cudaStream_t st);
cudaError_t conv_gemm_tc_f16(const ConvGemm& g, const RowMap& map, cudaStream_t st);
cudaError_t conv_gemm_tc_tf32x3(const ConvGemm& g, const RowMap& map, cudaStream_t st);
cudaError_t conv_gemm_tc_f16x3(const ConvGemm& g, const RowMap& map, cudaStream_t st);
cudaError_t wn_layer_fused_f16(const ConvGemm& gin, const ConvGemm& grs, const RowMap& map, cudaStream_t st);
void tc_map_cache_clear();

// ---- elementwise.cu
cudaError_t colstats(const float* x, int ld, int C, const RowMap& map, int mode, float eps, double* scratch,
                     float* out0, float* out1, cudaStream_t st);
cudaError_t colstats_partial_q(const float* x, int ld, int col0, int ncols, const RowMap& map, float* psum, float* psq,
                               int pld, cudaStream_t st);
cudaError_t colstats_final_q(const float* psum, const float* psq, int pld, int C, const RowMap& map, float eps,
                             float* mean, float* rstd, cudaStream_t st);
cudaError_t style_fc(const float* W, const float* bias, const float* s, int J, int K, int B, float* out,
                     cudaStream_t st);
cudaError_t adain_act(const float* x, int ldx, int C, const RowMap& map, const float* mean, const float* rstd,
                      const float* gb, int ldgb, float slope, float* y, int ldy, twin_t* yh, int ldyh,
                      cudaStream_t st);
cudaError_t expand_tokens(const float* tok, int C, const int* idx, const int* cu_f, const int* tok_off, int div,
                          const RowMap& map, float* out, int ldo, int col0, twin_t* outh, int ldh, int colh0,
                          cudaStream_t st);
cudaError_t upsample4(const float* a, const float* b, const int* cu_f, const RowMap& map, float* a4, float* b4,
                      cudaStream_t st);
struct F0NDst {  // up to three destinations for the (F0, N) channel pair
  float* f[3]; int ld[3]; int col[3];
  twin_t* h[3]; int ldh[3];
};
cudaError_t f0n_conv(const float* p4, const float* e4, const float* w, const RowMap& map, const F0NDst& d,
                     cudaStream_t st);
cudaError_t transpose_in(const float* src, const long long* off, const int* cs, int C, const RowMap& map, float* dst,
                         int ld, int col0, twin_t* dsth, int ldh, int colh0, cudaStream_t st);
cudaError_t transpose_out(const float* src, int ld, int col0, int C, const RowMap& map, float* dst,
                          const long long* off, const int* cs, int extra, cudaStream_t st);
cudaError_t rowvec_copy(float* rowv, float* dense, const int* cu, const RowMap& map, int to_rows, cudaStream_t st);
cudaError_t dwconv_adaln(const float* x, int ldx, const RowMap& map, int K, const float* wt, const float* bias,
                         const float* gb, int ldgb, float eps, float* y, int ldy, twin_t* yh, int ldyh,
                         cudaStream_t st);
cudaError_t adaln2(const float* x, int ldx, const RowMap& map, const float* gb1, const float* gb2, int ldgb, float eps,
                   float* y1, float* y2, int ldy, __half* y1g, __half* y2g, int ldyh, __half* y1l, __half* y2l,
                   cudaStream_t st);
cudaError_t grn_apply(float* x, int ldx, int C, const RowMap& map, const float* gx, const float* gamma,
                      const float* beta, float* scale_scratch, twin_t* xh, int ldxh, cudaStream_t st,
                      __half* xg = nullptr, __half* xl = nullptr);
cudaError_t grn_scale_partials(const float* colss, int ld, int C, const RowMap& map, const float* gamma, float* gx,
                               float* scale, cudaStream_t st);
cudaError_t grn_weights(const float* W, int N, int K, const float* scale, int n_seg, twin_t* Wu, cudaStream_t st);
cudaError_t rowdot(const float* x, int ldx, int C, const RowMap& map, const float* w, const float* bias, const int* cu,
                   float* dense, cudaStream_t st);
cudaError_t fill(float* p, size_t n, float v, cudaStream_t st);
cudaError_t saturation_count(const twin_t* h, size_t n, unsigned long long* count, cudaStream_t st);
cudaError_t twin_refresh(const float* f, int ld, int C, const RowMap& map, __half* h, __half* l, cudaStream_t st);

// ---- token_rate.cu
cudaError_t embed_tokens(const int* tok, const int* cu_tok, const float* table, int C, int n_vocab, const RowMap& map,
                         float* x, int ldx, cudaStream_t st, __half* xg = nullptr, __half* xl = nullptr);
cudaError_t row_ln(const float* x, int ldx, int C, const RowMap& map, const float* gamma, const float* beta,
                   int seg_stride, float eps, int relu, float* y, int ldy, cudaStream_t st, __half* yg = nullptr,
                   __half* yl = nullptr);
cudaError_t rope_inplace(float* x, int ldx, int col0, int blk, int n_blocks, int H, int D, int rd, const RowMap& map,
                         cudaStream_t st);
cudaError_t attention(const float* Q, int ldq, int qcol, const float* K, int ldk, int kcol, const float* V, int ldv,
                      int vcol, int H, int D, float scale, int rd, const float2* rope_tab, const RowMap& qmap,
                      const RowMap& kmap, int max_tq, int max_tk, const int* q_tok, const int* q_cu, int band, float* out,
                      int ldo, cudaStream_t st, __half* outg = nullptr, __half* outl = nullptr, int ldh = 0);
cudaError_t rope_table(float2* tab, int n_pos, int rd, cudaStream_t st);
cudaError_t dense_to_rows(const float* src, int C, const RowMap& map, const int* cu, float* dst, int ld, int col0,
                          cudaStream_t st, __half* dg = nullptr, __half* dl = nullptr);
cudaError_t seg_broadcast(const float* vec, int C, const RowMap& map, float* dst, int ld, int col0, cudaStream_t st,
                          __half* dg = nullptr, __half* dl = nullptr);
cudaError_t dwconv_rows(const float* x, int ldx, int C, int K, const float* w, const float* b, const RowMap& map,
                        int act, float* y, int ldy, cudaStream_t st, __half* yg = nullptr, __half* yl = nullptr);
cudaError_t rows_to_dense(const float* src, int ld, int C, const RowMap& map, const int* cu, float* dst,
                          cudaStream_t st);

// ---- length_reg.cu
cudaError_t durations_from_logits(const float* logits, int n_tokens, int* dur, cudaStream_t st);
cudaError_t frame_index(const int* dur, const int* cu_tok, int n_seg, int* cu_frames, int* idx, int idx_cap,
                        int* status, cudaStream_t st);
// allocation-free variant: cu_frames lives in a file-scope __device__ array (length_reg.cu); n_seg <= frame_index_static_max_seg()
int frame_index_static_max_seg();
cudaError_t frame_index_static(int device, const int* dur, const int* cu_tok, int n_seg, int* idx, int idx_cap,
                               int* status, cudaStream_t st);

// ---- source.cu
struct SourceMeta { int any_voiced; int max_h; int error; int pad; };
cudaError_t pcph_prefix(const float* p4, const RowMap& map, double* prefix, SourceMeta* meta, cudaStream_t st);
// src has (R + 16) rows of `ld` floats: logical hop row h of a segment lives at row seg_start + h + 8
cudaError_t pcph_source(const float* p4, const double* prefix, const SourceMeta* meta, const float* noise,
                        const float* init_phase, const int* cu4, const RowMap& map, float* src, int ld,
                        __half* srcg, __half* srcl, int ldg, float gscale, float* dense_out, cudaStream_t st);

// ---- stft_fft.cu
cudaError_t stft_twiddles(double2* tw, double2* tw2, cudaStream_t st);
// forward STFT of the hop-row source buffer -> |X| / atan2 phase rows [R, ldo] (fp32 masters and / or fp16 split twins)
cudaError_t stft_fft(const float* src, int lds, const float* window, const double2* tw, const double2* tw2,
                     const RowMap& map, float* mag, float* pha, int ldo, __half* magg, __half* magl, __half* phag,
                     __half* phal, int ldh, cudaStream_t st);

// ---- istft.cu
cudaError_t polar_spec(const float* logamp, const float* phase, int ld, int n_bins, const RowMap& map, float* spec,
                       int lds, __half* specg, int ldsh, __half* specl, float gscale, cudaStream_t st);
cudaError_t ola_tanh(const float* frames, int ldf, const float* wsq, const int* cu4, const RowMap& map, float* audio,
                     cudaStream_t st);
cudaError_t pcm16(const float* x, int16_t* y, long long n, cudaStream_t st);

}  // namespace stts
