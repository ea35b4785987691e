// HBM-bound kernels of the synthesis path: statistics, normalisation, activations, gathers,
// transposes at the module boundary.  All operate on the packed [R, ld] layout (common.cuh);
// guard rows and pad columns are never written, so they stay zero.
#include "common.cuh"
#include "kernels.h"

namespace stts {

// ============================================================================ column statistics
// AdaIN needs per-(utterance, channel) mean / variance over time (ada_norm.py:132,
// nn.InstanceNorm1d eps 1e-5, biased variance); GRN needs the L2 norm over time
// (generator.py:496-498).  Two deterministic stages (no float atomics): per-tile partial sums in
// fp64, then a per-segment reduction in tile order.
// block = 128 channels x 4 row groups (rows lr = g, g + 4, ...): a quarter of the serial chain of one thread per
// channel; the four fp64 partials are combined in fixed order
__global__ void __launch_bounds__(512) colstats_partial_kernel(const float* __restrict__ x, int ld, int C,
                                                               RowMap map, double* __restrict__ psum,
                                                               double* __restrict__ psq) {
  pdl_enter();
  __shared__ double sh_s[4][128], sh_q[4][128];
  const int tile = blockIdx.x;
  if (map.tile_seg[tile] < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int c = blockIdx.y * 128 + threadIdx.x;
  const int g = threadIdx.y;
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float* p = x + (size_t)tile * 128 * ld + c;
    int i = g;
    for (; i + 28 < nvalid; i += 32) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = p[(size_t)(i + 4 * j) * ld];
#pragma unroll
      for (int j = 0; j < 8; ++j) { const double d = v[j]; s += d; q = fma(d, d, q); }
    }
    for (; i < nvalid; i += 4) { const double d = p[(size_t)i * ld]; s += d; q = fma(d, d, q); }
  }
  sh_s[g][threadIdx.x] = s;
  sh_q[g][threadIdx.x] = q;
  __syncthreads();
  if (g == 0 && c < C) {
    for (int j = 1; j < 4; ++j) { s += sh_s[j][threadIdx.x]; q += sh_q[j][threadIdx.x]; }
    psum[(size_t)tile * C + c] = s;
    psq[(size_t)tile * C + c] = q;
  }
}

// mode 0: out0 = mean, out1 = 1/sqrt(var + eps)   mode 1: out0 = sqrt(sum x^2)
__global__ void __launch_bounds__(128) colstats_final_kernel(const double* __restrict__ psum,
                                                             const double* __restrict__ psq, int C, RowMap map,
                                                             int mode, float eps, float* __restrict__ out0,
                                                             float* __restrict__ out1) {
  pdl_enter();
  const int seg = blockIdx.x;
  const int c = blockIdx.y * 128 + threadIdx.x;
  if (c >= C) return;
  const int t0 = map.seg_start[seg] / 128;
  const int len = map.seg_len[seg];
  const int nt = (len + 127) / 128;
  double s = 0.0, q = 0.0;
  for (int t = 0; t < nt; ++t) { s += psum[(size_t)(t0 + t) * C + c]; q += psq[(size_t)(t0 + t) * C + c]; }
  if (mode == 0) {
    const double mean = s / len;
    double var = q / len - mean * mean;
    if (var < 0.0) var = 0.0;
    out0[(size_t)seg * C + c] = (float)mean;
    out1[(size_t)seg * C + c] = (float)(1.0 / sqrt(var + (double)eps));
  } else {
    out0[(size_t)seg * C + c] = (float)sqrt(q);
  }
}

cudaError_t colstats(const float* x, int ld, int C, const RowMap& map, int mode, float eps, double* scratch,
                     float* out0, float* out1, cudaStream_t st) {
  double* psum = scratch;
  double* psq = scratch + (size_t)map.n_tiles * C;
  dim3 g1(map.n_tiles, (C + 127) / 128);
  launch_k(colstats_partial_kernel, g1, dim3(128, 4), 0, st, x, ld, C, map, psum, psq);
  dim3 g2(map.n_seg, (C + 127) / 128);
  launch_k(colstats_final_kernel, g2, 128, 0, st, psum, psq, C, map, mode, eps, out0, out1);
  return cudaGetLastError();
}

// Quarter-tile form of the same statistics, fed by the GEMM epilogues (Epi::colsum / colsq: fp32 sums over the
// 32 rows of a (tile, quarter), 16-bit mode): colstats_partial_q fills the columns no GEMM produces (the
// asr_res / F0 / N side channels of the decoder's concatenated input), colstats_final_q reduces in fp64.
__global__ void __launch_bounds__(128) colstats_partial_q_kernel(const float* __restrict__ x, int ld, int col0, int ncols,
                                                                 RowMap map, float* __restrict__ psum,
                                                                 float* __restrict__ psq, int pld) {
  pdl_enter();
  const int tile = blockIdx.x, q = blockIdx.y;
  if (map.tile_seg[tile] < 0) return;
  int nrows = map.tile_nvalid[tile] - q * 32;
  nrows = nrows < 0 ? 0 : (nrows > 32 ? 32 : nrows);
  for (int c = threadIdx.x; c < ncols; c += blockDim.x) {
    const float* p = x + (size_t)(tile * 128 + q * 32) * ld + col0 + c;
    float s = 0.f, sq = 0.f;
    for (int i = 0; i < nrows; ++i) { const float v = p[(size_t)i * ld]; s += v; sq = fmaf(v, v, sq); }
    psum[(size_t)(tile * 4 + q) * pld + col0 + c] = s;
    psq[(size_t)(tile * 4 + q) * pld + col0 + c] = sq;
  }
}

// block = 32 channels x 8 slices of the utterance's quarters (a long utterance has hundreds of quarters: one
// thread per channel walking all of them is a serial chain of L2 latencies); fixed-order fp64 combine
__global__ void __launch_bounds__(256) colstats_final_q_kernel(const float* __restrict__ psum,
                                                               const float* __restrict__ psq, int pld, int C, RowMap map,
                                                               float eps, float* __restrict__ mean,
                                                               float* __restrict__ rstd) {
  pdl_enter();
  __shared__ double sh_s[8][32], sh_q[8][32];
  const int seg = blockIdx.x;
  const int c = blockIdx.y * 32 + threadIdx.x;
  const int sl = threadIdx.y;
  const int t0 = map.seg_start[seg] / 128;
  const int len = map.seg_len[seg];
  const int nq = (len + 31) / 32;              // quarters that hold valid rows (in tile order)
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float* ps = psum + (size_t)t0 * 4 * pld + c;
    const float* pq = psq + (size_t)t0 * 4 * pld + c;
    int k = sl;
    for (; k + 24 < nq; k += 32) {             // four independent loads in flight per array
      const float a0 = ps[(size_t)k * pld], a1 = ps[(size_t)(k + 8) * pld], a2 = ps[(size_t)(k + 16) * pld],
                  a3 = ps[(size_t)(k + 24) * pld];
      const float b0 = pq[(size_t)k * pld], b1 = pq[(size_t)(k + 8) * pld], b2 = pq[(size_t)(k + 16) * pld],
                  b3 = pq[(size_t)(k + 24) * pld];
      s += ((double)a0 + (double)a1) + ((double)a2 + (double)a3);
      q += ((double)b0 + (double)b1) + ((double)b2 + (double)b3);
    }
    for (; k < nq; k += 8) { s += (double)ps[(size_t)k * pld]; q += (double)pq[(size_t)k * pld]; }
  }
  sh_s[sl][threadIdx.x] = s;
  sh_q[sl][threadIdx.x] = q;
  __syncthreads();
  if (sl == 0 && c < C) {
#pragma unroll
    for (int j = 1; j < 8; ++j) { s += sh_s[j][threadIdx.x]; q += sh_q[j][threadIdx.x]; }
    const double m = s / len;
    double var = q / len - m * m;
    if (var < 0.0) var = 0.0;
    mean[(size_t)seg * C + c] = (float)m;
    rstd[(size_t)seg * C + c] = (float)(1.0 / sqrt(var + (double)eps));
  }
}

cudaError_t colstats_partial_q(const float* x, int ld, int col0, int ncols, const RowMap& map, float* psum, float* psq,
                               int pld, cudaStream_t st) {
  dim3 g(map.n_tiles, 4);
  launch_k(colstats_partial_q_kernel, g, 128, 0, st, x, ld, col0, ncols, map, psum, psq, pld);
  return cudaGetLastError();
}

cudaError_t colstats_final_q(const float* psum, const float* psq, int pld, int C, const RowMap& map, float eps,
                             float* mean, float* rstd, cudaStream_t st) {
  dim3 g(map.n_seg, (C + 31) / 32), b(32, 8);
  launch_k(colstats_final_q_kernel, g, b, 0, st, psum, psq, pld, C, map, eps, mean, rstd);
  return cudaGetLastError();
}

// ============================================================================ style FCs
// out[b, j] = bias[j] + sum_k W[j, k] * s[b, k]  -- all AdaIN / AdaLN / WN-cond Linear(style)
// layers of a stage concatenated along j (ada_norm.py:133,190; flow.py:39-40).
// grid (J / 128, B / 4): four utterances per CTA (the 64-FMA dot products of one thread are a serial chain; one CTA
// per 128 outputs looping over the whole batch took 27 us for 32 utterances)
constexpr int kFcBatch = 4;
__global__ void __launch_bounds__(128) style_fc_kernel(const float* __restrict__ W, const float* __restrict__ bias,
                                                       const float* __restrict__ s, int J, int K, int B,
                                                       float* __restrict__ out) {
  pdl_enter();
  extern __shared__ float sh[];  // W tile [128][K + 1] (coalesced load, conflict-free rows) + s [kFcBatch, K]
  float* sw = sh;
  float* ss = sh + 128 * (K + 1);
  const int j0 = blockIdx.x * 128;
  const int b0 = blockIdx.y * kFcBatch;
  const int nb = (B - b0 < kFcBatch) ? (B - b0) : kFcBatch;
  const int nj = (J - j0 < 128) ? (J - j0) : 128;
  for (int i = threadIdx.x; i < nj * K; i += 128) {
    const int row = i / K, col = i - row * K;
    sw[row * (K + 1) + col] = W[(size_t)j0 * K + i];
  }
  for (int i = threadIdx.x; i < nb * K; i += 128) ss[i] = s[(size_t)b0 * K + i];
  __syncthreads();
  const int j = j0 + threadIdx.x;
  if (j >= J) return;
  const float* w = sw + threadIdx.x * (K + 1);
  const float bj = bias[j];
  float acc[kFcBatch];
#pragma unroll
  for (int b = 0; b < kFcBatch; ++b) acc[b] = 0.f;
  for (int k = 0; k < K; ++k) {
    const float wk = w[k];
#pragma unroll
    for (int b = 0; b < kFcBatch; ++b) acc[b] = fmaf(wk, ss[b * K + k], acc[b]);   // rows >= nb: stale, never stored
  }
#pragma unroll
  for (int b = 0; b < kFcBatch; ++b)
    if (b < nb) out[(size_t)(b0 + b) * J + j] = acc[b] + bj;
}

cudaError_t style_fc(const float* W, const float* bias, const float* s, int J, int K, int B, float* out,
                     cudaStream_t st) {
  const size_t smem = (size_t)(128 * (K + 1) + kFcBatch * K) * sizeof(float);
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  dim3 grid((J + 127) / 128, (B + kFcBatch - 1) / kFcBatch);
  launch_k(style_fc_kernel, grid, 128, smem, st, W, bias, s, J, K, B, out);
  return cudaGetLastError();
}

// ============================================================================ AdaIN + LeakyReLU
// y = leaky_relu((1 + gamma) * (x - mean) * rstd + beta, 0.2)      ada_norm.py:135-139,170-176
// gb is the style-FC output row of this norm: gamma = gb[seg, 0:C], beta = gb[seg, C:2C].
// grid (tile, channel chunk of 128); thread = (row group 0..3, channel pair 0..63): every warp access
// is 256 contiguous bytes of one row; per-channel constants live in registers.  C is even.
__global__ void __launch_bounds__(256) adain_act_kernel(const float* __restrict__ x, int ldx, int C, RowMap map,
                                                        const float* __restrict__ mean, const float* __restrict__ rstd,
                                                        const float* __restrict__ gb, int ldgb, float slope,
                                                        float* __restrict__ y, int ldy, twin_t* __restrict__ yh,
                                                        int ldyh) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  // thread = (row group 0..7, 4 channels): a warp access is 512 contiguous bytes of one row (16-byte loads, 8-byte twin
  // stores); C is even, so a row's tail is either a full quad or one pair
  const int c = blockIdx.y * 128 + 4 * (threadIdx.x & 31);
  if (c >= C) return;
  const bool quad = c + 4 <= C;
  const int rg = threadIdx.x >> 5;
  const float* g = gb + (size_t)seg * ldgb;
  float mu[4], rs[4], ga[4], be[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ci = (quad || i < 2) ? c + i : c;
    mu[i] = mean[(size_t)seg * C + ci]; rs[i] = rstd[(size_t)seg * C + ci];
    ga[i] = 1.f + g[ci]; be[i] = g[C + ci];
  }
  const size_t base = (size_t)tile * 128;
  if (quad && ((ldx | ldy | ldyh) & 3) == 0) {
#pragma unroll 4
    for (int lr = rg; lr < nvalid; lr += 8) {
      const size_t r = base + lr;
      const float4 v = *reinterpret_cast<const float4*>(x + r * ldx + c);
      float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a = ga[i] * ((o[i] - mu[i]) * rs[i]) + be[i];
        o[i] = a > 0.f ? a : a * slope;
      }
      if (y != nullptr) *reinterpret_cast<float4*>(y + r * ldy + c) = make_float4(o[0], o[1], o[2], o[3]);
      if (yh != nullptr) {
        const twin2_t t0 = to_twin2(o[0], o[1]), t1 = to_twin2(o[2], o[3]);
        uint2 w;
        w.x = *reinterpret_cast<const uint32_t*>(&t0); w.y = *reinterpret_cast<const uint32_t*>(&t1);
        *reinterpret_cast<uint2*>(yh + r * ldyh + c) = w;
      }
    }
  } else {
    const int np = quad ? 2 : 1;                   // channel pairs of this thread
#pragma unroll 2
    for (int lr = rg; lr < nvalid; lr += 8) {
      const size_t r = base + lr;
      for (int h = 0; h < np; ++h) {
        const float2 v = *reinterpret_cast<const float2*>(x + r * ldx + c + 2 * h);
        float a = ga[2 * h] * ((v.x - mu[2 * h]) * rs[2 * h]) + be[2 * h];
        float b2 = ga[2 * h + 1] * ((v.y - mu[2 * h + 1]) * rs[2 * h + 1]) + be[2 * h + 1];
        a = a > 0.f ? a : a * slope;
        b2 = b2 > 0.f ? b2 : b2 * slope;
        if (y != nullptr) *reinterpret_cast<float2*>(y + r * ldy + c + 2 * h) = make_float2(a, b2);
        if (yh != nullptr) *reinterpret_cast<twin2_t*>(yh + r * ldyh + c + 2 * h) = to_twin2(a, b2);
      }
    }
  }
}

cudaError_t adain_act(const float* x, int ldx, int C, const RowMap& map, const float* mean, const float* rstd,
                      const float* gb, int ldgb, float slope, float* y, int ldy, twin_t* yh, int ldyh,
                      cudaStream_t st) {
  if ((C & 1) || (ldx & 1) || (ldy & 1) || (ldgb & 1)) return cudaErrorInvalidValue;
  dim3 grid(map.n_tiles, (C + 127) / 128);
  launch_k(adain_act_kernel, grid, 256, 0, st, x, ldx, C, map, mean, rstd, gb, ldgb, slope, y, ldy, yh, ldyh);
  return cudaGetLastError();
}

// ============================================================================ token -> frame expansion
// out[row(seg, f), c] = tok[(tok_off[seg] + idx[cu_f[seg] + f / div]) * C + c]
// Replaces the dense one-hot matmul of speech_predictor.py:88,93 / pitch_energy_predictor.py:91
// (exact: one non-zero per alignment column).
__global__ void __launch_bounds__(256) expand_tokens_kernel(const float* __restrict__ tok, int C,
                                                            const int* __restrict__ idx, const int* __restrict__ cu_f,
                                                            const int* __restrict__ tok_off, int div, RowMap map,
                                                            float* __restrict__ out, int ldo, int col0,
                                                            twin_t* __restrict__ outh, int ldh, int colh0) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int f_base = tile * 128 - map.seg_start[seg];
  const int total = nvalid * C;
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < total; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    const int f = (f_base + lr) / div;
    const int t = tok_off[seg] + idx[cu_f[seg] + f];
    const float v = tok[(size_t)t * C + c];
    const size_t r = (size_t)tile * 128 + lr;
    out[r * ldo + col0 + c] = v;
    if (outh != nullptr) outh[r * ldh + colh0 + c] = to_twin(v);
  }
}

cudaError_t expand_tokens(const float* tok, int C, const int* idx, const int* cu_f, const int* tok_off, int div,
                          const RowMap& map, float* out, int ldo, int col0, twin_t* outh, int ldh, int colh0,
                          cudaStream_t st) {
  launch_k(expand_tokens_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, 4, 64)), 256, 0, st, tok, C, idx, cu_f, tok_off, div, map, out, ldo, col0, outh, ldh,
                                                    colh0);
  return cudaGetLastError();
}

// ============================================================================ x4 linear upsample
// nn.Upsample(scale_factor=4, mode="linear"), align_corners=False   speech_predictor.py:76,89-90
// src = 0.25 * (i + 0.5) - 0.5 clamped at 0; i1 = min(i0 + 1, F - 1).
__global__ void __launch_bounds__(256) upsample4_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        const int* __restrict__ cu_f, RowMap map,
                                                        float* __restrict__ a4, float* __restrict__ b4) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int lr = threadIdx.x;
  if (lr >= 128 || lr >= map.tile_nvalid[tile]) return;
  const int i = tile * 128 + lr - map.seg_start[seg];
  const int F = map.seg_len[seg] >> 2;
  float src = 0.25f * ((float)i + 0.5f) - 0.5f;
  if (src < 0.f) src = 0.f;
  int i0 = (int)src;
  if (i0 > F - 1) i0 = F - 1;
  const int i1 = (i0 + 1 < F) ? i0 + 1 : F - 1;
  float l1 = src - (float)i0;
  l1 = fminf(fmaxf(l1, 0.f), 1.f);
  const float l0 = 1.f - l1;
  const int o = cu_f[seg];
  const size_t r = (size_t)tile * 128 + lr;
  // rounding order of ATen's CPU kernel (verified bit-exact on 16k samples): t = l1 * x1 rounded,
  // then fma(l0, x0, t).  The pitch curve feeds a phase integrator x 16 harmonics, so 1 ulp matters.
  a4[r] = __fmaf_rn(l0, a[o + i0], __fmul_rn(l1, a[o + i1]));
  b4[r] = __fmaf_rn(l0, b[o + i0], __fmul_rn(l1, b[o + i1]));
}

cudaError_t upsample4(const float* a, const float* b, const int* cu_f, const RowMap& map, float* a4, float* b4,
                      cudaStream_t st) {
  launch_k(upsample4_kernel, map.n_tiles, 128, 0, st, a, b, cu_f, map, a4, b4);
  return cudaGetLastError();
}

// ============================================================================ F0 / N 1->1 convs
// decoder.py:35-41,48-49: weight-normed Conv1d(1, 1, 3, padding=1) on the pitch and energy curves;
// results land in the channel slots of the concatenated decoder inputs (decoder.py:51,57).
// w = {f0_w0, f0_w1, f0_w2, f0_b, n_w0, n_w1, n_w2, n_b}
__global__ void __launch_bounds__(128) f0n_conv_kernel(const float* __restrict__ p4, const float* __restrict__ e4,
                                                       const float* __restrict__ w, RowMap map, F0NDst d) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int lr = threadIdx.x;
  const int nvalid = map.tile_nvalid[tile];
  if (lr >= nvalid) return;
  const int r = tile * 128 + lr;
  const int s0 = map.seg_start[seg], s1 = s0 + map.seg_len[seg];
  const float pm = (r - 1 >= s0) ? p4[r - 1] : 0.f, pc = p4[r], pp = (r + 1 < s1) ? p4[r + 1] : 0.f;
  const float em = (r - 1 >= s0) ? e4[r - 1] : 0.f, ec = e4[r], ep = (r + 1 < s1) ? e4[r + 1] : 0.f;
  const float f0 = w[3] + w[0] * pm + w[1] * pc + w[2] * pp;
  const float n = w[7] + w[4] * em + w[5] * ec + w[6] * ep;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (d.f[k] != nullptr) {
      d.f[k][(size_t)r * d.ld[k] + d.col[k]] = f0;
      d.f[k][(size_t)r * d.ld[k] + d.col[k] + 1] = n;
    }
    if (d.h[k] != nullptr) {
      d.h[k][(size_t)r * d.ldh[k] + d.col[k]] = to_twin(f0);
      d.h[k][(size_t)r * d.ldh[k] + d.col[k] + 1] = to_twin(n);
    }
  }
}

cudaError_t f0n_conv(const float* p4, const float* e4, const float* w, const RowMap& map, const F0NDst& d,
                     cudaStream_t st) {
  launch_k(f0n_conv_kernel, map.n_tiles, 128, 0, st, p4, e4, w, map, d);
  return cudaGetLastError();
}

// ============================================================================ boundary transposes
// Reference tensors are [C, L] per utterance (B = 1 slices of [B, C, L]); src element (c, t) of
// segment b is at src[off[b] + c * cs[b] + t].
__global__ void __launch_bounds__(256) transpose_in_kernel(const float* __restrict__ src,
                                                           const long long* __restrict__ off,
                                                           const int* __restrict__ cs, int C, RowMap map,
                                                           float* __restrict__ dst, int ld, int col0,
                                                           twin_t* __restrict__ dsth, int ldh, int colh0) {
  pdl_enter();
  __shared__ float t[32][33];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int t_base = tile * 128 - map.seg_start[seg];
  const float* s = src + off[seg];
  const int stride = cs[seg];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int c0 = blockIdx.y * 32; c0 < C; c0 += gridDim.y * 32) {
    for (int r0 = 0; r0 < nvalid; r0 += 32) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int c = c0 + ty + 8 * k, lr = r0 + tx;
        t[ty + 8 * k][tx] = (c < C && lr < nvalid) ? s[(size_t)c * stride + t_base + lr] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int lr = r0 + ty + 8 * k, c = c0 + tx;
        if (c < C && lr < nvalid) {
          const size_t r = (size_t)tile * 128 + lr;
          const float v = t[tx][ty + 8 * k];
          dst[r * ld + col0 + c] = v;
          if (dsth != nullptr) dsth[r * ldh + colh0 + c] = to_twin(v);
        }
      }
      __syncthreads();
    }
  }
}

cudaError_t transpose_in(const float* src, const long long* off, const int* cs, int C, const RowMap& map, float* dst,
                         int ld, int col0, twin_t* dsth, int ldh, int colh0, cudaStream_t st) {
  dim3 grid(map.n_tiles, (C + 31) / 32 < 8 ? (C + 31) / 32 : 8);
  launch_k(transpose_in_kernel, grid, 256, 0, st, src, off, cs, C, map, dst, ld, col0, dsth, ldh, colh0);
  return cudaGetLastError();
}

// dst element (c, t) of segment b at dst[off[b] + c * cs[b] + t]; `extra` replicated trailing
// frames (F.pad(..., mode="replicate"), generator.py:425-426) are appended after the last frame.
__global__ void __launch_bounds__(256) transpose_out_kernel(const float* __restrict__ src, int ld, int col0, int C,
                                                            RowMap map, float* __restrict__ dst,
                                                            const long long* __restrict__ off,
                                                            const int* __restrict__ cs, int extra) {
  pdl_enter();
  __shared__ float t[32][33];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int t_base = tile * 128 - map.seg_start[seg];
  const int len = map.seg_len[seg];
  float* d = dst + off[seg];
  const int stride = cs[seg];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int c0 = blockIdx.y * 32; c0 < C; c0 += gridDim.y * 32) {
    for (int r0 = 0; r0 < nvalid; r0 += 32) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int lr = r0 + ty + 8 * k, c = c0 + tx;
        t[ty + 8 * k][tx] = (c < C && lr < nvalid) ? src[((size_t)tile * 128 + lr) * ld + col0 + c] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int c = c0 + ty + 8 * k, lr = r0 + tx;
        if (c < C && lr < nvalid) {
          const float v = t[tx][ty + 8 * k];
          d[(size_t)c * stride + t_base + lr] = v;
          if (t_base + lr == len - 1)
            for (int e = 1; e <= extra; ++e) d[(size_t)c * stride + len - 1 + e] = v;
        }
      }
      __syncthreads();
    }
  }
}

cudaError_t transpose_out(const float* src, int ld, int col0, int C, const RowMap& map, float* dst,
                          const long long* off, const int* cs, int extra, cudaStream_t st) {
  dim3 grid(map.n_tiles, (C + 31) / 32 < 8 ? (C + 31) / 32 : 8);
  launch_k(transpose_out_kernel, grid, 256, 0, st, src, ld, col0, C, map, dst, off, cs, extra);
  return cudaGetLastError();
}

// dense per-row vector <-> concatenated per-utterance vector ([sum L])
__global__ void __launch_bounds__(128) rowvec_kernel(float* __restrict__ rowv, float* __restrict__ dense,
                                                     const int* __restrict__ cu, RowMap map, int to_rows) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int lr = threadIdx.x;
  if (lr >= map.tile_nvalid[tile]) return;
  const size_t r = (size_t)tile * 128 + lr;
  const int i = cu[seg] + (tile * 128 + lr - map.seg_start[seg]);
  if (to_rows) rowv[r] = dense[i]; else dense[i] = rowv[r];
}

cudaError_t rowvec_copy(float* rowv, float* dense, const int* cu, const RowMap& map, int to_rows, cudaStream_t st) {
  launch_k(rowvec_kernel, map.n_tiles, 128, 0, st, rowv, dense, cu, map, to_rows);
  return cudaGetLastError();
}

// ============================================================================ depthwise conv + AdaLN
// ConvNeXtBlock front half (generator.py:470-476): depthwise Conv1d(k, padding=(k-1)/2, groups=C)
// -> LayerNorm over channels (eps) -> (1 + gamma) * x + beta.  C = 512, 16 rows per CTA,
// 256 threads x 2 channels.  Weights are passed tap-major wt[k][C].
template <int K>
__global__ void __launch_bounds__(256) dwconv_adaln_kernel(const float* __restrict__ x, int ldx, RowMap map,
                                                           const float* __restrict__ wt, const float* __restrict__ bias,
                                                           const float* __restrict__ gb, int ldgb, float eps,
                                                           float* __restrict__ y, int ldy,
                                                           twin_t* __restrict__ yh, int ldyh) {
  pdl_enter();
  constexpr int C = 512, TR = 16, PAD = (K - 1) / 2;
  __shared__ float red[8][TR];
  __shared__ float stat[TR];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int lr0 = blockIdx.y * TR;
  if (lr0 >= nvalid) return;
  const int r0 = tile * 128 + lr0;
  const int c = threadIdx.x * 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  float w0[K], w1[K];
#pragma unroll
  for (int j = 0; j < K; ++j) {
    const float2 w = *reinterpret_cast<const float2*>(wt + (size_t)j * C + c);
    w0[j] = w.x; w1[j] = w.y;
  }
  float a0[TR], a1[TR];
  const float2 b = *reinterpret_cast<const float2*>(bias + c);
#pragma unroll
  for (int o = 0; o < TR; ++o) { a0[o] = b.x; a1[o] = b.y; }
  // k >= 15: the TR + K - 1 input rows go through shared memory with 16-byte cp.async copies, all in flight at once
  // (46 / 30 rows x 2 KB per CTA, two CTAs per SM), instead of one dependent 8-byte load per row and thread: the k = 31
  // layer was latency-bound at 1.1 TB/s.  Small k keeps the direct loads (7 / 3 extra rows).
  constexpr bool kSlab = K >= 15;
  extern __shared__ __align__(16) float slab[];
  if constexpr (kSlab) {
    for (int idx = threadIdx.x; idx < (TR + K - 1) * (C / 4); idx += 256) {
      const int row = idx / (C / 4), c4 = (idx - row * (C / 4)) * 4;
      const int r = r0 - PAD + row;
      float* d = slab + row * C + c4;
      if (r >= 0 && r < map.R) {
        const uint32_t da = (uint32_t)__cvta_generic_to_shared(d);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(x + (size_t)r * ldx + c4) : "memory");
      } else {
        *reinterpret_cast<float4*>(d) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TR + K - 1; ++i) {
    const int r = r0 - PAD + i;
    float2 v = make_float2(0.f, 0.f);
    if constexpr (kSlab) {
      v = *reinterpret_cast<const float2*>(slab + i * C + c);
    } else {
      if (r >= 0 && r < map.R) v = *reinterpret_cast<const float2*>(x + (size_t)r * ldx + c);
    }
#pragma unroll
    for (int o = 0; o < TR; ++o) {
      const int j = i - o;
      if (j >= 0 && j < K) { a0[o] = fmaf(w0[j], v.x, a0[o]); a1[o] = fmaf(w1[j], v.y, a1[o]); }
    }
  }
  // LayerNorm over the 512 channels of each row: mean, then variance of the centred values
  float m[TR];
#pragma unroll
  for (int o = 0; o < TR; ++o) {
    const float s = warp_sum(a0[o] + a1[o]);
    if (lane == 0) red[warp][o] = s;
  }
  __syncthreads();
  if (threadIdx.x < TR) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    stat[threadIdx.x] = s * (1.f / C);
  }
  __syncthreads();
#pragma unroll
  for (int o = 0; o < TR; ++o) m[o] = stat[o];
  __syncthreads();
#pragma unroll
  for (int o = 0; o < TR; ++o) {
    const float d0 = a0[o] - m[o], d1 = a1[o] - m[o];
    const float s = warp_sum(d0 * d0 + d1 * d1);
    if (lane == 0) red[warp][o] = s;
  }
  __syncthreads();
  if (threadIdx.x < TR) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    stat[threadIdx.x] = rsqrtf(s * (1.f / C) + eps);
  }
  __syncthreads();
  const float2 g = *reinterpret_cast<const float2*>(gb + (size_t)seg * ldgb + c);
  const float2 be = *reinterpret_cast<const float2*>(gb + (size_t)seg * ldgb + C + c);
#pragma unroll
  for (int o = 0; o < TR; ++o) {
    if (lr0 + o < nvalid) {
      const float rs = stat[o];
      const float v0 = (1.f + g.x) * ((a0[o] - m[o]) * rs) + be.x;
      const float v1 = (1.f + g.y) * ((a1[o] - m[o]) * rs) + be.y;
      const size_t r = (size_t)r0 + o;
      if (y != nullptr) *reinterpret_cast<float2*>(y + r * ldy + c) = make_float2(v0, v1);
      if (yh != nullptr) *reinterpret_cast<twin2_t*>(yh + r * ldyh + c) = to_twin2(v0, v1);
    }
  }
}

cudaError_t dwconv_adaln(const float* x, int ldx, const RowMap& map, int K, const float* wt, const float* bias,
                         const float* gb, int ldgb, float eps, float* y, int ldy, twin_t* yh, int ldyh,
                         cudaStream_t st) {
  dim3 grid(map.n_tiles, 8);
  switch (K) {
    case 31: {
      constexpr int smem = (16 + 30) * 512 * 4;
      static bool done = false;
      if (!done) { cudaFuncSetAttribute(dwconv_adaln_kernel<31>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); done = true; }
      launch_k(dwconv_adaln_kernel<31>, grid, 256, smem, st, x, ldx, map, wt, bias, gb, ldgb, eps, y, ldy, yh, ldyh);
      break;
    }
    case 15: {
      constexpr int smem = (16 + 14) * 512 * 4;
      static bool done = false;
      if (!done) { cudaFuncSetAttribute(dwconv_adaln_kernel<15>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); done = true; }
      launch_k(dwconv_adaln_kernel<15>, grid, 256, smem, st, x, ldx, map, wt, bias, gb, ldgb, eps, y, ldy, yh, ldyh);
      break;
    }
    case 7: launch_k(dwconv_adaln_kernel<7>, grid, 256, 0, st, x, ldx, map, wt, bias, gb, ldgb, eps, y, ldy, yh, ldyh); break;
    case 3: launch_k(dwconv_adaln_kernel<3>, grid, 256, 0, st, x, ldx, map, wt, bias, gb, ldgb, eps, y, ldy, yh, ldyh); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// ============================================================================ AdaLN on rows (two heads)
// amp/phase_final_layer_norm (generator.py:417-421): one LayerNorm (eps) statistic per row shared
// by two style-conditioned affine maps.  One warp per row, C = 512.
// four consecutive values -> fp16 twin g (8-byte store) and, optionally, the residual twin l
__device__ __forceinline__ void split4(const float* o, __half* g, __half* l) {
  float c[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) c[j] = fminf(fmaxf(o[j], -65504.f), 65504.f);
  const __half2 g0 = __floats2half2_rn(c[0], c[1]), g1 = __floats2half2_rn(c[2], c[3]);
  uint2 pk;
  pk.x = *reinterpret_cast<const uint32_t*>(&g0);
  pk.y = *reinterpret_cast<const uint32_t*>(&g1);
  *reinterpret_cast<uint2*>(g) = pk;
  if (l != nullptr) {
    const __half2 l0 = __floats2half2_rn(c[0] - __low2float(g0), c[1] - __high2float(g0));
    const __half2 l1 = __floats2half2_rn(c[2] - __low2float(g1), c[3] - __high2float(g1));
    pk.x = *reinterpret_cast<const uint32_t*>(&l0);
    pk.y = *reinterpret_cast<const uint32_t*>(&l1);
    *reinterpret_cast<uint2*>(l) = pk;
  }
}

__global__ void __launch_bounds__(256) adaln2_kernel(const float* __restrict__ x, int ldx, RowMap map,
                                                     const float* __restrict__ gb1, const float* __restrict__ gb2,
                                                     int ldgb, float eps, float* __restrict__ y1,
                                                     float* __restrict__ y2, int ldy, __half* __restrict__ y1g,
                                                     __half* __restrict__ y2g, int ldyh,
                                                     __half* __restrict__ y1l, __half* __restrict__ y2l) {
  pdl_enter();
  constexpr int C = 512;
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int lr = warp; lr < nvalid; lr += 8) {
    const size_t r = (size_t)tile * 128 + lr;
    float4 v[4];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[k] = *reinterpret_cast<const float4*>(x + r * ldx + k * 128 + lane * 4);
      s += v[k].x + v[k].y + v[k].z + v[k].w;
    }
    const float mean = warp_sum(s) * (1.f / C);
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v[k].x -= mean; v[k].y -= mean; v[k].z -= mean; v[k].w -= mean;
      q += v[k].x * v[k].x + v[k].y * v[k].y + v[k].z * v[k].z + v[k].w * v[k].w;
    }
    const float rs = rsqrtf(warp_sum(q) * (1.f / C) + eps);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = k * 128 + lane * 4;
      const float n[4] = {v[k].x * rs, v[k].y * rs, v[k].z * rs, v[k].w * rs};
      const float4 g1 = *reinterpret_cast<const float4*>(gb1 + (size_t)seg * ldgb + c);
      const float4 b1 = *reinterpret_cast<const float4*>(gb1 + (size_t)seg * ldgb + C + c);
      const float o1[4] = {(1.f + g1.x) * n[0] + b1.x, (1.f + g1.y) * n[1] + b1.y, (1.f + g1.z) * n[2] + b1.z,
                           (1.f + g1.w) * n[3] + b1.w};
      if (y1 != nullptr) *reinterpret_cast<float4*>(y1 + r * ldy + c) = make_float4(o1[0], o1[1], o1[2], o1[3]);
      if (y1g != nullptr) split4(o1, y1g + r * ldyh + c, y1l ? y1l + r * ldyh + c : nullptr);
      if (gb2 != nullptr) {
        const float4 g2 = *reinterpret_cast<const float4*>(gb2 + (size_t)seg * ldgb + c);
        const float4 b2 = *reinterpret_cast<const float4*>(gb2 + (size_t)seg * ldgb + C + c);
        const float o2[4] = {(1.f + g2.x) * n[0] + b2.x, (1.f + g2.y) * n[1] + b2.y, (1.f + g2.z) * n[2] + b2.z,
                             (1.f + g2.w) * n[3] + b2.w};
        if (y2 != nullptr) *reinterpret_cast<float4*>(y2 + r * ldy + c) = make_float4(o2[0], o2[1], o2[2], o2[3]);
        if (y2g != nullptr) split4(o2, y2g + r * ldyh + c, y2l ? y2l + r * ldyh + c : nullptr);
      }
    }
  }
}

cudaError_t adaln2(const float* x, int ldx, const RowMap& map, const float* gb1, const float* gb2, int ldgb, float eps,
                   float* y1, float* y2, int ldy, __half* y1g, __half* y2g, int ldyh, __half* y1l, __half* y2l,
                   cudaStream_t st) {
  launch_k(adaln2_kernel, map.n_tiles, 256, 0, st, x, ldx, map, gb1, gb2, ldgb, eps, y1, y2, ldy, y1g, y2g, ldyh, y1l, y2l);
  return cudaGetLastError();
}

// ============================================================================ GRN
// generator.py:496-499:  Gx = ||x||_2 over time, Nx = Gx / (mean_c Gx + 1e-6),
// y = gamma * (x * Nx) + beta + x  ==  x * (1 + gamma * Nx) + beta.
__global__ void __launch_bounds__(256) grn_scale_kernel(const float* __restrict__ gx, int C,
                                                        const float* __restrict__ gamma, float* __restrict__ scale) {
  pdl_enter();
  __shared__ float red[8];
  const int seg = blockIdx.x;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += gx[(size_t)seg * C + c];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float denom = tot / (float)C + 1e-6f;
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    scale[(size_t)seg * C + c] = 1.f + gamma[c] * (gx[(size_t)seg * C + c] / denom);
}

// GRN statistic from the partials the pw1 epilogue left (Epi::colss): Gx[c] = sqrt(sum over the
// utterance's (tile, quarter) rows); grn_scale_kernel then forms 1 + gamma * Gx / (mean_c Gx + 1e-6).
// Deterministic: fixed summation order, fp64 accumulation.  CTA = 64 channels x 4 interleaved row parts.
__global__ void __launch_bounds__(256) grn_gx_partials_kernel(const float* __restrict__ colss, int ld, int C,
                                                              RowMap map, float* __restrict__ gx) {
  pdl_enter();
  __shared__ double part[4][64];
  const int seg = blockIdx.x;
  const int cl = threadIdx.x & 63, pt = threadIdx.x >> 6;
  const int c = blockIdx.y * 64 + cl;
  const int t0 = 4 * (map.seg_start[seg] / 128), nt = 4 * ((map.seg_len[seg] + 127) / 128);
  double q = 0.0;
  if (c < C)
    for (int t = pt; t < nt; t += 4) q += (double)colss[(size_t)(t0 + t) * ld + c];
  part[pt][cl] = q;
  __syncthreads();
  if (pt == 0 && c < C) gx[(size_t)seg * C + c] = (float)sqrt(part[0][cl] + part[1][cl] + part[2][cl] + part[3][cl]);
}

cudaError_t grn_scale_partials(const float* colss, int ld, int C, const RowMap& map, const float* gamma, float* gx,
                               float* scale, cudaStream_t st) {
  launch_k(grn_gx_partials_kernel, dim3(map.n_seg, (C + 63) / 64), 256, 0, st, colss, ld, C, map, gx);
  launch_k(grn_scale_kernel, map.n_seg, 256, 0, st, gx, C, gamma, scale);
  return cudaGetLastError();
}

// Per-utterance pw2 weights with the GRN scale folded in: Wu[u, n, k] = fp16(W[n, k] * scale[u, k])
// (pwconv2(x * s + beta) = (W diag(s)) x + W beta: the W beta term is added to the bias at prepack).
__global__ void __launch_bounds__(256) grn_weights_kernel(const float* __restrict__ W, int N, int K,
                                                          const float* __restrict__ scale,
                                                          twin_t* __restrict__ Wu) {
  pdl_enter();
  const int u = blockIdx.y;
  const size_t total = (size_t)N * K / 2;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t e = 2 * i;
    const int k = (int)(e % K);
    const float2 w = *reinterpret_cast<const float2*>(W + e);
    const float2 sc = *reinterpret_cast<const float2*>(scale + (size_t)u * K + k);
    *reinterpret_cast<twin2_t*>(Wu + (size_t)u * N * K + e) = to_twin2(w.x * sc.x, w.y * sc.y);
  }
}

cudaError_t grn_weights(const float* W, int N, int K, const float* scale, int n_seg, twin_t* Wu, cudaStream_t st) {
  if (K % 2 != 0) return cudaErrorInvalidValue;
  launch_k(grn_weights_kernel, dim3(148, n_seg), 256, 0, st, W, N, K, scale, Wu);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) seg_affine_kernel(float* __restrict__ x, int ldx, int C, RowMap map,
                                                         const float* __restrict__ scale,
                                                         const float* __restrict__ shift,
                                                         twin_t* __restrict__ xh, int ldxh,
                                                         __half* __restrict__ xg, __half* __restrict__ xl) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int c4n = C / 4;
  const int total = nvalid * c4n;
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < total; i += gridDim.y * blockDim.x) {
    const int lr = i / c4n, c = (i - lr * c4n) * 4;
    const size_t r = (size_t)tile * 128 + lr;
    float4 v = *reinterpret_cast<float4*>(x + r * ldx + c);
    const float4 sc = *reinterpret_cast<const float4*>(scale + (size_t)seg * C + c);
    const float4 sf = *reinterpret_cast<const float4*>(shift + c);
    v.x = v.x * sc.x + sf.x; v.y = v.y * sc.y + sf.y; v.z = v.z * sc.z + sf.z; v.w = v.w * sc.w + sf.w;
    *reinterpret_cast<float4*>(x + r * ldx + c) = v;
    if (xh != nullptr) {
      *reinterpret_cast<twin2_t*>(xh + r * ldxh + c) = to_twin2(v.x, v.y);
      *reinterpret_cast<twin2_t*>(xh + r * ldxh + c + 2) = to_twin2(v.z, v.w);
    }
    if (xg != nullptr) {                         // fp16 split twins (same pitch as x): operand of an f16x3 GEMM
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) f16_split(vv[j], xg + r * ldx + c + j, xl + r * ldx + c + j);
    }
  }
}

cudaError_t grn_apply(float* x, int ldx, int C, const RowMap& map, const float* gx, const float* gamma,
                      const float* beta, float* scale_scratch, twin_t* xh, int ldxh, cudaStream_t st, __half* xg,
                      __half* xl) {
  launch_k(grn_scale_kernel, map.n_seg, 256, 0, st, gx, C, gamma, scale_scratch);
  dim3 grid(map.n_tiles, 4);
  launch_k(seg_affine_kernel, grid, 256, 0, st, x, ldx, C, map, scale_scratch, beta, xh, ldxh, xg, xl);
  return cudaGetLastError();
}

// ============================================================================ 1-channel projection
// F0_proj / N_proj: Conv1d(C, 1, 1)  pitch_energy_predictor.py:80-81,114,119.  One warp per row.
__global__ void __launch_bounds__(256) rowdot_kernel(const float* __restrict__ x, int ldx, int C, RowMap map,
                                                     const float* __restrict__ w, const float* __restrict__ bias,
                                                     const int* __restrict__ cu, float* __restrict__ dense) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int lr = warp + 8 * blockIdx.y; lr < nvalid; lr += 8 * gridDim.y) {
    const size_t r = (size_t)tile * 128 + lr;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(x[r * ldx + c], w[c], s);
    s = warp_sum(s);
    if (lane == 0) dense[cu[seg] + (tile * 128 + lr - map.seg_start[seg])] = s + bias[0];
  }
}

cudaError_t rowdot(const float* x, int ldx, int C, const RowMap& map, const float* w, const float* bias, const int* cu,
                   float* dense, cudaStream_t st) {
  launch_k(rowdot_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, 4, 16)), 256, 0, st, x, ldx, C, map, w, bias, cu, dense);
  return cudaGetLastError();
}

// fp16 split twins of the first C columns of the valid rows, recomputed from the fp32 master
__global__ void __launch_bounds__(256) twin_refresh_kernel(const float* __restrict__ f, int ld, int C, RowMap map,
                                                           __half* __restrict__ h, __half* __restrict__ l) {
  pdl_enter();
  const int tile = blockIdx.x;
  if (map.tile_seg[tile] < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    const size_t o = (size_t)(tile * 128 + lr) * ld + c;
    const float x = f[o];
    f16_split(x, h + o, l + o);
  }
}
cudaError_t twin_refresh(const float* f, int ld, int C, const RowMap& map, __half* h, __half* l, cudaStream_t st) {
  dim3 grid(map.n_tiles, 4);
  launch_k(twin_refresh_kernel, grid, 256, 0, st, f, ld, C, map, h, l);
  return cudaGetLastError();
}

// ============================================================================ saturation audit
// The fp16 operand twins saturate at +-65504 (to_twin / f16_split) where the reference's fp32 -- or a true
// bfloat16 path -- would not.  Counts the saturated elements of one twin buffer (not on the hot path:
// stts_saturation_count walks every twin of the workspace after a run).
__global__ void __launch_bounds__(256) saturation_kernel(const uint16_t* __restrict__ h, size_t n,
                                                         unsigned long long* __restrict__ count) {
  pdl_enter();
  unsigned long long c = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    c += ((h[i] & 0x7FFFu) == 0x7BFFu) ? 1ull : 0ull;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}
cudaError_t saturation_count(const twin_t* h, size_t n, unsigned long long* count, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  launch_k(saturation_kernel, 592, 256, 0, st, reinterpret_cast<const uint16_t*>(h), n, count);
  return cudaGetLastError();
}

// zero-fill of valid rows of a column block (WN "output" accumulator is rebuilt by EPI_RESSKIP's
// accumulate2 = 0 path, so this is only needed by tests)
__global__ void fill_kernel(float* p, size_t n, float v) {
  pdl_enter();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
cudaError_t fill(float* p, size_t n, float v, cudaStream_t st) {
  launch_k(fill_kernel, 592, 256, 0, st, p, n, v);
  return cudaGetLastError();
}

}  // namespace stts
