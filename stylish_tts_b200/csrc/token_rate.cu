// Token-rate kernels (<= 510 phonemes per utterance) on the packed token layout: embedding gather,
// row LayerNorm family, partial rotary embedding and multi-head attention.  Reference:
//   TextEncoder            train/models/text_encoder.py:396-462 (Encoder 332-393, MHA 171-296,
//                          RotaryPositionalEmbeddings 89-168, channel LayerNorm 15-33, FFN 305-329)
//   AdaptiveLayerNorm      train/models/ada_norm.py:185-201
// The dense contractions of these modules (1x1 / k3 / k5 convs) run on the conv-as-GEMM kernels;
// only what is not a GEMM lives here.  A packed token buffer has the same geometry as the
// frame-rate buffers (common.cuh): utterances start on multiples of 128 rows, guard rows stay zero,
// so `x * x_mask` of the reference is the identity and conv zero padding is free.
#include <cstdlib>

#include "kernels.h"

namespace stts {

// ------------------------------------------------------------------------------ embedding
// x[row(seg, t), :] = table[tok[cu_tok[seg] + t], :]   (table already scaled by sqrt(C))
// The copy / gather kernels of the token layout have only n_tiles (~50, or 1 for a single utterance) row tiles:
// kCopySplit CTAs share a tile's elements so that the launch is not one long serial loop per SM
constexpr int kCopySplit = 8;
__global__ void __launch_bounds__(128) embed_kernel(const int* __restrict__ tok, const int* __restrict__ cu_tok,
                                                    const float* __restrict__ table, int C, int n_vocab, RowMap map,
                                                    float* __restrict__ x, int ldx, __half* __restrict__ xg,
                                                    __half* __restrict__ xl) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int t0 = tile * 128 - map.seg_start[seg];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    int id = tok[cu_tok[seg] + t0 + lr];
    id = id < 0 ? 0 : (id >= n_vocab ? n_vocab - 1 : id);
    const float v = table[(size_t)id * C + c];
    const size_t o = (size_t)(tile * 128 + lr) * ldx + c;
    x[o] = v;
    if (xg != nullptr) f16_split(v, xg + o, xl + o);
  }
}

cudaError_t embed_tokens(const int* tok, const int* cu_tok, const float* table, int C, int n_vocab, const RowMap& map,
                         float* x, int ldx, cudaStream_t st, __half* xg, __half* xl) {
  launch_k(embed_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, kCopySplit, 64)), 128, 0, st, tok, cu_tok, table, C, n_vocab, map, x, ldx, xg, xl);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ row LayerNorm
// y = act(affine((x - mean) * rsqrt(var + eps)))  over the C channels of one row, biased variance,
// two-pass in registers (one warp per row, C <= 512).
//   seg_stride == 0 : affine = gamma[c] * n + beta[c]                (channel LayerNorm / nn.LayerNorm)
//   seg_stride  > 0 : affine = (1 + g[seg, c]) * n + g[seg, C + c]   (AdaptiveLayerNorm, g = style FC row)
__global__ void __launch_bounds__(256) row_ln_kernel(const float* x, int ldx, int C, RowMap map,
                                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                                     int seg_stride, float eps, int relu, float* y,
                                                     int ldy, __half* __restrict__ yg,
                                                     __half* __restrict__ yl) {
  pdl_enter();  // x may alias y (in place)
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* ga = seg_stride ? gamma + (size_t)seg * seg_stride : gamma;
  const float* be = seg_stride ? gamma + (size_t)seg * seg_stride + C : beta;
  for (int lr = warp + 8 * blockIdx.y; lr < nvalid; lr += 8 * gridDim.y) {
    const size_t r = (size_t)tile * 128 + lr;
    float v[16];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = lane + 32 * i;
      v[i] = (c < C) ? x[r * ldx + c] : 0.f;
      s += v[i];
    }
    const float mean = warp_sum(s) / (float)C;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = lane + 32 * i;
      const float d = (c < C) ? v[i] - mean : 0.f;
      q += d * d;
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = lane + 32 * i;
      if (c < C) {
        const float n = (v[i] - mean) * rstd;
        float o = seg_stride ? (1.f + ga[c]) * n + be[c] : n * ga[c] + be[c];
        if (relu) o = fmaxf(o, 0.f);
        y[r * ldy + c] = o;
        if (yg != nullptr) f16_split(o, yg + r * ldy + c, yl + r * ldy + c);   // operand twins of the next f16x3 GEMM
      }
    }
  }
}

cudaError_t row_ln(const float* x, int ldx, int C, const RowMap& map, const float* gamma, const float* beta,
                   int seg_stride, float eps, int relu, float* y, int ldy, cudaStream_t st, __half* yg, __half* yl) {
  if (C > 512) return cudaErrorInvalidValue;
  launch_k(row_ln_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, 4, 16)), 256, 0, st, x, ldx, C, map, gamma, beta, seg_stride, eps, relu, y, ldy, yg,
                                                      yl);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ partial RoPE
// In place on `n_blocks` column blocks of H heads x D features starting at columns col0 + b * blk:
// the first rd features of every head are rotated (half-split form, theta_i = 10000^(-2i/rd)),
// position = row index inside the utterance.  text_encoder.py:89-168,200-201.
__global__ void __launch_bounds__(256) rope_kernel(float* __restrict__ x, int ldx, int col0, int blk, int n_blocks,
                                                   int H, int D, int rd, RowMap map) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int h = rd >> 1;
  const int per_row = n_blocks * H * h;
  const int pos0 = tile * 128 - map.seg_start[seg];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * per_row; i += gridDim.y * blockDim.x) {
    const int lr = i / per_row;
    int k = i - lr * per_row;
    const int b = k / (H * h);
    k -= b * H * h;
    const int hd = k / h, j = k - hd * h;
    const float inv = 1.0f / powf(10000.f, (float)(2 * j) / (float)rd);
    const float ang = (float)(pos0 + lr) * inv;
    float sn, cs;
    sincosf(ang, &sn, &cs);
    float* p = x + (size_t)(tile * 128 + lr) * ldx + col0 + b * blk + hd * D;
    const float a = p[j], bb = p[j + h];
    p[j] = a * cs - bb * sn;
    p[j + h] = bb * cs + a * sn;
  }
}

cudaError_t rope_inplace(float* x, int ldx, int col0, int blk, int n_blocks, int H, int D, int rd, const RowMap& map,
                         cudaStream_t st) {
  launch_k(rope_kernel, dim3(map.n_tiles, 8), 256, 0, st, x, ldx, col0, blk, n_blocks, H, D, rd, map);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ attention
// out[q, hd*D : (hd+1)*D] = softmax_k(scale * <Q_q, K_k> + bias(q, k)) V_k over the keys of the same
// utterance.  One CTA = (16 queries, head, utterance), 128 threads, fp32 throughout.
//   band > 0: bias = -1e4 where |k - q_tok[q]| <= band  (the reference's inverted monotonic band,
//   pitch_energy_predictor.py:194-212 + text_encoder.py:255-262)
constexpr int kAttQ = 16;
__global__ void __launch_bounds__(128) attention_kernel(const float* __restrict__ Q, int ldq, int qcol,
                                                        const float* __restrict__ K, int ldk, int kcol,
                                                        const float* __restrict__ V, int ldv, int vcol, int D,
                                                        float scale, RowMap qmap, RowMap kmap,
                                                        const int* __restrict__ q_tok, const int* __restrict__ q_cu,
                                                        int band, float* __restrict__ out, int ldo) {
  pdl_enter();
  extern __shared__ float sh[];
  const int seg = blockIdx.z, hd = blockIdx.y;
  const int Tq = qmap.seg_len[seg], Tk = kmap.seg_len[seg];
  const int q0 = blockIdx.x * kAttQ;
  if (q0 >= Tq) return;
  const int nq = (Tq - q0 < kAttQ) ? (Tq - q0) : kAttQ;
  const int qrow0 = qmap.seg_start[seg] + q0, krow0 = kmap.seg_start[seg];
  const int Tkp = (Tk + 3) & ~3;
  float* sq = sh;                    // [kAttQ][D]
  float* sc = sh + kAttQ * D;        // [kAttQ][Tkp]
  for (int i = threadIdx.x; i < kAttQ * D; i += blockDim.x) {
    const int qi = i / D, d = i - qi * D;
    sq[i] = (qi < nq) ? Q[(size_t)(qrow0 + qi) * ldq + qcol + hd * D + d] * scale : 0.f;
  }
  __syncthreads();
  // ---- scores
  for (int j = threadIdx.x; j < Tk; j += blockDim.x) {
    const float* kp = K + (size_t)(krow0 + j) * ldk + kcol + hd * D;
    float acc[kAttQ];
#pragma unroll
    for (int i = 0; i < kAttQ; ++i) acc[i] = 0.f;
    for (int d = 0; d < D; d += 4) {
      const float4 kv = *reinterpret_cast<const float4*>(kp + d);
#pragma unroll
      for (int i = 0; i < kAttQ; ++i) {
        const float4 qv = *reinterpret_cast<const float4*>(sq + i * D + d);
        acc[i] = fmaf(qv.x, kv.x, acc[i]);
        acc[i] = fmaf(qv.y, kv.y, acc[i]);
        acc[i] = fmaf(qv.z, kv.z, acc[i]);
        acc[i] = fmaf(qv.w, kv.w, acc[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < kAttQ; ++i) {
      float b = 0.f;
      if (band > 0 && i < nq) {
        const int tau = q_tok[q_cu[seg] + q0 + i];
        if (j >= tau - band && j <= tau + band) b = -1e4f;
      }
      sc[i * Tkp + j] = acc[i] + b;
    }
  }
  __syncthreads();
  // ---- softmax (one warp per 4 queries)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < nq; i += 4) {
    float* row = sc + i * Tkp;
    float m = -3.0e38f;
    for (int j = lane; j < Tk; j += 32) m = fmaxf(m, row[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int j = lane; j < Tk; j += 32) {
      const float e = expf(row[j] - m);
      row[j] = e;
      s += e;
    }
    s = warp_sum(s);
    const float inv = 1.f / s;
    for (int j = lane; j < Tk; j += 32) row[j] *= inv;
  }
  __syncthreads();
  // ---- P V
  for (int idx = threadIdx.x; idx < nq * D; idx += blockDim.x) {
    const int i = idx / D, d = idx - i * D;
    const float* vp = V + (size_t)krow0 * ldv + vcol + hd * D + d;
    const float* pr = sc + i * Tkp;
    float a0 = 0.f, a1 = 0.f;
    int j = 0;
    for (; j + 2 <= Tk; j += 2) {
      a0 = fmaf(pr[j], vp[(size_t)j * ldv], a0);
      a1 = fmaf(pr[j + 1], vp[(size_t)(j + 1) * ldv], a1);
    }
    if (j < Tk) a0 = fmaf(pr[j], vp[(size_t)j * ldv], a0);
    out[(size_t)(qrow0 + i) * ldo + hd * D + d] = a0 + a1;
  }
}

// Tiled variant for the head sizes of the model (D = 16 text encoders, 40 cross attention, 160 prosody
// encoders).  One CTA = (16 queries, head, utterance); keys / values are staged through shared memory
// in chunks of 64 rows (coalesced), partial RoPE is applied while staging (rd > 0: the first rd
// features of the head, half-split form, position = row index inside the utterance; text_encoder.py:
// 89-168) so Q / K stay un-rotated in memory and the separate rotation pass disappears.
constexpr int kAttChunk = 64;
constexpr float kLog2e = 1.4426950408889634f;
// rope table: tab[pos * (rd / 2) + j] = (cos, sin)(pos * 10000^(-2 j / rd)), same fp32 formula as rope_kernel
// Stage ROWS rows x D features of a head into shared memory: dst[r][f] (or dst[f][r] when TRANSPOSE),
// zero beyond nrows.  float4 global loads, issued in independent batches (the kernel is latency-bound:
// few resident warps, so the loads of a batch must be in flight together).  128 threads.
template <int D, int ROWS, bool TRANSPOSE, int NT = 128>
__device__ __forceinline__ void stage_rows(const float* __restrict__ src, int ld, int row0, int nrows,
                                           float* __restrict__ dst, int dst_ld) {
  constexpr int D4 = D / 4;
  constexpr int V4 = (ROWS * D4 + NT - 1) / NT;                      // float4 per thread (guarded tail)
  constexpr int B = (V4 % 5 == 0) ? 5 : ((V4 % 4 == 0) ? 4 : ((V4 % 3 == 0) ? 3 : ((V4 % 2 == 0) ? 2 : 1)));
#pragma unroll
  for (int b0 = 0; b0 < V4; b0 += B) {
    float4 v[B];
#pragma unroll
    for (int u = 0; u < B; ++u) {
      const int i = threadIdx.x + (b0 + u) * NT;
      const int r = i / D4, f4 = i - r * D4;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < ROWS * D4 && r < nrows) v[u] = *reinterpret_cast<const float4*>(src + (size_t)(row0 + r) * ld + 4 * f4);
    }
#pragma unroll
    for (int u = 0; u < B; ++u) {
      const int i = threadIdx.x + (b0 + u) * NT;
      const int r = i / D4, f = 4 * (i - r * D4);
      if (i < ROWS * D4) {
        if (TRANSPOSE) {
          dst[(f + 0) * dst_ld + r] = v[u].x; dst[(f + 1) * dst_ld + r] = v[u].y;
          dst[(f + 2) * dst_ld + r] = v[u].z; dst[(f + 3) * dst_ld + r] = v[u].w;
        } else {
          dst[r * dst_ld + f + 0] = v[u].x; dst[r * dst_ld + f + 1] = v[u].y;
          dst[r * dst_ld + f + 2] = v[u].z; dst[r * dst_ld + f + 3] = v[u].w;
        }
      }
    }
  }
}

// partial RoPE in place on staged rows x[r][f] (row pitch ldx): the first rd features, half-split
// pairs (j, j + rd/2), position pos0 + r; then an optional scale of the whole row block
template <int NT = 128>
__device__ __forceinline__ void rope_staged(float* x, int ldx, int nrows, int rd, int pos0,
                                            const float2* __restrict__ tab) {
  const int h = rd >> 1;
  for (int i = threadIdx.x; i < nrows * h; i += NT) {
    const int r = i / h, j = i - r * h;
    const float2 cs = tab[(size_t)(pos0 + r) * h + j];
    const float a = x[r * ldx + j], b = x[r * ldx + j + h];
    x[r * ldx + j] = a * cs.x - b * cs.y;
    x[r * ldx + j + h] = b * cs.x + a * cs.y;
  }
}

__global__ void rope_table_kernel(float2* tab, int n_pos, int rd) {
  pdl_enter();
  const int h = rd >> 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pos * h; i += gridDim.x * blockDim.x) {
    const int pos = i / h, j = i - pos * h;
    const float inv = 1.0f / powf(10000.f, (float)(2 * j) / (float)rd);
    float sn, cs;
    sincosf((float)pos * inv, &sn, &cs);
    tab[i] = make_float2(cs, sn);
  }
}

cudaError_t rope_table(float2* tab, int n_pos, int rd, cudaStream_t st) {
  launch_k(rope_table_kernel, 64, 256, 0, st, tab, n_pos, rd);
  return cudaGetLastError();
}

template <int D, int QB>
__global__ void __launch_bounds__(128) attention_tiled_kernel(const float* __restrict__ Q, int ldq, int qcol,
                                                              const float* __restrict__ K, int ldk, int kcol,
                                                              const float* __restrict__ V, int ldv, int vcol,
                                                              float scale, int rd, const float2* __restrict__ tab,
                                                              RowMap qmap, RowMap kmap,
                                                              const int* __restrict__ q_tok,
                                                              const int* __restrict__ q_cu, int band,
                                                              float* __restrict__ out, int ldo) {
  pdl_enter();
  extern __shared__ float sh[];
  constexpr int KLD = D + 1;                  // padded: a column of kv[][] is read by consecutive threads
  constexpr int VLD = kAttChunk + 4;          // transposed V chunk vt[d][j]: float4 reads along j, conflict-free
  constexpr int NOUT = QB * D / 128;          // outputs per thread in the P V phase
  static_assert(QB % 16 == 0 && (QB * D) % 128 == 0 && NOUT <= 20, "attention tile shape");
  const int seg = blockIdx.z, hd = blockIdx.y;
  const int Tq = qmap.seg_len[seg], Tk = kmap.seg_len[seg];
  const int q0 = blockIdx.x * QB;
  if (q0 >= Tq) return;
  const int nq = (Tq - q0 < QB) ? (Tq - q0) : QB;
  const int qrow0 = qmap.seg_start[seg] + q0, krow0 = kmap.seg_start[seg];
  const int Tkp = (Tk + 3) & ~3;
  float* sq = sh;                              // [QB][D]  (rows >= nq zero)
  float* kv = sq + QB * D;                     // K chunk [kAttChunk][KLD] / V chunk transposed [D][VLD]
  float* sc = kv + (kAttChunk * KLD > D * VLD ? kAttChunk * KLD : D * VLD);   // [QB][Tkp]
  // ---- queries (rotated, then scaled)
  stage_rows<D, QB, false>(Q + qcol + hd * D, ldq, qrow0, nq, sq, D);
  __syncthreads();
  if (rd > 0) {
    rope_staged(sq, D, nq, rd, q0, tab);
    __syncthreads();
  }
  for (int i = threadIdx.x; i < QB * D; i += 128) sq[i] *= scale;
  const int jl = threadIdx.x & 63, ig = threadIdx.x >> 6;     // key inside the chunk, half of the query groups
  // ---- scores, chunk by chunk
  for (int c0 = 0; c0 < Tk; c0 += kAttChunk) {
    const int nk = (Tk - c0 < kAttChunk) ? (Tk - c0) : kAttChunk;
    __syncthreads();                           // previous chunk consumed (and sq visible)
    stage_rows<D, kAttChunk, false>(K + kcol + hd * D, ldk, krow0 + c0, nk, kv, KLD);
    __syncthreads();
    if (rd > 0) {
      rope_staged(kv, KLD, nk, rd, c0, tab);
      __syncthreads();
    }
    if (jl < nk) {
      const float* kr = kv + jl * KLD;
      const int j = c0 + jl;
#pragma unroll 1
      for (int g = ig; g < QB / 8; g += 2) {   // groups of 8 queries
        if (g * 8 >= nq) break;
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        const float* qr = sq + g * 8 * D;
#pragma unroll 2
        for (int d = 0; d < D; d += 4) {
          const float k0 = kr[d], k1 = kr[d + 1], k2 = kr[d + 2], k3 = kr[d + 3];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 qv = *reinterpret_cast<const float4*>(qr + i * D + d);     // warp-wide broadcast
            acc[i] = fmaf(qv.x, k0, acc[i]);
            acc[i] = fmaf(qv.y, k1, acc[i]);
            acc[i] = fmaf(qv.z, k2, acc[i]);
            acc[i] = fmaf(qv.w, k3, acc[i]);
          }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int qi = g * 8 + i;
          float b = 0.f;
          if (band > 0 && qi < nq) {
            const int tau = q_tok[q_cu[seg] + q0 + qi];
            if (j >= tau - band && j <= tau + band) b = -1e4f;
          }
          sc[qi * Tkp + j] = acc[i] + b;
        }
      }
    }
  }
  __syncthreads();
  // ---- softmax (one warp per query, round robin)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < nq; i += 4) {
    float* row = sc + i * Tkp;
    float m = -3.0e38f;
    for (int j = lane; j < Tk; j += 32) m = fmaxf(m, row[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int j = lane; j < Tk; j += 32) {
      const float e = expf(row[j] - m);
      row[j] = e;
      s += e;
    }
    s = warp_sum(s);
    const float inv = 1.f / s;
    for (int j = lane; j < Tk; j += 32) row[j] *= inv;
    for (int j = Tk + lane; j < Tkp; j += 32) row[j] = 0.f;    // pad read by the float4 P V loop
  }
  // rows of P that were never written (queries >= nq) must still be finite for the P V loop
  for (int i = nq * Tkp + threadIdx.x; i < QB * Tkp; i += 128) sc[i] = 0.f;
  // ---- P V, chunk by chunk; output element idx = threadIdx.x + 128 * o -> (query idx / D, feature idx % D)
  float acc[NOUT];
  int poff[NOUT], doff[NOUT];                  // row of P (query) and feature of every output of this thread
#pragma unroll
  for (int o = 0; o < NOUT; ++o) {
    const int idx = threadIdx.x + 128 * o;
    acc[o] = 0.f;
    poff[o] = (idx / D) * Tkp;
    doff[o] = idx % D;
  }
  for (int c0 = 0; c0 < Tk; c0 += kAttChunk) {
    const int nk = (Tk - c0 < kAttChunk) ? (Tk - c0) : kAttChunk;
    __syncthreads();                           // softmax done / previous chunk consumed
    stage_rows<D, kAttChunk, true>(V + vcol + hd * D, ldv, krow0 + c0, nk, kv, VLD);   // rows >= nk are zero
    __syncthreads();
    const int nk4 = (nk + 3) & ~3;
    for (int j = 0; j < nk4; j += 4) {         // NOUT independent accumulation chains per thread
#pragma unroll
      for (int o = 0; o < NOUT; ++o) {
        const float4 pv = *reinterpret_cast<const float4*>(sc + poff[o] + c0 + j);
        const float4 vv = *reinterpret_cast<const float4*>(kv + doff[o] * VLD + j);
        acc[o] = fmaf(pv.x, vv.x, acc[o]);
        acc[o] = fmaf(pv.y, vv.y, acc[o]);
        acc[o] = fmaf(pv.z, vv.z, acc[o]);
        acc[o] = fmaf(pv.w, vv.w, acc[o]);
      }
    }
  }
#pragma unroll
  for (int o = 0; o < NOUT; ++o) {
    const int idx = threadIdx.x + 128 * o;
    const int i = idx / D, d = idx - i * D;
    if (i < nq) out[(size_t)(qrow0 + i) * ldo + hd * D + d] = acc[o];
  }
}

// ------------------------------------------------------------------------------ attention, single pass
// One CTA = (128 / LPQ queries, head, utterance).  A query is owned by LPQ adjacent lanes, each holding D / LPQ
// features of q and of the output accumulator in registers; keys / values stream through shared memory in chunks of
// 64 rows (partial RoPE applied while staging, like above) and the softmax is the running-max form, so neither the
// score matrix nor a second pass over V exists: K and V are staged once per CTA instead of once per 16 queries, and
// the per-key work of a lane is 2 D / LPQ FMAs + one exp2 (the rescale of the accumulators runs only when the running
// maximum moves).  fp32 throughout; exp(x) is exp2f(x * log2 e) (2 ulp + 6e-8 |x| relative: |x| is the distance to
// the running maximum, so the weights that matter are exact to ~1e-7) -- a handful of instructions instead of expf.
// Folding log2 e into q instead costs parity: q * scale is exact for the text encoders (scale 2^-2) and the scores of
// the fixture weights reach ~1e3, where one more rounding of q is an ulp of the score.
//   band > 0: additive bias -1e4 where |key - q_tok[query]| <= band (inverted monotonic band).
// Optional fp16 split twins of the output (operand of the following f16x3 projection GEMM).
template <int D, int LPQ, int QPT, int NT, int KP, int CH>
__global__ void __launch_bounds__(NT) attention_stream_kernel(const float* __restrict__ Q, int ldq, int qcol,
                                                              const float* __restrict__ K, int ldk, int kcol,
                                                              const float* __restrict__ V, int ldv, int vcol,
                                                              float scale, int rd, const float2* __restrict__ tab,
                                                              RowMap qmap, RowMap kmap, const int* __restrict__ q_tok,
                                                              const int* __restrict__ q_cu, int band,
                                                              float* __restrict__ out, int ldo,
                                                              __half* __restrict__ outg, __half* __restrict__ outl,
                                                              int ldh) {
  pdl_enter();
  extern __shared__ float sh[];
  // A query group = LPQ x KP adjacent lanes: LPQ lanes split the head's features, KP lane sets split the keys of
  // every chunk (key j of the chunk goes to set j % KP) and are merged once at the end (flash-decoding style), and
  // the group owns QPT queries (register blocking: one shared-memory K / V read feeds QPT FMAs).  KP > 1 is for
  // the short-sequence / small-head case (text encoders: ~126 tokens x 16 features), which is otherwise one
  // dependent chain per thread on a handful of warps per SM.
  constexpr int QPB = NT / (LPQ * KP) * QPT;     // queries per CTA
  constexpr int DL = D / LPQ;
  constexpr int KLD = D + 4;                     // row pitch of the staged K / V chunks (16-byte aligned rows)
  static_assert(D % (4 * LPQ) == 0, "attention lane split");
  const int seg = blockIdx.z, hd = blockIdx.y;
  const int Tq = qmap.seg_len[seg], Tk = kmap.seg_len[seg];
  const int q0 = blockIdx.x * QPB;
  if (q0 >= Tq) return;
  const int nq = (Tq - q0 < QPB) ? (Tq - q0) : QPB;
  const int qrow0 = qmap.seg_start[seg] + q0, krow0 = kmap.seg_start[seg];
  constexpr int KROWS = CH + 4 * KP;      // rows past the chunk are read (and masked) by the last key group
  float* sq = sh;                                // [QPB][D]
  float* sk = sq + QPB * D;                      // [KROWS][KLD]
  float* sv = sk + KROWS * KLD;                  // [KROWS][KLD]
  for (int i = threadIdx.x; i < 4 * KP * KLD; i += NT) {      // the pad rows must hold finite numbers (0 * p = 0)
    sk[CH * KLD + i] = 0.f;
    sv[CH * KLD + i] = 0.f;
  }
  // ---- queries: stage, rotate, scale; then every lane keeps its slices in registers
  stage_rows<D, QPB, false, NT>(Q + qcol + hd * D, ldq, qrow0, nq, sq, D);
  __syncthreads();
  if (rd > 0) {
    rope_staged<NT>(sq, D, nq, rd, q0, tab);
    __syncthreads();
  }
  const int grp = threadIdx.x / (LPQ * KP), kp = (threadIdx.x / LPQ) % KP, part = threadIdx.x % LPQ;
  float q[QPT][DL], acc[QPT][DL], m[QPT], l[QPT];
  int tau[QPT];
#pragma unroll
  for (int a = 0; a < QPT; ++a) {
    const int qi = grp * QPT + a;
#pragma unroll
    for (int i = 0; i < DL; ++i) {
      q[a][i] = sq[qi * D + part * DL + i] * scale;
      acc[a][i] = 0.f;
    }
    tau[a] = (band > 0) ? q_tok[q_cu[seg] + q0 + (qi < nq ? qi : 0)] : 0;
    m[a] = -3.0e38f;
    l[a] = 0.f;
  }
  for (int c0 = 0; c0 < Tk; c0 += CH) {
    const int nk = (Tk - c0 < CH) ? (Tk - c0) : CH;
    __syncthreads();                             // previous chunk consumed
    stage_rows<D, CH, false, NT>(K + kcol + hd * D, ldk, krow0 + c0, nk, sk, KLD);
    stage_rows<D, CH, false, NT>(V + vcol + hd * D, ldv, krow0 + c0, nk, sv, KLD);
    __syncthreads();
    if (rd > 0) {
      rope_staged<NT>(sk, KLD, nk, rd, c0, tab);
      __syncthreads();
    }
#pragma unroll 1
    for (int jb = 0; jb < nk; jb += 4 * KP) {    // this lane set's keys of the group: jb + u * KP + kp, u = 0 .. 3
      float s[QPT][4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int jj = jb + u * KP + kp;         // < 64 + 4 KP: rows past the chunk are masked below, smem is padded
        const float* kr = sk + jj * KLD + part * DL;
        float d[QPT];
#pragma unroll
        for (int a = 0; a < QPT; ++a) d[a] = 0.f;
#pragma unroll
        for (int i = 0; i < DL; i += 4) {
          const float4 kv = *reinterpret_cast<const float4*>(kr + i);
#pragma unroll
          for (int a = 0; a < QPT; ++a) {
            d[a] = fmaf(q[a][i], kv.x, d[a]); d[a] = fmaf(q[a][i + 1], kv.y, d[a]);
            d[a] = fmaf(q[a][i + 2], kv.z, d[a]); d[a] = fmaf(q[a][i + 3], kv.w, d[a]);
          }
        }
        const int j = c0 + jj;
#pragma unroll
        for (int a = 0; a < QPT; ++a) {
#pragma unroll
          for (int o = 1; o < LPQ; o <<= 1) d[a] += __shfl_xor_sync(0xffffffffu, d[a], o);
          if (band > 0 && j >= tau[a] - band && j <= tau[a] + band) d[a] += -1e4f;
          s[a][u] = (jj < nk) ? d[a] : -3.0e38f;
        }
      }
      float p[QPT][4];
#pragma unroll
      for (int a = 0; a < QPT; ++a) {
        const float mn = fmaxf(fmaxf(fmaxf(s[a][0], s[a][1]), fmaxf(s[a][2], s[a][3])), m[a]);
        if (mn > m[a]) {                         // the running maximum moved (rare after the first groups): rescale
          const float corr = exp2f((m[a] - mn) * kLog2e);    // first group: exp(-3e38 - finite) = 0
          l[a] *= corr;
#pragma unroll
          for (int i = 0; i < DL; ++i) acc[a][i] *= corr;
          m[a] = mn;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) p[a][u] = (jb + u * KP + kp < nk) ? exp2f((s[a][u] - mn) * kLog2e) : 0.f;
        l[a] += (p[a][0] + p[a][1]) + (p[a][2] + p[a][3]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* vr = sv + (jb + u * KP + kp) * KLD + part * DL;
#pragma unroll
        for (int i = 0; i < DL; i += 4) {
          const float4 vv = *reinterpret_cast<const float4*>(vr + i);
#pragma unroll
          for (int a = 0; a < QPT; ++a) {
            acc[a][i] = fmaf(p[a][u], vv.x, acc[a][i]); acc[a][i + 1] = fmaf(p[a][u], vv.y, acc[a][i + 1]);
            acc[a][i + 2] = fmaf(p[a][u], vv.z, acc[a][i + 2]); acc[a][i + 3] = fmaf(p[a][u], vv.w, acc[a][i + 3]);
          }
        }
      }
    }
  }
  if constexpr (KP > 1) {                        // merge the KP partial softmaxes of every query
#pragma unroll
    for (int a = 0; a < QPT; ++a) {
      float mt = m[a];
#pragma unroll
      for (int o = LPQ; o < LPQ * KP; o <<= 1) mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, o));
      const float f = exp2f((m[a] - mt) * kLog2e);           // a set without keys has m = -3e38, l = 0: f = 0 (or 1 * 0)
      float lt = l[a] * f;
#pragma unroll
      for (int o = LPQ; o < LPQ * KP; o <<= 1) lt += __shfl_xor_sync(0xffffffffu, lt, o);
      l[a] = lt;
#pragma unroll
      for (int i = 0; i < DL; ++i) {
        float v = acc[a][i] * f;
#pragma unroll
        for (int o = LPQ; o < LPQ * KP; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        acc[a][i] = v;
      }
    }
  }
#pragma unroll
  for (int a = 0; a < QPT; ++a) {
    const int qi = grp * QPT + a;
    if (qi >= nq || kp != 0) continue;
    const float inv = 1.f / l[a];
    const size_t o = (size_t)(qrow0 + qi) * ldo + hd * D + part * DL;
#pragma unroll
    for (int i = 0; i < DL; i += 4) {
      const float y[4] = {acc[a][i] * inv, acc[a][i + 1] * inv, acc[a][i + 2] * inv, acc[a][i + 3] * inv};
      *reinterpret_cast<float4*>(out + o + i) = make_float4(y[0], y[1], y[2], y[3]);
      if (outg != nullptr) {
        const size_t og = (size_t)(qrow0 + qi) * ldh + hd * D + part * DL + i;
        __half2 g[2], lo[2];
        f16_split2(y[0], y[1], &g[0], &lo[0]);
        f16_split2(y[2], y[3], &g[1], &lo[1]);
        *reinterpret_cast<uint2*>(outg + og) = *reinterpret_cast<const uint2*>(g);
        if (outl != nullptr) *reinterpret_cast<uint2*>(outl + og) = *reinterpret_cast<const uint2*>(lo);
      }
    }
  }
}

template <int D, int LPQ, int QPT, int NT, int KP, int CH>
cudaError_t launch_attention_stream(const float* Q, int ldq, int qcol, const float* K, int ldk, int kcol, const float* V,
                                    int ldv, int vcol, int H, float scale, int rd, const float2* tab, const RowMap& qmap,
                                    const RowMap& kmap, int max_tq, const int* q_tok, const int* q_cu, int band,
                                    float* out, int ldo, __half* outg, __half* outl, int ldh, cudaStream_t st) {
  constexpr int QPB = NT / (LPQ * KP) * QPT;
  const size_t smem = (size_t)(QPB * D + 2 * (CH + 4 * KP) * (D + 4)) * sizeof(float);
  static bool configured = false;
  if (smem > 48 * 1024 && !configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_stream_kernel<D, LPQ, QPT, NT, KP, CH>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  dim3 grid((max_tq + QPB - 1) / QPB, H, qmap.n_seg);
  launch_k(attention_stream_kernel<D, LPQ, QPT, NT, KP, CH>, grid, NT, smem, st, Q, ldq, qcol, K, ldk, kcol, V, ldv, vcol, scale, rd,
                                                                   tab, qmap, kmap, q_tok, q_cu, band, out, ldo, outg,
                                                                   outl, ldh);
  return cudaGetLastError();
}

template <int D, int QB>
cudaError_t launch_attention_tiled(const float* Q, int ldq, int qcol, const float* K, int ldk, int kcol, const float* V,
                                   int ldv, int vcol, int H, float scale, int rd, const float2* tab, const RowMap& qmap,
                                   const RowMap& kmap, int max_tq, int max_tk, const int* q_tok, const int* q_cu,
                                   int band, float* out, int ldo, cudaStream_t st) {
  const int kvn = (kAttChunk * (D + 1) > D * (kAttChunk + 4)) ? kAttChunk * (D + 1) : D * (kAttChunk + 4);
  const size_t smem = (size_t)(QB * D + kvn + QB * ((max_tk + 3) & ~3)) * sizeof(float);
  if (smem > 227 * 1024) return cudaErrorInvalidValue;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_tiled_kernel<D, QB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  dim3 grid((max_tq + QB - 1) / QB, H, qmap.n_seg);
  launch_k(attention_tiled_kernel<D, QB>, grid, 128, smem, st, Q, ldq, qcol, K, ldk, kcol, V, ldv, vcol, scale, rd, tab, qmap,
                                                         kmap, q_tok, q_cu, band, out, ldo);
  return cudaGetLastError();
}

// rd: features of every head that still need the RoPE rotation (0: Q / K are used as stored).  The head sizes of
// the model (16 text encoders, 40 cross attention, 96 / 160 prosody encoders) run the single-pass streaming kernel
// (STTS_ATT_TILED=1 switches back to the two-pass tiled kernel for A/B runs); other head sizes take the generic
// kernel after an in-place rotation pass.  outg / outl: optional fp16 split twins of the output.
cudaError_t attention(const float* Q, int ldq, int qcol, const float* K, int ldk, int kcol, const float* V, int ldv,
                      int vcol, int H, int D, float scale, int rd, const float2* rope_tab, const RowMap& qmap,
                      const RowMap& kmap, int max_tq, int max_tk, const int* q_tok, const int* q_cu, int band, float* out,
                      int ldo, cudaStream_t st, __half* outg, __half* outl, int ldh) {
  if (D % 4 != 0) return cudaErrorInvalidValue;
  static const bool tiled = [] { const char* e = std::getenv("STTS_ATT_TILED"); return e != nullptr && e[0] == '1'; }();
  const bool twins = outg != nullptr;
  if (!tiled && (rd == 0 || rope_tab != nullptr)) {
    // last parameter: keys staged per round trip.  The token-rate sequences are <= ~256 keys: the small heads stage
    // them all at once (one global -> shared round trip and one barrier pair instead of four per CTA)
#define STTS_ATTS(DD, LL, QQ, NN, KK, CC)                                                                             \
  if (D == DD)                                                                                                        \
    return launch_attention_stream<DD, LL, QQ, NN, KK, CC>(Q, ldq, qcol, K, ldk, kcol, V, ldv, vcol, H, scale, rd,    \
                                                           rope_tab, qmap, kmap, max_tq, q_tok, q_cu, band, out, ldo, \
                                                           outg, outl, ldh, st);
    // Small batches (one utterance: 4 .. 32 CTAs with the throughput shapes below, each a serial walk over the keys):
    // one query per lane group and 4 .. 16 key sets, i.e. 8 .. 16 queries per CTA -- four to eight times as many CTAs,
    // each with a quarter of the serial chain.  STTS_ATT_SMALL=<n>: taken when the throughput shape would launch
    // fewer than n CTAs (default 74 = half the SMs; 0 = never).
    static const int small = [] { const char* e = std::getenv("STTS_ATT_SMALL"); return e != nullptr ? std::atoi(e) : 74; }();
    const int qpb_default = (D == 40) ? 128 : 64;
    if (small_shapes_enabled() && (long long)((max_tq + qpb_default - 1) / qpb_default) * H * qmap.n_seg < small) {
      STTS_ATTS(16, 1, 1, 128, 16, 128)  // 8 queries per CTA, 16 key sets
      STTS_ATTS(40, 1, 1, 128, 8, 64)    // 16 frames per CTA, 8 key sets
      STTS_ATTS(96, 4, 1, 256, 8, 64)    // 8 queries per CTA, 4 lanes x 24 features, 8 key sets
      STTS_ATTS(160, 8, 1, 256, 4, 64)   // 8 queries per CTA, 8 lanes x 20 features, 4 key sets
    }
    // measured per launch on the C2 batch (ncu): more key sets / queries per lane group than these are slower
    STTS_ATTS(16, 1, 2, 128, 4, 64)    // text encoders: 64 queries per CTA (2 per lane), 4 key sets          30.5 us
    STTS_ATTS(40, 1, 2, 128, 2, 64)    // cross attention: 128 frames per CTA, 2 key sets                     145 us
    STTS_ATTS(96, 4, 2, 256, 2, 128)   // duration prosody encoder: 64 queries per CTA, 2 key sets             57 us
    STTS_ATTS(160, 8, 4, 256, 2, 64)   // 64 queries per CTA, 8 lanes x 20 features, 4 queries per lane group   78 us
#undef STTS_ATTS
  }
  if (twins) return cudaErrorInvalidValue;       // the fallback kernels do not write twins
  // queries per CTA: as many as keep <= 20 outputs per thread (K / V are staged once per CTA)
#define STTS_ATT(DD, QQ)                                                                                              \
  if (D == DD && (rd == 0 || rope_tab != nullptr))                                                                    \
    return launch_attention_tiled<DD, QQ>(Q, ldq, qcol, K, ldk, kcol, V, ldv, vcol, H, scale, rd, rope_tab, qmap,     \
                                          kmap, max_tq, max_tk, q_tok, q_cu, band, out, ldo, st);
  STTS_ATT(16, 16)
  STTS_ATT(40, 32)
  STTS_ATT(96, 16)
  STTS_ATT(160, 16)
#undef STTS_ATT
  if (rd != 0) {   // generic head size: rotate q / k in place first
    cudaError_t e = rope_inplace(const_cast<float*>(Q), ldq, qcol, H * D, 1, H, D, rd, qmap, st);
    if (e == cudaSuccess) e = rope_inplace(const_cast<float*>(K), ldk, kcol, H * D, 1, H, D, rd, kmap, st);
    if (e != cudaSuccess) return e;
  }
  const size_t smem = (size_t)(kAttQ * D + kAttQ * ((max_tk + 3) & ~3)) * sizeof(float);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  dim3 grid((max_tq + kAttQ - 1) / kAttQ, H, qmap.n_seg);
  launch_k(attention_kernel, grid, 128, smem, st, Q, ldq, qcol, K, ldk, kcol, V, ldv, vcol, D, scale, qmap, kmap, q_tok, q_cu,
                                            band, out, ldo);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ packed rows -> token-major dense
// dst[(cu[seg] + t) * C + c] = src[row(seg, t), c]: the utterances' rows concatenated without gaps
__global__ void __launch_bounds__(256) rows_to_dense_kernel(const float* __restrict__ src, int ld, int C, RowMap map,
                                                            const int* __restrict__ cu, float* __restrict__ dst) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int t0 = tile * 128 - map.seg_start[seg];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    dst[(size_t)(cu[seg] + t0 + lr) * C + c] = src[(size_t)(tile * 128 + lr) * ld + c];
  }
}

cudaError_t rows_to_dense(const float* src, int ld, int C, const RowMap& map, const int* cu, float* dst,
                          cudaStream_t st) {
  launch_k(rows_to_dense_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, kCopySplit, 32)), 256, 0, st, src, ld, C, map, cu, dst);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------ depthwise conv over rows
// y[r, c] = act(b[c] + sum_j w[c, j] * x[r + j - K/2, c]), rows outside the utterance read as zero.
// BasicConvNeXtBlock.dwconv k7 (conv_next.py:18-51) and cross_post[0] k5 + SiLU
// (pitch_energy_predictor.py:56-78).  act: 0 = none, 1 = SiLU.
__global__ void __launch_bounds__(256) dwconv_rows_kernel(const float* __restrict__ x, int ldx, int C, int K,
                                                          const float* __restrict__ w, const float* __restrict__ b,
                                                          RowMap map, int act, float* __restrict__ y, int ldy,
                                                          __half* __restrict__ yg, __half* __restrict__ yl) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int s0 = map.seg_start[seg], s1 = s0 + map.seg_len[seg];
  const int half = K >> 1;
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    const int r = tile * 128 + lr;
    float acc = b[c];
    for (int j = 0; j < K; ++j) {
      const int rr = r + j - half;
      if (rr >= s0 && rr < s1) acc = fmaf(w[c * K + j], x[(size_t)rr * ldx + c], acc);
    }
    if (act == 1) acc = acc / (1.f + expf(-acc));
    y[(size_t)r * ldy + c] = acc;
    if (yg != nullptr) f16_split(acc, yg + (size_t)r * ldy + c, yl + (size_t)r * ldy + c);
  }
}

cudaError_t dwconv_rows(const float* x, int ldx, int C, int K, const float* w, const float* b, const RowMap& map,
                        int act, float* y, int ldy, cudaStream_t st, __half* yg, __half* yl) {
  launch_k(dwconv_rows_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, 4, 64)), 256, 0, st, x, ldx, C, K, w, b, map, act, y, ldy, yg, yl);
  return cudaGetLastError();
}

// token-major dense [sum T, C] -> packed rows, columns [col0, col0 + C)
__global__ void __launch_bounds__(256) dense_to_rows_kernel(const float* __restrict__ src, int C, RowMap map,
                                                            const int* __restrict__ cu, float* __restrict__ dst,
                                                            int ld, int col0, __half* __restrict__ dg,
                                                            __half* __restrict__ dl) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int t0 = tile * 128 - map.seg_start[seg];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    const float v = src[(size_t)(cu[seg] + t0 + lr) * C + c];
    const size_t o = (size_t)(tile * 128 + lr) * ld + col0 + c;
    dst[o] = v;
    if (dg != nullptr) f16_split(v, dg + o, dl + o);
  }
}

cudaError_t dense_to_rows(const float* src, int C, const RowMap& map, const int* cu, float* dst, int ld, int col0,
                          cudaStream_t st, __half* dg, __half* dl) {
  launch_k(dense_to_rows_kernel, dim3(map.n_tiles, split_rows(map.n_tiles, kCopySplit, 32)), 256, 0, st, src, C, map, cu, dst, ld, col0, dg, dl);
  return cudaGetLastError();
}

// dst[row, col0 + j] = vec[seg, j]  (style broadcast to every token: torch.cat([x, style.expand]), prosody_encoder.py:66-67)
__global__ void __launch_bounds__(256) seg_broadcast_kernel(const float* __restrict__ vec, int C, RowMap map,
                                                            float* __restrict__ dst, int ld, int col0,
                                                            __half* __restrict__ dg, __half* __restrict__ dl) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nvalid * C; i += gridDim.y * blockDim.x) {
    const int lr = i / C, c = i - lr * C;
    const float v = vec[(size_t)seg * C + c];
    const size_t o = (size_t)(tile * 128 + lr) * ld + col0 + c;
    dst[o] = v;
    if (dg != nullptr) f16_split(v, dg + o, dl + o);
  }
}

cudaError_t seg_broadcast(const float* vec, int C, const RowMap& map, float* dst, int ld, int col0, cudaStream_t st,
                          __half* dg, __half* dl) {
  launch_k(seg_broadcast_kernel, dim3(map.n_tiles, kCopySplit), 256, 0, st, vec, C, map, dst, ld, col0, dg, dl);
  return cudaGetLastError();
}

}  // namespace stts
