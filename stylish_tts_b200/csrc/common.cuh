// Shared declarations for the sm_100a synthesis kernels.
//
// Activation layout (DESIGN.md "Data layout in HBM"): every frame-rate tensor is a packed,
// time-major, channels-last matrix [R, ld] of fp32 (optionally with a 16-bit twin used as the
// tensor-core operand).  Utterances ("segments") are stacked along the row axis; every segment
// starts on a multiple of TILE_M = 128 rows and is followed by >= GUARD = 16 rows that are
// never written and therefore stay zero.  A conv1d over time is then a row-shifted implicit
// GEMM whose zero padding is provided by the guard rows (or by TMA / bounds-check zero fill at
// the two ends of the buffer), and every 128-row tile belongs to exactly one segment.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

// 16-bit twin of an activation / weight: the operand type of the one-pass tensor-core GEMMs.  IEEE
// fp16 (kind::f16 MMAs run fp16 and bf16 at the same rate; fp16 rounds at 2^-12 instead of 2^-9).
// Conversions saturate at +-65504 instead of producing inf; activations on this path are normalised
// or gated (|x| << 65504), weights are pre-scaled by a power of two per layer (prepack.cu).
typedef __half twin_t;
typedef __half2 twin2_t;
__host__ __device__ __forceinline__ float sat16(float x) { return x > 65504.f ? 65504.f : (x < -65504.f ? -65504.f : x); }
__host__ __device__ __forceinline__ twin_t to_twin(float x) { return __float2half_rn(sat16(x)); }
// pair conversion saturating at +-65504 in ONE instruction (F2FP.SATFINITE.F16.F32.PACK_AB) instead of four FMNMX + a
// pack: the twin stores are 2-3 of the ~12 instructions per element of the issue-bound GEMM epilogues
__device__ __forceinline__ twin2_t to_twin2(float x, float y) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y), "f"(x));
  return *reinterpret_cast<twin2_t*>(&r);
}
#include <stdint.h>

#define STTS_TILE_M 128
#define STTS_GUARD 16

namespace stts {

// ----------------------------------------------------------------------------- programmatic dependent launch
// Every kernel of the library is enqueued through launch_k with the programmatic-stream-serialization attribute and
// starts with pdl_enter(): `griddepcontrol.wait` blocks until the previous kernel of the stream has completed and its
// writes are visible (the ordinary stream guarantee), then `griddepcontrol.launch_dependents` lets the NEXT kernel's
// CTAs be scheduled while this one runs.  So the launch latency, CTA scheduling and pre-wait prologue (barrier init,
// TMEM allocation, tensor-map prefetch in the GEMM kernels) of kernel N + 1 overlap kernel N instead of following it;
// the look-ahead is exactly one kernel per stream (N + 2 cannot start before N + 1 has passed its own wait).  The
// step is ~500 dependent launches of 5-50 us each, also inside a CUDA graph (the capture records programmatic edges).
// No kernel reads or writes global memory that another kernel touches before its pdl_enter().  Measured (B200):
// one 5 s utterance, graph replay 3.64 -> 3.27 ms; with three batches in flight on three streams the look-ahead CTAs
// take SM slots from the other lanes (-1 .. -2 %), so the choice is per context (stts_set_pdl; SynthesizerPool turns it
// off for lanes > 1).  STTS_PDL=0 forces the attribute off (the two instructions are then no-ops).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_wait(); pdl_trigger(); }

bool pdl_enabled();     // pipeline.cu: the calling context's choice (stts_set_pdl), STTS_PDL=0 forces it off
void pdl_set(int on);   // -1: default (on)
// Small-batch kernel shapes (narrower GEMM column tiles, more key sets per attention query when the default shape
// would leave most SMs idle): the calling context's choice (stts_set_small_batch_shapes), default on.  The shapes
// differ in summation order, so results of a small batch and of the same utterances inside a large batch agree to
// rounding, not bit for bit; a caller that needs the latter turns them off.
bool small_shapes_enabled();
void small_shapes_set(int on);   // -1: default (on)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg;
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}

// the same with a thread-block cluster of `cluster_x` CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x,
                                    Args&&... args) {
  cudaLaunchConfig_t cfg;
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)cluster_x;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}

// grid.y of the row-loop kernels (CTAs sharing a 128-row tile through a blockIdx.y-strided loop): `base` for batches
// that fill the GPU anyway, more for small ones -- one utterance is 1-13 tiles, and `base` CTAs per tile made each
// thread walk a serial chain of dependent L2 round trips (depthwise conv of the cross attention: 37 us for 400 frames;
// polar: 45 us).  Same arithmetic per element whatever the split.
inline int split_rows(int n_tiles, int base, int cap) {
  if (n_tiles <= 0) return base;
  const int y = (1184 + n_tiles - 1) / n_tiles;          // ~8 CTAs of 256 threads per SM
  return y < base ? base : (y > cap ? cap : y);
}

// Device view of a packed layout (one per frame rate).
struct RowMap {
  int R;                 // rows (multiple of 128)
  int n_tiles;           // R / 128
  int n_seg;             // utterances
  const int* tile_seg;   // [n_tiles] segment of the tile, -1 = guard-only tile
  const int* tile_nvalid;// [n_tiles] valid rows in the tile (0..128)
  const int* seg_start;  // [n_seg] first row
  const int* seg_len;    // [n_seg] valid rows
};

// ----------------------------------------------------------------------------- epilogues
enum EpiKind : int {
  EPI_BIAS = 0,        // out = acc + bias
  EPI_RES_SCALE = 1,   // out = (acc + bias + res) * scale         (AdaptiveDecoderBlock, ada_norm.py:179-182)
  EPI_SILU = 2,        // out = silu(acc + bias)                    (ConvNeXtBlock pwconv1, generator.py:478-479)
  EPI_ADD = 3,         // out = res + acc + bias                    (ConvNeXt residual, generator.py:484)
  EPI_GATE = 4,        // out[i] = tanh(a_2i + g_i) * sigmoid(a_2i+1 + g_{H+i})   (flow.py:7-14)
  EPI_RESSKIP = 5,     // n<split: out[n] += v ; n>=split: out2[n-split] (+)= v    (flow.py:80-87)
  EPI_COUPLE = 6,      // out[i] = (out[i] - m_i) * exp(-logs_i), (m,logs) interleaved (flow.py:210-213)
  EPI_PRIOR = 7,       // out[i] = mean_i + noise[i] * exp(logstd_i)               (flow.py:311-315)
  EPI_STFT = 8,        // (re,im) interleaved -> out = |X|, out2 = atan2            (generator.py:32-44,407-410)
  EPI_RELU = 9,        // out = relu(acc + bias)                    (FFN conv_1, text_encoder.py:305-329)
  EPI_GELU = 10,       // out = gelu(acc + bias), exact erf form   (BasicConvNeXtBlock, conv_next.py:18-51)
  EPI_LN = 11,         // out = act(affine(LayerNorm_row(acc + bias + res)))  tcgen05 kernels only, N <= 256: the
                       //   channel LayerNorm / AdaptiveLayerNorm that follows the attention / FFN projections of the
                       //   token-rate encoders (text_encoder.py:171-221, prosody_encoder.py:33-60), see ln_epilogue
};

struct Epi {
  int kind;
  int N;                       // logical output columns of the GEMM (pairs count as 2)
  float* out; int ldo; int col0;
  twin_t* outh; int ldh; int colh0;          // optional 16-bit twin of `out`
  float* out2; int ldo2; int col20;
  twin_t* outh2; int ldh2; int colh20;       // optional 16-bit twin of `out2`
  // optional fp16 split twins of out / out2 (same ldh / colh geometry): g = fp16(x), l = fp16(x - g).
  // x = g + l to ~2^-22 relative: operands of the f16x3 island GEMMs (g alone: the one-pass fp16 GEMM)
  __half* outg; __half* outl; __half* outg2; __half* outl2;
  const float* bias;           // [N] (already permuted for the pair kinds) or null
  const float* res; int ldr; int rcol0;
  const float* segvec; int ldseg;            // per-utterance vector, row = segment
  float scale;
  int split;                   // EPI_RESSKIP
  int accumulate2;             // EPI_RESSKIP: out2 += (1) or = (0)
  const int* seg_start;        // EPI_STFT: first row of every segment (frame 0 is real by symmetry)
  // EPI_SILU, tcgen05 kernel only: per-(tile, row quarter) column sums of squares of the activated output,
  // [n_tiles * 4, colss_ld] fp32 -- the L2-over-time statistic of GRN without re-reading the tensor
  float* colss; int colss_ld;
  // EPI_BIAS / EPI_RES_SCALE, tcgen05 kernel only: per-(tile, row quarter) column sums and sums of squares of
  // the stored output, [n_tiles * 4, colst_ld] fp32 each -- InstanceNorm statistics of the next AdaIN without
  // re-reading the tensor (16-bit mode)
  float* colsum; float* colsq; int colst_ld;
  // EPI_LN: ln_seg_stride == 0: affine = ln_g[n] * x + ln_b[n]; > 0: (1 + g[seg, n]) * x + g[seg, N + n] with
  // g = ln_g + seg * ln_seg_stride (AdaptiveLayerNorm, the style FC row of the utterance)
  const float* ln_g; const float* ln_b; int ln_seg_stride; float ln_eps; int ln_relu;
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
// fp16 split of x: g = fp16(x), l = fp16(x - g); x = g + l to ~2^-22 relative (|x| < 65504; below
// 2^-3 the residual is an fp16 subnormal: absolute error <= 2^-25).  Operand split of the f16x3 GEMMs.
__device__ __forceinline__ void f16_split(float x, __half* g, __half* l) {
  x = fminf(fmaxf(x, -65504.f), 65504.f);    // saturate instead of inf (NaN passes through)
  const __half h = __float2half_rn(x);
  *g = h;
  *l = __float2half_rn(x - __half2float(h));
}

// the same for a pair, bit-identical to two f16_split calls: the packed conversion (F2FP) and the half2 -> float2
// unpack run on the main pipes, the scalar F2F conversions on the quarter-rate one (six of them per pair made the
// store loop of the LayerNorm epilogue conversion-bound)
__device__ __forceinline__ void f16_split2(float x0, float x1, __half2* g, __half2* l) {
  x0 = fminf(fmaxf(x0, -65504.f), 65504.f);
  x1 = fminf(fmaxf(x1, -65504.f), 65504.f);
  const __half2 h = __floats2half2_rn(x0, x1);
  const float2 hf = __half22float2(h);
  *g = h;
  *l = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
}

// One thread owns row r (valid, segment seg) and NV consecutive output columns starting at n0
// (n0 % NV == 0, NV even).  Shared by the SIMT and the tcgen05 kernels.
template <int KIND, int NV>
__device__ __forceinline__ void epi_apply(const Epi& e, int r, int seg, int n0, const float* acc) {
  if constexpr (KIND == EPI_BIAS || KIND == EPI_RES_SCALE || KIND == EPI_SILU || KIND == EPI_ADD ||
                KIND == EPI_RELU || KIND == EPI_GELU) {
    float v[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int n = n0 + i;
      float x = acc[i] + ((e.bias != nullptr && n < e.N) ? e.bias[n] : 0.f);
      if constexpr (KIND == EPI_RES_SCALE) {
        if (e.res != nullptr && n < e.N) x += e.res[(size_t)r * e.ldr + e.rcol0 + n];
        x *= e.scale;
      }
      if constexpr (KIND == EPI_SILU) x = x / (1.f + expf(-x));
      if constexpr (KIND == EPI_RELU) x = fmaxf(x, 0.f);
      if constexpr (KIND == EPI_GELU) x = 0.5f * x * (1.f + erff(x * 0.70710678118654752440f));
      if constexpr (KIND == EPI_ADD) {
        if (n < e.N) x += e.res[(size_t)r * e.ldr + e.rcol0 + n];
      }
      v[i] = x;
    }
    float* o = e.out + (size_t)r * e.ldo + e.col0 + n0;
    if (n0 + NV <= e.N && ((((size_t)r * e.ldo + e.col0 + n0) & 3) == 0) && (NV % 4 == 0)) {
#pragma unroll
      for (int i = 0; i < NV; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < NV; ++i) if (n0 + i < e.N) o[i] = v[i];
    }
    if (e.outh != nullptr) {
      twin_t* oh = e.outh + (size_t)r * e.ldh + e.colh0 + n0;
#pragma unroll
      for (int i = 0; i < NV; ++i) if (n0 + i < e.N) oh[i] = to_twin(v[i]);
    }
  } else if constexpr (KIND == EPI_GATE) {
    const int H = e.N >> 1;
#pragma unroll
    for (int i = 0; i < NV; i += 2) {
      const int n = n0 + i;
      if (n < e.N) {
        const int c = n >> 1;
        const float a = acc[i] + e.bias[n] + e.segvec[(size_t)seg * e.ldseg + c];
        const float b = acc[i + 1] + e.bias[n + 1] + e.segvec[(size_t)seg * e.ldseg + H + c];
        const float y = tanhf(a) * (1.f / (1.f + expf(-b)));
        e.out[(size_t)r * e.ldo + e.col0 + c] = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      }
    }
  } else if constexpr (KIND == EPI_RESSKIP) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int n = n0 + i;
      if (n < e.N) {
        const float v = acc[i] + e.bias[n];
        if (n < e.split) {
          float* p = e.out + (size_t)r * e.ldo + e.col0 + n;
          const float y = *p + v;
          *p = y;
          if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + n] = to_twin(y);
        } else {
          float* p = e.out2 + (size_t)r * e.ldo2 + e.col20 + (n - e.split);
          const float y = e.accumulate2 ? (*p + v) : v;
          *p = y;
          if (e.outh2 != nullptr) e.outh2[(size_t)r * e.ldh2 + e.colh20 + (n - e.split)] = to_twin(y);
        }
      }
    }
  } else if constexpr (KIND == EPI_COUPLE) {
#pragma unroll
    for (int i = 0; i < NV; i += 2) {
      const int n = n0 + i;
      if (n < e.N) {
        const int c = n >> 1;
        const float m = acc[i] + e.bias[n];
        const float logs = acc[i + 1] + e.bias[n + 1];
        float* p = e.out + (size_t)r * e.ldo + e.col0 + c;
        const float y = (*p - m) * expf(-logs);
        *p = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      }
    }
  } else if constexpr (KIND == EPI_PRIOR) {
#pragma unroll
    for (int i = 0; i < NV; i += 2) {
      const int n = n0 + i;
      if (n < e.N) {
        const int c = n >> 1;
        const float mean = acc[i] + e.bias[n];
        const float logstd = acc[i + 1] + e.bias[n + 1];
        const float nz = e.res[(size_t)r * e.ldr + e.rcol0 + c];
        const float y = mean + nz * expf(logstd);
        e.out[(size_t)r * e.ldo + e.col0 + c] = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      }
    }
  } else if constexpr (KIND == EPI_STFT) {
#pragma unroll
    for (int i = 0; i < NV; i += 2) {
      const int n = n0 + i;
      if (n < e.N) {
        const int c = n >> 1;
        // Frame 0 of an utterance is centred on sample 0 of the reflect-padded signal, so its
        // windowed frame is even about the centre and X_k is exactly real: the reference's Im is
        // pure FFT rounding noise there (random +-pi for Re < 0).  Return the exact value.
        const float re = acc[i], im = (e.seg_start != nullptr && r == e.seg_start[seg]) ? 0.f : acc[i + 1];
        const float mag = hypotf(re, im);
        const float d = mag + 1e-9f;
        const float ph = atan2f(im / d, re / d);
        e.out[(size_t)r * e.ldo + e.col0 + c] = mag;
        e.out2[(size_t)r * e.ldo2 + e.col20 + c] = ph;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(mag);
        if (e.outh2 != nullptr) e.outh2[(size_t)r * e.ldh2 + e.colh20 + c] = to_twin(ph);
      }
    }
  }
}

// ----------------------------------------------------------------------------- coalesced epilogue
// Lane-per-column-pair form used by the tcgen05 kernel after the accumulator tile has been
// transposed through shared memory: one lane owns accumulator columns (n, n + 1), n even, and
// walks down the rows of its warp's 32-row slice, so every global access of a warp is one
// contiguous 128 / 256-byte segment of a row.  Column constants (bias, style vector) are loaded
// once per 64-column slab.
template <int KIND, bool FAST>
struct EpiPair {
  static __device__ __forceinline__ float ex(float x) { if constexpr (FAST) return __expf(x); else return expf(x); }
  static __device__ __forceinline__ float th(float x) {
    if constexpr (FAST) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; } else return tanhf(x);
  }
  static __device__ __forceinline__ float dv(float a, float b) { if constexpr (FAST) return __fdividef(a, b); else return a / b; }
  float b0, b1, g0, g1;
  float ss0, ss1;  // EPI_SILU: running sum of squares of the two columns over the slab's rows
  int n;           // even accumulator column
  bool v0, v1;     // column n / n + 1 inside the logical N
  __device__ __forceinline__ void init(const Epi& e, int seg, int n_) {
    n = n_;
    v0 = n < e.N; v1 = n + 1 < e.N;
    b0 = (e.bias != nullptr && v0) ? e.bias[n] : 0.f;
    b1 = (e.bias != nullptr && v1) ? e.bias[n + 1] : 0.f;
    g0 = g1 = 0.f;
    ss0 = ss1 = 0.f;
    if constexpr (KIND == EPI_GATE) {
      if (v0) {
        g0 = e.segvec[(size_t)seg * e.ldseg + (n >> 1)];
        g1 = e.segvec[(size_t)seg * e.ldseg + (e.N >> 1) + (n >> 1)];
      }
    }
  }
  // Kinds that read memory (residual / read-modify-write) prefetch it for all rows of the slab
  // before the accumulators are touched: 32 independent 256-byte loads in flight per warp.
  static constexpr bool kHasPre = (KIND == EPI_RES_SCALE || KIND == EPI_ADD || KIND == EPI_RESSKIP ||
                                   KIND == EPI_COUPLE || KIND == EPI_PRIOR);
  __device__ __forceinline__ float2 pre(const Epi& e, int r) const {
    float2 t = make_float2(0.f, 0.f);
    if (!v0) return t;
    if constexpr (KIND == EPI_RES_SCALE || KIND == EPI_ADD) {
      if (e.res != nullptr) {
        const float* rp = e.res + (size_t)r * e.ldr + e.rcol0 + n;
        if (v1) t = *reinterpret_cast<const float2*>(rp); else t.x = rp[0];
      }
    } else if constexpr (KIND == EPI_RESSKIP) {
      if (n < e.split) {
        const float* p = e.out + (size_t)r * e.ldo + e.col0 + n;
        if (v1) t = *reinterpret_cast<const float2*>(p); else t.x = p[0];
      } else if (e.accumulate2) {
        const float* p = e.out2 + (size_t)r * e.ldo2 + e.col20 + (n - e.split);
        if (v1) t = *reinterpret_cast<const float2*>(p); else t.x = p[0];
      }
    } else if constexpr (KIND == EPI_COUPLE) {
      t.x = e.out[(size_t)r * e.ldo + e.col0 + (n >> 1)];
    } else if constexpr (KIND == EPI_PRIOR) {
      t.x = e.res[(size_t)r * e.ldr + e.rcol0 + (n >> 1)];
    }
    return t;
  }
  static __device__ __forceinline__ void store2(float* o, twin_t* oh, float x0, float x1, bool both,
                                                __half* og = nullptr, __half* ol = nullptr) {
    if (both) {
      if (o != nullptr) *reinterpret_cast<float2*>(o) = make_float2(x0, x1);
      if (oh != nullptr) *reinterpret_cast<twin2_t*>(oh) = to_twin2(x0, x1);
      if (og != nullptr) {
        const __half2 g = __floats2half2_rn(x0, x1);
        *reinterpret_cast<__half2*>(og) = g;
        if (ol != nullptr)
          *reinterpret_cast<__half2*>(ol) = __floats2half2_rn(x0 - __low2float(g), x1 - __high2float(g));
      }
    } else {
      if (o != nullptr) o[0] = x0;
      if (oh != nullptr) oh[0] = to_twin(x0);
      if (og != nullptr) { __half lo; f16_split(x0, og, ol != nullptr ? ol : &lo); }
    }
  }
  // after the last row of the slab: per-(tile, quarter) column statistics
  __device__ __forceinline__ void finish(const Epi& e, int tile, int q) const {
    if constexpr (KIND == EPI_SILU) {
      if (e.colss != nullptr && v0) {
        float* p = e.colss + (size_t)(tile * 4 + q) * e.colss_ld + n;
        p[0] = ss0;
        if (v1) p[1] = ss1;
      }
    }
  }
  __device__ __forceinline__ void apply(const Epi& e, int r, int seg, int seg_first_row, float a0, float a1,
                                        float2 t) {
    if (!v0) return;
    if constexpr (KIND == EPI_BIAS || KIND == EPI_RES_SCALE || KIND == EPI_SILU || KIND == EPI_ADD ||
                  KIND == EPI_RELU || KIND == EPI_GELU) {
      float x0 = a0 + b0, x1 = a1 + b1;
      if constexpr (KIND == EPI_RES_SCALE) {
        x0 = (x0 + t.x) * e.scale; x1 = (x1 + t.y) * e.scale;
      }
      if constexpr (KIND == EPI_SILU) {
        x0 = dv(x0, 1.f + ex(-x0)); x1 = dv(x1, 1.f + ex(-x1));
        ss0 = fmaf(x0, x0, ss0); ss1 = fmaf(x1, x1, ss1);
      }
      if constexpr (KIND == EPI_ADD) { x0 += t.x; x1 += t.y; }
      if constexpr (KIND == EPI_RELU) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); }
      if constexpr (KIND == EPI_GELU) {
        x0 = 0.5f * x0 * (1.f + erff(x0 * 0.70710678118654752440f));
        x1 = 0.5f * x1 * (1.f + erff(x1 * 0.70710678118654752440f));
      }
      store2(e.out ? e.out + (size_t)r * e.ldo + e.col0 + n : nullptr,
             e.outh ? e.outh + (size_t)r * e.ldh + e.colh0 + n : nullptr, x0, x1, v1,
             e.outg ? e.outg + (size_t)r * e.ldh + e.colh0 + n : nullptr,
             (e.outg && e.outl) ? e.outl + (size_t)r * e.ldh + e.colh0 + n : nullptr);
    } else if constexpr (KIND == EPI_RESSKIP) {
      const float x0 = a0 + b0, x1 = a1 + b1;
      if (n < e.split) {
        float* p = e.out + (size_t)r * e.ldo + e.col0 + n;
        const float y0 = t.x + x0, y1 = t.y + x1;
        store2(p, e.outh ? e.outh + (size_t)r * e.ldh + e.colh0 + n : nullptr, y0, y1, v1);
      } else {
        const int m = n - e.split;
        float* p = e.out2 + (size_t)r * e.ldo2 + e.col20 + m;
        const float y0 = x0 + t.x, y1 = x1 + t.y;   // t = 0 when not accumulating
        store2(p, e.outh2 ? e.outh2 + (size_t)r * e.ldh2 + e.colh20 + m : nullptr, y0, y1, v1);
      }
    } else {
      const int c = n >> 1;   // pair kinds: one output column per accumulator pair
      if constexpr (KIND == EPI_GATE) {
        const float a = a0 + b0 + g0, b = a1 + b1 + g1;
        const float y = th(a) * dv(1.f, 1.f + ex(-b));
        if (e.out != nullptr) e.out[(size_t)r * e.ldo + e.col0 + c] = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      } else if constexpr (KIND == EPI_COUPLE) {
        const float m = a0 + b0, logs = a1 + b1;
        float* p = e.out + (size_t)r * e.ldo + e.col0 + c;
        const float y = (t.x - m) * ex(-logs);
        *p = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      } else if constexpr (KIND == EPI_PRIOR) {
        const float mean = a0 + b0, logstd = a1 + b1;
        const float y = mean + t.x * ex(logstd);
        e.out[(size_t)r * e.ldo + e.col0 + c] = y;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(y);
      } else if constexpr (KIND == EPI_STFT) {
        // frame 0 of an utterance is exactly real (see epi_apply)
        const float re = a0, im = (r == seg_first_row) ? 0.f : a1;
        const float mag = hypotf(re, im);
        const float d = mag + 1e-9f;
        const float ph = atan2f(im / d, re / d);
        if (e.out != nullptr) e.out[(size_t)r * e.ldo + e.col0 + c] = mag;
        if (e.out2 != nullptr) e.out2[(size_t)r * e.ldo2 + e.col20 + c] = ph;
        if (e.outh != nullptr) e.outh[(size_t)r * e.ldh + e.colh0 + c] = to_twin(mag);
        if (e.outh2 != nullptr) e.outh2[(size_t)r * e.ldh2 + e.colh20 + c] = to_twin(ph);
        __half lo;
        if (e.outg != nullptr)
          f16_split(mag, e.outg + (size_t)r * e.ldh + e.colh0 + c,
                    e.outl != nullptr ? e.outl + (size_t)r * e.ldh + e.colh0 + c : &lo);
        if (e.outg2 != nullptr)
          f16_split(ph, e.outg2 + (size_t)r * e.ldh2 + e.colh20 + c,
                    e.outl2 != nullptr ? e.outl2 + (size_t)r * e.ldh2 + e.colh20 + c : &lo);
      }
    }
  }
};

// Lane-per-4-columns form of the "simple" kinds (bias / residual / activation): lanes 0-15 cover the 64
// columns of one row of the slab with 16-byte accesses, lanes 16-31 the next row, so one warp instruction
// moves two full row segments and the per-row pointer arithmetic is shared by four columns (the
// lane-per-pair form was instruction-issue-bound: ncu, profiles/README.md).  Falls back to guarded scalar
// accesses where the slab crosses the logical N or the destination is not 16-byte aligned.
template <int KIND, bool FAST>
struct EpiQuad {
  static constexpr bool kOk = (KIND == EPI_BIAS || KIND == EPI_RES_SCALE || KIND == EPI_SILU || KIND == EPI_ADD ||
                               KIND == EPI_RELU || KIND == EPI_GELU);
  static constexpr bool kHasPre = (KIND == EPI_RES_SCALE || KIND == EPI_ADD);
  using P = EpiPair<KIND, FAST>;
  static constexpr bool kStats = (KIND == EPI_BIAS || KIND == EPI_RES_SCALE);
  float bias[4];
  float ss[4];     // running sums of squares (EPI_SILU: GRN; kStats kinds: InstanceNorm)
  float sm[4];     // running sums (kStats kinds)
  bool st;         // statistics requested
  int n, nv;       // first column of this lane, number of its columns inside the logical N (0 .. 4)
  bool vec;        // all four columns valid and every destination / source row segment 16-byte aligned
  __device__ __forceinline__ void init(const Epi& e, int n_) {
    n = n_;
    nv = e.N - n; nv = nv < 0 ? 0 : (nv > 4 ? 4 : nv);
#pragma unroll
    for (int i = 0; i < 4; ++i) { bias[i] = (e.bias != nullptr && i < nv) ? e.bias[n + i] : 0.f; ss[i] = 0.f; sm[i] = 0.f; }
    st = kStats && e.colsum != nullptr;
    auto al = [](const void* base, int ld, int col, int elem) {
      return base == nullptr || ((((uintptr_t)base + (size_t)col * elem) & (elem * 4 - 1)) == 0 && (ld & 3) == 0);
    };
    vec = nv == 4 && al(e.out, e.ldo, e.col0 + n, 4) && al(e.outh, e.ldh, e.colh0 + n, 2) &&
          al(e.outg, e.ldh, e.colh0 + n, 2) && al(e.outl, e.ldh, e.colh0 + n, 2) &&
          (!kHasPre || al(e.res, e.ldr, e.rcol0 + n, 4));
  }
  // ---- vector path (vec == true): row pointers are set up once per slab (bind) and advanced by a constant
  // stride, so an iteration is one 16-byte staging read, the math and the stores -- no per-row 64-bit
  // pointer arithmetic, no null / range checks (they cost more than the math: ~150 instructions per row pair)
  float* p_o; twin_t* p_oh; __half* p_og; __half* p_ol; const float* p_r;
  long long s_o, s_h, s_r;     // element strides between consecutive iterations of this lane (2 rows)
  float scale_;
  __device__ __forceinline__ void bind(const Epi& e, int r_first) {
    p_o = e.out != nullptr ? e.out + (size_t)r_first * e.ldo + e.col0 + n : nullptr;
    p_oh = e.outh != nullptr ? e.outh + (size_t)r_first * e.ldh + e.colh0 + n : nullptr;
    p_og = e.outg != nullptr ? e.outg + (size_t)r_first * e.ldh + e.colh0 + n : nullptr;
    p_ol = (e.outg != nullptr && e.outl != nullptr) ? e.outl + (size_t)r_first * e.ldh + e.colh0 + n : nullptr;
    p_r = (kHasPre && e.res != nullptr) ? e.res + (size_t)r_first * e.ldr + e.rcol0 + n : nullptr;
    s_o = 2LL * e.ldo; s_h = 2LL * e.ldh; s_r = 2LL * e.ldr;
    scale_ = e.scale;
  }
  __device__ __forceinline__ float4 pre_vec(int k) const {
    if constexpr (kHasPre) {
      if (p_r != nullptr) return *reinterpret_cast<const float4*>(p_r + k * s_r);
    }
    return make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __device__ __forceinline__ void apply_vec(int k, float4 a, float4 t) {
    float x[4] = {a.x + bias[0], a.y + bias[1], a.z + bias[2], a.w + bias[3]};
    const float tt[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (KIND == EPI_RES_SCALE) x[i] = (x[i] + tt[i]) * scale_;
      if constexpr (KIND == EPI_SILU) { x[i] = P::dv(x[i], 1.f + P::ex(-x[i])); ss[i] = fmaf(x[i], x[i], ss[i]); }
      if constexpr (KIND == EPI_ADD) x[i] += tt[i];
      if constexpr (KIND == EPI_RELU) x[i] = fmaxf(x[i], 0.f);
      if constexpr (KIND == EPI_GELU) x[i] = 0.5f * x[i] * (1.f + erff(x[i] * 0.70710678118654752440f));
    }
    if constexpr (kStats) {
      if (st) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { sm[i] += x[i]; ss[i] = fmaf(x[i], x[i], ss[i]); }
      }
    }
    if (p_o != nullptr) *reinterpret_cast<float4*>(p_o + k * s_o) = make_float4(x[0], x[1], x[2], x[3]);
    if (p_oh != nullptr) {
      const twin2_t t0 = to_twin2(x[0], x[1]), t1 = to_twin2(x[2], x[3]);
      uint2 w;
      w.x = *reinterpret_cast<const uint32_t*>(&t0); w.y = *reinterpret_cast<const uint32_t*>(&t1);
      *reinterpret_cast<uint2*>(p_oh + k * s_h) = w;
    }
    if (p_og != nullptr) {
      const __half2 g0 = to_twin2(x[0], x[1]), g1 = to_twin2(x[2], x[3]);
      uint2 w;
      w.x = *reinterpret_cast<const uint32_t*>(&g0); w.y = *reinterpret_cast<const uint32_t*>(&g1);
      *reinterpret_cast<uint2*>(p_og + k * s_h) = w;
      if (p_ol != nullptr) {
        const __half2 l0 = __floats2half2_rn(sat16(x[0]) - __low2float(g0), sat16(x[1]) - __high2float(g0));
        const __half2 l1 = __floats2half2_rn(sat16(x[2]) - __low2float(g1), sat16(x[3]) - __high2float(g1));
        w.x = *reinterpret_cast<const uint32_t*>(&l0); w.y = *reinterpret_cast<const uint32_t*>(&l1);
        *reinterpret_cast<uint2*>(p_ol + k * s_h) = w;
      }
    }
  }
  __device__ __forceinline__ float4 pre(const Epi& e, int r) const {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if constexpr (kHasPre) {
      if (e.res == nullptr || nv == 0) return t;
      const float* rp = e.res + (size_t)r * e.ldr + e.rcol0 + n;
      if (vec) return *reinterpret_cast<const float4*>(rp);
      t.x = rp[0];
      if (nv > 1) t.y = rp[1];
      if (nv > 2) t.z = rp[2];
      if (nv > 3) t.w = rp[3];
    }
    return t;
  }
  __device__ __forceinline__ void apply(const Epi& e, int r, float4 a, float4 t) {
    if (nv == 0) return;
    float x[4] = {a.x + bias[0], a.y + bias[1], a.z + bias[2], a.w + bias[3]};
    const float tt[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if constexpr (KIND == EPI_RES_SCALE) x[i] = (x[i] + tt[i]) * e.scale;
      if constexpr (KIND == EPI_SILU) { x[i] = P::dv(x[i], 1.f + P::ex(-x[i])); ss[i] = fmaf(x[i], x[i], ss[i]); }
      if constexpr (KIND == EPI_ADD) x[i] += tt[i];
      if constexpr (KIND == EPI_RELU) x[i] = fmaxf(x[i], 0.f);
      if constexpr (KIND == EPI_GELU) x[i] = 0.5f * x[i] * (1.f + erff(x[i] * 0.70710678118654752440f));
    }
    if constexpr (kStats) {
      if (st) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { sm[i] += x[i]; ss[i] = fmaf(x[i], x[i], ss[i]); }
      }
    }
    float* o = e.out != nullptr ? e.out + (size_t)r * e.ldo + e.col0 + n : nullptr;
    twin_t* oh = e.outh != nullptr ? e.outh + (size_t)r * e.ldh + e.colh0 + n : nullptr;
    __half* og = e.outg != nullptr ? e.outg + (size_t)r * e.ldh + e.colh0 + n : nullptr;
    __half* ol = (e.outg != nullptr && e.outl != nullptr) ? e.outl + (size_t)r * e.ldh + e.colh0 + n : nullptr;
    if (vec) {
      if (o != nullptr) *reinterpret_cast<float4*>(o) = make_float4(x[0], x[1], x[2], x[3]);
      if (oh != nullptr) {
        const twin2_t t0 = to_twin2(x[0], x[1]), t1 = to_twin2(x[2], x[3]);
        uint2 w;
        w.x = *reinterpret_cast<const uint32_t*>(&t0); w.y = *reinterpret_cast<const uint32_t*>(&t1);
        *reinterpret_cast<uint2*>(oh) = w;
      }
      if (og != nullptr) {
        const __half2 g0 = to_twin2(x[0], x[1]), g1 = to_twin2(x[2], x[3]);
        uint2 w;
        w.x = *reinterpret_cast<const uint32_t*>(&g0); w.y = *reinterpret_cast<const uint32_t*>(&g1);
        *reinterpret_cast<uint2*>(og) = w;
        if (ol != nullptr) {
          const __half2 l0 = __floats2half2_rn(sat16(x[0]) - __low2float(g0), sat16(x[1]) - __high2float(g0));
          const __half2 l1 = __floats2half2_rn(sat16(x[2]) - __low2float(g1), sat16(x[3]) - __high2float(g1));
          w.x = *reinterpret_cast<const uint32_t*>(&l0); w.y = *reinterpret_cast<const uint32_t*>(&l1);
          *reinterpret_cast<uint2*>(ol) = w;
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (i < nv) {
          if (o != nullptr) o[i] = x[i];
          if (oh != nullptr) oh[i] = to_twin(x[i]);
          if (og != nullptr) { __half lo; f16_split(x[i], og + i, ol != nullptr ? ol + i : &lo); }
        }
      }
    }
  }
  // per-(tile, quarter) column sums of squares (EPI_SILU): the two row-interleaved halves of the warp are
  // combined with one shuffle; lanes 0-15 write
  __device__ __forceinline__ void finish(const Epi& e, int tile, int q, int lane) {
    if constexpr (kStats) {
      if (e.colsum == nullptr) return;          // warp-uniform
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        sm[i] += __shfl_xor_sync(0xffffffffu, sm[i], 16);
        ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 16);
      }
      if (lane < 16) {
        const size_t o = (size_t)(tile * 4 + q) * e.colst_ld + n;
#pragma unroll
        for (int i = 0; i < 4; ++i) if (i < nv) { e.colsum[o + i] = sm[i]; e.colsq[o + i] = ss[i]; }
      }
    }
    if constexpr (KIND == EPI_SILU) {
      if (e.colss == nullptr) return;
#pragma unroll
      for (int i = 0; i < 4; ++i) ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 16);
      if (lane < 16) {
        float* p = e.colss + (size_t)(tile * 4 + q) * e.colss_ld + n;
#pragma unroll
        for (int i = 0; i < 4; ++i) if (i < nv) p[i] = ss[i];
      }
    }
  }
};

// ----------------------------------------------------------------------------- conv-as-GEMM
// Y[r, n] = sum_s sum_j sum_c A_s[r + j * dil_s - pad_s, col0_s + c] * W[n, koff_s + j * cin_pad_s + c]
// Up to two K-segments (fused 1x1 shortcut, channel concatenation across buffers).
struct KSeg {
  const float* A; int lda; int col0;         // fp32 operand
  const twin_t* Ah; int ldah; int colh0;     // 16-bit twin (one-pass tensor-core mode)
  const __half* Ag; const __half* Al; int ldag;  // fp16 split twins g / l (f16 and f16x3 modes), columns as colh0
  int taps; int pad; int cin_pad;            // cin_pad % 64 == 0
  int dil;                                   // tap j reads row r + j * dil - pad; 0 is read as 1 (no dilation)
  int rows;                                  // rows allocated in A (bounds for zero fill)
};

struct ConvGemm {
  KSeg seg[2];
  int n_kseg;
  const float* W;                 // [Npad, Ktot] fp32, K-major
  const twin_t* Wh_seg;           // optional per-utterance fp16 weights [n_seg * Npad, Ktot], unscaled (GRN folded into pw2)
  const __half* Wg;               // fp16(W * wscale)                       (f16 / f16x3 modes)
  const __half* Wl;               // fp16(W * wscale - Wg)                  (f16x3 mode)
  float acc_scale;                // accumulators are multiplied by this before the epilogue (1 / wscale / a-scale)
  int f16;                        // one-pass GEMM on the island twins (Ag) instead of the body twins (Ah)
  int Ktot;                       // sum taps*cin_pad
  int Npad;                       // multiple of 128
  Epi epi;
};

// ----------------------------------------------------------------------------- small helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace stts
