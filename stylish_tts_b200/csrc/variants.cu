// the source / activation variants `north_star` names that sit next to the live path.
//   * Snake AdaptiveGeneratorBlock   train/models/ada_norm.py:11-127  (AdaIN -> x + sin^2(a x) / a -> dilated conv, x3)
//   * SineGenerator                  train/models/cfm/cfm_mel_decoder.py:16-101  (second half of this file)
// Both are stand-alone operators behind their own C-ABI entry (stts_generator_block_snake, stts_source_sinegen): the
// default model.yml never calls them, so they own a small layout + workspace (VariantWork) instead of a slice of the
// synthesis arena.  The block reuses the packed time-major layout and the conv-as-GEMM kernels of the live path
// (row step `dil` between taps); its zero guard after every utterance is 32 rows (>= the widest padding, k 11 d 5 = 25).
#include <algorithm>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "engine.h"

namespace stts {

struct VariantWork {
  // ---- snake block
  Layout L;
  std::vector<int> sig;           // n_seg, C, lengths
  char* arena = nullptr;
  size_t arena_bytes = 0;
  Act X[2], ACT, Y1;
  double* stats = nullptr;
  float *mean = nullptr, *rstd = nullptr, *gb = nullptr;
  long long* d_off = nullptr;
  int* d_cs = nullptr;
  // ---- sinegen
  char* sg = nullptr;
  size_t sg_bytes = 0;
  // ---- conv-DFT STFT modules by name
  std::map<std::string, struct StftConvPlan*> stft;
};
void free_stft_plans(VariantWork& w);

void free_variant_work(Context& c) {
  if (c.vwork == nullptr) return;
  if (c.vwork->arena) cudaFree(c.vwork->arena);
  if (c.vwork->sg) cudaFree(c.vwork->sg);
  free_stft_plans(*c.vwork);
  delete c.vwork;
  c.vwork = nullptr;
}

namespace {

constexpr int kSnakeGuard = 32;

VariantWork& work(Context& c) {
  if (c.vwork == nullptr) c.vwork = new VariantWork();
  return *c.vwork;
}

// AdaptiveInstance (ada_norm.py:130-140) followed by Snake1D (ada_norm.py:114, :117):
//   n = (1 + gamma) * ((x - mean) * rstd) + beta ;  y = n + (1 / a) * sin(a n)^2
// Same thread mapping as adain_act_kernel (elementwise.cu): warp accesses are 256 contiguous bytes of one row.
__global__ void __launch_bounds__(256) adain_snake_kernel(const float* __restrict__ x, int ldx, int C, RowMap map,
                                                          const float* __restrict__ mean, const float* __restrict__ rstd,
                                                          const float* __restrict__ gb, int ldgb,
                                                          const float* __restrict__ alpha, float* __restrict__ y, int ldy,
                                                          twin_t* __restrict__ yh, int ldyh) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int c = blockIdx.y * 128 + 2 * (threadIdx.x & 63);
  if (c >= C) return;
  const int rg = threadIdx.x >> 6;
  const float2 mu = *reinterpret_cast<const float2*>(mean + (size_t)seg * C + c);
  const float2 rs = *reinterpret_cast<const float2*>(rstd + (size_t)seg * C + c);
  const float* g = gb + (size_t)seg * ldgb;
  const float ga0 = 1.f + g[c], ga1 = 1.f + g[c + 1], be0 = g[C + c], be1 = g[C + c + 1];
  const float a0 = alpha[c], a1 = alpha[c + 1];
  const float ia0 = __fdiv_rn(1.f, a0), ia1 = __fdiv_rn(1.f, a1);
  const size_t base = (size_t)tile * 128;
#pragma unroll 2
  for (int lr = rg; lr < nvalid; lr += 4) {
    const size_t r = base + lr;
    const float2 v = *reinterpret_cast<const float2*>(x + r * ldx + c);
    const float n0 = ga0 * ((v.x - mu.x) * rs.x) + be0;
    const float n1 = ga1 * ((v.y - mu.y) * rs.y) + be1;
    const float s0 = sinf(a0 * n0), s1 = sinf(a1 * n1);
    const float o0 = n0 + ia0 * (s0 * s0), o1 = n1 + ia1 * (s1 * s1);
    if (y != nullptr) *reinterpret_cast<float2*>(y + r * ldy + c) = make_float2(o0, o1);
    if (yh != nullptr) *reinterpret_cast<twin2_t*>(yh + r * ldyh + c) = to_twin2(o0, o1);
  }
}

cudaError_t adain_snake(const float* x, int ldx, int C, const RowMap& map, const float* mean, const float* rstd,
                        const float* gb, int ldgb, const float* alpha, float* y, int ldy, twin_t* yh, int ldyh,
                        cudaStream_t st) {
  dim3 grid(map.n_tiles, (C + 127) / 128);
  launch_k(adain_snake_kernel, grid, 256, 0, st, x, ldx, C, map, mean, rstd, gb, ldgb, alpha, y, ldy, yh, ldyh);
  return cudaGetLastError();
}

// layout + workspace of the block for (n_seg, C, lengths); re-planned (cudaMalloc, synchronising) only when they change
int plan_snake(Context& c, VariantWork& w, int C, int J, int n_seg, const int* len) {
  std::vector<int> sig = {n_seg, C, J};
  sig.insert(sig.end(), len, len + n_seg);
  if (sig == w.sig && w.arena != nullptr) return 0;
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "snake block: sync before re-planning");
  tc_map_cache_clear();
  build_layout(w.L, n_seg, len, kSnakeGuard);
  Layout& L = w.L;
  size_t used = 0;
  auto take = [&](size_t bytes) { const size_t o = used; used += (bytes + 255) / 256 * 256; return o; };
  const size_t act_f = (size_t)L.R * C * 4, act_h = (size_t)L.R * C * 2;
  const size_t o_x0 = take(act_f), o_x1 = take(act_f), o_act = take(act_f), o_acth = take(act_h), o_y1 = take(act_f);
  const size_t o_stats = take((size_t)L.n_tiles * C * 2 * 8);
  const size_t o_mean = take((size_t)n_seg * C * 4), o_rstd = take((size_t)n_seg * C * 4), o_gb = take((size_t)n_seg * J * 4);
  const size_t o_ts = take((size_t)L.n_tiles * 4), o_tn = take((size_t)L.n_tiles * 4), o_ss = take((size_t)n_seg * 4);
  const size_t o_sl = take((size_t)n_seg * 4), o_off = take((size_t)n_seg * 8);
  if (used > w.arena_bytes) {
    if (w.arena) cudaFree(w.arena);
    w.arena = nullptr; w.arena_bytes = 0; w.sig.clear();
    e = cudaMalloc(&w.arena, used);
    if (e != cudaSuccess) return cuda_fail(c, e, "snake block: workspace cudaMalloc");
    w.arena_bytes = used;
  }
  e = cudaMemset(w.arena, 0, used);          // guard rows and channel padding stay zero from here on
  if (e != cudaSuccess) return cuda_fail(c, e, "snake block: workspace memset");
  auto act = [&](size_t of, size_t oh) {
    Act a;
    a.f = reinterpret_cast<float*>(w.arena + of);
    a.h = oh != (size_t)-1 ? reinterpret_cast<twin_t*>(w.arena + oh) : nullptr;
    a.ld = C; a.rows = L.R;
    return a;
  };
  w.X[0] = act(o_x0, (size_t)-1); w.X[1] = act(o_x1, (size_t)-1); w.ACT = act(o_act, o_acth); w.Y1 = act(o_y1, (size_t)-1);
  w.stats = reinterpret_cast<double*>(w.arena + o_stats);
  w.mean = reinterpret_cast<float*>(w.arena + o_mean);
  w.rstd = reinterpret_cast<float*>(w.arena + o_rstd);
  w.gb = reinterpret_cast<float*>(w.arena + o_gb);
  L.d_tile_seg = reinterpret_cast<int*>(w.arena + o_ts);
  L.d_tile_nvalid = reinterpret_cast<int*>(w.arena + o_tn);
  L.d_seg_start = reinterpret_cast<int*>(w.arena + o_ss);
  L.d_seg_len = reinterpret_cast<int*>(w.arena + o_sl);
  w.d_cs = L.d_seg_len;                                      // channel stride of the dense [C, L_b] tensors = L_b
  w.d_off = reinterpret_cast<long long*>(w.arena + o_off);
  std::vector<long long> off(n_seg);
  for (int b = 0; b < n_seg; ++b) off[b] = (long long)C * L.cu[b];
  cudaMemcpy(L.d_tile_seg, L.tile_seg.data(), (size_t)L.n_tiles * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_tile_nvalid, L.tile_nvalid.data(), (size_t)L.n_tiles * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_seg_start, L.seg_start.data(), (size_t)n_seg * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_seg_len, L.seg_len.data(), (size_t)n_seg * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(w.d_off, off.data(), (size_t)n_seg * 8, cudaMemcpyHostToDevice);
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "snake block: layout upload");
  w.sig = sig;
  return 0;
}

}  // namespace

// AdaptiveGeneratorBlock.forward (ada_norm.py:109-120) on a ragged batch: d_x / d_out are the utterances' dense
// [C, L_b] tensors back to back, d_style [n_seg, 64].  GEMM precision follows stts_set_precision (fp32 CUDA cores,
// one-pass fp16 operands with fp32 accumulation, or three-pass tf32).
int run_snake_block(Context& c, const std::string& module, int channels, int k, const int* dil, int n_seg,
                    const int* h_len, const float* d_x, const float* d_style, float* d_out, cudaStream_t st) {
  if (n_seg <= 0 || n_seg > 1024) return fail(c, -1, "snake block: n_seg must be in [1, 1024]");
  for (int b = 0; b < n_seg; ++b)
    if (h_len[b] <= 0) return fail(c, -2, "snake block: utterance length must be > 0");
  auto it = c.snake.find(module);
  if (it == c.snake.end() || it->second.C != channels || it->second.k != k || it->second.dil[0] != dil[0] ||
      it->second.dil[1] != dil[1] || it->second.dil[2] != dil[2]) {
    STTS_RET(pack_snake_block(c, module, channels, k, dil));
    it = c.snake.find(module);
  }
  const SnakeCfg& cfg = it->second;
  VariantWork& w = work(c);
  STTS_RET(plan_snake(c, w, channels, cfg.J, n_seg, h_len));
  const RowMap map = w.L.map();
  const int C = channels;
  const bool twins = c.precision == 1;
  STTS_CK(style_fc(c.vec[module + ".fc_W"], c.vec[module + ".fc_b"], d_style, cfg.J, 64, n_seg, w.gb, st));
  STTS_CK(transpose_in(d_x, w.d_off, w.d_cs, C, map, w.X[0].f, C, 0, nullptr, 0, 0, st));
  int cur = 0;
  for (int i = 0; i < 3; ++i) {
    const std::string si = std::to_string(i);
    const Act& x = w.X[cur];
    Act& xn = w.X[cur ^ 1];
    STTS_CK(colstats(x.f, C, C, map, 0, 1e-5f, w.stats, w.mean, w.rstd, st));
    c.launches++;
    STTS_CK(adain_snake(x.f, C, C, map, w.mean, w.rstd, w.gb + c.fc_off[module + ".n1." + si], cfg.J,
                        c.vec[module + ".a1." + si], twins ? nullptr : w.ACT.f, C, twins ? w.ACT.h : nullptr, C, st));
    {
      Epi e;
      std::memset(&e, 0, sizeof(e));
      e.kind = EPI_BIAS; e.out = w.Y1.f; e.ldo = C; e.scale = 1.f;
      STTS_RET(run_gemm(c, c.gemm[module + ".c1." + si], GC_BODY, &w.ACT, 0, nullptr, 0, e, map, st));
    }
    STTS_CK(colstats(w.Y1.f, C, C, map, 0, 1e-5f, w.stats, w.mean, w.rstd, st));
    c.launches++;
    STTS_CK(adain_snake(w.Y1.f, C, C, map, w.mean, w.rstd, w.gb + c.fc_off[module + ".n2." + si], cfg.J,
                        c.vec[module + ".a2." + si], twins ? nullptr : w.ACT.f, C, twins ? w.ACT.h : nullptr, C, st));
    {
      Epi e;
      std::memset(&e, 0, sizeof(e));
      e.kind = EPI_ADD; e.out = xn.f; e.ldo = C; e.scale = 1.f;
      e.res = x.f; e.ldr = C; e.rcol0 = 0;                       // x = xt + x, ada_norm.py:119
      STTS_RET(run_gemm(c, c.gemm[module + ".c2." + si], GC_BODY, &w.ACT, 0, nullptr, 0, e, map, st));
    }
    cur ^= 1;
  }
  STTS_CK(transpose_out(w.X[cur].f, C, 0, C, map, d_out, w.d_off, w.d_cs, 0, st));
  return 0;
}

// =====================================================================================================================
// SineGenerator (cfm_mel_decoder.py:16-101): per-sample harmonic sine source with the reference's wrap-corrected
// double cumulative sum.
//
//   rad[i, h]  = remainder(f0[i] * (h + 1) / sr, 1)  (+ rand_ini[h] at i = 0)                            :64-70
//   c1         = cumsum(rad)                          torch CPU: fp64 accumulator, every output rounded to fp32
//   shift[i]   = -1 if remainder(c1[i], 1) < remainder(c1[i - 1], 1) else 0                              :73-76
//   c2         = cumsum(rad + shift)                  same accumulation contract
//   sine       = sin(c2 * 2 * pi) * amp ;  out = tanh(merge . (sine * uv + noise_amp * noise))           :78, :93-101
//
// The two cumulative sums are chunked scans in fp64: (A) per-chunk totals, (B) exclusive scan of the totals of an
// utterance, (C) in-chunk scan with the chunk's prefix.  A chunk is 2048 samples (256 threads x 8 consecutive samples,
// all harmonics).  fp64 sums of fp32 addends are exact here (addends >= 2^-34 in their last bit, partial sums < 2^21),
// so the grouping of the parallel scan gives the same fp64 values as the reference's sequential loop, and c1 / c2 --
// hence every wrap decision -- are bit-identical to the reference (tests/test_variants_gpu.py checks the optional `phase`
// output = c2 against the golden vectors, bit for bit).  What remains floating-point-implementation-defined is sin (correctly rounded here via fp64, a
// 1-ulp SLEEF sinf in the reference), the 9-term merge dot product and tanh: <= 1e-6 on the output.
namespace {

constexpr int kSgThreads = 256, kSgPer = 8, kSgChunk = kSgThreads * kSgPer, kSgMaxDim = 16;

struct SgParams {
  int dim;
  float sr, sine_amp, noise_voiced, noise_unvoiced, thr;
  float w[kSgMaxDim];
};

__device__ __forceinline__ float rem1(float x) {                 // torch.remainder(x, 1) for fp32
  float m = fmodf(x, 1.f);
  if (m != 0.f && m < 0.f) m += 1.f;
  return m;
}

// exclusive scan of one double per thread over the 256-thread block; `total` = block sum.  ws: 8 doubles.
__device__ __forceinline__ double block_excl_scan(double v, double* ws, double& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();                                              // ws free again
  if (lane == 31) ws[warp] = inc;
  __syncthreads();
  double before = 0.0, tot = 0.0;
#pragma unroll
  for (int k = 0; k < kSgThreads / 32; ++k) { const double x = ws[k]; if (k < warp) before += x; tot += x; }
  total = tot;
  return before + inc - v;
}

// PASS 0: T1[chunk, h] = sum rad.  PASS 1: T2[chunk, h] = sum (rad + shift), needs P1.  PASS 2: outputs, needs P1, P2.
template <int PASS>
__global__ void __launch_bounds__(kSgThreads) sinegen_pass_kernel(
    const float* __restrict__ f0, const long long* __restrict__ cu, const int* __restrict__ chunk_seg,
    const int* __restrict__ chunk_start, const float* __restrict__ rand_ini, const double* __restrict__ P1,
    const double* __restrict__ P2, double* __restrict__ T, const SgParams p, const float* __restrict__ noise,
    float* __restrict__ out, float* __restrict__ phase) {
  pdl_enter();
  __shared__ double ws[kSgThreads / 32];
  const int chunk = blockIdx.x;
  const int seg = chunk_seg[chunk];
  const long long base = cu[seg];
  const int L = (int)(cu[seg + 1] - base);
  const int i0 = chunk_start[chunk] + threadIdx.x * kSgPer;      // first sample of this thread inside the utterance
  float f[kSgPer];
#pragma unroll
  for (int i = 0; i < kSgPer; ++i) f[i] = (i0 + i < L) ? f0[base + i0 + i] : 0.f;
  float merged[kSgPer];
#pragma unroll
  for (int i = 0; i < kSgPer; ++i) merged[i] = 0.f;
  for (int h = 0; h < p.dim; ++h) {
    const float mult = (float)(h + 1);
    const float ini = (h > 0) ? rand_ini[seg * p.dim + h] : 0.f;
    float rad[kSgPer];
    double s1 = 0.0;
#pragma unroll
    for (int i = 0; i < kSgPer; ++i) {
      float r = 0.f;
      if (i0 + i < L) {
        const float fh = (h == 0) ? f[i] : __fmul_rn(f[i], mult);                    // f0_buf[:, :, 0] * (idx + 2)
        r = rem1(__fdiv_rn(fh, p.sr));
        if (i0 + i == 0) r = __fadd_rn(r, ini);
      }
      rad[i] = r;
      s1 += (double)r;
    }
    double tot;
    const double e1 = block_excl_scan(s1, ws, tot);
    if (PASS == 0) {
      if (threadIdx.x == 0) T[(size_t)chunk * p.dim + h] = tot;
      continue;
    }
    double run = P1[(size_t)chunk * p.dim + h] + e1;
    float over_prev = rem1((float)run);                           // remainder(c1[i0 - 1], 1)
    float v[kSgPer];
    double s2 = 0.0;
#pragma unroll
    for (int i = 0; i < kSgPer; ++i) {
      run += (double)rad[i];
      const float over = rem1((float)run);
      const float shift = (i0 + i > 0 && __fsub_rn(over, over_prev) < 0.f) ? -1.f : 0.f;
      over_prev = over;
      v[i] = (i0 + i < L) ? __fadd_rn(rad[i], shift) : 0.f;
      s2 += (double)v[i];
    }
    const double e2 = block_excl_scan(s2, ws, tot);
    if (PASS == 1) {
      if (threadIdx.x == 0) T[(size_t)chunk * p.dim + h] = tot;
      continue;
    }
    double run2 = P2[(size_t)chunk * p.dim + h] + e2;
#pragma unroll
    for (int i = 0; i < kSgPer; ++i) {
      run2 += (double)v[i];
      if (i0 + i < L) {
        const float c2 = (float)run2;
        const float arg = __fmul_rn(__fmul_rn(c2, 2.f), 3.14159265358979323846f);
        const float sn = (float)sin((double)arg);
        const size_t gi = (size_t)(base + i0 + i);
        if (phase != nullptr) phase[gi * p.dim + h] = c2;
        const bool voiced = f[i] > p.thr;
        const float sw = __fmul_rn(sn, p.sine_amp);
        const float nz = __fmul_rn(voiced ? p.noise_voiced : p.noise_unvoiced, noise[gi * p.dim + h]);
        const float val = __fadd_rn(voiced ? sw : __fmul_rn(sw, 0.f), nz);
        merged[i] = fmaf(val, p.w[h], merged[i]);
      }
    }
  }
  if (PASS == 2) {
#pragma unroll
    for (int i = 0; i < kSgPer; ++i)
      if (i0 + i < L) out[base + i0 + i] = tanhf(merged[i]);
  }
}

// P[chunk, h] = sum of T over the earlier chunks of the same utterance
__global__ void sinegen_prefix_kernel(const double* __restrict__ T, double* __restrict__ P, const int* __restrict__ seg_chunk0,
                                      int dim) {
  pdl_enter();
  const int seg = blockIdx.x, h = threadIdx.x;
  if (h >= dim) return;
  double run = 0.0;
  for (int ch = seg_chunk0[seg]; ch < seg_chunk0[seg + 1]; ++ch) {
    P[(size_t)ch * dim + h] = run;
    run += T[(size_t)ch * dim + h];
  }
}

}  // namespace

int run_sinegen(Context& c, const float* d_f0, const int* h_len, int n_seg, const float* d_rand_ini, const float* d_noise,
                const float* h_merge_w, int harmonic_num, float samp_rate, float sine_amp, float noise_std,
                float voiced_threshold, float* d_out, float* d_phase, cudaStream_t st) {
  const int dim = harmonic_num + 1;
  if (n_seg <= 0 || n_seg > 1024) return fail(c, -1, "sinegen: n_seg must be in [1, 1024]");
  if (dim < 1 || dim > kSgMaxDim) return fail(c, -2, "sinegen: harmonic_num must be in [0, 15]");
  if (!(samp_rate > 0.f)) return fail(c, -2, "sinegen: samp_rate must be positive");
  std::vector<long long> cu(n_seg + 1, 0);
  std::vector<int> seg_chunk0(n_seg + 1, 0), chunk_seg, chunk_start;
  for (int b = 0; b < n_seg; ++b) {
    if (h_len[b] <= 0) return fail(c, -2, "sinegen: utterance length must be > 0");
    cu[b + 1] = cu[b] + h_len[b];
    for (int s0 = 0; s0 < h_len[b]; s0 += kSgChunk) { chunk_seg.push_back(b); chunk_start.push_back(s0); }
    seg_chunk0[b + 1] = (int)chunk_seg.size();
  }
  const int nch = (int)chunk_seg.size();
  VariantWork& w = work(c);
  size_t used = 0;
  auto take = [&](size_t bytes) { const size_t o = used; used += (bytes + 255) / 256 * 256; return o; };
  const size_t o_cu = take((size_t)(n_seg + 1) * 8), o_sc = take((size_t)(n_seg + 1) * 4), o_cs = take((size_t)nch * 4);
  const size_t o_ct = take((size_t)nch * 4), o_t = take((size_t)nch * dim * 8), o_p1 = take((size_t)nch * dim * 8);
  const size_t o_p2 = take((size_t)nch * dim * 8);
  if (used > w.sg_bytes) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return cuda_fail(c, e, "sinegen: sync before growing the scratch");
    if (w.sg) cudaFree(w.sg);
    w.sg = nullptr; w.sg_bytes = 0;
    e = cudaMalloc(&w.sg, used);
    if (e != cudaSuccess) return cuda_fail(c, e, "sinegen: scratch cudaMalloc");
    w.sg_bytes = used;
  }
  // pageable sources: cudaMemcpyAsync stages them before it returns
  STTS_CK(cudaMemcpyAsync(w.sg + o_cu, cu.data(), (size_t)(n_seg + 1) * 8, cudaMemcpyHostToDevice, st));
  STTS_CK(cudaMemcpyAsync(w.sg + o_sc, seg_chunk0.data(), (size_t)(n_seg + 1) * 4, cudaMemcpyHostToDevice, st));
  STTS_CK(cudaMemcpyAsync(w.sg + o_cs, chunk_seg.data(), (size_t)nch * 4, cudaMemcpyHostToDevice, st));
  STTS_CK(cudaMemcpyAsync(w.sg + o_ct, chunk_start.data(), (size_t)nch * 4, cudaMemcpyHostToDevice, st));
  c.launches -= 4;                                               // copies, not kernels
  const long long* d_cu = reinterpret_cast<const long long*>(w.sg + o_cu);
  const int* d_sc = reinterpret_cast<const int*>(w.sg + o_sc);
  const int* d_cs = reinterpret_cast<const int*>(w.sg + o_cs);
  const int* d_ct = reinterpret_cast<const int*>(w.sg + o_ct);
  double* T = reinterpret_cast<double*>(w.sg + o_t);
  double* P1 = reinterpret_cast<double*>(w.sg + o_p1);
  double* P2 = reinterpret_cast<double*>(w.sg + o_p2);
  SgParams p;
  std::memset(&p, 0, sizeof(p));
  p.dim = dim; p.sr = samp_rate; p.sine_amp = sine_amp; p.thr = voiced_threshold;
  p.noise_voiced = noise_std;                                    // uv * noise_std + (1 - uv) * sine_amp / 3, uv = 1
  p.noise_unvoiced = sine_amp / 3.f;                             // ... uv = 0 (fp32 division, like the tensor op)
  for (int h = 0; h < dim; ++h) p.w[h] = h_merge_w[h];
  launch_k(sinegen_pass_kernel<0>, nch, kSgThreads, 0, st, d_f0, d_cu, d_cs, d_ct, d_rand_ini, nullptr, nullptr, T, p, d_noise,
                                                      nullptr, nullptr);
  STTS_CK(cudaGetLastError());
  launch_k(sinegen_prefix_kernel, n_seg, 32, 0, st, T, P1, d_sc, dim);
  STTS_CK(cudaGetLastError());
  launch_k(sinegen_pass_kernel<1>, nch, kSgThreads, 0, st, d_f0, d_cu, d_cs, d_ct, d_rand_ini, P1, nullptr, T, p, d_noise,
                                                      nullptr, nullptr);
  STTS_CK(cudaGetLastError());
  launch_k(sinegen_prefix_kernel, n_seg, 32, 0, st, T, P2, d_sc, dim);
  STTS_CK(cudaGetLastError());
  launch_k(sinegen_pass_kernel<2>, nch, kSgThreads, 0, st, d_f0, d_cu, d_cs, d_ct, d_rand_ini, P1, P2, T, p, d_noise, d_out,
                                                      d_phase);
  STTS_CK(cudaGetLastError());
  return 0;
}

// =====================================================================================================================
// STFT / iSTFT as convolutions with a windowed DFT basis (train/models/stft.py:6-195, the module swapped in for the
// ONNX export, convert_to_onnx.py:31-36): SURVEY.md section 8 (f) rank 4.
//
//   transform: center padding n_fft / 2 (replicate | constant), frames = conv1d(stride hop) with the basis
//              window[n] * (cos, -sin)(2 pi k n / n_fft); returns |X| = sqrt(re^2 + im^2 + 1e-14), re / |X|, im / |X|
//   inverse:   conv_transpose1d(stride hop) with window[n] / n_fft * (cos, sin), real minus imaginary part, n_fft / 2
//              samples trimmed at both ends (the reference's "approximate" inverse: no envelope normalisation)
//
// The window is zero beyond win_length, so a frame is taps = ceil(win / hop) consecutive rows of `hop` samples: both
// directions are conv-as-GEMMs of the live path on a private packed layout (forward: K = taps x hop, N = 2 bins with
// (re, im) interleaved; inverse: K = 2 bins, N = taps x hop followed by an overlap-add gather).  Basis matrices are
// the module's own registered buffers (weight_forward_real / _imag, weight_backward_real / _imag), loaded with
// stts_load_tensor like any other state_dict entry.  fp32-grade GEMMs only (CUDA cores in fp32 mode, tf32x3 otherwise).
struct StftConvPlan {
  int n_fft = 0, hop = 0, win = 0, bins = 0, taps = 0, hop_pad = 0, ld_spec = 0, ld_y = 0;
  PackedGemm fwd, bwd;
  bool packed = false;
  // layout + workspace for (lengths)
  Layout L;
  std::vector<int> sig;
  char* arena = nullptr;
  size_t arena_bytes = 0;
  Act XR, SPEC, YF;
  long long* d_off_w = nullptr;      // dense sample offset of every utterance (waveform side)
  long long* d_off_s = nullptr;      // dense element offset of every utterance's [bins, frames] block
  int* d_len_w = nullptr;            // samples per utterance (waveform side)
};

namespace {

// rows of `hop` samples of the center-padded waveform: XR[row(seg, r), c] = padded[r * hop + c]
__global__ void __launch_bounds__(256) stftc_rowify_kernel(const float* __restrict__ wave, const long long* __restrict__ off,
                                                           const int* __restrict__ len, RowMap map, int hop, int pad_len,
                                                           int extra_rows, int pad_mode, float* __restrict__ xr, int ld) {
  pdl_enter();
  const int seg = blockIdx.y;
  const int T = len[seg];
  const int rows = map.seg_len[seg] + extra_rows;
  const float* w = wave + off[seg];
  const size_t r0 = map.seg_start[seg];
  const int total = rows * hop;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / hop, c = i - r * hop;
    int s = i - pad_len;                       // sample index in the unpadded waveform
    float v = 0.f;
    if (s >= -pad_len && s < T + pad_len) {
      if (s < 0) v = pad_mode == 0 ? w[0] : 0.f;
      else if (s >= T) v = pad_mode == 0 ? w[T - 1] : 0.f;
      else v = w[s];
    }
    xr[(r0 + r) * ld + c] = v;
  }
}

// (re, im) interleaved rows -> dense [bins, frames] magnitude, re / mag, im / mag      stft.py:127-136
__global__ void __launch_bounds__(256) stftc_finish_kernel(const float* __restrict__ spec, int ld, RowMap map, int bins,
                                                           const long long* __restrict__ off, float* __restrict__ mag,
                                                           float* __restrict__ xo, float* __restrict__ yo) {
  pdl_enter();
  __shared__ float2 t[32][33];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int f_base = tile * 128 - map.seg_start[seg];
  const int frames = map.seg_len[seg];
  const long long o = off[seg];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k0 = blockIdx.y * 32; k0 < bins; k0 += gridDim.y * 32) {
    for (int r0 = 0; r0 < nvalid; r0 += 32) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int lr = r0 + ty + 8 * q, k = k0 + tx;
        float2 v = make_float2(0.f, 0.f);
        if (lr < nvalid && k < bins) v = *reinterpret_cast<const float2*>(spec + (size_t)(tile * 128 + lr) * ld + 2 * k);
        t[ty + 8 * q][tx] = v;
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + ty + 8 * q, lr = r0 + tx;
        if (k < bins && lr < nvalid) {
          const float2 v = t[tx][ty + 8 * q];
          const float m = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)), 1e-14f));
          const size_t d = (size_t)o + (size_t)k * frames + f_base + lr;
          mag[d] = m;
          xo[d] = __fdiv_rn(v.x, m);
          yo[d] = __fdiv_rn(v.y, m);
        }
      }
      __syncthreads();
    }
  }
}

// dense [bins, frames] (mag, x, y) -> rows [frames, 2 bins]: (mag x, mag y) interleaved        stft.py:147-148
__global__ void __launch_bounds__(256) stftc_prep_kernel(const float* __restrict__ mag, const float* __restrict__ xi,
                                                         const float* __restrict__ yi, const long long* __restrict__ off,
                                                         RowMap map, int bins, float* __restrict__ a, int ld) {
  pdl_enter();
  __shared__ float2 t[32][33];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int f_base = tile * 128 - map.seg_start[seg];
  const int frames = map.seg_len[seg];
  const long long o = off[seg];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int k0 = blockIdx.y * 32; k0 < bins; k0 += gridDim.y * 32) {
    for (int r0 = 0; r0 < nvalid; r0 += 32) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + ty + 8 * q, lr = r0 + tx;
        float2 v = make_float2(0.f, 0.f);
        if (k < bins && lr < nvalid) {
          const size_t d = (size_t)o + (size_t)k * frames + f_base + lr;
          const float m = mag[d];
          v = make_float2(__fmul_rn(m, xi[d]), __fmul_rn(m, yi[d]));
        }
        t[ty + 8 * q][tx] = v;
      }
      __syncthreads();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int lr = r0 + ty + 8 * q, k = k0 + tx;
        if (lr < nvalid && k < bins)
          *reinterpret_cast<float2*>(a + (size_t)(tile * 128 + lr) * ld + 2 * k) = t[tx][ty + 8 * q];
      }
      __syncthreads();
    }
  }
}

// overlap-add of the per-frame segments (conv_transpose1d, stride hop) and the trim of pad_len samples    stft.py:155-181
__global__ void __launch_bounds__(256) stftc_ola_kernel(const float* __restrict__ yf, int ld, RowMap map, int hop, int taps,
                                                        int pad_len, const long long* __restrict__ off,
                                                        const int* __restrict__ len, float* __restrict__ wave) {
  pdl_enter();
  const int seg = blockIdx.y;
  const int T = len[seg];
  const int frames = map.seg_len[seg];
  const size_t r0 = map.seg_start[seg];
  float* w = wave + off[seg];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T; i += gridDim.x * blockDim.x) {
    const int s = i + pad_len;
    const int tq = s / hop, c = s - tq * hop;
    float acc = 0.f;
    for (int j = taps - 1; j >= 0; --j) {       // ascending frame index, like the scatter order of conv_transpose1d
      const int t = tq - j;
      if (t >= 0 && t < frames) acc += yf[(r0 + t) * ld + j * hop + c];
    }
    w[i] = acc;
  }
}

StftConvPlan& stft_plan(Context& c, const std::string& module) {
  VariantWork& w = work(c);
  auto it = w.stft.find(module);
  if (it == w.stft.end()) it = w.stft.emplace(module, new StftConvPlan()).first;
  return *it->second;
}

int pack_stft_conv(Context& c, StftConvPlan& pl, const std::string& module, int n_fft, int hop, int win) {
  if (pl.packed && pl.n_fft == n_fft && pl.hop == hop && pl.win == win) return 0;
  if (n_fft < 2 || (n_fft & 1) || hop < 1 || win < 1) return fail(c, -2, "stft conv: filter_length must be even, hop / win positive");
  const int weff = std::min(win, n_fft), taps = (weff + hop - 1) / hop, bins = n_fft / 2 + 1;
  if (taps - 1 > kSnakeGuard) return fail(c, -2, "stft conv: ceil(win_length / hop_length) must be <= 33");
  const char* keys[4] = {"weight_forward_real", "weight_forward_imag", "weight_backward_real", "weight_backward_imag"};
  const HostTensor* t[4];
  for (int i = 0; i < 4; ++i) {
    auto it = c.tensors.find(module + "/" + keys[i]);
    if (it == c.tensors.end()) return fail(c, -3, "missing weight " + module + "/" + keys[i]);
    t[i] = &it->second;
    if (t[i]->numel() != (int64_t)bins * n_fft) return fail(c, -2, std::string("stft conv: ") + keys[i] + " is not [n_fft / 2 + 1, 1, n_fft]");
  }
  if (pl.packed) { free_pack(pl.fwd); free_pack(pl.bwd); pl.packed = false; }
  const int span = taps * hop;
  {  // forward: conv weight [2 bins, hop, taps], w[2k + part][c][j] = basis_part[k][j hop + c]
    std::vector<float> w((size_t)2 * bins * hop * taps, 0.f);
    for (int k = 0; k < bins; ++k)
      for (int part = 0; part < 2; ++part) {
        const float* b = t[part]->data.data() + (size_t)k * n_fft;
        float* d = w.data() + (size_t)(2 * k + part) * hop * taps;
        for (int n = 0; n < span && n < n_fft; ++n) d[(size_t)(n % hop) * taps + n / hop] = b[n];
      }
    pl.fwd = pack_from_conv_weight(c, w.data(), nullptr, 2 * bins, hop, taps, 0);
    pl.fwd.name = module + ".fwd";
  }
  {  // inverse: 1x1 "conv" [span, 2 bins, 1], w[n][2k] = bwd_real[k][n], w[n][2k + 1] = -bwd_imag[k][n]
    std::vector<float> w((size_t)span * 2 * bins, 0.f);
    for (int n = 0; n < span && n < n_fft; ++n)
      for (int k = 0; k < bins; ++k) {
        w[(size_t)n * 2 * bins + 2 * k] = t[2]->data[(size_t)k * n_fft + n];
        w[(size_t)n * 2 * bins + 2 * k + 1] = -t[3]->data[(size_t)k * n_fft + n];
      }
    pl.bwd = pack_from_conv_weight(c, w.data(), nullptr, span, 2 * bins, 1, 0);
    pl.bwd.name = module + ".bwd";
  }
  if (cudaDeviceSynchronize() != cudaSuccess) return fail(c, -5, "cuda error while uploading the STFT basis");
  pl.n_fft = n_fft; pl.hop = hop; pl.win = win; pl.bins = bins; pl.taps = taps;
  pl.hop_pad = (hop + 63) / 64 * 64;
  pl.ld_spec = (2 * bins + 63) / 64 * 64;
  pl.ld_y = (span + 63) / 64 * 64;
  pl.packed = true;
  pl.sig.clear();
  // the window beyond taps * hop samples must be zero, or the frame would not fit its rows
  for (int part = 0; part < 2; ++part)
    for (int k = 0; k < bins; k += bins - 1)
      for (int n = span; n < n_fft; ++n)
        if (t[part]->data[(size_t)k * n_fft + n] != 0.f) return fail(c, -2, "stft conv: basis is not zero beyond win_length");
  return 0;
}

// layout (rows = frames per utterance) + workspace; len_w = samples per utterance on the waveform side
int plan_stft_conv(Context& c, StftConvPlan& pl, int n_seg, const int* frames, const int* len_w) {
  std::vector<int> sig = {n_seg};
  sig.insert(sig.end(), frames, frames + n_seg);
  sig.insert(sig.end(), len_w, len_w + n_seg);
  if (sig == pl.sig && pl.arena != nullptr) return 0;
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "stft conv: sync before re-planning");
  tc_map_cache_clear();
  build_layout(pl.L, n_seg, frames, kSnakeGuard);
  Layout& L = pl.L;
  size_t used = 0;
  auto take = [&](size_t bytes) { const size_t o = used; used += (bytes + 255) / 256 * 256; return o; };
  const size_t o_xr = take((size_t)L.R * pl.hop_pad * 4), o_sp = take((size_t)L.R * pl.ld_spec * 4);
  const size_t o_yf = take((size_t)L.R * pl.ld_y * 4);
  const size_t o_ts = take((size_t)L.n_tiles * 4), o_tn = take((size_t)L.n_tiles * 4), o_ss = take((size_t)n_seg * 4);
  const size_t o_sl = take((size_t)n_seg * 4), o_ow = take((size_t)n_seg * 8), o_os = take((size_t)n_seg * 8);
  const size_t o_lw = take((size_t)n_seg * 4);
  if (used > pl.arena_bytes) {
    if (pl.arena) cudaFree(pl.arena);
    pl.arena = nullptr; pl.arena_bytes = 0; pl.sig.clear();
    e = cudaMalloc(&pl.arena, used);
    if (e != cudaSuccess) return cuda_fail(c, e, "stft conv: workspace cudaMalloc");
    pl.arena_bytes = used;
  }
  e = cudaMemset(pl.arena, 0, used);
  if (e != cudaSuccess) return cuda_fail(c, e, "stft conv: workspace memset");
  auto act = [&](size_t of, int ld) { Act a; a.f = reinterpret_cast<float*>(pl.arena + of); a.ld = ld; a.rows = L.R; return a; };
  pl.XR = act(o_xr, pl.hop_pad); pl.SPEC = act(o_sp, pl.ld_spec); pl.YF = act(o_yf, pl.ld_y);
  L.d_tile_seg = reinterpret_cast<int*>(pl.arena + o_ts);
  L.d_tile_nvalid = reinterpret_cast<int*>(pl.arena + o_tn);
  L.d_seg_start = reinterpret_cast<int*>(pl.arena + o_ss);
  L.d_seg_len = reinterpret_cast<int*>(pl.arena + o_sl);
  pl.d_off_w = reinterpret_cast<long long*>(pl.arena + o_ow);
  pl.d_off_s = reinterpret_cast<long long*>(pl.arena + o_os);
  pl.d_len_w = reinterpret_cast<int*>(pl.arena + o_lw);
  std::vector<long long> ow(n_seg), os(n_seg);
  long long aw = 0, as = 0;
  for (int b = 0; b < n_seg; ++b) { ow[b] = aw; os[b] = as; aw += len_w[b]; as += (long long)pl.bins * frames[b]; }
  cudaMemcpy(L.d_tile_seg, L.tile_seg.data(), (size_t)L.n_tiles * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_tile_nvalid, L.tile_nvalid.data(), (size_t)L.n_tiles * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_seg_start, L.seg_start.data(), (size_t)n_seg * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(L.d_seg_len, L.seg_len.data(), (size_t)n_seg * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(pl.d_off_w, ow.data(), (size_t)n_seg * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(pl.d_off_s, os.data(), (size_t)n_seg * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(pl.d_len_w, len_w, (size_t)n_seg * 4, cudaMemcpyHostToDevice);
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(c, e, "stft conv: layout upload");
  pl.sig = sig;
  return 0;
}

Epi plain_epi(const Act& dst) {
  Epi e;
  std::memset(&e, 0, sizeof(e));
  e.kind = EPI_BIAS; e.out = dst.f; e.ldo = dst.ld; e.scale = 1.f;
  return e;
}

}  // namespace

void free_stft_plans(VariantWork& w) {
  for (auto& kv : w.stft) {
    StftConvPlan* pl = kv.second;
    if (pl->packed) { free_pack(pl->fwd); free_pack(pl->bwd); }
    if (pl->arena) cudaFree(pl->arena);
    delete pl;
  }
  w.stft.clear();
}

// STFT.transform (stft.py:98-136).  d_wave: utterances back to back, len [n_seg] samples (host).  Outputs: per
// utterance a dense [bins, frames_b] block, frames_b = len_b / hop + 1, blocks back to back.
int run_stft_conv_transform(Context& c, const std::string& module, int n_fft, int hop, int win, int pad_mode, int n_seg,
                            const int* h_len, const float* d_wave, float* d_mag, float* d_x, float* d_y, cudaStream_t st) {
  if (n_seg <= 0 || n_seg > 1024) return fail(c, -1, "stft conv: n_seg must be in [1, 1024]");
  if (pad_mode != 0 && pad_mode != 1) return fail(c, -2, "stft conv: pad_mode is 0 (replicate) or 1 (constant)");
  StftConvPlan& pl = stft_plan(c, module);
  STTS_RET(pack_stft_conv(c, pl, module, n_fft, hop, win));
  std::vector<int> frames(n_seg);
  for (int b = 0; b < n_seg; ++b) {
    if (h_len[b] <= 0) return fail(c, -2, "stft conv: utterance length must be > 0");
    frames[b] = h_len[b] / hop + 1;                       // (T + 2 (n_fft / 2) - n_fft) / hop + 1
  }
  STTS_RET(plan_stft_conv(c, pl, n_seg, frames.data(), h_len));
  const RowMap map = pl.L.map();
  STTS_CK(launch_k(stftc_rowify_kernel, dim3(64, n_seg), 256, 0, st, d_wave, pl.d_off_w, pl.d_len_w, map, hop, n_fft / 2,
                   pl.taps - 1, pad_mode, pl.XR.f, pl.XR.ld));
  STTS_RET(run_gemm(c, pl.fwd, GC_BODY, &pl.XR, 0, nullptr, 0, plain_epi(pl.SPEC), map, st, c.precision == 0 ? 0 : 2));
  STTS_CK(launch_k(stftc_finish_kernel, dim3(map.n_tiles, 8), 256, 0, st, pl.SPEC.f, pl.SPEC.ld, map, pl.bins, pl.d_off_s,
                   d_mag, d_x, d_y));
  return 0;
}

// STFT.inverse (stft.py:138-186).  Inputs: dense [bins, frames_b] blocks; h_out_len[b] <= (frames_b - 1) * hop samples
// are written per utterance (NULL: all of them), utterances back to back.
int run_stft_conv_inverse(Context& c, const std::string& module, int n_fft, int hop, int win, int n_seg,
                          const int* h_frames, const int* h_out_len, const float* d_mag, const float* d_x, const float* d_y,
                          float* d_wave, cudaStream_t st) {
  if (n_seg <= 0 || n_seg > 1024) return fail(c, -1, "stft conv: n_seg must be in [1, 1024]");
  StftConvPlan& pl = stft_plan(c, module);
  STTS_RET(pack_stft_conv(c, pl, module, n_fft, hop, win));
  std::vector<int> len(n_seg);
  for (int b = 0; b < n_seg; ++b) {
    if (h_frames[b] <= 0) return fail(c, -2, "stft conv: frame count must be > 0");
    const int full = (h_frames[b] - 1) * hop;             // (frames - 1) hop + n_fft - 2 (n_fft / 2)
    len[b] = h_out_len != nullptr ? h_out_len[b] : full;
    if (len[b] < 0 || len[b] > full) return fail(c, -2, "stft conv: requested length exceeds (frames - 1) * hop");
  }
  STTS_RET(plan_stft_conv(c, pl, n_seg, h_frames, len.data()));
  const RowMap map = pl.L.map();
  STTS_CK(launch_k(stftc_prep_kernel, dim3(map.n_tiles, 8), 256, 0, st, d_mag, d_x, d_y, pl.d_off_s, map, pl.bins, pl.SPEC.f,
                   pl.SPEC.ld));
  STTS_RET(run_gemm(c, pl.bwd, GC_BODY, &pl.SPEC, 0, nullptr, 0, plain_epi(pl.YF), map, st, c.precision == 0 ? 0 : 2));
  STTS_CK(launch_k(stftc_ola_kernel, dim3(64, n_seg), 256, 0, st, pl.YF.f, pl.YF.ld, map, hop, pl.taps, n_fft / 2, pl.d_off_w,
                   pl.d_len_w, d_wave));
  return 0;
}

}  // namespace stts
