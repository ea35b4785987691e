// F0-driven harmonic-plus-noise source: generate_pcph, train/models/generator.py:247-315
// (pseudo-constant-power harmonics, <= 16 harmonics, float64 running phase, + 0.01 * noise).
//
// The reference repeats f0 x75, takes a float64 cumsum over samples and materialises
// [B, 16, samples] in float64.  Here the phase of sample k of frame i is closed-form:
//     phase = init + sum_{j<i} 75 * f0_j / sr + (k + 1) * f0_i / sr          (float64)
// so one fp64 scan over FRAMES (pcph_prefix) plus an embarrassingly parallel sample kernel
// (pcph_source) replace the sample-rate scan; harmonics come from one fp64 sincospi and the
// angle-addition recurrence.  No host synchronisation: the data-dependent harmonic count and
// the reference's error condition (voiced frames but no f0 > 20 Hz -> torch.min of an empty
// tensor, generator.py:285) are written to a per-utterance SourceMeta record.
//
// Output layout: "hop rows" of 75 samples, [R + 16, ld]; logical hop h of a segment is stored at
// row seg_start + h + 8 and rows h in [-8, 0) and [L, L + 8) hold the reflect padding that
// torch.stft(center=True) would add (only 600 of the 1024 padded samples are touched by the
// 1200-tap window), so the STFT becomes a 16-tap row-shifted GEMM over this buffer.
#include "common.cuh"
#include "kernels.h"

namespace stts {

constexpr int kHop = 75;
constexpr double kSampleRate = 24000.0;

__global__ void __launch_bounds__(256) pcph_prefix_kernel(const float* __restrict__ p4, RowMap map,
                                                          double* __restrict__ prefix, SourceMeta* __restrict__ meta) {
  pdl_enter();
  __shared__ double part[256];
  __shared__ float mins[256];
  __shared__ int anyv[256];
  const int seg = blockIdx.x;
  const int s0 = map.seg_start[seg], len = map.seg_len[seg];
  const int chunk = (len + 255) / 256;
  const int b = threadIdx.x * chunk;
  const int e = (b + chunk < len) ? b + chunk : len;
  double s = 0.0;
  float mn = 3.0e38f;
  int av = 0;
  for (int i = b; i < e; ++i) {
    const float p = p4[s0 + i];
    s += (double)kHop * ((double)p / kSampleRate);
    if (p > 20.f) mn = fminf(mn, p);
    if (p > 10.f) av = 1;
  }
  part[threadIdx.x] = s;
  mins[threadIdx.x] = mn;
  anyv[threadIdx.x] = av;
  __syncthreads();
  if (threadIdx.x == 0) {
    double run = 0.0;
    float m = 3.0e38f;
    int a = 0;
    for (int i = 0; i < 256; ++i) {
      const double v = part[i];
      part[i] = run;
      run += v;
      m = fminf(m, mins[i]);
      a |= anyv[i];
    }
    SourceMeta r;
    r.any_voiced = a;
    r.error = (a && m > 1.0e38f) ? 1 : 0;
    int mh = 16;
    if (m < 1.0e38f) {
      const int q = (int)(12000.0 / (double)m);   // int(max_frequency / min_f0), generator.py:287
      mh = q < 16 ? q : 16;
    }
    r.max_h = mh;
    r.pad = 0;
    meta[seg] = r;
  }
  __syncthreads();
  double run = part[threadIdx.x];
  for (int i = b; i < e; ++i) {
    prefix[s0 + i] = run;
    run += (double)kHop * ((double)p4[s0 + i] / kSampleRate);
  }
}

cudaError_t pcph_prefix(const float* p4, const RowMap& map, double* prefix, SourceMeta* meta, cudaStream_t st) {
  launch_k(pcph_prefix_kernel, map.n_seg, 256, 0, st, p4, map, prefix, meta);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) pcph_source_kernel(const float* __restrict__ p4,
                                                          const double* __restrict__ prefix,
                                                          const SourceMeta* __restrict__ meta,
                                                          const float* __restrict__ noise,
                                                          const float* __restrict__ init_phase,
                                                          const int* __restrict__ cu4, RowMap map,
                                                          float* __restrict__ src, int ld,
                                                          __half* __restrict__ srcg, __half* __restrict__ srcl,
                                                          int ldg, float gscale, float* __restrict__ dense_out) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int s0 = map.seg_start[seg], len = map.seg_len[seg];
  const int h0 = tile * 128 - s0;
  const int lo = (h0 == 0) ? -8 : 0;
  const int hi = nvalid + ((h0 + nvalid == len) ? 8 : 0);
  const int L = len * kHop;
  const SourceMeta mt = meta[seg];
  const double init = (double)init_phase[seg];
  const size_t dense_base = (size_t)cu4[seg] * kHop;
  const int total = (hi - lo) * kHop;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int lrow = lo + i / kHop, c = i % kHop;
    const int h = h0 + lrow;
    int t = h * kHop + c;
    const bool interior = (t >= 0 && t < L);
    if (t < 0) t = -t;                     // reflect (no edge repeat), torch.stft pad_mode="reflect"
    if (t >= L) t = 2 * (L - 1) - t;
    const int fr = t / kHop, k = t - fr * kHop;
    const float p = p4[s0 + fr];
    float val = 0.f;
    if (mt.any_voiced && p > 10.f) {
      const double rad = (double)p / kSampleRate;
      const double phase = init + prefix[s0 + fr] + (double)(k + 1) * rad;
      double s1, c1;
      sincospi(2.0 * phase, &s1, &c1);
      double sh = s1, ch = c1;
      float sum = 0.f;
      for (int hh = 1; hh <= mt.max_h; ++hh) {
        if (__fmul_rn((float)hh, p) <= 12000.f) sum = __fadd_rn(sum, (float)sh);
        const double sn = sh * c1 + ch * s1;
        ch = ch * c1 - sh * s1;
        sh = sn;
      }
      const float nh = __fdiv_rn(12000.f, p);
      const float amp = __fmul_rn(0.1f, __fsqrt_rn(__fdiv_rn(2.f, nh)));
      val = __fmul_rn(amp, sum);
    }
    // harm + 0.01 * noise with both roundings of the reference (an FMA here is a 1-ulp difference of the source, which
    // the STFT phase of near-real bins then amplifies)
    val = __fadd_rn(val, __fmul_rn(0.01f, noise[dense_base + t]));
    src[(size_t)(s0 + h + 8) * ld + c] = val;
    if (srcg != nullptr) {
      // fp16 split twins of the STFT operand, written as explicit frames (im2col): frame m of the
      // utterance holds window tap n = 75 j + c <- hop m + j - 8, so this sample lands in the 16 frames
      // m = h + 8 - j.  K = 1200 (+16 pad) instead of 16 taps x 128 padded columns: 41 % fewer MMAs.
      __half g, l;
      f16_split(val * gscale, &g, &l);                   // pre-scaled: |val| < ~5
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int m = h + 8 - j;
        if (m >= 0 && m < len) {
          const size_t o = (size_t)(s0 + m) * ldg + j * kHop + c;
          srcg[o] = g;
          srcl[o] = l;
        }
      }
    }
    if (dense_out != nullptr && interior) dense_out[dense_base + t] = val;
  }
}

cudaError_t pcph_source(const float* p4, const double* prefix, const SourceMeta* meta, const float* noise,
                        const float* init_phase, const int* cu4, const RowMap& map, float* src, int ld,
                        __half* srcg, __half* srcl, int ldg, float gscale, float* dense_out, cudaStream_t st) {
  launch_k(pcph_source_kernel, map.n_tiles, 256, 0, st, p4, prefix, meta, noise, init_phase, cu4, map, src, ld, srcg, srcl,
                                                  ldg, gscale, dense_out);
  return cudaGetLastError();
}

}  // namespace stts
