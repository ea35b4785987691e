// iSTFT vocoder head, train/models/generator.py:425-433 + TorchSTFT.inverse 46-56:
//   replicate-pad one frame; spec = exp(logamp) * (cos(phase) + i sin(phase));
//   torch.istft(n_fft 2048, hop 75, win 1200 hann, center=True); tanh.
// Split in three:  polar_spec (this file) -> inverse-DFT contraction on the GEMM kernel
// (K = 2 * 1025 interleaved re/im, N = 1200 window taps, basis = irfft * window, prepack.cu)
// -> ola_tanh (this file): 16-way overlap-add, window-envelope divide, trim, tanh.
#include "common.cuh"
#include "kernels.h"

namespace stts {

// spec[r, 2k] = e^la cos(ph), spec[r, 2k+1] = e^la sin(ph); any of the fp32 master `spec`, the fp16
// twin `specg` and its residual `specl` may be null (only what the inverse-DFT GEMM of the current
// precision mode reads is written).  Row seg_end (one past the last
// frame, a guard row of the layout) receives the replicate-padded frame.
// FAST (16-bit mode, only the fp16 twin is written: 2^-12 operand rounding): __expf and MUFU sin / cos after a
// Cody-Waite reduction to [-pi, pi] (absolute error ~5e-7) instead of the libm paths.
template <bool FAST>
__global__ void __launch_bounds__(256) polar_kernel(const float* __restrict__ la, const float* __restrict__ ph, int ld,
                                                    int n_bins, RowMap map, float* __restrict__ spec, int lds,
                                                    __half* __restrict__ specg, int ldsh,
                                                    __half* __restrict__ specl, float gscale) {
  pdl_enter();
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int m0 = tile * 128;
  const bool last = (m0 + nvalid == map.seg_start[seg] + map.seg_len[seg]);
  const int rows = nvalid + (last ? 1 : 0);
  for (int lr = blockIdx.y; lr < rows; lr += gridDim.y) {
    const int srow = m0 + (lr < nvalid ? lr : nvalid - 1);
    const float* lap = la + (size_t)srow * ld;
    const float* php = ph + (size_t)srow * ld;
    if constexpr (FAST) {
      // 16-bit mode: four bins per thread -- two 16-byte loads, one 16-byte store of the (re, im) fp16 pairs; the
      // scalar loop below only takes what is left of the row (1 025 = 256 x 4 + 1).  HBM-bound: 8 B in, 4 B out per bin.
      const int n4 = n_bins & ~3;
      for (int k = threadIdx.x * 4; k < n4; k += blockDim.x * 4) {
        const float4 l4 = *reinterpret_cast<const float4*>(lap + k);
        const float4 p4 = *reinterpret_cast<const float4*>(php + k);
        const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, pv[4] = {p4.x, p4.y, p4.z, p4.w};
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float a = __expf(lv[i]) * gscale;
          const float n = rintf(pv[i] * 0.15915494309189535f);
          float r = fmaf(n, -6.28318548202514648f, pv[i]);
          r = fmaf(n, 1.74845553e-7f, r);
          float sn, cs;
          __sincosf(r, &sn, &cs);
          const twin2_t t = to_twin2(a * cs, a * sn);
          w[i] = *reinterpret_cast<const uint32_t*>(&t);
        }
        *reinterpret_cast<uint4*>(specg + (size_t)(m0 + lr) * ldsh + 2 * k) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
    for (int k = (FAST ? (n_bins & ~3) : 0) + threadIdx.x; k < n_bins; k += blockDim.x) {
      float a, sn, cs;
      const float p = php[k];
      if constexpr (FAST) {
        a = __expf(lap[k]);
        const float n = rintf(p * 0.15915494309189535f);
        float r = fmaf(n, -6.28318548202514648f, p);        // 2 pi = hi + lo (fp32 split)
        r = fmaf(n, 1.74845553e-7f, r);
        __sincosf(r, &sn, &cs);
      } else {
        a = expf(lap[k]);
        sincosf(p, &sn, &cs);
      }
      const float re = a * cs, im = a * sn;
      const size_t o = (size_t)(m0 + lr) * lds + 2 * k;
      if (spec != nullptr) *reinterpret_cast<float2*>(spec + o) = make_float2(re, im);
      if (specg != nullptr) {   // fp16 twins (pre-scaled by gscale: e^logamp may exceed the fp16 range)
        const size_t og = (size_t)(m0 + lr) * ldsh + 2 * k;
        const float sr = fminf(fmaxf(re * gscale, -65504.f), 65504.f), si = fminf(fmaxf(im * gscale, -65504.f), 65504.f);
        const __half2 g = __floats2half2_rn(sr, si);
        *reinterpret_cast<__half2*>(specg + og) = g;
        if (specl != nullptr)
          *reinterpret_cast<__half2*>(specl + og) = __floats2half2_rn(sr - __low2float(g), si - __high2float(g));
      }
    }
  }
}

cudaError_t polar_spec(const float* logamp, const float* phase, int ld, int n_bins, const RowMap& map, float* spec,
                       int lds, __half* specg, int ldsh, __half* specl, float gscale, cudaStream_t st) {
  dim3 grid(map.n_tiles, split_rows(map.n_tiles, 16, 129));
  if (spec == nullptr && specl == nullptr && specg != nullptr)
    launch_k(polar_kernel<true>, grid, 256, 0, st, logamp, phase, ld, n_bins, map, spec, lds, specg, ldsh, specl, gscale);
  else
    launch_k(polar_kernel<false>, grid, 256, 0, st, logamp, phase, ld, n_bins, map, spec, lds, specg, ldsh, specl, gscale);
  return cudaGetLastError();
}

// frames[row, n] = window[n] * irfft(spec[row])[424 + n], n in [0, 1200): the only non-zero part
// of every windowed frame.  Output sample t = 75 h + c of an utterance (after the n_fft/2 trim)
// receives frame f's tap n = 75 (h - f + 8) + c for f in [h - 7, h + 8], f in [0, L] (L + 1
// frames exist).  audio = tanh(sum / sum w^2)  with the same index set for the envelope.
__global__ void __launch_bounds__(256) ola_tanh_kernel(const float* __restrict__ frames, int ldf,
                                                       const float* __restrict__ wsq, const int* __restrict__ cu4,
                                                       RowMap map, float* __restrict__ audio) {
  pdl_enter();
  __shared__ float s_wsq[1200];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  for (int i = threadIdx.x; i < 1200; i += blockDim.x) s_wsq[i] = wsq[i];
  __syncthreads();
  const int nvalid = map.tile_nvalid[tile];
  const int s0 = map.seg_start[seg], len = map.seg_len[seg];
  const int h0 = tile * 128 - s0;
  const size_t base = (size_t)cu4[seg] * 75;
  const int total = nvalid * 75;
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < total; i += gridDim.y * blockDim.x) {
    const int lr = i / 75, c = i - lr * 75;
    const int h = h0 + lr;
    float acc = 0.f, env = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int f = h - 7 + j;              // frame index
      const int n = 75 * (15 - j) + c;      // its tap: 75 (h - f + 8) + c
      if (f >= 0 && f <= len) {
        acc += frames[(size_t)(s0 + f) * ldf + n];
        env += s_wsq[n];
      }
    }
    audio[base + (size_t)h * 75 + c] = tanhf(acc / env);
  }
}

cudaError_t ola_tanh(const float* frames, int ldf, const float* wsq, const int* cu4, const RowMap& map, float* audio,
                     cudaStream_t st) {
  dim3 grid(map.n_tiles, split_rows(map.n_tiles, 4, 38));
  launch_k(ola_tanh_kernel, grid, 256, 0, st, frames, ldf, wsq, cu4, map, audio);
  return cudaGetLastError();
}

// int16 PCM of a waveform: trunc(x * 32768) like numpy's `np.multiply(x, 32768).astype(np.int16)`
// (train/test_onnx.py:79), saturated at the int16 range (the reference cast is undefined for x = +1.0).
// 8 samples per thread: two float4 loads, one 16-byte store; the tail is scalar.
__global__ void __launch_bounds__(256) pcm16_kernel(const float* __restrict__ x, int16_t* __restrict__ y, long long n) {
  pdl_enter();
  const long long i8 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i8 >= n) return;
  auto q = [](float v) {
    const float t = truncf(v * 32768.f);
    return (int16_t)(int)fminf(fmaxf(t, -32768.f), 32767.f);
  };
  if (i8 + 8 <= n && ((reinterpret_cast<uintptr_t>(x + i8) & 15) == 0) && ((reinterpret_cast<uintptr_t>(y + i8) & 15) == 0)) {
    const float4 a = *reinterpret_cast<const float4*>(x + i8), b = *reinterpret_cast<const float4*>(x + i8 + 4);
    union { int16_t h[8]; uint4 v; } o;
    o.h[0] = q(a.x); o.h[1] = q(a.y); o.h[2] = q(a.z); o.h[3] = q(a.w);
    o.h[4] = q(b.x); o.h[5] = q(b.y); o.h[6] = q(b.z); o.h[7] = q(b.w);
    *reinterpret_cast<uint4*>(y + i8) = o.v;
  } else {
    for (long long i = i8; i < n && i < i8 + 8; ++i) y[i] = q(x[i]);
  }
}

cudaError_t pcm16(const float* x, int16_t* y, long long n, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  const long long threads = (n + 7) / 8;
  launch_k(pcm16_kernel, (unsigned)((threads + 255) / 256), 256, 0, st, x, y, n);
  return cudaGetLastError();
}

}  // namespace stts
