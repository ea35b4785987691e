// Forward STFT of the harmonic-plus-noise source as a shared-memory FFT: TorchSTFT.transform
// (train/models/generator.py:32-44: torch.stft n_fft 2048, hop 75, hann(1200), center / reflect, onesided) followed by
// |X| and atan2 phase with the last frame dropped (generator.py:407-410).
//
// Why an FFT and not the DFT-as-GEMM of round 1: the atan2 phase goes straight into phase_prior_conv.  A bin with
// Re < 0 and |Im| below the transform's own rounding noise flips by 2 pi and moves ~7 output frames by 5e-3 .. 8e-3.
// The reference's fp32 FFT has a noise floor of ~eps * log2(N) relative to the frame's rms (its fp32 and fp64 runs
// agree on every one of the 19.7 M bins of a 60 s utterance); a K = 1200 contraction on the tensor cores (three-pass
// tf32 / fp16 splits, 2^-21 per product) sits ~30x above that floor and flipped ~40 bins per minute of audio.  The
// radix-4 Stockham FFT below has the reference's noise class, needs 50x fewer FLOPs than the contraction and is
// bound by its 8 KB-per-frame output, not by the tensor pipe.
//
// The transform runs in fp64 on the fp32 products window * x (what the reference feeds its FFT): the result is the
// correctly rounded exact DFT of the reference's own input, i.e. as close to the reference's fp32 FFT as that FFT's
// own rounding allows (measured: a bin with Im / Re = 1.1e-5 flipped under an fp32 radix-4 FFT).  3 GFLOP of fp64 per
// 52 k-frame batch is ~0.1 ms on B200.
//
// One CTA (256 threads) transforms FRAMES consecutive frames of a tile, one after the other, sharing the twiddle
// tables in shared memory.  A frame is the 1200 window taps of the source (16 hop rows of 75 samples, reflect padding
// already materialised by pcph_source), centred in a zero-padded 2048-sample buffer (taps 424 .. 1623); the real
// transform is a 1024-point complex FFT of z[m] = y[2m] + i y[2m + 1] plus the usual split step.
#include "common.cuh"
#include "kernels.h"

namespace stts {

namespace {

constexpr int kN = 2048, kH = 1024, kWin = 1200, kOff = 424, kHopS = 75;
constexpr int kFramesPerCta = 4;

typedef double2 cplx;
__device__ __forceinline__ cplx mk(double x, double y) { return make_double2(x, y); }
__device__ __forceinline__ cplx cmul(cplx a, cplx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// smem: tw[1024] = exp(-2 pi i m / 1024), tw2[513] = exp(-2 pi i k / 2048), two ping-pong buffers of 1024 complex
struct Smem {
  cplx tw[kH];
  cplx tw2[kH / 2 + 1];
  cplx a[kH];
  cplx b[kH];
};

}  // namespace

// twiddle tables computed once in fp64 (exact octant symmetries are not needed: correctly rounded values)
__global__ void stft_twiddle_kernel(double2* __restrict__ tw, double2* __restrict__ tw2) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < kH) {
    double s, c;
    sincospi(-2.0 * (double)i / (double)kH, &s, &c);
    tw[i] = make_double2(c, s);
  }
  if (i <= kH / 2) {
    double s, c;
    sincospi(-2.0 * (double)i / (double)kN, &s, &c);
    tw2[i] = make_double2(c, s);
  }
}

cudaError_t stft_twiddles(double2* tw, double2* tw2, cudaStream_t st) {
  stft_twiddle_kernel<<<(kH + 255) / 256, 256, 0, st>>>(tw, tw2);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) stft_fft_kernel(const float* __restrict__ src, int lds,
                                                       const float* __restrict__ window,
                                                       const double2* __restrict__ g_tw, const double2* __restrict__ g_tw2,
                                                       RowMap map, float* __restrict__ mag, float* __restrict__ pha,
                                                       int ldo, __half* __restrict__ magg, __half* __restrict__ magl,
                                                       __half* __restrict__ phag, __half* __restrict__ phal, int ldh) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int lr0 = blockIdx.y * kFramesPerCta;
  if (lr0 >= nvalid) return;
  const int t = threadIdx.x;
  for (int i = t; i < kH; i += 256) sm.tw[i] = g_tw[i];
  for (int i = t; i <= kH / 2; i += 256) sm.tw2[i] = g_tw2[i];
  const int seg_first = map.seg_start[seg];

  for (int fi = 0; fi < kFramesPerCta && lr0 + fi < nvalid; ++fi) {
    const int row = tile * 128 + lr0 + fi;          // frame f of the utterance <-> packed row; its 16 hop rows start at `row`
    __syncthreads();                                 // previous frame's post-processing done with sm.a / sm.b (and tables loaded)
    // ---- load: z[m] = (y[2m], y[2m+1]), y[n] = window[n - 424] * x[n - 424], n in [424, 1624)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = t + 256 * j;
      cplx z = mk(0.0, 0.0);
      const int s = 2 * m - kOff;                    // window tap of y[2m] (even: 424 is even)
      if (s >= 0 && s < kWin) {
        const int r0 = s / kHopS, c0 = s - r0 * kHopS;
        const int s1 = s + 1, r1 = s1 / kHopS, c1 = s1 - r1 * kHopS;
        z.x = (double)__fmul_rn(window[s], src[(size_t)(row + r0) * lds + c0]);      // fp32 product, as torch.stft forms it
        z.y = (double)__fmul_rn(window[s1], src[(size_t)(row + r1) * lds + c1]);
      }
      sm.a[m] = z;
    }
    // ---- 1024-point complex FFT, radix-4 Stockham autosort: 5 stages, natural order in and out
    cplx* in = sm.a;
    cplx* out = sm.b;
#pragma unroll
    for (int stage = 0; stage < 5; ++stage) {
      const int Ns = 1 << (2 * stage);               // 1, 4, 16, 64, 256
      __syncthreads();
      const int k = t & (Ns - 1);
      const int tstep = 256 / Ns;                    // twiddle index step: exp(-2 pi i k / (4 Ns)) = tw[k * 256 / Ns]
      cplx x0 = in[t], x1 = in[t + 256], x2 = in[t + 512], x3 = in[t + 768];
      if (stage > 0) {
        x1 = cmul(x1, sm.tw[k * tstep]);
        x2 = cmul(x2, sm.tw[2 * k * tstep]);
        x3 = cmul(x3, sm.tw[3 * k * tstep]);
      }
      // radix-4 butterfly (forward: -i rotation)
      const cplx s02 = mk(x0.x + x2.x, x0.y + x2.y), d02 = mk(x0.x - x2.x, x0.y - x2.y);
      const cplx s13 = mk(x1.x + x3.x, x1.y + x3.y), d13 = mk(x1.x - x3.x, x1.y - x3.y);
      const cplx y0 = mk(s02.x + s13.x, s02.y + s13.y);
      const cplx y2 = mk(s02.x - s13.x, s02.y - s13.y);
      const cplx y1 = mk(d02.x + d13.y, d02.y - d13.x);      // d02 - i d13
      const cplx y3 = mk(d02.x - d13.y, d02.y + d13.x);      // d02 + i d13
      const int base = ((t - k) << 2) + k;           // (t / Ns) * 4 Ns + k
      out[base] = y0;
      out[base + Ns] = y1;
      out[base + 2 * Ns] = y2;
      out[base + 3 * Ns] = y3;
      cplx* tmp = in; in = out; out = tmp;
    }
    __syncthreads();
    const cplx* Z = in;                            // after 5 swaps: result buffer
    // ---- split step + |X|, atan2: X[k] = E + w^k O, E = (Z[k] + conj Z[N/2 - k]) / 2, O = -i (Z[k] - conj Z[N/2 - k]) / 2
    const bool real_frame = (row == seg_first);      // frame 0 is even about its centre: X is exactly real
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int k = t + 256 * j;
      if (k > kH) break;
      const cplx zk = Z[k & (kH - 1)];
      cplx zc = Z[(kH - k) & (kH - 1)];
      zc.y = -zc.y;
      const cplx e = mk(0.5 * (zk.x + zc.x), 0.5 * (zk.y + zc.y));
      const cplx d = mk(0.5 * (zk.x - zc.x), 0.5 * (zk.y - zc.y));
      const cplx o = mk(d.y, -d.x);                   // -i d
      cplx w;
      if (k <= kH / 2) w = sm.tw2[k];
      else { const cplx v = sm.tw2[kH - k]; w = mk(-v.x, v.y); }   // k = 1024 - q: exp(-i pi) * conj(tw2[q])
      const cplx wo = cmul(w, o);
      const float re = (float)(e.x + wo.x);           // one rounding of the exact value: the fp32 the reference's FFT aims at
      float im = (float)(e.y + wo.y);
      if (k == 0 || k == kH || real_frame) im = 0.f; // DC / Nyquist are real by construction: atan2(+0, -x) = +pi like the FFT
      const float mg = sqrtf(fmaf(re, re, im * im));
      const float ph = atan2f(im, re);
      const size_t o32 = (size_t)row * ldo + k;
      if (mag != nullptr) mag[o32] = mg;
      if (pha != nullptr) pha[o32] = ph;
      const size_t o16 = (size_t)row * ldh + k;
      if (magg != nullptr) {
        __half g, l;
        f16_split(mg, &g, &l);
        magg[o16] = g;
        if (magl != nullptr) magl[o16] = l;
      }
      if (phag != nullptr) {
        __half g, l;
        f16_split(ph, &g, &l);
        phag[o16] = g;
        if (phal != nullptr) phal[o16] = l;
      }
    }
  }
}

cudaError_t stft_fft(const float* src, int lds, const float* window, const double2* tw, const double2* tw2,
                     const RowMap& map, float* mag, float* pha, int ldo, __half* magg, __half* magl, __half* phag,
                     __half* phal, int ldh, cudaStream_t st) {
  dim3 grid(map.n_tiles, (128 + kFramesPerCta - 1) / kFramesPerCta);
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(stft_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  stft_fft_kernel<<<grid, 256, sizeof(Smem), st>>>(src, lds, window, tw, tw2, map, mag, pha, ldo, magg, magl, phag, phal,
                                                   ldh);
  return cudaGetLastError();
}

}  // namespace stts
