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
// Kernel shape: 64 threads per frame, 16 points per thread in registers, three Stockham passes of radix 16, 16, 4
// (1024 = 16 * 16 * 4) with two shared-memory exchanges; a 256-thread CTA runs four frames side by side (one named
// barrier per frame group) and walks 8 frames of a tile.  A frame is the 1200 window taps of the source (16 hop rows
// of 75 samples, reflect padding already materialised by pcph_source), centred in a zero-padded 2048-sample buffer
// (taps 424 .. 1623); the real transform is a 1024-point complex FFT of z[m] = y[2m] + i y[2m + 1] plus the split
// step.  Twiddles come from two fp64 tables through the read-only cache.
#include "common.cuh"
#include "kernels.h"

namespace stts {

namespace {

constexpr int kN = 2048, kH = 1024, kWin = 1200, kOff = 424, kHopS = 75;
constexpr int kGroups = 2;             // frames side by side in a CTA (64 threads each)
constexpr int kFramesPerCta = 8;
constexpr int kPadH = kH + kH / 16;    // smem exchange buffer: index i lives at i + (i >> 4) (bank-conflict-free passes)

typedef double2 cplx;
__device__ __forceinline__ cplx mk(double x, double y) { return make_double2(x, y); }
__device__ __forceinline__ cplx cmul(cplx a, cplx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ cplx cadd(cplx a, cplx b) { return mk(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ cplx csub(cplx a, cplx b) { return mk(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ int pad16(int i) { return i + (i >> 4); }

// forward 4-point DFT in place: (a0, a1, a2, a3) -> (X0, X1, X2, X3)
__device__ __forceinline__ void dft4(cplx& a0, cplx& a1, cplx& a2, cplx& a3) {
  const cplx s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = csub(a1, a3);
  a0 = cadd(s02, s13);
  a2 = csub(s02, s13);
  a1 = mk(d02.x + d13.y, d02.y - d13.x);      // d02 - i d13
  a3 = mk(d02.x - d13.y, d02.y + d13.x);      // d02 + i d13
}

// forward 16-point DFT of x[0..15] (natural order in); X[c + 4 d] is left at position 4 c + d
__device__ __forceinline__ void dft16(cplx* x) {
  constexpr double C1 = 0.92387953251128675613, S1 = 0.38268343236508977173, R2 = 0.70710678118654752440;
#pragma unroll
  for (int a = 0; a < 4; ++a) dft4(x[a], x[a + 4], x[a + 8], x[a + 12]);        // x[a + 4 c] = u[a][c]
  // u[a][c] *= W16^(a c), W16 = exp(-2 pi i / 16)
  x[5] = cmul(x[5], mk(C1, -S1));                                               // a = 1: W^1, W^2, W^3
  x[9] = mk(R2 * (x[9].x + x[9].y), R2 * (x[9].y - x[9].x));
  x[13] = cmul(x[13], mk(S1, -C1));
  x[6] = mk(R2 * (x[6].x + x[6].y), R2 * (x[6].y - x[6].x));                    // a = 2: W^2, W^4, W^6
  x[10] = mk(x[10].y, -x[10].x);
  x[14] = mk(R2 * (x[14].y - x[14].x), -R2 * (x[14].x + x[14].y));
  x[7] = cmul(x[7], mk(S1, -C1));                                               // a = 3: W^3, W^6, W^9
  x[11] = mk(R2 * (x[11].y - x[11].x), -R2 * (x[11].x + x[11].y));
  x[15] = cmul(x[15], mk(-C1, S1));
  // second layer over a for each c; inputs u[0..3][c] sit at x[c*? ...]: u[a][c] = x[a + 4 c]
#pragma unroll
  for (int c = 0; c < 4; ++c) dft4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);   // -> X[c + 4 d] at 4 c + d
}

}  // namespace

// twiddle tables computed once in fp64 (correctly rounded values)
__global__ void stft_twiddle_kernel(double2* __restrict__ tw, double2* __restrict__ tw2) {
  pdl_enter();
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
  launch_k(stft_twiddle_kernel, (kH + 255) / 256, 256, 0, st, tw, tw2);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(64 * kGroups, 4) stft_fft_kernel(
    const float* __restrict__ src, int lds, const float* __restrict__ window, const double2* __restrict__ g_tw,
    const double2* __restrict__ g_tw2, RowMap map, float* __restrict__ mag, float* __restrict__ pha, int ldo,
    __half* __restrict__ magg, __half* __restrict__ magl, __half* __restrict__ phag, __half* __restrict__ phal, int ldh) {
  pdl_enter();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int grp = threadIdx.x >> 6, j = threadIdx.x & 63;
  cplx* A = reinterpret_cast<cplx*>(smem_raw) + grp * kPadH;
  const int seg_first = map.seg_start[seg];
  const int bar_id = 1 + grp;
  auto gsync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory"); };

  for (int lr = blockIdx.y * kFramesPerCta + grp; lr < nvalid && lr < (blockIdx.y + 1) * kFramesPerCta; lr += kGroups) {
    const int row = tile * 128 + lr;                 // frame f of the utterance <-> packed row; its 16 hop rows start at `row`
    cplx x[16];
    // ---- pass 1 (radix 16, Ns = 1): inputs z[j + 64 r] straight from global memory
    //      z[m] = (y[2m], y[2m+1]), y[n] = window[n - 424] * x[n - 424] as an fp32 product (what torch.stft forms)
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int m = j + 64 * r;
      const int s = 2 * m - kOff;                    // window tap of y[2m] (even)
      cplx z = mk(0.0, 0.0);
      if (s >= 0 && s < kWin) {
        const int r0 = s / kHopS, c0 = s - r0 * kHopS;
        const int s1 = s + 1, r1 = s1 / kHopS, c1 = s1 - r1 * kHopS;
        const float2 w = *reinterpret_cast<const float2*>(window + s);
        z.x = (double)__fmul_rn(w.x, src[(size_t)(row + r0) * lds + c0]);
        z.y = (double)__fmul_rn(w.y, src[(size_t)(row + r1) * lds + c1]);
      }
      x[r] = z;
    }
    dft16(x);
    gsync();                                         // previous frame's split step is done with A
#pragma unroll
    for (int sI = 0; sI < 16; ++sI) A[pad16(16 * j + (sI >> 2) + 4 * (sI & 3))] = x[sI];   // y[16 j + p], p = c + 4 d
    gsync();
    // ---- pass 2 (radix 16, Ns = 16): inputs y[j + 64 r] * exp(-2 pi i r k / 256), k = j % 16
    {
      const int k = j & 15;
#pragma unroll
      for (int r = 0; r < 16; ++r) {
        const cplx v = A[pad16(j + 64 * r)];
        x[r] = (r == 0) ? v : cmul(v, __ldg(g_tw + 4 * r * k));
      }
      dft16(x);
      gsync();                                       // every thread of the frame has read its inputs
      const int base = ((j - k) << 4) + k;
#pragma unroll
      for (int sI = 0; sI < 16; ++sI) A[pad16(base + 16 * ((sI >> 2) + 4 * (sI & 3)))] = x[sI];
    }
    gsync();
    // ---- pass 3 (radix 4, Ns = 256): butterflies b = j + 64 c, in place (a thread owns the addresses it touches)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int b = j + 64 * c;
      cplx a0 = A[pad16(b)], a1 = A[pad16(b + 256)], a2 = A[pad16(b + 512)], a3 = A[pad16(b + 768)];
      a1 = cmul(a1, __ldg(g_tw + b));
      a2 = cmul(a2, __ldg(g_tw + 2 * b));
      a3 = cmul(a3, __ldg(g_tw + 3 * b));
      dft4(a0, a1, a2, a3);
      A[pad16(b)] = a0; A[pad16(b + 256)] = a1; A[pad16(b + 512)] = a2; A[pad16(b + 768)] = a3;
    }
    gsync();
    // ---- split step + |X|, atan2: X[k] = E + w^k O, E = (Z[k] + conj Z[N/2 - k]) / 2, O = -i (Z[k] - conj Z[N/2 - k]) / 2
    const bool real_frame = (row == seg_first);      // frame 0 is even about its centre: X is exactly real
#pragma unroll 1
    for (int k = j; k <= kH; k += 64) {
      const cplx zk = A[pad16(k & (kH - 1))];
      cplx zc = A[pad16((kH - k) & (kH - 1))];
      zc.y = -zc.y;
      const cplx e = mk(0.5 * (zk.x + zc.x), 0.5 * (zk.y + zc.y));
      const cplx d = mk(0.5 * (zk.x - zc.x), 0.5 * (zk.y - zc.y));
      const cplx o = mk(d.y, -d.x);                   // -i d
      cplx w;
      if (k <= kH / 2) w = __ldg(g_tw2 + k);
      else { const cplx v = __ldg(g_tw2 + (kH - k)); w = mk(-v.x, v.y); }   // k = 1024 - q: exp(-i pi) * conj(tw2[q])
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
  constexpr int kSmem = kGroups * kPadH * (int)sizeof(double2);     // 69 632 B
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(stft_fft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  launch_k(stft_fft_kernel, grid, 64 * kGroups, kSmem, st, src, lds, window, tw, tw2, map, mag, pha, ldo, magg, magl, phag, phal,
                                                   ldh);
  return cudaGetLastError();
}

}  // namespace stts
