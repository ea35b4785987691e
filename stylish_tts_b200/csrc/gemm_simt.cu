// fp32 CUDA-core implicit-GEMM conv1d (the "fp32" precision mode and the on-device cross-check
// for the tcgen05 kernels).  Covers every dense contraction of the path:
//   AdaptiveDecoderBlock convs k3 + 1x1 shortcut   ada_norm.py:157-182
//   WN in_layers k5, res_skip / pre / proj Linears flow.py:63-88,196-218
//   ConvNeXt pwconv1/2, projector, prior convs k7, output heads k7   generator.py:380-423
//   STFT / iSTFT written as DFT contractions        generator.py:32-56
// Tile 128 x BN x 16, 256 threads, register-prefetch double buffering, true fp32 FMA.
#include "common.cuh"

namespace stts {

namespace {

constexpr int BM = 128;
constexpr int BK = 16;
constexpr int APAD = 4;

template <int BN, int KIND>
__global__ void __launch_bounds__(256) conv_gemm_simt_kernel(const ConvGemm g, const RowMap map) {
  pdl_enter();
  constexpr int TN = BN / 16;           // columns per thread: 4 or 8
  constexpr int BLOADS = BN / 64;       // float4 B loads per thread: 1 or 2
  __shared__ __align__(16) float As[2][BK][BM + APAD];
  __shared__ __align__(16) float Bs[2][BK][BN + APAD];

  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int m0 = tile * BM;
  const int n0 = blockIdx.y * BN;
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;

  float acc[8][TN];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // chunk enumeration: (kseg, tap, c0)
  int n_chunks = 0;
  for (int s = 0; s < g.n_kseg; ++s) n_chunks += g.seg[s].taps * (g.seg[s].cin_pad / BK);

  float4 ra[2], rb[BLOADS];

  auto fetch = [&](int chunk) {
    // locate chunk
    int s = 0, c = chunk;
    int per = g.seg[0].taps * (g.seg[0].cin_pad / BK);
    int koff = 0;
    if (c >= per) { c -= per; koff = g.seg[0].taps * g.seg[0].cin_pad; s = 1; }
    const KSeg& ks = g.seg[s];
    const int cpb = ks.cin_pad / BK;
    const int tap = c / cpb;
    const int c0 = (c - tap * cpb) * BK;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 256;
      const int row = idx >> 2, kq = idx & 3;
      const int ar = m0 + row + tap * (ks.dil > 0 ? ks.dil : 1) - ks.pad;
      if (ar >= 0 && ar < ks.rows)
        ra[i] = *reinterpret_cast<const float4*>(ks.A + (size_t)ar * ks.lda + ks.col0 + c0 + kq * 4);
      else
        ra[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const int kcol = koff + tap * ks.cin_pad + c0;
#pragma unroll
    for (int i = 0; i < BLOADS; ++i) {
      const int idx = tid + i * 256;
      const int row = idx >> 2, kq = idx & 3;
      rb[i] = *reinterpret_cast<const float4*>(g.W + (size_t)(n0 + row) * g.Ktot + kcol + kq * 4);
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 256;
      const int row = idx >> 2, kq = idx & 3;
      As[buf][kq * 4 + 0][row] = ra[i].x;
      As[buf][kq * 4 + 1][row] = ra[i].y;
      As[buf][kq * 4 + 2][row] = ra[i].z;
      As[buf][kq * 4 + 3][row] = ra[i].w;
    }
#pragma unroll
    for (int i = 0; i < BLOADS; ++i) {
      const int idx = tid + i * 256;
      const int row = idx >> 2, kq = idx & 3;
      Bs[buf][kq * 4 + 0][row] = rb[i].x;
      Bs[buf][kq * 4 + 1][row] = rb[i].y;
      Bs[buf][kq * 4 + 2][row] = rb[i].z;
      Bs[buf][kq * 4 + 3][row] = rb[i].w;
    }
  };

  fetch(0);
  stash(0);
  __syncthreads();
  int cur = 0;
  for (int ch = 0; ch < n_chunks; ++ch) {
    if (ch + 1 < n_chunks) fetch(ch + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[8], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[cur][k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[cur][k][64 + ty * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
      a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[cur][k][tx * 4]);
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
      if constexpr (TN == 8) {
        const float4 b1 = *reinterpret_cast<const float4*>(&Bs[cur][k][64 + tx * 4]);
        b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (ch + 1 < n_chunks) stash(cur ^ 1);
    __syncthreads();
    cur ^= 1;
  }

  // epilogue: rows ty*4+{0..3}, 64+ty*4+{0..3}; column groups of 4
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int lr = (i < 4) ? (ty * 4 + i) : (64 + ty * 4 + (i - 4));
    if (lr >= nvalid) continue;
    const int r = m0 + lr;
    epi_apply<KIND, 4>(g.epi, r, seg, n0 + tx * 4, &acc[i][0]);
    if constexpr (TN == 8) epi_apply<KIND, 4>(g.epi, r, seg, n0 + 64 + tx * 4, &acc[i][4]);
  }
}

template <int KIND>
cudaError_t launch_kind(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  const int n_cols = g.epi.N;
  if (n_cols > 64) {
    dim3 grid(map.n_tiles, (n_cols + 127) / 128);
    launch_k(conv_gemm_simt_kernel<128, KIND>, grid, 256, 0, st, g, map);
  } else {
    dim3 grid(map.n_tiles, 1);
    launch_k(conv_gemm_simt_kernel<64, KIND>, grid, 256, 0, st, g, map);
  }
  return cudaGetLastError();
}

}  // namespace

cudaError_t conv_gemm_simt(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  switch (g.epi.kind) {
    case EPI_BIAS: return launch_kind<EPI_BIAS>(g, map, st);
    case EPI_RES_SCALE: return launch_kind<EPI_RES_SCALE>(g, map, st);
    case EPI_SILU: return launch_kind<EPI_SILU>(g, map, st);
    case EPI_ADD: return launch_kind<EPI_ADD>(g, map, st);
    case EPI_GATE: return launch_kind<EPI_GATE>(g, map, st);
    case EPI_RESSKIP: return launch_kind<EPI_RESSKIP>(g, map, st);
    case EPI_COUPLE: return launch_kind<EPI_COUPLE>(g, map, st);
    case EPI_PRIOR: return launch_kind<EPI_PRIOR>(g, map, st);
    case EPI_STFT: return launch_kind<EPI_STFT>(g, map, st);
    case EPI_RELU: return launch_kind<EPI_RELU>(g, map, st);
    case EPI_GELU: return launch_kind<EPI_GELU>(g, map, st);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace stts
