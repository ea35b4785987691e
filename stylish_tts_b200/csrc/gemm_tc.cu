// tcgen05 / TMA / TMEM implicit-GEMM conv1d for sm_100a.
//
// One CTA computes one 128 x BN output tile of  Y = sum_taps A[r + tap - pad, :] * W[:, tap, :]^T.
//   warp 0      : TMA producer  (cp.async.bulk.tensor.2d, 128B swizzle, mbarrier complete_tx)
//   warp 1      : TMEM allocator + MMA issuer (one elected lane, tcgen05.mma.cta_group::1)
//   warps 2..5  : operand converter during the main loop (fp32 mode only) and epilogue
//                 (tcgen05.ld 32x32b -> registers -> fused epilogue -> global)
// Two arithmetic modes:
//   MODE_F16   : kind::f16, fp16 operands (16-bit twins of the activations), fp32 accumulate, one pass.
//   MODE_TF32X3 : kind::tf32 with error-compensated operand splitting, x = hi + lo with hi the
//                 tf32 truncation of x: acc += hi*hi' + hi*lo' + lo*hi'  (~2^-21 relative, i.e.
//                 fp32-grade).  The activation split is done in shared memory
//                 by the converter warps (no extra HBM traffic); weight hi/lo are precomputed.
// The conv zero padding comes from the guard rows of the packed layout and from TMA
// out-of-bounds zero fill at the two ends of the buffer (signed row coordinate).
#include <cuda.h>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"

namespace stts {

namespace {

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// try_wait with a suspend-time hint: the hardware parks the thread until the phase completes or the hint expires,
// instead of returning after a few dozen cycles.  Used by the MANY-thread waits (8-16 epilogue warps waiting for the
// accumulator and, with two CTAs per SM, through the other CTA's whole main loop): without it more than half of the
// warp instructions of the short-K GEMMs were polls competing with the working epilogue warps for issue slots
// (persistent decoder convs -9 %, pwconv1 -6 %).  The single-thread pipeline waits of the TMA producer and the MMA
// issuer keep the plain poll: waking from a parked wait costs a little latency, which one 5 s utterance (one CTA
// per kernel) showed as +1.5 %.
constexpr uint32_t kMbarSuspendHint = 0x989680u;
template <bool PARK>
__device__ __forceinline__ bool mbar_try_wait_t(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  if constexpr (PARK) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(kMbarSuspendHint)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
  return ok != 0;
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) { return mbar_try_wait_t<false>(bar, parity); }
// Bounded wait: a protocol bug must trap (launch error) instead of hanging the GPU box.
// park: use the suspend-time hint (see above); a kernel-uniform runtime choice (STTS_MBAR_PARK, mbar_park_mode())
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag, bool park = false) {
  if (park ? mbar_try_wait_t<true>(bar, parity) : mbar_try_wait_t<false>(bar, parity)) return;
  uint64_t t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    if (park ? mbar_try_wait_t<true>(bar, parity) : mbar_try_wait_t<false>(bar, parity)) return;
    uint64_t t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > 2000000000ull) break;  // 2 s
  }
  printf("stts gemm_tc: mbarrier timeout tag=%d block=(%d,%d) thread=%d parity=%u\n", tag, blockIdx.x, blockIdx.y,
         threadIdx.x, parity);
  __trap();
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
template <bool TF32>
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  if constexpr (TF32) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
  }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 64 consecutive accumulator columns of this thread's TMEM lane (4 x 32x32b.x16, one wait)
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, float* v) {
  uint32_t r[64];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[16 * i + 0]), "=r"(r[16 * i + 1]), "=r"(r[16 * i + 2]), "=r"(r[16 * i + 3]), "=r"(r[16 * i + 4]),
          "=r"(r[16 * i + 5]), "=r"(r[16 * i + 6]), "=r"(r[16 * i + 7]), "=r"(r[16 * i + 8]), "=r"(r[16 * i + 9]),
          "=r"(r[16 * i + 10]), "=r"(r[16 * i + 11]), "=r"(r[16 * i + 12]), "=r"(r[16 * i + 13]),
          "=r"(r[16 * i + 14]), "=r"(r[16 * i + 15])
        : "r"(taddr + 16u * i));
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 64; ++i) v[i] = __uint_as_float(r[i]);
}
constexpr int kStgLd = 66;   // staging row stride in floats (even: float2 aligned; 66 % 32 = 2: conflict-free 64-bit)
constexpr int kStgLd4 = 68;  // staging row stride of the float4 (lane-per-4-columns) epilogues: 16-byte aligned rows

// K-major, 128-byte-swizzled shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout):
// start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) | version=1 [46,48) | layout SWIZZLE_128B=2 [61,64)
__device__ __forceinline__ uint64_t make_sdesc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                 // LBO (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;       // SBO: 8 rows * 128 B
  d |= (uint64_t)1 << 46;                 // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                 // SWIZZLE_128B
  return d;
}

// round-to-nearest tf32 (unbiased split: the 3-pass error then grows like sqrt(K), not K)
__device__ __forceinline__ float rna_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// Read-modify-write epilogue of one 32-row x 64-column slab, lane-per-4-columns: lanes 0-15 cover the
// 64 columns of one row with float4 accesses, lanes 16-31 the next row, so one warp instruction moves two
// full 256-byte row segments.  dst / dsth point at (first row of the slab, first column of the slab).
// (The lane-per-column-pair form with its per-row pointer arithmetic and null checks was issue-bound:
// ncu counted ~67 instructions per row and warp, 52 % of the fused layer's tile time.)
template <bool HAS_PRE, bool HAS_TWIN>
__device__ __forceinline__ void rmw_slab4(const float* __restrict__ stg, float* __restrict__ dst, int ld,
                                          twin_t* __restrict__ dsth, int ldh, float4 bias, int nrows, int lane) {
  const int sub = lane >> 4, c4 = (lane & 15) * 4;
  float* p = dst + (size_t)sub * ld + c4;
  twin_t* ph = HAS_TWIN ? dsth + (size_t)sub * ldh + c4 : nullptr;
  const float* sp = stg + sub * kStgLd4 + c4;
#pragma unroll 1
  for (int rb = 0; rb < 32; rb += 16) {
    float4 pre[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      pre[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (HAS_PRE && rb + 2 * i + sub < nrows) pre[i] = *reinterpret_cast<const float4*>(p + (size_t)(rb + 2 * i) * ld);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (rb + 2 * i + sub < nrows) {
        const float4 a = *reinterpret_cast<const float4*>(sp + (rb + 2 * i) * kStgLd4);
        float4 y;
        y.x = pre[i].x + (a.x + bias.x); y.y = pre[i].y + (a.y + bias.y);
        y.z = pre[i].z + (a.z + bias.z); y.w = pre[i].w + (a.w + bias.w);
        *reinterpret_cast<float4*>(p + (size_t)(rb + 2 * i) * ld) = y;
        if constexpr (HAS_TWIN) {
          const twin2_t t0 = to_twin2(y.x, y.y), t1 = to_twin2(y.z, y.w);
          uint2 w;
          w.x = *reinterpret_cast<const uint32_t*>(&t0);
          w.y = *reinterpret_cast<const uint32_t*>(&t1);
          *reinterpret_cast<uint2*>(ph + (size_t)(rb + 2 * i) * ldh) = w;
        }
      }
    }
  }
}

// One 32-row x 64-column slab of a "simple" epilogue in the lane-per-4-columns form (EpiQuad): NB
// iterations (2 rows each) of residual loads are issued before the stores of the batch.
template <int KIND, bool FAST, int NB>
__device__ __forceinline__ void quad_slab(const Epi& epi, const float* __restrict__ stg, int row0, int nrows, int col,
                                          int tile, int q, int lane) {
  EpiQuad<KIND, FAST> eq;
  eq.init(epi, col + (lane & 15) * 4);
  const int sub = lane >> 4;
  const float* sp = stg + sub * kStgLd4 + (lane & 15) * 4;
  if (eq.vec) {                                 // uniform for all slabs but the one that crosses the logical N
    eq.bind(epi, row0 + sub);
    const int nit = (nrows - sub + 1) >> 1;     // iterations (2 rows each) with a valid row for this lane
#pragma unroll 1
    for (int kb = 0; kb < 16; kb += NB) {
      float4 pre[NB];
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        pre[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (EpiQuad<KIND, FAST>::kHasPre && kb + i < nit) pre[i] = eq.pre_vec(kb + i);
      }
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        if (kb + i < nit) {
          const float4 a = *reinterpret_cast<const float4*>(sp + (kb + i) * (2 * kStgLd4));
          eq.apply_vec(kb + i, a, pre[i]);
        }
      }
    }
    eq.finish(epi, tile, q, lane);
    return;
  }
#pragma unroll 1
  for (int rb = 0; rb < 32; rb += 2 * NB) {
    float4 pre[NB];
    if constexpr (EpiQuad<KIND, FAST>::kHasPre) {
#pragma unroll
      for (int i = 0; i < NB; ++i) {
        pre[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rb + 2 * i + sub < nrows) pre[i] = eq.pre(epi, row0 + rb + 2 * i + sub);
      }
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      if (rb + 2 * i + sub < nrows) {
        const float4 a = *reinterpret_cast<const float4*>(sp + (rb + 2 * i) * kStgLd4);
        eq.apply(epi, row0 + rb + 2 * i + sub, a,
                 EpiQuad<KIND, FAST>::kHasPre ? pre[i] : make_float4(0.f, 0.f, 0.f, 0.f));
      }
    }
  }
  eq.finish(epi, tile, q, lane);
}

// EPI_STFT slab with row pointers bound once (lane = one (re, im) accumulator pair = one bin; staging in the
// lane-per-pair layout, stride kStgLd): |X| and atan2 phase to whichever of the fp32 masters / fp16 split twins
// the caller asked for.  Frame 0 of an utterance is exactly real (see epi_apply).
__device__ __forceinline__ void stft_slab(const Epi& e, const float* __restrict__ stg, int row0, int nrows, int col,
                                          int seg_first_row, int lane) {
  const int n = col + 2 * lane;
  if (n >= e.N) return;
  const int c = n >> 1;
  float* p_m = e.out != nullptr ? e.out + (size_t)row0 * e.ldo + e.col0 + c : nullptr;
  float* p_p = e.out2 != nullptr ? e.out2 + (size_t)row0 * e.ldo2 + e.col20 + c : nullptr;
  twin_t* p_mh = e.outh != nullptr ? e.outh + (size_t)row0 * e.ldh + e.colh0 + c : nullptr;
  twin_t* p_ph = e.outh2 != nullptr ? e.outh2 + (size_t)row0 * e.ldh2 + e.colh20 + c : nullptr;
  __half* p_mg = e.outg != nullptr ? e.outg + (size_t)row0 * e.ldh + e.colh0 + c : nullptr;
  __half* p_ml = (e.outg != nullptr && e.outl != nullptr) ? e.outl + (size_t)row0 * e.ldh + e.colh0 + c : nullptr;
  __half* p_pg = e.outg2 != nullptr ? e.outg2 + (size_t)row0 * e.ldh2 + e.colh20 + c : nullptr;
  __half* p_pl = (e.outg2 != nullptr && e.outl2 != nullptr) ? e.outl2 + (size_t)row0 * e.ldh2 + e.colh20 + c : nullptr;
  const long long so = e.ldo, so2 = e.ldo2, sh = e.ldh, sh2 = e.ldh2;
  const int r_real = seg_first_row - row0;      // row of the slab that is frame 0 of its utterance (or outside)
  const float* sp = stg + 2 * lane;
#pragma unroll 1
  for (int rb = 0; rb < nrows; rb += 8) {
    float2 av[8];
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) av[rr] = *reinterpret_cast<const float2*>(sp + (rb + rr) * kStgLd);
#pragma unroll
    for (int rr = 0; rr < 8; ++rr) {
      const int r = rb + rr;
      if (r < nrows) {
        const float re = av[rr].x, im = (r == r_real) ? 0.f : av[rr].y;
        // |X| without hypot's overflow guard (|X| < 1e5 here) and the phase without the reference's common
        // positive divisor (re, im) / (|X| + 1e-9): atan2 depends on the ratio and the signs only
        const float mag = sqrtf(fmaf(re, re, im * im));
        const float ph = atan2f(im, re);
        if (p_m != nullptr) p_m[r * so] = mag;
        if (p_p != nullptr) p_p[r * so2] = ph;
        if (p_mh != nullptr) p_mh[r * sh] = to_twin(mag);
        if (p_ph != nullptr) p_ph[r * sh2] = to_twin(ph);
        if (p_mg != nullptr) {
          __half g, l;
          f16_split(mag, &g, &l);
          p_mg[r * sh] = g;
          if (p_ml != nullptr) p_ml[r * sh] = l;
        }
        if (p_pg != nullptr) {
          __half g, l;
          f16_split(ph, &g, &l);
          p_pg[r * sh2] = g;
          if (p_pl != nullptr) p_pl[r * sh2] = l;
        }
      }
    }
  }
}

// EPI_LN: row LayerNorm fused into the epilogue.  A 128-row tile sees whole rows (N <= BN, one column tile), but a
// row's columns are spread over the N / 64 epilogue warps of its lane quarter: each thread (= one row of one 64-column
// slab, straight from TMEM) forms x = acc * scale + bias + res, the partial sums go through shared memory (lnbuf:
// [2][4 slabs][128 rows]) with a named barrier per quarter, two passes (mean, then centred squares: the same biased
// two-pass variance as row_ln_kernel).  The normalised values are transposed through the staging block so that the
// affine / ReLU / stores run lane-per-column: coalesced fp32 rows and fp16 split twins.  `res` may alias `out`
// (a row is read and written by this warp group only, writes follow the second barrier).
// LayerNorm epilogue of one warp = (32-row quarter q, 64-column slab).  The accumulators arrive one ROW per lane; global
// memory wants one COLUMN PAIR per lane (a row of the slab = 256 contiguous bytes per warp access).  Measured with cycle
// counters on the first version, which read the residual and the bias in the row-per-lane layout: 4.4 us in that pass
// (row-strided 16-byte loads: 32 cache lines per instruction, 64 generic loads of the bias) and 2.5 us in a
// not-unrolled store loop, 8.7 us per launch against 2.3 us for the K loop of a 128-channel projection.  Now:
//   ln_prefetch (BEFORE the accumulator wait, overlapping the K loop): residual rows, bias and the affine pair of this
//                lane's two columns, coalesced, into registers
//   ln_tail     raw accumulators -> staging (row per lane) | x = acc * scale + bias + res, back to staging (column
//                pair per lane) | row sums / centred squares (row per lane, partner slabs through lnbuf + a named
//                barrier per quarter) | normalise, affine, activation, twins, stores (column pair per lane, 4 rows in flight)
struct LnPre {
  float2 res[32];
  float bias0, bias1, g0, g1, b0, b1;
};
// PRE = false (the 256-column tile: 576 threads, 113 registers each): the residual rows are not held across the K loop
// but loaded, coalesced all the same, where they are added
template <bool PRE>
__device__ __forceinline__ void ln_prefetch(const Epi& e, LnPre& P, int slab, int lane, int row0, int nrows, int seg) {
  const int n = slab * 64 + 2 * lane;
  const bool colok = n < e.N;                          // N % 4 == 0 (launch guard): n + 1 < N as well
  P.bias0 = (colok && e.bias != nullptr) ? e.bias[n] : 0.f;
  P.bias1 = (colok && e.bias != nullptr) ? e.bias[n + 1] : 0.f;
  P.g0 = P.g1 = P.b0 = P.b1 = 0.f;
  if (colok) {
    if (e.ln_seg_stride > 0) {
      const float* g = e.ln_g + (size_t)seg * e.ln_seg_stride;
      P.g0 = 1.f + g[n]; P.g1 = 1.f + g[n + 1]; P.b0 = g[e.N + n]; P.b1 = g[e.N + n + 1];
    } else {
      P.g0 = e.ln_g[n]; P.g1 = e.ln_g[n + 1]; P.b0 = e.ln_b[n]; P.b1 = e.ln_b[n + 1];
    }
  }
  if constexpr (PRE) {
    const float* rp = (e.res != nullptr && colok) ? e.res + (size_t)row0 * e.ldr + e.rcol0 + n : nullptr;
#pragma unroll
    for (int r = 0; r < 32; ++r)
      P.res[r] = (rp != nullptr && r < nrows) ? *reinterpret_cast<const float2*>(rp + (size_t)r * e.ldr) : make_float2(0.f, 0.f);
  }
}

// v: this lane's row of raw accumulators, columns [64 slab, 64 slab + 64).  stg: the warp's 32 x kStgLd4 staging block
// (rows at pitch kStgLd; the 64 floats behind them hold the rows' mean and 1 / sigma)
template <bool PRE>
__device__ __forceinline__ void ln_tail(const Epi& e, float acc_scale, float (&v)[64], const LnPre& P, float* stg, float* lnbuf,
                                        int slab, int q, int lane, int row0, int nrows, int seg, int n_live) {
  static_assert(32 * kStgLd + 64 <= 32 * kStgLd4, "row statistics behind the staging rows");
  const int c0 = slab * 64;
  const int rl = q * 32 + lane;                       // row inside the tile
  const int n = c0 + 2 * lane;
  const bool colok = n < e.N;
  float* ms = stg + 32 * kStgLd;                      // [32] mean, [32] 1 / sigma
#pragma unroll
  for (int j = 0; j < 64; j += 2) *reinterpret_cast<float2*>(stg + lane * kStgLd + j) = make_float2(v[j], v[j + 1]);
  __syncwarp();
  if constexpr (PRE) {
#pragma unroll
    for (int r = 0; r < 32; ++r) {                    // column pair per lane
      float2* sp = reinterpret_cast<float2*>(stg + r * kStgLd + 2 * lane);
      const float2 a = *sp;
      float2 x;
      x.x = colok ? fmaf(a.x, acc_scale, P.bias0) + P.res[r].x : 0.f;
      x.y = colok ? fmaf(a.y, acc_scale, P.bias1) + P.res[r].y : 0.f;
      *sp = x;
    }
  } else {
    const float* rp = (e.res != nullptr && colok) ? e.res + (size_t)row0 * e.ldr + e.rcol0 + n : nullptr;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) {
      float2* sp = reinterpret_cast<float2*>(stg + r * kStgLd + 2 * lane);
      const float2 a = *sp;
      const float2 rr = (rp != nullptr && r < nrows) ? *reinterpret_cast<const float2*>(rp + (size_t)r * e.ldr) : make_float2(0.f, 0.f);
      float2 x;
      x.x = colok ? fmaf(a.x, acc_scale, P.bias0) + rr.x : 0.f;
      x.y = colok ? fmaf(a.y, acc_scale, P.bias1) + rr.y : 0.f;
      *sp = x;
    }
  }
  __syncwarp();
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 64; j += 2) {                   // row per lane
    const float2 t = *reinterpret_cast<const float2*>(stg + lane * kStgLd + j);
    v[j] = t.x; v[j + 1] = t.y;
    s += t.x;
    s += t.y;
  }
  const int bar = 1 + q, cnt = 32 * n_live;
  lnbuf[slab * 128 + rl] = s;
  asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(cnt) : "memory");
  float tot = 0.f;
  for (int k = 0; k < n_live; ++k) tot += lnbuf[k * 128 + rl];
  const float mean = tot / (float)e.N;
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < 64; ++j) {
    const float d = (c0 + j < e.N) ? v[j] - mean : 0.f;
    sq += d * d;
  }
  lnbuf[512 + slab * 128 + rl] = sq;
  asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(cnt) : "memory");
  float tq = 0.f;
  for (int k = 0; k < n_live; ++k) tq += lnbuf[512 + k * 128 + rl];
  ms[lane] = mean;
  ms[32 + lane] = rsqrtf(tq / (float)e.N + e.ln_eps);
  __syncwarp();
  if (colok) {
#pragma unroll 4
    for (int r = 0; r < nrows; ++r) {                 // column pair per lane: columns c0 + 2 lane, + 1
      const float2 a = *reinterpret_cast<const float2*>(stg + r * kStgLd + 2 * lane);
      const float m = ms[r], rs = ms[32 + r];
      float o0 = ((a.x - m) * rs) * P.g0 + P.b0, o1 = ((a.y - m) * rs) * P.g1 + P.b1;
      if (e.ln_relu) { o0 = fmaxf(o0, 0.f); o1 = fmaxf(o1, 0.f); }
      const size_t row = (size_t)(row0 + r);
      *reinterpret_cast<float2*>(e.out + row * e.ldo + e.col0 + n) = make_float2(o0, o1);
      if (e.outg != nullptr) {
        __half2 g2, l2;
        f16_split2(o0, o1, &g2, &l2);
        *reinterpret_cast<__half2*>(e.outg + row * e.ldh + e.colh0 + n) = g2;
        if (e.outl != nullptr) *reinterpret_cast<__half2*>(e.outl + row * e.ldh + e.colh0 + n) = l2;
      }
    }
  }
}

template <bool PRE>
__device__ __forceinline__ void ln_epilogue(const Epi& e, float acc_scale, uint32_t taddr, const LnPre& P, float* stg,
                                            float* lnbuf, int slab, int q, int lane, int row0, int nrows, int seg, int n_live) {
  float v[64];
  tmem_ld64(taddr, v);
  ln_tail<PRE>(e, acc_scale, v, P, stg, lnbuf, slab, q, lane, row0, nrows, seg, n_live);
}

// thread-block cluster helpers of the split-K LayerNorm GEMM (SK = 4 below)
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;" ::: "memory"); }
__device__ __forceinline__ uint32_t cluster_map(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
constexpr int kSkPitch = 132;   // floats per row of the exchanged partial accumulators (16-byte rows, 132 % 32 = 4)

// MODE_F16 doubles as the one-pass fp16 GEMM (runtime instruction-descriptor formats), MODE_F16X3 is
// the three-pass fp16 split (hi*hi + hi*lo + lo*hi, fp32-grade), MODE_TF32X3 the tf32 split on fp32 operands
enum { MODE_F16 = 0, MODE_TF32X3 = 1, MODE_F16X3 = 2 };

// OCC = 2 (16-bit mode, BN = 128 only): half the pipeline depth so that two CTAs share an SM (2 x 97 KB
// shared memory, 2 x 128 TMEM columns, 2 x 320 threads x <= 96 registers): the epilogue of one CTA
// overlaps the TMA / MMA main loop of the other -- pays for the short-K GEMMs, whose tiles spend as
// long in the epilogue as in the main loop.
template <int MODE, int BN, int OCC = 1>
struct Cfg {
  static constexpr bool kTF32 = (MODE == MODE_TF32X3);
  static constexpr bool kSplit = (MODE != MODE_F16);     // hi / lo operand pairs, 3 MMAs per k-step
  static constexpr int kElem = kTF32 ? 4 : 2;
  static constexpr int kBK = 128 / kElem;                 // elements per 128-byte swizzle row: 32 / 64
  static constexpr int kUmmaK = 32 / kElem;               // 8 (tf32) / 16 (fp16)
  static constexpr int kATile = 128 * 128;                // bytes
  static constexpr int kBTile = BN * 128;                 // bytes
  static constexpr int kStageBytes = kSplit ? (2 * kATile + 2 * kBTile) : (kATile + kBTile);
  static constexpr int kStages = kSplit ? (BN == 128 ? 3 : 2) : (OCC == 2 ? 3 : (BN == 128 ? 6 : 4));
  // tf32x3: the A residual is computed in shared memory by the converter warps; f16x3: it is loaded
  static constexpr int kTxBytes = kTF32 ? (kATile + 2 * kBTile) : (kSplit ? 2 * kATile + 2 * kBTile : kATile + kBTile);
  static constexpr int kEpiWarps = BN / 16;                // one 64-column slab x 32-row quarter per warp
  static constexpr int kThreads = 64 + 32 * kEpiWarps;
  static constexpr int kSmem = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  // instruction descriptor without the A / B format fields (bits 7-9, 10-12: 0 = f16, 1 = bf16, 2 = tf32)
  static constexpr uint32_t kIdescBase = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
};

struct TcParams {
  int n_kseg;
  int taps[2], pad[2], cin_chunks[2], col0[2];   // col0 in elements of the operand dtype
  int dil[2];           // row step between taps (dilated convs of the snake generator block), >= 1
  int n_chunks;
  int w_seg_rows;       // > 0: the weights are per utterance, utterance u at rows [u * w_seg_rows, ...)
  uint32_t ab_fmt;      // operand format of the MMAs: 0 = f16, 1 = bf16, 2 = tf32
  float acc_scale;      // accumulators * acc_scale before the epilogue (power-of-two operand scaling undone)
  int park;             // mbarrier waits with a suspend hint: 1 = the epilogue warps' accumulator waits, 2 = all
};

// SK = 4 (EPI_LN, BN = 128 only): split-K over a cluster of four CTAs.  The LayerNorm GEMMs have ONE column tile (the
// whole row is normalised), so a long-K layer (text-encoder ffn2: K = 1 536, three MMAs per k-step in the split modes)
// was one CTA per row tile streaming 1.5 MB of operands through one SM: 28 us per launch however small the batch.
// Here CTA r of the cluster runs chunks [r n / 4, (r + 1) n / 4) of the K loop; then the four 128 x 128 fp32 partial
// accumulators are reduce-scattered by row quarter through distributed shared memory (quarter q of every CTA goes to
// CTA q), and CTA q finishes its 32 rows (sum in rank order, bias, residual, LayerNorm, twins): K loop and epilogue are
// both four times shorter, no operand is loaded twice, and the result does not depend on timing.
template <int MODE, int BN, int KIND, int OCC = 1, int SK = 1>
__global__ void __launch_bounds__(64 + 2 * BN, OCC)
conv_gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                    const __grid_constant__ CUtensorMap mapA0lo, const __grid_constant__ CUtensorMap mapA1lo,
                    const __grid_constant__ CUtensorMap mapW, const __grid_constant__ CUtensorMap mapWlo,
                    const __grid_constant__ CUtensorMap mapWt, const __grid_constant__ CUtensorMap mapWtlo,
                    const TcParams p, const Epi epi, const RowMap map) {
  using C = Cfg<MODE, BN, OCC>;
  const int tile = blockIdx.y;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int m0 = tile * 128;
  static_assert(SK == 1 || (SK == 4 && KIND == EPI_LN && BN == 128 && OCC == 1), "split-K: LayerNorm epilogue only");
  const int n0 = (SK > 1) ? 0 : blockIdx.x * BN;
  const int krank = (SK > 1) ? blockIdx.x : 0;            // cluster (SK, 1, 1), gridDim.x == SK: rank in the cluster
  const int ch_begin = (SK > 1) ? (p.n_chunks * krank) / SK : 0;
  const int ch_end = (SK > 1) ? (p.n_chunks * (krank + 1)) / SK : p.n_chunks;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Tail column tile with <= 16 live columns (N = 1 025 output heads, N = 2 050 STFT): 16-row weight boxes and
  // N = 16 MMAs instead of a full BN-wide tile of padding (mapWt / mapWtlo: the same weights, box of 16 rows)
  const bool tail = !C::kTF32 && (epi.N - n0) <= 16;
  constexpr uint32_t kTailBytes = 16 * 128;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;                        // TMA -> consumers
  uint64_t* empty = bars + C::kStages;          // MMA commit -> TMA
  uint64_t* conv = bars + 2 * C::kStages;       // converter -> MMA (tf32x3 only)
  uint64_t* accum = bars + 3 * C::kStages;      // MMA commit -> epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * C::kStages + 1);

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
      mbar_init(&conv[s], C::kEpiWarps);
    }
    mbar_init(accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_enter();   // everything above (barrier init, TMEM allocation) overlaps the previous kernel of the stream
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc_full = C::kIdescBase | (p.ab_fmt << 7) | (p.ab_fmt << 10);
  const uint32_t idesc = tail ? ((idesc_full & ~(0x3Fu << 17)) | ((16u >> 3) << 17)) : idesc_full;

  auto stage_ptr = [&](int s) { return smem + (size_t)s * C::kStageBytes; };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int ks = 0, tap = 0, cc = 0;
      int kcol = 0;
      const int wrow = n0 + seg * p.w_seg_rows;
      const CUtensorMap* mw = tail ? &mapWt : &mapW;
      const CUtensorMap* mwlo = tail ? &mapWtlo : &mapWlo;
      const uint32_t tx = tail ? C::kTxBytes - (C::kSplit ? 2u : 1u) * ((uint32_t)C::kBTile - kTailBytes) : C::kTxBytes;
      for (int ch = 0; ch < ch_end; ++ch) {
        if (SK == 1 || ch >= ch_begin) {
        mbar_wait(&empty[stage], phase ^ 1, 1, p.park >= 2);
        uint8_t* sp = stage_ptr(stage);
        mbar_expect_tx(&full[stage], tx);
        const CUtensorMap* ma = (ks == 0) ? &mapA0 : &mapA1;
        tma_load_2d(sp, ma, &full[stage], p.col0[ks] + cc * C::kBK, m0 + tap * p.dil[ks] - p.pad[ks]);
        if constexpr (C::kSplit) {
          if constexpr (!C::kTF32) {
            const CUtensorMap* ml = (ks == 0) ? &mapA0lo : &mapA1lo;
            tma_load_2d(sp + C::kATile, ml, &full[stage], p.col0[ks] + cc * C::kBK, m0 + tap * p.dil[ks] - p.pad[ks]);
          }
          tma_load_2d(sp + 2 * C::kATile, mw, &full[stage], kcol, wrow);
          tma_load_2d(sp + 2 * C::kATile + C::kBTile, mwlo, &full[stage], kcol, wrow);
        } else {
          tma_load_2d(sp + C::kATile, mw, &full[stage], kcol, wrow);
        }
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        kcol += C::kBK;
        if (++cc == p.cin_chunks[ks]) { cc = 0; if (++tap == p.taps[ks]) { tap = 0; ++ks; } }
      }
    }
    if constexpr (SK > 1) { cluster_arrive(); cluster_wait(); cluster_arrive(); cluster_wait(); }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int ch = ch_begin; ch < ch_end; ++ch) {
        mbar_wait(&full[stage], phase, 2, p.park >= 2);
        if constexpr (C::kTF32) mbar_wait(&conv[stage], phase, 3, p.park >= 2);
        tc_fence_after();
        const uint32_t a_hi = smem_u32(stage_ptr(stage));
#pragma unroll
        for (int k = 0; k < C::kBK / C::kUmmaK; ++k) {
          const uint32_t koff = k * 32;  // bytes inside the 128B swizzle row
          if constexpr (C::kSplit) {
            const uint32_t a_lo = a_hi + C::kATile, b_hi = a_hi + 2 * C::kATile, b_lo = b_hi + C::kBTile;
            // small terms first, then the leading term
            tc_mma<C::kTF32>(tmem_base, make_sdesc(a_lo + koff), make_sdesc(b_hi + koff), idesc, ((ch - ch_begin) | k) != 0);
            tc_mma<C::kTF32>(tmem_base, make_sdesc(a_hi + koff), make_sdesc(b_lo + koff), idesc, 1);
            tc_mma<C::kTF32>(tmem_base, make_sdesc(a_hi + koff), make_sdesc(b_hi + koff), idesc, 1);
          } else {
            const uint32_t b = a_hi + C::kATile;
            tc_mma<false>(tmem_base, make_sdesc(a_hi + koff), make_sdesc(b + koff), idesc, ((ch - ch_begin) | k) != 0);
          }
        }
        tc_commit(&empty[stage]);
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      tc_commit(accum);
    }
    if constexpr (SK > 1) { cluster_arrive(); cluster_wait(); cluster_arrive(); cluster_wait(); }
  } else {
    // ------------------------------------------------------------ converter (tf32x3) + epilogue
    LnPre lnp;                               // EPI_LN: residual rows / bias / affine of this warp's block, loaded during the K loop
    if constexpr (KIND == EPI_LN) {
      const int ew_ = warp - 2;
      if constexpr (SK > 1) {                // the warps that finish this CTA's row quarter (= krank)
        int nr_ = nvalid - krank * 32;
        nr_ = nr_ < 0 ? 0 : (nr_ > 32 ? 32 : nr_);
        if (ew_ < (epi.N + 63) / 64) ln_prefetch<BN == 128>(epi, lnp, ew_, lane, m0 + krank * 32, nr_, seg);
      } else {
        int nr_ = nvalid - (warp & 3) * 32;
        nr_ = nr_ < 0 ? 0 : (nr_ > 32 ? 32 : nr_);
        if ((ew_ >> 2) * 64 < epi.N) ln_prefetch<BN == 128>(epi, lnp, ew_ >> 2, lane, m0 + (warp & 3) * 32, nr_, seg);
      }
    }
    if constexpr (C::kTF32) {
      const int ct = threadIdx.x - 64;  // 0 .. 32 * kEpiWarps - 1
      constexpr int kConvThreads = 32 * C::kEpiWarps;
      int stage = 0;
      uint32_t phase = 0;
      for (int ch = ch_begin; ch < ch_end; ++ch) {
        mbar_wait(&full[stage], phase, 4, p.park >= 1);
        float4* hi = reinterpret_cast<float4*>(stage_ptr(stage));
        float4* lo = reinterpret_cast<float4*>(stage_ptr(stage) + C::kATile);
#pragma unroll
        for (int i = 0; i < 1024 / kConvThreads; ++i) {
          const int idx = ct + i * kConvThreads;
          const float4 x = hi[idx];
          float4 h, l;
          h.x = rna_tf32(x.x); l.x = rna_tf32(x.x - h.x);
          h.y = rna_tf32(x.y); l.y = rna_tf32(x.y - h.y);
          h.z = rna_tf32(x.z); l.z = rna_tf32(x.z - h.z);
          h.w = rna_tf32(x.w); l.w = rna_tf32(x.w - h.w);
          hi[idx] = h;
          lo[idx] = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&conv[stage]);
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
    mbar_wait(accum, 0, 5, p.park >= 1);
    tc_fence_after();
    // All TMA writes and all MMAs of this CTA have completed: the pipeline stages are free and are
    // reused as the transpose staging area (one 32 x 66-float block per epilogue warp).  Warp ew
    // owns TMEM lane quarter q = warp % 4 (hardware rule) and the 64-column slab ew / 4.
    const int ew = warp - 2;
    const int q = warp & 3;
    using EQ = EpiQuad<KIND, MODE == MODE_F16>;
    constexpr int kLd = EQ::kOk ? kStgLd4 : kStgLd;
    float* stg = reinterpret_cast<float*>(smem) + ew * (32 * kLd);
    const int row0 = m0 + q * 32;
    int nrows = nvalid - q * 32;
    nrows = nrows < 0 ? 0 : (nrows > 32 ? 32 : nrows);
    const int seg_first = (epi.seg_start != nullptr) ? epi.seg_start[seg] : -1;
    const int c = (ew >> 2) * 64;
    if constexpr (KIND == EPI_LN && SK > 1) {
      // every CTA of the cluster has finished its K range: all four operand rings are free and take the partials
      cluster_arrive();
      cluster_wait();
      float* part = reinterpret_cast<float*>(smem);            // [SK source CTAs][32 rows][kSkPitch]
      const int slab = ew >> 2;
      if (c < epi.N) {
        float v[64];
        tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        const uint32_t dst = cluster_map(smem_u32(part + (krank * 32 + lane) * kSkPitch + c), (uint32_t)q);
#pragma unroll
        for (int j = 0; j < 64; j += 4) st_cluster_f4(dst + j * 4, v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
      cluster_arrive();
      cluster_wait();
      const int n_live = (epi.N + 63) / 64;
      if (ew < n_live) {                       // warps ew = 0 .. n_live - 1 finish this CTA's row quarter (= krank)
        const int c0 = ew * 64;
        float v[64];
#pragma unroll
        for (int j = 0; j < 64; j += 4) {
          float4 a = *reinterpret_cast<const float4*>(part + lane * kSkPitch + c0 + j);
#pragma unroll
          for (int sr = 1; sr < SK; ++sr) {
            const float4 b = *reinterpret_cast<const float4*>(part + (sr * 32 + lane) * kSkPitch + c0 + j);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
          }
          v[j] = a.x; v[j + 1] = a.y; v[j + 2] = a.z; v[j + 3] = a.w;
        }
        float* tail_smem = part + SK * 32 * kSkPitch;
        int nr = nvalid - krank * 32;
        nr = nr < 0 ? 0 : (nr > 32 ? 32 : nr);
        ln_tail<BN == 128>(epi, p.acc_scale, v, lnp, tail_smem + ew * (32 * kStgLd4), tail_smem + 2 * (32 * kStgLd4), ew, krank, lane,
                m0 + krank * 32, nr, seg, n_live);
      }
      (void)slab;
    } else
    if constexpr (KIND == EPI_LN) {
      if (c < epi.N) {                         // warp-uniform; one column tile (n0 == 0)
        float* lnbuf = reinterpret_cast<float*>(smem) + C::kEpiWarps * (32 * kStgLd4);
        ln_epilogue<BN == 128>(epi, p.acc_scale, tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, lnp,
                    reinterpret_cast<float*>(smem) + ew * (32 * kStgLd4), lnbuf, ew >> 2, q, lane, row0, nrows, seg,
                    (epi.N + 63) / 64);
      }
    } else
    if constexpr (EQ::kOk) {
      if (n0 + c < epi.N) {                    // warp-uniform
        {
          float v[64];
          tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 64; j += 4)
            *reinterpret_cast<float4*>(stg + lane * kStgLd4 + j) =
                make_float4(v[j] * p.acc_scale, v[j + 1] * p.acc_scale, v[j + 2] * p.acc_scale, v[j + 3] * p.acc_scale);
        }
        __syncwarp();
        quad_slab<KIND, MODE == MODE_F16, (BN == 256 || OCC == 2) ? 4 : 8>(epi, stg, row0, nrows, n0 + c, tile, q, lane);
      }
    } else
    if (n0 + c < epi.N) {                      // warp-uniform
      EpiPair<KIND, MODE == MODE_F16> ep;
      ep.init(epi, seg, n0 + c + 2 * lane);
      {
        float v[64];
        tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        __syncwarp();                            // previous slab fully consumed
#pragma unroll
        for (int j = 0; j < 64; j += 2)
          *reinterpret_cast<float2*>(stg + lane * kStgLd + j) = make_float2(v[j] * p.acc_scale, v[j + 1] * p.acc_scale);
      }
      __syncwarp();
      if constexpr (KIND == EPI_STFT) {
        stft_slab(epi, stg, row0, nrows, n0 + c, seg_first, lane);
      } else
      if constexpr (EpiPair<KIND, false>::kHasPre) {
        // batches of RB rows: RB independent residual loads in flight per warp, registers bounded
        // (576-thread CTAs of the 256-column tile have 113 registers per thread)
        constexpr int RB = (BN == 256 || OCC == 2) ? 4 : 16;
#pragma unroll 1
        for (int rb = 0; rb < 32; rb += RB) {
          float2 pre[RB];
#pragma unroll
          for (int rr = 0; rr < RB; ++rr)
            if (rb + rr < nrows) pre[rr] = ep.pre(epi, row0 + rb + rr);
          float2 av[RB];                               // staged accumulators read up front: the loads must
#pragma unroll                                         // not be serialised behind the global stores of apply()
          for (int rr = 0; rr < RB; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
#pragma unroll
          for (int rr = 0; rr < RB; ++rr)
            if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, pre[rr]);
        }
      } else {
#pragma unroll 1
        for (int rb = 0; rb < nrows; rb += 8) {        // 8 rows per batch: loads first, then the epilogue math + stores
          float2 av[8];
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
#pragma unroll
          for (int rr = 0; rr < 8; ++rr)
            if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, make_float2(0.f, 0.f));
        }
      }
      ep.finish(epi, tile, q);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
}

// ------------------------------------------------------------------ one-pass fp16 GEMM, 256 x 256 tiles
// TWO row tiles against the same weight stage: two M = 128 MMAs per k-step share one B descriptor, two 256-column
// accumulators fill the 512 TMEM columns, and the L2 -> SM operand traffic per output drops by a third
// ((256 + 256) / 65 536 vs (128 + 256) / 32 768 bytes per K-element).  One CTA per tile pair x column tile, 16 epilogue
// warps (no TMEM left to double-buffer, so no persistent overlap).  Row tiles are independent (own utterance, own
// valid-row count); per-utterance weight matrices (GRN-folded pwconv2) stay on the other kernels.  Written to test the
// hypothesis that the long-K GEMMs are L2-traffic-bound (their times match a 81 KB/us-per-SM model to 1 %): they are
// not -- the output heads already run at 84 % of the sustained cuBLAS bf16 rate (1 175 of 1 396 TF/s) and gained 3 %;
// the kernel pays where the 8-warp persistent epilogue or the 128-row tile was the limiter (see launch_bn).
struct M2Cfg {
  static constexpr int kBN = 256;
  static constexpr int kBK = 64, kUmmaK = 16;
  static constexpr int kATile = 128 * 128, kBTile = kBN * 128;           // bytes
  static constexpr int kStageBytes = 2 * kATile + kBTile;                // 64 KB
  static constexpr int kStages = 3;
  static constexpr int kEpiWarps = 16;
  static constexpr int kThreads = 64 + 32 * kEpiWarps;                   // 576
  static constexpr int kSmem = kStages * kStageBytes + 1024 + 256;
  static constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  static_assert(kStages * kStageBytes >= kEpiWarps * 32 * kStgLd4 * 4, "staging must fit into the stage ring");
};

template <int KIND>
__global__ void __launch_bounds__(M2Cfg::kThreads, 1)
conv_gemm_tc_m256_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                         const __grid_constant__ CUtensorMap mapW, const __grid_constant__ CUtensorMap mapWt,
                         const TcParams p, const Epi epi, const RowMap map) {
  using C = M2Cfg;
  const int pair = blockIdx.y;
  const int t0 = 2 * pair, t1 = 2 * pair + 1;
  const bool live0 = map.tile_seg[t0] >= 0, live1 = t1 < map.n_tiles && map.tile_seg[t1] >= 0;
  if (!live0 && !live1) return;
  const int n0 = blockIdx.x * C::kBN;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool tail = (epi.N - n0) <= 16;
  constexpr uint32_t kTailBytes = 16 * 128;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* accum = bars + 2 * C::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::kStages + 1);

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_enter();   // everything above (barrier init, TMEM allocation) overlaps the previous kernel of the stream
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc = tail ? ((C::kIdesc & ~(0x3Fu << 17)) | ((16u >> 3) << 17)) : C::kIdesc;
  auto stage_ptr = [&](int s) { return smem + (size_t)s * C::kStageBytes; };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int ks = 0, tap = 0, cc = 0, kcol = 0;
      const CUtensorMap* mw = tail ? &mapWt : &mapW;
      const uint32_t tx = tail ? 2 * C::kATile + kTailBytes : C::kStageBytes;
      for (int ch = 0; ch < p.n_chunks; ++ch) {
        mbar_wait(&empty[stage], phase ^ 1, 31, p.park >= 2);
        uint8_t* sp = stage_ptr(stage);
        mbar_expect_tx(&full[stage], tx);
        const CUtensorMap* ma = (ks == 0) ? &mapA0 : &mapA1;
        const int row = t0 * 128 + tap * p.dil[ks] - p.pad[ks];
        tma_load_2d(sp, ma, &full[stage], p.col0[ks] + cc * C::kBK, row);
        tma_load_2d(sp + C::kATile, ma, &full[stage], p.col0[ks] + cc * C::kBK, row + 128);   // past the buffer: zero fill
        tma_load_2d(sp + 2 * C::kATile, mw, &full[stage], kcol, n0);
        kcol += C::kBK;
        if (++cc == p.cin_chunks[ks]) { cc = 0; if (++tap == p.taps[ks]) { tap = 0; ++ks; } }
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer: two row tiles per weight stage
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int ch = 0; ch < p.n_chunks; ++ch) {
        mbar_wait(&full[stage], phase, 32, p.park >= 2);
        tc_fence_after();
        const uint32_t a0 = smem_u32(stage_ptr(stage)), a1 = a0 + C::kATile, b = a0 + 2 * C::kATile;
#pragma unroll
        for (int k = 0; k < C::kBK / C::kUmmaK; ++k) {
          tc_mma<false>(tmem_base, make_sdesc(a0 + k * 32), make_sdesc(b + k * 32), idesc, (ch | k) != 0);
          tc_mma<false>(tmem_base + (uint32_t)C::kBN, make_sdesc(a1 + k * 32), make_sdesc(b + k * 32), idesc, (ch | k) != 0);
        }
        tc_commit(&empty[stage]);
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      tc_commit(accum);
    }
  } else {
    // ------------------------------------------------------------ epilogue: accumulator 0, then accumulator 1
    mbar_wait(accum, 0, 35, p.park >= 1);
    tc_fence_after();
    const int ew = warp - 2;
    const int q = warp & 3;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {                // accumulator h = row tile 2 * pair + h
      const int tile = 2 * pair + h;
      if (tile >= map.n_tiles) break;
      const int seg = map.tile_seg[tile];
      if (seg < 0) continue;
      const int nvalid = map.tile_nvalid[tile];
      const int m0 = tile * 128;
      __syncwarp();
      using EQ = EpiQuad<KIND, true>;
      constexpr int kLd = EQ::kOk ? kStgLd4 : kStgLd;
      float* stg = reinterpret_cast<float*>(smem) + ew * (32 * kLd);
      const int row0 = m0 + q * 32;
      int nrows = nvalid - q * 32;
      nrows = nrows < 0 ? 0 : (nrows > 32 ? 32 : nrows);
      const int seg_first = (epi.seg_start != nullptr) ? epi.seg_start[seg] : -1;
      const int c = (ew >> 2) * 64;
      if constexpr (EQ::kOk) {
        if (n0 + c < epi.N) {                    // warp-uniform
          {
            float v[64];
            tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * C::kBN + c), v);
            __syncwarp();
  #pragma unroll
            for (int j = 0; j < 64; j += 4)
              *reinterpret_cast<float4*>(stg + lane * kStgLd4 + j) =
                  make_float4(v[j] * p.acc_scale, v[j + 1] * p.acc_scale, v[j + 2] * p.acc_scale, v[j + 3] * p.acc_scale);
          }
          __syncwarp();
          quad_slab<KIND, true, 4>(epi, stg, row0, nrows, n0 + c, tile, q, lane);
        }
      } else
      if (n0 + c < epi.N) {                      // warp-uniform
        EpiPair<KIND, true> ep;
        ep.init(epi, seg, n0 + c + 2 * lane);
        {
          float v[64];
          tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * C::kBN + c), v);
          __syncwarp();                            // previous slab fully consumed
  #pragma unroll
          for (int j = 0; j < 64; j += 2)
            *reinterpret_cast<float2*>(stg + lane * kStgLd + j) = make_float2(v[j] * p.acc_scale, v[j + 1] * p.acc_scale);
        }
        __syncwarp();
        if constexpr (KIND == EPI_STFT) {
          stft_slab(epi, stg, row0, nrows, n0 + c, seg_first, lane);
        } else
        if constexpr (EpiPair<KIND, false>::kHasPre) {
          // batches of RB rows: RB independent residual loads in flight per warp, registers bounded
          // (576-thread CTAs of the 256-column tile have 113 registers per thread)
          constexpr int RB = 4;
  #pragma unroll 1
          for (int rb = 0; rb < 32; rb += RB) {
            float2 pre[RB];
  #pragma unroll
            for (int rr = 0; rr < RB; ++rr)
              if (rb + rr < nrows) pre[rr] = ep.pre(epi, row0 + rb + rr);
            float2 av[RB];                               // staged accumulators read up front: the loads must
  #pragma unroll                                         // not be serialised behind the global stores of apply()
            for (int rr = 0; rr < RB; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
  #pragma unroll
            for (int rr = 0; rr < RB; ++rr)
              if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, pre[rr]);
          }
        } else {
  #pragma unroll 1
          for (int rb = 0; rb < nrows; rb += 8) {        // 8 rows per batch: loads first, then the epilogue math + stores
            float2 av[8];
  #pragma unroll
            for (int rr = 0; rr < 8; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
  #pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, make_float2(0.f, 0.f));
          }
        }
        ep.finish(epi, tile, q);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------ persistent one-pass fp16 GEMM, 128 x 256 tiles
// One CTA per SM loops over output tiles (tile id = m_tile * n_tiles_n + n_tile, grid-stride).  The
// TMA producer and the MMA issuer run ahead into the next tile while the 8 epilogue warps drain the
// previous accumulator: TMEM holds two 256-column accumulators (512 columns), handed over with the
// tmem_full / tmem_empty mbarrier pairs; the epilogue has its own transpose staging (the pipeline
// stages are never idle).  For the short-K layers (flow K 640, pwconv1 K 512) a tile spends as long in
// prologue + epilogue as in the main loop, so the overlap is worth up to 2x; long-K layers gain the
// removed per-CTA prologue (barrier init, TMEM alloc, first-load latency) of the ~3 waves.
struct PCfg {
  static constexpr int kBN = 256;
  static constexpr int kBK = 64, kUmmaK = 16;
  static constexpr int kATile = 128 * 128, kBTile = kBN * 128;           // bytes
  static constexpr int kStageBytes = kATile + kBTile;                    // 48 KB
  static constexpr int kStages = 3;
  static constexpr int kEpiWarps = 8;                                    // 4 lane quarters x 2 column halves
  static constexpr int kThreads = 64 + 32 * kEpiWarps;                   // 320
  static constexpr int kStagingBytes = kEpiWarps * 32 * kStgLd4 * 4;     // 69 632
  static constexpr int kSmem = kStages * kStageBytes + kStagingBytes + 1024 + 256;
  static constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);  // f16 x f16 -> f32
};

template <int KIND>
__global__ void __launch_bounds__(PCfg::kThreads, 1)
conv_gemm_tc_persistent_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                               const __grid_constant__ CUtensorMap mapW, const __grid_constant__ CUtensorMap mapWt,
                               const TcParams p, const Epi epi, const RowMap map, const int n_tiles_n) {
  using C = PCfg;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = map.n_tiles * n_tiles_n;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* staging = reinterpret_cast<float*>(smem + C::kStages * C::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes + C::kStagingBytes);
  uint64_t* full = bars;                         // TMA -> MMA
  uint64_t* empty = bars + C::kStages;           // MMA commit -> TMA
  uint64_t* tmem_full = bars + 2 * C::kStages;   // MMA commit -> epilogue      [2]
  uint64_t* tmem_empty = tmem_full + 2;          // epilogue -> MMA             [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], C::kEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_enter();   // everything above (barrier init, TMEM allocation) overlaps the previous kernel of the stream
  const uint32_t tmem_base = *tmem_slot;
  auto stage_ptr = [&](int s) { return smem + (size_t)s * C::kStageBytes; };

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int tile = t / n_tiles_n, n0 = (t - tile * n_tiles_n) * C::kBN;
        const int seg = map.tile_seg[tile];
        if (seg < 0) continue;
        const int m0 = tile * 128;
        const int wrow = n0 + seg * p.w_seg_rows;
        const bool tail = (epi.N - n0) <= 16;      // <= 16 live columns: 16-row weight box, N = 16 MMAs
        const CUtensorMap* mw = tail ? &mapWt : &mapW;
        const uint32_t tx = tail ? C::kATile + 16 * 128 : C::kStageBytes;
        int ks = 0, tap = 0, cc = 0, kcol = 0;
        for (int ch = 0; ch < p.n_chunks; ++ch) {
          mbar_wait(&empty[stage], phase ^ 1, 11, p.park >= 2);
          uint8_t* sp = stage_ptr(stage);
          mbar_expect_tx(&full[stage], tx);
          const CUtensorMap* ma = (ks == 0) ? &mapA0 : &mapA1;
          tma_load_2d(sp, ma, &full[stage], p.col0[ks] + cc * C::kBK, m0 + tap * p.dil[ks] - p.pad[ks]);
          tma_load_2d(sp + C::kATile, mw, &full[stage], kcol, wrow);
          kcol += C::kBK;
          if (++cc == p.cin_chunks[ks]) { cc = 0; if (++tap == p.taps[ks]) { tap = 0; ++ks; } }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;                                  // local tile counter (accumulator buffer = it & 1)
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int tile = t / n_tiles_n;
        if (map.tile_seg[tile] < 0) continue;
        const bool tail = (epi.N - (t - tile * n_tiles_n) * C::kBN) <= 16;
        const uint32_t idesc = tail ? ((C::kIdesc & ~(0x3Fu << 17)) | ((16u >> 3) << 17)) : C::kIdesc;
        const int buf = it & 1, use = it >> 1;
        if (use > 0) mbar_wait(&tmem_empty[buf], (uint32_t)((use - 1) & 1), 12);   // epilogue drained this buffer
        tc_fence_after();
        const uint32_t acc = tmem_base + (uint32_t)(buf * C::kBN);
        for (int ch = 0; ch < p.n_chunks; ++ch) {
          mbar_wait(&full[stage], phase, 13, p.park >= 2);
          tc_fence_after();
          const uint32_t a = smem_u32(stage_ptr(stage)), b = a + C::kATile;
#pragma unroll
          for (int k = 0; k < C::kBK / C::kUmmaK; ++k)
            tc_mma<false>(acc, make_sdesc(a + k * 32), make_sdesc(b + k * 32), idesc, (ch | k) != 0);
          tc_commit(&empty[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        tc_commit(&tmem_full[buf]);
        ++it;
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue (8 warps)
    const int ew = warp - 2;                       // 0 .. 7
    const int q = warp & 3;                        // TMEM lane quarter this warp may read (hardware rule)
    const int half = ew >> 2;                      // column half: slabs {0, 64} + 128 * half
    float* stg = staging + ew * (32 * kStgLd4);
    int it = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int tile = t / n_tiles_n, n0 = (t - tile * n_tiles_n) * C::kBN;
      const int seg = map.tile_seg[tile];
      if (seg < 0) continue;
      const int nvalid = map.tile_nvalid[tile];
      const int m0 = tile * 128;
      const int buf = it & 1, use = it >> 1;
      mbar_wait(&tmem_full[buf], (uint32_t)(use & 1), 14, p.park >= 1);
      tc_fence_after();
      const int row0 = m0 + q * 32;
      int nrows = nvalid - q * 32;
      nrows = nrows < 0 ? 0 : (nrows > 32 ? 32 : nrows);
      const int seg_first = (epi.seg_start != nullptr) ? epi.seg_start[seg] : -1;
#pragma unroll 1
      for (int sl = 0; sl < 2; ++sl) {
        const int c = half * 128 + sl * 64;
        const bool live = n0 + c < epi.N;          // warp-uniform
        if (live) {
          float v[64];
          tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * C::kBN + c), v);
          __syncwarp();                            // previous slab fully consumed
          if constexpr (EpiQuad<KIND, true>::kOk) {
#pragma unroll
            for (int j = 0; j < 64; j += 4)
              *reinterpret_cast<float4*>(stg + lane * kStgLd4 + j) =
                  make_float4(v[j] * p.acc_scale, v[j + 1] * p.acc_scale, v[j + 2] * p.acc_scale, v[j + 3] * p.acc_scale);
          } else {
#pragma unroll
            for (int j = 0; j < 64; j += 2)
              *reinterpret_cast<float2*>(stg + lane * kStgLd + j) = make_float2(v[j] * p.acc_scale, v[j + 1] * p.acc_scale);
          }
        }
        if (sl == 1) {                             // last TMEM read of this tile: hand the accumulator back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[buf]);
        }
        if (!live) continue;
        __syncwarp();
        if constexpr (EpiQuad<KIND, true>::kOk) {
          quad_slab<KIND, true, 8>(epi, stg, row0, nrows, n0 + c, tile, q, lane);
          continue;
        }
        EpiPair<KIND, true> ep;
        ep.init(epi, seg, n0 + c + 2 * lane);
        if constexpr (EpiPair<KIND, false>::kHasPre) {
#pragma unroll 1
          for (int rb = 0; rb < 32; rb += 16) {
            float2 pre[16];
#pragma unroll
            for (int rr = 0; rr < 16; ++rr)
              if (rb + rr < nrows) pre[rr] = ep.pre(epi, row0 + rb + rr);
            float2 av[16];
#pragma unroll
            for (int rr = 0; rr < 16; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
#pragma unroll
            for (int rr = 0; rr < 16; ++rr)
              if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, pre[rr]);
          }
        } else {
#pragma unroll 1
          for (int rb = 0; rb < nrows; rb += 8) {
            float2 av[8];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) av[rr] = *reinterpret_cast<const float2*>(stg + (rb + rr) * kStgLd + 2 * lane);
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              if (rb + rr < nrows) ep.apply(epi, row0 + rb + rr, seg, seg_first, av[rr].x, av[rr].y, make_float2(0.f, 0.f));
          }
        }
        ep.finish(epi, tile, q);
      }
      ++it;
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------ fused WaveNet layer (16-bit mode)
// One launch per WN layer of the reversed flow (flow.py:63-88) instead of two GEMMs:
//   acc1[128 x 256] = conv k5 (x16; 10 K chunks)                         -> TMEM columns [0, 256)
//   acts            = tanh(acc1_even + b + g) * sigmoid(acc1_odd + b + g) -> shared memory, fp16, laid out
//                     as the K-major 128B-swizzled A operand of the second contraction (never in HBM)
//   acc2[128 x NRS] = acts * Wrs^T (2 K chunks, weights streamed through the same stage ring while the
//                     gate epilogue runs)                                  -> TMEM columns [256, 256 + NRS)
//   x += res, skip (+)= skip                                               (EPI_RESSKIP, coalesced)
// The conv reads +-2 halo rows of neighbouring tiles, so the updated 16-bit twin of x goes to a second
// buffer (ping-pong per layer); the fp32 master is only touched by the tile that owns the rows.
// OCC = 2: two CTAs per SM, so that the gate / res-skip epilogues of one tile overlap the loads and MMAs of another
// (the phases of ONE tile are strictly sequential: ~5 us of loads + MMA, ~2 us gate, ~4 us read-modify-write of x / skip).
// What makes it fit: two stages instead of three (two CTAs = four in flight per SM), eight epilogue warps with two
// column slabs each (2 x 320 threads x <= 102 registers), the gated activations in the A slots of the two stages that
// only carry res/skip WEIGHTS at that point (all conv MMAs have completed), and the second accumulator over the
// columns of the first (the gate epilogue has read them): 98 KB of shared memory and 256 TMEM columns per CTA.
template <int OCC>
struct WnCfgT {
  static constexpr int kBK = 64, kUmmaK = 16;
  static constexpr int kATile = 128 * 128, kBTile = 256 * 128;            // bytes
  static constexpr int kStageBytes = kATile + kBTile;                     // 48 KB
  static constexpr int kStages = OCC == 2 ? 2 : 3;
  static constexpr int kEpiWarps = OCC == 2 ? 8 : 16;                     // 4 lane quarters x 2 or 4 slabs of 64 columns at a time
  static constexpr int kSlabStep = kEpiWarps / 4;
  static constexpr int kThreads = 64 + 32 * kEpiWarps;                    // 320 / 576
  static constexpr int kActsBytes = OCC == 2 ? 0 : 2 * kATile;            // gated activations: 2 K chunks
  static constexpr uint32_t kTmemCols = OCC == 2 ? 256u : 512u;
  static constexpr uint32_t kAcc2Col = OCC == 2 ? 0u : 256u;
  static constexpr int kSmem = kStages * kStageBytes + kActsBytes + 1024 /*gate constants*/ + 1024 /*align*/ + 256;
  static_assert(kStages * kStageBytes >= kEpiWarps * 32 * kStgLd4 * 4, "staging must fit into the stage ring");
};

struct WnParams {
  int col0;                    // first column of x16 (elements)
  float scale_in, scale_rs;    // accumulator scales (power-of-two weight pre-scaling undone)
  const float* bias_in;        // [256], (tanh, sigmoid) rows interleaved
  const float* segvec; int ldseg;   // style conditioning g: [n_seg, ldseg], tanh half then sigmoid half
  int park;                    // see TcParams::park
};

template <int NRS, int OCC>
__global__ void __launch_bounds__(WnCfgT<OCC>::kThreads, OCC)
wn_layer_fused_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapWin,
                      const __grid_constant__ CUtensorMap mapWrs, const WnParams p, const Epi epi, const RowMap map) {
  using C = WnCfgT<OCC>;
  constexpr int kInChunks = 10;      // 5 taps x 128 channels / 64
  constexpr int kRsChunks = 2;       // 128 gated channels / 64
  const int tile = blockIdx.x;
  const int seg = map.tile_seg[tile];
  if (seg < 0) return;
  const int nvalid = map.tile_nvalid[tile];
  const int m0 = tile * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* acts = smem + C::kStages * C::kStageBytes;
  float* gconst = reinterpret_cast<float*>(acts + C::kActsBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(acts + C::kActsBytes + 1024);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* accum1 = bars + 2 * C::kStages;
  uint64_t* acts_ready = accum1 + 1;
  uint64_t* accum2 = accum1 + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum1 + 3);

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum1, 1);
    mbar_init(acts_ready, C::kEpiWarps);
    mbar_init(accum2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  pdl_enter();   // everything above (barrier init, TMEM allocation) overlaps the previous kernel of the stream
  static_assert(C::kThreads >= 64 + 256, "gate constants: one thread per accumulator column");
  if (threadIdx.x >= 64 && threadIdx.x < 64 + 256) {   // gate constants per accumulator column: bias + style conditioning
    const int n = threadIdx.x - 64;
    gconst[n] = p.bias_in[n] + p.segvec[(size_t)seg * p.ldseg + ((n & 1) ? 128 : 0) + (n >> 1)];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  auto stage_ptr = [&](int s) { return smem + (size_t)s * C::kStageBytes; };
  // OCC 2: res/skip weight chunk ch arrives in the B slot of stage ch (10 conv chunks, 2 stages), its A slot is free
  static_assert(OCC != 2 || (kInChunks % C::kStages == 0 && kRsChunks <= C::kStages), "activation chunks alias the A slots");
  auto acts_chunk = [&](int ch) { return OCC == 2 ? stage_ptr(ch) : acts + (size_t)ch * C::kATile; };
  constexpr uint32_t kIdescIn = (1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  constexpr uint32_t kIdescRs = (1u << 4) | ((uint32_t)(NRS >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int tap = 0, cc = 0;
      for (int ch = 0; ch < kInChunks + kRsChunks; ++ch) {
        mbar_wait(&empty[stage], phase ^ 1, 21, p.park >= 2);
        uint8_t* sp = stage_ptr(stage);
        if (ch < kInChunks) {
          mbar_expect_tx(&full[stage], C::kStageBytes);
          tma_load_2d(sp, &mapX, &full[stage], p.col0 + cc * C::kBK, m0 + tap - 2);
          tma_load_2d(sp + C::kATile, &mapWin, &full[stage], ch * C::kBK, 0);
          if (++cc == 2) { cc = 0; ++tap; }
        } else {
          mbar_expect_tx(&full[stage], NRS * 128);
          tma_load_2d(sp + C::kATile, &mapWrs, &full[stage], (ch - kInChunks) * C::kBK, 0);
        }
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int ch = 0; ch < kInChunks; ++ch) {
        mbar_wait(&full[stage], phase, 22, p.park >= 2);
        tc_fence_after();
        const uint32_t a = smem_u32(stage_ptr(stage)), b = a + C::kATile;
#pragma unroll
        for (int k = 0; k < C::kBK / C::kUmmaK; ++k)
          tc_mma<false>(tmem_base, make_sdesc(a + k * 32), make_sdesc(b + k * 32), kIdescIn, (ch | k) != 0);
        tc_commit(&empty[stage]);
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      tc_commit(accum1);
      mbar_wait(acts_ready, 0, 23, p.park >= 2);
      tc_fence_after();
      for (int ch = 0; ch < kRsChunks; ++ch) {
        mbar_wait(&full[stage], phase, 24, p.park >= 2);
        tc_fence_after();
        const uint32_t a = smem_u32(acts_chunk(ch)), b = smem_u32(stage_ptr(stage)) + C::kATile;
#pragma unroll
        for (int k = 0; k < C::kBK / C::kUmmaK; ++k)
          tc_mma<false>(tmem_base + C::kAcc2Col, make_sdesc(a + k * 32), make_sdesc(b + k * 32), kIdescRs, (ch | k) != 0);
        tc_commit(&empty[stage]);
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      tc_commit(accum2);
    }
  } else {
    // ------------------------------------------------------------ gate epilogue -> shared-memory operand
    const int ew = warp - 2;
    const int q = warp & 3;            // TMEM lane quarter this warp may read (hardware rule)
    using EG = EpiPair<EPI_GATE, true>;
    mbar_wait(accum1, 0, 25, p.park >= 1);
    tc_fence_after();
#pragma unroll 1
    for (int slab = ew >> 2; slab < 4; slab += C::kSlabStep) {   // 64 accumulator columns = 32 gated channels
      float v[64];
      tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(slab * 64), v);
      const int r = q * 32 + lane;     // row of the tile = TMEM lane
      const float* gc = gconst + slab * 64;
      uint32_t w[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float y[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int j = 4 * i + 2 * h;
          const float a = fmaf(v[j], p.scale_in, gc[j]), b = fmaf(v[j + 1], p.scale_in, gc[j + 1]);
          y[h] = EG::th(a) * EG::dv(1.f, 1.f + EG::ex(-b));
        }
        const twin2_t t = to_twin2(y[0], y[1]);
        w[i] = *reinterpret_cast<const uint32_t*>(&t);
      }
      // K chunk slab / 2, 16-byte units (slab % 2) * 4 .. + 3 of this row, 128B swizzle: unit ^= row % 8
      uint8_t* rowp = acts_chunk(slab >> 1) + r * 128;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int unit = ((slab & 1) * 4 + u) ^ (r & 7);
        *reinterpret_cast<uint4*>(rowp + unit * 16) = make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(acts_ready);
    // ------------------------------------------------------------ res / skip epilogue
    mbar_wait(accum2, 0, 26, p.park >= 1);
    tc_fence_after();
    float* stg = reinterpret_cast<float*>(smem) + ew * (32 * kStgLd4);
    const int row0 = m0 + q * 32;
    int nrows = nvalid - q * 32;
    nrows = nrows < 0 ? 0 : (nrows > 32 ? 32 : nrows);
#pragma unroll 1
    for (int c = (ew >> 2) * 64; c < epi.N; c += C::kSlabStep * 64) {     // warp-uniform
      __syncwarp();                              // the staging block of the previous slab is consumed
      {
        float v[64];
        tmem_ld64(tmem_base + ((uint32_t)(q * 32) << 16) + C::kAcc2Col + (uint32_t)c, v);
#pragma unroll
        for (int j = 0; j < 64; j += 4)
          *reinterpret_cast<float4*>(stg + lane * kStgLd4 + j) =
              make_float4(v[j] * p.scale_rs, v[j + 1] * p.scale_rs, v[j + 2] * p.scale_rs, v[j + 3] * p.scale_rs);
      }
      __syncwarp();
      const float4 bias = *reinterpret_cast<const float4*>(epi.bias + c + (lane & 15) * 4);
      if (c < epi.split) {                       // x += res, 16-bit twin of the new x into the other buffer
        rmw_slab4<true, true>(stg, epi.out + (size_t)row0 * epi.ldo + epi.col0 + c, epi.ldo,
                              epi.outh + (size_t)row0 * epi.ldh + epi.colh0 + c, epi.ldh, bias, nrows, lane);
      } else {                                   // skip (+)= ; its twin only when the coupling projection follows
        const int m = c - epi.split;
        float* d = epi.out2 + (size_t)row0 * epi.ldo2 + epi.col20 + m;
        twin_t* dh = epi.outh2 != nullptr ? epi.outh2 + (size_t)row0 * epi.ldh2 + epi.colh20 + m : nullptr;
        if (epi.accumulate2) {
          if (dh != nullptr) rmw_slab4<true, true>(stg, d, epi.ldo2, dh, epi.ldh2, bias, nrows, lane);
          else rmw_slab4<true, false>(stg, d, epi.ldo2, nullptr, 0, bias, nrows, lane);
        } else {
          if (dh != nullptr) rmw_slab4<false, true>(stg, d, epi.ldo2, dh, epi.ldh2, bias, nrows, lane);
          else rmw_slab4<false, false>(stg, d, epi.ldo2, nullptr, 0, bias, nrows, lane);
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(C::kTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------ host: tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct MapKey {
  const void* base; int rows, ld, elem, box_rows;
  bool operator<(const MapKey& o) const {
    return std::tie(base, rows, ld, elem, box_rows) < std::tie(o.base, o.rows, o.ld, o.elem, o.box_rows);
  }
};
std::map<MapKey, CUtensorMap> g_map_cache;
std::mutex g_map_mutex;

// 2-D row-major [rows, ld] tensor of `elem`-byte elements, box = [128 bytes, box_rows], 128B swizzle.
bool get_map(const void* base, int rows, int ld, int elem, int box_rows, CUtensorMap* out) {
  MapKey key{base, rows, ld, elem, box_rows};
  std::lock_guard<std::mutex> lk(g_map_mutex);
  auto it = g_map_cache.find(key);
  if (it != g_map_cache.end()) { *out = it->second; return true; }
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * elem};
  cuuint32_t box[2] = {(cuuint32_t)(128 / elem), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  if (g_map_cache.size() > 4096) g_map_cache.clear();
  g_map_cache[key] = m;
  *out = m;
  return true;
}

int mbar_park_mode();
bool m256_fills_gpu(const ConvGemm& g, const RowMap& map);

template <int MODE, int BN, int KIND, int OCC = 1, int SK = 1>
cudaError_t launch_one(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  using C = Cfg<MODE, BN, OCC>;
  constexpr int elem = C::kElem;
  if constexpr (KIND == EPI_LN) {
    // whole rows in one column tile; float4 residual loads and half2 twin stores; staging + lnbuf inside the stage ring
    static_assert(C::kStages * C::kStageBytes >= C::kEpiWarps * 32 * kStgLd4 * 4 + 2 * 4 * 128 * 4, "EPI_LN staging");
    const Epi& e = g.epi;
    if (e.N > BN || (e.N & 3) || e.ln_g == nullptr || (e.ln_seg_stride == 0 && e.ln_b == nullptr) || (e.ldo & 1) ||
        (e.col0 & 1) || (e.res != nullptr && ((e.ldr & 3) || (e.rcol0 & 3))) ||
        (e.outg != nullptr && ((e.ldh & 1) || (e.colh0 & 1))))
      return cudaErrorInvalidValue;
  }
  CUtensorMap mA[2], mAlo[2], mW, mWlo;
  TcParams p{};
  p.n_kseg = g.n_kseg;
  p.park = mbar_park_mode();
  p.n_chunks = 0;
  for (int s = 0; s < 2; ++s) {
    const KSeg& ks = g.seg[s < g.n_kseg ? s : 0];
    const bool f16 = (MODE == MODE_F16X3) || (MODE == MODE_F16 && g.f16);
    const void* base = (MODE == MODE_TF32X3) ? (const void*)ks.A : (f16 ? (const void*)ks.Ag : (const void*)ks.Ah);
    const int ld = (MODE == MODE_TF32X3) ? ks.lda : (f16 ? ks.ldag : ks.ldah);
    if (base == nullptr) return cudaErrorInvalidValue;
    if (!get_map(base, ks.rows, ld, elem, 128, &mA[s])) return cudaErrorNotSupported;
    mAlo[s] = mA[s];
    if (MODE == MODE_F16X3) {
      if (ks.Al == nullptr) return cudaErrorInvalidValue;
      if (!get_map(ks.Al, ks.rows, ld, elem, 128, &mAlo[s])) return cudaErrorNotSupported;
    }
    if (s < g.n_kseg) {
      p.taps[s] = ks.taps;
      p.pad[s] = ks.pad;
      p.dil[s] = ks.dil > 0 ? ks.dil : 1;
      p.cin_chunks[s] = ks.cin_pad / C::kBK;
      p.col0[s] = (MODE == MODE_TF32X3) ? ks.col0 : ks.colh0;
      p.n_chunks += ks.taps * p.cin_chunks[s];
    }
  }
  p.acc_scale = 1.f;
  if (MODE == MODE_F16) {
    const bool per_seg = !g.f16 && g.Wh_seg != nullptr;
    const void* w = per_seg ? (const void*)g.Wh_seg : (const void*)g.Wg;
    if (w == nullptr) return cudaErrorInvalidValue;
    if (!get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, BN, &mW)) return cudaErrorNotSupported;
    p.w_seg_rows = per_seg ? g.Npad : 0;
    mWlo = mW;
    p.ab_fmt = 0u;                                     // fp16 operands
    p.acc_scale = per_seg ? 1.f : g.acc_scale;         // per-utterance weights are stored unscaled
  } else if (MODE == MODE_F16X3) {
    if (g.Wg == nullptr || g.Wl == nullptr) return cudaErrorInvalidValue;
    if (!get_map(g.Wg, g.Npad, g.Ktot, 2, BN, &mW)) return cudaErrorNotSupported;
    if (!get_map(g.Wl, g.Npad, g.Ktot, 2, BN, &mWlo)) return cudaErrorNotSupported;
    p.ab_fmt = 0u;
    p.acc_scale = g.acc_scale;
  } else {
    p.ab_fmt = 2u;
    // W holds hi (tf32-rounded) in rows [0, Npad) and lo in rows [Npad, 2*Npad)
    if (!get_map(g.W, g.Npad, g.Ktot, 4, BN, &mW)) return cudaErrorNotSupported;
    if (!get_map(g.W + (size_t)g.Npad * g.Ktot, g.Npad, g.Ktot, 4, BN, &mWlo)) return cudaErrorNotSupported;
  }
  // 16-row boxes of the same weights for a tail column tile (unused unless N leaves <= 16 columns in the last tile)
  CUtensorMap mWt = mW, mWtlo = mWlo;
  if (MODE != MODE_TF32X3 && g.epi.N % BN != 0 && g.epi.N % BN <= 16) {
    const bool per_seg = MODE == MODE_F16 && !g.f16 && g.Wh_seg != nullptr;
    const void* w = per_seg ? (const void*)g.Wh_seg : (const void*)g.Wg;
    if (!get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, 16, &mWt)) return cudaErrorNotSupported;
    mWtlo = mWt;
    if (MODE == MODE_F16X3 && !get_map(g.Wl, g.Npad, g.Ktot, 2, 16, &mWtlo)) return cudaErrorNotSupported;
  }
  auto kern = conv_gemm_tc_kernel<MODE, BN, KIND, OCC, SK>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  if constexpr (SK > 1) {
    static_assert(C::kStages * C::kStageBytes >= (SK * 32 * kSkPitch + 2 * 32 * kStgLd4 + 2 * 4 * 128) * 4, "split-K staging");
    if (p.n_chunks < 2 * SK || g.epi.N > BN) return cudaErrorInvalidValue;
    dim3 grid(SK, map.n_tiles);
    launch_k_cluster(kern, grid, C::kThreads, C::kSmem, st, SK, mA[0], mA[1], mAlo[0], mAlo[1], mW, mWlo, mWt, mWtlo, p, g.epi,
                     map);
    return cudaGetLastError();
  }
  dim3 grid((g.epi.N + BN - 1) / BN, map.n_tiles);
  launch_k(kern, grid, C::kThreads, C::kSmem, st, mA[0], mA[1], mAlo[0], mAlo[1], mW, mWlo, mWt, mWtlo, p, g.epi, map);
  return cudaGetLastError();
}

template <int KIND>
cudaError_t launch_persistent(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  using C = PCfg;
  CUtensorMap mA[2], mW;
  TcParams p{};
  p.n_kseg = g.n_kseg;
  p.park = mbar_park_mode();
  p.n_chunks = 0;
  for (int s = 0; s < 2; ++s) {
    const KSeg& ks = g.seg[s < g.n_kseg ? s : 0];
    const void* base = g.f16 ? (const void*)ks.Ag : (const void*)ks.Ah;
    const int ld = g.f16 ? ks.ldag : ks.ldah;
    if (base == nullptr) return cudaErrorInvalidValue;
    if (!get_map(base, ks.rows, ld, 2, 128, &mA[s])) return cudaErrorNotSupported;
    if (s < g.n_kseg) {
      p.taps[s] = ks.taps;
      p.pad[s] = ks.pad;
      p.dil[s] = ks.dil > 0 ? ks.dil : 1;
      p.cin_chunks[s] = ks.cin_pad / C::kBK;
      p.col0[s] = ks.colh0;
      p.n_chunks += ks.taps * p.cin_chunks[s];
    }
  }
  const bool per_seg = !g.f16 && g.Wh_seg != nullptr;
  const void* w = per_seg ? (const void*)g.Wh_seg : (const void*)g.Wg;
  if (w == nullptr) return cudaErrorInvalidValue;
  if (!get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, C::kBN, &mW)) return cudaErrorNotSupported;
  p.w_seg_rows = per_seg ? g.Npad : 0;
  p.ab_fmt = 0u;
  p.acc_scale = per_seg ? 1.f : g.acc_scale;
  CUtensorMap mWt = mW;
  if (g.epi.N % C::kBN != 0 && g.epi.N % C::kBN <= 16 &&
      !get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, 16, &mWt))
    return cudaErrorNotSupported;
  auto kern = conv_gemm_tc_persistent_kernel<KIND>;
  static bool attr_done = false;
  static int n_sm = 0;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem);
    if (e != cudaSuccess) return e;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    attr_done = true;
  }
  const int n_tiles_n = (g.epi.N + C::kBN - 1) / C::kBN;
  const int total = map.n_tiles * n_tiles_n;
  const int grid = total < n_sm ? total : n_sm;
  launch_k(kern, grid, C::kThreads, C::kSmem, st, mA[0], mA[1], mW, mWt, p, g.epi, map, n_tiles_n);
  return cudaGetLastError();
}

template <int KIND>
cudaError_t launch_m256(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  using C = M2Cfg;
  CUtensorMap mA[2], mW;
  TcParams p{};
  p.n_kseg = g.n_kseg;
  p.park = mbar_park_mode();
  p.n_chunks = 0;
  for (int s = 0; s < 2; ++s) {
    const KSeg& ks = g.seg[s < g.n_kseg ? s : 0];
    const void* base = g.f16 ? (const void*)ks.Ag : (const void*)ks.Ah;
    const int ld = g.f16 ? ks.ldag : ks.ldah;
    if (base == nullptr) return cudaErrorInvalidValue;
    if (!get_map(base, ks.rows, ld, 2, 128, &mA[s])) return cudaErrorNotSupported;
    if (s < g.n_kseg) {
      p.taps[s] = ks.taps;
      p.pad[s] = ks.pad;
      p.dil[s] = ks.dil > 0 ? ks.dil : 1;
      p.cin_chunks[s] = ks.cin_pad / C::kBK;
      p.col0[s] = ks.colh0;
      p.n_chunks += ks.taps * p.cin_chunks[s];
    }
  }
  const bool per_seg = !g.f16 && g.Wh_seg != nullptr;
  const void* w = per_seg ? (const void*)g.Wh_seg : (const void*)g.Wg;
  if (w == nullptr) return cudaErrorInvalidValue;
  if (!get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, C::kBN, &mW)) return cudaErrorNotSupported;
  p.w_seg_rows = per_seg ? g.Npad : 0;
  p.ab_fmt = 0u;
  p.acc_scale = per_seg ? 1.f : g.acc_scale;
  CUtensorMap mWt = mW;
  if (g.epi.N % C::kBN != 0 && g.epi.N % C::kBN <= 16 &&
      !get_map(w, per_seg ? g.Npad * map.n_seg : g.Npad, g.Ktot, 2, 16, &mWt))
    return cudaErrorNotSupported;
  auto kern = conv_gemm_tc_m256_kernel<KIND>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  if (per_seg) return cudaErrorInvalidValue;     // the two row tiles of a CTA share one weight stage
  const int n_tiles_n = (g.epi.N + C::kBN - 1) / C::kBN;
  dim3 grid(n_tiles_n, (map.n_tiles + 1) / 2);
  launch_k(kern, grid, C::kThreads, C::kSmem, st, mA[0], mA[1], mW, mWt, p, g.epi, map);
  return cudaGetLastError();
}

// STTS_PERSISTENT=0 falls back to one CTA per tile for the wide one-pass GEMMs
bool use_persistent() {
  static const bool v = [] { const char* e = std::getenv("STTS_PERSISTENT"); return e == nullptr || e[0] != '0'; }();
  return v;
}

// the 256-row tile halves the CTA count: only when the tile pairs still fill the machine (a single 5 s utterance has
// 13 row tiles; latency mode keeps the 128-row kernels)
bool m256_fills_gpu(const ConvGemm& g, const RowMap& map) {
  return (long long)((map.n_tiles + 1) / 2) * ((g.epi.N + 255) / 256) >= 148;
}

int mbar_park_mode() {
  static const int v = [] { const char* e = std::getenv("STTS_MBAR_PARK"); return e != nullptr ? std::atoi(e) : 1; }();
  return v;
}

// STTS_OCC2=<n>: 16-bit-mode GEMMs with at most n K chunks (64 columns each) run as BN = 128, two CTAs
// per SM (-1 = never; default 8: pwconv1, prior, the 1x1 flow / decoder projections).
int occ2_max_chunks() {
  static const int v = [] {
    const char* e = std::getenv("STTS_OCC2");
    return e != nullptr ? std::atoi(e) : 8;   // measured with the float4 epilogue: pwconv1 -5 %, flow proj -24 %, asr_res -21 %
  }();
  return v;
}

template <int MODE, int KIND>
cudaError_t launch_bn(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  if constexpr (KIND == EPI_LN) {            // one column tile must hold the whole row
    if constexpr (MODE != MODE_F16) {
      // long-K LayerNorm GEMMs (text-encoder ffn2 / conv prelude): split-K over a cluster of four CTAs.
      // STTS_LN_SPLITK=<min K>: smallest reduction length (elements) that takes it, 0 = never; default 512
      static const int min_k = [] { const char* e = std::getenv("STTS_LN_SPLITK"); return e != nullptr ? std::atoi(e) : 512; }();
      int ktot = 0;
      for (int s = 0; s < g.n_kseg; ++s) ktot += g.seg[s].taps * g.seg[s].cin_pad;
      // only while the four-fold grid still fits the SMs in one wave (a 32-utterance batch has 47 token tiles: 188 CTAs
      // of 193 KB shared memory = two waves, measured 30 -> 42 us per launch); a small-batch shape (summation order)
      if (min_k > 0 && ktot >= min_k && g.epi.N <= 128 && small_shapes_enabled() && map.n_tiles * 4 <= 148)
        return launch_one<MODE, 128, KIND, 1, 4>(g, map, st);
    }
    if (g.epi.N <= 128) return launch_one<MODE, 128, KIND>(g, map, st);
    if (g.epi.N <= 256 && g.Npad % 256 == 0) return launch_one<MODE, 256, KIND>(g, map, st);
    return cudaErrorInvalidValue;
  }
  if constexpr (MODE == MODE_F16) {
    int chunks = 0;
    for (int s = 0; s < g.n_kseg; ++s) chunks += g.seg[s].taps * (g.seg[s].cin_pad / 64);
    // very wide short-K layers (pwconv1: K 512, N 1 536): STTS_M256_WIDE=<min k-chunks> sends them to the 256 x 256 kernel
    static const int m256_wide = [] { const char* e = std::getenv("STTS_M256_WIDE"); return e != nullptr ? std::atoi(e) : 8; }();
    if (m256_wide >= 0 && chunks >= m256_wide && g.epi.N >= 1024 && g.Npad % 256 == 0 && (g.f16 || g.Wh_seg == nullptr) &&
        m256_fills_gpu(g, map))
      return launch_m256<KIND>(g, map, st);
    if (chunks <= occ2_max_chunks()) return launch_one<MODE, 128, KIND, 2>(g, map, st);
  }
  // BN = 256 whenever there is more than one 128-column tile (long-K one-pass GEMMs reach the tensor peak
  // only with the 128 x 256 tile: measured 64 % vs 36 % pipe-active); the split modes (three operand
  // tiles per stage) use it for the wide STFT only.
  static const bool split256 = [] { const char* e = std::getenv("STTS_SPLIT_BN256"); return e == nullptr || e[0] != '0'; }();
  const bool wide = g.epi.N > 128 && g.Npad % 256 == 0;
  if constexpr (MODE == MODE_F16) {
    // Small batches (one 5 s utterance = 13 row tiles): with 256-column tiles a long-K conv is 26 CTAs of ~13 us of
    // tensor work each on 148 SMs.  128-column tiles double the CTAs and halve the per-CTA main loop.
    // STTS_SMALL_GRID=<n>: grids of fewer than n 256-column CTAs take the 128-column kernel (0 = never)
    static const int small_grid = [] { const char* e = std::getenv("STTS_SMALL_GRID"); return e != nullptr ? std::atoi(e) : 74; }();
    if (wide && small_shapes_enabled() && (long long)map.n_tiles * ((g.epi.N + 255) / 256) < small_grid)
      return launch_one<MODE, 128, KIND>(g, map, st);
  }
  if constexpr (MODE == MODE_F16) {
    // persistent variant: measured faster for long-K, multi-column-tile layers (decoder convs -4 .. -17 %,
    // pwconv2 -7 %, inverse DFT -11 %, output heads -5 %) and slower for the short-K / epilogue-bound ones
    // (pwconv1 +54 %, flow res/skip +35 %: 8 instead of 16 epilogue warps), which keep one CTA per tile
    int chunks = 0;
    for (int s = 0; s < g.n_kseg; ++s) chunks += g.seg[s].taps * (g.seg[s].cin_pad / 64);
    // 256 x 256 tiles (two row tiles per weight stage) for the long-K wide GEMMs without per-utterance weights: they are
    // bound by L2 -> SM operand traffic, which this tile shape cuts by a third.  STTS_M256=<min k-chunks>, -1 = off.
    static const int m256_min = [] { const char* e = std::getenv("STTS_M256"); return e != nullptr ? std::atoi(e) : 16; }();
    // Measured per layer against the persistent kernel (same box): output heads -2 .. -3.5 %, inverse DFT -2 %, the
    // 1x1 projector (16 chunks, not persistent before) -20 %; prior convs (N = 512, 119 chunks) +12 %, decoder convs
    // +-1 %: so it takes the very wide outputs and the wide layers the persistent kernel does not.
    const bool pers = use_persistent() && chunks >= 24 && g.epi.N >= 512;
    if (wide && m256_min >= 0 && chunks >= m256_min && g.epi.N >= 256 && (g.f16 || g.Wh_seg == nullptr) && m256_fills_gpu(g, map) &&
        (g.epi.N >= 1024 || !pers))
      return launch_m256<KIND>(g, map, st);
    static const bool all = [] { const char* e = std::getenv("STTS_PERSISTENT"); return e != nullptr && e[0] == 'a'; }();
    if (wide && use_persistent() && (all || (chunks >= 24 && g.epi.N >= 512))) return launch_persistent<KIND>(g, map, st);
  }
  if (MODE == MODE_F16 && wide) return launch_one<MODE, 256, KIND>(g, map, st);
  if (MODE != MODE_F16 && wide && split256 && g.epi.N >= 1024) return launch_one<MODE, 256, KIND>(g, map, st);
  if constexpr (MODE != MODE_F16) {
    // Wave quantisation: the split-mode kernels are one CTA per SM (193 KB of shared memory).  A 32-utterance batch has
    // 47 token tiles: the text encoders' ffn1 (N = 512) is 188 CTAs with 128-column tiles -- two waves, the second a
    // quarter full -- and 94 with 256-column tiles: one wave.  STTS_SPLIT_WAVE=0 switches the rule off.
    static const bool wave = [] { const char* e = std::getenv("STTS_SPLIT_WAVE"); return e == nullptr || e[0] != '0'; }();
    const long long c128 = (long long)map.n_tiles * ((g.epi.N + 127) / 128), c256 = (long long)map.n_tiles * ((g.epi.N + 255) / 256);
    if (wave && wide && split256 && c128 > 148 && c256 <= 148) return launch_one<MODE, 256, KIND>(g, map, st);
  }
  return launch_one<MODE, 128, KIND>(g, map, st);
}

template <int MODE>
cudaError_t launch_mode(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  switch (g.epi.kind) {
    case EPI_BIAS: return launch_bn<MODE, EPI_BIAS>(g, map, st);
    case EPI_RES_SCALE: return launch_bn<MODE, EPI_RES_SCALE>(g, map, st);
    case EPI_SILU: return launch_bn<MODE, EPI_SILU>(g, map, st);
    case EPI_ADD: return launch_bn<MODE, EPI_ADD>(g, map, st);
    case EPI_GATE: return launch_bn<MODE, EPI_GATE>(g, map, st);
    case EPI_RESSKIP: return launch_bn<MODE, EPI_RESSKIP>(g, map, st);
    case EPI_COUPLE: return launch_bn<MODE, EPI_COUPLE>(g, map, st);
    case EPI_PRIOR: return launch_bn<MODE, EPI_PRIOR>(g, map, st);
    case EPI_STFT: return launch_bn<MODE, EPI_STFT>(g, map, st);
    case EPI_RELU: return launch_bn<MODE, EPI_RELU>(g, map, st);
    case EPI_GELU: return launch_bn<MODE, EPI_GELU>(g, map, st);
    case EPI_LN:
      if constexpr (MODE == MODE_F16) return cudaErrorInvalidValue;   // fp32-grade modes only (token-rate islands)
      else return launch_bn<MODE, EPI_LN>(g, map, st);
    default: return cudaErrorInvalidValue;
  }
}


template <int NRS, int OCC>
cudaError_t launch_wn(const ConvGemm& gin, const ConvGemm& grs, const RowMap& map, cudaStream_t st) {
  using C = WnCfgT<OCC>;
  const KSeg& ks = gin.seg[0];
  CUtensorMap mX, mWin, mWrs;
  if (ks.Ah == nullptr || gin.Wg == nullptr || grs.Wg == nullptr) return cudaErrorInvalidValue;
  if (!get_map(ks.Ah, ks.rows, ks.ldah, 2, 128, &mX)) return cudaErrorNotSupported;
  if (!get_map(gin.Wg, gin.Npad, gin.Ktot, 2, 256, &mWin)) return cudaErrorNotSupported;
  if (!get_map(grs.Wg, grs.Npad, grs.Ktot, 2, NRS, &mWrs)) return cudaErrorNotSupported;
  WnParams p;
  p.park = mbar_park_mode();
  p.col0 = ks.colh0;
  p.scale_in = gin.acc_scale;
  p.scale_rs = grs.acc_scale;
  p.bias_in = gin.epi.bias;
  p.segvec = gin.epi.segvec;
  p.ldseg = gin.epi.ldseg;
  auto kern = wn_layer_fused_kernel<NRS, OCC>;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  launch_k(kern, map.n_tiles, C::kThreads, C::kSmem, st, mX, mWin, mWrs, p, grs.epi, map);
  return cudaGetLastError();
}
}  // namespace

void tc_map_cache_clear() {
  std::lock_guard<std::mutex> lk(g_map_mutex);
  g_map_cache.clear();
}

cudaError_t conv_gemm_tc_f16(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  return launch_mode<MODE_F16>(g, map, st);
}
cudaError_t conv_gemm_tc_tf32x3(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  return launch_mode<MODE_TF32X3>(g, map, st);
}
cudaError_t conv_gemm_tc_f16x3(const ConvGemm& g, const RowMap& map, cudaStream_t st) {
  return launch_mode<MODE_F16X3>(g, map, st);
}


// Fused WN layer (16-bit mode): gin = the k5 in_layer (EPI_GATE constants in gin.epi), grs = the 1x1
// res/skip layer (EPI_RESSKIP epilogue in grs.epi).  Shapes are those of the default model (128
// channels, k5, 256 / 128 res-skip outputs); anything else is rejected.
cudaError_t wn_layer_fused_f16(const ConvGemm& gin, const ConvGemm& grs, const RowMap& map, cudaStream_t st) {
  const KSeg& k = gin.seg[0];
  if (gin.n_kseg != 1 || grs.n_kseg != 1 || k.taps != 5 || k.pad != 2 || k.cin_pad != 128 || gin.epi.N != 256 ||
      gin.Ktot != 640 || grs.Ktot != 128 || grs.seg[0].taps != 1 || grs.seg[0].cin_pad != 128)
    return cudaErrorInvalidValue;
  // STTS_WN_OCC=1: the one-CTA-per-SM variant (three stages, 16 epilogue warps) for A/B runs
  // (also taken when the tiles do not fill the SMs once: nothing to overlap with, and its 16 epilogue warps finish a
  // tile sooner -- one 5 s utterance 2.56 vs 2.65 ms; same arithmetic, bit-identical results)
  static const bool occ2_env = [] { const char* e = std::getenv("STTS_WN_OCC"); return e == nullptr || e[0] != '1'; }();
  const bool occ2 = occ2_env && map.n_tiles > 148;
  if (grs.epi.N == 256 && grs.Npad >= 256) return occ2 ? launch_wn<256, 2>(gin, grs, map, st) : launch_wn<256, 1>(gin, grs, map, st);
  if (grs.epi.N == 128) return occ2 ? launch_wn<128, 2>(gin, grs, map, st) : launch_wn<128, 1>(gin, grs, map, st);
  return cudaErrorInvalidValue;
}

}  // namespace stts
