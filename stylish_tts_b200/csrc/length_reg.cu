// Length regulation: duration logits -> integer durations -> frame -> token index.
//   DurationProcessor.prediction_to_duration   train/utils.py:468-474
//   DurationProcessor.duration_to_alignment    train/utils.py:476-489
// The reference materialises a dense one-hot [T, F] matrix and multiplies with it; here the
// alignment is the int32 index vector (its argmax over tokens), which is what every consumer
// needs.  Integer work: bit-exact by construction once the durations agree.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace stts {

__constant__ float kClassToDur[16] = {1, 2, 3, 4, 5, 6, 7, 9, 12, 15, 18, 22, 27, 32, 38, 46};  // utils.py:391-393

__global__ void __launch_bounds__(256) durations_kernel(const float* __restrict__ logits, int n, int* __restrict__ dur) {
  pdl_enter();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  float x[16];
  const float4* p = reinterpret_cast<const float4*>(logits + (size_t)t * 16);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 v = p[i];
    x[4 * i] = v.x; x[4 * i + 1] = v.y; x[4 * i + 2] = v.z; x[4 * i + 3] = v.w;
  }
  float m = x[0];
  int am = 0;
#pragma unroll
  for (int i = 1; i < 16; ++i) if (x[i] > m) { m = x[i]; am = i; }   // first maximum, like torch.argmax
  // fp32 softmax in ATen's CPU formulation: exp(x - max), sum, multiply by the reciprocal of the sum; then
  // p * table rounded, then added -- every product / sum rounded separately (no FMA contraction: the adversarial
  // tie rows of tests/golden/durations_adv.safetensors tell the formulations apart).  exp is evaluated in fp64 and rounded once:
  // the correctly rounded fp32 value, which is what a 1-ulp libm (the reference's CPU expf) returns for all but a
  // few percent of arguments -- the hardware ex2.approx path of expf() is 2 ulp.  What remains implementation-
  // defined is the last ulp of `soft` next to an x.5 tie (~2 rows per million, tests/test_durations_gpu.py).
  float e[16], s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { e[i] = (float)exp((double)(x[i] - m)); s = __fadd_rn(s, e[i]); }
  const float inv = __fdiv_rn(1.f, s);
  float soft = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) soft = __fadd_rn(soft, __fmul_rn(__fmul_rn(e[i], inv), kClassToDur[i]));
  soft = fmaxf(rintf(soft), 1.f);                 // torch.round = half to even
  const float arg = kClassToDur[am];
  dur[t] = (int)((arg < 7.f) ? arg : soft);
}

cudaError_t durations_from_logits(const float* logits, int n_tokens, int* dur, cudaStream_t st) {
  if (n_tokens <= 0) return cudaSuccess;
  launch_k(durations_kernel, (n_tokens + 255) / 256, 256, 0, st, logits, n_tokens, dur);
  return cudaGetLastError();
}

// cu_frames[b] = sum of the durations of utterances < b (exclusive scan over utterances)
__global__ void __launch_bounds__(1024) seg_totals_kernel(const int* __restrict__ dur, const int* __restrict__ cu_tok,
                                                          int n_seg, int* __restrict__ cu_frames) {
  pdl_enter();
  __shared__ int tot[1024];
  const int b = threadIdx.x;
  int s = 0;
  if (b < n_seg)
    for (int t = cu_tok[b]; t < cu_tok[b + 1]; ++t) s += dur[t];
  tot[b] = s;
  __syncthreads();
  if (b == 0) {
    int run = 0;
    for (int i = 0; i < n_seg; ++i) { const int v = tot[i]; cu_frames[i] = run; run += v; }
    cu_frames[n_seg] = run;
  }
}

// idx[cu_frames[b] + f] = token (local to the utterance) that owns frame f
__global__ void __launch_bounds__(512) frame_index_kernel(const int* __restrict__ dur, const int* __restrict__ cu_tok,
                                                          const int* __restrict__ cu_frames, int* __restrict__ idx,
                                                          int idx_cap, int* __restrict__ status) {
  pdl_enter();
  __shared__ int sc[512];
  __shared__ int carry;
  const int b = blockIdx.x;
  const int t0 = cu_tok[b], nt = cu_tok[b + 1] - t0;
  const int f0 = cu_frames[b];
  if (threadIdx.x == 0) {
    carry = 0;
    if (b == gridDim.x - 1 && cu_frames[gridDim.x] > idx_cap) atomicExch(status, 1);
  }
  __syncthreads();
  for (int base = 0; base < nt; base += 512) {
    const int t = base + threadIdx.x;
    const int d = (t < nt) ? dur[t0 + t] : 0;
    sc[threadIdx.x] = d;
    __syncthreads();
    for (int o = 1; o < 512; o <<= 1) {       // Hillis-Steele inclusive scan
      const int v = (threadIdx.x >= o) ? sc[threadIdx.x - o] : 0;
      __syncthreads();
      sc[threadIdx.x] += v;
      __syncthreads();
    }
    const int start = carry + sc[threadIdx.x] - d;
    for (int k = 0; k < d; ++k)
      if (f0 + start + k < idx_cap) idx[f0 + start + k] = t;
    __syncthreads();
    if (threadIdx.x == 511) carry += sc[511];
    __syncthreads();
  }
}

cudaError_t frame_index(const int* dur, const int* cu_tok, int n_seg, int* cu_frames, int* idx, int idx_cap,
                        int* status, cudaStream_t st) {
  if (n_seg <= 0 || n_seg > 1024) return cudaErrorInvalidValue;
  launch_k(seg_totals_kernel, 1, 1024, 0, st, dur, cu_tok, n_seg, cu_frames);
  launch_k(frame_index_kernel, n_seg, 512, 0, st, dur, cu_tok, cu_frames, idx, idx_cap, status);
  return cudaGetLastError();
}

// ---- allocation-free variant of frame_index (stts_frame_index_static) -------------------------------------------
// frame_index needs one piece of scratch, the exclusive scan of per-utterance frame totals (cu_frames).  Callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) get it from this
// file-scope array instead.  Every device that loads the library holds kStaticScratchBytes until exit.  The array is
// shared by every call on a device: calls that overlap on the GPU (different streams) race on it.  Nothing in it has
// to start at zero: seg_totals_kernel overwrites entries 0 .. n_seg before frame_index_kernel reads them (same
// stream), so the previous call's contents are never observed.
constexpr int kStaticMaxSeg = 1024;                                      // utterances per call (= frame_index's own limit)
constexpr size_t kStaticScratchBytes = 4352;                             // (1024 + 1) int32, rounded up to 256 B
static __device__ __align__(256) unsigned char g_frame_index_scratch[kStaticScratchBytes];
static_assert((kStaticMaxSeg + 1) * sizeof(int) <= kStaticScratchBytes, "scratch too small");

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
int* g_scratch_ptr[kMaxDevices] = {};   // device address of g_frame_index_scratch per device (the symbol's name is a
                                        // host-side stand-in, not a device address)
}  // namespace

// once per device, with that device current
cudaError_t frame_index_static_setup(int device, int** out) {
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_ptr[device] == nullptr) {
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e != cudaSuccess) return e;
    if (cur != device) return cudaErrorInvalidDevice;
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_frame_index_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_ptr[device] = static_cast<int*>(p);
  }
  *out = g_scratch_ptr[device];
  return cudaSuccess;
}

int frame_index_static_max_seg() { return kStaticMaxSeg; }

cudaError_t frame_index_static(int device, const int* dur, const int* cu_tok, int n_seg, int* idx, int idx_cap,
                               int* status, cudaStream_t st) {
  if (n_seg <= 0 || n_seg > kStaticMaxSeg) return cudaErrorInvalidValue;
  int* cu_frames = nullptr;
  const cudaError_t e = frame_index_static_setup(device, &cu_frames);
  if (e != cudaSuccess) return e;
  launch_k(seg_totals_kernel, 1, 1024, 0, st, dur, cu_tok, n_seg, cu_frames);
  launch_k(frame_index_kernel, n_seg, 512, 0, st, dur, cu_tok, cu_frames, idx, idx_cap, status);
  return cudaGetLastError();
}

}  // namespace stts
