// Stand-alone scale + mask + softmax over attention scores, forward and backward (sm_100a).
//
// replaces: Megatron's `scaled_upper_triang_masked_softmax_cuda`, `scaled_masked_softmax_cuda` and `scaled_softmax_cuda`
// (megatron/fused_kernels) for the UNFUSED attention path (scores -> softmax -> context, selected with
// --no-masked-softmax-fusion off / without flash attention).  The flagship path never materialises the scores: the
// softmax lives inside the tcgen05 flash-attention kernels (attention.cu); this kernel exists so the reference's
// unfused path has a native equivalent and so attention probabilities can be inspected.
//
// One CTA of 128 threads per score row [sk]: the row is read once (coalesced) into shared memory as fp32
// (sk <= 16384), reduced twice (max, sum of exponentials; warp shuffles + one smem exchange) and written once.
//   y[j] = softmax_j( scale * x[j]  with  x[j] := -10000 where mask[b, 0, i, j] != 0,  and  j > i + (sk - sq) excluded )
// exactly the conventions of the reference kernels (masked positions are filled with -10000 AFTER scaling; the causal
// variant gives future positions probability 0).  Backward: dx = scale * y * (dy - sum_j dy[j] y[j]).
#ifdef B2_EMU                      // CPU emulation of the CUDA execution model (tests/emu/cuda_emu.h): kernels only
#include "cuda_emu.h"
namespace b2 {}
#else
#include "common.cuh"
#endif
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int kThreads = 128;
constexpr int kMaxSk = 16384;

template <typename T> B2_DEVICE float to_f(T v);
template <> B2_DEVICE float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> B2_DEVICE float to_f<__half>(__half v) { return __half2float(v); }
template <> B2_DEVICE float to_f<float>(float v) { return v; }
template <typename T> B2_DEVICE T from_f(float v);
template <> B2_DEVICE __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> B2_DEVICE __half from_f<__half>(float v) { return __float2half_rn(v); }
template <> B2_DEVICE float from_f<float>(float v) { return v; }

B2_DEVICE float block_reduce(float v, bool is_max, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float other = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, other) : v + other;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();                       // red[] may still be read by the previous reduction
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
#pragma unroll
  for (int w = 1; w < kThreads / 32; ++w) r = is_max ? fmaxf(r, red[w]) : r + red[w];
  return r;
}

// rows = batches * heads * sq.  mask: uint8 [mask_batches, 1, sq, sk] (mask_batches == batches or 1) or NULL.
template <typename T>
__global__ void __launch_bounds__(kThreads) softmax_fwd_kernel(const T* __restrict__ x, const uint8_t* __restrict__ mask,
                                                               T* __restrict__ y, long long rows, int sq, int sk, int heads,
                                                               int mask_batches, float scale, int causal) {
  B2_DYN_SMEM(float, row);               // [sk]
  __shared__ float red[kThreads / 32];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const int i = static_cast<int>(r % sq);
    const long long b = r / (static_cast<long long>(sq) * heads);
    const T* xr = x + r * sk;
    const uint8_t* mr = mask ? mask + ((mask_batches == 1 ? 0 : b) * sq + i) * static_cast<long long>(sk) : nullptr;
    const int limit = causal ? min(sk, i + 1 + (sk - sq)) : sk;      // keys j < limit are visible
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < sk; j += kThreads) {
      float v = -INFINITY;
      if (j < limit) {
        v = to_f<T>(xr[j]) * scale;
        if (mr && mr[j]) v = -10000.f;
      }
      row[j] = v;
      mx = fmaxf(mx, v);
    }
    mx = block_reduce(mx, true, red);
    float sum = 0.f;
    for (int j = threadIdx.x; j < sk; j += kThreads) {
      const float e = (j < limit) ? __expf(row[j] - mx) : 0.f;
      row[j] = e;
      sum += e;
    }
    sum = block_reduce(sum, false, red);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    T* yr = y + r * sk;
    for (int j = threadIdx.x; j < sk; j += kThreads) yr[j] = from_f<T>(row[j] * inv);
    __syncthreads();                     // row[] is reused by the next row of this CTA
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) softmax_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                               T* __restrict__ dx, long long rows, int sk, float scale) {
  __shared__ float red[kThreads / 32];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* dyr = dy + r * sk;
    const T* yr = y + r * sk;
    float dot = 0.f;
    for (int j = threadIdx.x; j < sk; j += kThreads) dot += to_f<T>(dyr[j]) * to_f<T>(yr[j]);
    dot = block_reduce(dot, false, red);
    T* dxr = dx + r * sk;
    for (int j = threadIdx.x; j < sk; j += kThreads) {
      const float p = to_f<T>(yr[j]);
      dxr[j] = from_f<T>(scale * p * (to_f<T>(dyr[j]) - dot));
    }
  }
}

#ifndef B2_EMU
int grid_rows(long long rows) {
  long long cap = static_cast<long long>(b2host::num_sms()) * 8;
  return static_cast<int>(rows < cap ? (rows < 1 ? 1 : rows) : cap);
}

int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}

#endif  // !B2_EMU
}  // namespace

#ifndef B2_EMU
extern "C" int b2_softmax_fwd(const void* x, const unsigned char* mask, void* y, long long rows, int sq, int sk, int heads,
                              int mask_batches, float scale, int causal, int dtype, void* stream) {
  if (rows <= 0 || sq <= 0 || sk <= 0 || heads <= 0 || sk > kMaxSk) {
    b2host::set_error("softmax_fwd: need rows, sq, heads > 0 and 0 < sk <= %d", kMaxSk);
    return -1;
  }
  if (rows % (static_cast<long long>(sq) * heads) != 0) { b2host::set_error("softmax_fwd: rows must be batches * heads * sq"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = static_cast<size_t>(sk) * sizeof(float);
  const int grid = grid_rows(rows);
  if (dtype == B2_DTYPE_BF16) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(softmax_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    softmax_fwd_kernel<__nv_bfloat16><<<grid, kThreads, smem, st>>>(static_cast<const __nv_bfloat16*>(x), mask, static_cast<__nv_bfloat16*>(y),
                                                                    rows, sq, sk, heads, mask_batches, scale, causal);
  } else if (dtype == B2_DTYPE_FP16) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(softmax_fwd_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    softmax_fwd_kernel<__half><<<grid, kThreads, smem, st>>>(static_cast<const __half*>(x), mask, static_cast<__half*>(y), rows, sq, sk, heads,
                                                            mask_batches, scale, causal);
  } else {
    if (smem > 48 * 1024) cudaFuncSetAttribute(softmax_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    softmax_fwd_kernel<float><<<grid, kThreads, smem, st>>>(static_cast<const float*>(x), mask, static_cast<float*>(y), rows, sq, sk, heads,
                                                           mask_batches, scale, causal);
  }
  return check("softmax_fwd");
}

extern "C" int b2_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk, float scale, int dtype, void* stream) {
  if (rows <= 0 || sk <= 0) { b2host::set_error("softmax_bwd: empty problem"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_rows(rows);
  if (dtype == B2_DTYPE_BF16)
    softmax_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y),
                                                                 static_cast<__nv_bfloat16*>(dx), rows, sk, scale);
  else if (dtype == B2_DTYPE_FP16)
    softmax_bwd_kernel<__half><<<grid, kThreads, 0, st>>>(static_cast<const __half*>(dy), static_cast<const __half*>(y), static_cast<__half*>(dx),
                                                          rows, sk, scale);
  else
    softmax_bwd_kernel<float><<<grid, kThreads, 0, st>>>(static_cast<const float*>(dy), static_cast<const float*>(y), static_cast<float*>(dx), rows,
                                                         sk, scale);
  return check("softmax_bwd");
}
#endif  // !B2_EMU
