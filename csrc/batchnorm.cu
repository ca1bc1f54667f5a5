// BatchNorm2d (training and inference) forward / backward over NCHW tensors, fp32 statistics (sm_100a).
//
// replaces: cuDNN / ATen batch_norm under the torchvision ResNet that `pytorch_oxford_ddp.py --model_name resnet50` (the
// script's default, :123) builds, and apex `syncbn`: the per-channel sums are exposed between the two phases so
// the Python wrapper can all-reduce them over the data-parallel group (SyncBatchNorm) before statistics are finalised.
//
// Four memory-bound passes, each reading the activation once with coalesced accesses along H*W:
//   bn_reduce      grid (C, splits): per-channel sum(x), sum(x^2)            -> double atomics into sums[C][2]
//   bn_finalize    C threads: mean, rstd, scale = gamma*rstd, shift = beta - mean*scale, running-stat update
//   bn_apply       y = x*scale[c] + shift[c]  (+ ReLU)
//   bn_bwd_reduce  grid (C, splits): sum(g), sum(g * xhat)   (g = dy masked by y > 0 when ReLU is fused)
//   bn_bwd_apply   dx = scale * (g - mean_g - xhat * mean_gx);  dgamma = sum(g*xhat), dbeta = sum(g)
// Partial sums are fp32 per thread (a few thousand elements at most), combined in double across the CTA and across CTAs,
// so E[x^2] - mean^2 keeps its digits for activations with a large mean.
#ifdef B2_EMU                      // CPU emulation of the CUDA execution model (tests/emu/cuda_emu.h): kernels only
#include "cuda_emu.h"
namespace b2 {}
#else
#include "common.cuh"
#endif
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int kThreads = 256;

template <typename T> B2_DEVICE float ldf(const T* p, long long i);
template <> B2_DEVICE float ldf<float>(const float* p, long long i) { return p[i]; }
template <> B2_DEVICE float ldf<__nv_bfloat16>(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }
template <> B2_DEVICE float ldf<__half>(const __half* p, long long i) { return __half2float(p[i]); }
template <typename T> B2_DEVICE void stf(T* p, long long i, float v);
template <> B2_DEVICE void stf<float>(float* p, long long i, float v) { p[i] = v; }
template <> B2_DEVICE void stf<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16_rn(v); }
template <> B2_DEVICE void stf<__half>(__half* p, long long i, float v) { p[i] = __float2half_rn(v); }

B2_DEVICE void block_sum2_atomic(float a, float b, double* out2) {
  double da = a, db = b;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    da += __shfl_xor_sync(0xffffffffu, da, o);
    db += __shfl_xor_sync(0xffffffffu, db, o);
  }
  __shared__ double red[kThreads / 32][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { red[warp][0] = da; red[warp][1] = db; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double sa = 0, sb = 0;
    for (int w = 0; w < kThreads / 32; ++w) { sa += red[w][0]; sb += red[w][1]; }
    atomicAdd(out2, sa);
    atomicAdd(out2 + 1, sb);
  }
}

// element e of channel c, e in [0, N*HW): n = e / HW, p = e % HW  ->  offset (n*C + c)*HW + p
// I = unsigned (tensors below 2^31 elements: 32-bit index arithmetic, no software 64-bit division) or long long
template <typename T, typename I>
__global__ void __launch_bounds__(kThreads) bn_reduce_kernel(const T* __restrict__ x, double* __restrict__ sums, int N, int C, int HW) {
  const I c = blockIdx.x;
  const I M = static_cast<I>(N) * static_cast<I>(HW);
  float s = 0.f, q = 0.f;
  for (I e = static_cast<I>(blockIdx.y) * kThreads + threadIdx.x; e < M; e += static_cast<I>(gridDim.y) * kThreads) {
    const I n = e / static_cast<I>(HW), p = e - n * static_cast<I>(HW);
    const float v = ldf<T>(x, static_cast<long long>((n * static_cast<I>(C) + c) * static_cast<I>(HW) + p));
    s += v;
    q = fmaf(v, v, q);
  }
  block_sum2_atomic(s, q, sums + 2 * c);
}

// sums hold GLOBAL (already all-reduced for SyncBN) totals over `count` elements per channel
__global__ void bn_finalize_kernel(const double* __restrict__ sums, double count, const float* __restrict__ gamma, const float* __restrict__ beta,
                                   float eps, float momentum, float* __restrict__ running_mean, float* __restrict__ running_var,
                                   float* __restrict__ mean, float* __restrict__ rstd, float* __restrict__ scale, float* __restrict__ shift, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double m = sums[2 * c] / count;
  double var = sums[2 * c + 1] / count - m * m;
  if (var < 0) var = 0;
  const float r = rsqrtf(static_cast<float>(var) + eps);
  mean[c] = static_cast<float>(m);
  rstd[c] = r;
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * r;
  shift[c] = b - static_cast<float>(m) * g * r;
  if (running_mean) {
    const double unbiased = count > 1 ? var * count / (count - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * static_cast<float>(m);
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * static_cast<float>(unbiased);
  }
}

// inference: scale / shift from the running statistics
__global__ void bn_eval_coeffs_kernel(const float* __restrict__ running_mean, const float* __restrict__ running_var, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, float eps, float* __restrict__ scale, float* __restrict__ shift, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float r = rsqrtf(running_var[c] + eps), g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * r;
  shift[c] = b - running_mean[c] * g * r;
}

template <typename T, typename I>
__global__ void __launch_bounds__(kThreads) bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                                                            T* __restrict__ y, long long total_ll, int C, int HW, int relu) {
  const I total = static_cast<I>(total_ll);
  for (I i = static_cast<I>(blockIdx.x) * kThreads + threadIdx.x; i < total; i += static_cast<I>(gridDim.x) * kThreads) {
    const int c = static_cast<int>((i / static_cast<I>(HW)) % static_cast<I>(C));
    float v = fmaf(ldf<T>(x, i), scale[c], shift[c]);
    if (relu) v = fmaxf(v, 0.f);
    stf<T>(y, i, v);
  }
}

template <typename T, typename I>
__global__ void __launch_bounds__(kThreads) bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ y,
                                                                 const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                 double* __restrict__ sums, int N, int C, int HW, int relu) {
  const I c = blockIdx.x;
  const I M = static_cast<I>(N) * static_cast<I>(HW);
  const float mu = mean[c], r = rstd[c];
  float sg = 0.f, sgx = 0.f;
  for (I e = static_cast<I>(blockIdx.y) * kThreads + threadIdx.x; e < M; e += static_cast<I>(gridDim.y) * kThreads) {
    const I n = e / static_cast<I>(HW), p = e - n * static_cast<I>(HW);
    const long long off = static_cast<long long>((n * static_cast<I>(C) + c) * static_cast<I>(HW) + p);
    float g = ldf<T>(dy, off);
    if (relu && !(ldf<T>(y, off) > 0.f)) g = 0.f;
    sg += g;
    sgx = fmaf(g, (ldf<T>(x, off) - mu) * r, sgx);
  }
  block_sum2_atomic(sg, sgx, sums + 2 * c);
}

// sums: GLOBAL sum(g), sum(g*xhat) over `count` elements per channel
template <typename T, typename I>
__global__ void __launch_bounds__(kThreads) bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ y,
                                                                const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                const float* __restrict__ gamma, const double* __restrict__ sums, double inv_count,
                                                                T* __restrict__ dx, long long total_ll, int C, int HW, int relu) {
  const I total = static_cast<I>(total_ll);
  for (I i = static_cast<I>(blockIdx.x) * kThreads + threadIdx.x; i < total; i += static_cast<I>(gridDim.x) * kThreads) {
    const int c = static_cast<int>((i / static_cast<I>(HW)) % static_cast<I>(C));
    float g = ldf<T>(dy, i);
    if (relu && !(ldf<T>(y, i) > 0.f)) g = 0.f;
    const float r = rstd[c], xhat = (ldf<T>(x, i) - mean[c]) * r;
    const float mg = static_cast<float>(sums[2 * c] * inv_count), mgx = static_cast<float>(sums[2 * c + 1] * inv_count);
    stf<T>(dx, i, (gamma ? gamma[c] : 1.f) * r * (g - mg - xhat * mgx));
  }
}

#ifndef B2_EMU
int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}

int splits_for(int C, long long M) {
  // enough CTAs to cover the machine a few times over, but at least ~4096 elements per CTA
  long long want = (static_cast<long long>(b2host::num_sms()) * 4 + C - 1) / C;
  long long cap = (M + 4095) / 4096;
  long long s = want < cap ? want : cap;
  return static_cast<int>(s < 1 ? 1 : s);
}

int grid1d(long long n) {
  long long g = (n + kThreads - 1) / kThreads, cap = static_cast<long long>(b2host::num_sms()) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

// T from the dtype code, I = unsigned when every element offset (plus one grid stride) fits in 32 bits
#define BN_DISPATCH(dtype, total, CALL)                                                              \
  do {                                                                                               \
    if ((total) < (1LL << 31)) {                                                                     \
      using I = unsigned;                                                                            \
      if ((dtype) == B2_DTYPE_BF16) { using T = __nv_bfloat16; CALL; }                               \
      else if ((dtype) == B2_DTYPE_FP16) { using T = __half; CALL; }                                 \
      else { using T = float; CALL; }                                                                \
    } else {                                                                                         \
      using I = long long;                                                                           \
      if ((dtype) == B2_DTYPE_BF16) { using T = __nv_bfloat16; CALL; }                               \
      else if ((dtype) == B2_DTYPE_FP16) { using T = __half; CALL; }                                 \
      else { using T = float; CALL; }                                                                \
    }                                                                                                \
  } while (0)

#endif  // !B2_EMU
}  // namespace

#ifndef B2_EMU
// phase 1 of the training forward: sums[C][2] (double, ZERO on entry) += per-channel sum(x), sum(x^2)
extern "C" int b2_bn_reduce(const void* x, double* sums, int N, int C, int HW, int dtype, void* stream) {
  if (N <= 0 || C <= 0 || HW <= 0) { b2host::set_error("bn_reduce: empty tensor"); return -1; }
  dim3 grid(C, splits_for(C, static_cast<long long>(N) * HW));
  BN_DISPATCH(dtype, static_cast<long long>(N) * C * HW,
              (bn_reduce_kernel<T, I><<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const T*>(x), sums, N, C, HW)));
  return check("bn_reduce");
}

// phase 2: statistics + affine coefficients from the (global) sums over `count` elements per channel; updates the running
// statistics when given (momentum as in torch: new = (1 - momentum) * old + momentum * batch, unbiased variance)
extern "C" int b2_bn_finalize(const double* sums, double count, const float* gamma, const float* beta, float eps, float momentum,
                              float* running_mean, float* running_var, float* mean, float* rstd, float* scale, float* shift, int C, void* stream) {
  if (C <= 0 || count <= 0) { b2host::set_error("bn_finalize: empty"); return -1; }
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(sums, count, gamma, beta, eps, momentum, running_mean, running_var,
                                                                                     mean, rstd, scale, shift, C);
  return check("bn_finalize");
}

extern "C" int b2_bn_eval_coeffs(const float* running_mean, const float* running_var, const float* gamma, const float* beta, float eps,
                                 float* scale, float* shift, int C, void* stream) {
  if (C <= 0) { b2host::set_error("bn_eval_coeffs: empty"); return -1; }
  bn_eval_coeffs_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(running_mean, running_var, gamma, beta, eps, scale, shift, C);
  return check("bn_eval_coeffs");
}

// phase 3: y = x * scale[c] + shift[c] (+ ReLU)
extern "C" int b2_bn_apply(const void* x, const float* scale, const float* shift, void* y, int N, int C, int HW, int relu, int dtype, void* stream) {
  const long long total = static_cast<long long>(N) * C * HW;
  if (total <= 0) { b2host::set_error("bn_apply: empty tensor"); return -1; }
  BN_DISPATCH(dtype, total, (bn_apply_kernel<T, I><<<grid1d(total), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
                                static_cast<const T*>(x), scale, shift, static_cast<T*>(y), total, C, HW, relu)));
  return check("bn_apply");
}

// backward phase 1: sums[C][2] (double, ZERO on entry) += sum(g), sum(g * xhat); y is needed only when relu != 0
extern "C" int b2_bn_bwd_reduce(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, double* sums, int N, int C,
                                int HW, int relu, int dtype, void* stream) {
  if (N <= 0 || C <= 0 || HW <= 0) { b2host::set_error("bn_bwd_reduce: empty tensor"); return -1; }
  if (relu && !y) { b2host::set_error("bn_bwd_reduce: fused ReLU backward needs the forward output"); return -1; }
  dim3 grid(C, splits_for(C, static_cast<long long>(N) * HW));
  BN_DISPATCH(dtype, static_cast<long long>(N) * C * HW, (bn_bwd_reduce_kernel<T, I><<<grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
                         static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, sums, N, C, HW, relu)));
  return check("bn_bwd_reduce");
}

// backward phase 2: dx from the (global) sums; dgamma[c] = sums[c][1], dbeta[c] = sums[c][0] are read by the caller
extern "C" int b2_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const float* gamma,
                               const double* sums, double count, void* dx, int N, int C, int HW, int relu, int dtype, void* stream) {
  const long long total = static_cast<long long>(N) * C * HW;
  if (total <= 0 || count <= 0) { b2host::set_error("bn_bwd_apply: empty tensor"); return -1; }
  BN_DISPATCH(dtype, total, (bn_bwd_apply_kernel<T, I><<<grid1d(total), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
                         static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, gamma, sums, 1.0 / count,
                         static_cast<T*>(dx), total, C, HW, relu)));
  return check("bn_bwd_apply");
}
#endif  // !B2_EMU
