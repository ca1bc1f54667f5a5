// fp32 convolution / pooling / log-softmax kernels of the MNIST CNN (sm_100a).
//
// replaces: the cuDNN convolutions and ATen max_pool2d / log_softmax / nll_loss kernels under
// 1_training_mnist_ddp/model_def.py:22-45.  The network is tiny and fp32 by definition of the
// workload (the reference trains it in fp32), so these are CUDA-core kernels shaped for locality, not tensor-core
// kernels: a 3x3 valid convolution of a 28x28 / 26x26 image keeps the WHOLE input image of one sample (<= 32 x 26 x 26
// fp32 = 86.5 KB) and the filter slice of its output-channel group in shared memory (Blackwell: up to 227 KB per CTA), so
// every input value is read from HBM once per channel group and every multiply reads shared memory only.
//   * conv2d_fwd   : y = relu?(conv(x, w) + b)            grid (batch, Cout / 8), 256 threads, 8 accumulators per thread
//   * conv2d_dgrad : dx = conv_transpose(dy * relu'(y), w) grid (batch, Cin / 8 rounded up), dy image in shared memory
//   * conv2d_wgrad : dw, db                                 grid (Cout, Cin): each CTA reduces its 9 taps over batch x pixels
//   * maxpool2x2 fwd (+ argmax byte) / bwd, log_softmax + NLL fwd / bwd for <= 32 classes (one warp per row)
// Layout NCHW, stride 1, no padding, square kernels up to 5x5 (one instantiation per size: the tap loops are unrolled).
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
constexpr int kCoTile = 8;       // output channels per CTA in fwd, input channels per CTA in dgrad
constexpr int kMaxK = 5;

#ifndef B2_EMU
int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
#endif  // !B2_EMU

// x [B, Cin, H, W], w [Cout, Cin, K, K], bias [Cout] | NULL, y [B, Cout, Ho, Wo]  (Ho = H - K + 1)
template <int K>
__global__ void __launch_bounds__(kThreads) conv2d_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                              const float* __restrict__ bias, float* __restrict__ y, int Cin,
                                                              int H, int W, int Cout, int relu) {
  B2_DYN_SMEM(float, sm);
  constexpr int KK = K * K;
  const int Ho = H - K + 1, Wo = W - K + 1;
  float* sx = sm;                                    // [Cin][H][W]
  float* sw = sm + Cin * H * W;                      // [kCoTile][Cin][K][K]
  const int b = blockIdx.x, co0 = blockIdx.y * kCoTile;
  const int nco = min(kCoTile, Cout - co0);
  const float* xb = x + static_cast<long long>(b) * Cin * H * W;
  for (int i = threadIdx.x; i < Cin * H * W; i += kThreads) sx[i] = xb[i];
  for (int i = threadIdx.x; i < kCoTile * Cin * KK; i += kThreads) {
    const int c = i / (Cin * KK);
    sw[i] = c < nco ? w[static_cast<long long>(co0) * Cin * KK + i] : 0.f;
  }
  __syncthreads();
  for (int p = threadIdx.x; p < Ho * Wo; p += kThreads) {
    const int oy = p / Wo, ox = p % Wo;
    float acc[kCoTile];
#pragma unroll
    for (int c = 0; c < kCoTile; ++c) acc[c] = 0.f;
    for (int ci = 0; ci < Cin; ++ci) {
      const float* px = sx + (ci * H + oy) * W + ox;
      const float* pw = sw + ci * KK;
#pragma unroll
      for (int ky = 0; ky < K; ++ky)
#pragma unroll
        for (int kx = 0; kx < K; ++kx) {
          const float v = px[ky * W + kx];
#pragma unroll
          for (int c = 0; c < kCoTile; ++c) acc[c] = fmaf(v, pw[c * Cin * KK + ky * K + kx], acc[c]);
        }
    }
#pragma unroll
    for (int c = 0; c < kCoTile; ++c) {
      if (c < nco) {
        float v = acc[c] + (bias ? bias[co0 + c] : 0.f);
        if (relu) v = fmaxf(v, 0.f);
        y[((static_cast<long long>(b) * Cout + co0 + c) * Ho + oy) * Wo + ox] = v;
      }
    }
  }
}

// dx[b, ci, iy, ix] = sum_co sum_ky,kx g[b, co, iy - ky, ix - kx] * w[co, ci, ky, kx],   g = dy * (y > 0) when relu
template <int K>
__global__ void __launch_bounds__(kThreads) conv2d_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ yact,
                                                                const float* __restrict__ w, float* __restrict__ dx, int Cin,
                                                                int H, int W, int Cout, int relu) {
  B2_DYN_SMEM(float, sm);
  constexpr int KK = K * K;
  const int Ho = H - K + 1, Wo = W - K + 1;
  float* sg = sm;                                    // [Cout][Ho][Wo]
  float* sw = sm + Cout * Ho * Wo;                   // [Cout][kCoTile][K][K]  (this CTA's input-channel slice)
  const int b = blockIdx.x, ci0 = blockIdx.y * kCoTile;
  const int nci = min(kCoTile, Cin - ci0);
  const long long off = static_cast<long long>(b) * Cout * Ho * Wo;
  for (int i = threadIdx.x; i < Cout * Ho * Wo; i += kThreads) {
    float g = dy[off + i];
    if (relu && !(yact[off + i] > 0.f)) g = 0.f;
    sg[i] = g;
  }
  for (int i = threadIdx.x; i < Cout * kCoTile * KK; i += kThreads) {
    const int co = i / (kCoTile * KK), c = (i / KK) % kCoTile, t = i % KK;
    sw[i] = c < nci ? w[(static_cast<long long>(co) * Cin + ci0 + c) * KK + t] : 0.f;
  }
  __syncthreads();
  for (int p = threadIdx.x; p < H * W; p += kThreads) {
    const int iy = p / W, ix = p % W;
    float acc[kCoTile];
#pragma unroll
    for (int c = 0; c < kCoTile; ++c) acc[c] = 0.f;
    for (int co = 0; co < Cout; ++co) {
      const float* pg = sg + co * Ho * Wo;
      const float* pw = sw + co * kCoTile * KK;
#pragma unroll
      for (int ky = 0; ky < K; ++ky) {
        const int oy = iy - ky;
        if (oy < 0 || oy >= Ho) continue;
#pragma unroll
        for (int kx = 0; kx < K; ++kx) {
          const int ox = ix - kx;
          if (ox < 0 || ox >= Wo) continue;
          const float g = pg[oy * Wo + ox];
#pragma unroll
          for (int c = 0; c < kCoTile; ++c) acc[c] = fmaf(g, pw[c * KK + ky * K + kx], acc[c]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kCoTile; ++c)
      if (c < nci) dx[((static_cast<long long>(b) * Cin + ci0 + c) * H + iy) * W + ix] = acc[c];
  }
}

// dw[co, ci, ky, kx] = sum_b sum_oy,ox g[b, co, oy, ox] * x[b, ci, oy + ky, ox + kx];  db[co] = sum g  (written by ci == 0)
template <int K>
__global__ void __launch_bounds__(kThreads) conv2d_wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ yact,
                                                                const float* __restrict__ x, float* __restrict__ dw,
                                                                float* __restrict__ db, int B, int Cin, int H, int W, int Cout,
                                                                int relu, int accumulate) {
  constexpr int KK = K * K;
  const int Ho = H - K + 1, Wo = W - K + 1;
  const int co = blockIdx.x, ci = blockIdx.y;
  float acc[KK];
#pragma unroll
  for (int t = 0; t < KK; ++t) acc[t] = 0.f;
  float bsum = 0.f;
  const long long total = static_cast<long long>(B) * Ho * Wo;
  for (long long i = threadIdx.x; i < total; i += kThreads) {
    const int b = static_cast<int>(i / (Ho * Wo)), p = static_cast<int>(i % (Ho * Wo));
    const int oy = p / Wo, ox = p % Wo;
    const long long gi = ((static_cast<long long>(b) * Cout + co) * Ho + oy) * Wo + ox;
    float g = dy[gi];
    if (relu && !(yact[gi] > 0.f)) g = 0.f;
    bsum += g;
    const float* px = x + ((static_cast<long long>(b) * Cin + ci) * H + oy) * W + ox;
#pragma unroll
    for (int ky = 0; ky < K; ++ky)
#pragma unroll
      for (int kx = 0; kx < K; ++kx) acc[ky * K + kx] = fmaf(g, px[ky * W + kx], acc[ky * K + kx]);
  }
  __shared__ float red[kThreads / 32][KK + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t <= KK; ++t) {
    float v = t < KK ? acc[t < KK ? t : 0] : bsum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][t] = v;
  }
  __syncthreads();
  if (threadIdx.x <= KK) {
    float v = 0.f;
    for (int wq = 0; wq < kThreads / 32; ++wq) v += red[wq][threadIdx.x];
    if (threadIdx.x < KK) {
      float* o = dw + (static_cast<long long>(co) * Cin + ci) * KK + threadIdx.x;
      *o = accumulate ? *o + v : v;
    } else if (ci == 0 && db) {
      db[co] = accumulate ? db[co] + v : v;
    }
  }
}

// 2x2 max pooling, stride 2 (floor).  idx: which of the 4 taps won (first maximum, like ATen)
__global__ void maxpool2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, unsigned char* __restrict__ idx,
                                    long long planes, int H, int W) {
  const int Ho = H / 2, Wo = W / 2;
  const long long n = planes * Ho * Wo;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long pl = i / (Ho * Wo);
    const int p = static_cast<int>(i % (Ho * Wo)), oy = p / Wo, ox = p % Wo;
    const float* px = x + (pl * H + 2 * oy) * W + 2 * ox;
    float best = px[0];
    int bi = 0;
    const float c1 = px[1], c2 = px[W], c3 = px[W + 1];
    if (c1 > best || c1 != c1) { best = c1; bi = 1; }
    if (c2 > best || c2 != c2) { best = c2; bi = 2; }
    if (c3 > best || c3 != c3) { best = c3; bi = 3; }
    y[i] = best;
    idx[i] = static_cast<unsigned char>(bi);
  }
}

__global__ void maxpool2_bwd_kernel(const float* __restrict__ dy, const unsigned char* __restrict__ idx, float* __restrict__ dx,
                                    long long planes, int H, int W) {
  const int Ho = H / 2, Wo = W / 2;
  const long long n = planes * H * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long pl = i / (H * W);
    const int p = static_cast<int>(i % (H * W)), iy = p / W, ix = p % W;
    const int oy = iy / 2, ox = ix / 2;
    float g = 0.f;
    if (oy < Ho && ox < Wo) {
      const long long o = (pl * Ho + oy) * Wo + ox;
      if (idx[o] == ((iy & 1) * 2 + (ix & 1))) g = dy[o];
    }
    dx[i] = g;
  }
}

// one warp per row, C <= 32 classes: logp = x - logsumexp(x); loss[row] = -logp[target]
__global__ void log_softmax_nll_fwd_kernel(const float* __restrict__ x, const long long* __restrict__ target, float* __restrict__ logp,
                                           float* __restrict__ loss, long long rows, int C) {
  const long long row = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float v = lane < C ? x[row * C + lane] : -INFINITY;
  float mx = v;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float e = lane < C ? __expf(v - mx) : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
  const float lp = v - mx - __logf(e);
  if (lane < C) logp[row * C + lane] = lp;
  if (loss && target) {
    const int t = static_cast<int>(target[row]);
    const float lt = __shfl_sync(0xffffffffu, lp, t & 31);
    if (lane == 0) loss[row] = (t >= 0 && t < C) ? -lt : 0.f;
  }
}

// dx = exp(logp) * sum_j dlogp_j - dlogp   (general log_softmax backward);  nll: dlogp[row, target] = -gscale[row]
__global__ void log_softmax_nll_bwd_kernel(const float* __restrict__ logp, const float* __restrict__ dlogp,
                                           const long long* __restrict__ target, const float* __restrict__ gloss, float gscalar,
                                           float* __restrict__ dx, long long rows, int C) {
  const long long row = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float d = 0.f;
  if (lane < C) {
    if (dlogp) d = dlogp[row * C + lane];
    if (target && target[row] == lane) d -= (gloss ? gloss[row] : 1.f) * gscalar;
  }
  float s = d;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane < C) dx[row * C + lane] = d - __expf(logp[row * C + lane]) * s;
}

#ifndef B2_EMU
int grid1d(long long n, int threads) {
  long long g = (n + threads - 1) / threads, cap = static_cast<long long>(b2host::num_sms()) * 16;
  return static_cast<int>(g < 1 ? 1 : (g > cap ? cap : g));
}

bool conv_args_ok(const char* what, int B, int Cin, int H, int W, int Cout, int K) {
  if (B <= 0 || Cin <= 0 || Cout <= 0 || K < 1 || K > kMaxK || H < K || W < K) {
    b2host::set_error("%s: unsupported shape (kernel size 1..%d, valid convolution)", what, kMaxK);
    return false;
  }
  return true;
}

#endif  // !B2_EMU
}  // namespace

#ifndef B2_EMU
extern "C" int b2_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout, int K,
                             int relu, void* stream) {
  if (!conv_args_ok("conv2d_fwd", B, Cin, H, W, Cout, K)) return -1;
  const size_t smem = (static_cast<size_t>(Cin) * H * W + static_cast<size_t>(kCoTile) * Cin * K * K) * sizeof(float);
  if (smem > 200 * 1024) { b2host::set_error("conv2d_fwd: one input image (%zu bytes) does not fit in shared memory", smem); return -1; }
  dim3 grid(B, (Cout + kCoTile - 1) / kCoTile);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define B2_CONV_FWD(KV)                                                                                                        \
  case KV:                                                                                                                     \
    if (smem > 48 * 1024) cudaFuncSetAttribute(conv2d_fwd_kernel<KV>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)); \
    conv2d_fwd_kernel<KV><<<grid, kThreads, smem, st>>>(x, w, bias, y, Cin, H, W, Cout, relu);                                 \
    break;
  switch (K) { B2_CONV_FWD(1) B2_CONV_FWD(2) B2_CONV_FWD(3) B2_CONV_FWD(4) B2_CONV_FWD(5) }
#undef B2_CONV_FWD
  return check("conv2d_fwd");
}

extern "C" int b2_conv2d_bwd(const float* dy, const float* yact, const float* x, const float* w, float* dx, float* dw, float* db, int B,
                             int Cin, int H, int W, int Cout, int K, int relu, int accumulate, void* stream) {
  if (!conv_args_ok("conv2d_bwd", B, Cin, H, W, Cout, K)) return -1;
  if (relu && !yact) { b2host::set_error("conv2d_bwd: relu backward needs the forward output"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int Ho = H - K + 1, Wo = W - K + 1;
  if (dx) {
    const size_t smem = (static_cast<size_t>(Cout) * Ho * Wo + static_cast<size_t>(Cout) * kCoTile * K * K) * sizeof(float);
    if (smem > 200 * 1024) { b2host::set_error("conv2d_bwd: one gradient image (%zu bytes) does not fit in shared memory", smem); return -1; }
    dim3 grid(B, (Cin + kCoTile - 1) / kCoTile);
#define B2_CONV_DGRAD(KV)                                                                                                      \
  case KV:                                                                                                                     \
    if (smem > 48 * 1024) cudaFuncSetAttribute(conv2d_dgrad_kernel<KV>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)); \
    conv2d_dgrad_kernel<KV><<<grid, kThreads, smem, st>>>(dy, yact, w, dx, Cin, H, W, Cout, relu);                             \
    break;
    switch (K) { B2_CONV_DGRAD(1) B2_CONV_DGRAD(2) B2_CONV_DGRAD(3) B2_CONV_DGRAD(4) B2_CONV_DGRAD(5) }
#undef B2_CONV_DGRAD
    if (check("conv2d_dgrad")) return -1;
  }
  if (dw) {
    dim3 grid(Cout, Cin);
#define B2_CONV_WGRAD(KV) \
  case KV: conv2d_wgrad_kernel<KV><<<grid, kThreads, 0, st>>>(dy, yact, x, dw, db, B, Cin, H, W, Cout, relu, accumulate); break;
    switch (K) { B2_CONV_WGRAD(1) B2_CONV_WGRAD(2) B2_CONV_WGRAD(3) B2_CONV_WGRAD(4) B2_CONV_WGRAD(5) }
#undef B2_CONV_WGRAD
    if (check("conv2d_wgrad")) return -1;
  }
  return 0;
}

extern "C" int b2_maxpool2_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, void* stream) {
  if (planes <= 0 || H < 2 || W < 2) { b2host::set_error("maxpool2_fwd: bad shape"); return -1; }
  maxpool2_fwd_kernel<<<grid1d(planes * (H / 2) * (W / 2), 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, idx, planes, H, W);
  return check("maxpool2_fwd");
}

extern "C" int b2_maxpool2_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, void* stream) {
  if (planes <= 0 || H < 2 || W < 2) { b2host::set_error("maxpool2_bwd: bad shape"); return -1; }
  maxpool2_bwd_kernel<<<grid1d(planes * H * W, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, idx, dx, planes, H, W);
  return check("maxpool2_bwd");
}

extern "C" int b2_log_softmax_nll_fwd(const float* x, const long long* target, float* logp, float* loss, long long rows, int C, void* stream) {
  if (rows <= 0 || C < 1 || C > 32) { b2host::set_error("log_softmax_nll_fwd: 1..32 classes"); return -1; }
  const long long threads = rows * 32;
  log_softmax_nll_fwd_kernel<<<static_cast<int>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, target, logp, loss, rows, C);
  return check("log_softmax_nll_fwd");
}

extern "C" int b2_log_softmax_nll_bwd(const float* logp, const float* dlogp, const long long* target, const float* gloss, float gscalar,
                                      float* dx, long long rows, int C, void* stream) {
  if (rows <= 0 || C < 1 || C > 32) { b2host::set_error("log_softmax_nll_bwd: 1..32 classes"); return -1; }
  const long long threads = rows * 32;
  log_softmax_nll_bwd_kernel<<<static_cast<int>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(logp, dlogp, target, gloss, gscalar,
                                                                                                                    dx, rows, C);
  return check("log_softmax_nll_bwd");
}
#endif  // !B2_EMU
