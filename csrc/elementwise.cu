// Memory-bound fused elementwise / reduction kernels (sm_100a): 128-bit accesses, fp32 math.
//
// replaces: Megatron's torch.jit (NVFuser) bias_gelu / bias_dropout_add fusions,
// HF LLaMA's unfused SwiGLU and rotate_half RoPE (K24), ATen relu/max_pool/log_softmax glue of the
// MNIST model (K2), and the bias-gradient column reductions autograd runs after every Linear.
#include <type_traits>
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {

template <typename T> struct Cvt;
template <> struct Cvt<__nv_bfloat16> {
  static B2_DEVICE float2 up(uint32_t u) { return unpack_bf16x2(u); }
  static B2_DEVICE uint32_t down(float a, float b) { return pack_bf16x2(a, b); }
};
template <> struct Cvt<__half> {
  static B2_DEVICE float2 up(uint32_t u) { __half2 h = *reinterpret_cast<__half2*>(&u); return __half22float2(h); }
  static B2_DEVICE uint32_t down(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
};
template <typename T> B2_DEVICE void load8(const T* p, float* f) {
  uint4 q = *reinterpret_cast<const uint4*>(p);
  float2 a = Cvt<T>::up(q.x), b = Cvt<T>::up(q.y), c = Cvt<T>::up(q.z), d = Cvt<T>::up(q.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <typename T> B2_DEVICE void store8(T* p, const float* f) {
  *reinterpret_cast<uint4*>(p) = make_uint4(Cvt<T>::down(f[0], f[1]), Cvt<T>::down(f[2], f[3]),
                                            Cvt<T>::down(f[4], f[5]), Cvt<T>::down(f[6], f[7]));
}

// Philox4x32-10 counter RNG: the dropout mask is a pure function of (seed, vector index), so the
// backward pass regenerates it instead of storing it.
B2_DEVICE uint4 philox(uint64_t seed, uint64_t ctr) {
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
  uint32_t c0 = static_cast<uint32_t>(ctr), c1 = static_cast<uint32_t>(ctr >> 32), c2 = 0x2718u, c3 = 0x3141u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    uint32_t n0 = h1 ^ c1 ^ k0, n1 = l1, n2 = h0 ^ c3 ^ k1, n3 = l0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// 8 keep-decisions (16 random bits each) for one 8-element vector
B2_DEVICE void keep8(uint64_t seed, uint64_t vec_idx, float p, bool* keep) {
  uint4 r = philox(seed, vec_idx);
  const uint32_t thr = static_cast<uint32_t>(p * 65536.f);
  uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    keep[2 * i] = (w[i] & 0xFFFFu) >= thr;
    keep[2 * i + 1] = (w[i] >> 16) >= thr;
  }
}

B2_DEVICE float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * x * (1.f + k1 * x * x);
  float t = 1.f - 2.f / (1.f + __expf(2.f * u));
  return 0.5f * x * (1.f + t);
}
B2_DEVICE float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * x * (1.f + k1 * x * x);
  float t = 1.f - 2.f / (1.f + __expf(2.f * u));
  float du = k0 * (1.f + 3.f * k1 * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}
// erf through Abramowitz-Stegun 7.1.26 (absolute error < 6e-7 with fp32 arithmetic, far below 16-bit output rounding): one
// rcp + one ex2 and six FMAs, no branch - erff() costs about three times as many issue slots, which is what bounds the
// epilogue warps of the GELU GEMMs.  `e` returns exp(-z^2), which the derivative needs anyway.
B2_DEVICE float erf_as(float z, float& e) {
  const float az = fabsf(z);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, az, 1.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(az * az * -1.4426950408889634f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  return copysignf(fmaf(-p * t, e, 1.f), z);
}
B2_DEVICE float gelu_erf(float x) {
  float e;
  const float r = erf_as(x * 0.7071067811865476f, e);
  const float hx = 0.5f * x;
  return fmaf(hx, r, hx);
}
B2_DEVICE float gelu_erf_grad(float x) {       // Phi(x) + x phi(x);  exp(-x^2 / 2) is erf_as's by-product
  float e;
  const float r = erf_as(x * 0.7071067811865476f, e);
  return fmaf(x * 0.3989422804014327f, e, fmaf(0.5f, r, 0.5f));
}

// out = residual + dropout(x + bias)           (training: p > 0 uses the Philox mask)
template <typename T>
__global__ void bias_dropout_add_kernel(const T* __restrict__ x, const T* __restrict__ bias, const T* __restrict__ res,
                                        T* __restrict__ out, long long nvec, int H, float p, uint64_t seed,
                                        const unsigned long long* __restrict__ replay) {
  seed = replay_seed(seed, replay);
  const float inv = p > 0.f ? 1.f / (1.f - p) : 1.f;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], r[8], o[8];
    load8<T>(x + v * 8, a);
    if (bias) { load8<T>(bias + (v * 8) % H, b); }
    if (res) load8<T>(res + v * 8, r);
    bool keep[8];
    if (p > 0.f) keep8(seed, static_cast<uint64_t>(v), p, keep);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float t = a[e] + (bias ? b[e] : 0.f);
      if (p > 0.f) t = keep[e] ? t * inv : 0.f;
      o[e] = t + (res ? r[e] : 0.f);
    }
    store8<T>(out + v * 8, o);
  }
}
// dx = dropout_mask * dy / (1-p)
template <typename T>
__global__ void dropout_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, long long nvec, float p, uint64_t seed,
                                   const unsigned long long* __restrict__ replay) {
  seed = replay_seed(seed, replay);
  const float inv = 1.f / (1.f - p);
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    float a[8], o[8];
    load8<T>(dy + v * 8, a);
    bool keep[8];
    keep8(seed, static_cast<uint64_t>(v), p, keep);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = keep[e] ? a[e] * inv : 0.f;
    store8<T>(dx + v * 8, o);
  }
}

// mode 0: tanh GeLU, 1: erf GeLU, 2: ReLU, 3: SiLU
template <typename T>
__global__ void bias_act_fwd_kernel(const T* __restrict__ x, const T* __restrict__ bias, T* __restrict__ out,
                                    long long nvec, int H, int mode) {
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], o[8];
    load8<T>(x + v * 8, a);
    if (bias) load8<T>(bias + (v * 8) % H, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float z = a[e] + (bias ? b[e] : 0.f);
      o[e] = mode == 0 ? gelu_tanh(z) : mode == 1 ? gelu_erf(z) : mode == 2 ? fmaxf(z, 0.f) : z / (1.f + __expf(-z));
    }
    store8<T>(out + v * 8, o);
  }
}
template <typename T>
__global__ void bias_act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ bias,
                                    T* __restrict__ dx, long long nvec, int H, int mode) {
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    float g[8], a[8], b[8], o[8];
    load8<T>(dy + v * 8, g);
    load8<T>(x + v * 8, a);
    if (bias) load8<T>(bias + (v * 8) % H, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float z = a[e] + (bias ? b[e] : 0.f);
      float d;
      if (mode == 0) d = gelu_tanh_grad(z);
      else if (mode == 1) d = gelu_erf_grad(z);
      else if (mode == 2) d = z > 0.f ? 1.f : 0.f;
      else { float s = 1.f / (1.f + __expf(-z)); d = s * (1.f + z * (1.f - s)); }
      o[e] = g[e] * d;
    }
    store8<T>(dx + v * 8, o);
  }
}

// SwiGLU on a fused [rows, 2F] gate|up projection: out[r, f] = silu(gu[r, f]) * gu[r, F + f]
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ gu, T* __restrict__ out, long long rows, int F) {
  const int fv = F >> 3;
  const long long nvec = rows * fv;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    long long r = v / fv; int c = static_cast<int>(v - r * fv) * 8;
    float g[8], u[8], o[8];
    load8<T>(gu + r * 2 * F + c, g);
    load8<T>(gu + r * 2 * F + F + c, u);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = g[e] / (1.f + __expf(-g[e])) * u[e];
    store8<T>(out + r * F + c, o);
  }
}
template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ gu, T* __restrict__ dgu, long long rows, int F) {
  const int fv = F >> 3;
  const long long nvec = rows * fv;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    long long r = v / fv; int c = static_cast<int>(v - r * fv) * 8;
    float d[8], g[8], u[8], og[8], ou[8];
    load8<T>(dy + r * F + c, d);
    load8<T>(gu + r * 2 * F + c, g);
    load8<T>(gu + r * 2 * F + F + c, u);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float s = 1.f / (1.f + __expf(-g[e]));
      ou[e] = d[e] * g[e] * s;
      og[e] = d[e] * u[e] * s * (1.f + g[e] * (1.f - s));
    }
    store8<T>(dgu + r * 2 * F + c, og);
    store8<T>(dgu + r * 2 * F + F + c, ou);
  }
}

// dbias[n] (+)= sum_m dy[m, n]      fp32 output (main_grad), column stripes of 64, atomics per stripe
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ dy, float* __restrict__ out, int rows, int N, long long ld, int rows_per_cta) {
  __shared__ float red[32][65];
  const int cv = threadIdx.x & 7;         // 8 vectors of 8 = 64 columns
  const int rl = threadIdx.x >> 3;        // 32 row lanes
  const int col = blockIdx.x * 64 + cv * 8;
  const int r0 = blockIdx.y * rows_per_cta;
  const int r1 = min(rows, r0 + rows_per_cta);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < N) {
    for (int r = r0 + rl; r < r1; r += 32) {
      float a[8];
      load8<T>(dy + static_cast<long long>(r) * ld + col, a);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += a[e];
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[rl][cv * 8 + e] = acc[e];
  __syncthreads();
  if (threadIdx.x < 64) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < 32; ++r) s += red[r][threadIdx.x];
    int c = blockIdx.x * 64 + threadIdx.x;
    if (c < N) atomicAdd(out + c, s);
  }
}

// rotary embedding applied in place to the q and k slices of a packed activation.
// x viewed as [tokens, nheads, hd] with arbitrary token stride; rotate-half (HF LLaMA) convention.
template <typename T>
__global__ void rope_kernel(T* __restrict__ x, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                            const int* __restrict__ pos, long long tokens, int nheads, int hd, long long tok_stride,
                            int head_stride, float sign) {
  const int half = hd >> 1;
  const long long total = tokens * nheads * half;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int d = static_cast<int>(i % half);
    long long th = i / half;
    int h = static_cast<int>(th % nheads);
    long long t = th / nheads;
    int p = pos[t];
    float c = cos_t[static_cast<long long>(p) * half + d], s = sin_t[static_cast<long long>(p) * half + d] * sign;
    T* base = x + t * tok_stride + static_cast<long long>(h) * head_stride;
    float a = static_cast<float>(base[d]), b = static_cast<float>(base[d + half]);
    base[d] = static_cast<T>(a * c - b * s);
    base[d + half] = static_cast<T>(b * c + a * s);
  }
}

// out[i] = a[i] * alpha + b[i] * beta     (residual adds, gradient accumulation in low precision)
template <typename T>
__global__ void axpby_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, long long nvec, float alpha, float beta) {
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8], o[8];
    load8<T>(a + v * 8, x);
    if (b) load8<T>(b + v * 8, y);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = x[e] * alpha + (b ? y[e] * beta : 0.f);
    store8<T>(out + v * 8, o);
  }
}

// fp32 main_grad += low-precision grad  (bias / norm grads that arrive in bf16)
template <typename T>
__global__ void accum_fp32_kernel(float* __restrict__ dst, const T* __restrict__ src, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] += static_cast<float>(src[i]);
}
// dst(T) = src(fp32)   : main-param -> model-param copy-cast
template <typename T>
__global__ void cast_from_fp32_kernel(T* __restrict__ dst, const float* __restrict__ src, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = static_cast<T>(src[i]);
}

// dst[r, 0:cols] (T, row stride ld_dst) = src[r, 0:cols] (fp32, contiguous); cols % 8 == 0, 16-byte aligned rows.
// Used to drop an fp32 accumulator (attention dQ) into a slice of a packed 16-bit buffer.
template <typename T>
__global__ void cast_rows_from_fp32_kernel(T* __restrict__ dst, long long ld_dst, const float* __restrict__ src,
                                           long long rows, int cols) {
  const int vpr = cols >> 3;
  const long long total = rows * vpr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = static_cast<int>(i - r * vpr) * 8;
    const float4 lo = *reinterpret_cast<const float4*>(src + r * cols + c);
    const float4 hi = *reinterpret_cast<const float4*>(src + r * cols + c + 4);
    uint4 q;
    if constexpr (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) {
      q = make_uint4(pack_bf16x2(lo.x, lo.y), pack_bf16x2(lo.z, lo.w), pack_bf16x2(hi.x, hi.y), pack_bf16x2(hi.z, hi.w));
    } else {
      __half2 h0 = __floats2half2_rn(lo.x, lo.y), h1 = __floats2half2_rn(lo.z, lo.w), h2 = __floats2half2_rn(hi.x, hi.y),
              h3 = __floats2half2_rn(hi.z, hi.w);
      q = make_uint4(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2),
                     *reinterpret_cast<uint32_t*>(&h3));
    }
    *reinterpret_cast<uint4*>(dst + r * ld_dst + c) = q;
  }
}

int grid_for(long long work, int threads) {
  long long g = (work + threads - 1) / threads;
  long long cap = static_cast<long long>(b2host::num_sms()) * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}
int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
}  // namespace

#define DISPATCH_T(dtype, CALL)                                   \
  do {                                                            \
    if ((dtype) == B2_DTYPE_BF16) { using T = __nv_bfloat16; CALL; } \
    else { using T = __half; CALL; }                              \
  } while (0)

extern "C" int b2_bias_dropout_add(const void* x, const void* bias, const void* residual, void* out, long long n, int H,
                                   float p, unsigned long long seed, int dtype, void* stream) {
  if (n % 8 || H % 8) { b2host::set_error("bias_dropout_add: sizes must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long nvec = n / 8;
  DISPATCH_T(dtype, (bias_dropout_add_kernel<T><<<grid_for(nvec, 256), 256, 0, st>>>((const T*)x, (const T*)bias, (const T*)residual, (T*)out, nvec, H, p, seed, b2host::dropout_replay_ptr())));
  return check("bias_dropout_add");
}
extern "C" int b2_dropout_bwd(const void* dy, void* dx, long long n, float p, unsigned long long seed, int dtype, void* stream) {
  if (n % 8) { b2host::set_error("dropout_bwd: size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long nvec = n / 8;
  DISPATCH_T(dtype, (dropout_bwd_kernel<T><<<grid_for(nvec, 256), 256, 0, st>>>((const T*)dy, (T*)dx, nvec, p, seed, b2host::dropout_replay_ptr())));
  return check("dropout_bwd");
}
extern "C" int b2_bias_act_fwd(const void* x, const void* bias, void* out, long long n, int H, int mode, int dtype, void* stream) {
  if (n % 8 || H % 8) { b2host::set_error("bias_act: sizes must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long nvec = n / 8;
  DISPATCH_T(dtype, (bias_act_fwd_kernel<T><<<grid_for(nvec, 256), 256, 0, st>>>((const T*)x, (const T*)bias, (T*)out, nvec, H, mode)));
  return check("bias_act_fwd");
}
extern "C" int b2_bias_act_bwd(const void* dy, const void* x, const void* bias, void* dx, long long n, int H, int mode, int dtype, void* stream) {
  if (n % 8 || H % 8) { b2host::set_error("bias_act: sizes must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long nvec = n / 8;
  DISPATCH_T(dtype, (bias_act_bwd_kernel<T><<<grid_for(nvec, 256), 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)bias, (T*)dx, nvec, H, mode)));
  return check("bias_act_bwd");
}
extern "C" int b2_swiglu_fwd(const void* gu, void* out, long long rows, int F, int dtype, void* stream) {
  if (F % 8) { b2host::set_error("swiglu: F must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DISPATCH_T(dtype, (swiglu_fwd_kernel<T><<<grid_for(rows * (F / 8), 256), 256, 0, st>>>((const T*)gu, (T*)out, rows, F)));
  return check("swiglu_fwd");
}
extern "C" int b2_swiglu_bwd(const void* dy, const void* gu, void* dgu, long long rows, int F, int dtype, void* stream) {
  if (F % 8) { b2host::set_error("swiglu: F must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DISPATCH_T(dtype, (swiglu_bwd_kernel<T><<<grid_for(rows * (F / 8), 256), 256, 0, st>>>((const T*)dy, (const T*)gu, (T*)dgu, rows, F)));
  return check("swiglu_bwd");
}
extern "C" int b2_colsum(const void* dy, float* out, int rows, int N, long long ld, int dtype, void* stream) {
  if (N % 8 || ld % 8) { b2host::set_error("colsum: N and ld must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows_per_cta = 512;
  dim3 grid((N + 63) / 64, (rows + rows_per_cta - 1) / rows_per_cta);
  DISPATCH_T(dtype, (colsum_kernel<T><<<grid, 256, 0, st>>>((const T*)dy, out, rows, N, ld, rows_per_cta)));
  return check("colsum");
}
extern "C" int b2_rope(void* x, const float* cos_t, const float* sin_t, const int* pos, long long tokens, int nheads, int hd,
                       long long tok_stride, int head_stride, int backward, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long total = tokens * nheads * (hd / 2);
  DISPATCH_T(dtype, (rope_kernel<T><<<grid_for(total, 256), 256, 0, st>>>((T*)x, cos_t, sin_t, pos, tokens, nheads, hd, tok_stride, head_stride, backward ? -1.f : 1.f)));
  return check("rope");
}
extern "C" int b2_axpby(const void* a, const void* b, void* out, long long n, float alpha, float beta, int dtype, void* stream) {
  if (n % 8) { b2host::set_error("axpby: size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long nvec = n / 8;
  DISPATCH_T(dtype, (axpby_kernel<T><<<grid_for(nvec, 256), 256, 0, st>>>((const T*)a, (const T*)b, (T*)out, nvec, alpha, beta)));
  return check("axpby");
}
extern "C" int b2_accum_fp32(float* dst, const void* src, long long n, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DISPATCH_T(dtype, (accum_fp32_kernel<T><<<grid_for(n, 256), 256, 0, st>>>(dst, (const T*)src, n)));
  return check("accum_fp32");
}
extern "C" int b2_cast_rows_from_fp32(void* dst, long long ld_dst, const float* src, long long rows, int cols, int dtype,
                                      void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if ((cols % 8) || (ld_dst % 8)) { b2host::set_error("cast_rows_from_fp32: cols and ld_dst must be multiples of 8"); return -1; }
  DISPATCH_T(dtype, (cast_rows_from_fp32_kernel<T><<<grid_for(rows * (cols / 8), 256), 256, 0, st>>>((T*)dst, ld_dst, src, rows, cols)));
  return check("cast_rows_from_fp32");
}
extern "C" int b2_cast_from_fp32(void* dst, const float* src, long long n, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DISPATCH_T(dtype, (cast_from_fp32_kernel<T><<<grid_for(n, 256), 256, 0, st>>>((T*)dst, src, n)));
  return check("cast_from_fp32");
}
