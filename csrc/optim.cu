// Optimizer-side kernels over FLAT buffers (sm_100a, memory-bound, 128-bit accesses).
//
// The framework allocates parameters, gradients and optimizer state as contiguous arenas per dtype,
// so a whole model updates in one launch - there is no multi-tensor-apply chunk table.
//
// replaces: apex amp_C.multi_tensor_adam / multi_tensor_l2norm / multi_tensor_scale (SURVEY E25, K21,
// K22), DeepSpeed csrc/adam/multi_tensor_adam.cu (E30, K25), ATen foreach Adam (K10) and Adadelta (K3).
//
// All step-dependent scalars that are produced on the device (gradient sum-of-squares for clipping,
// found-inf flag of the fp16 loss scaler) are READ on the device, so an optimizer step never needs a
// host synchronisation and can be captured in a CUDA graph.
#include "common.cuh"
#include "b200dist.h"
#include <type_traits>

using namespace b2;

namespace {

constexpr int THREADS = 256;

B2_DEVICE float clip_coef(const float* sumsq, float max_norm) {
  if (sumsq == nullptr || max_norm <= 0.f) return 1.f;
  float nrm = sqrtf(*sumsq);
  float c = max_norm / (nrm + 1e-6f);
  return c < 1.f ? c : 1.f;
}

// AdamW / Adam(L2) on fp32 master weights; optionally emits the low-precision model copy.
template <typename PT, typename GT>
__global__ void __launch_bounds__(THREADS)
adam_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const GT* __restrict__ g,
            PT* __restrict__ p_lp, long long n, float lr, float beta1, float beta2, float eps, float wd, float bc1,
            float bc2, int adamw, float grad_scale, const float* __restrict__ sumsq, float max_norm,
            const float* __restrict__ found_inf) {
  if (found_inf != nullptr && *found_inf != 0.f) return;  // fp16 overflow: skip the step entirely
  const float gs = grad_scale * clip_coef(sumsq, max_norm);
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float gg[4];
    if constexpr (sizeof(GT) == 4) {
      float4 t = reinterpret_cast<const float4*>(g)[i];
      gg[0] = t.x; gg[1] = t.y; gg[2] = t.z; gg[3] = t.w;
    } else {
      uint2 t = reinterpret_cast<const uint2*>(g)[i];
      float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y);
      gg[0] = a.x; gg[1] = a.y; gg[2] = b.x; gg[3] = b.y;
    }
    float pa[4] = {pp.x, pp.y, pp.z, pp.w}, ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float gr = gg[e] * gs;
      if (!adamw) gr += wd * pa[e];
      ma[e] = beta1 * ma[e] + (1.f - beta1) * gr;
      va[e] = beta2 * va[e] + (1.f - beta2) * gr * gr;
      float upd = (ma[e] / bc1) / (sqrtf(va[e] / bc2) + eps);
      if (adamw) upd += wd * pa[e];
      pa[e] -= lr * upd;
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (p_lp != nullptr) {
      if constexpr (sizeof(PT) == 2) {
        uint2 o;
        if constexpr (std::is_same<PT, __nv_bfloat16>::value) { o.x = pack_bf16x2(pa[0], pa[1]); o.y = pack_bf16x2(pa[2], pa[3]); }
        else { __half2 h0 = __floats2half2_rn(pa[0], pa[1]), h1 = __floats2half2_rn(pa[2], pa[3]);
               o.x = *reinterpret_cast<uint32_t*>(&h0); o.y = *reinterpret_cast<uint32_t*>(&h1); }
        reinterpret_cast<uint2*>(p_lp)[i] = o;
      } else {
        reinterpret_cast<float4*>(p_lp)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
      }
    }
  }
  // tail (n % 4)
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    long long i = (n4 << 2) + threadIdx.x;
    float gr = static_cast<float>(g[i]) * gs;
    float pa = p[i], ma = m[i], va = v[i];
    if (!adamw) gr += wd * pa;
    ma = beta1 * ma + (1.f - beta1) * gr;
    va = beta2 * va + (1.f - beta2) * gr * gr;
    float upd = (ma / bc1) / (sqrtf(va / bc2) + eps);
    if (adamw) upd += wd * pa;
    pa -= lr * upd;
    p[i] = pa; m[i] = ma; v[i] = va;
    if (p_lp != nullptr) p_lp[i] = static_cast<PT>(pa);
  }
}

// Adadelta (torch semantics): sq = rho sq + (1-rho) g^2 ; d = sqrt(acc+eps)/sqrt(sq+eps) g ; acc = rho acc + (1-rho) d^2 ; p -= lr d
__global__ void __launch_bounds__(THREADS)
adadelta_kernel(float* __restrict__ p, float* __restrict__ sq, float* __restrict__ acc, const float* __restrict__ g,
                long long n, float lr, float rho, float eps, float wd) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gr = g[i] + wd * p[i];
    float s = rho * sq[i] + (1.f - rho) * gr * gr;
    float d = sqrtf(acc[i] + eps) / sqrtf(s + eps) * gr;
    acc[i] = rho * acc[i] + (1.f - rho) * d * d;
    sq[i] = s;
    p[i] -= lr * d;
  }
}

// SGD with momentum (Megatron --optimizer sgd)
__global__ void __launch_bounds__(THREADS)
sgd_kernel(float* __restrict__ p, float* __restrict__ mom, const float* __restrict__ g, long long n, float lr,
           float momentum, float wd, float grad_scale, int first_step) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gr = g[i] * grad_scale + wd * p[i];
    float b = first_step ? gr : momentum * mom[i] + gr;
    mom[i] = b;
    p[i] -= lr * (momentum != 0.f ? b : gr);
  }
}

// sum of squares of a flat buffer: per-CTA partials into `scratch`, the last CTA to finish (ticket in
// scratch_counter, which must be 0 on entry and is reset to 0 on exit) folds them in a fixed order,
// so the result is deterministic.  out[0] (+)= sum.
template <typename T>
__global__ void __launch_bounds__(THREADS)
sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ scratch, unsigned int* __restrict__ counter,
             float* __restrict__ out, int accumulate) {
  __shared__ float red[THREADS / 32];
  __shared__ bool is_last;
  float s = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float f = static_cast<float>(x[i]);
    s += f * f;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) t += red[w];
    scratch[blockIdx.x] = t;
    __threadfence();
    unsigned int ticket = atomicAdd(counter, 1u);
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float t = 0.f;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) t += reinterpret_cast<volatile float*>(scratch)[i];
    t = warp_sum(t);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < THREADS / 32; ++w) tot += red[w];
      if (accumulate) *out += tot; else *out = tot;
      *counter = 0u;
    }
  }
}

// x *= scale ; found_inf |= any(!isfinite(x))      (fp16 loss-scaler unscale + overflow check)
__global__ void __launch_bounds__(THREADS)
scale_check_kernel(float* __restrict__ x, long long n, float scale, float* __restrict__ found_inf) {
  bool bad = false;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float f = x[i] * scale;
    if (!isfinite(f)) bad = true;
    x[i] = f;
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0 && found_inf) *found_inf = 1.f;
}

int grid_for(long long work) {
  long long g = (work + THREADS - 1) / THREADS;
  long long cap = static_cast<long long>(b2host::num_sms()) * 8;
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

extern "C" int b2_adam(float* p, float* m, float* v, const void* g, void* p_lp, long long n, float lr, float beta1,
                       float beta2, float eps, float wd, int step, int adamw, float grad_scale, const float* sumsq,
                       float max_norm, const float* found_inf, int grad_dtype, int lp_dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2 = 1.f - powf(beta2, static_cast<float>(step));
  const int grid = grid_for((n + 3) / 4);
#define ADAM(PT, GT) adam_kernel<PT, GT><<<grid, THREADS, 0, st>>>(p, m, v, (const GT*)g, (PT*)p_lp, n, lr, beta1, beta2, eps, wd, bc1, bc2, adamw, grad_scale, sumsq, max_norm, found_inf)
  if (grad_dtype == B2_DTYPE_FP32) {
    if (lp_dtype == B2_DTYPE_BF16) ADAM(__nv_bfloat16, float);
    else if (lp_dtype == B2_DTYPE_FP16) ADAM(__half, float);
    else ADAM(float, float);
  } else if (grad_dtype == B2_DTYPE_BF16) {
    if (lp_dtype == B2_DTYPE_BF16) ADAM(__nv_bfloat16, __nv_bfloat16);
    else if (lp_dtype == B2_DTYPE_FP16) ADAM(__half, __nv_bfloat16);
    else ADAM(float, __nv_bfloat16);
  } else { b2host::set_error("adam: gradient dtype must be fp32 or bf16"); return -1; }
#undef ADAM
  return check("adam");
}

extern "C" int b2_adadelta(float* p, float* sq, float* acc, const float* g, long long n, float lr, float rho, float eps,
                           float wd, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  adadelta_kernel<<<grid_for(n), THREADS, 0, st>>>(p, sq, acc, g, n, lr, rho, eps, wd);
  return check("adadelta");
}
extern "C" int b2_sgd(float* p, float* mom, const float* g, long long n, float lr, float momentum, float wd,
                      float grad_scale, int first_step, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  sgd_kernel<<<grid_for(n), THREADS, 0, st>>>(p, mom, g, n, lr, momentum, wd, grad_scale, first_step);
  return check("sgd");
}

extern "C" int b2_sumsq_scratch_floats(void) { return 1024; }

// scratch: >= b2_sumsq_scratch_floats() floats; counter: one zero-initialised uint32 (left zero on exit)
extern "C" int b2_sumsq(const void* x, long long n, float* scratch, unsigned int* counter, float* out, int accumulate,
                        int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int grid = grid_for(n);
  if (grid > 1024) grid = 1024;
  if (dtype == B2_DTYPE_FP32) sumsq_kernel<float><<<grid, THREADS, 0, st>>>((const float*)x, n, scratch, counter, out, accumulate);
  else if (dtype == B2_DTYPE_BF16) sumsq_kernel<__nv_bfloat16><<<grid, THREADS, 0, st>>>((const __nv_bfloat16*)x, n, scratch, counter, out, accumulate);
  else sumsq_kernel<__half><<<grid, THREADS, 0, st>>>((const __half*)x, n, scratch, counter, out, accumulate);
  return check("sumsq");
}
extern "C" int b2_scale_check(float* x, long long n, float scale, float* found_inf, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  scale_check_kernel<<<grid_for(n), THREADS, 0, st>>>(x, n, scale, found_inf);
  return check("scale_check");
}
