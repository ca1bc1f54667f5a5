// Softmax cross-entropy over (optionally vocab-sharded) logits, fp32 math on bf16/fp16 logits.
//
// replaces: Megatron tensor_parallel.cross_entropy._VocabParallelCrossEntropy,
// ATen CrossEntropyLoss (K9) and HF's shifted CE with ignore_index -100 (K24).
//
// Protocol (vocab shard = columns [vocab_start, vocab_start + V) of the global vocabulary):
//   1. b2_xent_stats:  per row, ONE pass over the shard -> local max, local sum-exp (relative to the
//                      local max) and the target logit if the label falls in this shard (else 0).
//   2. (TP > 1) the three [rows] fp32 vectors are combined across the TP group (max, then rescaled
//      sums) - a 3-scalar-per-row exchange instead of three separate all-reduces over logits.
//   3. b2_xent_finish: loss[row] = log(sumexp) + max - target (0 where label == ignore_index).
//   4. b2_xent_bwd:    dlogits = (softmax - onehot) * row_scale, written in place over the logits.
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {
constexpr int THREADS = 256;

template <typename T> B2_DEVICE void load8f(const T* p, float* f);
template <> B2_DEVICE void load8f<__nv_bfloat16>(const __nv_bfloat16* p, float* f) {
  uint4 q = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16x2(q.x), b = unpack_bf16x2(q.y), c = unpack_bf16x2(q.z), d = unpack_bf16x2(q.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <> B2_DEVICE void load8f<__half>(const __half* p, float* f) {
  uint4 q = *reinterpret_cast<const uint4*>(p);
  const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
template <typename T> B2_DEVICE void store8f(T* p, const float* f);
template <> B2_DEVICE void store8f<__nv_bfloat16>(__nv_bfloat16* p, const float* f) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
template <> B2_DEVICE void store8f<__half>(__half* p, const float* f) {
  uint4 q; __half2* h = reinterpret_cast<__half2*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = q;
}

template <typename T>
__global__ void __launch_bounds__(THREADS)
xent_stats_kernel(const T* __restrict__ logits, const long long* __restrict__ labels, float* __restrict__ rmax,
                  float* __restrict__ rsum, float* __restrict__ rtgt, int V, long long ld, long long vocab_start) {
  __shared__ float sm_m[THREADS / 32], sm_s[THREADS / 32];
  const long long row = blockIdx.x;
  const T* lr = logits + row * ld;
  float m = -INFINITY, s = 0.f;
  for (int v = threadIdx.x * 8; v < V; v += THREADS * 8) {
    float f[8];
    load8f<T>(lr + v, f);
    float bm = f[0];
#pragma unroll
    for (int e = 1; e < 8; ++e) bm = fmaxf(bm, f[e]);
    float nm = fmaxf(m, bm);
    float acc = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) acc += __expf(f[e] - nm);
    s = s * __expf(m - nm) + acc;
    m = nm;
  }
  // warp then block combine of (m, s)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    float nm = fmaxf(m, om);
    s = (m == -INFINITY ? 0.f : s * __expf(m - nm)) + (om == -INFINITY ? 0.f : os * __expf(om - nm));
    m = nm;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { sm_m[warp] = m; sm_s[warp] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = sm_m[0], S = sm_s[0];
    for (int w = 1; w < THREADS / 32; ++w) {
      float nm = fmaxf(M, sm_m[w]);
      S = (M == -INFINITY ? 0.f : S * __expf(M - nm)) + (sm_m[w] == -INFINITY ? 0.f : sm_s[w] * __expf(sm_m[w] - nm));
      M = nm;
    }
    rmax[row] = M;
    rsum[row] = S;
    long long lab = labels[row] - vocab_start;
    rtgt[row] = (lab >= 0 && lab < V) ? static_cast<float>(lr[lab]) : 0.f;
  }
}

__global__ void xent_finish_kernel(const float* __restrict__ rmax, const float* __restrict__ rsum,
                                   const float* __restrict__ rtgt, const long long* __restrict__ labels,
                                   float* __restrict__ loss, long long rows, long long ignore_index) {
  long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (r >= rows) return;
  loss[r] = labels[r] == ignore_index ? 0.f : (__logf(rsum[r]) + rmax[r] - rtgt[r]);
}

// dlogits[r, v] = (exp(x - max) / sumexp - [v == label]) * scale[r]   (in place allowed)
template <typename T>
__global__ void __launch_bounds__(THREADS)
xent_bwd_kernel(const T* __restrict__ logits, T* __restrict__ dlogits, const long long* __restrict__ labels,
                const float* __restrict__ gmax, const float* __restrict__ gsum, const float* __restrict__ row_scale,
                float scalar_scale, int V, long long ld, long long vocab_start, long long ignore_index) {
  const long long row = blockIdx.x;
  const T* lr = logits + row * ld;
  T* dr = dlogits + row * ld;
  const long long lab_g = labels[row];
  const float sc = (lab_g == ignore_index ? 0.f : 1.f) * (row_scale ? row_scale[row] : 1.f) * scalar_scale;
  const float m = gmax[row], inv = 1.f / gsum[row];
  const long long lab = lab_g - vocab_start;
  for (int v = threadIdx.x * 8; v < V; v += THREADS * 8) {
    float f[8], o[8];
    load8f<T>(lr + v, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float p = __expf(f[e] - m) * inv;
      o[e] = (p - ((v + e) == lab ? 1.f : 0.f)) * sc;
    }
    store8f<T>(dr + v, o);
  }
}

int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
}  // namespace

extern "C" int b2_xent_stats(const void* logits, const long long* labels, float* rmax, float* rsum, float* rtgt,
                             long long rows, int V, long long ld, long long vocab_start, int dtype, void* stream) {
  if (V % 8 || ld % 8) { b2host::set_error("xent: vocab shard and row stride must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B2_DTYPE_BF16)
    xent_stats_kernel<__nv_bfloat16><<<static_cast<unsigned>(rows), THREADS, 0, st>>>((const __nv_bfloat16*)logits, labels, rmax, rsum, rtgt, V, ld, vocab_start);
  else
    xent_stats_kernel<__half><<<static_cast<unsigned>(rows), THREADS, 0, st>>>((const __half*)logits, labels, rmax, rsum, rtgt, V, ld, vocab_start);
  return check("xent_stats");
}
extern "C" int b2_xent_finish(const float* rmax, const float* rsum, const float* rtgt, const long long* labels,
                              float* loss, long long rows, long long ignore_index, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  xent_finish_kernel<<<static_cast<unsigned>((rows + 255) / 256), 256, 0, st>>>(rmax, rsum, rtgt, labels, loss, rows, ignore_index);
  return check("xent_finish");
}
extern "C" int b2_xent_bwd(const void* logits, void* dlogits, const long long* labels, const float* gmax,
                           const float* gsum, const float* row_scale, float scalar_scale, long long rows, int V,
                           long long ld, long long vocab_start, long long ignore_index, int dtype, void* stream) {
  if (V % 8 || ld % 8) { b2host::set_error("xent: vocab shard and row stride must be multiples of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B2_DTYPE_BF16)
    xent_bwd_kernel<__nv_bfloat16><<<static_cast<unsigned>(rows), THREADS, 0, st>>>((const __nv_bfloat16*)logits, (__nv_bfloat16*)dlogits, labels, gmax, gsum, row_scale, scalar_scale, V, ld, vocab_start, ignore_index);
  else
    xent_bwd_kernel<__half><<<static_cast<unsigned>(rows), THREADS, 0, st>>>((const __half*)logits, (__half*)dlogits, labels, gmax, gsum, row_scale, scalar_scale, V, ld, vocab_start, ignore_index);
  return check("xent_bwd");
}
