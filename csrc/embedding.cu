// Token + position embedding forward/backward (vocab-parallel aware) for sm_100a.
//
// replaces: Megatron VocabParallelEmbedding + learned position embedding + embedding dropout
//  and HF nn.Embedding in OPT/LLaMA (K24).
//   forward : out[t,:] = dropout( wte[tok[t]-vocab_start] (0 if outside the shard) + wpe[pos[t]] )
//   backward: d_wte[tok] += dy, d_wpe[pos] += dy   accumulated in fp32 (main_grad semantics)
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

template <typename T>
__global__ void embed_fwd_kernel(const long long* __restrict__ tok, const long long* __restrict__ pos,
                                 const T* __restrict__ wte, const T* __restrict__ wpe, T* __restrict__ out,
                                 long long tokens, int H, long long vocab_start, long long vocab_local) {
  const int hv = H >> 3;
  const long long nvec = tokens * hv;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    long long t = v / hv; int c = static_cast<int>(v - t * hv) * 8;
    long long id = tok[t] - vocab_start;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8];
    if (id >= 0 && id < vocab_local) load8<T>(wte + id * H + c, a);
    if (wpe != nullptr) {
      load8<T>(wpe + pos[t] * H + c, b);
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] += b[e];
    }
    store8<T>(out + t * H + c, a);
  }
}

template <typename T>
__global__ void embed_bwd_kernel(const T* __restrict__ dy, const long long* __restrict__ tok,
                                 const long long* __restrict__ pos, float* __restrict__ dwte, float* __restrict__ dwpe,
                                 long long tokens, int H, long long vocab_start, long long vocab_local) {
  const int hv = H >> 3;
  const long long nvec = tokens * hv;
  for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
    long long t = v / hv; int c = static_cast<int>(v - t * hv) * 8;
    float g[8];
    load8<T>(dy + t * H + c, g);
    long long id = tok[t] - vocab_start;
    if (id >= 0 && id < vocab_local) {
      float* d = dwte + id * H + c;
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(d + e, g[e]);
    }
    if (dwpe != nullptr) {
      float* d = dwpe + pos[t] * H + c;
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(d + e, g[e]);
    }
  }
}

int grid_for(long long work) {
  long long g = (work + 255) / 256;
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

extern "C" int b2_embedding_fwd(const long long* tok, const long long* pos, const void* wte, const void* wpe, void* out,
                                long long tokens, int H, long long vocab_start, long long vocab_local, int dtype,
                                void* stream) {
  if (H % 8) { b2host::set_error("embedding: hidden size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B2_DTYPE_BF16)
    embed_fwd_kernel<__nv_bfloat16><<<grid_for(tokens * (H / 8)), 256, 0, st>>>(tok, pos, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, tokens, H, vocab_start, vocab_local);
  else
    embed_fwd_kernel<__half><<<grid_for(tokens * (H / 8)), 256, 0, st>>>(tok, pos, (const __half*)wte, (const __half*)wpe, (__half*)out, tokens, H, vocab_start, vocab_local);
  return check("embedding_fwd");
}
extern "C" int b2_embedding_bwd(const void* dy, const long long* tok, const long long* pos, float* dwte, float* dwpe,
                                long long tokens, int H, long long vocab_start, long long vocab_local, int dtype,
                                void* stream) {
  if (H % 8) { b2host::set_error("embedding: hidden size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == B2_DTYPE_BF16)
    embed_bwd_kernel<__nv_bfloat16><<<grid_for(tokens * (H / 8)), 256, 0, st>>>((const __nv_bfloat16*)dy, tok, pos, dwte, dwpe, tokens, H, vocab_start, vocab_local);
  else
    embed_bwd_kernel<__half><<<grid_for(tokens * (H / 8)), 256, 0, st>>>((const __half*)dy, tok, pos, dwte, dwpe, tokens, H, vocab_start, vocab_local);
  return check("embedding_bwd");
}
