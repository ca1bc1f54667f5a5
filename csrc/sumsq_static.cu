// b2_sumsq_static: sum of squares of a flat buffer WITHOUT any caller- or library-side allocation.
//
// Who needs this: callers that must not (or cannot) obtain device memory at the point where they want a
// gradient / activation norm - a monitoring or watchdog thread in a training job whose framework allocator
// already owns nearly all of HBM (a fresh cudaMalloc can fail there, and cudaMalloc / cudaFree also stall
// every stream on the device while they run), C or C++ hosts that link this library without torch and have
// no workspace pool of their own, and latency-sensitive serving loops that freeze their memory map after
// start-up.  b2_sumsq() asks such callers for a scratch buffer and a zeroed counter; this variant brings
// its own: a fixed, file-scope __device__ array that every GPU which loads the library keeps until exit.
//
// The math is the same two-phase reduction as b2_sumsq (csrc/optim.cu): per-CTA partial sums, then the
// last CTA to finish (ticket counter) folds the partials in a fixed order, so the result is deterministic.
#include "common.cuh"
#include "b200dist.h"
#include <mutex>

using namespace b2;

namespace {

constexpr int THREADS = 256;
constexpr int MAX_CTAS = 1024;                 // one fp32 partial per CTA
constexpr int COUNTER_SLOT = MAX_CTAS;         // the ticket counter lives right after the partials
constexpr int SCRATCH_FLOATS = MAX_CTAS + 64;  // 4352 bytes per GPU

// The scratch itself.  In host code this NAME is only a host-side handle for the symbol - it is not a
// device address, and handing it to a kernel would make the kernel fault (or, on systems where the GPU
// can reach host memory, quietly use the wrong memory).  Host code below therefore asks the runtime for
// the real device address with cudaGetSymbolAddress() and passes that pointer to the kernel.
static __device__ __align__(256) float g_sumsq_scratch[SCRATCH_FLOATS];

template <typename T>
__global__ void __launch_bounds__(THREADS)
sumsq_static_kernel(const T* __restrict__ x, long long n, float* __restrict__ scratch, float* __restrict__ out,
                    int accumulate) {
  __shared__ float red[THREADS / 32];
  __shared__ bool is_last;
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + COUNTER_SLOT);
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

// Per-device set-up state: the resolved device address of g_sumsq_scratch on that device.
constexpr int MAX_DEVICES = 64;
float* g_dev_ptr[MAX_DEVICES] = {nullptr};
std::mutex g_setup_mu;

// Resolves (once per device, with that device current) the device address of the static scratch.
int resolve_scratch(float** out_ptr) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { b2host::set_error("sumsq_static: cudaGetDevice: %s", cudaGetErrorString(e)); return -1; }
  if (dev < 0 || dev >= MAX_DEVICES) { b2host::set_error("sumsq_static: device index %d out of range", dev); return -1; }
  std::lock_guard<std::mutex> lock(g_setup_mu);
  if (g_dev_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_sumsq_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      b2host::set_error("sumsq_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return -1;
    }
    g_dev_ptr[dev] = ptr;
  }
  *out_ptr = g_dev_ptr[dev];
  return 0;
}

}  // namespace

extern "C" long long b2_sumsq_static_max_elems(void) { return B2_SUMSQ_STATIC_MAX_ELEMS; }
extern "C" int b2_sumsq_static_scratch_bytes(void) { return SCRATCH_FLOATS * static_cast<int>(sizeof(float)); }

extern "C" int b2_sumsq_static(const void* x, long long n, float* out, int accumulate, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (x == nullptr || out == nullptr) { b2host::set_error("sumsq_static: null input or output pointer"); return -1; }
  if (n < 0 || n > B2_SUMSQ_STATIC_MAX_ELEMS) {
    b2host::set_error("sumsq_static: n = %lld is outside [0, %lld]; split the buffer and accumulate, or use b2_sumsq with "
                      "your own scratch", n, static_cast<long long>(B2_SUMSQ_STATIC_MAX_ELEMS));
    return -1;
  }
  if (dtype != B2_DTYPE_FP32 && dtype != B2_DTYPE_BF16 && dtype != B2_DTYPE_FP16) {
    b2host::set_error("sumsq_static: unsupported dtype %d", dtype);
    return -1;
  }
  float* scratch = nullptr;
  if (resolve_scratch(&scratch)) return -1;
  // The array keeps whatever the previous call left behind.  The partial slots are overwritten before they
  // are read; the ticket counter is the one word that has to start at zero, so clear it on the caller's
  // stream (a 4-byte memset, no allocation) - this also repairs it if an earlier launch was torn down midway.
  cudaError_t e = cudaMemsetAsync(scratch + COUNTER_SLOT, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) { b2host::set_error("sumsq_static: cudaMemsetAsync: %s", cudaGetErrorString(e)); return -1; }
  long long want = (n + THREADS - 1) / THREADS;
  long long cap = static_cast<long long>(b2host::num_sms()) * 8;
  if (want > cap) want = cap;
  if (want > MAX_CTAS) want = MAX_CTAS;
  if (want < 1) want = 1;
  const int grid = static_cast<int>(want);
  if (dtype == B2_DTYPE_FP32)
    sumsq_static_kernel<float><<<grid, THREADS, 0, st>>>(static_cast<const float*>(x), n, scratch, out, accumulate);
  else if (dtype == B2_DTYPE_BF16)
    sumsq_static_kernel<__nv_bfloat16><<<grid, THREADS, 0, st>>>(static_cast<const __nv_bfloat16*>(x), n, scratch, out, accumulate);
  else
    sumsq_static_kernel<__half><<<grid, THREADS, 0, st>>>(static_cast<const __half*>(x), n, scratch, out, accumulate);
  e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("sumsq_static launch: %s", cudaGetErrorString(e)); return -1; }
  return 0;
}
