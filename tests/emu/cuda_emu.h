// A small CPU emulation of the CUDA execution model, enough to run the SIMT kernels of csrc/softmax.cu, csrc/conv.cu and
// csrc/batchnorm.cu unchanged on a machine without a GPU (tests/test_kernel_emulation_cpu.py).
//
// Model: thread blocks run one after another; the threads of a block are real OS threads, so `__syncthreads()` is a
// barrier over the block, `__shared__` variables are function-level statics (one live block at a time) and dynamic shared
// memory is a per-block allocation.  Warp shuffles exchange values through a per-warp slot array between two warp-wide
// barriers - every lane of the warp must take part, which is what `*_sync(0xffffffff, ...)` demands on hardware too.
// What this checks: indexing, bounds, reductions, barrier placement, data races that a barrier was supposed to prevent
// (run under -fsanitize=thread if in doubt).  What it cannot check: performance, memory-model subtleties, launch limits.
#pragma once
#include <immintrin.h>
#include <pthread.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <mutex>
#include <vector>

struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};

namespace emu {
struct Warp {
  pthread_barrier_t bar;
  uint64_t slots[32];
  int lanes;
};
struct Block {
  pthread_barrier_t bar;
  std::vector<Warp> warps;
  std::vector<char> dyn;
};
inline Block*& cur() { static Block* b = nullptr; return b; }
inline std::mutex& atomic_mu() { static std::mutex m; return m; }
}  // namespace emu

inline thread_local dim3 threadIdx;
inline dim3 blockIdx, blockDim, gridDim;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define B2_DEVICE inline
#define B2_DYN_SMEM(type, name) type* name = reinterpret_cast<type*>(emu::cur()->dyn.data())

inline void __syncthreads() { pthread_barrier_wait(&emu::cur()->bar); }

template <typename T>
inline T __shfl_xor_sync(unsigned, T v, int lane_mask) {
  static_assert(sizeof(T) <= 8, "shuffle payload");
  const unsigned tid = threadIdx.x;
  emu::Warp& w = emu::cur()->warps[tid >> 5];
  const int lane = tid & 31;
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  w.slots[lane] = raw;
  pthread_barrier_wait(&w.bar);
  const int src = lane ^ lane_mask;
  uint64_t got = src < w.lanes ? w.slots[src] : raw;
  pthread_barrier_wait(&w.bar);
  T out;
  std::memcpy(&out, &got, sizeof(T));
  return out;
}

template <typename T>
inline T __shfl_sync(unsigned, T v, int src_lane) {
  const unsigned tid = threadIdx.x;
  emu::Warp& w = emu::cur()->warps[tid >> 5];
  const int lane = tid & 31;
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  w.slots[lane] = raw;
  pthread_barrier_wait(&w.bar);
  uint64_t got = (src_lane & 31) < w.lanes ? w.slots[src_lane & 31] : raw;
  pthread_barrier_wait(&w.bar);
  T out;
  std::memcpy(&out, &got, sizeof(T));
  return out;
}

inline double atomicAdd(double* p, double v) { std::lock_guard<std::mutex> g(emu::atomic_mu()); double o = *p; *p = o + v; return o; }
inline float atomicAdd(float* p, float v) { std::lock_guard<std::mutex> g(emu::atomic_mu()); float o = *p; *p = o + v; return o; }

inline float __expf(float x) { return std::exp(x); }
inline float __logf(float x) { return std::log(x); }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
using std::max;
using std::min;

// 16-bit float types (storage only; arithmetic happens in float like in the kernels)
struct __nv_bfloat16 { uint16_t v; };
struct __half { uint16_t v; };
inline float __bfloat162float(__nv_bfloat16 h) { uint32_t u = static_cast<uint32_t>(h.v) << 16; float f; std::memcpy(&f, &u, 4); return f; }
inline __nv_bfloat16 __float2bfloat16_rn(float f) {
  uint32_t u; std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return {static_cast<uint16_t>((u >> 16) | 0x40)};
  u += 0x7fffu + ((u >> 16) & 1u);
  return {static_cast<uint16_t>(u >> 16)};
}
inline float __half2float(__half h) { return _cvtsh_ss(h.v); }
inline __half __float2half_rn(float f) { return {static_cast<uint16_t>(_cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC))}; }

namespace emu {
struct ThreadArg { const std::function<void()>* fn; unsigned tx; };
inline void* thread_main(void* p) {
  ThreadArg* a = static_cast<ThreadArg*>(p);
  threadIdx = dim3(a->tx, 0, 0);
  (*a->fn)();
  return nullptr;
}

// kernel<<<grid, block, smem>>>(args...)  ->  emu::launch(grid, block, smem, [&] { kernel(args...); })   (1-D blocks)
inline void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& fn) {
  gridDim = grid;
  blockDim = block;
  const unsigned nt = block.x;
  for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        blockIdx = dim3(bx, by, bz);
        Block b;
        pthread_barrier_init(&b.bar, nullptr, nt);
        b.warps.resize((nt + 31) / 32);
        for (unsigned w = 0; w < b.warps.size(); ++w) {
          b.warps[w].lanes = static_cast<int>(std::min(32u, nt - 32 * w));
          pthread_barrier_init(&b.warps[w].bar, nullptr, b.warps[w].lanes);
        }
        b.dyn.assign(smem_bytes + 16, 0x7f);          // poison: kernels must not rely on zeroed shared memory
        cur() = &b;
        std::vector<pthread_t> th(nt);
        std::vector<ThreadArg> args(nt);
        for (unsigned t = 0; t < nt; ++t) {
          args[t] = {&fn, t};
          pthread_create(&th[t], nullptr, thread_main, &args[t]);
        }
        for (unsigned t = 0; t < nt; ++t) pthread_join(th[t], nullptr);
        pthread_barrier_destroy(&b.bar);
        for (auto& w : b.warps) pthread_barrier_destroy(&w.bar);
        cur() = nullptr;
      }
}
}  // namespace emu
