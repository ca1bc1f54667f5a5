// NVLink transfer-primitive probe (2 GPUs, one process): which in-kernel copy pattern reaches link rate with how many CTAs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bench/probes/p2p_probe bench/probes/p2p_probe.cu
// Every kernel runs on BOTH GPUs at the same time (each pulls from / pushes to the other), like the fused TP kernels do.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t ph) {
  uint32_t ok; asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory"); return ok != 0; }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) { while (!mbar_try(b, ph)) {} }
__device__ __forceinline__ void bulk_g2s(void* s, const void* g, uint32_t n, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(s)), "l"(g), "r"(n), "r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void bulk_s2g(void* g, const void* s, uint32_t n) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(n) : "memory"); }
__device__ __forceinline__ void commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

// one thread per CTA: src -> smem ring -> dst, pieces c, c + n, ...  (store=0: loads only)
__global__ void tma_copy_kernel(const uint8_t* src, uint8_t* dst, long long bytes, int piece, int stg, int store) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + (size_t)stg * piece);
  if (threadIdx.x != 0) return;
  for (int i = 0; i < stg; ++i) mbar_init(&bars[i], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const long long np = (bytes + piece - 1) / piece;
  long long ld = blockIdx.x, st = blockIdx.x;
  uint32_t issued = 0, stored = 0;
  while (st < np) {
    while (ld < np && issued - stored < (uint32_t)stg) {
      const uint32_t s = issued % stg;
      if (store && issued >= (uint32_t)stg) { if (issued - stg == stored - 1) wait_read<0>(); else wait_read<1>(); }   // the store that last used this slot has been read out
      const long long off = ld * piece;
      const uint32_t sz = (uint32_t)((bytes - off) < piece ? (bytes - off) : piece);
      mbar_expect_tx(&bars[s], sz);
      bulk_g2s(sm + (size_t)s * piece, src + off, sz, &bars[s]);
      ++issued; ld += gridDim.x;
    }
    const uint32_t s = stored % stg;
    mbar_wait(&bars[s], (stored / stg) & 1);
    if (store) {
      const long long off = st * piece;
      const uint32_t sz = (uint32_t)((bytes - off) < piece ? (bytes - off) : piece);
      bulk_s2g(dst + off, sm + (size_t)s * piece, sz);
      commit();
    }
    ++stored; st += gridDim.x;
  }
  if (store) wait_all<0>();
}

// all threads: 16-byte loads (U in flight per thread) then stores
template <int U>
__global__ void __launch_bounds__(512) ldst_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long nvec) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (; v + (U - 1) * stride < nvec; v += U * stride) {
    uint4 q[U];
#pragma unroll
    for (int j = 0; j < U; ++j) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q[j].x), "=r"(q[j].y), "=r"(q[j].z), "=r"(q[j].w) : "l"(src + v + j * stride));
#pragma unroll
    for (int j = 0; j < U; ++j) dst[v + j * stride] = q[j];
  }
  for (; v < nvec; v += stride) dst[v] = src[v];
}

// flag ping-pong: dev0 writes k to peer flag, waits for own flag == k, ...
__global__ void pingpong_kernel(volatile uint32_t* mine, uint32_t* peer, int iters, int first, unsigned long long* out_ns) {
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (int k = 1; k <= iters; ++k) {
    if (first) {
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(k) : "memory");
      unsigned long long s; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(s));
      while (true) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory"); if ((int)(v - k) >= 0) break;
        unsigned long long n; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(n)); if (n - s > 2000000000ull) { *out_ns = ~0ull; return; } }
    } else {
      unsigned long long s; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(s));
      while (true) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory"); if ((int)(v - k) >= 0) break;
        unsigned long long n; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(n)); if (n - s > 2000000000ull) { *out_ns = ~0ull; return; } }
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peer), "r"(k) : "memory");
    }
  }
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  *out_ns = t1 - t0;
}

struct Dev { uint8_t *a, *b; cudaStream_t st; cudaEvent_t e0, e1; };
Dev D[2];
template <typename F> double time_both(F launch, int iters = 20) {
  for (int w = 0; w < 3; ++w) for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); launch(d); }
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(D[d].st)); }
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(D[d].e0, D[d].st)); }
  for (int i = 0; i < iters; ++i) for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); launch(d); }
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(D[d].e1, D[d].st)); }
  double worst = 0;
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(D[d].st)); CK(cudaGetLastError()); float ms; CK(cudaEventElapsedTime(&ms, D[d].e0, D[d].e1)); if (ms / iters > worst) worst = ms / iters; }
  return worst;
}

int main() {
  int n = 0; CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("need 2 GPUs\n"); return 2; }
  const long long MAXB = 64ll << 20;
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    int can = 0; CK(cudaDeviceCanAccessPeer(&can, d, 1 - d)); if (!can) { printf("no peer access\n"); return 2; }
    CK(cudaDeviceEnablePeerAccess(1 - d, 0));
    CK(cudaMalloc(&D[d].a, MAXB)); CK(cudaMalloc(&D[d].b, MAXB)); CK(cudaMemset(D[d].a, d + 1, MAXB)); CK(cudaMemset(D[d].b, 0, MAXB));
    CK(cudaStreamCreate(&D[d].st)); CK(cudaEventCreate(&D[d].e0)); CK(cudaEventCreate(&D[d].e1));
    CK(cudaFuncSetAttribute(tma_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  }
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
  const long long sizes[2] = {6144ll * 768 * 2, 12288ll * 768 * 2};
  for (long long S : sizes) {
    printf("== %.1f MB per direction, both directions at once\n", S / 1e6);
    { double ms = time_both([&](int d) { CK(cudaMemcpyPeerAsync(D[d].b, d, D[1 - d].a, 1 - d, S, D[d].st)); });
      printf("memcpyPeerAsync (DMA pull)            : %7.1f us  %6.0f GB/s\n", ms * 1e3, S / ms / 1e6); }
    for (int mode = 0; mode < 4; ++mode) {       // 0 pull, 1 pull loads only, 2 push, 3 local copy
      const char* names[4] = {"tma pull (remote->smem->local)", "tma pull, loads only", "tma push (local->smem->remote)", "tma local copy"};
      for (int ctas : {8, 16, 24, 32, 48}) for (int piece : {16384, 32768, 65536}) {
        const int stg = (216 * 1024) / piece;
        double ms = time_both([&](int d) {
          const uint8_t* src = (mode == 2 || mode == 3) ? D[d].a : D[1 - d].a;
          uint8_t* dst = (mode == 2) ? D[1 - d].b : D[d].b;
          tma_copy_kernel<<<ctas, 32, stg * piece + 256, D[d].st>>>(src, dst, S, piece, stg, mode != 1);
        });
        printf("%-32s ctas %2d piece %2dK stg %2d: %7.1f us  %6.0f GB/s\n", names[mode], ctas, piece >> 10, stg, ms * 1e3, S / ms / 1e6);
      }
    }
    for (int mode = 0; mode < 2; ++mode) for (int ctas : {8, 16, 24, 32, 48, 148}) {
      double ms8 = time_both([&](int d) { ldst_copy_kernel<8><<<ctas, 512, 0, D[d].st>>>((const uint4*)(mode ? D[d].a : D[1 - d].a), (uint4*)(mode ? D[1 - d].b : D[d].b), S / 16); });
      double ms16 = time_both([&](int d) { ldst_copy_kernel<16><<<ctas, 512, 0, D[d].st>>>((const uint4*)(mode ? D[d].a : D[1 - d].a), (uint4*)(mode ? D[1 - d].b : D[d].b), S / 16); });
      printf("%-32s ctas %3d x 512 thr: U8 %7.1f us %6.0f GB/s | U16 %7.1f us %6.0f GB/s\n", mode ? "ld/st push (local ld, remote st)" : "ld/st pull (remote ld, local st)", ctas, ms8 * 1e3, S / ms8 / 1e6, ms16 * 1e3, S / ms16 / 1e6);
    }
  }
  // flag round trip
  uint32_t* fl[2]; unsigned long long* ns[2];
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaMalloc(&fl[d], 256)); CK(cudaMemset(fl[d], 0, 256)); CK(cudaMalloc(&ns[d], 8)); CK(cudaDeviceSynchronize()); }
  const int iters = 200;
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); pingpong_kernel<<<1, 1, 0, D[d].st>>>(fl[d], fl[1 - d], iters, d == 0, ns[d]); }
  for (int d = 0; d < 2; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(D[d].st)); }
  unsigned long long h; CK(cudaSetDevice(0)); CK(cudaMemcpy(&h, ns[0], 8, cudaMemcpyDeviceToHost));
  printf("flag round trip (st.release.sys -> peer spin ld.acquire.sys -> back): %.2f us\n", h == ~0ull ? -1.0 : h / 1e3 / iters);
  printf("probe done\n");
  return 0;
}
