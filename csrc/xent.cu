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

// one CTA: (max, sum-exp relative to it, target logit or 0) of one row of the shard; valid in thread 0
template <typename T>
B2_DEVICE void row_stats(const T* __restrict__ lr, long long label, int V, long long vocab_start, float& M, float& S, float& Tg,
                         float* sm_m, float* sm_s) {
  float m = -INFINITY, s = 0.f;
  // four 16-byte loads in flight per thread (one in flight left the kernel at half of the HBM rate), then one online-softmax
  // update over the 32 values
  constexpr int U = 4;
  int v = threadIdx.x * 8;
  for (; v + (U - 1) * THREADS * 8 < V; v += U * THREADS * 8) {
    uint4 q[U];
#pragma unroll
    for (int u = 0; u < U; ++u) q[u] = *reinterpret_cast<const uint4*>(lr + v + u * THREADS * 8);
    float f[U][8];
    float bm = -INFINITY;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      load8f<T>(reinterpret_cast<const T*>(&q[u]), f[u]);
#pragma unroll
      for (int e = 0; e < 8; ++e) bm = fmaxf(bm, f[u][e]);
    }
    const float nm = fmaxf(m, bm);
    float acc = 0.f;
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc += __expf(f[u][e] - nm);
    s = s * __expf(m - nm) + acc;
    m = nm;
  }
  for (; v < V; v += THREADS * 8) {
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
  __syncthreads();                       // the previous row's readers of sm_m / sm_s are done
  if (lane == 0) { sm_m[warp] = m; sm_s[warp] = s; }
  __syncthreads();
  if (threadIdx.x == 0) {
    M = sm_m[0]; S = sm_s[0];
    for (int w = 1; w < THREADS / 32; ++w) {
      float nm = fmaxf(M, sm_m[w]);
      S = (M == -INFINITY ? 0.f : S * __expf(M - nm)) + (sm_m[w] == -INFINITY ? 0.f : sm_s[w] * __expf(sm_m[w] - nm));
      M = nm;
    }
    const long long lab = label - vocab_start;
    Tg = (lab >= 0 && lab < V) ? static_cast<float>(lr[lab]) : 0.f;
  }
}

template <typename T>
__global__ void __launch_bounds__(THREADS)
xent_stats_kernel(const T* __restrict__ logits, const long long* __restrict__ labels, float* __restrict__ rmax,
                  float* __restrict__ rsum, float* __restrict__ rtgt, int V, long long ld, long long vocab_start) {
  __shared__ float sm_m[THREADS / 32], sm_s[THREADS / 32];
  const long long row = blockIdx.x;
  float M, S, Tg;
  row_stats<T>(logits + row * ld, labels[row], V, vocab_start, M, S, Tg, sm_m, sm_s);
  if (threadIdx.x == 0) { rmax[row] = M; rsum[row] = S; rtgt[row] = Tg; }
}

// Vocab-parallel softmax cross-entropy forward in ONE kernel: every rank reduces its vocab shard to
// (max, sum-exp, target logit) per row, the ranks exchange those three scalars per row through a symmetric buffer, and the
// same kernel finishes the loss - no all-gather, no host-side combine.
//   phase 1  persistent CTAs walk the rows of the local logits shard and write the row statistics into this rank's
//            symmetric stats buffer  [call parity][3][rows]  (parity: a peer may still read the previous call's values)
//   barrier  last CTA of the grid (ticket): system fence, tell every peer, wait for every peer, open the local gate
//   phase 2  one thread per row reads all ranks' statistics over NVLink (3 floats per rank and row), combines them and
//            writes the global max / sum-exp (kept for the backward) and the loss
// The grid is at most one wave (the CTAs wait for each other); ctrl = control block of the stats buffer.
struct XentPeers { float* stats[8]; uint32_t* ctrl[8]; };

template <typename T>
__global__ void __launch_bounds__(THREADS)
xent_vocab_parallel_kernel(const T* __restrict__ logits, const long long* __restrict__ labels, XentPeers P, int rank, int world,
                           float* __restrict__ gmax, float* __restrict__ gsum, float* __restrict__ loss, long long rows, int V,
                           long long ld, long long vocab_start, long long ignore_index) {
  __shared__ float sm_m[THREADS / 32], sm_s[THREADS / 32];
  __shared__ bool last;
  uint32_t* my_ctrl = P.ctrl[rank];
  const uint32_t epoch = my_ctrl[1] + 1;           // stable at entry: the previous call on this buffer has completed
  float* mine = P.stats[rank] + static_cast<long long>(epoch & 1) * 3 * rows;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    float M, S, Tg;
    row_stats<T>(logits + row * ld, labels[row], V, vocab_start, M, S, Tg, sm_m, sm_s);
    if (threadIdx.x == 0) { mine[row] = M; mine[rows + row] = S; mine[2 * rows + row] = Tg; }
  }
  // ---- grid-wide + cross-GPU barrier
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(my_ctrl + 2, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    if (threadIdx.x < world && static_cast<int>(threadIdx.x) != rank) {
      __threadfence_system();
      st_release_sys(P.ctrl[threadIdx.x] + 16 + rank, epoch);
      wait_ge_sys(my_ctrl + 16 + threadIdx.x, epoch, 0x301);
    }
    __syncthreads();
    if (threadIdx.x == 0) { my_ctrl[2] = 0u; my_ctrl[1] = epoch; __threadfence(); st_release_gpu(my_ctrl + 0, epoch); }
  } else if (threadIdx.x == 0) {
    wait_ge_gpu(my_ctrl + 0, epoch, 0x302);
  }
  __syncthreads();
  // ---- combine + finish, one thread per row
  for (long long row = blockIdx.x * static_cast<long long>(THREADS) + threadIdx.x; row < rows; row += static_cast<long long>(gridDim.x) * THREADS) {
    float m[8], sv[8], t = 0.f, gm = -INFINITY;
    for (int r = 0; r < world; ++r) {
      const float* st = P.stats[r] + static_cast<long long>(epoch & 1) * 3 * rows;
      // peer memory was written by another GPU during this kernel: volatile loads keep them out of the read-only path
      m[r] = *reinterpret_cast<const volatile float*>(st + row);
      sv[r] = *reinterpret_cast<const volatile float*>(st + rows + row);
      t += *reinterpret_cast<const volatile float*>(st + 2 * rows + row);
      gm = fmaxf(gm, m[r]);
    }
    float gs = 0.f;
    for (int r = 0; r < world; ++r) gs += (m[r] == -INFINITY) ? 0.f : sv[r] * __expf(m[r] - gm);
    gmax[row] = gm;
    gsum[row] = gs;
    loss[row] = labels[row] == ignore_index ? 0.f : (__logf(gs) + gm - t);
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

// stats_ptrs[r] / ctrls[r]: rank r's symmetric stats buffer (2 * 3 * rows floats, zero-initialised control block);
// gmax / gsum / loss: [rows] fp32 outputs on this rank.  replaces Megatron's three all-reduces over [s, b].
extern "C" int b2_xent_vocab_parallel(const void* logits, const long long* labels, void* const* stats_ptrs, void* const* ctrls,
                                      int rank, int world, float* gmax, float* gsum, float* loss, long long rows, int V,
                                      long long ld, long long vocab_start, long long ignore_index, int dtype, void* stream) {
  if (V % 8 || ld % 8) { b2host::set_error("xent: vocab shard and row stride must be multiples of 8"); return -1; }
  if (world < 2 || world > 8) { b2host::set_error("xent_vocab_parallel: world size %d unsupported", world); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  XentPeers P;
  for (int r = 0; r < 8; ++r) { P.stats[r] = r < world ? static_cast<float*>(stats_ptrs[r]) : nullptr; P.ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr; }
  // one full wave: every CTA must be resident (they wait for each other at the barrier), and the memory-bound first phase
  // wants all the warps an SM can hold
  static int occ_bf16 = 0, occ_fp16 = 0;
  int& occ = dtype == B2_DTYPE_BF16 ? occ_bf16 : occ_fp16;
  if (occ == 0) {
    cudaError_t e = dtype == B2_DTYPE_BF16
                        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, xent_vocab_parallel_kernel<__nv_bfloat16>, THREADS, 0)
                        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, xent_vocab_parallel_kernel<__half>, THREADS, 0);
    if (e != cudaSuccess || occ < 1) occ = 1;
  }
  long long g = static_cast<long long>(b2host::num_sms()) * occ;
  if (g > rows) g = rows;
  const unsigned grid = static_cast<unsigned>(g < 1 ? 1 : g);
  if (dtype == B2_DTYPE_BF16)
    xent_vocab_parallel_kernel<__nv_bfloat16><<<grid, THREADS, 0, st>>>((const __nv_bfloat16*)logits, labels, P, rank, world, gmax, gsum, loss, rows, V, ld, vocab_start, ignore_index);
  else
    xent_vocab_parallel_kernel<__half><<<grid, THREADS, 0, st>>>((const __half*)logits, labels, P, rank, world, gmax, gsum, loss, rows, V, ld, vocab_start, ignore_index);
  return check("xent_vocab_parallel");
}

