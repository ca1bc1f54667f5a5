// In-kernel collectives over NVLink 5 / NVSwitch (sm_100a): no NCCL on these paths.
//
// Every buffer that takes part is allocated as SYMMETRIC memory (same size on every rank, peer-mapped,
// and - when the switch supports it - bound to one multicast address).  Each symmetric buffer carries its
// own small control block (B2_CTRL_WORDS uint32 right after the payload): a call counter and one arrival
// slot per rank, written with system-scope release stores and polled with acquire loads, so consecutive
// kernels on a buffer (also when replayed from a CUDA graph) stay in lock-step without the host.
//
//   b2_allreduce : sum over ranks, scaled, in place:   multimem.ld_reduce (the switch adds) -> * scale ->
//                  multimem.st (the switch broadcasts); each rank owns 1/world of the vector.  Falls back to
//                  peer loads/stores when no multicast address exists.
//                  replaces: NCCL all-reduce of DDP buckets / the Megatron grad buffer + the separate
//                  div-by-world pass, fused as scale in the same kernel.
//   b2_allgather_pull / b2_signal_* : building blocks of the fused all-gather->GEMM (see gemm.cu).
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int MAX_RANKS = 8;

struct Peers {
  void* ptr[MAX_RANKS];        // payload base of every rank (peer-mapped virtual addresses)
  uint32_t* ctrl[MAX_RANKS];   // control block of every rank
};

// ctrl layout (uint32): [0..2] see enter(); [16 + r] arrival slot of rank r for phase A, [32 + r] phase B
constexpr int CTRL_A = 16, CTRL_B = 32;

// All ranks arrive and wait.  One CTA-wide call; `epoch` must be identical on all ranks.
B2_DEVICE void fabric_barrier(const Peers& P, int rank, int world, uint32_t epoch, int slot_base) {
  __syncthreads();
  if (threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(P.ctrl[threadIdx.x] + slot_base + rank, epoch);
    const uint32_t* mine = P.ctrl[rank] + slot_base + threadIdx.x;
    wait_ge_sys(mine, epoch, 0x201);
  }
  __syncthreads();
}

B2_DEVICE void mm_ld_reduce_f32x4(const void* mc, float& a, float& b, float& c, float& d) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(mc) : "memory");
}
B2_DEVICE void mm_st_f32x4(void* mc, float a, float b, float c, float d) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
B2_DEVICE void mm_ld_reduce_bf16x8(const void* mc, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(mc) : "memory");
}
B2_DEVICE void mm_st_b32x4(void* mc, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ctrl words: [0] gate (epoch published to the CTAs of this grid), [1] epoch of the last COMPLETED call,
// [2] CTA completion ticket.  Calls on one buffer are stream-ordered, so word [1] is stable at kernel entry.
B2_DEVICE uint32_t enter(const Peers& P, int rank, int world, uint32_t* my_ctrl) {
  const uint32_t epoch = my_ctrl[1] + 1;
  if (blockIdx.x == 0) {
    fabric_barrier(P, rank, world, epoch, CTRL_A);         // every rank has launched and its input is visible
    if (threadIdx.x == 0) st_release_gpu(my_ctrl + 0, epoch);
  } else {
    if (threadIdx.x == 0) {
      wait_ge_gpu(my_ctrl + 0, epoch, 0x202);
    }
    __syncthreads();
  }
  return epoch;
}

// MODE 0: all-reduce in place.  MODE 1: reduce-scatter (rank r's reduced slice -> local `out`, payload untouched).
// MODE 2: all-gather (rank r's slice of the payload is broadcast into every rank's payload; no arithmetic).
template <typename T, bool MULTICAST, int MODE>
__global__ void __launch_bounds__(512)
allreduce_kernel(Peers P, void* mc_base, long long n_elems, float scale, int rank, int world, void* out) {
  uint32_t* my_ctrl = P.ctrl[rank];
  const uint32_t epoch = enter(P, rank, world, my_ctrl);
  // my slice, in 16-byte vectors
  constexpr int EPV = 16 / sizeof(T);
  const long long nvec = n_elems / EPV;
  const long long per = (nvec + world - 1) / world;
  const long long v0 = per * rank, v1 = min(nvec, v0 + per);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  long long v = v0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if constexpr (MULTICAST && MODE != 2) {
    // four independent 16-byte multimem.ld_reduce per thread in flight: the switch round trip (~2-3 us) needs depth, not
    // more CTAs - one load per iteration left this kernel behind NCCL from 25 MiB up
    constexpr int U = 4;
    for (; v + (U - 1) * stride < v1; v += U * stride) {
      uint32_t r[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint8_t* mc = static_cast<const uint8_t*>(mc_base) + (v + u * stride) * 16;
        if constexpr (sizeof(T) == 4) {
          float a, b, c, d;
          mm_ld_reduce_f32x4(mc, a, b, c, d);
          r[u][0] = __float_as_uint(a); r[u][1] = __float_as_uint(b); r[u][2] = __float_as_uint(c); r[u][3] = __float_as_uint(d);
        } else {
          mm_ld_reduce_bf16x8(mc, r[u][0], r[u][1], r[u][2], r[u][3]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if constexpr (sizeof(T) == 4) {
#pragma unroll
          for (int i = 0; i < 4; ++i) r[u][i] = __float_as_uint(__uint_as_float(r[u][i]) * scale);
        } else if (scale != 1.f) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { const float2 f = unpack_bf16x2(r[u][i]); r[u][i] = pack_bf16x2(f.x * scale, f.y * scale); }
        }
        if constexpr (MODE == 1) {
          *reinterpret_cast<uint4*>(static_cast<uint8_t*>(out) + (v + u * stride - v0) * 16) = make_uint4(r[u][0], r[u][1], r[u][2], r[u][3]);
        } else {
          mm_st_b32x4(static_cast<uint8_t*>(mc_base) + (v + u * stride) * 16, r[u][0], r[u][1], r[u][2], r[u][3]);
        }
      }
    }
  }
  if constexpr (!MULTICAST && MODE != 2) {
    // two ranks: plain peer loads / stores move n bytes per direction where the multicast path moves 1.5 n (the switch
    // also fetches the requester's own copy); four independent remote loads per thread hide the NVLink round trip
    if (world == 2) {
      constexpr int U = 4;
      const uint8_t* mine = static_cast<const uint8_t*>(P.ptr[rank]);
      const uint8_t* peer = static_cast<const uint8_t*>(P.ptr[rank ^ 1]);
      for (; v + (U - 1) * stride < v1; v += U * stride) {
        uint4 a[U], b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) b[u] = *reinterpret_cast<const uint4*>(peer + (v + u * stride) * 16);
#pragma unroll
        for (int u = 0; u < U; ++u) a[u] = *reinterpret_cast<const uint4*>(mine + (v + u * stride) * 16);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          uint4 o;
          if constexpr (sizeof(T) == 4) {
            o = make_uint4(__float_as_uint((__uint_as_float(a[u].x) + __uint_as_float(b[u].x)) * scale), __float_as_uint((__uint_as_float(a[u].y) + __uint_as_float(b[u].y)) * scale),
                           __float_as_uint((__uint_as_float(a[u].z) + __uint_as_float(b[u].z)) * scale), __float_as_uint((__uint_as_float(a[u].w) + __uint_as_float(b[u].w)) * scale));
          } else {
            const uint32_t* pa = reinterpret_cast<const uint32_t*>(&a[u]);
            const uint32_t* pb = reinterpret_cast<const uint32_t*>(&b[u]);
            uint32_t po[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 fa = unpack_bf16x2(pa[i]), fb = unpack_bf16x2(pb[i]);
              po[i] = pack_bf16x2((fa.x + fb.x) * scale, (fa.y + fb.y) * scale);
            }
            o = make_uint4(po[0], po[1], po[2], po[3]);
          }
          if constexpr (MODE == 1) {
            *reinterpret_cast<uint4*>(static_cast<uint8_t*>(out) + (v + u * stride - v0) * 16) = o;
          } else {
            *reinterpret_cast<uint4*>(static_cast<uint8_t*>(P.ptr[rank]) + (v + u * stride) * 16) = o;
            *reinterpret_cast<uint4*>(static_cast<uint8_t*>(P.ptr[rank ^ 1]) + (v + u * stride) * 16) = o;
          }
        }
      }
    }
  }
  for (; v < v1; v += stride) {
    if constexpr (MODE == 2) {
      const uint4 q = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(P.ptr[rank]) + v * 16);
      if constexpr (MULTICAST) {
        mm_st_b32x4(static_cast<uint8_t*>(mc_base) + v * 16, q.x, q.y, q.z, q.w);
      } else {
        for (int r = 0; r < world; ++r)
          if (r != rank) *reinterpret_cast<uint4*>(static_cast<uint8_t*>(P.ptr[r]) + v * 16) = q;
      }
    } else if constexpr (MULTICAST) {
      uint8_t* mc = static_cast<uint8_t*>(mc_base) + v * 16;
      uint32_t a, b, c, d;
      if constexpr (sizeof(T) == 4) {
        float fa, fb, fc, fd;
        mm_ld_reduce_f32x4(mc, fa, fb, fc, fd);
        a = __float_as_uint(fa * scale); b = __float_as_uint(fb * scale); c = __float_as_uint(fc * scale); d = __float_as_uint(fd * scale);
      } else {
        mm_ld_reduce_bf16x8(mc, a, b, c, d);
        if (scale != 1.f) {
          float2 f;
          f = unpack_bf16x2(a); a = pack_bf16x2(f.x * scale, f.y * scale);
          f = unpack_bf16x2(b); b = pack_bf16x2(f.x * scale, f.y * scale);
          f = unpack_bf16x2(c); c = pack_bf16x2(f.x * scale, f.y * scale);
          f = unpack_bf16x2(d); d = pack_bf16x2(f.x * scale, f.y * scale);
        }
      }
      if constexpr (MODE == 1) *reinterpret_cast<uint4*>(static_cast<uint8_t*>(out) + (v - v0) * 16) = make_uint4(a, b, c, d);
      else                     mm_st_b32x4(mc, a, b, c, d);
    } else {
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
      for (int r = 0; r < world; ++r) {
        uint4 q = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(P.ptr[r]) + v * 16);
        if constexpr (sizeof(T) == 4) {
          acc[0] += __uint_as_float(q.x); acc[1] += __uint_as_float(q.y); acc[2] += __uint_as_float(q.z); acc[3] += __uint_as_float(q.w);
        } else {
          float2 f;
          f = unpack_bf16x2(q.x); acc[0] += f.x; acc[1] += f.y;
          f = unpack_bf16x2(q.y); acc[2] += f.x; acc[3] += f.y;
          f = unpack_bf16x2(q.z); acc[4] += f.x; acc[5] += f.y;
          f = unpack_bf16x2(q.w); acc[6] += f.x; acc[7] += f.y;
        }
      }
      uint4 o;
      if constexpr (sizeof(T) == 4) {
        o = make_uint4(__float_as_uint(acc[0] * scale), __float_as_uint(acc[1] * scale), __float_as_uint(acc[2] * scale), __float_as_uint(acc[3] * scale));
      } else {
        o = make_uint4(pack_bf16x2(acc[0] * scale, acc[1] * scale), pack_bf16x2(acc[2] * scale, acc[3] * scale),
                       pack_bf16x2(acc[4] * scale, acc[5] * scale), pack_bf16x2(acc[6] * scale, acc[7] * scale));
      }
      if constexpr (MODE == 1) {
        *reinterpret_cast<uint4*>(static_cast<uint8_t*>(out) + (v - v0) * 16) = o;
      } else {
        for (int r = 0; r < world; ++r)
          *reinterpret_cast<uint4*>(static_cast<uint8_t*>(P.ptr[r]) + v * 16) = o;
      }
    }
  }
  // all CTAs of this grid done -> one cross-GPU barrier -> result complete everywhere
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(my_ctrl + 2, 1u);           // word 2 = CTA completion ticket
    last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    fabric_barrier(P, rank, world, epoch, CTRL_B);
    if (threadIdx.x == 0) {
      my_ctrl[2] = 0u;
      my_ctrl[1] = epoch;                                  // this call is complete (next launch's epoch base)
      __threadfence();
    }
  }
}

// ZeRO gradient path: reduce-scatter of a symmetric gradient buffer fused with what follows it.
//   slice `rank` of sum_r grad_r  * gscale  ->  (+)= fp32 gradient shard `gacc`   (gradient accumulation steps)
//   and, on the optimizer boundary (ADAM), straight into AdamW on the fp32 master / m / v shard, emitting the
//   low-precision parameter shard - the reduced gradient never round-trips through memory between the two.
// replaces: DeepSpeed stage3 reduce_scatter_coalesced + FusedAdam multi_tensor_adam.
template <typename T, bool MULTICAST, bool ADAM>
__global__ void __launch_bounds__(512)
rs_update_kernel(Peers P, void* mc_base, long long n_elems, int rank, int world, float gscale, float* __restrict__ gacc,
                 int accumulate, float* __restrict__ sumsq, float* __restrict__ p32, float* __restrict__ m,
                 float* __restrict__ v, __nv_bfloat16* __restrict__ p_lp, float lr, float beta1, float beta2, float eps,
                 float wd, float bc1, float bc2) {
  uint32_t* my_ctrl = P.ctrl[rank];
  const uint32_t epoch = enter(P, rank, world, my_ctrl);
  constexpr int EPV = 16 / sizeof(T);
  const long long nvec = n_elems / EPV;
  const long long per = (nvec + world - 1) / world;
  const long long v0 = per * rank, v1 = min(nvec, v0 + per);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  float ss = 0.f;
  for (long long vi = v0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; vi < v1; vi += stride) {
    float g[EPV];
    if constexpr (MULTICAST) {
      uint8_t* mc = static_cast<uint8_t*>(mc_base) + vi * 16;
      if constexpr (sizeof(T) == 4) {
        mm_ld_reduce_f32x4(mc, g[0], g[1], g[2], g[3]);
      } else {
        uint32_t a, b, c, d;
        mm_ld_reduce_bf16x8(mc, a, b, c, d);
        float2 f;
        f = unpack_bf16x2(a); g[0] = f.x; g[1] = f.y; f = unpack_bf16x2(b); g[2] = f.x; g[3] = f.y;
        f = unpack_bf16x2(c); g[4] = f.x; g[5] = f.y; f = unpack_bf16x2(d); g[6] = f.x; g[7] = f.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < EPV; ++i) g[i] = 0.f;
      for (int r = 0; r < world; ++r) {
        const uint4 q = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(P.ptr[r]) + vi * 16);
        if constexpr (sizeof(T) == 4) {
          g[0] += __uint_as_float(q.x); g[1] += __uint_as_float(q.y); g[2] += __uint_as_float(q.z); g[3] += __uint_as_float(q.w);
        } else {
          float2 f;
          f = unpack_bf16x2(q.x); g[0] += f.x; g[1] += f.y; f = unpack_bf16x2(q.y); g[2] += f.x; g[3] += f.y;
          f = unpack_bf16x2(q.z); g[4] += f.x; g[5] += f.y; f = unpack_bf16x2(q.w); g[6] += f.x; g[7] += f.y;
        }
      }
    }
    const long long e0 = (vi - v0) * EPV;       // element offset inside this rank's shard
#pragma unroll
    for (int i = 0; i < EPV; ++i) {
      float gr = g[i] * gscale;
      if (gacc != nullptr) {
        if (accumulate) gr += gacc[e0 + i];
        gacc[e0 + i] = gr;
      }
      ss += gr * gr;
      if constexpr (ADAM) {
        float pa = p32[e0 + i], ma = m[e0 + i], va = v[e0 + i];
        ma = beta1 * ma + (1.f - beta1) * gr;
        va = beta2 * va + (1.f - beta2) * gr * gr;
        pa -= lr * ((ma / bc1) / (sqrtf(va / bc2) + eps) + wd * pa);
        p32[e0 + i] = pa; m[e0 + i] = ma; v[e0 + i] = va;
        if (p_lp != nullptr) p_lp[e0 + i] = __float2bfloat16(pa);
      }
    }
  }
  if (sumsq != nullptr) {
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0 && ss != 0.f) atomicAdd(sumsq, ss);
  }
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) last = (atomicAdd(my_ctrl + 2, 1u) == gridDim.x - 1);
  __syncthreads();
  if (last) {
    fabric_barrier(P, rank, world, epoch, CTRL_B);      // nobody may overwrite its gradient buffer before all have read it
    if (threadIdx.x == 0) { my_ctrl[2] = 0u; my_ctrl[1] = epoch; __threadfence(); }
  }
}

int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}

}  // namespace

extern "C" int b2_ctrl_words(void) { return B2_CTRL_WORDS; }

// ptrs / ctrls: `world` peer-mapped addresses each (host arrays).  mc_ptr: multicast address of the payload or NULL.
// n_elems * sizeof(dtype) must be a multiple of 16.  In place on every rank's payload.
static bool g_force_no_mc = false;
// A/B switch for measurements: 1 = never use the multicast (multimem) instructions, 0 = default selection
extern "C" void b2_collective_force_peer_path(int on) { g_force_no_mc = on != 0; }
static int run_collective(int mode, void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, float scale,
                          int rank, int world, int max_ctas, void* out, cudaStream_t st) {
  if (world < 1 || world > MAX_RANKS) { b2host::set_error("collective: world size %d unsupported (max %d)", world, MAX_RANKS); return -1; }
  const int esz = dtype == B2_DTYPE_FP32 ? 4 : 2;
  if (dtype == B2_DTYPE_FP16 && mode != 2) { b2host::set_error("collective: fp16 reductions are not supported (use bf16 or fp32)"); return -1; }
  if ((n_elems * esz) % (16LL * world)) { b2host::set_error("collective: payload must be a multiple of 16 bytes per rank"); return -1; }
  Peers P;
  for (int r = 0; r < MAX_RANKS; ++r) { P.ptr[r] = r < world ? ptrs[r] : nullptr; P.ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr; }
  long long nvec = n_elems * esz / 16;
  long long per = (nvec + world - 1) / world;
  int threads = 512;
  long long want = (per + threads - 1) / threads;
  // every thread keeps 4 vectors in flight: payloads above ~16 MiB per rank get up to 128 CTAs, small ones stay small so
  // that a bucket reduced on the communication stream takes few SMs from the backward pass it overlaps
  want = (want + 3) / 4;
  int cap = max_ctas > 0 ? max_ctas : (per * 16 >= (16LL << 20) ? 128 : 64);
  if (cap > b2host::num_sms()) cap = b2host::num_sms();
  int grid = static_cast<int>(want < 1 ? 1 : (want > cap ? cap : want));
#define B2_LAUNCH(T_, MC_, MODE_) allreduce_kernel<T_, MC_, MODE_><<<grid, threads, 0, st>>>(P, mc_ptr, n_elems, scale, rank, world, out)
#define B2_BY_MODE(T_, MC_) do { if (mode == 0) B2_LAUNCH(T_, MC_, 0); else if (mode == 1) B2_LAUNCH(T_, MC_, 1); else B2_LAUNCH(T_, MC_, 2); } while (0)
  // reductions between TWO ranks take the peer load/store path (less NVLink traffic than a round trip through the switch's
  // multicast unit, see the kernel); the all-gather (mode 2) and larger groups use multimem when it is available
  const bool use_mc = mc_ptr != nullptr && (world > 2 || mode == 2) && !g_force_no_mc;
  if (use_mc) { if (esz == 4) B2_BY_MODE(float, true); else B2_BY_MODE(__nv_bfloat16, true); }
  else                   { if (esz == 4) B2_BY_MODE(float, false); else B2_BY_MODE(__nv_bfloat16, false); }
#undef B2_BY_MODE
#undef B2_LAUNCH
  return check("collective");
}

extern "C" int b2_allreduce(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, float scale,
                            int rank, int world, int max_ctas, void* stream) {
  return run_collective(0, ptrs, ctrls, mc_ptr, n_elems, dtype, scale, rank, world, max_ctas, nullptr, static_cast<cudaStream_t>(stream));
}
// out (local, n_elems / world elements) = scale * sum over ranks of slice `rank` of the symmetric payload
extern "C" int b2_reduce_scatter(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, float scale,
                                 int rank, int world, int max_ctas, void* out, void* stream) {
  return run_collective(1, ptrs, ctrls, mc_ptr, n_elems, dtype, scale, rank, world, max_ctas, out, static_cast<cudaStream_t>(stream));
}
// every rank's slice `r` of the symmetric payload (n_elems total) is replicated into all ranks' payloads
extern "C" int b2_allgather(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, int rank, int world,
                            int max_ctas, void* stream) {
  return run_collective(2, ptrs, ctrls, mc_ptr, n_elems, dtype, 1.f, rank, world, max_ctas, nullptr, static_cast<cudaStream_t>(stream));
}

// ZeRO: reduce-scatter (+ accumulate into the fp32 gradient shard) (+ AdamW on the shard when `adam` != 0).
// gacc may be NULL when adam != 0 and no accumulation is wanted; sumsq (device, optional) receives += sum g^2.
extern "C" int b2_rs_update(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, int rank, int world,
                            float gscale, float* gacc, int accumulate, float* sumsq, int adam, float* p32, float* m, float* v,
                            void* p_lp, float lr, float beta1, float beta2, float eps, float wd, int step, int max_ctas, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (world < 1 || world > MAX_RANKS) { b2host::set_error("rs_update: world size %d unsupported", world); return -1; }
  if (dtype != B2_DTYPE_FP32 && dtype != B2_DTYPE_BF16) { b2host::set_error("rs_update: gradients must be fp32 or bf16"); return -1; }
  const int esz = dtype == B2_DTYPE_FP32 ? 4 : 2;
  if ((n_elems * esz) % (16LL * world)) { b2host::set_error("rs_update: payload must be a multiple of 16 bytes per rank"); return -1; }
  Peers P;
  for (int r = 0; r < MAX_RANKS; ++r) { P.ptr[r] = r < world ? ptrs[r] : nullptr; P.ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr; }
  const long long per = n_elems * esz / 16 / world;
  const int threads = 512;
  long long want = (per + threads - 1) / threads;
  const int cap = max_ctas > 0 ? max_ctas : 96;
  const int grid = static_cast<int>(want < 1 ? 1 : (want > cap ? cap : want));
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step > 0 ? step : 1));
  const float bc2 = 1.f - powf(beta2, static_cast<float>(step > 0 ? step : 1));
  __nv_bfloat16* lp = static_cast<__nv_bfloat16*>(p_lp);
#define B2_RS(T_, MC_, AD_) rs_update_kernel<T_, MC_, AD_><<<grid, threads, 0, st>>>(P, mc_ptr, n_elems, rank, world, gscale, gacc, accumulate, sumsq, p32, m, v, lp, lr, beta1, beta2, eps, wd, bc1, bc2)
  if (mc_ptr != nullptr) {
    if (esz == 4) { if (adam) B2_RS(float, true, true); else B2_RS(float, true, false); }
    else          { if (adam) B2_RS(__nv_bfloat16, true, true); else B2_RS(__nv_bfloat16, true, false); }
  } else {
    if (esz == 4) { if (adam) B2_RS(float, false, true); else B2_RS(float, false, false); }
    else          { if (adam) B2_RS(__nv_bfloat16, false, true); else B2_RS(__nv_bfloat16, false, false); }
  }
#undef B2_RS
  return check("rs_update");
}

namespace b2host {
int set_spin_timeout_comm(unsigned long long ns) {
  return cudaMemcpyToSymbol(b2::g_b2_spin_timeout_ns, &ns, sizeof(ns)) == cudaSuccess ? 0 : -1;
}
}  // namespace b2host
