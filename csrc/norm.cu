// LayerNorm / RMSNorm forward and backward (sm_100a, memory-bound: 128-bit loads, 8 / 16 / 32 lanes or a whole CTA per
// row depending on its width, row held in registers, fp32 math, bf16/fp16 I/O).
//
// replaces: apex fused_layer_norm_cuda (used by Megatron MixedFusedLayerNorm), ATen
// layer_norm under swin_b / OPT (K6, K24) and the unfused HF LlamaRMSNorm (K24).
//
// Backward produces dx in one pass and accumulates dgamma/dbeta partials per CTA into a small fp32
// workspace [ctas, 2, H] that a second kernel folds (optionally accumulating into fp32 main_grad).
#ifdef B2_EMU                      // CPU emulation of the CUDA execution model (tests/emu): kernels only
#include "vec_emu.h"
namespace b2 {}
#else
#include "common.cuh"
#endif
#include "b200dist.h"

using namespace b2;

namespace {

template <typename T> struct Cvt;
#ifdef B2_EMU
template <typename T> struct Cvt {
  static B2_DEVICE float2 up(uint32_t u) { return emu_unpack2<T>(u); }
  static B2_DEVICE uint32_t down(float a, float b) { return emu_pack2<T>(a, b); }
};
#else
template <> struct Cvt<__nv_bfloat16> {
  static B2_DEVICE float2 up(uint32_t u) { return unpack_bf16x2(u); }
  static B2_DEVICE uint32_t down(float a, float b) { return pack_bf16x2(a, b); }
};
template <> struct Cvt<__half> {
  static B2_DEVICE float2 up(uint32_t u) { __half2 h = *reinterpret_cast<__half2*>(&u); return __half22float2(h); }
  static B2_DEVICE uint32_t down(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
};
#endif

template <typename T>
B2_DEVICE void load8(const T* p, float* f) {
  uint4 q = *reinterpret_cast<const uint4*>(p);
  float2 a = Cvt<T>::up(q.x), b = Cvt<T>::up(q.y), c = Cvt<T>::up(q.z), d = Cvt<T>::up(q.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <typename T>
B2_DEVICE void store8(T* p, const float* f) {
  uint4 q = make_uint4(Cvt<T>::down(f[0], f[1]), Cvt<T>::down(f[2], f[3]), Cvt<T>::down(f[4], f[5]),
                       Cvt<T>::down(f[6], f[7]));
  *reinterpret_cast<uint4*>(p) = q;
}

constexpr int CTA_THREADS = 128;
constexpr int MAXV = 4;   // 8-element vectors per thread: TEAM * MAXV * 8 >= H

// TEAM threads cooperate on one row.  H <= 1024: TEAM = 8 / 16 / 32 lanes of a warp (team_for below) - narrow rows (Swin's 128 /
// 256 channels) then keep every lane busy with 4 / 2 rows per warp instead of one half-used warp per row, which is what a
// latency-bound pass over hundreds of thousands of short rows needs; CTA_THREADS / TEAM rows per CTA.
// H <= 8192: TEAM = 256, the whole CTA on one row.
// Teams that share a warp run the same number of loop iterations (their reductions are full-mask shuffles): the loop is over
// the warp's first row and a team whose row lies past the end works on zeros and stores nothing.
template <int TEAM>
B2_DEVICE float team_sum(float v, float* red) {
  if constexpr (TEAM < 32) {
#pragma unroll
    for (int o = TEAM / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  }
  v = warp_sum(v);
  if constexpr (TEAM > 32) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < TEAM / 32; ++w) t += red[w];
    v = t;
  }
  return v;
}

template <typename T, int TEAM, bool RMS>
__global__ void __launch_bounds__(TEAM <= 32 ? CTA_THREADS : TEAM)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma, const T* __restrict__ beta, T* __restrict__ y,
                float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int H, float eps,
                const int* __restrict__ row_map /* optional: y row of x row r is row_map[r] (a permutation) */) {
  __shared__ float red[8];
  constexpr int ROWS_PER_CTA = TEAM <= 32 ? CTA_THREADS / TEAM : 1;
  constexpr int TPW = TEAM < 32 ? 32 / TEAM : 1;                 // teams per warp
  const int tid = TEAM <= 32 ? static_cast<int>(threadIdx.x % TEAM) : static_cast<int>(threadIdx.x);
  const int sub = TEAM <= 32 ? static_cast<int>(threadIdx.x / TEAM) : 0;
  const int nvec = H >> 3;
  for (int base = blockIdx.x * ROWS_PER_CTA + (sub / TPW) * TPW; base < rows; base += gridDim.x * ROWS_PER_CTA) {
    const int row = base + sub % TPW;
    const bool live = row < rows;
    const T* xr = x + static_cast<size_t>(live ? row : rows - 1) * H;
    float v[MAXV][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
        load8<T>(xr + vi * 8, v[i]);
#pragma unroll
        for (int e = 0; e < 8; ++e) s += v[i][e];
      }
    }
    float mean = 0.f;
    if (!RMS) { s = team_sum<TEAM>(s, red); mean = s / H; }
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { float d = v[i][e] - mean; ss += d * d; }
      }
    }
    ss = team_sum<TEAM>(ss, red);
    const float rstd = rsqrtf(ss / H + eps);
    if (!live) continue;                       // (after the team reductions: every lane of the warp took part in them)
    if (tid == 0) {
      if (!RMS && mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + static_cast<size_t>(row_map ? row_map[row] : row) * H;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
        float g[8], b[8], o[8];
        load8<T>(gamma + vi * 8, g);
        if (!RMS && beta) load8<T>(beta + vi * 8, b);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float xh = (v[i][e] - mean) * rstd;
          o[e] = xh * g[e] + ((!RMS && beta) ? b[e] : 0.f);
        }
        store8<T>(yr + vi * 8, o);
      }
    }
  }
}

template <typename T, int TEAM, bool RMS, int NV /* 8-element vectors per thread: TEAM * NV * 8 >= H */>
__global__ void __launch_bounds__(TEAM <= 32 ? CTA_THREADS : TEAM, (TEAM <= 32 && NV <= 3) ? 4 : 1)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ gamma,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                const T* __restrict__ dres /* optional: dx = dres + d(norm) (residual branch gradient) */,
                T* __restrict__ dx, float* __restrict__ ws /* [gridDim.x, 2, H] */, int rows, int H,
                const int* __restrict__ row_map /* optional: the dy row of x row r is row_map[r] (forward wrote y through it) */) {
  B2_DYN_SMEM(float, sm);        // TEAM <= 32: [CTA_THREADS / TEAM][2][H]
  __shared__ float red[8];
  constexpr int ROWS_PER_CTA = TEAM <= 32 ? CTA_THREADS / TEAM : 1;
  constexpr int TPW = TEAM < 32 ? 32 / TEAM : 1;                 // teams per warp
  const int tid = TEAM <= 32 ? static_cast<int>(threadIdx.x % TEAM) : static_cast<int>(threadIdx.x);
  const int sub = TEAM <= 32 ? static_cast<int>(threadIdx.x / TEAM) : 0;
  const int nvec = H >> 3;
  float dg[NV][8], db[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) { dg[i][e] = 0.f; db[i][e] = 0.f; }

  // TEAM > 32 uses __syncthreads inside team_sum and teams that share a warp use full-mask shuffles: every thread of a CTA /
  // warp runs the same trip count
  for (int base = blockIdx.x * ROWS_PER_CTA + (sub / TPW) * TPW; base < rows; base += gridDim.x * ROWS_PER_CTA) {
    const int row_ = base + sub % TPW;
    const bool live = row_ < rows;
    const int row = live ? row_ : rows - 1;          // a team past the end reads a valid row, treats dy as 0 and stores nothing
    const T* xr = x + static_cast<size_t>(row) * H;
    const T* dyr = dy + static_cast<size_t>(row_map ? row_map[row] : row) * H;
    const float mean = RMS ? 0.f : mean_in[row];
    const float rstd = rstd_in[row];
    // the row stays in registers in its packed 16-bit form (half the registers of fp32 copies -> more CTAs per SM,
    // more loads in flight); the second pass re-derives xhat and dy*gamma from it
    uint4 xq[NV], dq[NV], rq[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
        xq[i] = *reinterpret_cast<const uint4*>(xr + vi * 8);
        dq[i] = live ? *reinterpret_cast<const uint4*>(dyr + vi * 8) : make_uint4(0u, 0u, 0u, 0u);
        if (dres != nullptr) rq[i] = *reinterpret_cast<const uint4*>(dres + static_cast<size_t>(row) * H + vi * 8);
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
        float xv[8], dv[8], g[8];
        load8<T>(reinterpret_cast<const T*>(&xq[i]), xv);
        load8<T>(reinterpret_cast<const T*>(&dq[i]), dv);
        load8<T>(gamma + vi * 8, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float h = (xv[e] - mean) * rstd;
          float t = dv[e] * g[e];
          s1 += t; s2 += t * h;
          dg[i][e] += dv[e] * h;
          db[i][e] += dv[e];
        }
      }
    }
    s1 = team_sum<TEAM>(s1, red); s2 = team_sum<TEAM>(s2, red);
    const float m1 = RMS ? 0.f : s1 / H, m2 = s2 / H;
    if (!live) continue;
    T* dxr = dx + static_cast<size_t>(row) * H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
        float xv[8], dv[8], g[8], o[8];
        load8<T>(reinterpret_cast<const T*>(&xq[i]), xv);
        load8<T>(reinterpret_cast<const T*>(&dq[i]), dv);
        load8<T>(gamma + vi * 8, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = rstd * (dv[e] * g[e] - m1 - (xv[e] - mean) * rstd * m2);
        if (dres != nullptr) {
          float r[8];
          load8<T>(reinterpret_cast<const T*>(&rq[i]), r);
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] += r[e];
        }
        store8<T>(dxr + vi * 8, o);
      }
    }
  }
  float* wrow = ws + static_cast<size_t>(blockIdx.x) * 2 * H;
  if constexpr (TEAM <= 32) {
    // fold the per-team partials through smem, one workspace row per CTA
    float* mine = sm + sub * 2 * H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { mine[vi * 8 + e] = dg[i][e]; mine[H + vi * 8 + e] = db[i][e]; }
      }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < 2 * H; j += blockDim.x) {
      float a = 0.f;
#pragma unroll
      for (int w = 0; w < ROWS_PER_CTA; ++w) a += sm[w * 2 * H + j];
      wrow[j] = a;
    }
  } else {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      int vi = tid + i * TEAM;
      if (vi < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { wrow[vi * 8 + e] = dg[i][e]; wrow[H + vi * 8 + e] = db[i][e]; }
      }
    }
  }
}

// fold workspace rows -> dgamma / dbeta (fp32, optionally accumulated: main_grad semantics).
// One 1024-thread CTA per 32 columns: warp w sums parts w, w+32, ... with 8 independent loads in flight (the loop is
// latency-bound: a few MB spread over many short strided reads), then the 32 warps fold through shared memory.
__global__ void __launch_bounds__(1024)
norm_bwd_fold_kernel(const float* __restrict__ ws, int nparts, int H, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, int accumulate) {
  __shared__ float part[32][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (j < 2 * H) {
    const float* col = ws + j;
    const size_t ld = static_cast<size_t>(2) * H;
    int p = w;
    for (; p + 7 * 32 < nparts; p += 8 * 32) {
#pragma unroll
      for (int u = 0; u < 8; ++u) a[u] += col[(p + u * 32) * ld];
    }
    for (; p < nparts; p += 32) a[0] += col[p * ld];
  }
  part[w][lane] = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
  __syncthreads();
  if (w == 0 && j < 2 * H) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) t += part[k][lane];
    float* dst = j < H ? (dgamma + j) : (dbeta ? dbeta + (j - H) : nullptr);
    if (dst) { if (accumulate) *dst += t; else *dst = t; }
  }
}

// lanes per row for H <= 1024: two vectors per thread for narrow rows.  Registers bound the loads in flight per SM (a vector slot of
// the backward costs ~28 registers: packed x / dy / dres and 16 fp32 dgamma / dbeta partials); two to three slots per thread at
// four to five CTAs per SM is the maximum, one slot with half the lanes idle (a warp on a 128-wide row) a quarter of it.
int team_for(int H) {
  const int nvec = H >> 3;
  return nvec <= 16 ? 8 : (nvec <= 32 ? 16 : 32);
}

#ifndef B2_EMU
template <typename T, bool RMS>
int fwd_dispatch(const void* x, const void* g, const void* b, void* y, float* mean, float* rstd, int rows, int H,
                 float eps, const int* row_map, cudaStream_t st) {
  if (H <= 1024) {
    const int team = team_for(H), rpc = CTA_THREADS / team;
    int grid = (rows + rpc - 1) / rpc;
    int cap = b2host::num_sms() * 16;
    if (grid > cap) grid = cap;
#define B2_NF(TEAM_) norm_fwd_kernel<T, TEAM_, RMS><<<grid, CTA_THREADS, 0, st>>>((const T*)x, (const T*)g, (const T*)b, (T*)y, mean, rstd, rows, H, eps, row_map)
    if (team == 8) B2_NF(8); else if (team == 16) B2_NF(16); else B2_NF(32);
#undef B2_NF
  } else if (H <= 8192) {
    int grid = rows;
    int cap = b2host::num_sms() * 8;
    if (grid > cap) grid = cap;
    norm_fwd_kernel<T, 256, RMS><<<grid, 256, 0, st>>>((const T*)x, (const T*)g, (const T*)b, (T*)y, mean, rstd, rows, H, eps, row_map);
  } else { b2host::set_error("norm: hidden size %d > 8192 unsupported", H); return -1; }
  return 0;
}

int bwd_parts(int rows, int H) {
  const int rpc = H <= 1024 ? CTA_THREADS / team_for(H) : 1;
  int grid = (rows + rpc - 1) / rpc;
  int cap = b2host::num_sms() * (H <= 1024 ? 4 : 2);
  return grid < cap ? grid : cap;
}

template <typename T, bool RMS>
int bwd_dispatch(const void* dy, const void* x, const void* g, const float* mean, const float* rstd, const void* dres,
                 void* dx, float* ws, int grid, int rows, int H, const int* row_map, cudaStream_t st) {
  if (H <= 1024) {
    const int team = team_for(H), nvec = H >> 3;
    size_t smem = static_cast<size_t>(CTA_THREADS / team) * 2 * H * sizeof(float);      // <= 32 KB (16 x 2 x 128, 8 x 2 x 256, 4 x 2 x 1024 floats)
#define B2_NB(TEAM_, NV_) norm_bwd_kernel<T, TEAM_, RMS, NV_><<<grid, CTA_THREADS, smem, st>>>((const T*)dy, (const T*)x, (const T*)g, mean, rstd, (const T*)dres, (T*)dx, ws, rows, H, row_map)
    const int nv = (nvec + team - 1) / team;
    if (team == 8) { if (nv <= 1) B2_NB(8, 1); else B2_NB(8, 2); }
    else if (team == 16) B2_NB(16, 2);
    else { if (nv <= 2) B2_NB(32, 2); else if (nv == 3) B2_NB(32, 3); else B2_NB(32, 4); }
#undef B2_NB
  } else if (H <= 8192) {
#define B2_NB(NV_) norm_bwd_kernel<T, 256, RMS, NV_><<<grid, 256, 0, st>>>((const T*)dy, (const T*)x, (const T*)g, mean, rstd, (const T*)dres, (T*)dx, ws, rows, H, row_map)
    const int nv = (H + 2047) / 2048;
    if (nv == 1) B2_NB(1); else if (nv == 2) B2_NB(2); else if (nv == 3) B2_NB(3); else B2_NB(4);
#undef B2_NB
  } else { b2host::set_error("norm bwd: hidden size %d > 8192 unsupported", H); return -1; }
  return 0;
}

int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}

#endif  // !B2_EMU

}  // namespace

#ifndef B2_EMU
extern "C" int b2_norm_bwd_num_parts(int rows, int H) { return bwd_parts(rows, H); }

extern "C" int b2_layernorm_fwd_map(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                    int rows, int H, float eps, const int* row_map, int dtype, int rms, void* stream) {
  if (H % 8) { b2host::set_error("norm: hidden size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int r;
  if (dtype == B2_DTYPE_BF16) r = rms ? fwd_dispatch<__nv_bfloat16, true>(x, gamma, beta, y, mean, rstd, rows, H, eps, row_map, st)
                                      : fwd_dispatch<__nv_bfloat16, false>(x, gamma, beta, y, mean, rstd, rows, H, eps, row_map, st);
  else r = rms ? fwd_dispatch<__half, true>(x, gamma, beta, y, mean, rstd, rows, H, eps, row_map, st)
               : fwd_dispatch<__half, false>(x, gamma, beta, y, mean, rstd, rows, H, eps, row_map, st);
  return r ? r : check("layernorm_fwd");
}

extern "C" int b2_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                int rows, int H, float eps, int dtype, int rms, void* stream) {
  return b2_layernorm_fwd_map(x, gamma, beta, y, mean, rstd, rows, H, eps, nullptr, dtype, rms, stream);
}

// ws: fp32 workspace of b2_norm_bwd_num_parts(rows, H) * 2 * H elements
extern "C" int b2_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                                void* dx, float* dgamma, float* dbeta, float* ws, int rows, int H, int dtype, int rms,
                                int accumulate, void* stream) {
  return b2_layernorm_bwd_res(dy, x, gamma, mean, rstd, nullptr, dx, dgamma, dbeta, ws, rows, H, dtype, rms, accumulate, stream);
}

// same, with the residual-branch gradient folded in: dx = dres + d(norm)   (dres may be NULL; dres may alias dx)
extern "C" int b2_layernorm_bwd_res(const void* dy, const void* x, const void* gamma, const float* mean,
                                    const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta, float* ws,
                                    int rows, int H, int dtype, int rms, int accumulate, void* stream) {
  return b2_layernorm_bwd_map(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, ws, rows, H, nullptr, dtype, rms, accumulate, stream);
}

// same, for a forward that wrote its rows through `row_map` (b2_layernorm_fwd_map): dy is read through the same permutation, so a
// row shuffle that follows the norm (Swin: cyclic shift + window partition) costs no pass of its own in either direction
extern "C" int b2_layernorm_bwd_map(const void* dy, const void* x, const void* gamma, const float* mean,
                                    const float* rstd, const void* dres, void* dx, float* dgamma, float* dbeta, float* ws,
                                    int rows, int H, const int* row_map, int dtype, int rms, int accumulate, void* stream) {
  if (H % 8) { b2host::set_error("norm: hidden size must be a multiple of 8"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = bwd_parts(rows, H);
  int r;
  if (dtype == B2_DTYPE_BF16) r = rms ? bwd_dispatch<__nv_bfloat16, true>(dy, x, gamma, mean, rstd, dres, dx, ws, grid, rows, H, row_map, st)
                                      : bwd_dispatch<__nv_bfloat16, false>(dy, x, gamma, mean, rstd, dres, dx, ws, grid, rows, H, row_map, st);
  else r = rms ? bwd_dispatch<__half, true>(dy, x, gamma, mean, rstd, dres, dx, ws, grid, rows, H, row_map, st)
               : bwd_dispatch<__half, false>(dy, x, gamma, mean, rstd, dres, dx, ws, grid, rows, H, row_map, st);
  if (r) return r;
  norm_bwd_fold_kernel<<<(2 * H + 31) / 32, 1024, 0, st>>>(ws, grid, H, dgamma, rms ? nullptr : dbeta, accumulate);
  return check("layernorm_bwd");
}
#endif  // !B2_EMU
