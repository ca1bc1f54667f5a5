// Swin shifted-window attention, forward and backward, fused (sm_100a): entry points, the CUDA-core kernels (kept as the
// second implementation: b2_window_attn_set_impl(0)) and the shift / partition gather.  The default kernels are the
// tensor-core ones of window_attn_tc.cu.
//
//   per (window, head):  S = (q * scale) k^T + rel_pos_bias[head] (+ shift mask[window % nW]);  P = softmax(S);  O = P v
//
// replaces: torchvision ShiftedWindowAttention's eager chain (matmul -> +bias -> +mask -> fp32 softmax -> matmul,
// each step a round trip of the [windows, heads, 49, 49] score tensor through HBM).  Windows are 7 x 7 = 49 tokens and
// head_dim is 32 in every torchvision Swin stage, so one thread owns one query row: its 49 scores, 32 q and 32 output
// values live in registers, K / V sit in shared memory as fp32 and are read as warp broadcasts.  The arithmetic is a few
// per cent of a Swin step; the point of the kernel is that the scores never leave the SM.
//
// Layouts: qkv [Bw, 49, 3, heads, 32] (the fused projection's output, Bw = batch * nW windows), out [Bw, 49, heads * 32],
// bias fp32 [heads, 49, 49], mask fp32 [nW, 49, 49] or NULL, lse fp32 [Bw, heads, 49].
// Backward: one CTA walks a range of windows of ONE head, so d(bias) accumulates in registers and reaches memory with
// 49 x 49 atomics per CTA instead of per window.
#include <type_traits>
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int WT = 49;      // tokens per window
constexpr int WD = 32;      // head dim
constexpr int WG = 64;      // threads per (window, head): 49 active rows

template <typename T> struct WCvt;
template <> struct WCvt<__nv_bfloat16> {
  static B2_DEVICE float2 up(uint32_t u) { return unpack_bf16x2(u); }
  static B2_DEVICE uint32_t down(float a, float b) { return pack_bf16x2(a, b); }
};
template <> struct WCvt<__half> {
  static B2_DEVICE float2 up(uint32_t u) { __half2 h = *reinterpret_cast<__half2*>(&u); return __half22float2(h); }
  static B2_DEVICE uint32_t down(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
};

// 32 contiguous 16-bit values -> fp32
template <typename T>
B2_DEVICE void load_row32(const T* p, float* f) {
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const uint4 q = *reinterpret_cast<const uint4*>(p + c * 8);
    float2 a = WCvt<T>::up(q.x), b = WCvt<T>::up(q.y), cc = WCvt<T>::up(q.z), d = WCvt<T>::up(q.w);
    f[c * 8 + 0] = a.x; f[c * 8 + 1] = a.y; f[c * 8 + 2] = b.x; f[c * 8 + 3] = b.y;
    f[c * 8 + 4] = cc.x; f[c * 8 + 5] = cc.y; f[c * 8 + 6] = d.x; f[c * 8 + 7] = d.y;
  }
}
template <typename T>
B2_DEVICE void store_row32(T* p, const float* f) {
#pragma unroll
  for (int c = 0; c < 4; ++c)
    *reinterpret_cast<uint4*>(p + c * 8) = make_uint4(WCvt<T>::down(f[c * 8], f[c * 8 + 1]), WCvt<T>::down(f[c * 8 + 2], f[c * 8 + 3]),
                                                      WCvt<T>::down(f[c * 8 + 4], f[c * 8 + 5]), WCvt<T>::down(f[c * 8 + 6], f[c * 8 + 7]));
}
// cooperative: [49, 32] rows of one (window, part, head) -> fp32 smem tile; `t` in [0, WG)
template <typename T>
B2_DEVICE void stage_tile(const T* base, long long row_stride, float* dst, int t) {
  for (int v = t; v < WT * 4; v += WG) {
    const int j = v >> 2, c = v & 3;
    const uint4 q = *reinterpret_cast<const uint4*>(base + j * row_stride + c * 8);
    float2 a = WCvt<T>::up(q.x), b = WCvt<T>::up(q.y), cc = WCvt<T>::up(q.z), d = WCvt<T>::up(q.w);
    float* o = dst + j * WD + c * 8;
    *reinterpret_cast<float4*>(o) = make_float4(a.x, a.y, b.x, b.y);
    *reinterpret_cast<float4*>(o + 4) = make_float4(cc.x, cc.y, d.x, d.y);
  }
}
B2_DEVICE float dot32(const float* a, const float* s /* smem, warp-uniform address */) {
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;       // four independent chains: the FMA latency, not the count, bounds this
#pragma unroll
  for (int d = 0; d < WD; d += 4) {
    const float4 k = *reinterpret_cast<const float4*>(s + d);
    a0 = fmaf(a[d], k.x, a0); a1 = fmaf(a[d + 1], k.y, a1); a2 = fmaf(a[d + 2], k.z, a2); a3 = fmaf(a[d + 3], k.w, a3);
  }
  return (a0 + a1) + (a2 + a3);
}

// ------------------------------------------------------------------------------------------ forward
// CTA = 2 groups of 64 threads, one (window, head) pair each
template <typename T>
__global__ void __launch_bounds__(2 * WG)
wattn_fwd_kernel(const T* __restrict__ qkv, const float* __restrict__ bias, const float* __restrict__ mask, T* __restrict__ out,
                 float* __restrict__ lse, long long pairs, int nW, int heads, float scale) {
  __shared__ __align__(16) float sK[2][WT * WD];
  __shared__ __align__(16) float sV[2][WT * WD];
  const int grp = threadIdx.x / WG, i = threadIdx.x % WG;
  const long long pair = blockIdx.x * 2LL + grp;
  const bool live = pair < pairs;
  const long long w = live ? pair / heads : 0;
  const int h = live ? static_cast<int>(pair % heads) : 0;
  const int C = heads * WD;
  const long long rs = 3LL * C;
  const T* base = qkv + w * WT * rs + h * WD;
  if (live) {
    stage_tile<T>(base + C, rs, sK[grp], i);
    stage_tile<T>(base + 2 * C, rs, sV[grp], i);
  }
  __syncthreads();
  if (!live || i >= WT) return;
  float q[WD];
  load_row32<T>(base + i * rs, q);
#pragma unroll
  for (int d = 0; d < WD; ++d) q[d] *= scale;
  const float* brow = bias + (static_cast<long long>(h) * WT + i) * WT;
  const float* mrow = mask ? mask + ((w % nW) * WT + i) * WT : nullptr;
  float s[WT];
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < WT; ++j) {
    float v = dot32(q, sK[grp] + j * WD) + __ldg(brow + j);
    if (mrow) v += __ldg(mrow + j);
    s[j] = v;
    m = fmaxf(m, v);
  }
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < WT; ++j) { s[j] = __expf(s[j] - m); l += s[j]; }
  const float inv = 1.f / l;
  float o[WD];
#pragma unroll
  for (int d = 0; d < WD; ++d) o[d] = 0.f;
#pragma unroll
  for (int j = 0; j < WT; ++j) {
    const float p = s[j] * inv;
    const float* vr = sV[grp] + j * WD;
#pragma unroll
    for (int d = 0; d < WD; d += 4) {
      const float4 vv = *reinterpret_cast<const float4*>(vr + d);
      o[d] = fmaf(p, vv.x, o[d]); o[d + 1] = fmaf(p, vv.y, o[d + 1]); o[d + 2] = fmaf(p, vv.z, o[d + 2]); o[d + 3] = fmaf(p, vv.w, o[d + 3]);
    }
  }
  store_row32<T>(out + (w * WT + i) * C + h * WD, o);
  lse[(w * heads + h) * WT + i] = m + __logf(l);
}

// ------------------------------------------------------------------------------------------ backward
// CTA = 64 threads, head h = blockIdx.y, windows blockIdx.x, blockIdx.x + gridDim.x, ...
template <typename T>
__global__ void __launch_bounds__(WG)
wattn_bwd_kernel(const T* __restrict__ qkv, const float* __restrict__ bias, const float* __restrict__ mask, const T* __restrict__ out,
                 const T* __restrict__ dout, const float* __restrict__ lse, T* __restrict__ dqkv, float* __restrict__ dbias,
                 long long Bw, int nW, int heads, float scale) {
  extern __shared__ __align__(16) float sm[];
  float* sK = sm;                      // [49][32]
  float* sV = sK + WT * WD;
  float* sQ = sV + WT * WD;
  float* sDO = sQ + WT * WD;
  float* sP = sDO + WT * WD;           // [49][49]  P   (row = query)
  float* sDS = sP + WT * WT;           // [49][49]  dS
  const int i = threadIdx.x;
  const int h = blockIdx.y;
  const int C = heads * WD;
  const long long rs = 3LL * C;
  float dbacc[WT];
#pragma unroll
  for (int j = 0; j < WT; ++j) dbacc[j] = 0.f;
  const float* brow = bias + (static_cast<long long>(h) * WT + (i < WT ? i : 0)) * WT;

  for (long long w = blockIdx.x; w < Bw; w += gridDim.x) {
    const T* base = qkv + w * WT * rs + h * WD;
    stage_tile<T>(base, rs, sQ, i);
    stage_tile<T>(base + C, rs, sK, i);
    stage_tile<T>(base + 2 * C, rs, sV, i);
    stage_tile<T>(dout + w * WT * C + h * WD, C, sDO, i);
    __syncthreads();
    if (i < WT) {
      float q[WD], g[WD], o[WD], dq[WD];
#pragma unroll
      for (int d = 0; d < WD; ++d) { q[d] = sQ[i * WD + d] * scale; g[d] = sDO[i * WD + d]; dq[d] = 0.f; }
      load_row32<T>(out + (w * WT + i) * C + h * WD, o);
      float delta = 0.f;
#pragma unroll
      for (int d = 0; d < WD; ++d) delta = fmaf(g[d], o[d], delta);
      const float li = lse[(w * heads + h) * WT + i];
      const float* mrow = mask ? mask + ((w % nW) * WT + i) * WT : nullptr;
#pragma unroll
      for (int j = 0; j < WT; ++j) {
        float sv = dot32(q, sK + j * WD) + __ldg(brow + j);
        if (mrow) sv += __ldg(mrow + j);
        const float p = __expf(sv - li);
        const float dp = dot32(g, sV + j * WD);
        const float ds = p * (dp - delta);
        dbacc[j] += ds;
        sP[i * WT + j] = p;
        sDS[i * WT + j] = ds;
        const float* kr = sK + j * WD;
#pragma unroll
        for (int d = 0; d < WD; d += 4) {
          const float4 kk = *reinterpret_cast<const float4*>(kr + d);
          dq[d] = fmaf(ds, kk.x, dq[d]); dq[d + 1] = fmaf(ds, kk.y, dq[d + 1]); dq[d + 2] = fmaf(ds, kk.z, dq[d + 2]); dq[d + 3] = fmaf(ds, kk.w, dq[d + 3]);
        }
      }
#pragma unroll
      for (int d = 0; d < WD; ++d) dq[d] *= scale;
      store_row32<T>(dqkv + (w * WT + i) * rs + h * WD, dq);
    }
    __syncthreads();
    if (i < WT) {      // thread = key row j: dK_j = scale * sum_i dS_ij q_i,  dV_j = sum_i P_ij dO_i
      float dk[WD], dv[WD];
#pragma unroll
      for (int d = 0; d < WD; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
#pragma unroll 7
      for (int r = 0; r < WT; ++r) {
        const float ds = sDS[r * WT + i], p = sP[r * WT + i];
        const float* qr = sQ + r * WD;
        const float* gr = sDO + r * WD;
#pragma unroll
        for (int d = 0; d < WD; d += 4) {
          const float4 qq = *reinterpret_cast<const float4*>(qr + d);
          const float4 gg = *reinterpret_cast<const float4*>(gr + d);
          dk[d] = fmaf(ds, qq.x, dk[d]); dk[d + 1] = fmaf(ds, qq.y, dk[d + 1]); dk[d + 2] = fmaf(ds, qq.z, dk[d + 2]); dk[d + 3] = fmaf(ds, qq.w, dk[d + 3]);
          dv[d] = fmaf(p, gg.x, dv[d]); dv[d + 1] = fmaf(p, gg.y, dv[d + 1]); dv[d + 2] = fmaf(p, gg.z, dv[d + 2]); dv[d + 3] = fmaf(p, gg.w, dv[d + 3]);
        }
      }
#pragma unroll
      for (int d = 0; d < WD; ++d) dk[d] *= scale;
      T* drow = dqkv + (w * WT + i) * rs + h * WD;
      store_row32<T>(drow + C, dk);
      store_row32<T>(drow + 2 * C, dv);
    }
    __syncthreads();
  }
  if (i < WT) {
    float* db = dbias + (static_cast<long long>(h) * WT + i) * WT;
#pragma unroll
    for (int j = 0; j < WT; ++j) atomicAdd(db + j, dbacc[j]);
  }
}

int wcheck(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
constexpr int BWD_SMEM = (4 * WT * WD + 2 * WT * WT) * static_cast<int>(sizeof(float));

}  // namespace

// 1 (default): tensor-core kernels of window_attn_tc.cu;  0: the one-thread-per-row CUDA-core kernels of this file
static int g_wattn_impl = 1;
extern "C" int b2_window_attn_set_impl(int impl) {
  const int prev = g_wattn_impl;
  if (impl == 0 || impl == 1) g_wattn_impl = impl;
  return prev;
}

extern "C" int b2_window_attn_fwd(const void* qkv, const float* bias, const float* mask, void* out, float* lse, long long Bw, int nW,
                                  int heads, int tokens, int head_dim, float scale, int dtype, void* stream) {
  if (tokens != WT || head_dim != WD) { b2host::set_error("window_attn: only 7x7 windows with head_dim 32 (got %d tokens, head_dim %d)", tokens, head_dim); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g_wattn_impl == 1) return b2host::wattn_tc_fwd(qkv, bias, mask, out, lse, Bw, nW, heads, scale, dtype, st);
  const long long pairs = Bw * heads;
  const unsigned grid = static_cast<unsigned>((pairs + 1) / 2);
  if (dtype == B2_DTYPE_BF16)
    wattn_fwd_kernel<__nv_bfloat16><<<grid, 2 * WG, 0, st>>>((const __nv_bfloat16*)qkv, bias, mask, (__nv_bfloat16*)out, lse, pairs, nW > 0 ? nW : 1, heads, scale);
  else
    wattn_fwd_kernel<__half><<<grid, 2 * WG, 0, st>>>((const __half*)qkv, bias, mask, (__half*)out, lse, pairs, nW > 0 ? nW : 1, heads, scale);
  return wcheck("window_attn_fwd");
}

// dbias (fp32 [heads, 49, 49]) is ACCUMULATED into (zero it for a plain gradient); dqkv has the layout of qkv
extern "C" int b2_window_attn_bwd(const void* qkv, const float* bias, const float* mask, const void* out, const void* dout,
                                  const float* lse, void* dqkv, float* dbias, long long Bw, int nW, int heads, int tokens, int head_dim,
                                  float scale, int dtype, void* stream) {
  if (tokens != WT || head_dim != WD) { b2host::set_error("window_attn: only 7x7 windows with head_dim 32 (got %d tokens, head_dim %d)", tokens, head_dim); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g_wattn_impl == 1) return b2host::wattn_tc_bwd(qkv, bias, mask, out, dout, lse, dqkv, dbias, Bw, nW, heads, scale, dtype, st);
  long long per_head = (static_cast<long long>(b2host::num_sms()) * 8 + heads - 1) / heads;
  if (per_head > Bw) per_head = Bw;
  if (per_head < 1) per_head = 1;
  dim3 grid(static_cast<unsigned>(per_head), static_cast<unsigned>(heads));
  static bool cfg = false;
  if (!cfg) {
    cudaFuncSetAttribute(wattn_bwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    cudaFuncSetAttribute(wattn_bwd_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM);
    cfg = true;
  }
  if (dtype == B2_DTYPE_BF16)
    wattn_bwd_kernel<__nv_bfloat16><<<grid, WG, BWD_SMEM, st>>>((const __nv_bfloat16*)qkv, bias, mask, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout,
                                                                lse, (__nv_bfloat16*)dqkv, dbias, Bw, nW > 0 ? nW : 1, heads, scale);
  else
    wattn_bwd_kernel<__half><<<grid, WG, BWD_SMEM, st>>>((const __half*)qkv, bias, mask, (const __half*)out, (const __half*)dout, lse, (__half*)dqkv,
                                                         dbias, Bw, nW > 0 ? nW : 1, heads, scale);
  return wcheck("window_attn_bwd");
}

// ------------------------------------------------------------------------------------------
// cyclic shift + window partition (and its inverse) in one pass.
//   partition: out[b, wy, wx, ty, tx, :] = in[b, (wy*ws + ty + sh) % H, (wx*ws + tx + sw) % W, :]     (= roll(-s) then partition)
//   reverse  : out[b, y, x, :]           = in[b, wy, wx, ty, tx, :]  with (y', x') = ((y - sh) % H, (x - sw) % W)   (= unpartition then roll(+s))
// replaces: torch.roll + view/permute/reshape copies around every Swin attention (two passes over the activations each way);
// the two maps are inverse permutations, so each is the other's backward.
// ------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256)
window_shuffle_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int B, int H, int W, int vecC, int ws, int sh, int sw,
                      int reverse) {
  const int nwx = W / ws, nwy = H / ws;
  const long long total = static_cast<long long>(B) * H * W * vecC;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % vecC);
    long long row = idx / vecC;                       // OUTPUT row
    long long src;
    if (!reverse) {
      const int tx = static_cast<int>(row % ws); row /= ws;
      const int ty = static_cast<int>(row % ws); row /= ws;
      const int wx = static_cast<int>(row % nwx); row /= nwx;
      const int wy = static_cast<int>(row % nwy); row /= nwy;      // row = b
      int y = wy * ws + ty + sh; if (y >= H) y -= H;
      int x = wx * ws + tx + sw; if (x >= W) x -= W;
      src = (row * H + y) * W + x;
    } else {
      const int x = static_cast<int>(row % W); row /= W;
      const int y = static_cast<int>(row % H); row /= H;           // row = b
      int yp = y - sh; if (yp < 0) yp += H;
      int xp = x - sw; if (xp < 0) xp += W;
      src = (((row * nwy + yp / ws) * nwx + xp / ws) * ws + yp % ws) * ws + xp % ws;
    }
    out[idx] = in[src * vecC + v];
  }
}
}  // namespace

// ------------------------------------------------------------------------------------------
// residual add with a per-sample scale (stochastic depth) and, optionally, the window maps above folded in:
//   out[r, :] = res[r, :] + scale[b(r)] * x[src(r), :]          r = OUTPUT row, b(r) = r / (H * W) = its sample
//   map 0: src(r) = r;   map 1: out in window order, x in image order (partition);   map 2: out in image order, x in window
//   order (reverse).  res == NULL: no residual;  scale == NULL: 1.
// replaces: torchvision's `x + stochastic_depth(branch)` (a broadcast multiply that ATen runs unvectorised, then an add) and
// the window_reverse pass before it; its backward (d branch = scale * dout through the inverse map) is the same kernel.
// ------------------------------------------------------------------------------------------
namespace {
template <typename T>
__global__ void __launch_bounds__(256)
rows_scale_add_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res, const float* __restrict__ scale, uint4* __restrict__ out,
                      int B, int H, int W, int vecC, int ws, int sh, int sw, int map) {
  const int nwx = W / ws, nwy = H / ws;
  const long long per_sample = static_cast<long long>(H) * W;
  const long long total = static_cast<long long>(B) * per_sample * vecC;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % vecC);
    const long long orow = idx / vecC;                // OUTPUT row
    long long row = orow;
    long long src = orow;
    if (map == 1) {
      const int tx = static_cast<int>(row % ws); row /= ws;
      const int ty = static_cast<int>(row % ws); row /= ws;
      const int wx = static_cast<int>(row % nwx); row /= nwx;
      const int wy = static_cast<int>(row % nwy); row /= nwy;      // row = b
      int y = wy * ws + ty + sh; if (y >= H) y -= H;
      int xx = wx * ws + tx + sw; if (xx >= W) xx -= W;
      src = (row * H + y) * W + xx;
    } else if (map == 2) {
      const int xx = static_cast<int>(row % W); row /= W;
      const int y = static_cast<int>(row % H); row /= H;           // row = b
      int yp = y - sh; if (yp < 0) yp += H;
      int xp = xx - sw; if (xp < 0) xp += W;
      src = (((row * nwy + yp / ws) * nwx + xp / ws) * ws + yp % ws) * ws + xp % ws;
    }
    const float sc = scale ? __ldg(scale + orow / per_sample) : 1.f;
    const uint4 q = x[src * vecC + v];
    const uint4 r = res ? res[idx] : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t qi[4] = {q.x, q.y, q.z, q.w}, ri[4] = {r.x, r.y, r.z, r.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 a = WCvt<T>::up(qi[k]), b = WCvt<T>::up(ri[k]);
      o[k] = WCvt<T>::down(fmaf(sc, a.x, b.x), fmaf(sc, a.y, b.y));
    }
    out[idx] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
}  // namespace

// x / res / out: 16-bit rows of C elements (C % 8 == 0), B * H * W rows; scale fp32 [B] or NULL; map 0 / 1 / 2 as above
// (maps 1 and 2 need H and W multiples of ws and 0 <= shift < extent; map 0 ignores ws and the shifts)
extern "C" int b2_rows_scale_add(const void* x, const void* res, const float* scale, void* out, int B, int H, int W, int C, int ws,
                                 int shift_h, int shift_w, int map, int dtype, void* stream) {
  if ((C % 8) || map < 0 || map > 2 || B <= 0 || H <= 0 || W <= 0) { b2host::set_error("rows_scale_add: need C %% 8 == 0 and map in 0..2"); return -1; }
  if (map != 0 && (ws <= 0 || (H % ws) || (W % ws) || shift_h < 0 || shift_w < 0 || shift_h >= H || shift_w >= W)) {
    b2host::set_error("rows_scale_add: window maps need H and W multiples of the window size, 0 <= shift < extent"); return -1;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * H * W * (C / 8);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(b2host::num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  if (map == 0) ws = 1;
  if (dtype == B2_DTYPE_BF16)
    rows_scale_add_kernel<__nv_bfloat16><<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<const uint4*>(x), static_cast<const uint4*>(res), scale, static_cast<uint4*>(out), B, H, W, C / 8, ws, shift_h, shift_w, map);
  else
    rows_scale_add_kernel<__half><<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<const uint4*>(x), static_cast<const uint4*>(res), scale, static_cast<uint4*>(out), B, H, W, C / 8, ws, shift_h, shift_w, map);
  return wcheck("rows_scale_add");
}

// ------------------------------------------------------------------------------------------
// non-overlapping 4 x 4 patches of an NCHW image as GEMM rows (the patch-embedding convolution, kernel = stride = 4, as one GEMM):
//   out[(b, py, px), (c, i, j)] = x[b, c, 4 py + i, 4 px + j]        16-bit elements, one 8-byte vector (j = 0..3) per thread
// reads are fully coalesced along the image row; the 96-byte output rows fill up in L2 before they are written back.
// replaces: view / permute / contiguous (ATen's generic strided copy, one 2-byte element per thread with 6-D index math)
// ------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256)
patchify4_kernel(const uint2* __restrict__ x, uint2* __restrict__ out, int B, int C, int H, int W) {
  const int pw = W >> 2, ph = H >> 2;
  const long long total = static_cast<long long>(B) * C * H * pw;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    long long t = idx;
    const int px = static_cast<int>(t % pw); t /= pw;
    const int y = static_cast<int>(t % H); t /= H;
    const int c = static_cast<int>(t % C); t /= C;                     // t = b
    const long long row = (t * ph + (y >> 2)) * pw + px;
    out[(row * C + c) * 4 + (y & 3)] = x[idx];
  }
}
}  // namespace

// x [B, C, H, W] -> out [B * (H/4) * (W/4), C * 16]; 16-bit elements, H and W multiples of 4, both 8-byte aligned
extern "C" int b2_patchify4(const void* x, void* out, int B, int C, int H, int W, void* stream) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || (H % 4) || (W % 4)) { b2host::set_error("patchify4: H and W must be multiples of 4"); return -1; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * C * H * (W / 4);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(b2host::num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  patchify4_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(static_cast<const uint2*>(x), static_cast<uint2*>(out), B, C, H, W);
  return wcheck("patchify4");
}

// in / out: 16-bit elements, C % 8 == 0, H and W multiples of ws, 0 <= shift < ws
extern "C" int b2_window_shuffle(const void* in, void* out, int B, int H, int W, int C, int ws, int shift_h, int shift_w, int reverse,
                                 void* stream) {
  if ((C % 8) || ws <= 0 || (H % ws) || (W % ws) || shift_h < 0 || shift_w < 0 || shift_h >= H || shift_w >= W) {
    b2host::set_error("window_shuffle: need C %% 8 == 0, H and W multiples of the window size, 0 <= shift < extent"); return -1;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * H * W * (C / 8);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(b2host::num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  window_shuffle_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(static_cast<const uint4*>(in), static_cast<uint4*>(out), B, H, W, C / 8,
                                                                        ws, shift_h, shift_w, reverse);
  return wcheck("window_shuffle");
}
