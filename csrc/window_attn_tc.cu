// Swin shifted-window attention on the tensor cores, forward and backward (sm_100a).
//
//   per (window, head):  S = scale * q k^T + rel_pos_bias[head] (+ shift mask[window % nW]);  P = softmax(S);  O = P v
//
// replaces: torchvision ShiftedWindowAttention's eager chain and this repo's first CUDA-core kernel
// (window_attn.cu, one thread per query row), which ran at ~13 TFLOP/s - a third of a Swin-B step.
//
// A window is 49 tokens x head_dim 32: one (window, head) problem is a 64 x 64 x 32 matrix product at most, far below a
// tcgen05 tile, and thousands of them are independent.  So a CTA of four warps works on ONE pair at a time with warp-level
// mma.sync (m16n8k16, fp32 accumulate): warp w owns query rows 16 w .. 16 w + 15 (and, in the backward, key rows 16 w ..).
// The q / k / v (/ dO / O) rows of the next window of the same head stream into the other half of a double buffer with
// cp.async while the current one is computed; operands reach the registers with ldmatrix (.trans where the contraction runs
// over rows), scores / probabilities / dS never leave the registers: the C fragments of S are repacked as the A fragments of
// the second product.  The backward needs dS both as [query, key] (for dQ) and as [key, query] (for dK, dV); instead of
// transposing through shared memory it recomputes the 49 x 49 scores in both orientations, which costs 28 more MMAs a warp.
// The relative-position bias of the CTA's head sits in shared memory; d(bias) accumulates in registers over all windows the
// CTA walks and reaches memory with 49 x 49 atomics per CTA.
//
// Layouts: qkv [Bw, 49, 3, heads, 32] (the fused projection's output, Bw = batch * nW windows), out [Bw, 49, heads * 32],
// bias fp32 [heads, 49, 49], mask fp32 [nW, 49, 49] or NULL, lse fp32 [Bw, heads, 49].
#ifdef B2_EMU                      // CPU emulation (tests/emu): kernels only, primitives from mma_emu.h
#include "mma_emu.h"
namespace b2 {}
#else
#include "common.cuh"
#endif
#include "b200dist.h"

using namespace b2;

namespace wtc {

#ifndef B2_EMU
template <typename T> B2_DEVICE void mma16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1);
template <> B2_DEVICE void mma16816<__nv_bfloat16>(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> B2_DEVICE void mma16816<__half>(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// lane l passes the address of row l % 8 of 8x8 matrix l / 8; register i comes from matrix i
B2_DEVICE void ldsm_x4(uint32_t r[4], const void* row_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(row_ptr)) : "memory");
}
B2_DEVICE void ldsm_x4_t(uint32_t r[4], const void* row_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(row_ptr)) : "memory");
}
B2_DEVICE void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
B2_DEVICE void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
B2_DEVICE void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> B2_DEVICE void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <typename T> B2_DEVICE uint32_t pack2(float lo, float hi);
template <> B2_DEVICE uint32_t pack2<__nv_bfloat16>(float lo, float hi) { return pack_bf16x2(lo, hi); }
template <> B2_DEVICE uint32_t pack2<__half>(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <typename T> B2_DEVICE void unpack2(uint32_t u, float& lo, float& hi);
template <> B2_DEVICE void unpack2<__nv_bfloat16>(uint32_t u, float& lo, float& hi) {
  lo = __uint_as_float(u << 16);
  hi = __uint_as_float(u & 0xffff0000u);
}
template <> B2_DEVICE void unpack2<__half>(uint32_t u, float& lo, float& hi) {
  const float2 f = __half22float2(*reinterpret_cast<__half2*>(&u));
  lo = f.x; hi = f.y;
}
#else
using ::mma16816; using ::ldsm_x4; using ::ldsm_x4_t; using ::cp_async16; using ::cp_async4; using ::cp_async_commit; using ::cp_async_wait;
using ::pack2; using ::unpack2;
#endif

constexpr int WT = 49;             // tokens per window
constexpr int WD = 32;             // head dim
constexpr int ROWS = 64;           // tile rows in shared memory (49 real + zero rows up to the MMA tile edge)
constexpr int RSTRIDE = 80;        // bytes per tile row: 64 of data + 16 of padding, so the 8 rows of an ldmatrix hit 8 bank groups
constexpr int TILE_BYTES = ROWS * RSTRIDE;
constexpr int BSTRIDE = 56;        // floats per row of the bias tile
constexpr int BIAS_BYTES = WT * BSTRIDE * 4;
constexpr int THREADS = 128;
constexpr int FWD_TILES = 3, BWD_TILES = 5;       // q k v | q k v dO O
constexpr int FWD_SMEM = BIAS_BYTES + 2 * FWD_TILES * TILE_BYTES;
constexpr int BWD_SMEM = BIAS_BYTES + 3 * ROWS * 4 + 2 * BWD_TILES * TILE_BYTES;     // + lse [2 stages][64], delta [64]

// address lane `lane` passes to ldmatrix for a 16 x 16 block at (row0, col0) of a tile: matrices are
// (rows 0-7, cols 0-7), (rows 8-15, cols 0-7), (rows 0-7, cols 8-15), (rows 8-15, cols 8-15).
// Plain: the A fragment of rows row0.. / k-columns col0..;   .trans: the B fragments (b0, b1) of n-columns col0..col0+7
// (registers 0, 1) and col0+8..col0+15 (registers 2, 3) for a contraction over rows row0..row0+15.
B2_DEVICE const unsigned char* blk16(const unsigned char* tile, int row0, int col0, int lane) {
  return tile + (row0 + (lane & 7) + ((lane >> 3) & 1) * 8) * RSTRIDE + (col0 + (lane >> 4) * 8) * 2;
}
// address for an 8-row x 32-column block at row0: matrices are columns 0-7, 8-15, 16-23, 24-31 of rows row0..row0+7.
// Plain: the B fragments of n = rows row0..row0+7 for a contraction over the 32 columns: (r0, r1) for k = 0..15, (r2, r3) for 16..31
B2_DEVICE const unsigned char* row8(const unsigned char* tile, int row0, int lane) {
  return tile + (row0 + (lane & 7)) * RSTRIDE + (lane >> 3) * 16;
}

// rows [49, 64) of `ntiles` tiles are never written by the copies: zero them once
B2_DEVICE void zero_pad_rows(unsigned char* tiles, int ntiles, int tid) {
  constexpr int WORDS = (ROWS - WT) * RSTRIDE / 4;
  for (int i = tid; i < ntiles * WORDS; i += THREADS) {
    const int t = i / WORDS, o = i - t * WORDS;
    reinterpret_cast<uint32_t*>(tiles + t * TILE_BYTES + WT * RSTRIDE)[o] = 0u;
  }
}
// bias[head] (fp32 [49][49]) -> shared [49][56], padding columns zero.  4-byte asynchronous copies (rows of 49 floats are only
// 4-byte aligned): they join the caller's first copy group, so the prologue costs one memory latency instead of twenty
B2_DEVICE void stage_bias(float* sBias, const float* bias_h, int tid) {
  for (int i = tid; i < WT * BSTRIDE; i += THREADS) {
    const int r = i / BSTRIDE, c = i - r * BSTRIDE;
    if (c < WT) cp_async4(sBias + i, bias_h + r * WT + c);
    else sBias[i] = 0.f;
  }
}
// 49 rows x 64 bytes of one (window, part, head) -> tile rows (16-byte asynchronous copies, 4 per row)
template <typename T>
B2_DEVICE void issue_rows(unsigned char* tile, const T* src, long long row_stride, int tid) {
  for (int v = tid; v < WT * 4; v += THREADS) {
    const int row = v >> 2, c = v & 3;
    cp_async16(tile + row * RSTRIDE + c * 16, src + row * row_stride + c * 8);
  }
}
B2_DEVICE float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
B2_DEVICE float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// ------------------------------------------------------------------------------------------ forward
// grid (x, heads): CTA (x, h) walks windows x, x + gridDim.x, ... of head h
template <typename T>
__global__ void __launch_bounds__(THREADS, 4)
wattn_tc_fwd_kernel(const T* __restrict__ qkv, const float* __restrict__ bias, const float* __restrict__ mask, T* __restrict__ out,
                    float* __restrict__ lse, long long Bw, int nW, int heads, float scale) {
#ifdef B2_EMU
  B2_DYN_SMEM(unsigned char, smem);
#else
  extern __shared__ __align__(128) unsigned char smem[];
#endif
  float* sBias = reinterpret_cast<float*>(smem);
  unsigned char* tiles = smem + BIAS_BYTES;                    // [2 stages][q k v]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int h = blockIdx.y;
  const int C = heads * WD;
  const long long rs = 3LL * C;

  auto issue = [&](long long w, int s) {
    const T* base = qkv + w * WT * rs + h * WD;
    unsigned char* st = tiles + s * FWD_TILES * TILE_BYTES;
#pragma unroll
    for (int part = 0; part < 3; ++part) issue_rows<T>(st + part * TILE_BYTES, base + part * C, rs, tid);
    cp_async_commit();
  };
  long long w = blockIdx.x;
  stage_bias(sBias, bias + static_cast<long long>(h) * WT * WT, tid);      // part of the first copy group
  if (w < Bw) issue(w, 0);
  zero_pad_rows(tiles, 2 * FWD_TILES, tid);
  const float inv_scale = 1.f / scale;

  const int r0 = warp * 16 + g, r1 = r0 + 8;                   // this thread's two query rows
  const bool v0 = r0 < WT, v1 = r1 < WT;
  const float* b0p = sBias + (v0 ? r0 : WT - 1) * BSTRIDE;
  const float* b1p = sBias + (v1 ? r1 : WT - 1) * BSTRIDE;

  for (int s = 0; w < Bw; w += gridDim.x, s ^= 1) {
    const long long wn = w + gridDim.x;
    if (wn < Bw) { issue(wn, s ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const unsigned char* sQ = tiles + s * FWD_TILES * TILE_BYTES;
    const unsigned char* sK = sQ + TILE_BYTES;
    const unsigned char* sV = sK + TILE_BYTES;

    // S / scale = (bias + mask) / scale + q k^T for this warp's 16 rows: the accumulators START at the additive terms (their
    // loads run ahead of the MMAs), 7 key tiles of 8, contraction over the 32 dims in two k-steps
    const float* m0p = mask ? mask + ((w % nW) * WT + (v0 ? r0 : WT - 1)) * WT : nullptr;
    const float* m1p = mask ? mask + ((w % nW) * WT + (v1 ? r1 : WT - 1)) * WT : nullptr;
    float sc[7][4];
#pragma unroll
    for (int nt = 0; nt < 7; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int c = nt * 8 + t4 * 2 + e;
        float a0 = b0p[c], a1 = b1p[c];
        if (mask && c < WT) { a0 += __ldg(m0p + c); a1 += __ldg(m1p + c); }
        sc[nt][e] = a0 * inv_scale; sc[nt][2 + e] = a1 * inv_scale;
      }
    }
    uint32_t qa[2][4];
    ldsm_x4(qa[0], blk16(sQ, warp * 16, 0, lane));
    ldsm_x4(qa[1], blk16(sQ, warp * 16, 16, lane));
#pragma unroll
    for (int nt = 0; nt < 7; ++nt) {
      uint32_t kb[4];
      ldsm_x4(kb, row8(sK, nt * 8, lane));
      mma16816<T>(sc[nt], qa[0], kb[0], kb[1]);
      mma16816<T>(sc[nt], qa[1], kb[2], kb[3]);
    }
    // scale, row max (key columns 49..55 do not exist)
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 7; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int c = nt * 8 + t4 * 2 + e;
        const float s0 = c < WT ? sc[nt][e] * scale : -INFINITY;
        const float s1 = c < WT ? sc[nt][2 + e] * scale : -INFINITY;
        sc[nt][e] = s0; sc[nt][2 + e] = s1;
        mx0 = fmaxf(mx0, s0); mx1 = fmaxf(mx1, s1);
      }
    }
    mx0 = quad_max(mx0); mx1 = quad_max(mx1);
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 7; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float p0 = __expf(sc[nt][e] - mx0), p1 = __expf(sc[nt][2 + e] - mx1);
        sc[nt][e] = p0; sc[nt][2 + e] = p1;
        l0 += p0; l1 += p1;
      }
    }
    l0 = quad_sum(l0); l1 = quad_sum(l1);

    // O = P v: the C fragments of key tiles 2 kt, 2 kt + 1 are the A fragment of key block kt (keys 56..63 do not exist)
    float o[4][4];
#pragma unroll
    for (int dn = 0; dn < 4; ++dn) o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f;
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
      uint32_t pa[4];
      pa[0] = pack2<T>(sc[2 * kt][0], sc[2 * kt][1]);
      pa[1] = pack2<T>(sc[2 * kt][2], sc[2 * kt][3]);
      if (kt < 3) {
        pa[2] = pack2<T>(sc[2 * kt + 1][0], sc[2 * kt + 1][1]);
        pa[3] = pack2<T>(sc[2 * kt + 1][2], sc[2 * kt + 1][3]);
      } else {
        pa[2] = 0u; pa[3] = 0u;
      }
#pragma unroll
      for (int dp = 0; dp < 2; ++dp) {
        uint32_t vb[4];
        ldsm_x4_t(vb, blk16(sV, kt * 16, dp * 16, lane));
        mma16816<T>(o[2 * dp], pa, vb[0], vb[1]);
        mma16816<T>(o[2 * dp + 1], pa, vb[2], vb[3]);
      }
    }
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    T* orow0 = out + (w * WT + r0) * C + h * WD + t4 * 2;
    T* orow1 = out + (w * WT + r1) * C + h * WD + t4 * 2;
#pragma unroll
    for (int dn = 0; dn < 4; ++dn) {
      if (v0) *reinterpret_cast<uint32_t*>(orow0 + dn * 8) = pack2<T>(o[dn][0] * i0, o[dn][1] * i0);
      if (v1) *reinterpret_cast<uint32_t*>(orow1 + dn * 8) = pack2<T>(o[dn][2] * i1, o[dn][3] * i1);
    }
    if (t4 == 0) {
      float* lp = lse + (w * heads + h) * WT;
      if (v0) lp[r0] = mx0 + __logf(l0);
      if (v1) lp[r1] = mx1 + __logf(l1);
    }
    __syncthreads();             // every warp is done with stage s before the next iteration's copies land in it
  }
}

// ------------------------------------------------------------------------------------------ backward
//   P = exp(S - lse),  dP = dO v^T,  delta_i = dO_i . O_i,  dS = P o (dP - delta),
//   dQ = scale dS k,   dK = scale dS^T q,   dV = P^T dO,   d(bias) += dS
// grid (x, heads) as in the forward.  dbias is accumulated INTO (atomics).
template <typename T>
__global__ void __launch_bounds__(THREADS, 3)
wattn_tc_bwd_kernel(const T* __restrict__ qkv, const float* __restrict__ bias, const float* __restrict__ mask, const T* __restrict__ out,
                    const T* __restrict__ dout, const float* __restrict__ lse, T* __restrict__ dqkv, float* __restrict__ dbias,
                    long long Bw, int nW, int heads, float scale) {
#ifdef B2_EMU
  B2_DYN_SMEM(unsigned char, smem);
#else
  extern __shared__ __align__(128) unsigned char smem[];
#endif
  float* sBias = reinterpret_cast<float*>(smem);
  float* sLse2 = reinterpret_cast<float*>(smem + BIAS_BYTES);          // [2 stages][64] (0 beyond row 48)
  float* sDelta = sLse2 + 2 * ROWS;                                    // [64]
  unsigned char* tiles = smem + BIAS_BYTES + 3 * ROWS * 4;             // [2 stages][q k v dO O]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int h = blockIdx.y;
  const int C = heads * WD;
  const long long rs = 3LL * C;

  auto issue = [&](long long w, int s) {
    const T* base = qkv + w * WT * rs + h * WD;
    unsigned char* st = tiles + s * BWD_TILES * TILE_BYTES;
#pragma unroll
    for (int part = 0; part < 3; ++part) issue_rows<T>(st + part * TILE_BYTES, base + part * C, rs, tid);
    issue_rows<T>(st + 3 * TILE_BYTES, dout + w * WT * C + h * WD, C, tid);
    issue_rows<T>(st + 4 * TILE_BYTES, out + w * WT * C + h * WD, C, tid);
    if (tid < WT) cp_async4(sLse2 + s * ROWS + tid, lse + (w * heads + h) * WT + tid);
    cp_async_commit();
  };
  long long w = blockIdx.x;
  stage_bias(sBias, bias + static_cast<long long>(h) * WT * WT, tid);      // part of the first copy group
  if (w < Bw) issue(w, 0);
  zero_pad_rows(tiles, 2 * BWD_TILES, tid);
  if (tid >= WT && tid < ROWS) { sLse2[tid] = 0.f; sLse2[ROWS + tid] = 0.f; }
  const float inv_scale = 1.f / scale;

  const int r0 = warp * 16 + g, r1 = r0 + 8;       // this thread's two rows: query rows in the first half, key rows in the second
  const bool v0 = r0 < WT, v1 = r1 < WT;
  const int rc0 = v0 ? r0 : WT - 1, rc1 = v1 ? r1 : WT - 1;            // clamped, for addresses only
  float dbacc[7][4];
#pragma unroll
  for (int nt = 0; nt < 7; ++nt) dbacc[nt][0] = dbacc[nt][1] = dbacc[nt][2] = dbacc[nt][3] = 0.f;

  for (int s = 0; w < Bw; w += gridDim.x, s ^= 1) {
    const long long wn = w + gridDim.x;
    if (wn < Bw) { issue(wn, s ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const unsigned char* sQ = tiles + s * BWD_TILES * TILE_BYTES;
    const unsigned char* sK = sQ + TILE_BYTES;
    const unsigned char* sV = sK + TILE_BYTES;
    const unsigned char* sG = sV + TILE_BYTES;      // dO
    const unsigned char* sO = sG + TILE_BYTES;
    const float* sLse = sLse2 + s * ROWS;

    // delta_i = dO_i . O_i (two threads per row, 16 dims each) -> shared
    {
      const int row = tid >> 1, half = tid & 1;
      float part = 0.f;
      if (row < WT) {
        const uint32_t* gp = reinterpret_cast<const uint32_t*>(sG + row * RSTRIDE + half * 32);
        const uint32_t* op = reinterpret_cast<const uint32_t*>(sO + row * RSTRIDE + half * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float g0, g1, o0, o1;
          unpack2<T>(gp[i], g0, g1);
          unpack2<T>(op[i], o0, o1);
          part = fmaf(g0, o0, part);
          part = fmaf(g1, o1, part);
        }
      }
      part += __shfl_xor_sync(0xffffffffu, part, 1);
      if (half == 0) sDelta[row] = part;                               // rows 49..63: 0
    }
    __syncthreads();
    const float* mbase = mask ? mask + (w % nW) * WT * WT : nullptr;

    // ---- query orientation: rows = queries 16 warp ..; dS -> dQ, d(bias)
    {
      uint32_t qa[2][4], ga[2][4];
      ldsm_x4(qa[0], blk16(sQ, warp * 16, 0, lane));
      ldsm_x4(qa[1], blk16(sQ, warp * 16, 16, lane));
      ldsm_x4(ga[0], blk16(sG, warp * 16, 0, lane));
      ldsm_x4(ga[1], blk16(sG, warp * 16, 16, lane));
      const float lse0 = sLse[r0], lse1 = sLse[r1], dl0 = sDelta[r0], dl1 = sDelta[r1];
      const float* b0p = sBias + rc0 * BSTRIDE;
      const float* b1p = sBias + rc1 * BSTRIDE;
      const float* m0p = mbase ? mbase + rc0 * WT : nullptr;
      const float* m1p = mbase ? mbase + rc1 * WT : nullptr;
      float dq[4][4];
#pragma unroll
      for (int dn = 0; dn < 4; ++dn) dq[dn][0] = dq[dn][1] = dq[dn][2] = dq[dn][3] = 0.f;
#pragma unroll
      for (int kt = 0; kt < 4; ++kt) {             // key block of 16 = key tiles 2 kt, 2 kt + 1
        uint32_t da[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int nt = 2 * kt + j;
          if (nt < 7) {
            float sa[4], pa[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int e = 0; e < 2; ++e) {          // S / scale starts at (bias + mask) / scale
              const int c = nt * 8 + t4 * 2 + e;
              float a0 = b0p[c], a1 = b1p[c];
              if (mbase && c < WT) { a0 += __ldg(m0p + c); a1 += __ldg(m1p + c); }
              sa[e] = a0 * inv_scale; sa[2 + e] = a1 * inv_scale;
            }
            uint32_t kb[4], vb[4];
            ldsm_x4(kb, row8(sK, nt * 8, lane));
            mma16816<T>(sa, qa[0], kb[0], kb[1]);
            mma16816<T>(sa, qa[1], kb[2], kb[3]);
            ldsm_x4(vb, row8(sV, nt * 8, lane));
            mma16816<T>(pa, ga[0], vb[0], vb[1]);
            mma16816<T>(pa, ga[1], vb[2], vb[3]);
            float ds[4];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int c = nt * 8 + t4 * 2 + e;
              const float d0 = __expf(fmaf(sa[e], scale, -lse0)) * (pa[e] - dl0);
              const float d1 = __expf(fmaf(sa[2 + e], scale, -lse1)) * (pa[2 + e] - dl1);
              ds[e] = (c < WT && v0) ? d0 : 0.f;
              ds[2 + e] = (c < WT && v1) ? d1 : 0.f;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) dbacc[nt][e] += ds[e];
            da[2 * j] = pack2<T>(ds[0], ds[1]);
            da[2 * j + 1] = pack2<T>(ds[2], ds[3]);
          }
        }
#pragma unroll
        for (int dp = 0; dp < 2; ++dp) {           // dQ += dS[:, key block] k[key block, :]
          uint32_t kb[4];
          ldsm_x4_t(kb, blk16(sK, kt * 16, dp * 16, lane));
          mma16816<T>(dq[2 * dp], da, kb[0], kb[1]);
          mma16816<T>(dq[2 * dp + 1], da, kb[2], kb[3]);
        }
      }
      T* d0p = dqkv + (w * WT + r0) * rs + h * WD + t4 * 2;
      T* d1p = dqkv + (w * WT + r1) * rs + h * WD + t4 * 2;
#pragma unroll
      for (int dn = 0; dn < 4; ++dn) {
        if (v0) *reinterpret_cast<uint32_t*>(d0p + dn * 8) = pack2<T>(dq[dn][0] * scale, dq[dn][1] * scale);
        if (v1) *reinterpret_cast<uint32_t*>(d1p + dn * 8) = pack2<T>(dq[dn][2] * scale, dq[dn][3] * scale);
      }
    }

    // ---- key orientation: rows = keys 16 warp ..; columns = queries.  S^T = k q^T, dP^T = v dO^T -> dK, dV
    {
      uint32_t ka[2][4], va[2][4];
      ldsm_x4(ka[0], blk16(sK, warp * 16, 0, lane));
      ldsm_x4(ka[1], blk16(sK, warp * 16, 16, lane));
      ldsm_x4(va[0], blk16(sV, warp * 16, 0, lane));
      ldsm_x4(va[1], blk16(sV, warp * 16, 16, lane));
      float dk[4][4], dv[4][4];
#pragma unroll
      for (int dn = 0; dn < 4; ++dn) {
        dk[dn][0] = dk[dn][1] = dk[dn][2] = dk[dn][3] = 0.f;
        dv[dn][0] = dv[dn][1] = dv[dn][2] = dv[dn][3] = 0.f;
      }
#pragma unroll
      for (int qt = 0; qt < 4; ++qt) {             // query block of 16 = query tiles 2 qt, 2 qt + 1
        uint32_t pt[4] = {0u, 0u, 0u, 0u}, dt[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int nt = 2 * qt + j;
          if (nt < 7) {
            float sa[4], pa[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int e = 0; e < 2; ++e) {          // S^T / scale starts at (bias + mask)[query][key] / scale
              const int q = nt * 8 + t4 * 2 + e;
              const int qc = q < WT ? q : WT - 1;
              float a0 = sBias[qc * BSTRIDE + rc0], a1 = sBias[qc * BSTRIDE + rc1];
              if (mbase) { a0 += __ldg(mbase + qc * WT + rc0); a1 += __ldg(mbase + qc * WT + rc1); }
              sa[e] = a0 * inv_scale; sa[2 + e] = a1 * inv_scale;
            }
            uint32_t qb[4], gb[4];
            ldsm_x4(qb, row8(sQ, nt * 8, lane));
            mma16816<T>(sa, ka[0], qb[0], qb[1]);
            mma16816<T>(sa, ka[1], qb[2], qb[3]);
            ldsm_x4(gb, row8(sG, nt * 8, lane));
            mma16816<T>(pa, va[0], gb[0], gb[1]);
            mma16816<T>(pa, va[1], gb[2], gb[3]);
            float p[4], ds[4];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int q = nt * 8 + t4 * 2 + e;                       // query index (column)
              const float lq = sLse[q], dq_ = sDelta[q];
              const float p0 = __expf(fmaf(sa[e], scale, -lq)), p1 = __expf(fmaf(sa[2 + e], scale, -lq));
              const bool ok0 = q < WT && v0, ok1 = q < WT && v1;
              p[e] = ok0 ? p0 : 0.f;
              p[2 + e] = ok1 ? p1 : 0.f;
              ds[e] = ok0 ? p0 * (pa[e] - dq_) : 0.f;
              ds[2 + e] = ok1 ? p1 * (pa[2 + e] - dq_) : 0.f;
            }
            pt[2 * j] = pack2<T>(p[0], p[1]);
            pt[2 * j + 1] = pack2<T>(p[2], p[3]);
            dt[2 * j] = pack2<T>(ds[0], ds[1]);
            dt[2 * j + 1] = pack2<T>(ds[2], ds[3]);
          }
        }
#pragma unroll
        for (int dp = 0; dp < 2; ++dp) {           // dV += P^T[:, query block] dO[query block, :],  dK += dS^T[...] q[...]
          uint32_t gb[4], qb[4];
          ldsm_x4_t(gb, blk16(sG, qt * 16, dp * 16, lane));
          mma16816<T>(dv[2 * dp], pt, gb[0], gb[1]);
          mma16816<T>(dv[2 * dp + 1], pt, gb[2], gb[3]);
          ldsm_x4_t(qb, blk16(sQ, qt * 16, dp * 16, lane));
          mma16816<T>(dk[2 * dp], dt, qb[0], qb[1]);
          mma16816<T>(dk[2 * dp + 1], dt, qb[2], qb[3]);
        }
      }
      T* d0p = dqkv + (w * WT + r0) * rs + h * WD + t4 * 2;
      T* d1p = dqkv + (w * WT + r1) * rs + h * WD + t4 * 2;
#pragma unroll
      for (int dn = 0; dn < 4; ++dn) {
        if (v0) {
          *reinterpret_cast<uint32_t*>(d0p + C + dn * 8) = pack2<T>(dk[dn][0] * scale, dk[dn][1] * scale);
          *reinterpret_cast<uint32_t*>(d0p + 2 * C + dn * 8) = pack2<T>(dv[dn][0], dv[dn][1]);
        }
        if (v1) {
          *reinterpret_cast<uint32_t*>(d1p + C + dn * 8) = pack2<T>(dk[dn][2] * scale, dk[dn][3] * scale);
          *reinterpret_cast<uint32_t*>(d1p + 2 * C + dn * 8) = pack2<T>(dv[dn][2], dv[dn][3]);
        }
      }
    }
    __syncthreads();             // stage s (and sLse / sDelta) are free for the next iteration
  }

  float* db = dbias + static_cast<long long>(h) * WT * WT;
#pragma unroll
  for (int nt = 0; nt < 7; ++nt) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = nt * 8 + t4 * 2 + e;
      if (c < WT) {
        if (v0) atomicAdd(db + r0 * WT + c, dbacc[nt][e]);
        if (v1) atomicAdd(db + r1 * WT + c, dbacc[nt][2 + e]);
      }
    }
  }
}

}  // namespace wtc

#ifndef B2_EMU
namespace b2host {

namespace {
// CTAs per head: one wave - as many as are resident at once (`per_sm` from the occupancy query) and no more, so that no CTA
// starts after the others have finished; never more than there are windows
dim3 wattn_grid(long long Bw, int heads, int per_sm) {
  long long per_head = static_cast<long long>(num_sms()) * per_sm / heads;
  if (per_head > Bw) per_head = Bw;
  if (per_head < 1) per_head = 1;
  return dim3(static_cast<unsigned>(per_head), static_cast<unsigned>(heads));
}
int g_grid_override[2] = {0, 0};          // CTAs per SM the grids are sized for (forward, backward); 0 = what the occupancy query says
template <typename K>
int resident_ctas(K kernel, int smem_bytes) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, wtc::THREADS, smem_bytes) != cudaSuccess || n < 1) { cudaGetLastError(); n = 1; }
  return n;
}
int wattn_check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
}  // namespace

int wattn_tc_fwd(const void* qkv, const float* bias, const float* mask, void* out, float* lse, long long Bw, int nW, int heads, float scale,
                 int dtype, cudaStream_t st) {
  static bool cfg = false;
  static int per_sm = 1;
  if (!cfg) {
    cudaFuncSetAttribute(wtc::wattn_tc_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, wtc::FWD_SMEM);
    cudaFuncSetAttribute(wtc::wattn_tc_fwd_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, wtc::FWD_SMEM);
    per_sm = resident_ctas(wtc::wattn_tc_fwd_kernel<__nv_bfloat16>, wtc::FWD_SMEM);
    cfg = true;
  }
  const dim3 grid = wattn_grid(Bw, heads, g_grid_override[0] > 0 ? g_grid_override[0] : per_sm);
  if (dtype == B2_DTYPE_BF16)
    wtc::wattn_tc_fwd_kernel<__nv_bfloat16><<<grid, wtc::THREADS, wtc::FWD_SMEM, st>>>(
        static_cast<const __nv_bfloat16*>(qkv), bias, mask, static_cast<__nv_bfloat16*>(out), lse, Bw, nW > 0 ? nW : 1, heads, scale);
  else
    wtc::wattn_tc_fwd_kernel<__half><<<grid, wtc::THREADS, wtc::FWD_SMEM, st>>>(
        static_cast<const __half*>(qkv), bias, mask, static_cast<__half*>(out), lse, Bw, nW > 0 ? nW : 1, heads, scale);
  return wattn_check("window_attn_fwd (tensor core)");
}

int wattn_tc_bwd(const void* qkv, const float* bias, const float* mask, const void* out, const void* dout, const float* lse, void* dqkv,
                 float* dbias, long long Bw, int nW, int heads, float scale, int dtype, cudaStream_t st) {
  static bool cfg = false;
  static int per_sm = 1;
  if (!cfg) {
    cudaFuncSetAttribute(wtc::wattn_tc_bwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, wtc::BWD_SMEM);
    cudaFuncSetAttribute(wtc::wattn_tc_bwd_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, wtc::BWD_SMEM);
    per_sm = resident_ctas(wtc::wattn_tc_bwd_kernel<__nv_bfloat16>, wtc::BWD_SMEM);
    cfg = true;
  }
  const dim3 grid = wattn_grid(Bw, heads, g_grid_override[1] > 0 ? g_grid_override[1] : per_sm);
  if (dtype == B2_DTYPE_BF16)
    wtc::wattn_tc_bwd_kernel<__nv_bfloat16><<<grid, wtc::THREADS, wtc::BWD_SMEM, st>>>(
        static_cast<const __nv_bfloat16*>(qkv), bias, mask, static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout), lse,
        static_cast<__nv_bfloat16*>(dqkv), dbias, Bw, nW > 0 ? nW : 1, heads, scale);
  else
    wtc::wattn_tc_bwd_kernel<__half><<<grid, wtc::THREADS, wtc::BWD_SMEM, st>>>(
        static_cast<const __half*>(qkv), bias, mask, static_cast<const __half*>(out), static_cast<const __half*>(dout), lse,
        static_cast<__half*>(dqkv), dbias, Bw, nW > 0 ? nW : 1, heads, scale);
  return wattn_check("window_attn_bwd (tensor core)");
}

}  // namespace b2host

// tuning knob: size the forward / backward grids for this many CTAs per SM (0 = the occupancy query's answer, the default)
extern "C" void b2_window_attn_set_grid(int fwd_ctas_per_sm, int bwd_ctas_per_sm) {
  b2host::g_grid_override[0] = fwd_ctas_per_sm;
  b2host::g_grid_override[1] = bwd_ctas_per_sm;
}
#endif  // !B2_EMU
