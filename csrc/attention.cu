// Causal / full flash attention forward + backward on tcgen05 (sm_100a), head_dim 64 or 128.
//
// replaces: Megatron CoreAttention (torch.baddbmm + bmm through cuBLAS) together with the
// fused_kernels scaled_upper_triang_masked_softmax CUDA extension (E19, K15) - the [b*np, s, s] score
// tensor is never materialised - and HF eager attention (K24).
//
// Tensors are addressed through 4-D TMA maps (d, head, batch, seq) with arbitrary strides, so the
// kernels read Q/K/V straight out of the packed [s, b, heads, 3*d] QKV projection and write O in the
// [s, b, heads*d] layout the output projection consumes (no permutes / contiguous copies).
//
// Forward CTA = 128 query rows of one (batch, head):
//   warp 0 TMA producer (Q once, K/V ring) | warp 1 MMA issuer | warp 2 TMEM alloc | warps 4-11 softmax
//   (two warps per 32 query rows, each owning half of the score / output columns).
//   S = Q K^T        tcgen05.mma 128x128x16 (K-major A and B)            -> TMEM  (double buffered)
//   P = softmax      tcgen05.ld (no shuffles; the two half-row maxima meet in smem), exp2, bf16 -> smem (128B swizzle)
//   O_j = P V        tcgen05.mma 128xDx16 (A K-major from smem, B = V MN-major) -> TMEM (double buffered)
//   O  = O*alpha + O_j in registers; the MMA warp runs S_{j+1} and P V_{j-1} while softmax works on block j.
//
// Backward CTA = 128 key/value rows of one (batch, head), looping over query blocks:
//   S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (smem, bf16) -> dV += P^T dO, dK += dS^T Q (TMEM
//   accumulators), dQ_i = dS K (TMEM -> smem -> TMA reduce-add into an fp32 dQ buffer).
//   dS^T is written once and read under two descriptor interpretations (K-major A for dK, MN-major A
//   for dQ) - the two canonical 128B-swizzle layouts coincide byte for byte.
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int BLK = 128;        // query rows / key rows per tile
constexpr int NTHREADS = 384;   // warps 0-3: TMA / MMA / TMEM alloc / idle, warps 4-11: softmax (two per row)
constexpr int SM_WARP0 = 4;
constexpr float LOG2E = 1.4426950408889634f;

// single MUFU.EX2 (no denormal/range branches): inputs here are <= 0 after max subtraction
B2_DEVICE float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct AttnParams {
  int seqlen, batch, heads;
  int causal;
  float scale;          // softmax scale (1/sqrt(d))
  float* lse;           // [batch, heads, seqlen] (natural log units)
  const float* delta;   // bwd: [batch, heads, seqlen]
  // attention dropout (DROP kernels): ONE 32-bit hash decides a PAIR of adjacent keys (k even: low 16 bits, k odd: high 16
  // bits), each half compared with the 16-bit threshold drop_thresh = round(p * 65536) - the forward, whose threads walk
  // along k, hashes half as often (the hash cost 60 % of the forward's time per element); element (b, h, q, k) is kept iff
  // half(lowbias32(seed + q*CQ + (k>>1)*CK + (b*H+h)*CB), k&1) >= drop_thresh; kept values are
  // scaled by inv_keep.  The row sum / LSE use the un-dropped probabilities (dropout acts on softmax output).
  uint32_t drop_thresh;  // round(p * 2^16)
  uint32_t drop_seed;
  const unsigned long long* drop_replay;   // device replay counter mixed into drop_seed (NULL = off), see host.cu
  float inv_keep;        // 1 / (1 - p)
  int dbg_skip;          // profiling hook (b2_attn_set_debug): bit 0 = skip the dQ reduce-add, bit 1 = skip the softmax math
};

// counter-based dropout decision shared by forward and backward (and restated in ops/native.py for the CPU oracle):
// lowbias32 integer hash of a linear combination of the element's coordinates
constexpr uint32_t DROP_CQ = 0x9E3779B1u, DROP_CK = 0x85EBCA77u, DROP_CB = 0xC2B2AE3Du;
B2_DEVICE uint32_t drop_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
  return x;
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <int D>
struct FwdSmem {
  static constexpr int SLABS = D / 64;
  static constexpr int TILE = BLK * D * 2;              // one Q/K/V tile (bytes)
  static constexpr int KV_STAGES = D == 64 ? 3 : 2;
  static constexpr int P_BYTES = BLK * BLK * 2;         // 32 KB
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_KV = TILE;
  static constexpr int OFF_P = OFF_KV + KV_STAGES * 2 * TILE;
  static constexpr int OFF_XCH = OFF_P + 2 * P_BYTES;    // float [2 parity][2 halves][128]: row max / row sum exchange
  static constexpr int OFF_BAR = OFF_XCH + 2 * 2 * BLK * 4;
  // + up to 1 KB of slack for the manual 1024 B round-up of the base.  The dynamic array is declared
  // __align__(1024), which the toolchain honours (the round-up is then a no-op), so the D = 128 layout may trim the
  // slack to stay inside the 227 KB per-CTA limit.
  static constexpr int MAX_SMEM = 227 * 1024;
  static constexpr int TOTAL = (OFF_BAR + 256 + 1024) < MAX_SMEM ? (OFF_BAR + 256 + 1024) : MAX_SMEM;
  static_assert(OFF_BAR + 256 <= MAX_SMEM, "attention forward tile does not fit in shared memory");
};

template <int D, bool DROP>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                const AttnParams p) {
  using L = FwdSmem<D>;
  constexpr int SLABS = L::SLABS;
  constexpr int KV_STAGES = L::KV_STAGES;
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t S_COL = 0;          // 2 x 128
  constexpr uint32_t O_COL = 256;        // 2 x D

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + L::OFF_Q;
  uint8_t* sKV = smem + L::OFF_KV;
  uint8_t* sP = smem + L::OFF_P;
  float* sXch = reinterpret_cast<float*>(smem + L::OFF_XCH);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::OFF_BAR);
  uint64_t* q_full = bars;                   // 1
  uint64_t* kv_full = bars + 1;              // KV_STAGES
  uint64_t* kv_empty = kv_full + KV_STAGES;  // KV_STAGES
  uint64_t* s_full = kv_empty + KV_STAGES;   // 2
  uint64_t* s_empty = s_full + 2;            // 2
  uint64_t* p_full = s_empty + 2;            // 2
  uint64_t* o_full = p_full + 2;             // 2
  uint64_t* o_empty = o_full + 2;            // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nqb = (p.seqlen + BLK - 1) / BLK;
  const int qb = nqb - 1 - blockIdx.x;       // heavy (late) query blocks first
  const int head = blockIdx.y, bat = blockIdx.z;
  const int q0 = qb * BLK;
  const int nkv = p.causal ? (qb + 1) : nqb;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], 256); mbar_init(&p_full[i], 256);
      mbar_init(&o_full[i], 1); mbar_init(&o_empty[i], 256);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, L::TILE);
#pragma unroll
      for (int s = 0; s < SLABS; ++s) tma_load_4d(sQ + s * (BLK * 128), &tmQ, q_full, s * 64, head, bat, q0);
      int stage = 0; uint32_t phase = 0;
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(&kv_empty[stage], phase ^ 1);
        uint8_t* sK = sKV + stage * 2 * L::TILE;
        uint8_t* sV = sK + L::TILE;
        mbar_expect_tx(&kv_full[stage], 2 * L::TILE);
#pragma unroll
        for (int s = 0; s < SLABS; ++s) {
          tma_load_4d(sK + s * (BLK * 128), &tmK, &kv_full[stage], s * 64, head, bat, j * BLK);
          tma_load_4d(sV + s * (BLK * 128), &tmV, &kv_full[stage], s * 64, head, bat, j * BLK);
        }
        if (++stage == KV_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer: S_0, then per j: S_{j+1} followed by P V_j ----
    const uint32_t idesc_s = make_idesc_f16(BLK, BLK, 1, 0, 0);
    const uint32_t idesc_o = make_idesc_f16(BLK, D, 1, 0, 1);
    mbar_wait(q_full, 0);
    auto issue_S = [&](int j) {
      const int stage = j % KV_STAGES;
      const uint32_t ph = (j / KV_STAGES) & 1;
      mbar_wait(&kv_full[stage], ph);
      mbar_wait(&s_empty[j & 1], ((j >> 1) & 1) ^ 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sKV + stage * 2 * L::TILE);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + S_COL + (j & 1) * BLK, make_smem_desc(aQ + off, 16, 1024),
                      make_smem_desc(aK + off, 16, 1024), idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(&s_full[j & 1]);
      }
      __syncwarp();
    };
    auto issue_PV = [&](int j) {
      const int stage = j % KV_STAGES;
      mbar_wait(&p_full[j & 1], (j >> 1) & 1);
      mbar_wait(&o_empty[j & 1], ((j >> 1) & 1) ^ 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t aP = smem_u32(sP + (j & 1) * L::P_BYTES);
        const uint32_t aV = smem_u32(sKV + stage * 2 * L::TILE + L::TILE);
#pragma unroll
        for (int kk = 0; kk < BLK / 16; ++kk) {
          const uint32_t offP = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + O_COL + (j & 1) * D, make_smem_desc(aP + offP, 16, 1024),
                      make_smem_desc(aV + kk * 2048, BLK * 128, 1024), idesc_o, kk > 0 ? 1u : 0u);
        }
        umma_commit(&o_full[j & 1]);
        umma_commit(&kv_empty[stage]);
      }
      __syncwarp();
    };
    issue_S(0);
    for (int j = 0; j < nkv; ++j) {
      if (j + 1 < nkv) issue_S(j + 1);
      issue_PV(j);
    }
  } else if (warp >= SM_WARP0) {
    // ---- softmax / accumulate / epilogue: two warps per query row (column halves) ----
    // Warps 4-7 (g = 0) own score columns 0-63 and output columns 0..D/2-1 of their 32 rows, warps 8-11 (g = 1) the
    // other halves: two softmax warps per SM sub-partition hide each other's TMEM / MUFU latency, and each thread
    // keeps its 64 scores in registers between the max and the exp pass (one TMEM read of S instead of two).
    const int g = (warp - SM_WARP0) >> 2;
    const int ew = (warp - SM_WARP0) & 3;
    const int r = ew * 32 + lane;                         // query row inside the tile
    const uint32_t lane_off = static_cast<uint32_t>(ew * 32) << 16;
    const int qpos = q0 + r;
    const float sl2 = p.scale * LOG2E;
    [[maybe_unused]] const uint32_t drop_seed_r = p.drop_replay == nullptr ? p.drop_seed : p.drop_seed ^ static_cast<uint32_t>((*p.drop_replay * 0x9E3779B97F4A7C15ull) >> 32);
    [[maybe_unused]] const uint32_t drop_base = drop_seed_r + static_cast<uint32_t>(qpos) * DROP_CQ +
                                                static_cast<uint32_t>(bat * p.heads + head) * DROP_CB;
    constexpr int DH = D / 2;                             // output columns per thread
    float m_run = -INFINITY, l_run = 0.f, alpha_prev = 1.f;   // l_run: partial sum over this thread's columns
    float o_acc[DH];
#pragma unroll
    for (int i = 0; i < DH; ++i) o_acc[i] = 0.f;

    auto absorb_O = [&](int j, float alpha) {
      mbar_wait(&o_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < DH / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + lane_off + O_COL + (j & 1) * D + g * DH + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = o_acc[c * 32 + i] * alpha + __uint_as_float(v[i]);
      }
      tc_fence_before();
      mbar_arrive(&o_empty[j & 1]);
    };

    for (int j = 0; j < nkv; ++j) {
      mbar_wait(&s_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t tS = tmem_base + lane_off + S_COL + (j & 1) * BLK + g * 64;
      const int k0 = j * BLK;
      const bool need_mask = (p.causal && j == qb) || (k0 + BLK > p.seqlen);
      uint32_t v[64];
      tmem_ld_32x32b_x32(tS, v);
      tmem_ld_32x32b_x32(tS + 32, v + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_empty[j & 1]);                       // scores are in registers: S_{j+2} may overwrite the buffer
      if (need_mask) {
        // last visible column of this block, relative to this thread's first column
        const int lim = min(p.causal ? qpos : 0x7fffffff, p.seqlen - 1) - k0 - g * 64;
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = (i <= lim) ? v[i] : 0xff800000u;     // -inf
      }
      float m0 = __uint_as_float(v[0]), m1 = __uint_as_float(v[1]), m2 = __uint_as_float(v[2]), m3 = __uint_as_float(v[3]);
#pragma unroll
      for (int i = 4; i < 64; i += 4) {
        m0 = fmaxf(m0, __uint_as_float(v[i])); m1 = fmaxf(m1, __uint_as_float(v[i + 1]));
        m2 = fmaxf(m2, __uint_as_float(v[i + 2])); m3 = fmaxf(m3, __uint_as_float(v[i + 3]));
      }
      // combine the two column halves of the row through shared memory (parity-buffered per block)
      float* xch = sXch + (j & 1) * 2 * BLK;
      xch[g * BLK + r] = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
      named_bar_sync(2, 256);
      const float m_new = fmaxf(m_run, fmaxf(xch[r], xch[BLK + r]));
      const float msc = (m_new == -INFINITY) ? 0.f : m_new * sl2;
      const float alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run * sl2 - msc);
      // probabilities -> bf16 -> swizzled smem (A operand of P V); masked entries are exp2(-inf) = 0
      float rs0 = 0.f, rs1 = 0.f;
      uint8_t* rowp = sP + (j & 1) * L::P_BYTES + g * (BLK * 128) + r * 128;     // 64 columns = one 128 B slab row
      // dropout hash input of this thread's first column; + DROP_CK per column
      [[maybe_unused]] uint32_t hx = drop_base + static_cast<uint32_t>((k0 + g * 64) >> 1) * DROP_CK;     // + DROP_CK per key PAIR
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float e0 = fast_exp2(fmaf(__uint_as_float(v[q * 8 + 2 * e]), sl2, -msc));
          float e1 = fast_exp2(fmaf(__uint_as_float(v[q * 8 + 2 * e + 1]), sl2, -msc));
          rs0 += e0; rs1 += e1;
          if constexpr (DROP) {
            const uint32_t hh = drop_mix(hx);
            e0 = (hh & 0xFFFFu) >= p.drop_thresh ? e0 * p.inv_keep : 0.f;
            e1 = (hh >> 16) >= p.drop_thresh ? e1 * p.inv_keep : 0.f;
            hx += DROP_CK;
          }
          pk[e] = pack_bf16x2(e0, e1);
        }
        *reinterpret_cast<uint4*>(rowp + ((q ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      fence_proxy_async_smem();
      mbar_arrive(&p_full[j & 1]);
      l_run = l_run * alpha + (rs0 + rs1);
      m_run = m_new;
      if (j > 0) absorb_O(j - 1, alpha_prev);
      alpha_prev = alpha;
    }
    absorb_O(nkv - 1, alpha_prev);

    // ---- epilogue: row sum across the two halves, normalise, stage through the (dead) Q tile, TMA store; LSE ----
    float* xch = sXch + (nkv & 1) * 2 * BLK;
    xch[g * BLK + r] = l_run;
    named_bar_sync(2, 256);
    const float l_tot = xch[r] + xch[BLK + r];
    const float inv_l = l_tot > 0.f ? 1.f / l_tot : 0.f;
    if (g == 0 && qpos < p.seqlen && p.lse != nullptr)
      p.lse[(static_cast<size_t>(bat) * p.heads + head) * p.seqlen + qpos] =
          (m_run == -INFINITY) ? -INFINITY : (m_run * p.scale + __logf(l_tot));
    {
      const int col0 = g * DH;                                       // first output column of this thread
      uint8_t* rowp = sQ + (col0 >> 6) * (BLK * 128) + r * 128;
      const int chunk0 = (col0 & 63) >> 3;
#pragma unroll
      for (int q = 0; q < DH / 8; ++q) {
        const float* o = o_acc + q * 8;
        *reinterpret_cast<uint4*>(rowp + (((chunk0 + q) ^ (r & 7)) << 4)) =
            make_uint4(pack_bf16x2(o[0] * inv_l, o[1] * inv_l), pack_bf16x2(o[2] * inv_l, o[3] * inv_l),
                       pack_bf16x2(o[4] * inv_l, o[5] * inv_l), pack_bf16x2(o[6] * inv_l, o[7] * inv_l));
      }
    }
    fence_proxy_async_smem();
    named_bar_sync(1, 256);
    if (g == 0 && r == 0) {
#pragma unroll
      for (int s = 0; s < SLABS; ++s) tma_store_4d(&tmO, sQ + s * (BLK * 128), s * 64, head, bat, q0);
      tma_store_commit();
      tma_store_wait_all<0>();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------
// backward preprocess: delta[b,h,s] = sum_d dO * O
// ------------------------------------------------------------------------------------------
// D / 8 lanes per (s, b, h) row, one 16-byte load of O and of dO per lane (rows of neighbouring heads are
// contiguous, so a warp reads 512 B runs); memory-level parallelism comes from the resident CTAs.
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                  float* __restrict__ delta, int seqlen, int batch, int heads, int D,
                  long long o_ss, long long o_sb, long long o_sh, long long do_ss, long long do_sb,
                  long long do_sh) {
  const int lpr = D >> 3;                                   // lanes per row: 8 (D = 64) or 16 (D = 128)
  const long long gt = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long row = gt / lpr;
  const int sub = static_cast<int>(gt % lpr);
  const long long total = static_cast<long long>(seqlen) * batch * heads;
  float acc = 0.f;
  int h = 0, b = 0, s = 0;
  if (row < total) {
    h = static_cast<int>(row % heads);
    b = static_cast<int>((row / heads) % batch);
    s = static_cast<int>(row / (static_cast<long long>(heads) * batch));
    const uint4 qa = *reinterpret_cast<const uint4*>(o + s * o_ss + b * o_sb + h * o_sh + sub * 8);
    const uint4 qg = *reinterpret_cast<const uint4*>(dout + s * do_ss + b * do_sb + h * do_sh + sub * 8);
    const uint32_t a4[4] = {qa.x, qa.y, qa.z, qa.w}, g4[4] = {qg.x, qg.y, qg.z, qg.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 a = unpack_bf16x2(a4[i]), g = unpack_bf16x2(g4[i]);
      acc = fmaf(a.x, g.x, fmaf(a.y, g.y, acc));
    }
  }
  for (int off = lpr >> 1; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (row < total && sub == 0) delta[(static_cast<size_t>(b) * heads + h) * seqlen + s] = acc;
}

// ------------------------------------------------------------------------------------------
// backward main (head_dim 64 or 128)
// ------------------------------------------------------------------------------------------
// D = 64 : Q/dO double-buffered, dQ has its own TMEM columns and smem staging (scores of block i+1 overlap the dQ drain).
// D = 128: 512 TMEM columns and 227 KB of smem force sharing - Q/dO single-buffered, the dQ accumulator aliases the
//          dP^T columns and its fp32 staging aliases the P^T / dS^T tiles, so one query block is processed strictly
//          after the previous one (attention is a few % of a LLaMA-7B step at seq 512; correctness first).
template <int D>
struct BwdSmem {
  static constexpr int SLABS = D / 64;
  static constexpr int TILE = BLK * D * 2;             // one K/V/Q/dO tile: 16 KB (D = 64) / 32 KB (D = 128)
  static constexpr int QST = D == 64 ? 2 : 1;          // Q / dO stages
  static constexpr bool DQ_ALIAS = D != 64;
  static constexpr int PT_BYTES = BLK * BLK * 2;       // 32 KB
  static constexpr int OFF_K = 0;
  static constexpr int OFF_V = TILE;
  static constexpr int OFF_Q = 2 * TILE;
  static constexpr int OFF_DO = OFF_Q + QST * TILE;
  static constexpr int OFF_PT = OFF_DO + QST * TILE;
  static constexpr int OFF_DST = OFF_PT + PT_BYTES;
  static constexpr int OFF_DQ = DQ_ALIAS ? OFF_PT : OFF_DST + PT_BYTES;          // fp32 staging: 128 x D x 4
  static constexpr int OFF_STAT = DQ_ALIAS ? OFF_DST + PT_BYTES : OFF_DQ + BLK * D * 4;   // lse[2][128], delta[2][128]
  static constexpr int OFF_BAR = OFF_STAT + 4 * BLK * 4;
  static constexpr int TOTAL = OFF_BAR + 256 + 1024;
  static_assert(!DQ_ALIAS || BLK * D * 4 <= 2 * PT_BYTES, "dQ staging must fit in the P^T + dS^T tiles");
  static_assert(TOTAL <= 227 * 1024, "attention backward tile does not fit in shared memory");
};

template <int D, bool DROP>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQ /* fp32, box {32,1,1,128} */,
                const __grid_constant__ CUtensorMap tmDK, const __grid_constant__ CUtensorMap tmDV,
                const AttnParams p) {
  static_assert(D == 64 || D == 128, "head_dim 64 or 128");
  using L = BwdSmem<D>;
  constexpr int SLABS = L::SLABS;
  constexpr int QST = L::QST;
  constexpr bool DQ_ALIAS = L::DQ_ALIAS;
  constexpr uint32_t TMEM_COLS = 512;
  constexpr uint32_t ST_COL = 0, DPT_COL = 128, DV_COL = 256, DK_COL = 256 + D;
  constexpr uint32_t DQ_COL = DQ_ALIAS ? DPT_COL : 256 + 2 * D;
  static_assert(DK_COL + D <= 512 && (DQ_ALIAS || DQ_COL + D <= 512), "TMEM budget");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + L::OFF_K;
  uint8_t* sV = smem + L::OFF_V;
  uint8_t* sQ = smem + L::OFF_Q;
  uint8_t* sDO = smem + L::OFF_DO;
  uint8_t* sPT = smem + L::OFF_PT;
  uint8_t* sDST = smem + L::OFF_DST;
  uint8_t* sDQ = smem + L::OFF_DQ;
  float* sLse = reinterpret_cast<float*>(smem + L::OFF_STAT);   // [2][128]
  float* sDelta = sLse + 2 * BLK;                               // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::OFF_BAR);
  uint64_t* kv_full = bars;          // 1
  uint64_t* q_full = bars + 1;       // 2
  uint64_t* q_empty = bars + 3;      // 2
  uint64_t* st_full = bars + 5;      // 1   S^T and dP^T ready
  uint64_t* pds_full = bars + 6;     // 1   P^T / dS^T written to smem (256 arrivals)
  uint64_t* dq_full = bars + 7;      // 1   dQ_i in TMEM (also: dV/dK MMAs of this iteration done)
  uint64_t* dq_empty = bars + 8;     // 1   dQ TMEM drained + staging free (256 arrivals)
  uint64_t* acc_full = bars + 9;     // 1   final dK/dV
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nqb = (p.seqlen + BLK - 1) / BLK;
  const int kb = blockIdx.x;                 // early kv blocks have the most query blocks: heavy first
  const int head = blockIdx.y, bat = blockIdx.z;
  const int k0 = kb * BLK;
  const int qb_first = p.causal ? kb : 0;
  const int niter = nqb - qb_first;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ); tma_prefetch_desc(&tmDK); tma_prefetch_desc(&tmDV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    mbar_init(st_full, 1); mbar_init(pds_full, 256); mbar_init(dq_full, 1); mbar_init(dq_empty, 256);
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * L::TILE);
#pragma unroll
      for (int s = 0; s < SLABS; ++s) {
        tma_load_4d(sK + s * (BLK * 128), &tmK, kv_full, s * 64, head, bat, k0);
        tma_load_4d(sV + s * (BLK * 128), &tmV, kv_full, s * 64, head, bat, k0);
      }
      for (int it = 0; it < niter; ++it) {
        const int st = it % QST;
        mbar_wait(&q_empty[st], ((it / QST) & 1) ^ 1);
        mbar_expect_tx(&q_full[st], 2 * L::TILE);
#pragma unroll
        for (int s = 0; s < SLABS; ++s) {
          tma_load_4d(sQ + st * L::TILE + s * (BLK * 128), &tmQ, &q_full[st], s * 64, head, bat, (qb_first + it) * BLK);
          tma_load_4d(sDO + st * L::TILE + s * (BLK * 128), &tmDO, &q_full[st], s * 64, head, bat, (qb_first + it) * BLK);
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t id_kk = make_idesc_f16(BLK, BLK, 1, 0, 0);   // S^T, dP^T : K-major x K-major, N = 128
    const uint32_t id_kn = make_idesc_f16(BLK, D, 1, 0, 1);     // dV, dK   : A K-major, B MN-major, N = D
    const uint32_t id_nn = make_idesc_f16(BLK, D, 1, 1, 1);     // dQ       : A MN-major, B MN-major
    mbar_wait(kv_full, 0);
    const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aPT = smem_u32(sPT), aDST = smem_u32(sDST);
    auto issue_scores = [&](int it) {
      const int st = it % QST;
      mbar_wait(&q_full[st], (it / QST) & 1);
      if (DQ_ALIAS && it > 0) mbar_wait(dq_empty, (it - 1) & 1);   // dP^T columns still hold dQ of the previous block
      tc_fence_after();
      if (lane == 0) {
        const uint32_t aQ = smem_u32(sQ + st * L::TILE), aDO = smem_u32(sDO + st * L::TILE);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + ST_COL, make_smem_desc(aK + off, 16, 1024), make_smem_desc(aQ + off, 16, 1024),
                      id_kk, kk > 0 ? 1u : 0u);
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + DPT_COL, make_smem_desc(aV + off, 16, 1024), make_smem_desc(aDO + off, 16, 1024),
                      id_kk, kk > 0 ? 1u : 0u);
        }
        umma_commit(st_full);
      }
      __syncwarp();
    };
    issue_scores(0);
    for (int it = 0; it < niter; ++it) {
      const int st = it % QST;
      mbar_wait(pds_full, it & 1);
      mbar_wait(dq_empty, (it & 1) ^ 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t aQ = smem_u32(sQ + st * L::TILE), aDO = smem_u32(sDO + st * L::TILE);
        // dV += P^T dO ; dK += dS^T Q      (reduction over the 128 query rows)
#pragma unroll
        for (int kk = 0; kk < BLK / 16; ++kk) {
          const uint32_t offA = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + DV_COL, make_smem_desc(aPT + offA, 16, 1024),
                      make_smem_desc(aDO + kk * 2048, BLK * 128, 1024), id_kn, (it > 0 || kk > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int kk = 0; kk < BLK / 16; ++kk) {
          const uint32_t offA = (kk >> 2) * (BLK * 128) + (kk & 3) * 32;
          umma_f16_ss(tmem_base + DK_COL, make_smem_desc(aDST + offA, 16, 1024),
                      make_smem_desc(aQ + kk * 2048, BLK * 128, 1024), id_kn, (it > 0 || kk > 0) ? 1u : 0u);
        }
        // dQ_i = dS K   (M = query, reduction over the 128 key rows; dS^T buffer read MN-major)
#pragma unroll
        for (int kk = 0; kk < BLK / 16; ++kk) {
          umma_f16_ss(tmem_base + DQ_COL, make_smem_desc(aDST + kk * 2048, BLK * 128, 1024),
                      make_smem_desc(aK + kk * 2048, BLK * 128, 1024), id_nn, kk > 0 ? 1u : 0u);
        }
        umma_commit(dq_full);
        umma_commit(&q_empty[st]);
        if (it == niter - 1) umma_commit(acc_full);
      }
      __syncwarp();
      if (it + 1 < niter) issue_scores(it + 1);
    }
  } else if (warp >= SM_WARP0) {
    // two warps per key row: group g owns query columns [64g, 64g+64) of S^T / dP^T, half of the dQ columns
    // and one of the two final accumulators (g = 0: dV, g = 1: dK)
    const int g = (warp - SM_WARP0) >> 2;
    const int ew = (warp - SM_WARP0) & 3;
    const int r = ew * 32 + lane;                  // key row for S^T / dP^T, query row for dQ
    const uint32_t lane_off = static_cast<uint32_t>(ew * 32) << 16;
    const int kpos = k0 + r;
    const float sl2 = p.scale * LOG2E;
    [[maybe_unused]] const uint32_t drop_seed_r = p.drop_replay == nullptr ? p.drop_seed : p.drop_seed ^ static_cast<uint32_t>((*p.drop_replay * 0x9E3779B97F4A7C15ull) >> 32);
    [[maybe_unused]] const uint32_t drop_base = drop_seed_r + static_cast<uint32_t>(kpos >> 1) * DROP_CK +
                                                static_cast<uint32_t>(bat * p.heads + head) * DROP_CB;
    const size_t stat_base = (static_cast<size_t>(bat) * p.heads + head) * p.seqlen;

    auto drain_dq = [&](int it) {
      // dQ_i (TMEM, fp32) -> swizzled fp32 staging (one 128 B-row slab per 32 columns) -> TMA reduce-add
      mbar_wait(dq_full, it & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 64; ++c) {
        const int cc = g * (D / 64) + c;             // 32-column chunk of dQ handled by this thread
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + lane_off + DQ_COL + cc * 32, v);
        tmem_ld_wait();
        uint8_t* rowp = sDQ + cc * (BLK * 128) + r * 128;
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<uint4*>(rowp + ((q ^ (r & 7)) << 4)) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      named_bar_sync(1, 256);
      if (g == 0 && r == 0) {
        if (!(p.dbg_skip & 1)) {
#pragma unroll
          for (int c = 0; c < D / 32; ++c) tma_reduce_add_4d(&tmDQ, sDQ + c * (BLK * 128), c * 32, head, bat, (qb_first + it) * BLK);
        }
        tma_store_commit();
        tma_store_wait_read<0>();
      }
      named_bar_sync(1, 256);
      mbar_arrive(dq_empty);
    };

    for (int it = 0; it < niter; ++it) {
      const int qb = qb_first + it;
      const int qbase = qb * BLK;
      // stage lse / delta of this query block (parity buffer: the previous iteration may still be read)
      if (g == 0) {
        const int qp = qbase + r;
        sLse[(it & 1) * BLK + r] = qp < p.seqlen ? p.lse[stat_base + qp] * LOG2E : INFINITY;
        sDelta[(it & 1) * BLK + r] = qp < p.seqlen ? p.delta[stat_base + qp] : 0.f;
      }
      named_bar_sync(2, 256);
      mbar_wait(st_full, it & 1);
      tc_fence_after();
      const bool diag = p.causal && qb == kb;
      const int qmin = diag ? (kpos - qbase) : 0;     // first visible query column of this block (branch-free mask)
      const float* lse_s = sLse + (it & 1) * BLK;
      const float* del_s = sDelta + (it & 1) * BLK;
#pragma unroll 1
      for (int cc = 0; cc < ((p.dbg_skip & 2) ? 0 : 2); ++cc) {
        const int c = g * 2 + cc;
        uint32_t sv[32], dv[32];
        tmem_ld_32x32b_x32(tmem_base + lane_off + ST_COL + c * 32, sv);
        tmem_ld_32x32b_x32(tmem_base + lane_off + DPT_COL + c * 32, dv);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float pr[2], ds[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int qi = c * 32 + i + e;
            float pe = fast_exp2(fmaf(__uint_as_float(sv[i + e]), sl2, -lse_s[qi]));
            pe = (qi >= qmin) ? pe : 0.f;
            float dp = __uint_as_float(dv[i + e]);
            if constexpr (DROP) {
              // same decision as the forward made for element (query qbase + qi, key kpos)
              const uint32_t hh = drop_mix(drop_base + static_cast<uint32_t>(qbase + qi) * DROP_CQ);
              const float mk = ((kpos & 1) ? (hh >> 16) : (hh & 0xFFFFu)) >= p.drop_thresh ? p.inv_keep : 0.f;
              pr[e] = pe * mk;                       // dropped probabilities feed dV = P~^T dO
              dp *= mk;                              // dP = dP~ * mask / keep
            } else {
              pr[e] = pe;
            }
            ds[e] = pe * (dp - del_s[qi]) * p.scale;
          }
          pk[i >> 1] = pack_bf16x2(pr[0], pr[1]);
          dk[i >> 1] = pack_bf16x2(ds[0], ds[1]);
        }
        const int slab_off = g * (BLK * 128) + r * 128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int chunk = (cc * 4 + q) ^ (r & 7);
          *reinterpret_cast<uint4*>(sPT + slab_off + (chunk << 4)) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
          *reinterpret_cast<uint4*>(sDST + slab_off + (chunk << 4)) = make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(pds_full);
      drain_dq(it);
    }

    // ---- final dK / dV: TMEM -> bf16 -> smem (P^T / dS^T tiles are dead) -> TMA store ----
    mbar_wait(acc_full, 0);
    tc_fence_after();
    {
      uint8_t* stg = g == 0 ? sPT : sDST;
      const uint32_t col = g == 0 ? DV_COL : DK_COL;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + lane_off + col + c * 32, v);
        tmem_ld_wait();
        uint8_t* rowp = stg + (c >> 1) * (BLK * 128) + r * 128;       // 64 bf16 columns per 128 B-row slab
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int chunk = ((c & 1) * 4 + q) ^ (r & 7);
          *reinterpret_cast<uint4*>(rowp + (chunk << 4)) =
              make_uint4(pack_bf16x2(__uint_as_float(v[8 * q]), __uint_as_float(v[8 * q + 1])),
                         pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3])),
                         pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5])),
                         pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7])));
        }
      }
    }
    tc_fence_before();
    fence_proxy_async_smem();
    named_bar_sync(1, 256);
    if (g == 0 && r == 0) {
#pragma unroll
      for (int s = 0; s < SLABS; ++s) {
        tma_store_4d(&tmDV, sPT + s * (BLK * 128), s * 64, head, bat, k0);
        tma_store_4d(&tmDK, sDST + s * (BLK * 128), s * 64, head, bat, k0);
      }
      tma_store_commit();
      tma_store_wait_all<0>();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

int make_map4(CUtensorMap* m, CUtensorMapDataType dt, int esize, const void* base, int D, int heads, int batch,
              int seqlen, long long ss, long long sb, long long sh, int box_inner) {
  uint64_t dims[4] = {static_cast<uint64_t>(D), static_cast<uint64_t>(heads), static_cast<uint64_t>(batch),
                      static_cast<uint64_t>(seqlen)};
  uint64_t strides[4] = {0, static_cast<uint64_t>(sh) * esize, static_cast<uint64_t>(sb) * esize,
                         static_cast<uint64_t>(ss) * esize};
  uint32_t box[4] = {static_cast<uint32_t>(box_inner), 1, 1, BLK};
  return b2host::encode_tmap(m, dt, 4, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
  return 0;
}
int g_attn_dbg = 0;
void set_dropout(AttnParams& p, float p_drop, unsigned long long seed) {
  p.dbg_skip = g_attn_dbg;
  long long t = static_cast<long long>(static_cast<double>(p_drop) * 65536.0 + 0.5);
  if (t > 65535) t = 65535;
  if (t < 0) t = 0;
  p.drop_thresh = static_cast<uint32_t>(t);
  p.drop_seed = static_cast<uint32_t>(seed ^ (seed >> 32));
  p.inv_keep = 65536.f / static_cast<float>(65536 - t);        // 1 / (realised keep probability): unbiased for every p
  p.drop_replay = p_drop > 0.f ? b2host::dropout_replay_ptr() : nullptr;
}
}  // namespace

// strides are in ELEMENTS: (seq, batch, head); head_dim contiguous.  bf16 only.
// profiling hook: results are WRONG while a bit is set (timing experiments only)
extern "C" void b2_attn_set_debug(int bits) { g_attn_dbg = bits; }

extern "C" int b2_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int seqlen, int batch,
                           int heads, int D, long long q_ss, long long q_sb, long long q_sh, long long k_ss,
                           long long k_sb, long long k_sh, long long v_ss, long long v_sb, long long v_sh,
                           long long o_ss, long long o_sb, long long o_sh, float scale, int causal, float p_drop,
                           unsigned long long seed, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p_drop < 0.f || p_drop >= 1.f) { b2host::set_error("attn_fwd: dropout probability must be in [0, 1)"); return -1; }
  if (D != 64 && D != 128) { b2host::set_error("attn_fwd: head_dim must be 64 or 128 (got %d)", D); return -1; }
  CUtensorMap mq, mk, mv, mo;
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (make_map4(&mq, dt, 2, q, D, heads, batch, seqlen, q_ss, q_sb, q_sh, 64)) return -1;
  if (make_map4(&mk, dt, 2, k, D, heads, batch, seqlen, k_ss, k_sb, k_sh, 64)) return -1;
  if (make_map4(&mv, dt, 2, v, D, heads, batch, seqlen, v_ss, v_sb, v_sh, 64)) return -1;
  if (make_map4(&mo, dt, 2, o, D, heads, batch, seqlen, o_ss, o_sb, o_sh, 64)) return -1;
  AttnParams p;
  p.seqlen = seqlen; p.batch = batch; p.heads = heads; p.causal = causal; p.scale = scale; p.lse = lse; p.delta = nullptr;
  set_dropout(p, p_drop, seed);
  const bool drop = p_drop > 0.f;
  dim3 grid((seqlen + BLK - 1) / BLK, heads, batch);
#define B2_ATTN_FWD(D_, DROP_)                                                                                     \
  do {                                                                                                            \
    auto kern = attn_fwd_kernel<D_, DROP_>;                                                                       \
    static bool cfg = false;                                                                                      \
    if (!cfg) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, FwdSmem<D_>::TOTAL); cfg = true; } \
    kern<<<grid, NTHREADS, FwdSmem<D_>::TOTAL, st>>>(mq, mk, mv, mo, p);                                          \
  } while (0)
  if (D == 64) { if (drop) B2_ATTN_FWD(64, true); else B2_ATTN_FWD(64, false); }
  else         { if (drop) B2_ATTN_FWD(128, true); else B2_ATTN_FWD(128, false); }
#undef B2_ATTN_FWD
  return check("attn_fwd");
}

// dq_accum: fp32 [seq, batch, heads, D] contiguous, must be ZERO on entry (reduce-add target).
// delta: fp32 scratch [batch, heads, seqlen] (written here).
extern "C" int b2_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                           const float* lse, float* delta, float* dq_accum, void* dk, void* dv, int seqlen, int batch,
                           int heads, int D, long long q_ss, long long q_sb, long long q_sh, long long k_ss,
                           long long k_sb, long long k_sh, long long v_ss, long long v_sb, long long v_sh,
                           long long o_ss, long long o_sb, long long o_sh, long long do_ss, long long do_sb,
                           long long do_sh, long long dk_ss, long long dk_sb, long long dk_sh, long long dv_ss,
                           long long dv_sb, long long dv_sh, float scale, int causal, float p_drop,
                           unsigned long long seed, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p_drop < 0.f || p_drop >= 1.f) { b2host::set_error("attn_bwd: dropout probability must be in [0, 1)"); return -1; }
  if (D != 64 && D != 128) { b2host::set_error("attn_bwd: head_dim must be 64 or 128 (got %d)", D); return -1; }
  {
    long long rows = static_cast<long long>(seqlen) * batch * heads;
    long long blocks = (rows * (D / 8) + 255) / 256;
    attn_delta_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(o), static_cast<const __nv_bfloat16*>(dout), delta, seqlen, batch, heads, D,
        o_ss, o_sb, o_sh, do_ss, do_sb, do_sh);
  }
  CUtensorMap mq, mk, mv, mdo, mdq, mdk, mdv;
  const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (make_map4(&mq, dt, 2, q, D, heads, batch, seqlen, q_ss, q_sb, q_sh, 64)) return -1;
  if (make_map4(&mk, dt, 2, k, D, heads, batch, seqlen, k_ss, k_sb, k_sh, 64)) return -1;
  if (make_map4(&mv, dt, 2, v, D, heads, batch, seqlen, v_ss, v_sb, v_sh, 64)) return -1;
  if (make_map4(&mdo, dt, 2, dout, D, heads, batch, seqlen, do_ss, do_sb, do_sh, 64)) return -1;
  if (make_map4(&mdk, dt, 2, dk, D, heads, batch, seqlen, dk_ss, dk_sb, dk_sh, 64)) return -1;
  if (make_map4(&mdv, dt, 2, dv, D, heads, batch, seqlen, dv_ss, dv_sb, dv_sh, 64)) return -1;
  if (make_map4(&mdq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dq_accum, D, heads, batch, seqlen,
                static_cast<long long>(batch) * heads * D, static_cast<long long>(heads) * D, D, 32)) return -1;
  AttnParams p;
  p.seqlen = seqlen; p.batch = batch; p.heads = heads; p.causal = causal; p.scale = scale;
  p.lse = const_cast<float*>(lse); p.delta = delta;
  set_dropout(p, p_drop, seed);
  const bool drop = p_drop > 0.f;
  dim3 grid((seqlen + BLK - 1) / BLK, heads, batch);
#define B2_ATTN_BWD(D_, DROP_)                                                                                     \
  do {                                                                                                            \
    auto kern = attn_bwd_kernel<D_, DROP_>;                                                                       \
    static bool cfg = false;                                                                                      \
    if (!cfg) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, BwdSmem<D_>::TOTAL); cfg = true; } \
    kern<<<grid, NTHREADS, BwdSmem<D_>::TOTAL, st>>>(mq, mk, mv, mdo, mdq, mdk, mdv, p);                          \
  } while (0)
  if (D == 64) { if (drop) B2_ATTN_BWD(64, true); else B2_ATTN_BWD(64, false); }
  else         { if (drop) B2_ATTN_BWD(128, true); else B2_ATTN_BWD(128, false); }
#undef B2_ATTN_BWD
  return check("attn_bwd");
}
