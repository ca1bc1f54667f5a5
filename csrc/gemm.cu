// tcgen05 / TMEM / TMA GEMM for sm_100a.
//
//   C[M,N] (+)= op(A)[M,K] * op(B)[K,N]      bf16 or fp16 operands, fp32 accumulate in TMEM
//
// One persistent, warp-specialised kernel (1 CTA per SM):
//   warp 0    : TMA producer  (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx-count)
//   warp 1    : MMA issuer    (one thread issues tcgen05.mma kind::f16 128xBNx16, commits to mbarriers)
//   warp 2    : TMEM allocator
//   warps 4-11: epilogue      two independent 128-thread teams (warps 4-7 / 8-11) take alternating 64-column
//                              chunks of the accumulator: tcgen05.ld 32x32b -> fused epilogue -> swizzled smem ->
//                              TMA store / TMA reduce-add.  Two teams = two warps per SM sub-partition, so one
//                              team's TMEM / global-load / TMA-store latency hides under the other's math; the
//                              whole epilogue overlaps the next tile's mainloop through a double-buffered TMEM
//                              accumulator.
// Operand majors are expressed purely through the UMMA descriptors, so forward (x W^T), dgrad
// (dy W) and wgrad (dy^T x) all read the tensors in place - no transposes are materialised.
//
// Replaces the cuBLAS calls under torch.matmul / F.linear on every path of the reference
//  and apex's fused_weight_gradient_mlp_cuda (wgrad accumulate into
// fp32 main_grad, here as split-K TMA reduce-add).
#include "common.cuh"
#include "b200dist.h"

using namespace b2;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;          // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 384;
constexpr int EPI_THREADS = 128;   // per team
constexpr int EPI_WARP0 = 4;
constexpr int STAGE_CHUNK_BYTES = 128 * 128;  // 128 rows x 128 B staging tile
constexpr int SMEM_MAIN_BYTES = 192 * 1024;

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, split_k, kb_per_split, num_kb;
  int epilogue;
  int act_erf;      // GELU epilogues: 0 = tanh approximation (Megatron / GPT-2), 1 = exact erf form (Swin, torch default)
  int fmt;          // 0 fp16, 1 bf16
  int out_fp32;
  int reduce_add;   // fp32 output accumulated with TMA reduce-add
  const void* bias;     // [N] (operand dtype)
  const void* aux;      // [M, ld_aux] (operand dtype): residual or pre-activation
  long long ld_aux;
  float alpha;
  // ---- fused collectives (FUSE != 0); M is split into `f_world` contiguous shards of f_bpr 128-row blocks
  int f_rank, f_world, f_ncomm, f_bpr, f_shard_rows;
  int m_rot;                       // tile order starts at row-block m_rot (overlap: see ag / rs comments)
  long long f_row_bytes;           // AG: bytes per row of the gathered operand
  void* f_peer[8];                 // AG: every rank's shard (symmetric)   RS: unused
  uint32_t* f_ctrl[8];             // control block of the symmetric buffer on every rank
  uint8_t* f_xg;                   // AG: local gathered operand [M, K]
  uint32_t* f_flags;               // AG: local per-row-block arrival counters (pieces landed, cumulative over calls)
  int f_npb;                       // AG: 32 KB pieces per 128-row block
  int f_last_rows, f_npb_last;     // AG: rows / pieces of a shard's LAST block (= 128 / f_npb unless the shard is not a multiple of 128 rows)
  int f_nmajor, n_rot;             // AGB: tile order is N-major and starts at column tile n_rot (this rank's own rows of B)
  int f_self_clean;                // AGB: the arrival counters are zeroed at the end of the launch (they count from 0 every call)
  uint32_t* f_cnt[8];              // RS: per-row-block arrival counters on every rank (symmetric)
  uint8_t* f_stage_peer[8];        // RS: every rank's staging buffer (symmetric): [2 parity][world sources][slot]
  const void* f_stage;             // RS: this rank's staging [2][world][shard_rows, N]
  void* f_out;                     // RS: reduced output shard [shard_rows, N]
  long long f_ld_out;
  const void* f_bias;              // RS: bias [N] added once, after the reduction
  const void* f_res;               // RS: residual shard [shard_rows, f_ld_res]
  long long f_ld_res;
  const CUtensorMap* f_armaps;     // GEMM->all-reduce: device array [world] of maps over every rank's full [M, N] output (NULL: reduce-scatter)
  unsigned long long* f_dbg;       // optional [gridDim.x][8] globaltimer stamps (profiling builds of the caller)
};

B2_DEVICE void dbg_stamp(const GemmParams& p, int k) {
  if (p.f_dbg != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory");
    p.f_dbg[blockIdx.x * 8 + k] = t;
  }
}

constexpr int FUSE_NONE = 0, FUSE_AG = 1, FUSE_RS = 2, FUSE_AGB = 3;     // AGB: the gathered operand is B (a row-sharded weight)
constexpr int FCTRL_A = 16, FCTRL_B = 32;   // arrival slots inside a control block (same layout as comm.cu)

// ---- 1-D bulk copies (TMA without a tensor map) used by the all-gather CTAs
B2_DEVICE void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
B2_DEVICE void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

// All-gather CTAs of the fused AG->GEMM kernel.  The remote shards are cut into 32 KB pieces that are dealt round-robin
// over the `f_ncomm` all-gather CTAs IN ROW ORDER (piece g goes to CTA g % f_ncomm), so the 128-row blocks of the gathered
// operand complete one after the other, in exactly the order in which the GEMM CTAs consume them - not in f_ncomm-wide
// waves.  One thread per CTA drives a 7-stage ring: 1-D bulk copy peer shard -> shared memory (NVLink is read exactly
// once), bulk copy shared memory -> local gathered operand; a piece that has landed adds 1 to its block's arrival counter
// (release, device scope); a GEMM CTA may load a block once its counter has reached epoch * pieces-per-block.
// This rank's OWN shard reaches the gathered operand (the backward's wgrad reads it there; this kernel's GEMM reads own
// rows in place) through the two otherwise idle warps of every GEMM CTA - a slow background copy spread over the whole
// kernel, so its HBM traffic never competes with the NVLink stream in bursts (copying it from the all-gather CTAs cut the
// pull rate from ~600 to ~470 GB/s; copying it after the pulls left a 20-30 us tail).
B2_DEVICE void ag_comm_cta(const GemmParams& p, uint8_t* smem, int c, uint32_t epoch) {
  constexpr int STG = 7, LA = 6, PUB = 3;
  constexpr uint32_t PIECE = 32 * 1024;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STG * PIECE);
  uint32_t* my_ctrl = p.f_ctrl[p.f_rank];
  if (threadIdx.x == 0) {
    for (int i = 0; i < STG; ++i) mbar_init(&bars[i], 1);
    mbar_fence_init();
    if (c == 0) {                                   // my shard is complete (stream order): tell the peers
      __threadfence_system();
      for (int r = 0; r < p.f_world; ++r)
        if (r != p.f_rank) st_release_sys(p.f_ctrl[r] + FCTRL_A + p.f_rank, epoch);
    }
    dbg_stamp(p, 0);
    const long long blk_bytes = 128LL * p.f_row_bytes;
    const int npb = p.f_npb;
    const long long pps = static_cast<long long>(p.f_bpr - 1) * npb + p.f_npb_last;     // pieces per shard
    const long long total = static_cast<long long>(p.f_world - 1) * pps;               // pieces of all remote shards
    const long long mine = total > c ? (total - c + p.f_ncomm - 1) / p.f_ncomm : 0;    // pieces c, c + ncomm, ...
    long long issued = 0, stored = 0, published = 0;
    int ready_upto = 0;                              // sources (rank+1 ... rank+ready_upto) known to be complete
    uint32_t* pub[PUB];
    auto locate = [&](long long k, int& src, long long& off, uint32_t& sz, uint32_t*& flag) {
      const long long g = c + k * p.f_ncomm;
      const int si = static_cast<int>(g / pps);
      const long long rem = g - si * pps;
      int blk = static_cast<int>(rem / npb);
      if (blk > p.f_bpr - 1) blk = p.f_bpr - 1;
      const int pc = static_cast<int>(rem - static_cast<long long>(blk) * npb);
      src = (p.f_rank + 1 + si) % p.f_world;
      off = blk * blk_bytes + static_cast<long long>(pc) * PIECE;
      const long long this_bytes = (blk == p.f_bpr - 1) ? static_cast<long long>(p.f_last_rows) * p.f_row_bytes : blk_bytes;
      const long long left = this_bytes - static_cast<long long>(pc) * PIECE;
      sz = left < static_cast<long long>(PIECE) ? static_cast<uint32_t>(left) : PIECE;
      flag = p.f_flags + src * p.f_bpr + blk;
    };
    // the piece's bulk store has COMPLETED (cp.async.bulk.wait_group, not .read): its bytes are in L2, which is where the GEMM
    // CTAs' TMA loads read them.  A relaxed add is enough - a release here costs a MEMBAR that also waits for this thread's
    // outstanding NVLink loads (measured: 2.7 us per piece, pull rate 196 GB/s instead of ~460)
    auto publish = [&](uint32_t* flag) { red_relaxed_gpu_add(flag, 1u); };
    while (stored < mine) {
      while (issued < mine && issued - stored < LA) {
        int src; long long off; uint32_t sz; uint32_t* flag;
        locate(issued, src, off, sz, flag);
        const int si = (src - p.f_rank - 1 + p.f_world) % p.f_world;
        while (ready_upto <= si) {                   // the source's shard must be complete before the first pull from it
          const int r = (p.f_rank + 1 + ready_upto) % p.f_world;
          wait_ge_sys(my_ctrl + FCTRL_A + r, epoch, 0x101);
          if (ready_upto == 0) dbg_stamp(p, 1);
          ++ready_upto;
        }
        const uint32_t s = static_cast<uint32_t>(issued % STG);
        if (issued >= STG) tma_store_wait_read<STG - LA - 1>();     // the store that last used this slot has been read out
        mbar_expect_tx(&bars[s], sz);
        bulk_g2s(smem + s * PIECE, static_cast<const uint8_t*>(p.f_peer[src]) + off, sz, &bars[s]);
        ++issued;
      }
      {
        int src; long long off; uint32_t sz; uint32_t* flag;
        locate(stored, src, off, sz, flag);
        const uint32_t s = static_cast<uint32_t>(stored % STG);
        mbar_wait(&bars[s], static_cast<uint32_t>(stored / STG) & 1);
        bulk_s2g(p.f_xg + static_cast<long long>(src) * p.f_shard_rows * p.f_row_bytes + off, smem + s * PIECE, sz);
        tma_store_commit();
        pub[stored % PUB] = flag;
        ++stored;
        if (stored - published >= PUB) {             // the store issued PUB - 1 groups ago has certainly completed
          tma_store_wait_all<PUB - 1>();
          publish(pub[published % PUB]);
          ++published;
        }
      }
    }
    tma_store_wait_all<0>();
    for (; published < stored; ++published) publish(pub[published % PUB]);
    // all of this CTA's pulls are done; the last all-gather CTA tells every peer "I no longer read your shard"
    dbg_stamp(p, 2);
    __threadfence_system();
    const unsigned int t = atomicAdd(my_ctrl + 3, 1u);
    if (t == static_cast<unsigned int>(p.f_ncomm - 1)) {
      my_ctrl[3] = 0u;
      for (int r = 0; r < p.f_world; ++r)
        if (r != p.f_rank) st_release_sys(p.f_ctrl[r] + FCTRL_B + p.f_rank, epoch);
    }
  }
}

// Last CTA of a fused launch: wait until no peer still depends on this rank's symmetric data, then mark
// the call complete (ctrl[1] = epoch) so the next launch on this buffer derives the next epoch.
// announce: first tell every peer that this rank's stores of the call have all completed (GEMM->all-reduce: together with
// the wait this is the closing cross-GPU barrier - every rank's broadcast rows have landed everywhere)
B2_DEVICE void fused_finish(const GemmParams& p, uint32_t epoch, bool wait_peers_done, bool announce = false) {
  __syncthreads();
  __shared__ bool last;
  uint32_t* my_ctrl = p.f_ctrl[p.f_rank];
  if (threadIdx.x == 0) {
    __threadfence();
    last = (atomicAdd(my_ctrl + 2, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    if (announce) {
      __threadfence_system();
      for (int r = 0; r < p.f_world; ++r)
        if (r != p.f_rank) st_release_sys(p.f_ctrl[r] + FCTRL_B + p.f_rank, epoch);
    }
    if (wait_peers_done) {
      for (int r = 0; r < p.f_world; ++r)
        if (r != p.f_rank)
          wait_ge_sys(my_ctrl + FCTRL_B + r, epoch, 0x102);
    }
    my_ctrl[2] = 0u;
    my_ctrl[1] = epoch;
    __threadfence();
    dbg_stamp(p, 5);
  }
  if (last && p.f_self_clean) {      // every CTA of the grid has passed its last wait: the counters restart from zero next call
    for (int i = threadIdx.x; i < p.f_world * p.f_bpr; i += blockDim.x) p.f_flags[i] = 0u;
  }
}

B2_DEVICE float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
B2_DEVICE float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * x * (1.f + k1 * x * x);
  float t = tanh_approx(u);                        // one MUFU op; 2^-11 relative error, below bf16/fp16 rounding
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
}
B2_DEVICE float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float x2 = x * x;
  float u = k0 * x * fmaf(k1, x2, 1.f);
  float t = tanh_approx(u);
  float du = fmaf(3.f * k0 * k1, x2, k0);
  return fmaf(0.5f * x * fmaf(-t, t, 1.f), du, fmaf(0.5f, t, 0.5f));
}
// erf through Abramowitz-Stegun 7.1.26 (absolute error < 6e-7 with fp32 arithmetic, far below 16-bit output rounding): one
// rcp + one ex2 and six FMAs, no branch - erff() costs about three times as many issue slots, which is what bounds the
// epilogue warps of the GELU GEMMs.  `e` returns exp(-z^2), which the derivative needs anyway.
B2_DEVICE float erf_as(float z, float& e) {
  const float az = fabsf(z);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, az, 1.f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(az * az * -1.4426950408889634f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  return copysignf(fmaf(-p * t, e, 1.f), z);
}
B2_DEVICE float gelu_erf(float x) {
  float e;
  const float r = erf_as(x * 0.7071067811865476f, e);
  const float hx = 0.5f * x;
  return fmaf(hx, r, hx);
}
B2_DEVICE float gelu_erf_grad(float x) {       // Phi(x) + x phi(x);  exp(-x^2 / 2) is erf_as's by-product
  float e;
  const float r = erf_as(x * 0.7071067811865476f, e);
  return fmaf(x * 0.3989422804014327f, e, fmaf(0.5f, r, 0.5f));
}
B2_DEVICE uint32_t pack2(float lo, float hi, int fmt) {
  if (fmt == 1) return pack_bf16x2(lo, hi);
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
B2_DEVICE float2 unpack2(uint32_t u, int fmt) {
  if (fmt == 1) return unpack_bf16x2(u);
  __half2 h = *reinterpret_cast<__half2*>(&u);
  return __half22float2(h);
}

B2_DEVICE void decode_tile(const GemmParams& p, int t, int& tm, int& tn, int& ks) {
  // grouped rasterisation: 8 M-tiles share each sweep over N so A and B tiles are re-used from L2
  int mn_tiles = p.tiles_m * p.tiles_n;
  ks = t / mn_tiles;
  int r = t - ks * mn_tiles;
  const int G = 8;
  int group_sz = G * p.tiles_n;
  int g = r / group_sz;
  int first_m = g * G;
  int gm = min(p.tiles_m - first_m, G);
  int in_g = r - g * group_sz;
  tm = first_m + in_g % gm;
  tn = in_g / gm;
  if (p.f_nmajor) {                   // all-gathered B operand: N-major (a column tile's row tiles are consecutive), own columns first
    tn = r / p.tiles_m + p.n_rot;
    if (tn >= p.tiles_n) tn -= p.tiles_n;
    tm = r - (r / p.tiles_m) * p.tiles_m;
  }
  if (p.m_rot) { tm += p.m_rot; if (tm >= p.tiles_m) tm -= p.tiles_m; }
}

// CTA2: two CTAs of a cluster (one TPC) work on one 256 x BN tile with tcgen05.mma.cta_group::2: each CTA stages its own
// 128 rows of A and HALF of the B tile (the pair shares B, so operand traffic per MMA halves and the smem ring gets
// six stages instead of four), the leader CTA issues the MMAs for both, every CTA runs the epilogue of its own rows.
// ERF: the GELU epilogues use the exact erf form (compile-time: a run-time switch inside the unrolled epilogue loops
// doubled the cost of the tanh path that every GPT layer takes)
template <int BN, bool A_MN, bool B_MN, int FUSE = FUSE_NONE, bool CTA2 = false, bool ERF = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2,
                    const GemmParams p, const __grid_constant__ CUtensorMap tmA2 /* AG: this rank's own shard */) {
  constexpr int BN_LOAD = CTA2 ? BN / 2 : BN;     // B rows staged by this CTA
  constexpr int BM_TILE = CTA2 ? 2 * BM : BM;     // rows per scheduled tile
  constexpr int A_BYTES = BM * BK * 2;
  constexpr int B_BYTES = BN_LOAD * BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // two 16 KB tiles of the operand ring go to the epilogue's aux prefetch (residual / pre-activation tiles arrive by TMA
  // one chunk ahead instead of row-strided global loads); GEMM->RS adds its residual after the reduction and keeps the full ring
  constexpr bool AUX_TMA = (FUSE != FUSE_RS);
  // GEMM->RS: four 16 KB push buffers (two per epilogue team) between the epilogue and the push warp come out of the ring
  constexpr int RS_PUSH_BUFS = 4;
  // pair kernels (except GEMM->RS) double-buffer the epilogue staging: a second 16 KB tile per team, also out of the ring,
  // so a chunk's shared-memory write never waits for the previous chunk's TMA store to be read out (for K = 768 the
  // epilogue, not the mainloop, paces the kernel: ncu shows the tensor pipe active 64 % of the time)
  constexpr bool EPI_DB = CTA2 && (FUSE != FUSE_RS);
  constexpr int RING_BYTES = AUX_TMA ? SMEM_MAIN_BYTES - (EPI_DB ? 4 : 2) * STAGE_CHUNK_BYTES : SMEM_MAIN_BYTES - RS_PUSH_BUFS * STAGE_CHUNK_BYTES;
  constexpr int STAGES = RING_BYTES / STAGE_BYTES;
  constexpr uint32_t TMEM_COLS = 512;
  // two epilogue teams everywhere; in GEMM->RS each team leader signals its own half of every tile, so a row block
  // is complete once (world - 1) * tiles_n * TEAMS arrivals have been counted
  constexpr int TEAMS = 2;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  [[maybe_unused]] uint8_t* aux_tiles = smem + RING_BYTES;   // AUX_TMA: 2 x 16 KB (one per epilogue team)
  uint8_t* epi_buf = smem + SMEM_MAIN_BYTES;       // 2 x 16 KB
  [[maybe_unused]] uint8_t* epi_buf2 = smem + RING_BYTES + 2 * STAGE_CHUNK_BYTES;     // EPI_DB: the teams' second staging tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_buf + 2 * STAGE_CHUNK_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  [[maybe_unused]] uint64_t* aux_bar = bars + 2 * STAGES + 5;   // [2]
  // GEMM->RS push queue: epilogue team T fills push buffer 2T + (its push count & 1), the push warp sends it over NVLink
  [[maybe_unused]] uint8_t* push_bufs = smem + RING_BYTES;
  [[maybe_unused]] uint64_t* pfull = bars + 2 * STAGES + 7;      // [RS_PUSH_BUFS]
  [[maybe_unused]] uint64_t* pempty = pfull + RS_PUSH_BUFS;      // [RS_PUSH_BUFS]
  struct PushMeta { uint8_t* dst; uint32_t* cnt; };
  [[maybe_unused]] PushMeta* pmeta = reinterpret_cast<PushMeta*>(pempty + RS_PUSH_BUFS);   // [RS_PUSH_BUFS]
  // once a CTA has reached its OWN rows nothing is pushed any more: the push buffers become landing tiles for the peers'
  // staged partials (buffer 2T) and the residual tile (buffer 2T + 1) of team T, filled by TMA and tracked by `sbar`
  [[maybe_unused]] uint64_t* sbar = reinterpret_cast<uint64_t*>(pmeta + RS_PUSH_BUFS);     // [RS_PUSH_BUFS]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = CTA2 ? cluster_ctarank() : 0u;     // 0 = leader of the pair

  // fused collectives: the call epoch comes from the buffer's control block (stable at kernel entry)
  uint32_t f_epoch = 0;
  int nctas = CTA2 ? gridDim.x / 2 : gridDim.x, cta0 = CTA2 ? blockIdx.x / 2 : blockIdx.x;   // tile-loop stride / start
  if constexpr (FUSE != FUSE_NONE) {
    f_epoch = p.f_ctrl[p.f_rank][1] + 1;
    if constexpr (FUSE == FUSE_AG || FUSE == FUSE_AGB) {
      // pair kernels: f_ncomm is even, so a cluster is either two all-gather CTAs or one GEMM pair
      nctas = (static_cast<int>(gridDim.x) - p.f_ncomm) / (CTA2 ? 2 : 1);
      cta0 = (static_cast<int>(blockIdx.x) - p.f_ncomm) / (CTA2 ? 2 : 1);
      if (static_cast<int>(blockIdx.x) < p.f_ncomm) {     // all-gather CTAs (lowest block ids: scheduled first);
        ag_comm_cta(p, smem, blockIdx.x, f_epoch);         // they never touch TMEM / the MMA pipeline
        fused_finish(p, f_epoch, true);
        return;
      }
    }
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (p.epilogue == B2_EPI_BIAS_GELU_DUAL) tma_prefetch_desc(&tmC2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], CTA2 ? 2 : 1);      // pair: the leader's expect_tx arrive + the peer's remote arrive
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], EPI_THREADS * TEAMS * (CTA2 ? 2 : 1));   // pair: both CTAs' epilogue threads (leader's copy)
    }
    if constexpr (AUX_TMA) { mbar_init(&aux_bar[0], 1); mbar_init(&aux_bar[1], 1); }
    if constexpr (FUSE == FUSE_RS) { for (int i = 0; i < RS_PUSH_BUFS; ++i) { mbar_init(&pfull[i], 1); mbar_init(&pempty[i], 1); mbar_init(&sbar[i], 1); } }
    mbar_fence_init();
  }
  if (warp == 2) { if constexpr (CTA2) tmem_alloc_2sm<TMEM_COLS>(tmem_slot); else tmem_alloc<TMEM_COLS>(tmem_slot); }
  tc_fence_before();
  if constexpr (CTA2) cluster_sync(); else __syncthreads();     // pair: barriers of both CTAs are live before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if constexpr (FUSE != FUSE_NONE) { if (threadIdx.x == 0) dbg_stamp(p, 0); }

  const int total_tiles = p.tiles_m * p.tiles_n * p.split_k;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cta0; t < total_tiles; t += nctas) {
        int tm, tn, ks;
        decode_tile(p, t, tm, tn, ks);
        const int m0 = tm * BM_TILE + static_cast<int>(cta_rank) * BM, n0 = tn * BN + static_cast<int>(cta_rank) * (CTA2 ? BN_LOAD : 0);
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
        [[maybe_unused]] bool a_local = false;
        if constexpr (FUSE == FUSE_AG) {
          const int rb = m0 / BM;                        // the 128-row block this CTA stages (pair: each CTA its own half)
          a_local = (rb / p.f_bpr == p.f_rank);          // own rows: read in place from the shard, no waiting
          if (!a_local) {                                // remote rows must have landed in the gathered operand
            wait_ge_gpu(p.f_flags + rb, f_epoch * static_cast<uint32_t>(p.f_npb), 0x103);
            fence_proxy_async_all();
          }
        }
        if constexpr (FUSE == FUSE_AGB) {
          // rows [nb0, nb1) of the gathered B operand that this CTA stages: every 128-row block of every shard they touch
          // must have landed (own shard: in place since before the launch)
          const int nb0 = n0, nb1 = min(p.N, n0 + BN_LOAD);
          for (int rr = nb0; rr < nb1;) {
            const int sh = rr / p.f_shard_rows, lr = rr - sh * p.f_shard_rows, blk = lr >> 7;
            if (sh != p.f_rank)
              wait_ge_gpu(p.f_flags + sh * p.f_bpr + blk, static_cast<uint32_t>(blk == p.f_bpr - 1 ? p.f_npb_last : p.f_npb), 0x105);
            rr = sh * p.f_shard_rows + min(p.f_shard_rows, (blk + 1) << 7);
          }
          fence_proxy_async_all();
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sA = stage_base + stage * STAGE_BYTES;
          uint8_t* sB = sA + A_BYTES;
          if constexpr (CTA2) {
            // both CTAs' bytes complete on the LEADER's barrier (the leader issues the MMAs for the pair)
            const uint32_t lbar = mapa_u32(smem_u32(&full_bar[stage]), 0);
            if (cta_rank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
            else               mbar_arrive_cluster(lbar);
            if constexpr (!A_MN) {
              if constexpr (FUSE == FUSE_AG) {
                if (a_local) tma_load_2d_2sm(sA, &tmA2, lbar, kb * BK, m0 - p.f_rank * p.f_shard_rows);
                else         tma_load_2d_2sm(sA, &tmA, lbar, kb * BK, m0);
              } else {
                tma_load_2d_2sm(sA, &tmA, lbar, kb * BK, m0);
              }
            } else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) tma_load_2d_2sm(sA + j * (BK * 128), &tmA, lbar, m0 + j * 64, kb * BK);
            }
            if constexpr (!B_MN) {
              tma_load_2d_2sm(sB, &tmB, lbar, kb * BK, n0);
            } else {
#pragma unroll
              for (int j = 0; j < BN_LOAD / 64; ++j) tma_load_2d_2sm(sB + j * (BK * 128), &tmB, lbar, n0 + j * 64, kb * BK);
            }
          } else {
          mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          if constexpr (!A_MN) {
            if constexpr (FUSE == FUSE_AG) {
              if (a_local) tma_load_2d(sA, &tmA2, &full_bar[stage], kb * BK, m0 - p.f_rank * p.f_shard_rows);
              else         tma_load_2d(sA, &tmA, &full_bar[stage], kb * BK, m0);
            } else {
              tma_load_2d(sA, &tmA, &full_bar[stage], kb * BK, m0);
            }
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sA + j * (BK * 128), &tmA, &full_bar[stage], m0 + j * 64, kb * BK);
          }
          if constexpr (!B_MN) {
            tma_load_2d(sB, &tmB, &full_bar[stage], kb * BK, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sB + j * (BK * 128), &tmB, &full_bar[stage], n0 + j * 64, kb * BK);
          }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && cta_rank == 0) {
    // ===================== MMA issuer (pair: the leader CTA only) =====================
    const uint32_t idesc = make_idesc_f16(BM_TILE, BN, p.fmt, A_MN ? 1 : 0, B_MN ? 1 : 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int t = cta0; t < total_tiles; t += nctas, ++it) {
      int tm, tn, ks;
      decode_tile(p, t, tm, tn, ks);
      const int kb0 = ks * p.kb_per_split;
      const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      const int buf = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_empty[buf], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sA = smem_u32(stage_base + stage * STAGE_BYTES);
          const uint32_t sB = sA + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            uint64_t adesc, bdesc;
            if constexpr (!A_MN) adesc = make_smem_desc(sA + k * 32, 16, 1024);
            else                 adesc = make_smem_desc(sA + k * 2048, BK * 128, 1024);
            if constexpr (!B_MN) bdesc = make_smem_desc(sB + k * 32, 16, 1024);
            else                 bdesc = make_smem_desc(sB + k * 2048, BK * 128, 1024);
            if constexpr (CTA2) umma_f16_ss_2sm(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            else                umma_f16_ss(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          if constexpr (CTA2) {           // frees the stage in BOTH CTAs / hands the accumulator to BOTH epilogues
            umma_commit_2sm(&empty_bar[stage], 3);
            if (kb == kb1 - 1) umma_commit_2sm(&tmem_full[buf], 3);
          } else {
            umma_commit(&empty_bar[stage]);
            if (kb == kb1 - 1) umma_commit(&tmem_full[buf]);
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (FUSE == FUSE_RS && warp == 3) {
    // ===================== GEMM->RS push warp =====================
    // One thread sends the partial-tile chunks that the epilogue teams park in the push buffers to the rank that owns the
    // rows (ONE 16 KB bulk copy each, the swizzled tile verbatim), keeps at most four in flight, and - as each copy
    // COMPLETES - adds 1 to the owner's arrival counter of that row block and hands the buffer back.  The epilogue teams
    // therefore never wait for NVLink, only for a free buffer, and nobody on this GPU ever waits for a peer while a signal
    // is still owed (the pushes depend on nothing remote).
    if (lane == 0) {
      constexpr int NCH = BN / 64;
      uint32_t q[TEAMS] = {0u, 0u};
      uint32_t issued = 0, retired = 0;
      int fifo[4];
      // the chunk's bulk copy has COMPLETED (cp.async.bulk.wait_group without .read: the write is performed at the owner),
      // so a relaxed add that is issued afterwards cannot overtake it; a system-scope fence here would also wait for the
      // copies still in flight and cut the push rate in half (measured)
      auto retire = [&]() {
        const int slot = fifo[retired & 3];
        red_relaxed_sys_add(pmeta[slot].cnt, 1u);
        mbar_arrive(&pempty[slot]);
        ++retired;
      };
      for (int t = cta0; t < total_tiles; t += nctas) {
        int tm, tn, ks;
        decode_tile(p, t, tm, tn, ks);
        const int rb = (tm * BM_TILE + static_cast<int>(cta_rank) * BM) / BM;
        if (rb / p.f_bpr == p.f_rank) continue;            // own rows: reduced and stored by the epilogue itself
        for (int c = 0; c < NCH; ++c) {
          if (tn * BN + c * 64 >= p.N) break;
          const int team = c % TEAMS;
          const int slot = team * 2 + static_cast<int>(q[team] & 1u);
          mbar_wait(&pfull[slot], (q[team] >> 1) & 1u);
          ++q[team];
          bulk_s2g(pmeta[slot].dst, push_bufs + slot * STAGE_CHUNK_BYTES, STAGE_CHUNK_BYTES);
          tma_store_commit();
          fifo[issued & 3] = slot;
          ++issued;
          if (issued - retired > 3) { tma_store_wait_all<3>(); retire(); }
        }
      }
      tma_store_wait_all<0>();
      while (retired < issued) retire();
      dbg_stamp(p, 2);
    }
  } else if (FUSE == FUSE_AG && p.f_xg != nullptr && (warp == 2 || warp == 3)) {
    // ===================== own shard -> gathered operand (background copy, see ag_comm_cta) =====================
    const long long nthr = static_cast<long long>(gridDim.x - p.f_ncomm) * 64;
    const long long tid = static_cast<long long>(blockIdx.x - p.f_ncomm) * 64 + (warp - 2) * 32 + lane;
    const uint4* src = static_cast<const uint4*>(p.f_peer[p.f_rank]);
    uint4* dst = reinterpret_cast<uint4*>(p.f_xg + static_cast<long long>(p.f_rank) * p.f_shard_rows * p.f_row_bytes);
    const long long total = static_cast<long long>(p.f_shard_rows) * p.f_row_bytes / 16;
    long long v = tid;
    for (; v + 7 * nthr < total; v += 8 * nthr) {
      uint4 q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) q[j] = __ldg(src + v + j * nthr);
#pragma unroll
      for (int j = 0; j < 8; ++j) dst[v + j * nthr] = q[j];
    }
    for (; v < total; v += nthr) dst[v] = __ldg(src + v);
  } else if (warp >= EPI_WARP0 && warp < EPI_WARP0 + 4 * TEAMS) {
    // ===================== epilogue =====================
    const int team = (warp - EPI_WARP0) >> 2;      // chunks team, team + TEAMS, ... of every tile
    const int ew = (warp - EPI_WARP0) & 3;         // TMEM lane quarter (a warp may only touch lanes 32*(warp%4)..+31)
    const int et = ew * 32 + lane;                 // 0..127 == accumulator row
    const bool leader = (et == 0);
    const int bar_id = 1 + team;
    uint8_t* team_buf = epi_buf + team * STAGE_CHUNK_BYTES;   // TEAMS == 2: one staging tile per team (two with EPI_DB)
    [[maybe_unused]] uint8_t* team_buf2 = epi_buf2 + team * STAGE_CHUNK_BYTES;
    // staging tile of store group number n of this team, and the wait that makes it reusable
    auto stage_of = [&](int n) -> uint8_t* { return (EPI_DB && (n & 1)) ? team_buf2 : team_buf; };
    auto stage_wait = [&]() { if constexpr (EPI_DB) tma_store_wait_read<1>(); else tma_store_wait_read<0>(); };
    const int fmt = p.fmt;
    int it = 0;
    int sctr = 0;  // staging-buffer ring counter (one increment per issued TMA store group)
    [[maybe_unused]] uint32_t rs_pq = 0;       // GEMM->RS: chunks this team has handed to the push warp
    [[maybe_unused]] uint32_t rs_phA = 0, rs_phB = 0;     // phases of the two landing-tile barriers
    [[maybe_unused]] bool rs_landing = false;             // the team's push buffers have been taken over as landing tiles
    [[maybe_unused]] uint8_t* landA = push_bufs + (team * 2) * STAGE_CHUNK_BYTES;
    [[maybe_unused]] uint8_t* landB = push_bufs + (team * 2 + 1) * STAGE_CHUNK_BYTES;
    // leader: start the TMA loads of one own-row chunk: first peer's staged partial -> landA, residual tile -> landB
    auto rs_issue = [&](int lblk, int col0) {
      if constexpr (FUSE == FUSE_RS) {
        const int ncc = (p.N + 63) >> 6;
        const long long slot_bytes = static_cast<long long>(p.f_shard_rows) * ncc * 128;
        const int s0 = p.f_rank == 0 ? 1 : 0;
        const uint8_t* src = static_cast<const uint8_t*>(p.f_stage) + (static_cast<long long>(f_epoch & 1) * p.f_world + s0) * slot_bytes +
                             (static_cast<long long>(lblk) * ncc + (col0 >> 6)) * STAGE_CHUNK_BYTES;
        mbar_expect_tx(&sbar[team * 2], STAGE_CHUNK_BYTES);
        bulk_g2s(landA, src, STAGE_CHUNK_BYTES, &sbar[team * 2]);
        if (p.f_res != nullptr) {
          mbar_expect_tx(&sbar[team * 2 + 1], STAGE_CHUNK_BYTES);
          tma_load_2d(landB, &tmC2, &sbar[team * 2 + 1], col0, lblk * BM);
        }
      }
    };
    // ---- aux tile prefetch (AUX_TMA): the team leader keeps ONE 16 KB TMA load in flight, for the next chunk this team
    // will process; (aux_t, aux_c) is that chunk.  The consumer waits on the team's mbarrier, copies its 128-byte row
    // out of the swizzled tile, and after the team barrier the leader issues the following chunk's load.
    [[maybe_unused]] const bool aux_on = AUX_TMA && !p.out_fp32 && (p.epilogue == B2_EPI_BIAS_RESIDUAL || p.epilogue == B2_EPI_DGELU);
    [[maybe_unused]] uint8_t* aux_tile = aux_tiles + team * STAGE_CHUNK_BYTES;
    [[maybe_unused]] uint32_t aux_phase = 0;
    [[maybe_unused]] int aux_t = cta0, aux_c = team - TEAMS;
    [[maybe_unused]] bool aux_more = true;
    auto aux_issue_next = [&]() {           // leader only
      if constexpr (AUX_TMA) {
        while (aux_more) {
          aux_c += TEAMS;
          if (aux_c >= BN / 64) { aux_t += nctas; aux_c = team; }
          if (aux_t >= total_tiles) { aux_more = false; break; }
          int tm2, tn2, ks2;
          decode_tile(p, aux_t, tm2, tn2, ks2);
          const int c0 = tn2 * BN + aux_c * 64;
          if (c0 >= p.N) { aux_c = BN / 64; continue; }         // rest of this tile is past N: next tile
          mbar_expect_tx(&aux_bar[team], STAGE_CHUNK_BYTES);
          tma_load_2d(aux_tile, &tmC2, &aux_bar[team], c0, tm2 * BM_TILE + static_cast<int>(cta_rank) * BM);
          break;
        }
      }
    };
    if (aux_on && leader) aux_issue_next();
    // accumulator buffer drained: tell the MMA issuer (pair: the leader CTA's barrier, from both CTAs)
    auto acc_release = [&](uint64_t* bar) {
      if constexpr (CTA2) mbar_arrive_cluster(mapa_u32(smem_u32(bar), 0));
      else                mbar_arrive(bar);
    };
    for (int t = cta0; t < total_tiles; t += nctas, ++it) {
      int tm, tn, ks;
      decode_tile(p, t, tm, tn, ks);
      const int m0 = tm * BM_TILE + static_cast<int>(cta_rank) * BM, n0 = tn * BN;
      const int buf = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[buf], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * BN + (static_cast<uint32_t>(ew * 32) << 16);
      const int row = m0 + et;
      const bool row_ok = row < p.M;

      if (p.out_fp32) {
        // 32-column chunks of fp32 (128 B rows)
        constexpr int NCH = BN / 32;
#pragma unroll 1
        for (int c = team; c < NCH; c += TEAMS) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c * 32, v);
          tmem_ld_wait();
          if (c + TEAMS >= NCH) { tc_fence_before(); acc_release(&tmem_empty[buf]); }
          if (n0 + c * 32 >= p.N) continue;   // uniform across the team
          // fp32 output never prefetches aux tiles, so in the pair kernels the team's aux tile serves as a third staging
          // tile: two reduce-adds stay in flight behind the one being staged
          constexpr bool TRI = EPI_DB && AUX_TMA;
          uint8_t* sb;
          if constexpr (TRI) { const int s3 = sctr % 3; sb = s3 == 0 ? team_buf : (s3 == 1 ? team_buf2 : aux_tile); }
          else sb = stage_of(sctr);
          ++sctr;
          if (leader) { if constexpr (TRI) tma_store_wait_read<2>(); else stage_wait(); }
          named_bar_sync(bar_id, EPI_THREADS);
          if (p.alpha != 1.f) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * p.alpha);
          }
          uint8_t* rowp = sb + et * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 q = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            *reinterpret_cast<uint4*>(rowp + ((j ^ (et & 7)) << 4)) = q;
          }
          fence_proxy_async_smem();
          named_bar_sync(bar_id, EPI_THREADS);
          if (leader) {
            if (p.reduce_add) tma_reduce_add_2d(&tmC, sb, n0 + c * 32, m0);
            else              tma_store_2d(&tmC, sb, n0 + c * 32, m0);
            tma_store_commit();
          }
        }
      } else {
        constexpr int NCH = BN / 64;
        const bool dual = (p.epilogue == B2_EPI_BIAS_GELU_DUAL);
        [[maybe_unused]] bool rs_own = false;
        [[maybe_unused]] const int rb = m0 / BM;                 // this CTA's 128-row block (pair: its half of the 256-row tile)
        [[maybe_unused]] const int rs_owner = (FUSE == FUSE_RS) ? rb / p.f_bpr : 0;
        [[maybe_unused]] const int rs_lblk = (FUSE == FUSE_RS) ? rb - rs_owner * p.f_bpr : 0;
        if constexpr (FUSE == FUSE_RS) {
          rs_own = (rs_owner == p.f_rank);
          if (rs_own) {          // every peer's partial tiles of this row block must have arrived
            if (leader) {
              // one arrival per 64-column chunk and peer (sent by the peers' push warps, which wait for nothing remote)
              const uint32_t expected = f_epoch * static_cast<uint32_t>((p.f_world - 1) * ((p.N + 63) >> 6));
              const uint32_t* cnt = p.f_cnt[p.f_rank] + rs_lblk;
              if (team == 0) dbg_stamp(p, 3);
              wait_ge_sys(cnt, expected, 0x104);
              if (team == 0) dbg_stamp(p, 6);
              if (!rs_landing) {            // both push buffers of this team have been retired by the push warp
                mbar_wait(&pempty[team * 2 + static_cast<int>(rs_pq & 1u)], ((rs_pq >> 1) & 1u) ^ 1u);
                mbar_wait(&pempty[team * 2 + static_cast<int>((rs_pq + 1) & 1u)], (((rs_pq + 1) >> 1) & 1u) ^ 1u);
              }
              if (n0 + team * 64 < p.N) rs_issue(rs_lblk, n0 + team * 64);
            }
            rs_landing = true;
            named_bar_sync(bar_id, EPI_THREADS);
          }
        }
        // (GEMM->RS applies bias / residual after the reduction, on the owner: no per-tile aux there)
        const bool use_aux = (FUSE != FUSE_RS) && (p.epilogue == B2_EPI_BIAS_RESIDUAL || p.epilogue == B2_EPI_DGELU);
#pragma unroll 1
        for (int c = team; c < NCH; c += TEAMS) {
          const int col0 = n0 + c * 64;
          // this thread's 128 B of the aux tile (residual / pre-activation): requested before the TMEM load so
          // the global-memory latency runs under it (and under the other team's math)
          uint4 aq[8];
          if (use_aux && !aux_on) {
            const uint4* ap = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.aux) +
                                                             static_cast<long long>(row) * p.ld_aux + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              aq[j] = (row_ok && col0 + j * 8 < p.N) ? __ldg(ap + j) : make_uint4(0u, 0u, 0u, 0u);
          }
          uint32_t v[64];
          tmem_ld_32x32b_x32(taddr + c * 64, v);
          tmem_ld_32x32b_x32(taddr + c * 64 + 32, v + 32);
          tmem_ld_wait();
          if (c + TEAMS >= NCH) { tc_fence_before(); acc_release(&tmem_empty[buf]); }
          if (col0 >= p.N) continue;          // uniform across the team
          if constexpr (AUX_TMA) {
            if (aux_on) {                     // the prefetched tile of exactly this chunk (same iteration order as aux_issue_next)
              mbar_wait(&aux_bar[team], aux_phase);
              aux_phase ^= 1;
              const uint8_t* arow = aux_tile + et * 128;
#pragma unroll
              for (int j = 0; j < 8; ++j) aq[j] = *reinterpret_cast<const uint4*>(arow + ((j ^ (et & 7)) << 4));
            }
          }
          float* f = reinterpret_cast<float*>(v);
          if (p.alpha != 1.f) {
#pragma unroll
            for (int i = 0; i < 64; ++i) f[i] *= p.alpha;
          }
          // ---- bias ----
          if (p.bias != nullptr) {
            const uint4* bp = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.bias) + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (col0 + j * 8 < p.N) {
                uint4 q = __ldg(bp + j);
                float2 a = unpack2(q.x, fmt), b = unpack2(q.y, fmt), cc = unpack2(q.z, fmt), d = unpack2(q.w, fmt);
                f[8 * j + 0] += a.x; f[8 * j + 1] += a.y; f[8 * j + 2] += b.x; f[8 * j + 3] += b.y;
                f[8 * j + 4] += cc.x; f[8 * j + 5] += cc.y; f[8 * j + 6] += d.x; f[8 * j + 7] += d.y;
              }
            }
          }
          if constexpr (FUSE == FUSE_RS) {
            if (rs_own) {
              // rows owned by this rank are computed LAST: the peers' partial tiles already sit in the local
              // staging slots, so the reduction (+ bias + residual) happens right here and the result goes
              // straight to the output shard - no separate reduce pass, no staging round trip of the own partial
              // staging slots are chunk-major: [parity][source rank][row block][64-column chunk][128 rows][64], and every
              // 16 KB chunk is the sender's 128B-SWIZZLED shared-memory tile verbatim (it left there as one 1-D bulk copy):
              // 16-byte unit j of row r sits at unit j ^ (r & 7).  The chunk arrives in the landing tile by ONE bulk copy
              // (issued a chunk ahead), so every thread reads its 128-byte row from shared memory without bank conflicts.
              auto add_row = [&](const uint8_t* tile) {
                const uint8_t* rowp = tile + et * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  if (col0 + j * 8 < p.N) {
                    const uint4 q = *reinterpret_cast<const uint4*>(rowp + ((j ^ (et & 7)) << 4));
                    float2 a = unpack2(q.x, fmt), b = unpack2(q.y, fmt), cc = unpack2(q.z, fmt), d = unpack2(q.w, fmt);
                    f[8 * j + 0] += a.x; f[8 * j + 1] += a.y; f[8 * j + 2] += b.x; f[8 * j + 3] += b.y;
                    f[8 * j + 4] += cc.x; f[8 * j + 5] += cc.y; f[8 * j + 6] += d.x; f[8 * j + 7] += d.y;
                  }
                }
              };
              mbar_wait(&sbar[team * 2], rs_phA); rs_phA ^= 1u;
              add_row(landA);
              {                                   // further peers (world > 2): one after the other through the same tile
                const int ncc = (p.N + 63) >> 6;
                const long long slot_bytes = static_cast<long long>(p.f_shard_rows) * ncc * 128;
                const int s0 = p.f_rank == 0 ? 1 : 0;
                for (int sidx = s0 + 1; sidx < p.f_world; ++sidx) {
                  if (sidx == p.f_rank) continue;
                  named_bar_sync(bar_id, EPI_THREADS);        // everybody has read the previous partial
                  if (leader) {
                    const uint8_t* src = static_cast<const uint8_t*>(p.f_stage) + (static_cast<long long>(f_epoch & 1) * p.f_world + sidx) * slot_bytes +
                                         (static_cast<long long>(rs_lblk) * ncc + (col0 >> 6)) * STAGE_CHUNK_BYTES;
                    mbar_expect_tx(&sbar[team * 2], STAGE_CHUNK_BYTES);
                    bulk_g2s(landA, src, STAGE_CHUNK_BYTES, &sbar[team * 2]);
                  }
                  mbar_wait(&sbar[team * 2], rs_phA); rs_phA ^= 1u;
                  add_row(landA);
                }
              }
              if (p.f_res != nullptr) {
                mbar_wait(&sbar[team * 2 + 1], rs_phB); rs_phB ^= 1u;
                add_row(landB);
              }
              if (p.f_bias != nullptr) {           // same 128 bytes for every row: L1 hits
                const uint4* bp2 = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.f_bias) + col0);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  if (col0 + j * 8 < p.N) {
                    const uint4 q = __ldg(bp2 + j);
                    float2 a = unpack2(q.x, fmt), b = unpack2(q.y, fmt), cc = unpack2(q.z, fmt), d = unpack2(q.w, fmt);
                    f[8 * j + 0] += a.x; f[8 * j + 1] += a.y; f[8 * j + 2] += b.x; f[8 * j + 3] += b.y;
                    f[8 * j + 4] += cc.x; f[8 * j + 5] += cc.y; f[8 * j + 6] += d.x; f[8 * j + 7] += d.y;
                  }
                }
              }
            }
          }
          // staging: TEAMS == 2 -> the team's own tile (reused once its previous TMA store has read it);
          //          TEAMS == 1 -> ring of two tiles
          uint8_t* sb0 = stage_of(dual ? sctr + 1 : sctr);      // dual: z goes through tile n, gelu(z) through tile n + 1
          sctr += dual ? 2 : 1;
          if (dual) {
            // pre-activation z -> C2 first, gelu(z) -> C second (through the same staging tile when TEAMS == 2)
            uint8_t* sb1 = stage_of(sctr - 2);
            if (leader) stage_wait();
            named_bar_sync(bar_id, EPI_THREADS);
            uint8_t* rowp = sb1 + et * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              uint4 q = make_uint4(pack2(f[8 * j], f[8 * j + 1], fmt), pack2(f[8 * j + 2], f[8 * j + 3], fmt),
                                   pack2(f[8 * j + 4], f[8 * j + 5], fmt), pack2(f[8 * j + 6], f[8 * j + 7], fmt));
              *reinterpret_cast<uint4*>(rowp + ((j ^ (et & 7)) << 4)) = q;
            }
            if (TEAMS == 2) {
              fence_proxy_async_smem();
              named_bar_sync(bar_id, EPI_THREADS);
              if (leader) { tma_store_2d(&tmC2, sb1, col0, m0); tma_store_commit(); }
            }
            if constexpr (ERF) {
#pragma unroll
              for (int i = 0; i < 64; ++i) f[i] = gelu_erf(f[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 64; ++i) f[i] = gelu_tanh(f[i]);     // runs while the TMA engine reads the z tile
            }
            if (TEAMS == 2) {
              if (leader) stage_wait();           // EPI_DB: the other tile (last used two groups ago) is free; else wait for the z store
              named_bar_sync(bar_id, EPI_THREADS);
            }
          } else {
            if (p.epilogue == B2_EPI_BIAS_GELU) {
              if constexpr (ERF) {
#pragma unroll
                for (int i = 0; i < 64; ++i) f[i] = gelu_erf(f[i]);
              } else {
#pragma unroll
                for (int i = 0; i < 64; ++i) f[i] = gelu_tanh(f[i]);
              }
            } else if (use_aux) {
              const bool is_res = (p.epilogue == B2_EPI_BIAS_RESIDUAL);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint4 q = aq[j];
                float2 a = unpack2(q.x, fmt), b = unpack2(q.y, fmt), cc = unpack2(q.z, fmt), d = unpack2(q.w, fmt);
                float z[8] = {a.x, a.y, b.x, b.y, cc.x, cc.y, d.x, d.y};
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  if (is_res)            f[8 * j + e] += z[e];
                  else if constexpr (ERF) f[8 * j + e] *= gelu_erf_grad(z[e]);
                  else                   f[8 * j + e] *= gelu_tanh_grad(z[e]);
                }
              }
            }
            [[maybe_unused]] const int pslot = team * 2 + static_cast<int>(rs_pq & 1u);
            if constexpr (FUSE == FUSE_RS) {
              if (!rs_own) {                // remote rows: the chunk goes into one of this team's two push buffers
                if (leader) mbar_wait(&pempty[pslot], ((rs_pq >> 1) & 1u) ^ 1u);
                sb0 = push_bufs + pslot * STAGE_CHUNK_BYTES;
              } else if (leader) {
                tma_store_wait_read<0>();
              }
            } else {
              if (leader) stage_wait();
            }
            named_bar_sync(bar_id, EPI_THREADS);
            if (aux_on && leader) aux_issue_next();       // every thread of the team has copied its aux row out
            if constexpr (FUSE == FUSE_RS) {              // ... and its landing-tile rows: fetch this tile's next chunk
              if (rs_own && leader && c + TEAMS < NCH && col0 + TEAMS * 64 < p.N) rs_issue(rs_lblk, col0 + TEAMS * 64);
            }
          }
          {
            uint8_t* rowp = sb0 + et * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              uint4 q = make_uint4(pack2(f[8 * j], f[8 * j + 1], fmt), pack2(f[8 * j + 2], f[8 * j + 3], fmt),
                                   pack2(f[8 * j + 4], f[8 * j + 5], fmt), pack2(f[8 * j + 6], f[8 * j + 7], fmt));
              *reinterpret_cast<uint4*>(rowp + ((j ^ (et & 7)) << 4)) = q;
            }
          }
          fence_proxy_async_smem();
          named_bar_sync(bar_id, EPI_THREADS);
          if (leader) {
            if constexpr (FUSE == FUSE_RS) {
              // partial tile -> staging slot [epoch parity][this rank] on the OWNER of these rows: the swizzled 16 KB tile
              // goes out verbatim as ONE 1-D bulk copy through the peer mapping (large NVLink writes; the owner undoes the
              // swizzle when it reads), and the transfer overlaps the following tiles' mainloops
              if (rs_own) {
                if (p.f_armaps != nullptr) {      // all-reduce: the reduced rows go to EVERY rank's full output (TMA stores, peers included)
                  for (int r = 0; r < p.f_world; ++r) tma_store_2d(p.f_armaps + r, sb0, col0, m0);
                } else {
                  tma_store_2d(&tmC, sb0, col0, rs_lblk * BM);          // final rows of this rank
                }
                tma_store_commit();
              } else {
                const int ncc = (p.N + 63) >> 6;
                const long long slot_bytes = static_cast<long long>(p.f_shard_rows) * ncc * 128;
                const int pslot = team * 2 + static_cast<int>(rs_pq & 1u);
                pmeta[pslot].dst = p.f_stage_peer[rs_owner] + (static_cast<long long>(f_epoch & 1) * p.f_world + p.f_rank) * slot_bytes +
                                   (static_cast<long long>(rs_lblk) * ncc + (col0 >> 6)) * STAGE_CHUNK_BYTES;
                pmeta[pslot].cnt = p.f_cnt[rs_owner] + rs_lblk;
                mbar_arrive(&pfull[pslot]);           // release: the push warp sees the meta words and (through the team barrier) the tile
              }
            } else {
              tma_store_2d(&tmC, sb0, col0, m0);
              tma_store_commit();
            }
          }
          if constexpr (FUSE == FUSE_RS) { if (!rs_own) ++rs_pq; }
        }
      }
    }
    if constexpr (FUSE != FUSE_NONE) { if (leader && team == 0) dbg_stamp(p, 1); }
    if (leader) tma_store_wait_all<0>();
  }

  tc_fence_before();
  if constexpr (CTA2) cluster_sync(); else __syncthreads();     // pair: neither CTA may free TMEM / exit while the other still uses it
  if constexpr (FUSE != FUSE_NONE) { if (threadIdx.x == 0) dbg_stamp(p, 4); }
  if (warp == 2) { if constexpr (CTA2) tmem_dealloc_2sm<TMEM_COLS>(tmem_base); else tmem_dealloc<TMEM_COLS>(tmem_base); }
  if constexpr (FUSE == FUSE_AG || FUSE == FUSE_AGB) fused_finish(p, f_epoch, true);
  if constexpr (FUSE == FUSE_RS) {
    // (every epilogue leader has waited for the completion of its TMA stores before the CTA-wide barrier above)
    const bool ar = p.f_armaps != nullptr;
    fused_finish(p, f_epoch, ar, ar);
  }
}

template <int BN, bool A_MN, bool B_MN, bool ERF = false>
int launch(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const CUtensorMap& tC2,
           const GemmParams& p, int max_ctas, cudaStream_t stream) {
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, FUSE_NONE, false, ERF>;
  const int smem = SMEM_MAIN_BYTES + 2 * STAGE_CHUNK_BYTES + 1024 /*align slack*/ + 512 /*barriers*/;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { b2host::set_error("gemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    configured = true;
  }
  int total = p.tiles_m * p.tiles_n * p.split_k;
  int grid = total < max_ctas ? total : max_ctas;
  kern<<<grid, NUM_THREADS, smem, stream>>>(tA, tB, tC, tC2, p, tA);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("gemm launch: %s", cudaGetErrorString(e)); return -1; }
  return 0;
}

int g_resident_pairs = 0;

// CTA-pair launch: grid = 2 x (number of 256 x BN tiles, capped at SMs / 2 clusters), cluster dimension (2, 1, 1)
template <int BN, bool A_MN, bool B_MN, bool ERF = false>
int launch_pair(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const CUtensorMap& tC2,
                const GemmParams& p, int max_ctas, cudaStream_t stream) {
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, FUSE_NONE, true, ERF>;
  const int smem = SMEM_MAIN_BYTES + 2 * STAGE_CHUNK_BYTES + 1024 /*align slack*/ + 512 /*barriers*/;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { b2host::set_error("gemm(pair): cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    configured = true;
  }
  const int total = p.tiles_m * p.tiles_n * p.split_k;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NUM_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  // the kernel is persistent: launch exactly as many pairs as can be CO-RESIDENT (a pair needs both SMs of one TPC;
  // parts with an odd number of enabled SMs in some GPCs host fewer than SMs/2 pairs), never a second wave
  static int resident_pairs = 0;
  if (resident_pairs == 0) {
    cfg.gridDim = dim3(2 * (b2host::num_sms() / 2), 1, 1);
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    if (e != cudaSuccess || n <= 0) { b2host::set_error("gemm(pair): cudaOccupancyMaxActiveClusters: %s (n=%d)", cudaGetErrorString(e), n); return -1; }
    resident_pairs = n;
    g_resident_pairs = n;
  }
  int max_clusters = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
  if (max_clusters > resident_pairs) max_clusters = resident_pairs;
  const int clusters = total < max_clusters ? total : max_clusters;
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tA, tB, tC, tC2, p, tA);
  if (e != cudaSuccess) { b2host::set_error("gemm(pair) launch: %s", cudaGetErrorString(e)); return -1; }
  return 0;
}

// 0 = never, 1 = heuristic (default), 2 = whenever the shape allows; set through b2_gemm_set_pair_mode (tuning / A-B tests)
int g_pair_mode = 1;

}  // namespace

extern "C" void b2_gemm_set_pair_mode(int mode) { g_pair_mode = mode; }
// number of CTA pairs of the GEMM kernel that can be co-resident on this GPU (0 before the first pair launch)
extern "C" int b2_gemm_resident_pairs(void) { return g_resident_pairs; }

// Tile shape, CTA pairing and split-K of one plain GEMM: pure host arithmetic (b2_gemm_plan exposes it to the CPU tests)
struct GemmPlan { int bn, pair, tiles_m, tiles_n, num_kb, split_k, kb_per_split; };
static GemmPlan plan_gemm(int M, int N, int K, int out_fp32, int accumulate, int split_k, int sms, int pair_mode) {
  // Tile-N selection: prefer 256 wide tiles unless that leaves most SMs idle.
  const int tiles_m = (M + BM - 1) / BM;
  int bn = 256;
  {
    long long t256 = 1LL * tiles_m * ((N + 255) / 256);
    long long t128 = 1LL * tiles_m * ((N + 127) / 128);
    // fp32-accumulate outputs (wgrad) fill the SMs through split-K instead, so they keep the wide tile
    // (128x256 halves the operand traffic per MMA); everything else drops to 128 when 256 would idle SMs
    const bool can_split = out_fp32 && accumulate && split_k <= 0;
    if (N <= 128 || (!can_split && t256 * (split_k > 0 ? split_k : 1) < sms && t128 > t256)) bn = 128;
  }
  const int tiles_n = (N + bn - 1) / bn;
  const int num_kb = (K + BK - 1) / BK;
  // CTA pairs (cta_group::2, 256 x bn tiles): when there are enough 256-row tiles to keep every pair busy
  bool pair = false;
  if (pair_mode > 0 && M >= 256 && bn >= 128) {
    const long long t2 = 1LL * ((M + 255) / 256) * tiles_n;
    pair = pair_mode == 2 || t2 * (out_fp32 && accumulate && split_k <= 0 ? 4 : 1) >= sms / 2;
  }
  const int tiles_m_eff = pair ? (M + 255) / 256 : tiles_m;

  if (split_k <= 0) {  // auto: only meaningful for fp32 accumulate outputs
    split_k = 1;
    if (out_fp32 && accumulate) {
      // cost in k-block units: rounds over the resident CTAs (or pairs) x (k-blocks per item + ramp), floored by the
      // fp32 reduce epilogue a round has to hide; the smallest split within 4 % of the best keeps the reduce traffic low
      const int tiles = tiles_m_eff * tiles_n;
      const int slots = std::max(1, pair ? sms / 2 : sms);
      auto cost = [&](int s) {
        const int kbs = (num_kb + s - 1) / s;
        const int items = tiles * ((num_kb + kbs - 1) / kbs);
        return 1LL * ((items + slots - 1) / slots) * std::max(kbs + 4, 32);
      };
      long long best = cost(1);
      for (int s = 2; s <= 32 && num_kb / s >= 4; ++s) best = std::min(best, cost(s));
      for (int s = 1; s <= 32 && (s == 1 || num_kb / s >= 4); ++s)
        if (cost(s) * 100 <= best * 104) { split_k = s; break; }
    }
  }
  GemmPlan g;
  g.bn = bn; g.pair = pair ? 1 : 0; g.tiles_m = tiles_m_eff; g.tiles_n = tiles_n; g.num_kb = num_kb;
  g.kb_per_split = (num_kb + split_k - 1) / split_k;
  g.split_k = (num_kb + g.kb_per_split - 1) / g.kb_per_split;
  return g;
}

// out[7] = {bn, pair, tiles_m, tiles_n, num_kb, split_k, kb_per_split} for the given problem on `sms` SMs (pair_mode as in
// b2_gemm_set_pair_mode; split_k <= 0 = automatic).  No device needed.
extern "C" int b2_gemm_plan(int M, int N, int K, int out_fp32, int accumulate, int split_k, int sms, int pair_mode, int* out) {
  if (M <= 0 || N <= 0 || K <= 0 || sms <= 0 || out == nullptr) return -1;
  const GemmPlan g = plan_gemm(M, N, K, out_fp32, accumulate, split_k, sms, pair_mode);
  out[0] = g.bn; out[1] = g.pair; out[2] = g.tiles_m; out[3] = g.tiles_n; out[4] = g.num_kb; out[5] = g.split_k; out[6] = g.kb_per_split;
  return 0;
}

extern "C" int b2_gemm(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                       long long ldc, int a_mn_major, int b_mn_major, int dtype, int epilogue, const void* bias,
                       const void* aux, long long ld_aux, void* C2, long long ldc2, int out_fp32, int accumulate,
                       int split_k, float alpha, int max_ctas, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int act_erf = (epilogue & B2_EPI_ERF) ? 1 : 0;
  epilogue &= ~B2_EPI_ERF;
  if (M <= 0 || N <= 0 || K <= 0) { b2host::set_error("gemm: empty problem"); return -1; }
  if ((lda % 8) || (ldb % 8) || (out_fp32 ? (ldc % 4) : (ldc % 8)) || (N % 8)) {
    b2host::set_error("gemm: leading dimensions / N must keep 16-byte alignment (lda=%lld ldb=%lld ldc=%lld N=%d)", lda,
                      ldb, ldc, N);
    return -1;
  }
  if (out_fp32 && epilogue != B2_EPI_NONE) { b2host::set_error("gemm: fp32 output supports no fused epilogue"); return -1; }
  if (accumulate && !out_fp32) { b2host::set_error("gemm: accumulate requires fp32 output"); return -1; }
  if (split_k > 1 && !(out_fp32 && accumulate)) { b2host::set_error("gemm: split-K requires fp32 accumulate output"); return -1; }
  const CUtensorMapDataType dt = dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;

  const int sms = max_ctas > 0 ? max_ctas : b2host::num_sms();
  const GemmPlan plan = plan_gemm(M, N, K, out_fp32, accumulate, split_k, sms, g_pair_mode);
  const int bn = plan.bn;
  const bool pair = plan.pair != 0;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = plan.tiles_m; p.tiles_n = plan.tiles_n;
  p.num_kb = plan.num_kb;
  p.kb_per_split = plan.kb_per_split;
  p.split_k = plan.split_k;
  p.epilogue = epilogue; p.act_erf = act_erf; p.fmt = dtype == B2_DTYPE_BF16 ? 1 : 0;
  p.out_fp32 = out_fp32; p.reduce_add = accumulate;
  p.bias = bias; p.aux = aux; p.ld_aux = ld_aux; p.alpha = alpha;
  if ((epilogue == B2_EPI_BIAS_RESIDUAL || epilogue == B2_EPI_DGELU) && (aux == nullptr || (ld_aux % 8))) {
    b2host::set_error("gemm: epilogue needs a 16-byte aligned aux tensor"); return -1;
  }
  if (epilogue == B2_EPI_BIAS_GELU_DUAL && (C2 == nullptr || (ldc2 % 8))) {
    b2host::set_error("gemm: dual-output epilogue needs C2"); return -1;
  }

  CUtensorMap tA, tB, tC, tC2;
  {
    uint64_t dims[2], strides[2]; uint32_t box[2];
    if (!a_mn_major) { dims[0] = K; dims[1] = M; box[0] = BK; box[1] = BM; }
    else             { dims[0] = M; dims[1] = K; box[0] = 64; box[1] = BK; }
    strides[0] = 0; strides[1] = static_cast<uint64_t>(lda) * 2;
    if (b2host::encode_tmap(&tA, dt, 2, A, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
    if (!b_mn_major) { dims[0] = K; dims[1] = N; box[0] = BK; box[1] = pair ? bn / 2 : bn; }   // pair: each CTA stages half of B
    else             { dims[0] = N; dims[1] = K; box[0] = 64; box[1] = BK; }
    strides[1] = static_cast<uint64_t>(ldb) * 2;
    if (b2host::encode_tmap(&tB, dt, 2, B, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
    dims[0] = N; dims[1] = M;
    if (out_fp32) {
      box[0] = 32; box[1] = BM; strides[1] = static_cast<uint64_t>(ldc) * 4;
      if (b2host::encode_tmap(&tC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, C, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
      tC2 = tC;
    } else {
      box[0] = 64; box[1] = BM; strides[1] = static_cast<uint64_t>(ldc) * 2;
      if (b2host::encode_tmap(&tC, dt, 2, C, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
      if (epilogue == B2_EPI_BIAS_GELU_DUAL) {
        strides[1] = static_cast<uint64_t>(ldc2) * 2;
        if (b2host::encode_tmap(&tC2, dt, 2, C2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
      } else if (epilogue == B2_EPI_BIAS_RESIDUAL || epilogue == B2_EPI_DGELU) {
        // the aux operand (residual / pre-activation) is prefetched tile by tile with TMA through the C2 map slot
        strides[1] = static_cast<uint64_t>(ld_aux) * 2;
        if (b2host::encode_tmap(&tC2, dt, 2, aux, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
      } else {
        tC2 = tC;
      }
    }
  }
  const int mc = sms;
#define B2_DISPATCH(BN_)                                                                         \
  do {                                                                                           \
    if (!a_mn_major && !b_mn_major) return launch<BN_, false, false>(tA, tB, tC, tC2, p, mc, stream); \
    if (!a_mn_major && b_mn_major) return launch<BN_, false, true>(tA, tB, tC, tC2, p, mc, stream);   \
    if (a_mn_major && b_mn_major) return launch<BN_, true, true>(tA, tB, tC, tC2, p, mc, stream);     \
    return launch<BN_, true, false>(tA, tB, tC, tC2, p, mc, stream);                                  \
  } while (0)
#define B2_DISPATCH_PAIR(BN_)                                                                         \
  do {                                                                                                \
    if (!a_mn_major && !b_mn_major) return launch_pair<BN_, false, false>(tA, tB, tC, tC2, p, mc, stream); \
    if (!a_mn_major && b_mn_major) return launch_pair<BN_, false, true>(tA, tB, tC, tC2, p, mc, stream);   \
    if (a_mn_major && b_mn_major) return launch_pair<BN_, true, true>(tA, tB, tC, tC2, p, mc, stream);     \
    return launch_pair<BN_, true, false>(tA, tB, tC, tC2, p, mc, stream);                                  \
  } while (0)
  if (act_erf) {       // exact-erf GELU epilogues: forward / data-gradient layouts only (A K-major)
    if (a_mn_major) { b2host::set_error("gemm: erf GELU epilogues need a K-major A operand"); return -1; }
    if (pair) {
      if (bn == 256) return b_mn_major ? launch_pair<256, false, true, true>(tA, tB, tC, tC2, p, mc, stream)
                                       : launch_pair<256, false, false, true>(tA, tB, tC, tC2, p, mc, stream);
      return b_mn_major ? launch_pair<128, false, true, true>(tA, tB, tC, tC2, p, mc, stream)
                        : launch_pair<128, false, false, true>(tA, tB, tC, tC2, p, mc, stream);
    }
    if (bn == 256) return b_mn_major ? launch<256, false, true, true>(tA, tB, tC, tC2, p, mc, stream)
                                     : launch<256, false, false, true>(tA, tB, tC, tC2, p, mc, stream);
    return b_mn_major ? launch<128, false, true, true>(tA, tB, tC, tC2, p, mc, stream)
                      : launch<128, false, false, true>(tA, tB, tC, tC2, p, mc, stream);
  }
  if (pair) {
    if (bn == 256) B2_DISPATCH_PAIR(256);
    B2_DISPATCH_PAIR(128);
  }
#undef B2_DISPATCH_PAIR
  if (bn == 256) B2_DISPATCH(256);
  B2_DISPATCH(128);
#undef B2_DISPATCH
}


static unsigned long long* g_fused_dbg = nullptr;
// profiling hook: subsequent fused launches write globaltimer stamps into buf ([SMs][8] uint64), NULL disables
extern "C" void b2_fused_set_debug(void* buf) { g_fused_dbg = static_cast<unsigned long long*>(buf); }

// =============================================================================================
// Fused compute + collective entry points (tensor / sequence parallelism, ZeRO-3 parameter gather)
// =============================================================================================
namespace {

int g_fused_pair_mode = 1;     // 0: single-CTA tiles in the fused kernels, 1: CTA pairs whenever the shard geometry allows
int g_fused_ncomm = 24;        // default number of all-gather CTAs (even); measured on 2 x B200: 24 CTAs x 6 x 32 KB in flight pull ~630 GB/s, 16 CTAs ~400

// grid: number of CTAs (a multiple of 2 when PAIR); PAIR launches clusters of two
template <int BN, bool B_MN, int FUSE, bool PAIR>
int launch_fused(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const CUtensorMap& tC2,
                 const GemmParams& p, int grid, cudaStream_t stream, const CUtensorMap& tA2) {
  auto kern = gemm_tcgen05_kernel<BN, false, B_MN, FUSE, PAIR, false>;
  const int smem = SMEM_MAIN_BYTES + 2 * STAGE_CHUNK_BYTES + 1024 + 512;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { b2host::set_error("fused gemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return -1; }
    configured = true;
  }
  if constexpr (!PAIR) {
    kern<<<grid, NUM_THREADS, smem, stream>>>(tA, tB, tC, tC2, p, tA2);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { b2host::set_error("fused gemm launch: %s", cudaGetErrorString(e)); return -1; }
    return 0;
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(NUM_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tA, tB, tC, tC2, p, tA2);
    if (e != cudaSuccess) { b2host::set_error("fused gemm(pair) launch: %s", cudaGetErrorString(e)); return -1; }
    return 0;
  }
}

// CTA pairs that can be co-resident (every CTA of a fused launch must be: the GEMM CTAs wait for the all-gather CTAs / for
// peers).  Uses the plain pair kernel's answer - same block size, same dynamic shared memory, same cluster shape.
int fused_resident_pairs() {
  static int cached = 0;
  if (cached > 0) return cached;
  auto kern = gemm_tcgen05_kernel<256, false, false, FUSE_AG, true, false>;
  const int smem = SMEM_MAIN_BYTES + 2 * STAGE_CHUNK_BYTES + 1024 + 512;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (b2host::num_sms() / 2), 1, 1);
  cfg.blockDim = dim3(NUM_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
  if (e != cudaSuccess || n <= 0) { cudaGetLastError(); n = 0; }
  cached = n > 0 ? n : -1;
  return cached;
}

int make_ab_maps(CUtensorMap* tA, CUtensorMap* tB, const void* A, const void* B, int M, int N, int K, long long lda,
                 long long ldb, int b_mn, int bn_box, CUtensorMapDataType dt) {
  uint64_t dims[2], strides[2]; uint32_t box[2];
  dims[0] = K; dims[1] = M; box[0] = BK; box[1] = BM; strides[0] = 0; strides[1] = static_cast<uint64_t>(lda) * 2;
  if (b2host::encode_tmap(tA, dt, 2, A, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  if (!b_mn) { dims[0] = K; dims[1] = N; box[0] = BK; box[1] = bn_box; }
  else       { dims[0] = N; dims[1] = K; box[0] = 64; box[1] = BK; }
  strides[1] = static_cast<uint64_t>(ldb) * 2;
  return b2host::encode_tmap(tB, dt, 2, B, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

#define B2_FUSED_DISPATCH(FUSE_, PAIR_)                                                                                  \
  do {                                                                                                                   \
    if (bn == 256) return b_mn_major ? launch_fused<256, true, FUSE_, PAIR_>(tA, tB, tC, tC2, p, grid, stream, tA2)      \
                                     : launch_fused<256, false, FUSE_, PAIR_>(tA, tB, tC, tC2, p, grid, stream, tA2);    \
    return b_mn_major ? launch_fused<128, true, FUSE_, PAIR_>(tA, tB, tC, tC2, p, grid, stream, tA2)                     \
                      : launch_fused<128, false, FUSE_, PAIR_>(tA, tB, tC, tC2, p, grid, stream, tA2);                   \
  } while (0)

}  // namespace

// tuning knobs of the fused kernels: pair_mode 0 = single-CTA tiles, 1 = CTA pairs when the geometry allows (default);
// n_comm = default number of all-gather CTAs (rounded to an even number), <= 0 keeps the current value
extern "C" void b2_fused_config(int pair_mode, int n_comm) {
  if (pair_mode >= 0) g_fused_pair_mode = pair_mode;
  if (n_comm > 0) g_fused_ncomm = (n_comm + 1) & ~1;
}

// All-gather -> GEMM in ONE kernel.  Every rank holds a [shard_rows, K] shard of A in symmetric memory
// (shard_ptrs[r], control blocks ctrls[r]); `n_comm` CTAs pull the shards over NVLink into the local gathered
// operand xg [world*shard_rows, K] (returned to the caller: backward needs it) while the remaining CTAs run
// the tcgen05 GEMM (CTA pairs, 256-row tiles) tile by tile as soon as the rows of a tile have arrived (own shard first).
//   C[M, N] = epilogue(xg @ op(B)),  M = world * shard_rows.
// flags: world * shard_rows / 128 uint32 arrival counters owned by this (buffer, K) pair; zero before the first call and
//        never touched by the caller afterwards (they count pieces cumulatively, in step with the buffer's call epoch).
// replaces: NCCL all_gather_into_tensor + cuBLAS GEMM of Megatron's ColumnParallelLinear forward and
//           RowParallelLinear backward under sequence parallelism, and DeepSpeed's
//           all_gather_coalesced + GEMM when the gathered operand is a ZeRO-3 weight shard (C19).
extern "C" int b2_ag_gemm(void* const* shard_ptrs, void* const* ctrls, void* xg, unsigned int* flags, const void* B, void* C,
                          int shard_rows, int N, int K, long long ldb, long long ldc, int b_mn_major, int dtype, int epilogue,
                          const void* bias, const void* aux, long long ld_aux, void* C2, long long ldc2, int rank, int world,
                          int n_comm, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (world < 2 || world > 8) { b2host::set_error("ag_gemm: world size %d unsupported", world); return -1; }
  if (shard_rows % BM) { b2host::set_error("ag_gemm: rows per rank (%d) must be a multiple of %d", shard_rows, BM); return -1; }
  if ((K % 8) || (N % 8) || (ldb % 8) || (ldc % 8)) { b2host::set_error("ag_gemm: K, N and leading dimensions must be multiples of 8"); return -1; }
  const int M = shard_rows * world;
  const CUtensorMapDataType dt = dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int bn = N > 128 ? 256 : 128;
  const int sms = b2host::num_sms();
  if (n_comm <= 0) n_comm = g_fused_ncomm;
  if (n_comm > sms / 4) n_comm = sms / 4;
  n_comm = (n_comm + 1) & ~1;
  // CTA pairs need 256-row tiles that do not straddle two ranks' shards, and every CTA of the grid co-resident
  const int res_pairs = g_fused_pair_mode ? fused_resident_pairs() : 0;
  const bool pair = g_fused_pair_mode && (shard_rows % (2 * BM) == 0) && res_pairs > n_comm / 2 + 1;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = pair ? M / (2 * BM) : M / BM; p.tiles_n = (N + bn - 1) / bn; p.split_k = 1;
  p.num_kb = (K + BK - 1) / BK; p.kb_per_split = p.num_kb;
  if (epilogue & B2_EPI_ERF) { b2host::set_error("ag_gemm: erf GELU epilogues are not built for the fused kernels"); return -1; }
  p.epilogue = epilogue; p.fmt = dtype == B2_DTYPE_BF16 ? 1 : 0; p.out_fp32 = 0; p.reduce_add = 0;
  p.bias = bias; p.aux = aux; p.ld_aux = ld_aux; p.alpha = 1.f;
  if ((epilogue == B2_EPI_BIAS_RESIDUAL || epilogue == B2_EPI_DGELU) && (aux == nullptr || (ld_aux % 8))) {
    b2host::set_error("ag_gemm: epilogue needs a 16-byte aligned aux tensor"); return -1;
  }
  p.f_rank = rank; p.f_world = world; p.f_bpr = shard_rows / BM; p.f_shard_rows = shard_rows;
  p.f_ncomm = n_comm;
  p.m_rot = pair ? rank * (p.f_bpr / 2) : rank * p.f_bpr;
  p.f_row_bytes = static_cast<long long>(K) * 2;
  p.f_npb = static_cast<int>((128LL * p.f_row_bytes + 32767) / 32768);
  p.f_last_rows = BM; p.f_npb_last = p.f_npb;
  for (int r = 0; r < 8; ++r) { p.f_peer[r] = r < world ? shard_ptrs[r] : nullptr; p.f_ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr; }
  p.f_xg = static_cast<uint8_t*>(xg); p.f_flags = flags; p.f_dbg = g_fused_dbg;
  CUtensorMap tA, tB, tC, tC2;
  if (make_ab_maps(&tA, &tB, xg, B, M, N, K, K, ldb, b_mn_major, pair ? bn / 2 : bn, dt)) return -1;
  uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(M)}, strides[2] = {0, static_cast<uint64_t>(ldc) * 2};
  uint32_t box[2] = {64, BM};
  if (b2host::encode_tmap(&tC, dt, 2, C, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  tC2 = tC;
  if (epilogue == B2_EPI_BIAS_GELU_DUAL) {
    if (C2 == nullptr) { b2host::set_error("ag_gemm: dual epilogue needs C2"); return -1; }
    strides[1] = static_cast<uint64_t>(ldc2) * 2;
    if (b2host::encode_tmap(&tC2, dt, 2, C2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  } else if (epilogue == B2_EPI_BIAS_RESIDUAL || epilogue == B2_EPI_DGELU) {
    strides[1] = static_cast<uint64_t>(ld_aux) * 2;        // aux tiles are prefetched by TMA through the C2 map slot
    if (b2host::encode_tmap(&tC2, dt, 2, aux, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  }
  CUtensorMap tA2;
  {
    uint64_t d2[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(shard_rows)}, s2[2] = {0, static_cast<uint64_t>(K) * 2};
    uint32_t b2[2] = {BK, BM};
    if (b2host::encode_tmap(&tA2, dt, 2, shard_ptrs[rank], d2, s2, b2, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  }
  // one CTA per SM: all-gather CTAs and GEMM CTAs must be co-resident
  if (pair) {
    int pairs = sms / 2 < res_pairs ? sms / 2 : res_pairs;
    const int grid = 2 * pairs;
    B2_FUSED_DISPATCH(FUSE_AG, true);
  }
  const int grid = sms;
  B2_FUSED_DISPATCH(FUSE_AG, false);
}

// All-gather of a ROW-SHARDED WEIGHT fused into the GEMM that consumes it (ZeRO-3 parameter fetch):
//   C[M, N] = A[M, K] @ Wfull^T,   Wfull [N, K] = the `world` shards [shard_rows, K] stacked, N = world * shard_rows.
// `wg` is this rank's full-size gathered weight in symmetric memory with its OWN shard already in place at rows
// [rank * shard_rows, ...); shard_ptrs[r] = address of rank r's own shard inside rank r's copy of that buffer.  The
// all-gather CTAs pull the world - 1 remote shards (own region never moves) in row order into `wg`; the GEMM CTAs walk the
// column tiles N-major, starting with this rank's own rows of W, and wait per 128-row block of W for the counters.  After
// the launch `wg` holds the complete weight (the backward and later calls read it there).  shard_rows need not be a
// multiple of 128 (LLaMA's 32000-row head over 8 ranks: 4000); flags: world * ceil(shard_rows / 128) uint32, zero before
// the first call (the launch leaves them zero).
extern "C" int b2_agw_gemm(void* const* shard_ptrs, void* const* ctrls, void* wg, unsigned int* flags, const void* A, void* C,
                           int M, int shard_rows, int K, long long lda, long long ldc, int dtype, int rank, int world,
                           int n_comm, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (world < 2 || world > 8) { b2host::set_error("agw_gemm: world size %d unsupported", world); return -1; }
  if ((K % 8) || (lda % 8) || (ldc % 8) || (shard_rows % 8) || M <= 0) { b2host::set_error("agw_gemm: K, shard_rows and leading dimensions must be multiples of 8"); return -1; }
  const int N = shard_rows * world;
  const CUtensorMapDataType dt = dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int sms = b2host::num_sms();
  if (n_comm <= 0) n_comm = g_fused_ncomm;
  if (n_comm > sms / 4) n_comm = sms / 4;
  n_comm = (n_comm + 1) & ~1;
  const int res_pairs = g_fused_pair_mode ? fused_resident_pairs() : 0;
  const bool pair = g_fused_pair_mode && M >= 2 * BM && res_pairs > n_comm / 2 + 1;
  const int bn = pair ? 256 : 128;          // single-CTA: 128-wide tiles (a CTA then stages exactly one 128-row block of W)
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = pair ? (M + 2 * BM - 1) / (2 * BM) : (M + BM - 1) / BM; p.tiles_n = (N + bn - 1) / bn; p.split_k = 1;
  p.num_kb = (K + BK - 1) / BK; p.kb_per_split = p.num_kb;
  p.epilogue = B2_EPI_NONE; p.fmt = dtype == B2_DTYPE_BF16 ? 1 : 0; p.alpha = 1.f;
  p.f_rank = rank; p.f_world = world; p.f_shard_rows = shard_rows; p.f_bpr = (shard_rows + BM - 1) / BM;
  p.f_ncomm = n_comm;
  p.f_row_bytes = static_cast<long long>(K) * 2;
  p.f_npb = static_cast<int>((128LL * p.f_row_bytes + 32767) / 32768);
  p.f_last_rows = shard_rows - (p.f_bpr - 1) * BM;
  p.f_npb_last = static_cast<int>((static_cast<long long>(p.f_last_rows) * p.f_row_bytes + 32767) / 32768);
  p.f_nmajor = 1; p.n_rot = (rank * shard_rows) / bn;      // the column tile that holds the first row of the own shard
  p.f_self_clean = 1;
  for (int r = 0; r < 8; ++r) { p.f_peer[r] = r < world ? shard_ptrs[r] : nullptr; p.f_ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr; }
  p.f_xg = static_cast<uint8_t*>(wg); p.f_flags = flags; p.f_dbg = g_fused_dbg;
  CUtensorMap tA, tB, tC;
  if (make_ab_maps(&tA, &tB, A, wg, M, N, K, lda, K, 0, pair ? bn / 2 : bn, dt)) return -1;
  uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(M)}, strides[2] = {0, static_cast<uint64_t>(ldc) * 2};
  uint32_t box[2] = {64, BM};
  if (b2host::encode_tmap(&tC, dt, 2, C, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  const CUtensorMap& tC2 = tC;
  const CUtensorMap& tA2 = tA;
  const int b_mn_major = 0;
  if (pair) {
    int pairs = sms / 2 < res_pairs ? sms / 2 : res_pairs;
    const int grid = 2 * pairs;
    if (bn == 256) return launch_fused<256, false, FUSE_AGB, true>(tA, tB, tC, tC2, p, grid, stream, tA2);
  }
  (void)b_mn_major;
  const int grid = sms;
  return launch_fused<128, false, FUSE_AGB, false>(tA, tB, tC, tC2, p, grid, stream, tA2);
}

// GEMM -> reduce-scatter in ONE kernel.  Each rank multiplies its K-slice (partial sums for ALL M rows); the
// epilogue sends every partial tile straight into the staging slot of the rank that OWNS those rows (remote rows first, so
// NVLink traffic overlaps the remaining tiles; a tile chunk leaves as one 16 KB bulk copy of the swizzled shared-memory
// tile), bumps an arrival counter there, and the same kernel then folds the `world` partials of its own rows with bias
// and residual into the output shard.
//   out[shard_rows, N] = residual + bias + sum_ranks (A_r @ op(B_r))[rows of this rank]
// stage_ptrs[r]: rank r's staging buffer (symmetric), 2 * world * shard_rows * ceil(N/64) * 64 elements:
//                [call parity][source rank][row block][64-column chunk][128 rows x 128 B, swizzled]
// cnt_ptrs[r]:   rank r's per-row-block arrival counters (symmetric, zero before the first call)
// replaces: cuBLAS GEMM + NCCL reduce_scatter_tensor (+ bias-dropout-add) of RowParallelLinear forward and
//           ColumnParallelLinear backward under sequence parallelism.
static int gemm_rs_impl(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, int b_mn_major,
                        int dtype, void* const* cnt_ptrs, void* const* ctrls, void* const* stage_ptrs,
                        void* out, long long ld_out, const void* bias, const void* residual, long long ld_res, int rank,
                        int world, const void* armaps_dev, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (world < 2 || world > 8) { b2host::set_error("gemm_rs: world size %d unsupported", world); return -1; }
  if (M % (world * BM)) { b2host::set_error("gemm_rs: M (%d) must be a multiple of world * %d", M, BM); return -1; }
  if ((K % 8) || (N % 8) || (lda % 8) || (ldb % 8) || (ld_out % 8) || (residual && (ld_res % 8))) {
    b2host::set_error("gemm_rs: K, N and leading dimensions must be multiples of 8"); return -1;
  }
  const CUtensorMapDataType dt = dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int sms = b2host::num_sms();
  const int shard_rows = M / world;
  const int res_pairs = g_fused_pair_mode ? fused_resident_pairs() : 0;
  const bool pair = g_fused_pair_mode && (shard_rows % (2 * BM) == 0) && res_pairs > 0;
  // the push buffers take 64 KB of the operand ring: single-CTA 128 x 256 tiles (48 KB stages) would be left with two stages
  const int bn = (N > 128 && pair) ? 256 : 128;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = pair ? M / (2 * BM) : M / BM; p.tiles_n = (N + bn - 1) / bn; p.split_k = 1;
  p.num_kb = (K + BK - 1) / BK; p.kb_per_split = p.num_kb;
  p.epilogue = B2_EPI_NONE; p.fmt = dtype == B2_DTYPE_BF16 ? 1 : 0; p.alpha = 1.f;
  p.f_rank = rank; p.f_world = world; p.f_shard_rows = shard_rows; p.f_bpr = shard_rows / BM; p.f_ncomm = 0;
  p.m_rot = ((rank + 1) % world) * (pair ? p.f_bpr / 2 : p.f_bpr);
  for (int r = 0; r < 8; ++r) {
    p.f_cnt[r] = r < world ? static_cast<uint32_t*>(cnt_ptrs[r]) : nullptr;
    p.f_ctrl[r] = r < world ? static_cast<uint32_t*>(ctrls[r]) : nullptr;
    p.f_stage_peer[r] = r < world ? static_cast<uint8_t*>(stage_ptrs[r]) : nullptr;
  }
  p.f_dbg = g_fused_dbg;
  p.f_armaps = static_cast<const CUtensorMap*>(armaps_dev);
  p.f_stage = stage_ptrs[rank]; p.f_out = out; p.f_ld_out = ld_out; p.f_bias = bias; p.f_res = residual; p.f_ld_res = ld_res;
  CUtensorMap tA, tB, tC;
  if (make_ab_maps(&tA, &tB, A, B, M, N, K, lda, ldb, b_mn_major, pair ? bn / 2 : bn, dt)) return -1;
  {
    uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(p.f_shard_rows)}, strides[2] = {0, static_cast<uint64_t>(ld_out) * 2};
    uint32_t box[2] = {64, BM};
    if (b2host::encode_tmap(&tC, dt, 2, out, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  }
  CUtensorMap tC2 = tC;
  if (residual != nullptr) {       // the owner's reduce epilogue fetches residual tiles by TMA through the C2 map slot
    if (armaps_dev != nullptr)     // all-reduce: `residual` is the full [M, N] tensor; this rank reduces rows [rank * shard_rows, ...)
      residual = static_cast<const uint8_t*>(residual) + static_cast<long long>(rank) * p.f_shard_rows * ld_res * 2;
    uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(p.f_shard_rows)}, strides[2] = {0, static_cast<uint64_t>(ld_res) * 2};
    uint32_t box[2] = {64, BM};
    if (b2host::encode_tmap(&tC2, dt, 2, residual, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  }
  const CUtensorMap& tA2 = tA;
  const int total = p.tiles_m * p.tiles_n;
  if (pair) {
    int pairs = sms / 2 < res_pairs ? sms / 2 : res_pairs;
    if (total < pairs) pairs = total;
    const int grid = 2 * pairs;
    B2_FUSED_DISPATCH(FUSE_RS, true);
  }
  const int grid = total < sms ? total : sms;
  B2_FUSED_DISPATCH(FUSE_RS, false);
}
#undef B2_FUSED_DISPATCH

extern "C" int b2_gemm_rs(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, int b_mn_major,
                          int dtype, void* const* cnt_ptrs, void* const* ctrls, void* const* stage_ptrs,
                          void* out, long long ld_out, const void* bias, const void* residual, long long ld_res, int rank,
                          int world, void* stream_) {
  return gemm_rs_impl(A, B, M, N, K, lda, ldb, b_mn_major, dtype, cnt_ptrs, ctrls, stage_ptrs, out, ld_out, bias, residual, ld_res, rank,
                      world, nullptr, stream_);
}

// Uploads (synchronously, once per output buffer / shape) the `world` tensor maps of the GEMM->all-reduce kernel:
// map[r] = rank r's full [M, N] output (row stride ld_out elements) in symmetric memory, boxes of 128 rows x 64 columns.
extern "C" int b2_make_ar_maps(void* dev_out, void* const* out_ptrs, int world, int M, int N, long long ld_out, int dtype) {
  CUtensorMap maps[8];
  const CUtensorMapDataType dt = dtype == B2_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  if (world < 2 || world > 8) { b2host::set_error("make_ar_maps: world size %d unsupported", world); return -1; }
  for (int r = 0; r < world; ++r) {
    uint64_t dims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(M)}, strides[2] = {0, static_cast<uint64_t>(ld_out) * 2};
    uint32_t box[2] = {64, BM};
    if (b2host::encode_tmap(&maps[r], dt, 2, out_ptrs[r], dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)) return -1;
  }
  cudaError_t e = cudaMemcpy(dev_out, maps, sizeof(CUtensorMap) * world, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { b2host::set_error("make_ar_maps: %s", cudaGetErrorString(e)); return -1; }
  return 0;
}

// GEMM -> all-reduce in ONE kernel (tensor parallelism WITHOUT sequence parallelism: RowParallelLinear forward and
// ColumnParallelLinear backward all-reduce their [tokens, hidden] result over the TP group).  Same kernel as b2_gemm_rs -
// partial tiles of rows another rank reduces are pushed to it while later tiles are still being multiplied - except that
// the rank that has reduced a row block (+ bias + residual) TMA-stores it into EVERY rank's full output, and the launch
// ends with a cross-GPU barrier, after which out_ptrs[rank] holds all M rows on every rank.
//   out[M, N] (on all ranks) = residual + bias + sum_ranks (A_r @ op(B_r));   residual: the full [M, N] tensor or NULL
// armaps_dev: from b2_make_ar_maps for these out_ptrs.  out buffers are symmetric memory.
// replaces: cuBLAS GEMM + NCCL all_reduce (Megatron `reduce_from_tensor_model_parallel_region`, the logged TP=4 run with
//           sequence_parallel False, 3_training_megatron-lm.ipynb:1162).
extern "C" int b2_gemm_ar(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, int b_mn_major,
                          int dtype, void* const* cnt_ptrs, void* const* ctrls, void* const* stage_ptrs, const void* armaps_dev,
                          void* out_local, long long ld_out, const void* bias, const void* residual, long long ld_res, int rank,
                          int world, void* stream_) {
  if (armaps_dev == nullptr) { b2host::set_error("gemm_ar: output tensor maps missing (b2_make_ar_maps)"); return -1; }
  return gemm_rs_impl(A, B, M, N, K, lda, ldb, b_mn_major, dtype, cnt_ptrs, ctrls, stage_ptrs, out_local, ld_out, bias, residual, ld_res,
                      rank, world, armaps_dev, stream_);
}

namespace b2host {
int set_spin_timeout_gemm(unsigned long long ns) {
  return cudaMemcpyToSymbol(b2::g_b2_spin_timeout_ns, &ns, sizeof(ns)) == cudaSuccess ? 0 : -1;
}
}  // namespace b2host
