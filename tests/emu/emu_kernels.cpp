// Driver that runs the SIMT kernels of csrc/{softmax,conv,batchnorm}.cu under the CPU emulation in cuda_emu.h.
// Built by tests/test_kernel_emulation_cpu.py with g++ (no nvcc, no GPU); each entry point mirrors the launch of the
// corresponding extern "C" host wrapper in the .cu file (same grid shape rules, smaller caps).
#define B2_EMU 1
#include "cuda_emu.h"
#include "b200dist.h"

namespace k_softmax {
#include "softmax.cu"
}
namespace k_conv {
#include "conv.cu"
}
namespace k_bn {
#include "batchnorm.cu"
}
#include "vec_emu.h"
namespace k_norm {
#include "norm.cu"
}
#include "mma_emu.h"
namespace k_wtc {
#include "window_attn_tc.cu"
}

namespace {
template <typename F>
void with_dtype(int dtype, F f) {            // 0 fp16, 1 bf16, 2 fp32 (B2_DTYPE_*)
  if (dtype == 1) f(static_cast<__nv_bfloat16*>(nullptr));
  else if (dtype == 0) f(static_cast<__half*>(nullptr));
  else f(static_cast<float*>(nullptr));
}
}  // namespace

extern "C" {

void emu_softmax_fwd(const void* x, const unsigned char* mask, void* y, long long rows, int sq, int sk, int heads, int mask_batches,
                     float scale, int causal, int dtype, int grid) {
  with_dtype(dtype, [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    emu::launch(dim3(grid), dim3(128), sizeof(float) * sk, [&] {
      k_softmax::softmax_fwd_kernel<T>(static_cast<const T*>(x), mask, static_cast<T*>(y), rows, sq, sk, heads, mask_batches, scale, causal);
    });
  });
}

void emu_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk, float scale, int dtype, int grid) {
  with_dtype(dtype, [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    emu::launch(dim3(grid), dim3(128), 0, [&] {
      k_softmax::softmax_bwd_kernel<T>(static_cast<const T*>(dy), static_cast<const T*>(y), static_cast<T*>(dx), rows, sk, scale);
    });
  });
}

#define EMU_K_SWITCH(K, BODY)                                             \
  switch (K) {                                                            \
    case 1: { constexpr int KV = 1; BODY; break; }                        \
    case 2: { constexpr int KV = 2; BODY; break; }                        \
    case 3: { constexpr int KV = 3; BODY; break; }                        \
    case 4: { constexpr int KV = 4; BODY; break; }                        \
    default: { constexpr int KV = 5; BODY; break; }                       \
  }

void emu_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout, int K, int relu) {
  const size_t smem = (static_cast<size_t>(Cin) * H * W + 8u * Cin * K * K) * sizeof(float);
  dim3 grid(B, (Cout + 7) / 8);
  EMU_K_SWITCH(K, emu::launch(grid, dim3(256), smem, [&] { k_conv::conv2d_fwd_kernel<KV>(x, w, bias, y, Cin, H, W, Cout, relu); }));
}

void emu_conv2d_bwd(const float* dy, const float* yact, const float* x, const float* w, float* dx, float* dw, float* db, int B, int Cin, int H,
                    int W, int Cout, int K, int relu, int accumulate) {
  const int Ho = H - K + 1, Wo = W - K + 1;
  if (dx) {
    const size_t smem = (static_cast<size_t>(Cout) * Ho * Wo + static_cast<size_t>(Cout) * 8 * K * K) * sizeof(float);
    dim3 grid(B, (Cin + 7) / 8);
    EMU_K_SWITCH(K, emu::launch(grid, dim3(256), smem, [&] { k_conv::conv2d_dgrad_kernel<KV>(dy, yact, w, dx, Cin, H, W, Cout, relu); }));
  }
  if (dw) {
    dim3 grid(Cout, Cin);
    EMU_K_SWITCH(K, emu::launch(grid, dim3(256), 0, [&] { k_conv::conv2d_wgrad_kernel<KV>(dy, yact, x, dw, db, B, Cin, H, W, Cout, relu, accumulate); }));
  }
}

void emu_maxpool2_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, int grid) {
  emu::launch(dim3(grid), dim3(256), 0, [&] { k_conv::maxpool2_fwd_kernel(x, y, idx, planes, H, W); });
}

void emu_maxpool2_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, int grid) {
  emu::launch(dim3(grid), dim3(256), 0, [&] { k_conv::maxpool2_bwd_kernel(dy, idx, dx, planes, H, W); });
}

void emu_log_softmax_nll_fwd(const float* x, const long long* target, float* logp, float* loss, long long rows, int C) {
  emu::launch(dim3(static_cast<unsigned>((rows * 32 + 255) / 256)), dim3(256), 0,
              [&] { k_conv::log_softmax_nll_fwd_kernel(x, target, logp, loss, rows, C); });
}

void emu_log_softmax_nll_bwd(const float* logp, const float* dlogp, const long long* target, const float* gloss, float gscalar, float* dx,
                             long long rows, int C) {
  emu::launch(dim3(static_cast<unsigned>((rows * 32 + 255) / 256)), dim3(256), 0,
              [&] { k_conv::log_softmax_nll_bwd_kernel(logp, dlogp, target, gloss, gscalar, dx, rows, C); });
}

// BatchNorm: `wide` != 0 instantiates the 64-bit index variants (used on the GPU only for tensors of >= 2^31 elements)
void emu_bn_train_fwd(const void* x, const float* gamma, const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                      float* mean, float* rstd, void* y, double* sums, int N, int C, int HW, int relu, int dtype, int splits, int wide) {
  const long long total = static_cast<long long>(N) * C * HW;
  std::vector<float> scale(C), shift(C);
  with_dtype(dtype, [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    auto run = [&](auto idx_tag) {
      using I = decltype(idx_tag);
      emu::launch(dim3(C, splits), dim3(256), 0, [&] { k_bn::bn_reduce_kernel<T, I>(static_cast<const T*>(x), sums, N, C, HW); });
      emu::launch(dim3((C + 127) / 128), dim3(128), 0, [&] {
        k_bn::bn_finalize_kernel(sums, static_cast<double>(N) * HW, gamma, beta, eps, momentum, running_mean, running_var, mean, rstd, scale.data(),
                                 shift.data(), C);
      });
      emu::launch(dim3(3), dim3(256), 0, [&] {
        k_bn::bn_apply_kernel<T, I>(static_cast<const T*>(x), scale.data(), shift.data(), static_cast<T*>(y), total, C, HW, relu);
      });
    };
    if (wide) run(static_cast<long long>(0)); else run(0u);
  });
}

void emu_bn_bwd(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const float* gamma, double* sums, void* dx,
                int N, int C, int HW, int relu, int dtype, int splits, int wide) {
  const long long total = static_cast<long long>(N) * C * HW;
  with_dtype(dtype, [&](auto* tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    auto run = [&](auto idx_tag) {
      using I = decltype(idx_tag);
      emu::launch(dim3(C, splits), dim3(256), 0, [&] {
        k_bn::bn_bwd_reduce_kernel<T, I>(static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, sums, N, C, HW, relu);
      });
      emu::launch(dim3(3), dim3(256), 0, [&] {
        k_bn::bn_bwd_apply_kernel<T, I>(static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, gamma, sums,
                                        1.0 / (static_cast<double>(N) * HW), static_cast<T*>(dx), total, C, HW, relu);
      });
    };
    if (wide) run(static_cast<long long>(0)); else run(0u);
  });
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// The same entry-point NAMES and SIGNATURES as include/b200dist.h, so that the Python bindings (ops/native.py: argument
// order, ctypes signature table) can be exercised end to end against the emulated kernels.  `stream` is ignored.
// ---------------------------------------------------------------------------------------------------------------------
extern "C" {

int b2_softmax_fwd(const void* x, const unsigned char* mask, void* y, long long rows, int sq, int sk, int heads, int mask_batches, float scale,
                   int causal, int dtype, void*) {
  if (rows <= 0 || sk <= 0 || rows % (static_cast<long long>(sq) * heads) != 0) return -1;
  emu_softmax_fwd(x, mask, y, rows, sq, sk, heads, mask_batches, scale, causal, dtype, static_cast<int>(std::min<long long>(rows, 3)));
  return 0;
}
int b2_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk, float scale, int dtype, void*) {
  emu_softmax_bwd(dy, y, dx, rows, sk, scale, dtype, static_cast<int>(std::min<long long>(rows, 3)));
  return 0;
}
int b2_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout, int K, int relu, void*) {
  emu_conv2d_fwd(x, w, bias, y, B, Cin, H, W, Cout, K, relu);
  return 0;
}
int b2_conv2d_bwd(const float* dy, const float* yact, const float* x, const float* w, float* dx, float* dw, float* db, int B, int Cin, int H, int W,
                  int Cout, int K, int relu, int accumulate, void*) {
  emu_conv2d_bwd(dy, yact, x, w, dx, dw, db, B, Cin, H, W, Cout, K, relu, accumulate);
  return 0;
}
int b2_maxpool2_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, void*) { emu_maxpool2_fwd(x, y, idx, planes, H, W, 2); return 0; }
int b2_maxpool2_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, void*) { emu_maxpool2_bwd(dy, idx, dx, planes, H, W, 2); return 0; }
int b2_log_softmax_nll_fwd(const float* x, const long long* target, float* logp, float* loss, long long rows, int C, void*) {
  emu_log_softmax_nll_fwd(x, target, logp, loss, rows, C);
  return 0;
}
int b2_log_softmax_nll_bwd(const float* logp, const float* dlogp, const long long* target, const float* gloss, float gscalar, float* dx, long long rows,
                           int C, void*) {
  emu_log_softmax_nll_bwd(logp, dlogp, target, gloss, gscalar, dx, rows, C);
  return 0;
}

#define EMU_BN(dtype, BODY)                                                                         \
  with_dtype(dtype, [&](auto* tag) { using T = std::remove_pointer_t<decltype(tag)>; using I = unsigned; BODY; })

int b2_bn_reduce(const void* x, double* sums, int N, int C, int HW, int dtype, void*) {
  EMU_BN(dtype, emu::launch(dim3(C, 2), dim3(256), 0, [&] { k_bn::bn_reduce_kernel<T, I>(static_cast<const T*>(x), sums, N, C, HW); }));
  return 0;
}
int b2_bn_finalize(const double* sums, double count, const float* gamma, const float* beta, float eps, float momentum, float* running_mean,
                   float* running_var, float* mean, float* rstd, float* scale, float* shift, int C, void*) {
  emu::launch(dim3((C + 127) / 128), dim3(128), 0, [&] {
    k_bn::bn_finalize_kernel(sums, count, gamma, beta, eps, momentum, running_mean, running_var, mean, rstd, scale, shift, C);
  });
  return 0;
}
int b2_bn_eval_coeffs(const float* running_mean, const float* running_var, const float* gamma, const float* beta, float eps, float* scale, float* shift,
                      int C, void*) {
  emu::launch(dim3((C + 127) / 128), dim3(128), 0, [&] { k_bn::bn_eval_coeffs_kernel(running_mean, running_var, gamma, beta, eps, scale, shift, C); });
  return 0;
}
int b2_bn_apply(const void* x, const float* scale, const float* shift, void* y, int N, int C, int HW, int relu, int dtype, void*) {
  const long long total = static_cast<long long>(N) * C * HW;
  EMU_BN(dtype, emu::launch(dim3(3), dim3(256), 0, [&] { k_bn::bn_apply_kernel<T, I>(static_cast<const T*>(x), scale, shift, static_cast<T*>(y), total, C, HW, relu); }));
  return 0;
}
int b2_bn_bwd_reduce(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, double* sums, int N, int C, int HW, int relu,
                     int dtype, void*) {
  if (relu && !y) return -1;
  EMU_BN(dtype, emu::launch(dim3(C, 2), dim3(256), 0, [&] {
           k_bn::bn_bwd_reduce_kernel<T, I>(static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, sums, N, C, HW, relu);
         }));
  return 0;
}
int b2_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const float* gamma, const double* sums,
                    double count, void* dx, int N, int C, int HW, int relu, int dtype, void*) {
  const long long total = static_cast<long long>(N) * C * HW;
  EMU_BN(dtype, emu::launch(dim3(3), dim3(256), 0, [&] {
           k_bn::bn_bwd_apply_kernel<T, I>(static_cast<const T*>(dy), static_cast<const T*>(x), static_cast<const T*>(y), mean, rstd, gamma, sums, 1.0 / count,
                                           static_cast<T*>(dx), total, C, HW, relu);
         }));
  return 0;
}

}  // extern "C"

// tensor-core window attention (mma.sync / ldmatrix / cp.async emulated by mma_emu.h); grid (grid_x, heads) like the host launchers
extern "C" {
void emu_wattn_tc_fwd(const void* qkv, const float* bias, const float* mask, void* out, float* lse, long long Bw, int nW, int heads,
                      float scale, int dtype, int grid_x) {
  namespace W = k_wtc::wtc;
  if (dtype == 1)
    emu::launch(dim3(grid_x, heads), dim3(W::THREADS), W::FWD_SMEM, [&] {
      W::wattn_tc_fwd_kernel<__nv_bfloat16>(static_cast<const __nv_bfloat16*>(qkv), bias, mask, static_cast<__nv_bfloat16*>(out), lse, Bw, nW, heads, scale);
    });
  else
    emu::launch(dim3(grid_x, heads), dim3(W::THREADS), W::FWD_SMEM, [&] {
      W::wattn_tc_fwd_kernel<__half>(static_cast<const __half*>(qkv), bias, mask, static_cast<__half*>(out), lse, Bw, nW, heads, scale);
    });
}

void emu_wattn_tc_bwd(const void* qkv, const float* bias, const float* mask, const void* out, const void* dout, const float* lse, void* dqkv,
                      float* dbias, long long Bw, int nW, int heads, float scale, int dtype, int grid_x) {
  namespace W = k_wtc::wtc;
  if (dtype == 1)
    emu::launch(dim3(grid_x, heads), dim3(W::THREADS), W::BWD_SMEM, [&] {
      W::wattn_tc_bwd_kernel<__nv_bfloat16>(static_cast<const __nv_bfloat16*>(qkv), bias, mask, static_cast<const __nv_bfloat16*>(out),
                                            static_cast<const __nv_bfloat16*>(dout), lse, static_cast<__nv_bfloat16*>(dqkv), dbias, Bw, nW, heads, scale);
    });
  else
    emu::launch(dim3(grid_x, heads), dim3(W::THREADS), W::BWD_SMEM, [&] {
      W::wattn_tc_bwd_kernel<__half>(static_cast<const __half*>(qkv), bias, mask, static_cast<const __half*>(out), static_cast<const __half*>(dout), lse,
                                     static_cast<__half*>(dqkv), dbias, Bw, nW, heads, scale);
    });
}
}  // extern "C"

// LayerNorm / RMSNorm (csrc/norm.cu) with the launch rules of its host dispatch (team by row width, rows per CTA, workspace fold);
// `grid_cap` stands in for the SM-count caps so that small inputs still exercise the row loops
extern "C" {
int emu_norm_team(int H) { return k_norm::team_for(H); }

void emu_norm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int rows, int H, float eps,
                  const int* row_map, int dtype, int rms, int grid_cap) {
  auto run = [&](auto* tag, auto rms_tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    constexpr bool RMS = decltype(rms_tag)::value;
    const T *xp = static_cast<const T*>(x), *gp = static_cast<const T*>(gamma), *bp = static_cast<const T*>(beta);
    T* yp = static_cast<T*>(y);
    if (H <= 1024) {
      const int team = k_norm::team_for(H), rpc = 128 / team;
      const int grid = std::min((rows + rpc - 1) / rpc, grid_cap);
      if (team == 8) emu::launch(dim3(grid), dim3(128), 0, [&] { k_norm::norm_fwd_kernel<T, 8, RMS>(xp, gp, bp, yp, mean, rstd, rows, H, eps, row_map); });
      else if (team == 16) emu::launch(dim3(grid), dim3(128), 0, [&] { k_norm::norm_fwd_kernel<T, 16, RMS>(xp, gp, bp, yp, mean, rstd, rows, H, eps, row_map); });
      else emu::launch(dim3(grid), dim3(128), 0, [&] { k_norm::norm_fwd_kernel<T, 32, RMS>(xp, gp, bp, yp, mean, rstd, rows, H, eps, row_map); });
    } else {
      emu::launch(dim3(std::min(rows, grid_cap)), dim3(256), 0, [&] { k_norm::norm_fwd_kernel<T, 256, RMS>(xp, gp, bp, yp, mean, rstd, rows, H, eps, row_map); });
    }
  };
  auto with_rms = [&](auto* tag) { if (rms) run(tag, std::true_type{}); else run(tag, std::false_type{}); };
  if (dtype == 1) with_rms(static_cast<__nv_bfloat16*>(nullptr)); else with_rms(static_cast<__half*>(nullptr));
}

void emu_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd, const void* dres, void* dx,
                  float* dgamma, float* dbeta, int rows, int H, const int* row_map, int dtype, int rms, int accumulate, int grid_cap) {
  auto run = [&](auto* tag, auto rms_tag) {
    using T = std::remove_pointer_t<decltype(tag)>;
    constexpr bool RMS = decltype(rms_tag)::value;
    const T *dyp = static_cast<const T*>(dy), *xp = static_cast<const T*>(x), *gp = static_cast<const T*>(gamma), *rp = static_cast<const T*>(dres);
    T* dxp = static_cast<T*>(dx);
    int grid;
    std::vector<float> ws;
#define EMU_NB(TEAM_, NV_, THREADS_, SMEM_) emu::launch(dim3(grid), dim3(THREADS_), SMEM_, [&] { \
      k_norm::norm_bwd_kernel<T, TEAM_, RMS, NV_>(dyp, xp, gp, mean, rstd, rp, dxp, ws.data(), rows, H, row_map); })
    if (H <= 1024) {
      const int team = k_norm::team_for(H), rpc = 128 / team, nvec = H >> 3, nv = (nvec + team - 1) / team;
      grid = std::min((rows + rpc - 1) / rpc, grid_cap);
      ws.assign(static_cast<size_t>(grid) * 2 * H, std::nanf(""));
      const size_t smem = static_cast<size_t>(rpc) * 2 * H * sizeof(float);
      if (team == 8) { if (nv <= 1) EMU_NB(8, 1, 128, smem); else EMU_NB(8, 2, 128, smem); }
      else if (team == 16) EMU_NB(16, 2, 128, smem);
      else { if (nv <= 2) EMU_NB(32, 2, 128, smem); else if (nv == 3) EMU_NB(32, 3, 128, smem); else EMU_NB(32, 4, 128, smem); }
    } else {
      grid = std::min(rows, grid_cap);
      ws.assign(static_cast<size_t>(grid) * 2 * H, std::nanf(""));
      const int nv = (H + 2047) / 2048;
      if (nv == 1) EMU_NB(256, 1, 256, 0); else if (nv == 2) EMU_NB(256, 2, 256, 0); else if (nv == 3) EMU_NB(256, 3, 256, 0); else EMU_NB(256, 4, 256, 0);
    }
#undef EMU_NB
    const int nparts = grid;
    emu::launch(dim3((2 * H + 31) / 32), dim3(1024), 0, [&] { k_norm::norm_bwd_fold_kernel(ws.data(), nparts, H, dgamma, rms ? nullptr : dbeta, accumulate); });
  };
  auto with_rms = [&](auto* tag) { if (rms) run(tag, std::true_type{}); else run(tag, std::false_type{}); };
  if (dtype == 1) with_rms(static_cast<__nv_bfloat16*>(nullptr)); else with_rms(static_cast<__half*>(nullptr));
}
}  // extern "C"
