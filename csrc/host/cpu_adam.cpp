// CPU Adam / AdamW over flat fp32 shards in (pinned) host memory, with the low-precision parameter copy written in the
// same pass.  C ABI, called through ctypes.
//
// Native equivalent of DeepSpeed's `csrc/adam/cpu_adam.cpp` + `includes/simd.h` (DeepSpeedCPUAdam), the optimizer of
// ZeRO-Offload that the reference's `configs/default_offload_opt_param-original.json:24-27`
// (`offload_optimizer.device = cpu`) selects.
//
// Layout of the work: the shard is cut into tiles of 16 Ki elements; worker threads take tiles from an atomic counter,
// so the pass is bandwidth-balanced whatever the shard size.  Per element one pass reads p, m, v, g and writes p, m, v
// and the bf16 / fp16 copy that the H2D copy of the next all-gather ships to the GPU.  AVX-512 path (runtime-detected),
// otherwise a loop the compiler vectorises for x86-64-v3 (AVX2 + FMA).
#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct AdamArgs {
  float* p; float* m; float* v; const float* g; void* lp; int lp_dtype;   // lp_dtype: 0 none, 1 bf16, 2 fp16
  float lr, beta1, beta2, eps, wd, grad_scale, bc1, bc2_rsqrt; int adamw;
};

inline uint16_t f32_to_bf16(float f) {
  uint32_t u; std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);   // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                                        // round to nearest even
  return static_cast<uint16_t>(u >> 16);
}

inline uint16_t f32_to_f16(float f) {
  return static_cast<uint16_t>(_cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
}

// p <- p - lr * ( m_hat / (sqrt(v_hat) + eps) + wd * p )   (AdamW)   |   g <- g + wd * p first (Adam with L2)
void adam_range_generic(const AdamArgs& a, int64_t lo, int64_t hi) {
  const float step_size = a.lr / a.bc1;
  for (int64_t i = lo; i < hi; ++i) {
    float g = a.g[i] * a.grad_scale;
    float p = a.p[i];
    if (!a.adamw && a.wd != 0.f) g += a.wd * p;
    const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * g;
    const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * g * g;
    const float denom = std::sqrt(v) * a.bc2_rsqrt + a.eps;
    if (a.adamw && a.wd != 0.f) p -= a.lr * a.wd * p;
    p -= step_size * (m / denom);
    a.p[i] = p; a.m[i] = m; a.v[i] = v;
  }
  if (a.lp_dtype == 1) {
    uint16_t* o = static_cast<uint16_t*>(a.lp);
    for (int64_t i = lo; i < hi; ++i) o[i] = f32_to_bf16(a.p[i]);
  } else if (a.lp_dtype == 2) {
    uint16_t* o = static_cast<uint16_t*>(a.lp);
    for (int64_t i = lo; i < hi; ++i) o[i] = f32_to_f16(a.p[i]);
  }
}

__attribute__((target("avx512f,avx512bw,avx512vl")))
void adam_range_avx512(const AdamArgs& a, int64_t lo, int64_t hi) {
  const __m512 b1 = _mm512_set1_ps(a.beta1), b2 = _mm512_set1_ps(a.beta2);
  const __m512 ob1 = _mm512_set1_ps(1.f - a.beta1), ob2 = _mm512_set1_ps(1.f - a.beta2);
  const __m512 eps = _mm512_set1_ps(a.eps), gs = _mm512_set1_ps(a.grad_scale), wd = _mm512_set1_ps(a.wd);
  const __m512 step_size = _mm512_set1_ps(a.lr / a.bc1), bc2 = _mm512_set1_ps(a.bc2_rsqrt);
  const __m512 decay = _mm512_set1_ps(1.f - a.lr * a.wd);
  const bool l2 = !a.adamw && a.wd != 0.f, dw = a.adamw && a.wd != 0.f;
  int64_t i = lo;
  for (; i + 16 <= hi; i += 16) {
    __m512 g = _mm512_mul_ps(_mm512_loadu_ps(a.g + i), gs);
    __m512 p = _mm512_loadu_ps(a.p + i);
    if (l2) g = _mm512_fmadd_ps(wd, p, g);
    const __m512 m = _mm512_fmadd_ps(b1, _mm512_loadu_ps(a.m + i), _mm512_mul_ps(ob1, g));
    const __m512 v = _mm512_fmadd_ps(b2, _mm512_loadu_ps(a.v + i), _mm512_mul_ps(ob2, _mm512_mul_ps(g, g)));
    const __m512 denom = _mm512_fmadd_ps(_mm512_sqrt_ps(v), bc2, eps);
    if (dw) p = _mm512_mul_ps(p, decay);
    p = _mm512_fnmadd_ps(step_size, _mm512_div_ps(m, denom), p);
    _mm512_storeu_ps(a.p + i, p); _mm512_storeu_ps(a.m + i, m); _mm512_storeu_ps(a.v + i, v);
    if (a.lp_dtype == 1) {
      // round-to-nearest-even bf16: add 0x7fff + lsb, take the high half (NaNs are not expected in parameters)
      const __m512i u = _mm512_castps_si512(p);
      const __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
      const __m512i r = _mm512_srli_epi32(_mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff))), 16);
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(static_cast<uint16_t*>(a.lp) + i), _mm512_cvtepi32_epi16(r));
    } else if (a.lp_dtype == 2) {
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(static_cast<uint16_t*>(a.lp) + i),
                          _mm512_cvtps_ph(p, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
    }
  }
  if (i < hi) adam_range_generic(a, i, hi);
}

bool has_avx512() {
  static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl");
  return ok;
}

}  // namespace

extern "C" {

// 1 when the AVX-512 path is in use on this machine (tests print it; both paths are checked against the oracle)
int b2h_cpu_adam_simd_width() { return has_avx512() ? 16 : 8; }

// One Adam(W) step over n elements.  p, m, v, g: fp32 host arrays.  lp (may be NULL): low-precision copy of the updated
// parameters (lp_dtype 1 = bf16, 2 = fp16).  step >= 1 is the step being taken (bias correction).  threads <= 0: hardware
// concurrency (capped at 32).  force_generic != 0 disables the AVX-512 path.  Returns 0, or -1 on bad arguments.
int b2h_cpu_adam(float* p, float* m, float* v, const float* g, void* lp, int lp_dtype, int64_t n, float lr, float beta1,
                 float beta2, float eps, float wd, int step, int adamw, float grad_scale, int bias_correction, int threads,
                 int force_generic) {
  if (n < 0 || step < 1 || !p || !m || !v || !g || (lp_dtype != 0 && !lp) || lp_dtype < 0 || lp_dtype > 2) return -1;
  AdamArgs a{p, m, v, g, lp, lp ? lp_dtype : 0, lr, beta1, beta2, eps, wd, grad_scale, 1.f, 1.f, adamw};
  if (bias_correction) {
    a.bc1 = 1.f - std::pow(beta1, static_cast<float>(step));
    a.bc2_rsqrt = 1.f / std::sqrt(1.f - std::pow(beta2, static_cast<float>(step)));
  }
  const bool simd = has_avx512() && !force_generic;
  constexpr int64_t TILE = 1 << 14;
  const int64_t tiles = (n + TILE - 1) / TILE;
  int nt = threads > 0 ? threads : static_cast<int>(std::min<unsigned>(32u, std::max(1u, std::thread::hardware_concurrency())));
  nt = static_cast<int>(std::min<int64_t>(nt, std::max<int64_t>(tiles, 1)));
  std::atomic<int64_t> next{0};
  auto work = [&]() {
    for (;;) {
      const int64_t t = next.fetch_add(1, std::memory_order_relaxed);
      if (t >= tiles) return;
      const int64_t lo = t * TILE, hi = std::min(n, lo + TILE);
      if (simd) adam_range_avx512(a, lo, hi); else adam_range_generic(a, lo, hi);
    }
  };
  if (nt <= 1) { work(); return 0; }
  std::vector<std::thread> pool;
  pool.reserve(nt - 1);
  for (int t = 0; t < nt - 1; ++t) pool.emplace_back(work);
  work();
  for (auto& th : pool) th.join();
  return 0;
}

// sum of squares of an fp32 host array (gradient norm of an offloaded shard), double accumulation
double b2h_cpu_sumsq(const float* g, int64_t n) {
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t i = 0;
  for (; i + 8 <= n; i += 8)
    for (int k = 0; k < 8; ++k) acc[k] += static_cast<double>(g[i + k]) * g[i + k];
  double s = 0;
  for (; i < n; ++i) s += static_cast<double>(g[i]) * g[i];
  for (int k = 0; k < 8; ++k) s += acc[k];
  return s;
}

}  // extern "C"
