// ThreadSanitizer pass over the emulated kernels: every __shared__ access that is not ordered by a barrier shows up as a data
// race.  g++ -fsanitize=thread -O1 -g -pthread -mf16c -mavx2 -Itests/emu -Iinclude -Icsrc tests/emu/tsan_main.cpp tests/emu/emu_kernels.cpp
#include <cstdio>
#include <cstdlib>
#include <vector>
extern "C" {
void emu_softmax_fwd(const void*, const unsigned char*, void*, long long, int, int, int, int, float, int, int, int);
void emu_softmax_bwd(const void*, const void*, void*, long long, int, float, int, int);
void emu_conv2d_fwd(const float*, const float*, const float*, float*, int, int, int, int, int, int, int);
void emu_conv2d_bwd(const float*, const float*, const float*, const float*, float*, float*, float*, int, int, int, int, int, int, int, int);
void emu_bn_train_fwd(const void*, const float*, const float*, float, float, float*, float*, float*, float*, void*, double*, int, int, int, int, int, int, int);
void emu_bn_bwd(const void*, const void*, const void*, const float*, const float*, const float*, double*, void*, int, int, int, int, int, int, int);
void emu_log_softmax_nll_fwd(const float*, const long long*, float*, float*, long long, int);
void emu_wattn_tc_fwd(const void*, const float*, const float*, void*, float*, long long, int, int, float, int, int);
void emu_wattn_tc_bwd(const void*, const float*, const float*, const void*, const void*, const float*, void*, float*, long long, int, int, float, int, int);
void emu_norm_fwd(const void*, const void*, const void*, void*, float*, float*, int, int, float, const int*, int, int, int);
void emu_norm_bwd(const void*, const void*, const void*, const float*, const float*, const void*, void*, float*, float*, int, int, const int*, int, int, int, int);
}
static std::vector<unsigned short> rnd_bf16(size_t n) {          // bf16 bit patterns of values in [-2, 2]
  std::vector<unsigned short> v(n);
  for (auto& x : v) { float f = (rand() % 2001 - 1000) / 500.f; unsigned u; __builtin_memcpy(&u, &f, 4); x = static_cast<unsigned short>(u >> 16); }
  return v;
}
static std::vector<float> rnd(size_t n) { std::vector<float> v(n); for (auto& x : v) x = (rand() % 2001 - 1000) / 500.f; return v; }
int main() {
  {  // softmax: 5 rows on 2 CTAs -> the per-CTA row loop reuses row[] and red[]
    auto x = rnd(2 * 5 * 70); std::vector<float> y(x.size()), dx(x.size());
    emu_softmax_fwd(x.data(), nullptr, y.data(), 10, 5, 70, 2, 1, 0.5f, 1, 2, 2);
    emu_softmax_bwd(x.data(), y.data(), dx.data(), 10, 70, 0.5f, 2, 2);
  }
  {  // conv 3x3, 2 samples, 3 -> 9 channels
    int B = 2, Cin = 3, H = 7, W = 6, Cout = 9, K = 3, Ho = H - K + 1, Wo = W - K + 1;
    auto x = rnd(B * Cin * H * W), w = rnd(Cout * Cin * K * K), b = rnd(Cout), dy = rnd(B * Cout * Ho * Wo);
    std::vector<float> y(B * Cout * Ho * Wo), dx(x.size()), dw(w.size()), db(Cout);
    emu_conv2d_fwd(x.data(), w.data(), b.data(), y.data(), B, Cin, H, W, Cout, K, 1);
    emu_conv2d_bwd(dy.data(), y.data(), x.data(), w.data(), dx.data(), dw.data(), db.data(), B, Cin, H, W, Cout, K, 1, 0);
  }
  {  // batch norm
    int N = 3, C = 4, HW = 30;
    auto x = rnd(N * C * HW), g = rnd(C), be = rnd(C), dy = rnd(N * C * HW);
    std::vector<float> rm(C), rv(C, 1.f), mean(C), rstd(C), y(x.size()), dx(x.size());
    std::vector<double> sums(2 * C, 0.0);
    emu_bn_train_fwd(x.data(), g.data(), be.data(), 1e-5f, 0.1f, rm.data(), rv.data(), mean.data(), rstd.data(), y.data(), sums.data(), N, C, HW, 1, 2, 2, 0);
    std::fill(sums.begin(), sums.end(), 0.0);
    emu_bn_bwd(dy.data(), x.data(), y.data(), mean.data(), rstd.data(), g.data(), sums.data(), dx.data(), N, C, HW, 1, 2, 2, 0);
  }
  {  // log-softmax
    auto z = rnd(19 * 10); std::vector<long long> t(19, 3); std::vector<float> lp(z.size()), loss(19);
    emu_log_softmax_nll_fwd(z.data(), t.data(), lp.data(), loss.data(), 19, 10);
  }
  {  // tensor-core window attention: 3 windows on 1 CTA per head -> both halves of the cp.async double buffer are reused
    const int Bw = 3, heads = 2, C = heads * 32;
    auto qkv = rnd_bf16(static_cast<size_t>(Bw) * 49 * 3 * C), dout = rnd_bf16(static_cast<size_t>(Bw) * 49 * C);
    auto bias = rnd(heads * 49 * 49), mask = rnd(49 * 49);
    std::vector<unsigned short> out(static_cast<size_t>(Bw) * 49 * C), dqkv(qkv.size());
    std::vector<float> lse(Bw * heads * 49), dbias(heads * 49 * 49, 0.f);
    emu_wattn_tc_fwd(qkv.data(), bias.data(), mask.data(), out.data(), lse.data(), Bw, 1, heads, 0.17677669f, 1, 1);
    emu_wattn_tc_bwd(qkv.data(), bias.data(), mask.data(), out.data(), dout.data(), lse.data(), dqkv.data(), dbias.data(), Bw, 1, heads, 0.17677669f, 1, 1);
  }
  {  // LayerNorm: 8 lanes per row (16 rows per CTA, the last warp's teams partly without a row), row loop, workspace fold
    const int rows = 37, H = 128;
    auto x = rnd_bf16(static_cast<size_t>(rows) * H), g = rnd_bf16(H), b = rnd_bf16(H), dy = rnd_bf16(static_cast<size_t>(rows) * H);
    std::vector<unsigned short> y(x.size()), dx(x.size());
    std::vector<float> mean(rows), rstd(rows), dg(H, 0.f), db(H, 0.f);
    emu_norm_fwd(x.data(), g.data(), b.data(), y.data(), mean.data(), rstd.data(), rows, H, 1e-5f, nullptr, 1, 0, 2);
    emu_norm_bwd(dy.data(), x.data(), g.data(), mean.data(), rstd.data(), nullptr, dx.data(), dg.data(), db.data(), rows, H, nullptr, 1, 0, 0, 2);
  }
  std::puts("tsan pass done");
  return 0;
}
