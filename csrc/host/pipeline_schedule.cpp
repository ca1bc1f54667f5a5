// Pipeline-parallel schedule generator (C ABI): the per-rank COMPUTE order of forward / backward
// micro-batch steps for
//   * no pipelining               (F0 B0 F1 B1 ...)
//   * 1F1B without interleaving   (warm-up = P - rank - 1 forwards, steady 1F1B, cool-down)
//   * interleaved 1F1B            (V model chunks per rank, micro-batches advance in groups of P;
//                                  defined here for every P >= 2 - the reference refuses P == 2)
// It stands in for the control flow of megatron.core.pipeline_parallel.schedules.  The
// executor (parallel/pipeline.py) derives all p2p traffic from this order, so the same list is what
// the tests check for dependency-safety and for the 1F1B in-flight-activation bound.
#include <algorithm>
#include <cstdint>

extern "C" {

// op encoding: ops[3*i] = 0 forward | 1 backward ; ops[3*i+1] = model chunk ; ops[3*i+2] = micro-batch
// returns the number of ops written, or -1 (bad arguments) / -2 (capacity too small).
int b2h_pp_schedule(int32_t P, int32_t rank, int32_t V, int32_t M, int32_t forward_only, int32_t* ops, int32_t cap) {
  if (P < 1 || rank < 0 || rank >= P || V < 1 || M < 1) return -1;
  int n = 0;
  auto push = [&](int t, int c, int m) {
    if (n < cap) { ops[3 * n] = t; ops[3 * n + 1] = c; ops[3 * n + 2] = m; }
    ++n;
  };
  if (P == 1 && V == 1) {
    for (int m = 0; m < M; ++m) { push(0, 0, m); if (!forward_only) push(1, 0, m); }
    return n <= cap ? n : -2;
  }
  if (V == 1) {
    int warm = std::min(P - rank - 1, M);
    int fwd = 0, bwd = 0;
    for (int i = 0; i < warm; ++i) push(0, 0, fwd++);
    int steady = M - warm;
    for (int i = 0; i < steady; ++i) {
      push(0, 0, fwd++);
      if (!forward_only) push(1, 0, bwd++);
    }
    if (!forward_only) while (bwd < M) push(1, 0, bwd++);
    return n <= cap ? n : -2;
  }
  // interleaved
  if (M % P != 0) return -1;
  const int total = M * V;
  auto chunk_of = [&](int k, bool fwd_dir) {
    int in_group = k % (P * V);
    int c = in_group / P;
    return fwd_dir ? c : (V - 1 - c);
  };
  auto mb_of = [&](int k) { return (k / (P * V)) * P + (k % P); };
  int warm;
  if (forward_only) warm = total;
  else if (M == P) warm = total;
  else warm = std::min((P - rank - 1) * 2 + (V - 1) * P, total);
  int f = 0, b = 0;
  for (int i = 0; i < warm; ++i, ++f) push(0, chunk_of(f, true), mb_of(f));
  int steady = total - warm;
  for (int i = 0; i < steady; ++i) {
    push(0, chunk_of(f, true), mb_of(f)); ++f;
    push(1, chunk_of(b, false), mb_of(b)); ++b;
  }
  if (!forward_only) for (; b < total; ++b) push(1, chunk_of(b, false), mb_of(b));
  return n <= cap ? n : -2;
}

}  // extern "C"
