// Host-side support shared by every kernel file: error string, SM count, TMA descriptor encoding
// through the driver entry point (no link-time dependency on libcuda).
#include "common.cuh"
#include "b200dist.h"
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace b2host {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

static int g_reserved_sms = 0;
// SMs the persistent kernels leave free (b2_set_reserved_sms): kernels that spin on a peer GPU (fused collectives) occupy
// every SM they are given; a library kernel the PEER is waiting for (an NCCL send of pipeline parallelism) must still
// find an SM on this GPU, or the two GPUs wait on each other forever
void set_reserved_sms(int n) { g_reserved_sms = n < 0 ? 0 : n; }
int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  const int n = cached[dev] - g_reserved_sms;
  return n < 8 ? 8 : n;
}

// Device word that every dropout kernel mixes into its seed (NULL = off).  A CUDA graph bakes the seeds that the host drew at
// capture time into its kernel nodes; a captured one-thread kernel bumps this word at the start of every replay, so each
// replay draws fresh masks while forward and backward of the same replay still see the same value.
static const unsigned long long* g_dropout_replay = nullptr;
const unsigned long long* dropout_replay_ptr() { return g_dropout_replay; }

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int encode_tmap(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled driver entry point not available");
    return -1;
  }
  cuuint64_t gdims[5];
  cuuint64_t gstrides[4];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstrides[i - 1] = strides_bytes[i];
  }
  CUresult r = fn(out, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): rank=%d base=%p dims=[%llu,%llu,..] stride1=%llu box=[%u,%u,..]",
              static_cast<int>(r), rank, base, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 1 ? strides_bytes[1] : 0), box[0], rank > 1 ? box[1] : 0);
    return -1;
  }
  return 0;
}

}  // namespace b2host

namespace {
__global__ void counter_add_kernel(unsigned long long* p, unsigned long long inc) { *p += inc; }
}  // namespace
extern "C" void b2_set_dropout_replay_counter(const void* dev_word) { b2host::g_dropout_replay = static_cast<const unsigned long long*>(dev_word); }
extern "C" int b2_counter_add(void* dev_word, unsigned long long inc, void* stream) {
  counter_add_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<unsigned long long*>(dev_word), inc);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { b2host::set_error("counter_add: %s", cudaGetErrorString(e)); return -1; }
  return 0;
}
// bound of every cross-GPU / cross-CTA flag wait of the fused kernels on the CURRENT device (default 4000 ms; a run under
// compute-sanitizer, where a peer can be seconds late, wants more)
extern "C" int b2_set_spin_timeout_ms(long long ms) {
  const unsigned long long ns = static_cast<unsigned long long>(ms < 1 ? 1 : ms) * 1000000ull;
  if (b2host::set_spin_timeout_gemm(ns) || b2host::set_spin_timeout_comm(ns)) { b2host::set_error("set_spin_timeout: cudaMemcpyToSymbol failed"); return -1; }
  return 0;
}
extern "C" const char* b2_last_error(void) { return b2host::last_error(); }
extern "C" int b2_num_sms(void) { return b2host::num_sms(); }
extern "C" void b2_set_reserved_sms(int n) { b2host::set_reserved_sms(n); }
