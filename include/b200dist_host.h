/* C ABI of lib/libb200dist_host.so - the CPU-side runtime of the framework (no CUDA dependency).
 * Sources: csrc/host/*.cpp.  Python reaches these through ctypes (utils/hostlib.py).
 */
#ifndef B200DIST_HOST_H_
#define B200DIST_HOST_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * DDP reducer core (csrc/host/ddp_reducer.cpp).  replaces: c10d Reducer bucket assignment / ready tracking
 * (torch/csrc/distributed/c10d/reducer.cpp).
 * assign_buckets: parameters arrive in REVERSE registration order; `key` separates dtypes/devices; the first
 * bucket closes at first_cap bytes, later ones at cap bytes.  Returns the number of buckets.
 */
int b2h_ddp_assign_buckets(const int64_t* nbytes, const int32_t* key, int32_t n, int64_t first_cap, int64_t cap,
                           int32_t* bucket_of);
void* b2h_reducer_create(const int32_t* bucket_of, int32_t nparams, int32_t nbuckets);
void b2h_reducer_destroy(void* h);
void b2h_reducer_reset(void* h);
/* marks one parameter's gradient ready; launch[] receives the buckets that became launchable (in bucket order);
 * returns their count, -1 if the parameter was already marked in this iteration */
int b2h_reducer_mark_ready(void* h, int32_t param, int32_t* launch);
/* end of backward: buckets not yet launched (unused parameters) and the list of parameters that never fired */
int b2h_reducer_finalize(void* h, int32_t* launch, int32_t* unused, int32_t* nunused);
/* order in which gradients became ready in the last iteration (bucket rebuild after iteration 1) */
int b2h_reducer_ready_order(void* h, int32_t* out);

/* ------------------------------------------------------------------------------------------
 * Pipeline schedule generator (csrc/host/pipeline_schedule.cpp).  replaces: the control flow of
 * megatron.core.pipeline_parallel.schedules (1F1B and interleaved 1F1B).
 * ops: int32 [cap, 3] rows (op 0 = forward / 1 = backward, model chunk, micro-batch).  Returns the row count (rows beyond
 * `cap` are counted but not written), -1 on an invalid request (interleaving needs M % P == 0).
 */
int b2h_pp_schedule(int32_t P, int32_t rank, int32_t V, int32_t M, int32_t forward_only, int32_t* ops, int32_t cap);

/* ------------------------------------------------------------------------------------------
 * Dataset index builders (csrc/host/data_helpers.cpp).  replaces: Megatron's pybind11 helpers.cpp
 * (build_sample_idx, build_blending_indices, build_mapping, build_blocks_mapping).
 */
int64_t b2h_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx, int64_t n_doc_idx, int32_t seq_length,
                             int32_t num_epochs, int64_t tokens_per_epoch, int32_t* out, int64_t out_rows);
void b2h_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index, const double* weights,
                                int32_t num_datasets, int64_t size);
/* out == NULL: count only.  Rows: (first sentence, one-past-last sentence, target length). */
int64_t b2h_build_mapping(const int64_t* docs, int64_t n_docs, const int32_t* sizes, int32_t num_epochs, int64_t max_num_samples,
                          int32_t max_seq_length, double short_seq_prob, int32_t seed, int32_t min_num_sent, int64_t* out,
                          int64_t out_rows, int32_t shuffle);
/* Rows: (first sentence, one-past-last sentence, document, block id within the epoch). */
int64_t b2h_build_blocks_mapping(const int64_t* docs, int64_t n_docs, const int32_t* sizes, const int32_t* titles_sizes,
                                 int32_t num_epochs, int64_t max_num_samples, int32_t max_seq_length, int32_t seed,
                                 int32_t use_one_sent_blocks, int64_t* out, int64_t out_rows, int32_t shuffle);

/* ------------------------------------------------------------------------------------------
 * CPU Adam / AdamW for offloaded optimizer state (csrc/host/cpu_adam.cpp).  replaces: DeepSpeed csrc/adam/cpu_adam.cpp +
 * includes/simd.h (DeepSpeedCPUAdam).  p, m, v, g: fp32 host arrays of n elements; lp (may be NULL):
 * low-precision copy of the updated parameters, lp_dtype 1 = bf16 (round to nearest even), 2 = fp16.
 * threads <= 0: hardware concurrency (at most 32).  Returns 0, -1 on bad arguments.
 */
int b2h_cpu_adam(float* p, float* m, float* v, const float* g, void* lp, int lp_dtype, int64_t n, float lr, float beta1,
                 float beta2, float eps, float wd, int step, int adamw, float grad_scale, int bias_correction, int threads,
                 int force_generic);
int b2h_cpu_adam_simd_width(void);           /* 16: AVX-512 path in use, 8: AVX2 */
double b2h_cpu_sumsq(const float* g, int64_t n);

#ifdef __cplusplus
}
#endif
#endif /* B200DIST_HOST_H_ */
