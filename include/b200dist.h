/* b200dist native library - C ABI.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (a cudaStream_t
 * passed as void*), launches asynchronously, and returns 0 on success or -1 on error (call
 * b2_last_error() for the message).  Nothing here allocates device memory unless it says so.
 * Python reaches these through ctypes (ops/_native.py); C/C++ callers can link libb200dist.so.
 *
 * The "replaces" notes cite the reference components of SURVEY.md section 2 that each kernel
 * stands in for (the reference tree itself contains no native code; these are the external
 * libraries its four entry scripts drive).
 */
#ifndef B200DIST_H_
#define B200DIST_H_

#ifdef __cplusplus
extern "C" {
#endif

#define B2_DTYPE_FP16 0
#define B2_DTYPE_BF16 1
#define B2_DTYPE_FP32 2

/* GEMM epilogues */
#define B2_EPI_NONE 0
#define B2_EPI_BIAS 1            /* C = acc + bias[n]                                   */
#define B2_EPI_BIAS_GELU 2       /* C = gelu_tanh(acc + bias[n])                        */
#define B2_EPI_BIAS_GELU_DUAL 3  /* C2 = acc + bias[n];  C = gelu_tanh(C2)              */
#define B2_EPI_BIAS_RESIDUAL 4   /* C = aux[m,n] + acc + bias[n]                        */
#define B2_EPI_DGELU 5           /* C = acc * gelu_tanh'(aux[m,n])                      */
#define B2_EPI_ERF 0x100         /* OR into a GELU epilogue: exact erf GELU instead of the tanh approximation */

const char* b2_last_error(void);
int b2_num_sms(void);
/* Dropout under CUDA graphs: a captured kernel node carries the seed the host drew at capture time.  Install a device
 * word with b2_set_dropout_replay_counter (NULL = off) and capture b2_counter_add(word, 1) at the head of the graph: every
 * dropout kernel (bias-dropout-add, dropout backward, attention dropout) scrambles its seed with the word, so each replay
 * draws new masks and forward / backward of one replay agree.  (The reference draws from the CUDA RNG tracker instead:
 * megatron/arguments.py:687-690 hidden_dropout / attention_dropout 0.1.) */
void b2_set_dropout_replay_counter(const void* dev_word);
int b2_counter_add(void* dev_word, unsigned long long inc, void* stream);

/* ------------------------------------------------------------------------------------------
 * GEMM (tcgen05 + TMEM + TMA).  C[M,N] (+)= alpha * op(A)[M,K] . op(B)[K,N]
 *   a_mn_major = 0: A stored [M,K] row-major (lda elements between rows); 1: A stored [K,M].
 *   b_mn_major = 0: B stored [N,K] row-major (an nn.Linear weight);       1: B stored [K,N].
 *   out_fp32/accumulate: fp32 C, optionally C += (TMA reduce-add; split-K allowed, 0 = auto).
 *   max_ctas: 0 = all SMs; smaller values leave SMs free for a concurrent communication kernel.
 * replaces: cuBLAS under torch.matmul/F.linear; apex
 *           fused_weight_gradient_mlp_cuda.wgrad_gemm_accum_fp32 (K18); NVFuser bias-GeLU /
 *           bias-dropout-add epilogues (K16, K17).
 */
/* Persistent kernels size their grids to (SM count - n): leave SMs free for library kernels (NCCL p2p of pipeline
 * parallelism) that a peer GPU may be waiting for while one of the spin-waiting fused kernels runs here. */
void b2_set_reserved_sms(int n);

int b2_gemm(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
            long long ldc, int a_mn_major, int b_mn_major, int dtype, int epilogue, const void* bias,
            const void* aux, long long ld_aux, void* C2, long long ldc2, int out_fp32, int accumulate,
            int split_k, float alpha, int max_ctas, void* stream);
/* CTA-pair (tcgen05 cta_group::2, 256-row tiles) selection for b2_gemm: 0 never, 1 heuristic (default), 2 whenever the shape allows */
void b2_gemm_set_pair_mode(int mode);
/* the tiling b2_gemm would choose: out[7] = {bn, pair, tiles_m, tiles_n, num_kb, split_k, kb_per_split} for an M x N x K problem on
 * `sms` SMs (pair_mode as above; split_k <= 0 = automatic, only fp32-accumulate outputs split).  Pure host arithmetic, no device. */
int b2_gemm_plan(int M, int N, int K, int out_fp32, int accumulate, int split_k, int sms, int pair_mode, int* out);
int b2_gemm_resident_pairs(void);   /* CTA pairs of the GEMM kernel that can be co-resident (0 before the first pair launch) */

/* ------------------------------------------------------------------------------------------
 * Swin shifted-window attention (7x7 windows = 49 tokens, head_dim 32), fused forward / backward.
 * qkv [Bw, 49, 3, heads, 32] (Bw = batch * nW windows) -> out [Bw, 49, heads*32]; bias fp32 [heads, 49, 49]
 * (relative-position bias gathered from its table), mask fp32 [nW, 49, 49] or NULL (shifted windows; window w uses
 * mask[w % nW]); lse fp32 [Bw, heads, 49].  Backward writes dqkv (layout of qkv) and ACCUMULATES d(bias) into dbias.
 * replaces: torchvision ShiftedWindowAttention's matmul / +bias / +mask / softmax / matmul chain.
 */
/* 1 (default): tensor-core window attention (mma.sync, csrc/window_attn_tc.cu); 0: CUDA-core kernels.  Returns the previous value. */
int b2_window_attn_set_impl(int impl);
/* tuning knob of the tensor-core kernels: grids sized for this many CTAs per SM (0 = occupancy query, the default) */
void b2_window_attn_set_grid(int fwd_ctas_per_sm, int bwd_ctas_per_sm);
int b2_window_attn_fwd(const void* qkv, const float* bias, const float* mask, void* out, float* lse, long long Bw, int nW,
                       int heads, int tokens, int head_dim, float scale, int dtype, void* stream);
int b2_window_attn_bwd(const void* qkv, const float* bias, const float* mask, const void* out, const void* dout,
                       const float* lse, void* dqkv, float* dbias, long long Bw, int nW, int heads, int tokens, int head_dim,
                       float scale, int dtype, void* stream);
/* cyclic shift + window partition in one pass (reverse != 0: the inverse, un-partition + shift back); 16-bit elements,
 * in [B, H, W, C] <-> out [B * (H/ws) * (W/ws), ws*ws, C].  replaces torch.roll + permute copies. */
int b2_window_shuffle(const void* in, void* out, int B, int H, int W, int C, int ws, int shift_h, int shift_w, int reverse,
                      void* stream);
/* out[r,:] = res[r,:] + scale[r / (H*W)] * x[src(r),:] over B*H*W rows of C 16-bit elements (C % 8 == 0); res / scale may be NULL
 * (0 / 1).  map 0: src(r) = r; 1: window partition of x; 2: window reverse of x (the maps of b2_window_shuffle).  The residual add
 * with stochastic depth, with Swin's window_reverse folded in; the backward is the same kernel with the inverse map. */
int b2_rows_scale_add(const void* x, const void* res, const float* scale, void* out, int B, int H, int W, int C, int ws,
                      int shift_h, int shift_w, int map, int dtype, void* stream);
/* x [B, C, H, W] -> out [B * (H/4) * (W/4), C * 16]: the non-overlapping 4 x 4 patches as GEMM rows (patch-embedding convolution with
 * kernel = stride = 4 as one GEMM); 16-bit elements, H and W multiples of 4.  replaces view / permute / contiguous. */
int b2_patchify4(const void* x, void* out, int B, int C, int H, int W, void* stream);

/* ------------------------------------------------------------------------------------------
 * LayerNorm / RMSNorm (rms != 0).  x,y,gamma,beta: [rows,H] / [H] in `dtype`; mean,rstd fp32 [rows].
 * Backward needs an fp32 workspace of b2_norm_bwd_num_parts(rows,H) * 2 * H floats (every element is
 * written before it is read); dgamma/dbeta are fp32 and are accumulated when accumulate != 0.
 * replaces: apex fused_layer_norm_cuda, ATen layer_norm (K6), HF LlamaRMSNorm (K24).
 */
int b2_norm_bwd_num_parts(int rows, int H);
int b2_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                     int rows, int H, float eps, int dtype, int rms, void* stream);
int b2_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                     void* dx, float* dgamma, float* dbeta, float* ws, int rows, int H, int dtype, int rms,
                     int accumulate, void* stream);
/* same, with the gradient of the residual branch folded in: dx = dres + d(norm)/dx (dres may be NULL or alias dx) */
int b2_layernorm_bwd_res(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                         const void* dres, void* dx, float* dgamma, float* dbeta, float* ws, int rows, int H, int dtype,
                         int rms, int accumulate, void* stream);
/* the same pair with a row permutation folded in: the forward writes the normalised row r to y[row_map[r]], the backward reads
 * the gradient of row r from dy[row_map[r]] (row_map: int32 [rows], a permutation, or NULL).  A row shuffle that follows the
 * norm - Swin's cyclic shift + window partition - then costs no pass of its own in either direction. */
int b2_layernorm_fwd_map(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                         int rows, int H, float eps, const int* row_map, int dtype, int rms, void* stream);
int b2_layernorm_bwd_map(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                         const void* dres, void* dx, float* dgamma, float* dbeta, float* ws, int rows, int H,
                         const int* row_map, int dtype, int rms, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused elementwise ops.  replaces: Megatron jit bias_gelu / bias_dropout_add (K16, K17), HF SwiGLU
 * and RoPE (K24), autograd's bias-gradient reductions.
 * act modes: 0 tanh-GeLU, 1 erf-GeLU, 2 ReLU, 3 SiLU.
 */
int b2_bias_dropout_add(const void* x, const void* bias, const void* residual, void* out, long long n, int H,
                        float p, unsigned long long seed, int dtype, void* stream);
int b2_dropout_bwd(const void* dy, void* dx, long long n, float p, unsigned long long seed, int dtype, void* stream);
int b2_bias_act_fwd(const void* x, const void* bias, void* out, long long n, int H, int mode, int dtype, void* stream);
int b2_bias_act_bwd(const void* dy, const void* x, const void* bias, void* dx, long long n, int H, int mode, int dtype, void* stream);
int b2_swiglu_fwd(const void* gate_up, void* out, long long rows, int F, int dtype, void* stream);
int b2_swiglu_bwd(const void* dy, const void* gate_up, void* d_gate_up, long long rows, int F, int dtype, void* stream);
int b2_colsum(const void* dy, float* out_accum, int rows, int N, long long ld, int dtype, void* stream);
int b2_rope(void* x, const float* cos_t, const float* sin_t, const int* pos, long long tokens, int nheads, int hd,
            long long tok_stride, int head_stride, int backward, int dtype, void* stream);
int b2_axpby(const void* a, const void* b, void* out, long long n, float alpha, float beta, int dtype, void* stream);
int b2_accum_fp32(float* dst, const void* src, long long n, int dtype, void* stream);
int b2_cast_from_fp32(void* dst, const float* src, long long n, int dtype, void* stream);
/* dst[r, 0:cols] (dtype, row stride ld_dst elements) = src[r, 0:cols] (fp32, contiguous); cols, ld_dst % 8 == 0 */
int b2_cast_rows_from_fp32(void* dst, long long ld_dst, const float* src, long long rows, int cols, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Softmax cross-entropy on (vocab-sharded) logits - see csrc/xent.cu for the 4-step protocol.
 * replaces: Megatron _VocabParallelCrossEntropy (E12/K20/C13), ATen CrossEntropyLoss (K9).
 */
int b2_xent_stats(const void* logits, const long long* labels, float* rmax, float* rsum, float* rtgt,
                  long long rows, int V, long long ld, long long vocab_start, int dtype, void* stream);
int b2_xent_finish(const float* rmax, const float* rsum, const float* rtgt, const long long* labels, float* loss,
                   long long rows, long long ignore_index, void* stream);
/* Vocab-parallel forward in ONE kernel: row statistics of this rank's shard -> exchange through the symmetric stats buffers
 * (stats_ptrs[r]: 2 * 3 * rows floats + control block on every rank) -> global max / sum-exp / loss on every rank.
 * replaces the three NCCL all-reduces of Megatron's _VocabParallelCrossEntropy. */
int b2_xent_vocab_parallel(const void* logits, const long long* labels, void* const* stats_ptrs, void* const* ctrls, int rank,
                           int world, float* gmax, float* gsum, float* loss, long long rows, int V, long long ld,
                           long long vocab_start, long long ignore_index, int dtype, void* stream);
int b2_xent_bwd(const void* logits, void* dlogits, const long long* labels, const float* gmax, const float* gsum,
                const float* row_scale, float scalar_scale, long long rows, int V, long long ld,
                long long vocab_start, long long ignore_index, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Optimizers over flat fp32 arenas.  sumsq / found_inf are DEVICE pointers (may be NULL): the clip
 * coefficient min(1, max_norm / (sqrt(*sumsq) + 1e-6)) and the overflow skip are evaluated on the GPU.
 * replaces: apex amp_C multi_tensor_{adam,l2norm,scale} (E25/K21/K22), DeepSpeed FusedAdam (E30/K25),
 *           ATen Adam / Adadelta (K10, K3).
 */
int b2_adam(float* p, float* m, float* v, const void* g, void* p_lp, long long n, float lr, float beta1, float beta2,
            float eps, float wd, int step, int adamw, float grad_scale, const float* sumsq, float max_norm,
            const float* found_inf, int grad_dtype, int lp_dtype, void* stream);
int b2_adadelta(float* p, float* sq, float* acc, const float* g, long long n, float lr, float rho, float eps, float wd,
                void* stream);
int b2_sgd(float* p, float* mom, const float* g, long long n, float lr, float momentum, float wd, float grad_scale,
           int first_step, void* stream);
/* sum of squares: scratch >= b2_sumsq_scratch_floats() floats, counter = one uint32 that is 0 on entry
 * (the kernel leaves it 0 again).  out[0] = (or +=) sum x^2. */
int b2_sumsq_scratch_floats(void);
int b2_sumsq(const void* x, long long n, float* scratch, unsigned int* counter, float* out, int accumulate, int dtype,
             void* stream);
int b2_scale_check(float* x, long long n, float scale, float* found_inf, void* stream);

/* Allocation-free sum of squares: out[0] = (or +=) sum x^2, same result as b2_sumsq, but the caller passes
 * only the input and the output.  (No counterpart in the reference: this entry point exists for users of
 * THIS library.)  It is meant for callers that cannot get device memory when they need a norm: a
 * monitoring thread in a job whose framework allocator already holds almost all of HBM (a new cudaMalloc
 * may fail there, and cudaMalloc/cudaFree stall all streams of the device), plain C/C++ hosts without a
 * workspace pool, services that freeze their memory map after start-up.
 *   - Scratch: a file-scope __device__ array of b2_sumsq_static_scratch_bytes() = 4352 bytes that every GPU
 *     loading the library holds until the process exits.  This call performs no cudaMalloc/cudaFree and
 *     calls no allocating library routine.
 *   - Limit: n <= B2_SUMSQ_STATIC_MAX_ELEMS (2^31 elements; the 1024 partial-sum slots bound the grid and,
 *     with it, how many values one fp32 lane accumulates).  Larger n is rejected with an error; split the
 *     buffer and pass accumulate = 1, or use b2_sumsq with your own scratch.
 *   - The first call on each device (with that device current) looks up the array's device address; make
 *     that first call before starting any stream capture.
 *   - All calls on a device share the one array (it keeps the previous call's contents; the ticket word that
 *     must start at zero is cleared on `stream` by every call).  Two calls whose kernels overlap on the GPU
 *     (different streams, no ordering between them) race on it: order them.
 */
#define B2_SUMSQ_STATIC_MAX_ELEMS (1LL << 31)
long long b2_sumsq_static_max_elems(void);
int b2_sumsq_static_scratch_bytes(void);
int b2_sumsq_static(const void* x, long long n, float* out, int accumulate, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Embedding (vocab-parallel aware).  replaces: Megatron VocabParallelEmbedding (E10/K11), nn.Embedding.
 */
int b2_embedding_fwd(const long long* tok, const long long* pos, const void* wte, const void* wpe, void* out,
                     long long tokens, int H, long long vocab_start, long long vocab_local, int dtype, void* stream);
int b2_embedding_bwd(const void* dy, const long long* tok, const long long* pos, float* dwte, float* dwpe,
                     long long tokens, int H, long long vocab_start, long long vocab_local, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * In-kernel collectives over NVLink / NVSwitch on SYMMETRIC buffers (csrc/comm.cu).  ptrs[r] / ctrls[r] are the
 * peer-mapped addresses of rank r's payload and of its control block (B2_CTRL_WORDS zero-initialised uint32
 * owned by that buffer); mc_ptr is the NVSwitch multicast address of the payload or NULL (peer load/store path).
 * All ranks of the group must launch the same calls on a buffer in the same order.
 * replaces: NCCL all-reduce + separate scale of DDP buckets / Megatron's grad buffer.
 */
#define B2_CTRL_WORDS 64
int b2_ctrl_words(void);
int b2_allreduce(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, float scale,
                 int rank, int world, int max_ctas, void* stream);
/* out (local, n_elems/world) = scale * sum over ranks of slice `rank` of the symmetric payload (payload unchanged) */
int b2_reduce_scatter(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, float scale,
                      int rank, int world, int max_ctas, void* out, void* stream);
/* slice r of rank r's payload is replicated into every rank's payload (multimem.st broadcast / peer stores) */
int b2_allgather(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, int rank, int world,
                 int max_ctas, void* stream);

/* ZeRO gradient path in one kernel: reduce-scatter of the symmetric gradient buffer (multimem.ld_reduce of this
 * rank's slice) * gscale, accumulated into the fp32 gradient shard `gacc` (may be NULL) and - when adam != 0 -
 * consumed directly by AdamW on the fp32 master/m/v shard, writing the bf16 parameter shard `p_lp`.
 * sumsq (device, optional) += sum of squares of the scaled reduced gradient (for global-norm clipping).
 * replaces: DeepSpeed reduce_scatter_coalesced + FusedAdam. */
int b2_rs_update(void* const* ptrs, void* const* ctrls, void* mc_ptr, long long n_elems, int dtype, int rank, int world,
                 float gscale, float* gacc, int accumulate, float* sumsq, int adam, float* p32, float* m, float* v,
                 void* p_lp, float lr, float beta1, float beta2, float eps, float wd, int step, int max_ctas, void* stream);

/* Fused compute + collective kernels (csrc/gemm.cu).  A operand is [rows, K] row-major; B as in b2_gemm.
 *  b2_ag_gemm : all-gather(A shards over the group) -> GEMM, one kernel; xg receives the gathered A.  `flags` = world *
 *               shard_rows / 128 uint32 arrival counters that belong to the shard buffer (zero before its first call, owned
 *               by the kernels afterwards).  n_comm <= 0: the default number of all-gather CTAs.
 *  b2_gemm_rs : GEMM -> reduce-scatter(+bias+residual), one kernel.  stage_ptrs[r] = rank r's symmetric staging buffer of
 *               2 * world * shard_rows * ceil(N/64) * 64 elements, cnt_ptrs[r] = rank r's shard_rows / 128 arrival counters.
 *  b2_fused_config : pair_mode 0/1 (CTA-pair 256-row tiles in the fused kernels, default 1; < 0 keeps), default n_comm (<= 0 keeps).
 * Every cross-GPU wait in these kernels is bounded (B2_SPIN_TIMEOUT_NS, 4 s): a peer that never arrives turns into a
 * launch failure (trap), not a hung GPU.
 * replaces: NCCL all-gather / reduce-scatter around cuBLAS GEMMs in Megatron's Column/RowParallelLinear with
 *           sequence parallelism and DeepSpeed ZeRO-3 parameter gathers (C19).
 */
int b2_ag_gemm(void* const* shard_ptrs, void* const* ctrls, void* xg, unsigned int* flags, const void* B, void* C,
               int shard_rows, int N, int K, long long ldb, long long ldc, int b_mn_major, int dtype, int epilogue,
               const void* bias, const void* aux, long long ld_aux, void* C2, long long ldc2, int rank, int world,
               int n_comm, void* stream);
int b2_gemm_rs(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, int b_mn_major, int dtype,
               void* const* cnt_ptrs, void* const* ctrls, void* const* stage_ptrs, void* out,
               long long ld_out, const void* bias, const void* residual, long long ld_res, int rank, int world,
               void* stream);
/* All-gather of a row-sharded WEIGHT fused into its consuming GEMM (ZeRO-3 parameter fetch): C[M, world*shard_rows] =
 * A[M, K] @ Wfull^T.  wg = this rank's full gathered weight (symmetric memory, own shard already in place at row
 * rank * shard_rows); shard_ptrs[r] = rank r's own shard inside rank r's copy of wg; flags = world * ceil(shard_rows/128)
 * uint32, zero before the first call.  After the launch wg is complete.  replaces DeepSpeed stage3 all_gather_coalesced +
 * the cuBLAS GEMM that follows it. */
int b2_agw_gemm(void* const* shard_ptrs, void* const* ctrls, void* wg, unsigned int* flags, const void* A, void* C, int M,
                int shard_rows, int K, long long lda, long long ldc, int dtype, int rank, int world, int n_comm, void* stream);
/* GEMM -> all-reduce (tensor parallelism without sequence parallelism): the b2_gemm_rs kernel whose reducing rank stores
 * each finished row block into EVERY rank's full [M, N] output and which ends with a cross-GPU barrier.  out buffers are
 * symmetric; armaps_dev = `world` tensor maps uploaded once by b2_make_ar_maps (a synchronous cudaMemcpy: set-up time only).
 * residual: full [M, N] or NULL.  replaces: cuBLAS GEMM + NCCL all_reduce of Row/ColumnParallelLinear with
 * sequence_parallel False (3_training_megatron-lm.ipynb:1162). */
int b2_make_ar_maps(void* dev_out, void* const* out_ptrs, int world, int M, int N, long long ld_out, int dtype);
int b2_gemm_ar(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, int b_mn_major, int dtype,
               void* const* cnt_ptrs, void* const* ctrls, void* const* stage_ptrs, const void* armaps_dev, void* out_local,
               long long ld_out, const void* bias, const void* residual, long long ld_res, int rank, int world, void* stream);
void b2_fused_config(int pair_mode, int n_comm);
int b2_set_spin_timeout_ms(long long ms);   /* bound of the cross-GPU waits on the current device (default 4000 ms) */

/* ------------------------------------------------------------------------------------------
 * BatchNorm2d over NCHW tensors (csrc/batchnorm.cu), fp32 / bf16 / fp16 activations, fp32 statistics.  Training forward =
 * b2_bn_reduce (sums[C][2] double, zero on entry) -> [all-reduce sums and count over the group for SyncBatchNorm] ->
 * b2_bn_finalize (mean, rstd, scale, shift, running statistics) -> b2_bn_apply (y = x*scale + shift, optional ReLU).
 * Inference = b2_bn_eval_coeffs -> b2_bn_apply.  Backward = b2_bn_bwd_reduce (sum g, sum g*xhat) -> [all-reduce] ->
 * b2_bn_bwd_apply; dbeta[c] = sums[c][0], dgamma[c] = sums[c][1].
 * replaces: cuDNN / ATen batch_norm (torchvision ResNet under pytorch_oxford_ddp.py), apex syncbn (E25).
 */
int b2_bn_reduce(const void* x, double* sums, int N, int C, int HW, int dtype, void* stream);
int b2_bn_finalize(const double* sums, double count, const float* gamma, const float* beta, float eps, float momentum,
                   float* running_mean, float* running_var, float* mean, float* rstd, float* scale, float* shift, int C, void* stream);
int b2_bn_eval_coeffs(const float* running_mean, const float* running_var, const float* gamma, const float* beta, float eps,
                      float* scale, float* shift, int C, void* stream);
int b2_bn_apply(const void* x, const float* scale, const float* shift, void* y, int N, int C, int HW, int relu, int dtype, void* stream);
int b2_bn_bwd_reduce(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, double* sums, int N, int C,
                     int HW, int relu, int dtype, void* stream);
int b2_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* rstd, const float* gamma,
                    const double* sums, double count, void* dx, int N, int C, int HW, int relu, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * fp32 kernels of the MNIST CNN (csrc/conv.cu): valid stride-1 convolution with K x K filters (K <= 5), NCHW.
 * fwd: y = relu?(conv(x, w) + bias).  bwd: g = dy * (yact > 0) when relu (yact = the forward output); dx (may be NULL),
 * dw / db (may be NULL; accumulate != 0 adds into them).  One input / gradient image must fit in shared memory
 * (<= 200 KB: e.g. 32 x 26 x 26 fp32).  maxpool2: 2 x 2, stride 2, idx = winning tap (uint8) for the backward.
 * log_softmax_nll: C <= 32 classes; loss[row] = -logp[row, target[row]]; bwd adds the NLL gradient
 * (gloss[row] (or 1) * gscalar) to dlogp (may be NULL) and applies the log-softmax Jacobian.
 * replaces: cuDNN convolution fwd/bwd, ATen max_pool2d, log_softmax, nll_loss (model_def.py:22-45; K1, K2).
 */
int b2_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout, int K,
                  int relu, void* stream);
int b2_conv2d_bwd(const float* dy, const float* yact, const float* x, const float* w, float* dx, float* dw, float* db, int B,
                  int Cin, int H, int W, int Cout, int K, int relu, int accumulate, void* stream);
int b2_maxpool2_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, void* stream);
int b2_maxpool2_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, void* stream);
int b2_log_softmax_nll_fwd(const float* x, const long long* target, float* logp, float* loss, long long rows, int C, void* stream);
int b2_log_softmax_nll_bwd(const float* logp, const float* dlogp, const long long* target, const float* gloss, float gscalar,
                           float* dx, long long rows, int C, void* stream);

/* ------------------------------------------------------------------------------------------
 * Stand-alone scale + mask + softmax over attention scores (the unfused attention path).  x, y: [rows, sk] with
 * rows = batches * heads * sq; mask: uint8 [mask_batches, 1, sq, sk] (non-zero = masked, filled with -10000 after
 * scaling; mask_batches == batches or 1) or NULL; causal != 0 hides keys j > i + (sk - sq).  sk <= 16384.
 * Backward: dx = scale * y * (dy - sum_j dy_j y_j).  dtype: B2_DTYPE_{FP16,BF16,FP32}.
 * replaces: Megatron scaled_upper_triang_masked_softmax_cuda / scaled_masked_softmax_cuda / scaled_softmax_cuda (E19).
 */
int b2_softmax_fwd(const void* x, const unsigned char* mask, void* y, long long rows, int sq, int sk, int heads,
                   int mask_batches, float scale, int causal, int dtype, void* stream);
int b2_softmax_bwd(const void* dy, const void* y, void* dx, long long rows, int sk, float scale, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * Flash attention (tcgen05), bf16, head_dim 64 or 128.  Strides are in
 * elements for (seq, batch, head); head_dim is contiguous.  lse: fp32 [batch, heads, seqlen].
 * Backward: delta is fp32 scratch [batch, heads, seqlen]; dq_accum is fp32 [seq, batch, heads, D],
 * contiguous and ZERO on entry (dQ tiles are TMA reduce-added into it).
 * p_drop > 0 applies attention dropout to the softmax output: element (b, h, q, k) is kept iff a counter-based
 * hash of (seed, b*heads+h, q, k) >= p_drop * 2^32 (pass the same p_drop / seed to the backward).
 * replaces: Megatron CoreAttention baddbmm/bmm + scaled_upper_triang_masked_softmax_cuda (E17/E19,
 *           K14/K15), HF eager attention (K24).
 */
int b2_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int seqlen, int batch, int heads,
                int D, long long q_ss, long long q_sb, long long q_sh, long long k_ss, long long k_sb,
                long long k_sh, long long v_ss, long long v_sb, long long v_sh, long long o_ss, long long o_sb,
                long long o_sh, float scale, int causal, float p_drop, unsigned long long seed, void* stream);
int b2_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                float* delta, float* dq_accum, void* dk, void* dv, int seqlen, int batch, int heads, int D,
                long long q_ss, long long q_sb, long long q_sh, long long k_ss, long long k_sb, long long k_sh,
                long long v_ss, long long v_sb, long long v_sh, long long o_ss, long long o_sb, long long o_sh,
                long long do_ss, long long do_sb, long long do_sh, long long dk_ss, long long dk_sb,
                long long dk_sh, long long dv_ss, long long dv_sb, long long dv_sh, float scale, int causal, float p_drop,
                unsigned long long seed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DIST_H_ */
