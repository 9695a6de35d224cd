/*
 * detb200 — C ABI of the B200 (sm_100a) native kernel library used by determined_b200.
 *
 * Conventions
 *   - every function returns 0 on success, a positive cudaError_t or a negative DETB200_ERR_* code;
 *     detb200_last_error() returns a human-readable description (thread local);
 *   - all pointers are DEVICE pointers unless a parameter is documented as host memory;
 *   - `stream` is the CUDA stream to launch on; nothing synchronises unless documented;
 *   - step-varying scalars (learning rate, step count, scale factors, skip flags) are read from
 *     device memory so that every launch can be captured in a CUDA graph;
 *   - dtype codes: 0 = float32, 1 = bfloat16.
 *
 * Each entry cites the reference call site it replaces (paths relative to the reference tree).
 */
#ifndef DETB200_H
#define DETB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* detb200_stream_t;

#define DETB200_ERR_BAD_ARG (-1)
#define DETB200_ERR_TOO_LARGE (-2)
#define DETB200_ERR_UNSUPPORTED (-3)
#define DETB200_MAX_RANKS 16

const char* detb200_last_error(void);
int detb200_abi_version(void);

/* ---------------------------------------------------------------------------------------------
 * Fused flat-buffer optimizers (csrc/optim.cu).
 * Reference: optimizer.step() at harness/determined/pytorch/_pytorch_context.py:899-907,
 * per-parameter grad.div_ :805-812, optimizer.zero_grad() :909-910.
 * One launch reads p,g,(state) once and writes p,(state) once and, if zero_grad!=0, g=0.
 * effective grad = g * static_grad_scale * (*grad_scale if non-NULL); update skipped when
 * skip!=NULL and *skip!=0 (AMP found-inf), gradients are still zeroed.
 * ------------------------------------------------------------------------------------------- */
int detb200_optim_tick(int* step, const int* skip, detb200_stream_t stream);
int detb200_adam_flat(float* p, float* g, float* m, float* v, void* p_lowp_bf16 /* nullable */, int64_t n,
                      const float* lr, const int* step, const float* grad_scale /* nullable */,
                      const int* skip /* nullable */, float beta1, float beta2, float eps, float weight_decay,
                      int decoupled_wd /* 1 = AdamW */, float static_grad_scale, int zero_grad, int maximize,
                      detb200_stream_t stream);
int detb200_sgd_flat(float* p, float* g, float* momentum_buf /* nullable */, int64_t n, const float* lr,
                     const int* step /* required with momentum */, const float* grad_scale, const int* skip,
                     float momentum, float dampening, float weight_decay, int nesterov, float static_grad_scale,
                     int zero_grad, int maximize, detb200_stream_t stream);
int detb200_adadelta_flat(float* p, float* g, float* square_avg, float* acc_delta, int64_t n, const float* lr,
                          const float* grad_scale, const int* skip, float rho, float eps, float weight_decay,
                          float static_grad_scale, int zero_grad, int maximize, detb200_stream_t stream);

/* Gradient-norm pieces of clip_grad_norm_ (reference _pytorch_context.py:885-888).
 * detb200_sumsq: *out (+)= sum(x^2); deterministic two-stage reduction.  `scratch` is
 * detb200_sumsq_scratch_bytes() bytes of device memory, zero-initialised once by the caller and
 * owned by one stream at a time.  nonfinite_flag (nullable) is OR-ed with 1 if the sum is not finite. */
int64_t detb200_sumsq_scratch_bytes(void);
int detb200_sumsq(const float* x, int64_t n, void* scratch, float* out, int* nonfinite_flag, int accumulate,
                  detb200_stream_t stream);
/* Same reduction with the scratch built into the library (a file-scope __device__ array of 4864
 * bytes per GPU, held until exit), for callers that have no device memory to give: processes whose
 * HBM is entirely owned by a pool/caching allocator, C callers without an allocator, and hot paths
 * that cannot afford cudaMalloc's device-wide sync.  No counterpart in the reference (it has no
 * native code at all).  n <= detb200_sumsq_noalloc_max_elems() (2^30) or DETB200_ERR_TOO_LARGE.
 * Allocates nothing.  Calls on one device share the array: order them on the GPU, overlapping
 * calls race.  The first call on a device resolves the array's address; do it before any stream
 * capture.  The device that owns `stream` must be the current device (DETB200_ERR_BAD_ARG
 * otherwise): the array is per GPU and the address is looked up for the current one. */
int64_t detb200_sumsq_noalloc_max_elems(void);
int detb200_sumsq_noalloc(const float* x, int64_t n, float* out, int* nonfinite_flag /* nullable */, int accumulate,
                          detb200_stream_t stream);
/* coef = extra_scale * min(1, max_norm / (sqrt(sum(terms)) + 1e-6)); norm_out nullable. */
int detb200_clip_coef(const float* sumsq_terms, int nterms, float max_norm, float extra_scale, float* coef_out,
                      float* norm_out, detb200_stream_t stream);
int detb200_scale_flat(float* x, int64_t n, const float* dyn_scale /* nullable */, float static_scale,
                       detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * In-kernel collectives over NVLink peer memory (csrc/comm.cu).
 * Reference: DDP all-reduce installed at _pytorch_context.py:296-299, Horovod DistributedOptimizer
 * :357-368, GradScaler.unscale_ :886-887, hvd.broadcast_parameters _pytorch_trial.py:602-606,
 * DeepSpeed ZeRO step (user engine.step(), pytorch/deepspeed/_deepspeed_trial.py:364-499).
 * peer_bufs / peer_ctrls are HOST arrays of `world` device pointers (own buffer at index `rank`,
 * peers' buffers as mapped with detb200_ipc_open).  phases: 1 start barrier | 2 work | 4 end
 * barrier; pass 7 on real multi-GPU runs.  timeout_s<=0 disables the in-kernel deadline.
 * ------------------------------------------------------------------------------------------- */
int64_t detb200_ctrl_bytes(void);
int detb200_ipc_handle_bytes(void);
int detb200_symm_alloc(void** ptr, int64_t bytes);     /* cudaMalloc + zero fill (synchronises) */
int detb200_symm_free(void* ptr);
int detb200_ipc_export(void* ptr, void* handle_out /* host, detb200_ipc_handle_bytes() */);
int detb200_ipc_open(const void* handle /* host */, void** ptr_out);
int detb200_ipc_close(void* ptr);
int detb200_enable_peer_access(int peer_device);
void* detb200_comm_stats_ptr(void* ctrl);               /* device float[2][DETB200_MAX_RANKS]: sumsq per rank, then nonfinite per rank */
int detb200_comm_read_stats(const void* ctrl, int world, float* stats_out /* host float[2] */,
                            detb200_stream_t stream);   /* synchronises the stream */
int detb200_allreduce_fused(void* const* peer_bufs, void* const* peer_ctrls, int rank, int world,
                            int64_t offset_elems, int64_t n_elems, float scale,
                            const float* inv_loss_scale /* nullable */, int* found_inf /* nullable */,
                            int first_of_step, int phases, double timeout_s, detb200_stream_t stream);
/* mode 0: reduce + AdamW + weight push in one pass; mode 1: reduce into grad_shard (+stats);
 * mode 2: AdamW from grad_shard + weight push. */
int detb200_rs_adam_ag(void* const* peer_grad_bufs, void* const* peer_ctrls, void* const* peer_param_bufs,
                       int param_is_bf16, int rank, int world, int64_t offset_elems, int64_t n_elems,
                       float* master_shard, float* m_shard, float* v_shard, float* grad_shard /* nullable in mode 0 */,
                       float scale, const float* lr, const int* step, const float* grad_scale, const int* skip,
                       float beta1, float beta2, float eps, float weight_decay, int mode, int first_of_step,
                       int phases, double timeout_s, detb200_stream_t stream);
/* Extended forms.  mc_buf / mc_grad / mc_param: this rank's mapping of an NVSwitch multicast object bound to every
 * rank's buffer (detb200_mc_*), or NULL for peer-to-peer loads and stores; with it the reduction happens inside the
 * switch (multimem.ld_reduce) and the result is written to all ranks by one multimem.st.  max_ctas <= 0: default
 * budget (32 CTAs; the rest of the GPU stays with the backward pass the collective overlaps). */
int detb200_allreduce_fused_ex(void* const* peer_bufs, void* const* peer_ctrls, void* mc_buf, int rank, int world,
                               int64_t offset_elems, int64_t n_elems, float scale, const float* inv_loss_scale,
                               int* found_inf, int first_of_step, int phases, int max_ctas, double timeout_s,
                               detb200_stream_t stream);
/* Test entry point: all `world` ranks of the all-reduce in one cooperative launch on ONE GPU (bufs / ctrls are `world`
 * allocations of that GPU; found_inf is int[world] or NULL): the flag / epoch / barrier protocol without a second GPU. */
int detb200_allreduce_fused_emulated(void* const* bufs, void* const* ctrls, int world, int64_t offset_elems,
                                     int64_t n_elems, float scale, const float* inv_loss_scale, int* found_inf,
                                     int first_of_step, int ctas_per_rank, double timeout_s, detb200_stream_t stream);
/* 16-bit wire format (reference hvd.Compression.fp16, _pytorch_context.py:367; expconf optimizations.gradient_compression):
 * narrow (local) -> allreduce_bf16 over the symmetric bf16 staging buffers -> widen (local).  Lengths multiples of 8. */
int detb200_narrow_bf16(const float* x, void* y_bf16, int64_t n, detb200_stream_t stream);
int detb200_widen_bf16(const void* x_bf16, float* y, int64_t n, detb200_stream_t stream);
int detb200_allreduce_bf16(void* const* peer_stage, void* const* peer_ctrls, void* mc_stage, int rank, int world,
                           int64_t offset_elems, int64_t n_elems, float scale, const float* inv_loss_scale,
                           int* found_inf, int first_of_step, int phases, int max_ctas, double timeout_s,
                           detb200_stream_t stream);
/* rs_adam_ag with separate gradient / parameter offsets (ZeRO-2 reduces out of a landing buffer), gradient
 * accumulation into grad_shard in mode 1 (accumulate != 0) and the multicast variants.  mode 3: reduce into grad_shard like
 * mode 1 but leave the step's statistics (sum of squares, non-finite count) alone and skip their exchange: for the
 * non-final micro-steps of gradient accumulation, whose partial sums are not what the optimizer will see. */
int detb200_rs_adam_ag_ex(void* const* peer_grad_bufs, void* const* peer_ctrls, void* const* peer_param_bufs /* nullable in mode 1 */,
                          void* mc_grad, void* mc_param, int param_is_bf16, int rank, int world,
                          int64_t grad_offset_elems, int64_t param_offset_elems, int64_t n_elems,
                          float* master_shard, float* m_shard, float* v_shard, float* grad_shard,
                          float scale, const float* lr, const int* step, const float* grad_scale, const int* skip,
                          float beta1, float beta2, float eps, float weight_decay, int mode, int accumulate,
                          int first_of_step, int phases, int max_ctas, double timeout_s, detb200_stream_t stream);
int detb200_broadcast(void* const* peer_bufs, void* const* peer_ctrls, int rank, int world, int root,
                      int64_t offset_elems, int64_t n_elems, int phases, double timeout_s,
                      detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Symmetric memory on the CUDA virtual-memory API and NVSwitch multicast objects (csrc/symm.cpp).
 * Reference: none of its own; NCCL's NVLS algorithm is what its DDP / DeepSpeed paths use on an NVSwitch box.
 * File descriptors returned here are passed to the peer processes over a unix socket (SCM_RIGHTS) by the caller.
 * A multicast mapping is for multimem.* instructions only; never build a tensor map over it.
 * ------------------------------------------------------------------------------------------- */
int detb200_vmm_query(int* caps /* host int[2]: exportable VMM allocations, multicast objects */);
int detb200_vmm_granularity(int world, int64_t* bytes_out);
int detb200_vmm_alloc(int64_t bytes, void** ptr, int* fd_out);          /* zero-filled; synchronises */
int detb200_vmm_import(int fd, int64_t bytes, void** ptr_out);          /* map a peer's allocation */
int detb200_vmm_free(void* ptr);
int detb200_mc_create(int world, int64_t bytes, int64_t* mc_out, int* fd_out);   /* one rank */
int detb200_mc_import(int fd, int64_t* mc_out);                                   /* the others */
int detb200_mc_add_device(int64_t mc);                                            /* all, before any bind */
int detb200_mc_bind(int64_t mc, void* local_ptr /* from detb200_vmm_alloc */, int64_t bytes);
int detb200_mc_map(int64_t mc, int64_t bytes, void** mc_ptr_out);
int detb200_mc_free(int64_t mc, void* mc_ptr, int64_t bytes);
int detb200_close_fd(int fd);

/* ---------------------------------------------------------------------------------------------
 * LayerNorm / RMSNorm (csrc/norm.cu).  Reference: ATen/apex/Megatron fused LN used by the GPT-2 and
 * GPT-NeoX examples (examples/deepspeed/gpt_neox/Dockerfile:11).  cols even, <= 8192.
 * ------------------------------------------------------------------------------------------- */
int detb200_norm_fwd(const void* x, const void* gamma, const void* beta /* nullable */, void* y,
                     float* mean /* nullable, rows */, float* rstd /* rows */, int64_t rows, int cols, float eps,
                     int dtype, int rms, detb200_stream_t stream);
int64_t detb200_norm_bwd_workspace_elems(int64_t rows, int cols, int rms);
int detb200_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                     void* dx, void* dgamma, void* dbeta /* nullable */, float* workspace, int64_t rows, int cols,
                     int dtype, int rms, detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * BatchNorm, channels-last [M = N*H*W, C] (csrc/batchnorm.cu).  Reference: nn.BatchNorm2d in the user's
 * model reaches ATen batch_norm_{collect_statistics,transform_input,backward_reduce,backward_elemt}
 * _channels_last kernels.  C must be a multiple of 8 (bf16) / 4 (fp32) with C/vec dividing 256.
 * `workspace`: detb200_bn_workspace_bytes(C) bytes, zero-initialised once, one stream at a time.
 * ------------------------------------------------------------------------------------------- */
int64_t detb200_bn_workspace_bytes(int C);
int detb200_bn_fwd_train(const void* x, void* y, int64_t M, int C, const float* gamma, const float* beta,
                         float* running_mean, float* running_var, long long* num_batches_tracked,
                         float* mean, float* invstd, float* scale, float* shift, float momentum, float eps,
                         int relu, void* workspace, int dtype, detb200_stream_t stream);
int detb200_bn_apply(const void* x, void* y, int64_t M, int C, const float* scale, const float* shift, int relu,
                     int dtype, detb200_stream_t stream);
int detb200_bn_bwd(const void* dy, const void* x, void* dx /* may alias dy */, int64_t M, int C, const float* gamma,
                   const float* mean, const float* invstd, float* dgamma, float* dbeta, float* coef /* 3*C */,
                   void* workspace, int dtype, detb200_stream_t stream);
/* Fused variants: y = relu?(bn(x) + residual?); the backward gates dy by (y > 0) when `y` is given and writes the gated dy
 * (gradient of the residual branch) to `dres` when non-null.  Replaces the separate add / ReLU / threshold_backward kernels
 * of a ResNet basic block. */
int detb200_bn_fwd_train_fused(const void* x, const void* residual /* nullable */, void* y, int64_t M, int C, const float* gamma,
                               const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                               float* mean, float* invstd, float* scale, float* shift, float momentum, float eps, int relu,
                               void* workspace, int dtype, detb200_stream_t stream);
int detb200_bn_bwd_fused(const void* dy, const void* x, const void* y /* nullable */, void* dx, void* dres /* nullable */, int64_t M, int C,
                         const float* gamma, const float* mean, const float* invstd, float* dgamma, float* dbeta, float* coef,
                         void* workspace, int dtype, detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Softmax cross-entropy (csrc/xent.cu).  Reference: log_softmax + nll_loss in
 * examples/tutorials/mnist_pytorch/model.py:34, train.py:43 and the HF/NeoX LM losses.
 * ------------------------------------------------------------------------------------------- */
int detb200_xent_fwd(const void* logits, const int64_t* target, float* loss /* rows */, float* lse /* rows */,
                     int64_t rows, int vocab, int64_t ignore_index, int dtype, detb200_stream_t stream);
int detb200_xent_bwd(const void* logits, const int64_t* target, const float* lse, const float* grad_scale,
                     void* dlogits /* may alias logits */, int64_t rows, int vocab, int64_t ignore_index,
                     int dtype, detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * bf16 GEMM on tcgen05 tensor cores with TMEM accumulators and TMA-fed operands (csrc/gemm_sm100.cu).
 * C[M,N] (+)= op(A) * op(B)^T with fp32 accumulation.  Reference: every nn.Linear / matmul the user's
 * model runs under DDP or DeepSpeed (harness/determined/pytorch/_pytorch_context.py:296-299 wraps the
 * model, the GEMMs themselves are library calls there).
 *   a_mn_major = 0: A is [M,K] row-major (lda = row stride);  1: A is stored [K,M] (lda = stride of k)
 *   b_mn_major = 0: B is [N,K] row-major (nn.Linear weight);  1: B is stored [K,N]
 *   epilogue: 0 none, 1 GELU(erf), 2 GELU(tanh), 3 ReLU  -> `aux` (nullable) receives the pre-activation (bf16);
 *             4/5/6: multiply by GELU'(erf) / GELU'(tanh) / ReLU' of the pre-activation read from `aux`.
 *   out_fp32: C is fp32 (else bf16); accumulate: C += result (e.g. weight gradients straight into a flat fp32 buffer).
 * All operand bases 16-byte aligned, all row strides multiples of 8 elements.
 * ------------------------------------------------------------------------------------------- */
int detb200_gemm_bf16(const void* A, const void* B, void* C, const void* bias /* nullable, [N] */, void* aux /* nullable */,
                      int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t ldaux,
                      int a_mn_major, int b_mn_major, int out_fp32, int bias_fp32, int epilogue, int accumulate,
                      detb200_stream_t stream);

/* GEMMs fused with their collective over NVLink peer memory (same file).  `peer_*` are host arrays of
 * `world` device pointers (own pointer at [rank], peers' IPC-mapped pointers elsewhere); `peer_ctrl`
 * points at detb200_gemm_ctrl_bytes() bytes of zero-initialised symmetric memory per rank.  `epoch` is the
 * 1-based index of this fused call on this ctrl block (every rank issues the same sequence).
 *   ag_gemm:  A is the [T, lda] gather buffer; this rank's rows [rank*T/world, ...) must be final in stream
 *             order.  n_push_ctas CTAs push them to the peers while the others run the GEMM; `ready_value` is
 *             the running total of n_push_ctas over all ag_gemm calls including this one.
 *   gemm_rs:  out[T/world, N] = sum over ranks of (A_r * B_r^T) rows of this rank (+bias), partial tiles travel
 *             as bf16; peer_recv[r] = [world][T/world][N] bf16 receive slots on rank r; `count_base` is the
 *             running total of detb200_gemm_rs_arrivals() over earlier gemm_rs calls.  world <= 8.
 * Reference: Megatron mpu all-gather / reduce-scatter around cuBLAS GEMMs in examples/deepspeed/gpt_neox. */
int64_t detb200_gemm_ctrl_bytes(void);
int detb200_gemm_rs_arrivals(int64_t T, int64_t N, int world);
int detb200_ag_gemm_bf16(void* const* peer_a, void* const* peer_ctrl, int rank, int world, const void* B, void* C, const void* bias,
                         void* aux, int64_t T, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t ldaux,
                         int b_mn_major, int out_fp32, int bias_fp32, int epilogue, unsigned int epoch, unsigned int ready_value,
                         int n_push_ctas, detb200_stream_t stream);
/* Generalised gather-fused GEMM: gather_b = 0 gathers A (as above); gather_b = 1 gathers B, the ZeRO-3 weight case:
 * gathered_mn_major = 0 -> B is [N, K] sharded over N (forward), 1 -> B is stored [K, N], sharded over the reduction (dgrad).
 * `other` is the local operand (A must be K-major). */
int detb200_ag_gemm_bf16_ex(void* const* peer_g, void* const* peer_ctrl, int rank, int world, const void* other, void* C, const void* bias,
                            void* aux, int64_t M, int64_t N, int64_t K, int64_t ld_g, int64_t ld_other, int64_t ldc, int64_t ldaux,
                            int gather_b, int gathered_mn_major, int other_mn_major, int out_fp32, int bias_fp32, int epilogue,
                            unsigned int epoch, unsigned int ready_value, int n_push_ctas, detb200_stream_t stream);
int detb200_gemm_rs_bf16(const void* A, const void* B, void* const* peer_recv, void* const* peer_ctrl, int rank, int world, void* out,
                         const void* bias, int64_t T, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldo, int a_mn_major,
                         int b_mn_major, int out_fp32, int bias_fp32, unsigned int epoch, unsigned int count_base,
                         detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Flash attention forward on tcgen05 (csrc/attention.cu): S = QK^T and O += PV on the tensor cores, S/P/O in
 * tensor memory, K/V streamed by TMA, online softmax in registers.  q/k/v/o are bf16 with logical shape
 * (B, S, H, D), element strides {batch, seq, head} and D contiguous (64 or 128); H_kv divides H (GQA).
 * lse (nullable): fp32 [B, H, S_q].  Reference: attention of the GPT-2 / GPT-NeoX examples
 * (examples/deepspeed/gpt_neox, Megatron fused softmax + bmm).
 * ------------------------------------------------------------------------------------------- */
int detb200_attn_debug_delay_v(int ns);   /* test hook: hold back every V tile of the forward kernel (0 = off) */
int detb200_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int H_kv, int S_q, int S_kv, int D,
                     const int64_t* q_strides, const int64_t* k_strides, const int64_t* v_strides, const int64_t* o_strides,
                     float scale, int causal, detb200_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Transformer glue (csrc/elementwise.cu).  Reference: the ATen elementwise / cat / reduce kernels the
 * GPT-2 / GPT-NeoX examples launch around attention and the linear layers.
 * ------------------------------------------------------------------------------------------- */
int detb200_rotary_qkv(void* qkv, const float* cos_t, const float* sin_t, int64_t B, int S, int H, int D, int rot, int64_t sb, int64_t ss,
                       int64_t sw, int64_t sh, int conj, detb200_stream_t stream);
int detb200_colsum_bf16(const void* x, float* out, int64_t M, int N, int64_t ld, int accumulate, detb200_stream_t stream);
int detb200_cast_bf16(const float* x, void* y, int64_t n, detb200_stream_t stream);
int detb200_add_bf16(const void* a, const void* b, const void* c /* nullable */, void* y, int64_t n, detb200_stream_t stream);
/* Multi-tensor gradient accumulation: dsts[k][0:numels[k]) (fp32) += srcs[k] (bf16 where is_bf16[k], else fp32), all k in
 * one launch per 48 tensors; the four tables are HOST arrays.  Reference: autograd's one-add-per-parameter accumulation
 * plus autocast's bf16 -> fp32 gradient casts inside loss.backward() (_pytorch_context.py:786-803). */
int detb200_accum_multi(const void* const* srcs, float* const* dsts, const int64_t* numels, const unsigned char* is_bf16,
                        int count, detb200_stream_t stream);

/* Backward of detb200_attn_fwd on tcgen05 (H_kv == H, S_q == S_kv, D == 128): five GEMMs per (key block, query block)
 * pair, dK/dV accumulated in tensor memory, dQ through fp32 atomics into `dq_acc` ([B,S,H,D], zeroed here) and converted
 * to bf16.  `delta`: fp32 [2,B,H,ceil64(S)] scratch.  dq/dk/dv may be the three slices of one packed [B,S,3,H,D] buffer. */
int detb200_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse, float* delta,
                     float* dq_acc, void* dq, void* dk, void* dv, int B, int H, int S, int D, const int64_t* q_strides,
                     const int64_t* k_strides, const int64_t* v_strides, const int64_t* o_strides, const int64_t* do_strides,
                     const int64_t* dq_strides, const int64_t* dk_strides, const int64_t* dv_strides, float scale, int causal,
                     detb200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DETB200_H */
