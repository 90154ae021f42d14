/* dptb200 — C ABI of the sm_100a kernel library (libdptb200.so).
 *
 * Every entry point returns 0 on success, a positive cudaError_t value when a CUDA runtime
 * call failed, or one of the negative DPT_ERR_* codes below.  `stream` is a cudaStream_t
 * passed as void*.  Nothing here allocates device memory unless the comment says so.
 *
 * Tensors: activations are NHWC bf16, conv weights are [Cout][R][S][Cin] ("KRSC") bf16 for
 * compute with fp32 masters/gradients of the same physical layout.
 */
#ifndef DPTB200_H_
#define DPTB200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPT_ERR_SHAPE (-1)  /* unsupported shape / alignment */
#define DPT_ERR_DRIVER (-2) /* driver entry point unavailable */
#define DPT_ERR_TMAP (-3)   /* cuTensorMapEncodeTiled rejected the descriptor */
#define DPT_ERR_ARG (-4)    /* invalid argument */
#define DPT_ERR_TIMEOUT (-5)
#define DPT_ERR_TOO_LARGE (-6) /* size beyond a built-in fixed-capacity buffer */

/* 16-bit element type of activations / shadow weights for every entry point below: 0 = bf16
 * (default), 1 = fp16 (the reference's autocast default, train_ddp.py:204).  Process-wide. */
int dpt_set_lowp_dtype(int is_f16);
int dpt_get_lowp_dtype(void);

/* ---- tcgen05 / TMA implicit-GEMM convolution (csrc/ops/igemm_tcgen05.cu) ----------------
 * Stand in for the cuDNN convolution forward / backward-data / backward-filter calls that
 * torchvision's resnet18 makes on the reference path (train_ddp.py:153-156, 205-213). */
/* EXPERIMENTAL two-SM GEMM (thread-block cluster of 2 CTAs, tcgen05.mma.cta_group::2, 256x256
 * tiles, persistent): C[M,N] = A[M,K] * B[N,K]^T, 16-bit row-major operands (lda/ldb multiples of
 * 8 elements), C 16-bit or fp32.  Not used by the model paths yet. */
int dpt_gemm2sm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                     long long ldb, long long ldc, int out_f32, int f16, void* stream);
/* Persistent-tile variant of the conv / GEMM kernels (CTAs loop over output tiles with a
 * double-buffered TMEM accumulator): used from `min_tiles` tiles upwards with `grid_ctas` CTAs;
 * <= 0 restores the defaults (4 waves, 2 CTAs per SM). */
int dpt_igemm_set_persist(long long min_tiles, int grid_ctas);
/* EXPERIMENTAL 128x256 output tiles (25 % less operand traffic per FLOP) where the column count
 * is a multiple of 256; also DPT_IGEMM_BN256=1.  Off by default. */
int dpt_igemm_set_bn256(int on);
int dpt_conv_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, float* stat_sum, float* stat_sq,
                   int out_f32_atomic, int splits, void* stream);
int dpt_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, const void* dy2, const void* w2, int R2,
                   int S2, int pad2, void* stream);
/* split-K variant: out_f32_atomic = 1 accumulates into a pre-zeroed fp32 dx with `splits` CTAs per tile */
int dpt_conv_dgrad_ex(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin,
                      int Cout, int R, int S, int stride, int pad, const void* dy2, const void* w2,
                      int R2, int S2, int pad2, int out_f32_atomic, int splits, void* stream);
int dpt_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, int splits, void* stream);
/* op: 0 = A[M,K]*B[N,K]^T, 1 = A[M,K]*B[K,N], 2 = A[K,M]^T*B[K,N]; out_mode 0 bf16, 1 fp32,
 * 2 fp32 atomic accumulate (cuBLAS addmm/mm on the reference path, train_ddp.py:205). */
int dpt_gemm_bf16(int op, const void* A, const void* B, void* C, int M, int N, int K,
                  long long lda, long long ldb, long long ldc, int out_mode, const float* bias,
                  float* stat_sum, float* stat_sq, int splits, void* stream);


/* ---- BatchNorm / pooling / head / loss (csrc/ops/norm_pool_loss.cu) --------------------------
 * Stand in for ATen/cuDNN native_batch_norm(+backward), relu_/threshold_backward, add_,
 * max_pool2d_with_indices(+backward), mean.dim + addmm (+mm x2), _log_softmax + nll_loss
 * (+backward) as dispatched by torchvision's resnet18 and nn.CrossEntropyLoss on the reference
 * path (train_ddp.py:205-207, 211-213, 338).  x/y are NHWC bf16 viewed as [M, C]. */
int dpt_bn_stats(const void* x, float* sum, float* sq, long long M, int C, void* stream);
int dpt_bn_fwd_train(const void* x, const float* sum, const float* sq, const float* gamma,
                     const float* beta, const void* residual, void* y, float* save_mean,
                     float* save_invstd, float* running_mean, float* running_var,
                     long long* num_batches_tracked, long long M, int C, float momentum, float eps,
                     int relu, void* stream);
/* single-launch statistics + normalisation for small tensors produced by a split-K conv (fp32 x) */
int dpt_bn_fwd_train_small_f32(const float* x, const float* gamma, const float* beta,
                               const void* residual, void* y, void* x_bf16, float* save_mean,
                               float* save_invstd, float* running_mean, float* running_var,
                               long long* num_batches_tracked, long long M, int C, float momentum,
                               float eps, int relu, void* stream);
int dpt_bn_fwd_eval(const void* x, const float* gamma, const float* beta, const float* running_mean,
                    const float* running_var, const void* residual, void* y, long long M, int C,
                    float eps, int relu, void* stream);
int dpt_bn_bwd_reduce(const void* dy1, const void* dy2, const void* y, const void* x,
                      const float* save_mean, const float* save_invstd, void* g_out, float* dgamma,
                      float* dbeta, long long M, int C, void* stream);
int dpt_bn_bwd_reduce_ex(const void* dy1, int dy1_is_f32, const void* dy2, const void* y,
                         const void* x, const float* save_mean, const float* save_invstd,
                         void* g_out, float* dgamma, float* dbeta, long long M, int C, void* stream);
/* both passes in one launch (grid barrier); scratch = 2*C zeroed floats, counter = zeroed u32 */
int dpt_bn_bwd_fused(const void* dy1, int dy1_is_f32, const void* dy2, const void* y, const void* x,
                     const float* save_mean, const float* save_invstd, const float* gamma,
                     void* g_out, float* dgamma, float* dbeta, float* scratch,
                     unsigned int* counter, void* dx, long long M, int C, void* stream);
int dpt_bn_bwd_apply(const void* g, const void* x, const float* save_mean, const float* save_invstd,
                     const float* gamma, const float* dgamma, const float* dbeta, void* dx,
                     long long M, int C, void* stream);
/* BatchNorm training forward for callers that may not allocate (fixed start-up memory plans,
 * loops that cannot take cudaMalloc's device-wide sync, plugins handed only pointers + a stream).
 * This entry point has no counterpart in the reference; it exists for those callers only.
 * Same result as dpt_bn_stats followed by dpt_bn_fwd_train (and, like them, on activations of
 * the element type set with dpt_set_lowp_dtype: bf16, fp16 or fp32), but the sum / sum-of-squares
 * scratch is a 16 KiB device array built into the library (held by every GPU that loads it
 * until exit), so nothing is allocated.  Limits: C <= DPT_BN_NOALLOC_MAX_CHANNELS, otherwise
 * DPT_ERR_TOO_LARGE; within the limit C must be 8 * 2^k like for the other BatchNorm entry
 * points (DPT_ERR_SHAPE otherwise), so the maximum itself is usable.  The first call on a device
 * resolves the array's address for that device (make it before any stream capture there).
 * Calls on one device share the array: it is cleared at the start of each call, and calls that
 * overlap on the GPU race — serialise them. */
#define DPT_BN_NOALLOC_MAX_CHANNELS 2048
int dpt_bn_fwd_train_noalloc(const void* x, const float* gamma, const float* beta,
                             const void* residual, void* y, float* save_mean, float* save_invstd,
                             float* running_mean, float* running_var,
                             long long* num_batches_tracked, long long M, int C, float momentum,
                             float eps, int relu, void* stream);
int dpt_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream);
int dpt_maxpool3x3s2_bwd(const void* dy, const void* dy2, const void* idx, void* dx, int N, int H,
                         int W, int C, void* stream);
int dpt_head_fwd(const void* x, const float* w, const float* b, float* pooled, float* logits, int B,
                 int HW, int C, int K, void* stream);
int dpt_head_bwd(const float* dlogits, const float* w, const float* pooled, void* dx, float* dw,
                 float* db, int B, int HW, int C, int K, void* stream);
/* loss = mean CE; dlogits = (softmax - onehot) * (*loss_scale) / B; metrics[3] (f64) accumulate
 * loss*B, #top-1 correct, B (the host-synchronising .item() calls of train_ddp.py:216-222). */
int dpt_cross_entropy(const float* logits, const long long* target, float* loss, float* dlogits,
                      const float* loss_scale, double* metrics, int B, int K, void* stream);
int dpt_im2col_nchw_f32(const float* x, void* out, int N, int C, int H, int W, int R, int S,
                        int stride, int pad, int ldk, void* stream);
int dpt_pad_rows_f32_to_bf16(const float* src, void* dst, int rows, int cols, int ld, void* stream);
/* RandomCrop(pad) + RandomHorizontalFlip + ToTensor + Normalize on the device (the reference does
 * it per sample on the host, train_ddp.py:91-96); mean3/std3 are HOST pointers to 3 floats. */
int dpt_augment_crop_flip_norm(const void* in_u8, const int* params, float* out, int N, int C, int S,
                               int pad, const float* mean3, const float* std3, void* stream);
int dpt_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, void* stream);

/* ---- flat-arena optimizer / AMP (csrc/ops/optim.cu) ------------------------------------------
 * torch.optim.SGD foreach path (train_ddp.py:339-344, 214), GradScaler unscale_/step/update
 * (train_ddp.py:207-209, 346) without the found_inf.item() host sync. */
int dpt_amp_unscale(float* g, long long n, const float* inv_scale, float* found_inf, void* stream);
int dpt_amp_update_scale(float* scale, int* growth_tracker, float* found_inf, float* inv_scale_out,
                         float growth_factor, float backoff_factor, int growth_interval,
                         void* stream);
int dpt_sgd_flat(float* p, const float* g, float* m, void* shadow_bf16, long long n, float lr,
                 float momentum, float dampening, float wd, int nesterov, const float* grad_scale,
                 const float* found_inf, void* stream);
int dpt_adam_flat(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                  float lr, float beta1, float beta2, float eps, float wd, int decoupled,
                  int* step_count, const float* grad_scale, const float* found_inf, void* stream);
int dpt_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
/* Kernels here are launched with programmatic dependent launch (they overlap their prologue with
 * the previous kernel of the same stream).  Call this right after making a stream wait on another
 * stream's event: the next launch is then an ordinary, fully ordered one. */
int dpt_pdl_skip_next(void);

/* ---- NVLink peer-memory collectives (csrc/comm/p2p_collectives.cu) ---------------------------
 * Stand in for ProcessGroupNCCL allreduce / broadcast / barrier as driven by torch DDP on the
 * reference path (train_ddp.py:65, 112, 166, 305-310).  dpt_symm_alloc (cudaMalloc + IPC export)
 * and dpt_comm_create (512 bytes of device counters + one pinned host word) are the only calls
 * in this header that allocate memory. */
typedef struct dpt_comm dpt_comm;
int dpt_symm_alloc(size_t bytes, void** ptr, void* ipc_handle_64);
int dpt_symm_open(const void* ipc_handle_64, void** ptr);
int dpt_symm_close(void* ptr);
int dpt_symm_free(void* ptr);
size_t dpt_comm_flag_bytes(void);
int dpt_comm_create(int rank, int world, void* const* flags, double timeout_s, dpt_comm** out);
int dpt_comm_destroy(dpt_comm* c);
unsigned dpt_comm_error(const dpt_comm* c);
/* Debug: device-synchronising copy of the per-CTA barrier epoch counters (<= 64 words) to host
 * memory; returns the number of words written (negative on error).  Equal on all ranks after a
 * completed step. */
int dpt_comm_epochs(const dpt_comm* c, unsigned* out, int n);
int dpt_allreduce_fused(dpt_comm* c, void* const* bufs, long long n, float scale,
                        const float* inv_scale, void* const* found_inf, void* const* param,
                        void* const* shadow, float* momentum, float lr, float mu, float wd,
                        int nesterov, int ctas, void* stream);
/* EXPERIMENTAL: fused-SGD two-shot all-reduce whose publish phase is one multicast store
 * (param_mc / shadow_mc: multicast mappings of the parameter / shadow arenas at this bucket). */
int dpt_allreduce_fused_mcpush(dpt_comm* c, void* const* bufs, long long n, float scale,
                               const float* inv_scale, void* const* found_inf, void* const* param,
                               void* const* shadow, float* momentum, void* param_mc, void* shadow_mc,
                               float lr, float mu, float wd, int nesterov, int ctas, void* stream);
int dpt_allreduce_oneshot(dpt_comm* c, void* const* in, float* out, long long n, float scale,
                          const float* inv_scale, void* const* found_inf, int ctas, void* stream);
/* The next dpt_allreduce_fused* / dpt_allreduce_nvls* launch made by this thread returns without
 * its closing cross-GPU barrier.  Legal only when another collective of the same communicator
 * follows on the same stream on every rank before the result is consumed (its opening barrier
 * then covers this kernel's peer stores); the reducer uses it for every bucket but the last.
 * (NCCL pipelines successive all-reduces of a stream the same way under torch DDP,
 * train_ddp.py:305-310.) */
int dpt_comm_skip_closing_barrier_next(void);
/* One-shot "push" all-reduce for small buckets (the exposed tail of backward): every rank stores
 * its n floats into slot [rank] of every peer's staging area (`staging[r]` = rank r's area of
 * 2 * world * cap floats in symmetric memory, double buffered), one cross-GPU barrier, then every
 * rank reduces from its own memory and applies the epilogue to its own copy.  found_inf / param /
 * shadow are this rank's LOCAL pointers (the SGD update is replicated, so momentum is complete on
 * every rank); param == NULL leaves the averaged gradient in `grads`.  Returns DPT_ERR_TOO_LARGE
 * when n > cap.  Stands in for NCCL's low-latency small-message all-reduce under torch DDP
 * (train_ddp.py:305-310) + GradScaler.unscale_ (:208) + optimizer.step (:214). */
int dpt_allreduce_oneshot_push(dpt_comm* c, float* grads, void* const* staging, long long cap,
                               long long n, float scale, const float* inv_scale, float* found_inf,
                               float* param, void* shadow, float* momentum, float lr, float mu,
                               float wd, int nesterov, int ctas, void* stream);
/* Step timeline (benchmarks/step_trace.py; the role DDP's Logger/Timer plays in the reference's
 * stack, torch reducer_timer.hpp:26-32, but on the device clock and valid under CUDA-graph
 * replay): `buf` = caller-zeroed device memory of (1 + records) * 32 bytes.  After
 * dpt_comm_set_trace every collective kernel appends {kind<<56 | bytes, t_start, t_after_first_
 * barrier, t_end} (globaltimer ns); dpt_trace_stamp appends {255<<56 | tag, t} from any stream. */
int dpt_comm_set_trace(dpt_comm* c, void* buf, int records);
int dpt_trace_stamp(void* buf, int records, int tag, void* stream);
/* NVLS (NVSwitch multicast) variant: multimem.ld_reduce + multimem.st through `mc`, this rank's
 * mapping of the multicast object bound to every rank's buffer (parallel/nvls.py sets it up). */
int dpt_allreduce_nvls(dpt_comm* c, void* mc, long long n, float scale, const float* inv_scale,
                       void* const* found_inf, int ctas, void* stream);
int dpt_allreduce_nvls_sgd(dpt_comm* c, void* mc, long long n, float scale, const float* inv_scale,
                           void* const* found_inf, float* param_local, float* momentum,
                           void* param_mc, void* shadow_mc, float lr, float mu, float wd,
                           int nesterov, int ctas, void* stream);
int dpt_broadcast(dpt_comm* c, void* const* bufs, long long bytes, int src, int ctas, void* stream);
int dpt_barrier(dpt_comm* c, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPTB200_H_ */
