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

#ifdef __cplusplus
extern "C" {
#endif

#define DPT_ERR_SHAPE (-1)  /* unsupported shape / alignment */
#define DPT_ERR_DRIVER (-2) /* driver entry point unavailable */
#define DPT_ERR_TMAP (-3)   /* cuTensorMapEncodeTiled rejected the descriptor */
#define DPT_ERR_ARG (-4)    /* invalid argument */
#define DPT_ERR_TIMEOUT (-5)

/* ---- tcgen05 / TMA implicit-GEMM convolution (csrc/ops/igemm_tcgen05.cu) ----------------
 * Stand in for the cuDNN convolution forward / backward-data / backward-filter calls that
 * torchvision's resnet18 makes on the reference path (train_ddp.py:153-156, 205-213). */
int dpt_conv_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, float* stat_sum, float* stat_sq,
                   int out_f32_atomic, int splits, void* stream);
int dpt_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, const void* dy2, const void* w2, int R2,
                   int S2, int pad2, void* stream);
int dpt_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin, int Cout,
                   int R, int S, int stride, int pad, int splits, void* stream);
/* op: 0 = A[M,K]*B[N,K]^T, 1 = A[M,K]*B[K,N], 2 = A[K,M]^T*B[K,N]; out_mode 0 bf16, 1 fp32,
 * 2 fp32 atomic accumulate (cuBLAS addmm/mm on the reference path, train_ddp.py:205). */
int dpt_gemm_bf16(int op, const void* A, const void* B, void* C, int M, int N, int K,
                  long long lda, long long ldb, long long ldc, int out_mode, const float* bias,
                  float* stat_sum, float* stat_sq, int splits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPTB200_H_ */
