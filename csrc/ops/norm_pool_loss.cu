// Bandwidth-bound model kernels for sm_100a: BatchNorm (training fwd / bwd with fused ReLU and
// residual add, eval fold), 3x3/2 max-pool, global-average-pool + linear head, fused
// cross-entropy (+loss scale, +top-1 count), im2col for the RGB stem.
//
// All activations are NHWC bf16 with C % 8 == 0, processed as 16-byte vectors; statistics and
// parameters are fp32.  These kernels replace cuDNN/ATen batch_norm(+backward), relu_/
// threshold_backward, add_, max_pool2d_with_indices(+backward), mean.dim, addmm/mm,
// _log_softmax/nll_loss(+backward) on the reference's path (train_ddp.py:205-207, 211-213 via
// torchvision resnet._forward_impl).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/dptb200.h"
#include "../common/ptx.cuh"


// ---- element type plumbing -----------------------------------------------------------------
// Activations are 16-bit: bf16 (default) or fp16 (the reference's literal autocast default,
// train_ddp.py:204).  The implementation header is instantiated once per type; the C ABI entry
// points dispatch on the process-wide setting made with dpt_set_lowp_dtype().
namespace dpt {
inline int& lowp_dtype_flag() {
    static int f = 0;  // 0 = bf16, 1 = fp16, 2 = fp32 (TF32 tensor-core math)
    return f;
}
inline bool lowp_is_f16() { return lowp_dtype_flag() == 1; }
}  // namespace dpt

extern "C" int dpt_igemm_set_lowp_dtype(int is_f16);
extern "C" int dpt_optim_set_lowp_dtype(int is_f16);
extern "C" int dpt_comm_set_lowp_dtype(int is_f16);
extern "C" int dpt_set_lowp_dtype(int is_f16) {
    if (is_f16 < 0 || is_f16 > 2) is_f16 = 0;
    dpt::lowp_dtype_flag() = is_f16;
    dpt_igemm_set_lowp_dtype(is_f16);
    dpt_optim_set_lowp_dtype(is_f16);
    dpt_comm_set_lowp_dtype(is_f16);
    return 0;
}
extern "C" int dpt_get_lowp_dtype() { return dpt::lowp_dtype_flag(); }

// ------------------------------------------------------------------------------------------------
// Built-in scratch of the allocation-free BatchNorm entry point (dpt_bn_fwd_train_noalloc, at
// the end of this file).  One array per loaded library and GPU, shared by all element types.
constexpr int kBnNoAllocMaxChannels = 2048;   // == DPT_BN_NOALLOC_MAX_CHANNELS (dptb200.h)
static __device__ __align__(256) float g_bn_noalloc_scratch[2 * kBnNoAllocMaxChannels];

namespace {
constexpr int kMaxDevices = 64;
float* g_bn_scratch_dev_ptr[kMaxDevices] = {nullptr};
std::mutex g_bn_scratch_mutex;

// In host code `g_bn_noalloc_scratch` is only a handle for the symbol, not a device address:
// resolve it through the runtime for the current device and cache the answer per device.
int bn_noalloc_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (dev < 0 || dev >= kMaxDevices) return DPT_ERR_ARG;
    std::lock_guard<std::mutex> lock(g_bn_scratch_mutex);
    if (g_bn_scratch_dev_ptr[dev] == nullptr) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_bn_noalloc_scratch);
        if (e != cudaSuccess) return static_cast<int>(e);
        g_bn_scratch_dev_ptr[dev] = ptr;
    }
    *out = g_bn_scratch_dev_ptr[dev];
    return 0;
}
}  // namespace

namespace dpt_lp_bf16 {
using lowp = __nv_bfloat16;
using lowp2 = __nv_bfloat162;
__device__ __forceinline__ float2 lp2_to_float2(lowp2 v) { return __bfloat1622float2(v); }
__device__ __forceinline__ lowp2 floats_to_lp2(float a, float b) { return __floats2bfloat162_rn(a, b); }
__device__ __forceinline__ float lp_to_float(lowp v) { return __bfloat162float(v); }
__device__ __forceinline__ lowp float_to_lp(float v) { return __float2bfloat16_rn(v); }
}  // namespace dpt_lp_bf16
#define DPT_NS dpt_lp_bf16
#include "norm_pool_loss_impl.cuh"
#undef DPT_NS

namespace dpt_lp_f16 {
using lowp = __half;
using lowp2 = __half2;
__device__ __forceinline__ float2 lp2_to_float2(lowp2 v) { return __half22float2(v); }
__device__ __forceinline__ lowp2 floats_to_lp2(float a, float b) { return __floats2half2_rn(a, b); }
__device__ __forceinline__ float lp_to_float(lowp v) { return __half2float(v); }
__device__ __forceinline__ lowp float_to_lp(float v) { return __float2half_rn(v); }
}  // namespace dpt_lp_f16
#define DPT_NS dpt_lp_f16
#include "norm_pool_loss_impl.cuh"
#undef DPT_NS

// fp32 activations (the reference's FP32 mode; convolutions then run as TF32 on the tensor cores)
namespace dpt_lp_f32 {
using lowp = float;
using lowp2 = float2;
__device__ __forceinline__ float2 lp2_to_float2(lowp2 v) { return v; }
__device__ __forceinline__ lowp2 floats_to_lp2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float lp_to_float(lowp v) { return v; }
__device__ __forceinline__ lowp float_to_lp(float v) { return v; }
}  // namespace dpt_lp_f32
#define DPT_NS dpt_lp_f32
#include "norm_pool_loss_impl.cuh"
#undef DPT_NS

// Every entry point below forwards to the instantiation selected by dpt_set_lowp_dtype().
#define DPT_DISPATCH(name, params, ...)                                        \
    extern "C" int name params {                                               \
        switch (dpt::lowp_dtype_flag()) {                                      \
            case 1: return dpt_lp_f16::name(__VA_ARGS__);                      \
            case 2: return dpt_lp_f32::name(__VA_ARGS__);                      \
            default: return dpt_lp_bf16::name(__VA_ARGS__);                    \
        }                                                                      \
    }

DPT_DISPATCH(dpt_bn_stats,
             (const void* x, float* sum, float* sq, long long M, int C, void* stream),
             x, sum, sq, M, C, stream)
DPT_DISPATCH(dpt_bn_fwd_train,
             (const void* x, const float* sum, const float* sq, const float* gamma, const float*
              beta, const void* residual, void* y, float* save_mean, float* save_invstd, float*
              running_mean, float* running_var, long long* num_batches_tracked, long long M, int C,
              float momentum, float eps, int relu, void* stream),
             x, sum, sq, gamma, beta, residual, y, save_mean, save_invstd, running_mean,
             running_var, num_batches_tracked, M, C, momentum, eps, relu, stream)
DPT_DISPATCH(dpt_bn_fwd_train_small_f32,
             (const float* x, const float* gamma, const float* beta, const void* residual, void* y,
              void* x_bf16, float* save_mean, float* save_invstd, float* running_mean, float*
              running_var, long long* num_batches_tracked, long long M, int C, float momentum, float
              eps, int relu, void* stream),
             x, gamma, beta, residual, y, x_bf16, save_mean, save_invstd, running_mean, running_var,
             num_batches_tracked, M, C, momentum, eps, relu, stream)
DPT_DISPATCH(dpt_bn_fwd_eval,
             (const void* x, const float* gamma, const float* beta, const float* running_mean, const
              float* running_var, const void* residual, void* y, long long M, int C, float eps, int
              relu, void* stream),
             x, gamma, beta, running_mean, running_var, residual, y, M, C, eps, relu, stream)
DPT_DISPATCH(dpt_bn_bwd_reduce_ex,
             (const void* dy1, int dy1_is_f32, const void* dy2, const void* y, const void* x, const
              float* save_mean, const float* save_invstd, void* g_out, float* dgamma, float* dbeta,
              long long M, int C, void* stream),
             dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream)
DPT_DISPATCH(dpt_bn_bwd_reduce,
             (const void* dy1, const void* dy2, const void* y, const void* x, const float*
              save_mean, const float* save_invstd, void* g_out, float* dgamma, float* dbeta, long
              long M, int C, void* stream),
             dy1, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream)
DPT_DISPATCH(dpt_bn_bwd_fused,
             (const void* dy1, int dy1_is_f32, const void* dy2, const void* y, const void* x, const
              float* save_mean, const float* save_invstd, const float* gamma, void* g_out, float*
              dgamma, float* dbeta, float* scratch, unsigned int* counter, void* dx, long long M,
              int C, void* stream),
             dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, gamma, g_out, dgamma, dbeta,
             scratch, counter, dx, M, C, stream)
DPT_DISPATCH(dpt_bn_bwd_apply,
             (const void* g, const void* x, const float* save_mean, const float* save_invstd, const
              float* gamma, const float* dgamma, const float* dbeta, void* dx, long long M, int C,
              void* stream),
             g, x, save_mean, save_invstd, gamma, dgamma, dbeta, dx, M, C, stream)
DPT_DISPATCH(dpt_maxpool3x3s2_fwd,
             (const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream),
             x, y, idx, N, H, W, C, stream)
DPT_DISPATCH(dpt_maxpool3x3s2_bwd,
             (const void* dy, const void* dy2, const void* idx, void* dx, int N, int H, int W, int
              C, void* stream),
             dy, dy2, idx, dx, N, H, W, C, stream)
DPT_DISPATCH(dpt_head_fwd,
             (const void* x, const float* w, const float* b, float* pooled, float* logits, int B,
              int HW, int C, int K, void* stream),
             x, w, b, pooled, logits, B, HW, C, K, stream)
DPT_DISPATCH(dpt_head_bwd,
             (const float* dlogits, const float* w, const float* pooled, void* dx, float* dw, float*
              db, int B, int HW, int C, int K, void* stream),
             dlogits, w, pooled, dx, dw, db, B, HW, C, K, stream)
DPT_DISPATCH(dpt_cross_entropy,
             (const float* logits, const long long* target, float* loss, float* dlogits, const
              float* loss_scale, double* metrics, int B, int K, void* stream),
             logits, target, loss, dlogits, loss_scale, metrics, B, K, stream)
DPT_DISPATCH(dpt_im2col_nchw_f32,
             (const float* x, void* out, int N, int C, int H, int W, int R, int S, int stride, int
              pad, int ldk, void* stream),
             x, out, N, C, H, W, R, S, stride, pad, ldk, stream)
DPT_DISPATCH(dpt_pad_rows_f32_to_bf16,
             (const float* src, void* dst, int rows, int cols, int ld, void* stream),
             src, dst, rows, cols, ld, stream)
DPT_DISPATCH(dpt_nchw_f32_to_nhwc_bf16,
             (const float* x, void* y, int N, int C, int H, int W, void* stream),
             x, y, N, C, H, W, stream)
DPT_DISPATCH(dpt_augment_crop_flip_norm,
             (const void* in_u8, const int* params, float* out, int N, int C, int S, int pad, const
              float* mean3, const float* std3, void* stream),
             in_u8, params, out, N, C, S, pad, mean3, std3, stream)
#undef DPT_DISPATCH

// ------------------------------------------------------------------------------------------------
// Allocation-free BatchNorm training forward.
//
// Who needs this: a BatchNorm forward in training mode needs a little scratch between its two
// passes (per-channel sum and sum of squares).  The regular entry points (dpt_bn_stats +
// dpt_bn_fwd_train) leave that scratch to the caller, who normally gets it from cudaMalloc or
// from a framework allocator.  Some callers must not do either: a host that has handed its whole
// memory budget to a fixed plan at start-up and forbids later allocations, a latency-critical
// loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply, or a plugin
// that is called with nothing but raw pointers and a stream and has no allocator to ask.  For
// them this entry point carries its own scratch: a file-scope device array (top of this file), so
// every GPU that loads this library holds 2 * DPT_BN_NOALLOC_MAX_CHANNELS floats (16 KiB) until
// the process exits.  It calls no cudaMalloc / cudaFree and no allocating library routine; the
// caller still passes every input and output.
//
// Rules of use:
//  * Activations have the element type set with dpt_set_lowp_dtype (bf16, fp16 or fp32), like
//    every other entry point.
//  * C must be <= DPT_BN_NOALLOC_MAX_CHANNELS (2048); larger C is rejected with
//    DPT_ERR_TOO_LARGE and a message on stderr.  Within the limit C must have the form the other
//    BatchNorm entry points accept, 8 * 2^k (otherwise DPT_ERR_SHAPE) — 2048 is reachable.
//  * The first call on each device looks up the device address of the scratch array (once per
//    device, with that device current) — make that first call before starting any stream
//    capture on that device.
//  * All calls on a device share the one array, and it keeps whatever the previous call left;
//    the entry point clears the part it uses before accumulating.  Two calls that overlap on the
//    GPU (different streams without ordering between them) race on it: serialise them.
extern "C" int dpt_bn_fwd_train_noalloc(const void* x, const float* gamma, const float* beta,
                                        const void* residual, void* y, float* save_mean,
                                        float* save_invstd, float* running_mean,
                                        float* running_var, long long* num_batches_tracked,
                                        long long M, int C, float momentum, float eps, int relu,
                                        void* stream) {
    if (C > kBnNoAllocMaxChannels) {
        fprintf(stderr,
                "[dptb200] dpt_bn_fwd_train_noalloc: C = %d exceeds the built-in scratch "
                "(max %d channels); use dpt_bn_stats + dpt_bn_fwd_train with your own buffer\n",
                C, kBnNoAllocMaxChannels);
        return DPT_ERR_TOO_LARGE;
    }
    if (C <= 0 || C % 8 != 0 || (256 % (C / 8)) != 0 || M <= 0) return DPT_ERR_SHAPE;
    float* scratch = nullptr;
    const int rc = bn_noalloc_scratch(&scratch);
    if (rc != 0) return rc;
    // the array still holds the previous call's sums: clear the 2*C floats used here
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(float) * 2 * C,
                                    static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return static_cast<int>(e);
    switch (dpt::lowp_dtype_flag()) {
        case 1:
            return dpt_lp_f16::bn_fwd_train_with_scratch(scratch, x, gamma, beta, residual, y,
                                                         save_mean, save_invstd, running_mean,
                                                         running_var, num_batches_tracked, M, C,
                                                         momentum, eps, relu, stream);
        case 2:
            return dpt_lp_f32::bn_fwd_train_with_scratch(scratch, x, gamma, beta, residual, y,
                                                         save_mean, save_invstd, running_mean,
                                                         running_var, num_batches_tracked, M, C,
                                                         momentum, eps, relu, stream);
        default:
            return dpt_lp_bf16::bn_fwd_train_with_scratch(scratch, x, gamma, beta, residual, y,
                                                          save_mean, save_invstd, running_mean,
                                                          running_var, num_batches_tracked, M, C,
                                                          momentum, eps, relu, stream);
    }
}
