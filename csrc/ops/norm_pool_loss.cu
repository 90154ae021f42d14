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
#define DPT_LOWP_IS_F16 1
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
#undef DPT_LOWP_IS_F16
#undef DPT_NS

extern "C" int dpt_bn_stats(const void* x, float* sum, float* sq, long long M, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_stats(x, sum, sq, M, C, stream);
        case 2: return dpt_lp_f32::dpt_bn_stats(x, sum, sq, M, C, stream);
        default: return dpt_lp_bf16::dpt_bn_stats(x, sum, sq, M, C, stream);
    }
}

extern "C" int dpt_bn_fwd_train(const void* x, const float* sum, const float* sq, const float* gamma, const float* beta, const void* residual, void* y, float* save_mean, float* save_invstd, float* running_mean, float* running_var, long long* num_batches_tracked, long long M, int C, float momentum, float eps, int relu, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_fwd_train(x, sum, sq, gamma, beta, residual, y, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
        case 2: return dpt_lp_f32::dpt_bn_fwd_train(x, sum, sq, gamma, beta, residual, y, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
        default: return dpt_lp_bf16::dpt_bn_fwd_train(x, sum, sq, gamma, beta, residual, y, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
    }
}

extern "C" int dpt_bn_fwd_train_small_f32(const float* x, const float* gamma, const float* beta, const void* residual, void* y, void* x_bf16, float* save_mean, float* save_invstd, float* running_mean, float* running_var, long long* num_batches_tracked, long long M, int C, float momentum, float eps, int relu, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_fwd_train_small_f32(x, gamma, beta, residual, y, x_bf16, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
        case 2: return dpt_lp_f32::dpt_bn_fwd_train_small_f32(x, gamma, beta, residual, y, x_bf16, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
        default: return dpt_lp_bf16::dpt_bn_fwd_train_small_f32(x, gamma, beta, residual, y, x_bf16, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
    }
}

extern "C" int dpt_bn_fwd_eval(const void* x, const float* gamma, const float* beta, const float* running_mean, const float* running_var, const void* residual, void* y, long long M, int C, float eps, int relu, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_fwd_eval(x, gamma, beta, running_mean, running_var, residual, y, M, C, eps, relu, stream);
        case 2: return dpt_lp_f32::dpt_bn_fwd_eval(x, gamma, beta, running_mean, running_var, residual, y, M, C, eps, relu, stream);
        default: return dpt_lp_bf16::dpt_bn_fwd_eval(x, gamma, beta, running_mean, running_var, residual, y, M, C, eps, relu, stream);
    }
}

extern "C" int dpt_bn_bwd_reduce_ex(const void* dy1, int dy1_is_f32, const void* dy2, const void* y, const void* x, const float* save_mean, const float* save_invstd, void* g_out, float* dgamma, float* dbeta, long long M, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_bwd_reduce_ex(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
        case 2: return dpt_lp_f32::dpt_bn_bwd_reduce_ex(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
        default: return dpt_lp_bf16::dpt_bn_bwd_reduce_ex(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
    }
}

extern "C" int dpt_bn_bwd_reduce(const void* dy1, const void* dy2, const void* y, const void* x, const float* save_mean, const float* save_invstd, void* g_out, float* dgamma, float* dbeta, long long M, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_bwd_reduce(dy1, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
        case 2: return dpt_lp_f32::dpt_bn_bwd_reduce(dy1, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
        default: return dpt_lp_bf16::dpt_bn_bwd_reduce(dy1, dy2, y, x, save_mean, save_invstd, g_out, dgamma, dbeta, M, C, stream);
    }
}

extern "C" int dpt_bn_bwd_fused(const void* dy1, int dy1_is_f32, const void* dy2, const void* y, const void* x, const float* save_mean, const float* save_invstd, const float* gamma, void* g_out, float* dgamma, float* dbeta, float* scratch, unsigned int* counter, void* dx, long long M, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_bwd_fused(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, gamma, g_out, dgamma, dbeta, scratch, counter, dx, M, C, stream);
        case 2: return dpt_lp_f32::dpt_bn_bwd_fused(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, gamma, g_out, dgamma, dbeta, scratch, counter, dx, M, C, stream);
        default: return dpt_lp_bf16::dpt_bn_bwd_fused(dy1, dy1_is_f32, dy2, y, x, save_mean, save_invstd, gamma, g_out, dgamma, dbeta, scratch, counter, dx, M, C, stream);
    }
}

extern "C" int dpt_bn_bwd_apply(const void* g, const void* x, const float* save_mean, const float* save_invstd, const float* gamma, const float* dgamma, const float* dbeta, void* dx, long long M, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_bn_bwd_apply(g, x, save_mean, save_invstd, gamma, dgamma, dbeta, dx, M, C, stream);
        case 2: return dpt_lp_f32::dpt_bn_bwd_apply(g, x, save_mean, save_invstd, gamma, dgamma, dbeta, dx, M, C, stream);
        default: return dpt_lp_bf16::dpt_bn_bwd_apply(g, x, save_mean, save_invstd, gamma, dgamma, dbeta, dx, M, C, stream);
    }
}

extern "C" int dpt_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_maxpool3x3s2_fwd(x, y, idx, N, H, W, C, stream);
        case 2: return dpt_lp_f32::dpt_maxpool3x3s2_fwd(x, y, idx, N, H, W, C, stream);
        default: return dpt_lp_bf16::dpt_maxpool3x3s2_fwd(x, y, idx, N, H, W, C, stream);
    }
}

extern "C" int dpt_maxpool3x3s2_bwd(const void* dy, const void* dy2, const void* idx, void* dx, int N, int H, int W, int C, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_maxpool3x3s2_bwd(dy, dy2, idx, dx, N, H, W, C, stream);
        case 2: return dpt_lp_f32::dpt_maxpool3x3s2_bwd(dy, dy2, idx, dx, N, H, W, C, stream);
        default: return dpt_lp_bf16::dpt_maxpool3x3s2_bwd(dy, dy2, idx, dx, N, H, W, C, stream);
    }
}

extern "C" int dpt_head_fwd(const void* x, const float* w, const float* b, float* pooled, float* logits, int B, int HW, int C, int K, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_head_fwd(x, w, b, pooled, logits, B, HW, C, K, stream);
        case 2: return dpt_lp_f32::dpt_head_fwd(x, w, b, pooled, logits, B, HW, C, K, stream);
        default: return dpt_lp_bf16::dpt_head_fwd(x, w, b, pooled, logits, B, HW, C, K, stream);
    }
}

extern "C" int dpt_head_bwd(const float* dlogits, const float* w, const float* pooled, void* dx, float* dw, float* db, int B, int HW, int C, int K, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_head_bwd(dlogits, w, pooled, dx, dw, db, B, HW, C, K, stream);
        case 2: return dpt_lp_f32::dpt_head_bwd(dlogits, w, pooled, dx, dw, db, B, HW, C, K, stream);
        default: return dpt_lp_bf16::dpt_head_bwd(dlogits, w, pooled, dx, dw, db, B, HW, C, K, stream);
    }
}

extern "C" int dpt_cross_entropy(const float* logits, const long long* target, float* loss, float* dlogits, const float* loss_scale, double* metrics, int B, int K, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_cross_entropy(logits, target, loss, dlogits, loss_scale, metrics, B, K, stream);
        case 2: return dpt_lp_f32::dpt_cross_entropy(logits, target, loss, dlogits, loss_scale, metrics, B, K, stream);
        default: return dpt_lp_bf16::dpt_cross_entropy(logits, target, loss, dlogits, loss_scale, metrics, B, K, stream);
    }
}

extern "C" int dpt_im2col_nchw_f32(const float* x, void* out, int N, int C, int H, int W, int R, int S, int stride, int pad, int ldk, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_im2col_nchw_f32(x, out, N, C, H, W, R, S, stride, pad, ldk, stream);
        case 2: return dpt_lp_f32::dpt_im2col_nchw_f32(x, out, N, C, H, W, R, S, stride, pad, ldk, stream);
        default: return dpt_lp_bf16::dpt_im2col_nchw_f32(x, out, N, C, H, W, R, S, stride, pad, ldk, stream);
    }
}

extern "C" int dpt_pad_rows_f32_to_bf16(const float* src, void* dst, int rows, int cols, int ld, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_pad_rows_f32_to_bf16(src, dst, rows, cols, ld, stream);
        case 2: return dpt_lp_f32::dpt_pad_rows_f32_to_bf16(src, dst, rows, cols, ld, stream);
        default: return dpt_lp_bf16::dpt_pad_rows_f32_to_bf16(src, dst, rows, cols, ld, stream);
    }
}

extern "C" int dpt_nchw_f32_to_nhwc_bf16(const float* x, void* y, int N, int C, int H, int W, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_nchw_f32_to_nhwc_bf16(x, y, N, C, H, W, stream);
        case 2: return dpt_lp_f32::dpt_nchw_f32_to_nhwc_bf16(x, y, N, C, H, W, stream);
        default: return dpt_lp_bf16::dpt_nchw_f32_to_nhwc_bf16(x, y, N, C, H, W, stream);
    }
}

extern "C" int dpt_bn_fwd_train_noalloc(const void* x, const float* gamma, const float* beta, const void* residual, void* y, float* save_mean, float* save_invstd, float* running_mean, float* running_var, long long* num_batches_tracked, long long M, int C, float momentum, float eps, int relu, void* stream) {
    return dpt_lp_bf16::dpt_bn_fwd_train_noalloc(x, gamma, beta, residual, y, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, M, C, momentum, eps, relu, stream);
}

extern "C" int dpt_augment_crop_flip_norm(const void* in_u8, const int* params, float* out, int N, int C, int S, int pad, const float* mean3, const float* std3, void* stream) {
    switch (dpt::lowp_dtype_flag()) {
        case 1: return dpt_lp_f16::dpt_augment_crop_flip_norm(in_u8, params, out, N, C, S, pad, mean3, std3, stream);
        case 2: return dpt_lp_f32::dpt_augment_crop_flip_norm(in_u8, params, out, N, C, S, pad, mean3, std3, stream);
        default: return dpt_lp_bf16::dpt_augment_crop_flip_norm(in_u8, params, out, N, C, S, pad, mean3, std3, stream);
    }
}
