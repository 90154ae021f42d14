"""Raw (non-autograd) Python entry points of the sm_100a kernels.

Tensors are NHWC bf16 activations ``[N, H, W, C]`` and KRSC conv weights ``[Cout, R, S, Cin]``
unless stated otherwise.  Every function enqueues on the current CUDA stream and returns
immediately; outputs may be passed in (``out=``) so that callers with a static memory plan
(CUDA-graph capture) allocate nothing.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call, count, ptr, stream

BF16 = torch.bfloat16


def set_lowp_dtype(dtype) -> None:
    """Select the 16-bit element type (torch.bfloat16 or torch.float16) for subsequent launches."""
    _lib.set_lowp_dtype(dtype)


def conv_out_hw(H: int, W: int, R: int, S: int, stride: int, pad: int) -> tuple[int, int]:
    return (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1


def conv_fprop(x, w, stride=1, pad=0, *, out=None, stats=None, splits=1):
    """y = conv(x, w).  ``stats=(sum, sq)`` (fp32 [Cout], pre-zeroed) receives the per-channel
    sum / sum of squares of y for the following BatchNorm.  ``splits>1`` accumulates into an
    fp32 ``out`` (pre-zeroed) with split-K atomics instead of storing bf16."""
    N, H, W, Cin = x.shape
    Cout, R, S, _ = w.shape
    P, Q = conv_out_hw(H, W, R, S, stride, pad)
    atomic = splits > 1
    if out is None:
        out = (torch.zeros if atomic else torch.empty)(
            (N, P, Q, Cout), device=x.device, dtype=torch.float32 if atomic else x.dtype)
    s1, s2 = stats if stats is not None else (None, None)
    call("dpt_conv_fprop", ptr(x), ptr(w), ptr(out), N, H, W, Cin, Cout, R, S, stride, pad,
         ptr(s1), ptr(s2), int(atomic), splits, stream())
    count()
    return out


def conv_dgrad(dy, w, x_shape, stride=1, pad=0, *, out=None, dy2=None, w2=None, pad2=0, splits=1):
    """dx = conv_backward_data(dy, w) [+ conv_backward_data(dy2, w2)] for convs sharing x.
    ``splits > 1``: ``out`` is a pre-zeroed fp32 tensor accumulated with split-K atomics."""
    N, H, W, Cin = x_shape
    Cout, R, S, _ = w.shape
    atomic = splits > 1
    if out is None:
        out = (torch.zeros if atomic else torch.empty)(
            (N, H, W, Cin), device=dy.device, dtype=torch.float32 if atomic else dy.dtype)
    R2, S2 = (w2.shape[1], w2.shape[2]) if w2 is not None else (0, 0)
    call("dpt_conv_dgrad_ex", ptr(dy), ptr(w), ptr(out), N, H, W, Cin, Cout, R, S, stride, pad,
         ptr(dy2), ptr(w2), R2, S2, pad2, int(atomic), splits, stream())
    count()
    return out


def conv_wgrad(dy, x, w_shape, stride=1, pad=0, *, out=None, splits=0):
    """dw (fp32, KRSC) += conv_backward_filter(dy, x).  ``out`` must be zeroed (or hold a
    gradient to accumulate into) before the call."""
    N, H, W, Cin = x.shape
    Cout, R, S, _ = w_shape
    if out is None:
        out = torch.zeros((Cout, R, S, Cin), device=dy.device, dtype=torch.float32)
    call("dpt_conv_wgrad", ptr(dy), ptr(x), ptr(out), N, H, W, Cin, Cout, R, S, stride, pad,
         splits, stream())
    count()
    return out


def gemm(op, A, B, C, M, N, K, lda, ldb, ldc, *, out_mode=0, bias=None, stats=None, splits=1):
    s1, s2 = stats if stats is not None else (None, None)
    call("dpt_gemm_bf16", op, ptr(A), ptr(B), ptr(C), M, N, K, lda, ldb, ldc, out_mode, ptr(bias),
         ptr(s1), ptr(s2), splits, stream())
    count()
    return C


def bn_stats(x, sum_, sq):
    C_ = x.shape[-1]
    call("dpt_bn_stats", ptr(x), ptr(sum_), ptr(sq), x.numel() // C_, C_, stream())
    count()


def bn_fwd_train(x, sum_, sq, gamma, beta, *, residual=None, relu=False, out=None, save_mean=None,
                 save_invstd=None, running_mean=None, running_var=None, num_batches_tracked=None,
                 momentum=0.1, eps=1e-5):
    C_ = x.shape[-1]
    M = x.numel() // C_
    if out is None:
        out = torch.empty_like(x)
    if save_mean is None:
        save_mean = torch.empty(C_, device=x.device, dtype=torch.float32)
        save_invstd = torch.empty(C_, device=x.device, dtype=torch.float32)
    call("dpt_bn_fwd_train", ptr(x), ptr(sum_), ptr(sq), ptr(gamma), ptr(beta), ptr(residual),
         ptr(out), ptr(save_mean), ptr(save_invstd), ptr(running_mean), ptr(running_var),
         ptr(num_batches_tracked), M, C_, momentum, eps, int(relu), stream())
    count()
    return out, save_mean, save_invstd


def bn_fwd_train_small_f32(x32, gamma, beta, *, residual=None, relu=False, out, x_bf16, save_mean,
                           save_invstd, running_mean=None, running_var=None,
                           num_batches_tracked=None, momentum=0.1, eps=1e-5):
    """Statistics + normalisation in one launch for a small fp32 tensor (split-K conv output)."""
    C_ = x32.shape[-1]
    M = x32.numel() // C_
    call("dpt_bn_fwd_train_small_f32", ptr(x32), ptr(gamma), ptr(beta), ptr(residual), ptr(out),
         ptr(x_bf16), ptr(save_mean), ptr(save_invstd), ptr(running_mean), ptr(running_var),
         ptr(num_batches_tracked), M, C_, momentum, eps, int(relu), stream())
    count()
    return out


def bn_fwd_eval(x, gamma, beta, running_mean, running_var, *, residual=None, relu=False, out=None,
                eps=1e-5):
    C_ = x.shape[-1]
    if out is None:
        out = torch.empty_like(x)
    call("dpt_bn_fwd_eval", ptr(x), ptr(gamma), ptr(beta), ptr(running_mean), ptr(running_var),
         ptr(residual), ptr(out), x.numel() // C_, C_, eps, int(relu), stream())
    count()
    return out


def bn_bwd(dy, x, save_mean, save_invstd, gamma, dgamma, dbeta, *, dy2=None, y=None, g_out=None,
           dx=None):
    """Fused (residual-sum + ReLU mask +) BatchNorm backward.  ``dgamma``/``dbeta`` are
    accumulated into (zero them first).  Returns (dx, g) where g is the masked incoming gradient
    (what flows into the residual branch)."""
    C_ = x.shape[-1]
    M = x.numel() // C_
    need_g = (dy2 is not None) or (y is not None)
    if need_g and g_out is None:
        g_out = torch.empty_like(x)
    f32 = dy.dtype == torch.float32
    need_g = need_g or f32
    if need_g and g_out is None:
        g_out = torch.empty_like(x)
    call("dpt_bn_bwd_reduce_ex", ptr(dy), int(f32), ptr(dy2), ptr(y), ptr(x), ptr(save_mean),
         ptr(save_invstd), ptr(g_out), ptr(dgamma), ptr(dbeta), M, C_, stream())
    g = g_out if need_g else dy
    if dx is None:
        dx = torch.empty_like(x)
    call("dpt_bn_bwd_apply", ptr(g), ptr(x), ptr(save_mean), ptr(save_invstd), ptr(gamma),
         ptr(dgamma), ptr(dbeta), ptr(dx), M, C_, stream())
    count(2)
    return dx, g


def bn_bwd_fused(dy, x, save_mean, save_invstd, gamma, dgamma, dbeta, scratch, counter, *, dy2=None,
                 y=None, g_out, dx):
    """Single-launch BatchNorm backward.  ``scratch`` (2*C fp32) and ``counter`` (int32 scalar)
    must be zero on entry; dgamma / dbeta are accumulated into."""
    C_ = x.shape[-1]
    M = x.numel() // C_
    call("dpt_bn_bwd_fused", ptr(dy), int(dy.dtype == torch.float32), ptr(dy2), ptr(y), ptr(x),
         ptr(save_mean), ptr(save_invstd), ptr(gamma), ptr(g_out), ptr(dgamma), ptr(dbeta),
         ptr(scratch), ptr(counter), ptr(dx), M, C_, stream())
    count()
    return dx, g_out


def maxpool_fwd(x, *, out=None, idx=None):
    N, H, W, C_ = x.shape
    P, Q = conv_out_hw(H, W, 3, 3, 2, 1)
    if out is None:
        out = torch.empty((N, P, Q, C_), device=x.device, dtype=x.dtype)
    if idx is None:
        idx = torch.empty((N, P, Q, C_), device=x.device, dtype=torch.uint8)
    call("dpt_maxpool3x3s2_fwd", ptr(x), ptr(out), ptr(idx), N, H, W, C_, stream())
    count()
    return out, idx


def maxpool_bwd(dy, idx, x_shape, *, out=None, dy2=None):
    N, H, W, C_ = x_shape
    if out is None:
        out = torch.empty(x_shape, device=dy.device, dtype=dy.dtype)
    call("dpt_maxpool3x3s2_bwd", ptr(dy), ptr(dy2), ptr(idx), ptr(out), N, H, W, C_, stream())
    count()
    return out


def head_fwd(x, w, b, *, pooled=None, logits=None):
    """Global average pool over HW + linear: x [N,H,W,C] bf16, w [K,C] fp32, b [K] fp32."""
    N, H, W, C_ = x.shape
    K = w.shape[0]
    if pooled is None:
        pooled = torch.empty((N, C_), device=x.device, dtype=torch.float32)
    if logits is None:
        logits = torch.empty((N, K), device=x.device, dtype=torch.float32)
    call("dpt_head_fwd", ptr(x), ptr(w), ptr(b), ptr(pooled), ptr(logits), N, H * W, C_, K, stream())
    count()
    return logits, pooled


def head_bwd(dlogits, w, pooled, x_shape, dw, db, *, dx=None, dtype=BF16, part="both"):
    """part: "both" | "dx" (input gradient only) | "dw" (weight / bias gradients only) — the two
    kernels are independent, so an engine can put the weight gradient on its side stream."""
    N, H, W, C_ = x_shape
    K = w.shape[0]
    if dx is None and part != "dw":
        dx = torch.empty(x_shape, device=dlogits.device, dtype=dtype)
    call("dpt_head_bwd", ptr(dlogits), ptr(w), ptr(pooled), ptr(dx) if part != "dw" else None,
         ptr(dw) if part != "dx" else None, ptr(db) if part != "dx" else None, N, H * W, C_, K,
         stream())
    count(2 if part == "both" else 1)
    return dx


def cross_entropy(logits, target, *, loss=None, dlogits=None, loss_scale=None, metrics=None,
                  need_grad=True):
    B, K = logits.shape
    if loss is None:
        loss = torch.empty((), device=logits.device, dtype=torch.float32)
    if need_grad and dlogits is None:
        dlogits = torch.empty_like(logits)
    call("dpt_cross_entropy", ptr(logits), ptr(target), ptr(loss), ptr(dlogits) if need_grad else None,
         ptr(loss_scale), ptr(metrics), B, K, stream())
    count()
    return loss, dlogits


def im2col_nchw(x, R, S, stride, pad, ldk, *, out=None, dtype=BF16):
    N, C_, H, W = x.shape
    P, Q = conv_out_hw(H, W, R, S, stride, pad)
    if out is None:
        out = torch.empty((N * P * Q, ldk), device=x.device, dtype=dtype)
    call("dpt_im2col_nchw_f32", ptr(x), ptr(out), N, C_, H, W, R, S, stride, pad, ldk, stream())
    count()
    return out


def pad_rows_to_bf16(src, rows, cols, ld, *, out=None, dtype=BF16):
    if out is None:
        out = torch.empty((rows, ld), device=src.device, dtype=dtype)
    call("dpt_pad_rows_f32_to_bf16", ptr(src), ptr(out), rows, cols, ld, stream())
    count()
    return out


def nchw_to_nhwc_bf16(x, *, out=None, dtype=BF16):
    N, C_, H, W = x.shape
    if out is None:
        out = torch.empty((N, H, W, C_), device=x.device, dtype=dtype)
    call("dpt_nchw_f32_to_nhwc_bf16", ptr(x), ptr(out), N, C_, H, W, stream())
    count()
    return out


def sgd_flat(p, g, m, shadow, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
             grad_scale=None, found_inf=None):
    call("dpt_sgd_flat", ptr(p), ptr(g), ptr(m), ptr(shadow), p.numel(), lr, momentum, dampening,
         weight_decay, int(nesterov), ptr(grad_scale), ptr(found_inf), stream())
    count()


def adam_flat(p, g, m, v, shadow, step_count, lr, beta1=0.9, beta2=0.999, eps=1e-8,
              weight_decay=0.0, decoupled=False, grad_scale=None, found_inf=None):
    call("dpt_adam_flat", ptr(p), ptr(g), ptr(m), ptr(v), ptr(shadow), p.numel(), lr, beta1, beta2,
         eps, weight_decay, int(decoupled), ptr(step_count), ptr(grad_scale), ptr(found_inf),
         stream())
    count(2)


def amp_unscale(g, inv_scale, found_inf):
    call("dpt_amp_unscale", ptr(g), g.numel(), ptr(inv_scale), ptr(found_inf), stream())
    count()


def amp_update_scale(scale, growth_tracker, found_inf, inv_scale_out, growth_factor=2.0,
                     backoff_factor=0.5, growth_interval=2000):
    call("dpt_amp_update_scale", ptr(scale), ptr(growth_tracker), ptr(found_inf),
         ptr(inv_scale_out), growth_factor, backoff_factor, growth_interval, stream())
    count()


def cast_to_bf16(src, dst):
    call("dpt_cast_f32_to_bf16", ptr(src), ptr(dst), src.numel(), stream())
    count()


def augment_crop_flip_norm(images_u8, params, mean, std, pad=4, *, out=None):
    """Device-side CIFAR train transform.  images_u8: uint8 [N,C,S,S] (device); params: int32
    [N,3] = (dy, dx, flip) on the device; mean/std: sequences of C floats."""
    import ctypes
    N, C_, S, _ = images_u8.shape
    if out is None:
        out = torch.empty((N, C_, S, S), device=images_u8.device, dtype=torch.float32)
    m = (ctypes.c_float * 3)(*(list(mean) + [0.0] * 3)[:3])
    sd = (ctypes.c_float * 3)(*(list(std) + [1.0] * 3)[:3])
    call("dpt_augment_crop_flip_norm", ptr(images_u8), ptr(params), ptr(out), N, C_, S, pad,
         ctypes.cast(m, ctypes.c_void_p), ctypes.cast(sd, ctypes.c_void_p), stream())
    count()
    return out
