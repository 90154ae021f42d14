"""Cross-entropy criterion backed by the fused CE kernel (forward + backward in one launch)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        from . import functional as Fn
        loss, dlogits = Fn.cross_entropy(logits.contiguous(), target.contiguous())
        ctx.save_for_backward(dlogits)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits * grad_out, None


class CrossEntropyLoss(nn.Module):
    """``nn.CrossEntropyLoss()`` (mean reduction, no smoothing — reference train_ddp.py:338).
    fp32 CUDA logits use the fused kernel; anything else goes through ATen."""

    def forward(self, logits, target):
        if logits.is_cuda and logits.dtype == torch.float32 and logits.dim() == 2:
            from . import _lib
            if _lib.available():
                return _CEFn.apply(logits, target)
        return F.cross_entropy(logits.float(), target)
