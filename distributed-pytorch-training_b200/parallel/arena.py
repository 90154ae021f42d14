"""Flat parameter / gradient / shadow arenas.

All trainable tensors of a model are re-pointed to views of ONE contiguous fp32 buffer, with
matching flat buffers for gradients and for the bf16 "shadow" weights the tensor-core kernels
read.  Consequences:
  * the optimizer is a single kernel over the arena (csrc/ops/optim.cu) instead of torch's
    4 multi-tensor launches (torch/optim/sgd.py:425-473);
  * a gradient bucket is a contiguous slice, so the reducer never copies gradients in or out
    (torch's Reducer copies both ways with gradient_as_bucket_view=False, reducer.hpp:276-334);
  * the wgrad kernels write straight into their slice of the gradient arena;
  * when a Communicator is given the arenas are symmetric memory that peers map over NVLink.

Layout = reversed registration order (the order gradients become ready in backward), each tensor
aligned to 8 elements so every bf16 shadow slice is a legal TMA base address.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

ALIGN = 8


@dataclass
class Slot:
    name: str
    offset: int   # in elements
    numel: int
    shape: tuple
    stride: tuple


def _dense_strides(p: torch.Tensor) -> tuple:
    """Strides that keep the parameter's memory format (channels_last conv weights stay KRSC)."""
    if p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous():
        return p.stride()
    if p.is_contiguous():
        return p.stride()
    return torch.empty(p.shape).stride()


class ParamArena:
    def __init__(self, named_params, device, comm=None, with_shadow=True, multicast=False) -> None:
        named_params = [(n, p) for n, p in named_params if p.requires_grad]
        self.device = torch.device(device)
        self.comm = comm
        self.slots: dict[int, Slot] = {}
        self.order: list[torch.nn.Parameter] = []
        off = 0
        for name, p in reversed(named_params):
            if p.dtype != torch.float32:
                raise TypeError(f"arena parameters must be fp32 masters ({name}: {p.dtype})")
            self.slots[id(p)] = Slot(name, off, p.numel(), tuple(p.shape), tuple(_dense_strides(p)))
            self.order.append(p)
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self.numel = max(off, ALIGN)
        self.symmetric = comm is not None and self.device.type == "cuda"
        if self.symmetric:
            self.param_buf = (comm.alloc_multicast(self.numel * 4) if multicast else None) \
                or comm.alloc(self.numel * 4)
            # gradients may additionally be bound to an NVSwitch multicast object (NVLS)
            self.grad_buf = (comm.alloc_multicast(self.numel * 4) if multicast else None) \
                or comm.alloc(self.numel * 4)
            self.shadow_buf = ((comm.alloc_multicast(self.numel * 2) if multicast else None)
                               or comm.alloc(self.numel * 2)) if with_shadow else None
            self.param = self.param_buf.view(torch.float32, 0, self.numel)
            self.grad = self.grad_buf.view(torch.float32, 0, self.numel)
            self.shadow = self.shadow_buf.view(torch.bfloat16, 0, self.numel) if with_shadow else None
        else:
            self.param_buf = self.grad_buf = self.shadow_buf = None
            self.param = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            self.grad = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            self.shadow = (torch.zeros(self.numel, dtype=torch.bfloat16, device=self.device)
                           if with_shadow and self.device.type == "cuda" else None)
        with torch.no_grad():
            for p in self.order:
                s = self.slots[id(p)]
                view = self.param.as_strided(s.shape, s.stride, s.offset)
                view.copy_(p.detach().to(self.device))
                p.data = view
                p.grad = self.grad.as_strided(s.shape, s.stride, s.offset)
                p._dpt_arena = self
        self.refresh_shadow()

    # ------------------------------------------------------------------ views
    def slot(self, p) -> Slot:
        return self.slots[id(p)]

    def grad_view(self, p) -> torch.Tensor:
        s = self.slots[id(p)]
        return self.grad.as_strided(s.shape, s.stride, s.offset)

    def shadow_view(self, p) -> torch.Tensor:
        """bf16 copy of ``p`` in physical (storage) order, flattened."""
        s = self.slots[id(p)]
        return self.shadow[s.offset: s.offset + s.numel]

    def flat_grad_view(self, p) -> torch.Tensor:
        s = self.slots[id(p)]
        return self.grad[s.offset: s.offset + s.numel]

    def flat_param_view(self, p) -> torch.Tensor:
        s = self.slots[id(p)]
        return self.param[s.offset: s.offset + s.numel]

    def set_shadow_dtype(self, dtype) -> None:
        """Re-type the 16-bit shadow arena (bf16 <-> fp16) and refill it from the fp32 masters."""
        if self.shadow is None or self.shadow.dtype == dtype or dtype == torch.float32:
            return
        self.shadow = self.shadow.view(dtype)
        self.refresh_shadow()

    def refresh_shadow(self) -> None:
        if self.shadow is None:
            return
        if self.device.type == "cuda":
            from ..ops import functional as Fn
            Fn.set_lowp_dtype(self.shadow.dtype)
            Fn.cast_to_bf16(self.param, self.shadow)
        else:
            self.shadow.copy_(self.param)

    def rebind_grads(self) -> None:
        """Re-attach ``.grad`` views (e.g. after a ``zero_grad(set_to_none=True)``)."""
        for p in self.order:
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + self.slots[id(p)].offset * 4:
                p.grad = self.grad_view(p)

    def zero_grad(self) -> None:
        self.grad.zero_()


def arena_of(params) -> "ParamArena | None":
    """The arena shared by all ``params`` (None if any is outside an arena or they differ)."""
    arena = None
    for p in params:
        a = getattr(p, "_dpt_arena", None)
        if a is None or (arena is not None and a is not arena):
            return None
        arena = a
    return arena
