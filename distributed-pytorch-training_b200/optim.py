"""Optimizers over the flat arenas: one kernel per step, AMP-aware, host-sync free.

``SGD`` mirrors ``torch.optim.SGD`` as the reference configures it (train_ddp.py:339-344:
lr 0.1, momentum 0.9, weight_decay 5e-4 applied to every parameter, no scheduler) including the
``state_dict`` layout (``momentum_buffer`` per parameter index).  ``Adam`` / ``AdamW`` are the extra
fused variants.  On CPU tensors the same formulas run through torch ops so the gloo
plumbing tests exercise identical semantics.
"""
from __future__ import annotations

import torch
from torch.optim import Optimizer

from .parallel.arena import ParamArena, arena_of


class _FlatOptimizer(Optimizer):
    supports_found_inf = True  # GradScaler.step passes the device word instead of syncing

    def _setup_arena(self) -> None:
        params = [p for g in self.param_groups for p in g["params"]]
        self._params = params
        self.arena = arena_of(params)
        dev = params[0].device
        self._cuda = dev.type == "cuda"
        if self.arena is None and self._cuda and len(self.param_groups) == 1:
            named = [(str(i), p) for i, p in enumerate(params)]
            self.arena = ParamArena(named, dev)
        full = self.arena is not None and len(self.arena.order) == len(params) \
            and len(self.param_groups) == 1
        self._flat = full and self._cuda
        self.grads_already_unscaled = False

    @property
    def flat_grad(self):
        return self.arena.grad if self._flat else None

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self.arena is not None:
            # gradients live in the arena: clearing is one memset and the views stay attached
            self.arena.zero_grad()
            self.arena.rebind_grads()
        else:
            super().zero_grad(set_to_none=set_to_none)

    def _state_buf(self, p, key):
        st = self.state[p]
        if key not in st:
            st[key] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st[key]


class SGD(_FlatOptimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0,
                 nesterov=False) -> None:
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                        nesterov=nesterov, maximize=False, foreach=None, differentiable=False,
                        fused=None)
        super().__init__(params, defaults)
        self._setup_arena()
        self._momentum_flat = None
        self.skip_update = False  # set by the wrapper when the all-reduce already applied SGD

    def _flat_momentum(self):
        if self._momentum_flat is None:
            self._momentum_flat = torch.zeros_like(self.arena.param)
            for p in self._params:
                s = self.arena.slot(p)
                self.state[p]["momentum_buffer"] = self._momentum_flat.as_strided(
                    s.shape, s.stride, s.offset)
        return self._momentum_flat

    @torch.no_grad()
    def step(self, closure=None, *, found_inf=None, grad_scale=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.skip_update:
            return loss
        g0 = self.param_groups[0]
        if self._flat:
            from .ops import functional as Fn
            m = self._flat_momentum() if g0["momentum"] != 0 else self.arena.param  # unused if 0
            Fn.sgd_flat(self.arena.param, self.arena.grad, m, self.arena.shadow, g0["lr"],
                        g0["momentum"], g0["dampening"], g0["weight_decay"], g0["nesterov"],
                        grad_scale=grad_scale, found_inf=found_inf)
            return loss
        if found_inf is not None and float(found_inf.item()) > 0:
            return loss
        for group in self.param_groups:
            lr, mu, damp, wd, nest = (group["lr"], group["momentum"], group["dampening"],
                                      group["weight_decay"], group["nesterov"])
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.is_cuda:
                    from .ops import functional as Fn
                    m = self._state_buf(p, "momentum_buffer") if mu != 0 else p
                    Fn.sgd_flat(p.data.view(-1), p.grad.view(-1), m.view(-1), None, lr, mu, damp,
                                wd, nest, grad_scale=grad_scale)
                    continue
                d = p.grad if grad_scale is None else p.grad * grad_scale
                if wd != 0:
                    d = d.add(p, alpha=wd)
                if mu != 0:
                    st = self.state[p]
                    if "momentum_buffer" not in st:
                        buf = st["momentum_buffer"] = torch.clone(d).detach()
                    else:
                        buf = st["momentum_buffer"]
                        buf.mul_(mu).add_(d, alpha=1 - damp)
                    d = d.add(buf, alpha=mu) if nest else buf
                p.add_(d, alpha=-lr)
        arena = getattr(self, "arena", None)
        if arena is not None:
            arena.refresh_shadow()
        return loss

    def load_state_dict(self, state_dict) -> None:
        if self._flat:
            # copy momentum buffers into the flat arena instead of replacing the views
            flat = self._flat_momentum()
            for idx, st in state_dict.get("state", {}).items():
                p = self._params[int(idx)]
                if "momentum_buffer" in st and st["momentum_buffer"] is not None:
                    s = self.arena.slot(p)
                    flat.as_strided(s.shape, s.stride, s.offset).copy_(st["momentum_buffer"])
            for g, sg in zip(self.param_groups, state_dict["param_groups"]):
                for k, v in sg.items():
                    if k != "params":
                        g[k] = v
            return
        super().load_state_dict(state_dict)


class Adam(_FlatOptimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                 decoupled_weight_decay=False) -> None:
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                        decoupled_weight_decay=decoupled_weight_decay)
        super().__init__(params, defaults)
        self._setup_arena()
        self._m = self._v = self._step = None

    def _flat_state(self):
        if self._m is None:
            self._m = torch.zeros_like(self.arena.param)
            self._v = torch.zeros_like(self.arena.param)
            self._step = torch.zeros((), dtype=torch.int32, device=self.arena.param.device)
            for p in self._params:
                s = self.arena.slot(p)
                self.state[p]["exp_avg"] = self._m.as_strided(s.shape, s.stride, s.offset)
                self.state[p]["exp_avg_sq"] = self._v.as_strided(s.shape, s.stride, s.offset)
                self.state[p]["step"] = self._step
        return self._m, self._v, self._step

    @torch.no_grad()
    def step(self, closure=None, *, found_inf=None, grad_scale=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g0 = self.param_groups[0]
        if self._flat:
            from .ops import functional as Fn
            m, v, step = self._flat_state()
            Fn.adam_flat(self.arena.param, self.arena.grad, m, v, self.arena.shadow, step, g0["lr"],
                         g0["betas"][0], g0["betas"][1], g0["eps"], g0["weight_decay"],
                         g0["decoupled_weight_decay"], grad_scale=grad_scale, found_inf=found_inf)
            return loss
        if found_inf is not None and float(found_inf.item()) > 0:
            return loss
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if "step" not in st:
                    st["step"] = torch.zeros((), dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                t = float(st["step"])
                g = p.grad if grad_scale is None else p.grad * grad_scale
                if group["decoupled_weight_decay"]:
                    p.mul_(1 - group["lr"] * group["weight_decay"])
                elif group["weight_decay"] != 0:
                    g = g.add(p, alpha=group["weight_decay"])
                st["exp_avg"].mul_(b1).add_(g, alpha=1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
                bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
                denom = (st["exp_avg_sq"].sqrt() / (bc2 ** 0.5)).add_(group["eps"])
                p.addcdiv_(st["exp_avg"], denom, value=-group["lr"] / bc1)
        arena = getattr(self, "arena", None)
        if arena is not None:
            arena.refresh_shadow()
        return loss

    def load_state_dict(self, state_dict) -> None:
        """Flat mode: copy the saved moments INTO the flat arenas the kernel updates (the base
        class would replace state[p][...] with fresh tensors the flat step never reads, silently
        restarting from zero moments and step 0)."""
        if not self._flat:
            return super().load_state_dict(state_dict)
        m, v, step = self._flat_state()
        saved_step = None
        for idx, st in state_dict.get("state", {}).items():
            p = self._params[int(idx)]
            s = self.arena.slot(p)
            if st.get("exp_avg") is not None:
                m.as_strided(s.shape, s.stride, s.offset).copy_(st["exp_avg"])
            if st.get("exp_avg_sq") is not None:
                v.as_strided(s.shape, s.stride, s.offset).copy_(st["exp_avg_sq"])
            if st.get("step") is not None:
                saved_step = int(float(st["step"]))
        if saved_step is not None:
            step.fill_(saved_step)
        for g, sg in zip(self.param_groups, state_dict["param_groups"]):
            for k, val in sg.items():
                if k != "params":
                    g[k] = val


class AdamW(Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2) -> None:
        super().__init__(params, lr, betas, eps, weight_decay, decoupled_weight_decay=True)


def build_optimizer(name: str, params, lr, momentum=0.9, weight_decay=5e-4):
    if name == "sgd":
        return SGD(params, lr=lr, momentum=momentum, weight_decay=weight_decay)
    if name == "adam":
        return Adam(params, lr=lr, weight_decay=weight_decay)
    if name == "adamw":
        return AdamW(params, lr=lr, weight_decay=weight_decay)
    raise ValueError(name)
