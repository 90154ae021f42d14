"""GradScaler-compatible loss scaler backed by device scalars (no host round-trip per step).

API / state machine parity with torch.amp.GradScaler (torch/amp/grad_scaler.py:53-714) as the
reference uses it (train_ddp.py:207-209, 346): ``scale() -> unscale_() -> step() -> update()``,
init_scale 65536, growth 2.0, backoff 0.5, interval 2000, ``state_dict()`` with the same five keys.

What changes underneath:
  * ``found_inf`` is a device word that every fused all-reduce epilogue ORs into (csrc/comm);
    ``step()`` never calls ``.item()`` — the flat optimizer kernel reads the word itself and skips
    the update (torch syncs the host at grad_scaler.py:371);
  * ``update()`` is one tiny kernel that also emits ``1/scale`` for the next step's epilogue;
  * with bf16 (or when disabled) everything is a no-op and ``scale()`` returns the loss unchanged.
On CPU the same semantics are implemented with torch ops (used by the gloo plumbing tests).
"""
from __future__ import annotations

from enum import Enum

import torch


class OptState(Enum):
    READY = 0
    UNSCALED = 1
    STEPPED = 2


class GradScaler:
    def __init__(self, device: str | torch.device = "cuda", init_scale: float = 2.0 ** 16,
                 growth_factor: float = 2.0, backoff_factor: float = 0.5,
                 growth_interval: int = 2000, enabled: bool = True) -> None:
        self._device = torch.device(device)
        if enabled and self._device.type == "cuda" and not torch.cuda.is_available():
            enabled = False  # same self-disabling behaviour as torch (grad_scaler.py:134-140)
        self._enabled = enabled
        self._init_scale = float(init_scale)
        self._growth_factor = float(growth_factor)
        self._backoff_factor = float(backoff_factor)
        self._growth_interval = int(growth_interval)
        self._scale = None          # f32 device scalar, created lazily like torch
        self._growth_tracker = None  # i32 device scalar
        self._inv_scale = None
        self._found_inf = None
        self._state: dict[int, OptState] = {}
        # set by DistributedDataParallel.bind_scaler: the all-reduce epilogue multiplies every
        # gradient by inv_scale and raises found_inf, so unscale_() must not do it a second time
        # (eager steps after a TrainStep bound the scaler, e.g. the partial batch ending an epoch)
        self.unscaled_by_reducer = False

    # ------------------------------------------------------------------ device state
    def _lazy_init(self, device=None) -> None:
        if self._scale is not None:
            return
        dev = torch.device(device) if device is not None else self._device
        if dev.type == "cuda" and dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self._device = dev
        self._scale = torch.full((), self._init_scale, dtype=torch.float32, device=dev)
        self._growth_tracker = torch.zeros((), dtype=torch.int32, device=dev)
        self._inv_scale = torch.full((), 1.0 / self._init_scale, dtype=torch.float32, device=dev)
        self._found_inf = torch.zeros((), dtype=torch.float32, device=dev)

    def bind_found_inf(self, found_inf: torch.Tensor) -> None:
        """Use an externally owned found_inf word (the symmetric one peers write into)."""
        self._lazy_init(found_inf.device)
        self._found_inf = found_inf

    @property
    def found_inf(self):
        self._lazy_init()
        return self._found_inf

    @property
    def inv_scale(self):
        self._lazy_init()
        return self._inv_scale

    @property
    def scale_tensor(self):
        self._lazy_init()
        return self._scale

    def is_enabled(self) -> bool:
        return self._enabled

    # ------------------------------------------------------------------ API
    def scale(self, outputs):
        if not self._enabled:
            return outputs
        self._lazy_init(outputs.device)
        return outputs * self._scale.to(outputs.dtype)

    def _native(self) -> bool:
        return self._device.type == "cuda"

    def unscale_(self, optimizer) -> None:
        if not self._enabled:
            return
        self._lazy_init()
        st = self._state.get(id(optimizer), OptState.READY)
        if st is OptState.UNSCALED:
            raise RuntimeError("unscale_() has already been called on this optimizer since the last update().")
        if st is OptState.STEPPED:
            raise RuntimeError("unscale_() is being called after step().")
        if getattr(optimizer, "grads_already_unscaled", False) or self.unscaled_by_reducer:
            pass  # the fused all-reduce epilogue applied inv_scale and the inf check
        elif hasattr(optimizer, "flat_grad") and optimizer.flat_grad is not None and self._native():
            from .ops import functional as Fn
            Fn.amp_unscale(optimizer.flat_grad, self._inv_scale, self._found_inf)
        else:
            inv = self._inv_scale
            for group in optimizer.param_groups:
                for p in group["params"]:
                    if p.grad is None:
                        continue
                    bad = (~torch.isfinite(p.grad)).any().to(torch.float32)
                    self._found_inf.copy_(torch.maximum(self._found_inf, bad.to(self._found_inf.device)))
                    p.grad.mul_(inv.to(p.grad.device))
        self._state[id(optimizer)] = OptState.UNSCALED

    def step(self, optimizer, *args, **kwargs):
        if not self._enabled:
            return optimizer.step(*args, **kwargs)
        self._lazy_init()
        st = self._state.get(id(optimizer), OptState.READY)
        if st is OptState.STEPPED:
            raise RuntimeError("step() has already been called since the last update().")
        if st is OptState.READY:
            self.unscale_(optimizer)
        if getattr(optimizer, "supports_found_inf", False):
            # device-side skip: the kernel reads found_inf, the host never waits
            ret = optimizer.step(*args, found_inf=self._found_inf, **kwargs)
        else:
            ret = None
            if float(self._found_inf.item()) == 0.0:  # generic torch optimizers need the sync
                ret = optimizer.step(*args, **kwargs)
        self._state[id(optimizer)] = OptState.STEPPED
        return ret

    def update(self, new_scale=None) -> None:
        if not self._enabled:
            return
        self._lazy_init()
        if new_scale is not None:
            val = float(new_scale) if not torch.is_tensor(new_scale) else float(new_scale.item())
            self._scale.fill_(val)
            self._inv_scale.fill_(1.0 / val)
            self._found_inf.zero_()
        elif self._native():
            from .ops import functional as Fn
            Fn.amp_update_scale(self._scale, self._growth_tracker, self._found_inf, self._inv_scale,
                                self._growth_factor, self._backoff_factor, self._growth_interval)
        else:
            if float(self._found_inf.item()) > 0:
                self._scale.mul_(self._backoff_factor)
                self._growth_tracker.zero_()
            else:
                self._growth_tracker.add_(1)
                if int(self._growth_tracker.item()) == self._growth_interval:
                    self._scale.mul_(self._growth_factor)
                    self._growth_tracker.zero_()
            self._inv_scale.copy_(1.0 / self._scale)
            self._found_inf.zero_()
        self._state.clear()

    # ------------------------------------------------------------------ accessors
    def get_scale(self) -> float:
        if not self._enabled:
            return 1.0
        return self._init_scale if self._scale is None else float(self._scale.item())

    def get_growth_factor(self) -> float:
        return self._growth_factor

    def get_backoff_factor(self) -> float:
        return self._backoff_factor

    def get_growth_interval(self) -> int:
        return self._growth_interval

    def state_dict(self) -> dict:
        """Same keys as torch (grad_scaler.py:627-652)."""
        if not self._enabled:
            return {}
        return {
            "scale": self.get_scale(),
            "growth_factor": self._growth_factor,
            "backoff_factor": self._backoff_factor,
            "growth_interval": self._growth_interval,
            "_growth_tracker": 0 if self._growth_tracker is None else int(self._growth_tracker.item()),
        }

    def load_state_dict(self, state_dict: dict) -> None:
        if not self._enabled:
            return
        if len(state_dict) == 0:
            raise RuntimeError("The source state dict is empty, possibly because it was saved "
                               "from a disabled instance of GradScaler.")
        self._init_scale = float(state_dict["scale"])
        self._growth_factor = float(state_dict["growth_factor"])
        self._backoff_factor = float(state_dict["backoff_factor"])
        self._growth_interval = int(state_dict["growth_interval"])
        if self._scale is not None:
            self._scale.fill_(self._init_scale)
            self._inv_scale.fill_(1.0 / self._init_scale)
        self._lazy_init()
        self._growth_tracker.fill_(int(state_dict["_growth_tracker"]))
