"""Checkpoint save / resume in torch ``state_dict`` format.

The reference has no checkpointing (no torch.save in train_ddp.py); BASELINE.json asks for the
stock format, so a checkpoint written here loads into torchvision ResNet + torch DDP + torch
SGD + torch GradScaler objects and vice versa:
  * ``model``     — ``state_dict()`` of the (wrapped) model: torchvision key names, ``module.``
                    prefix when wrapped; conv weights are logical ``[Cout, Cin, R, S]``;
  * ``optimizer`` — ``{"state": {i: {"momentum_buffer"}}, "param_groups": [...]}``;
  * ``scaler``    — the five GradScaler keys;
  * ``epoch`` / ``args``.
Tensors are cloned to contiguous host memory on save (the live ones are views of flat arenas).
"""
from __future__ import annotations

import os
from typing import Any

import torch


def _to_host(obj: Any) -> Any:
    if torch.is_tensor(obj):
        return obj.detach().to("cpu").contiguous().clone()
    if isinstance(obj, dict):
        return {k: _to_host(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_host(v) for v in obj)
    return obj


def strip_prefix(sd: dict, prefix: str = "module.") -> dict:
    if all(k.startswith(prefix) for k in sd):
        return {k[len(prefix):]: v for k, v in sd.items()}
    return sd


def add_prefix(sd: dict, prefix: str = "module.") -> dict:
    if all(k.startswith(prefix) for k in sd):
        return sd
    return {prefix + k: v for k, v in sd.items()}


def save_checkpoint(path: str, model, optimizer=None, scaler=None, epoch: int = 0,
                    args=None, rank: int = 0) -> None:
    """Rank 0 writes (all replicas are identical).  EVERY rank must call this: with the SGD
    update fused into the two-shot all-reduce each rank only holds the momentum of the slices
    it owns, and the shards are gathered collectively before rank 0 serialises the state."""
    if optimizer is not None and hasattr(model, "gather_optimizer_state"):
        model.gather_optimizer_state(optimizer)
    if rank != 0:
        return
    payload = {
        "model": _to_host(model.state_dict()),
        "optimizer": _to_host(optimizer.state_dict()) if optimizer is not None else None,
        "scaler": scaler.state_dict() if scaler is not None else None,
        "epoch": int(epoch),
        "args": dict(vars(args)) if args is not None else None,
    }
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(payload, tmp)
    os.replace(tmp, path)


def load_checkpoint(path: str, model, optimizer=None, scaler=None, map_location="cpu") -> dict:
    """Loads into live objects in place (copying into the arenas); returns the payload.
    Accepts checkpoints with or without the DDP ``module.`` prefix regardless of whether
    ``model`` is wrapped."""
    payload = torch.load(path, map_location=map_location, weights_only=False)
    sd = payload["model"] if "model" in payload else payload
    wants_prefix = all(k.startswith("module.") for k in model.state_dict().keys())
    sd = add_prefix(sd) if wants_prefix else strip_prefix(sd)
    model.load_state_dict(sd)
    if optimizer is not None and payload.get("optimizer") is not None:
        optimizer.load_state_dict(payload["optimizer"])
    if scaler is not None and payload.get("scaler"):
        scaler.load_state_dict(payload["scaler"])
    if hasattr(model, "sync_after_load"):
        model.sync_after_load()
    return payload
