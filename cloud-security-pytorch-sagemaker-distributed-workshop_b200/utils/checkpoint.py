"""Checkpoint contract.

The reference saves once, at the end, on rank 0: ``torch.save(state_dict, $SM_MODEL_DIR/model.pth)``
(``notebooks/code/cifar10-distributed-native-cpu.py:196-199``) and serves it with ``model_fn``
(``inference.py:28-34``).  We emit the same file with *un-prefixed* keys and ordinary contiguous CPU
tensors (flat-arena views and channels-last strides are an internal detail), so the reference's
unmodified ``inference.py`` loads it.  Beyond the reference: an optional full checkpoint
(model + optimizer + epoch) for resume.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.nn as nn


def unwrap(model: nn.Module) -> nn.Module:
    return model.module if hasattr(model, "module") and isinstance(model.module, nn.Module) else model


def export_state_dict(model: nn.Module) -> Dict[str, torch.Tensor]:
    sd = unwrap(model).state_dict()
    return {k: v.detach().to("cpu").contiguous().clone() for k, v in sd.items()}


def save_model(model: nn.Module, model_dir: str, filename: str = "model.pth") -> str:
    os.makedirs(model_dir, exist_ok=True)
    path = os.path.join(model_dir, filename)
    tmp = path + ".tmp"
    torch.save(export_state_dict(model), tmp)
    os.replace(tmp, path)
    return path


def load_model_state(model: nn.Module, path: str, strict: bool = True):
    sd = torch.load(path, map_location="cpu", weights_only=True)
    sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
    target = unwrap(model)
    out = target.load_state_dict(sd, strict=strict)
    if hasattr(model, "refresh_shadow"):
        model.refresh_shadow()
    return out


def save_full(model: nn.Module, optimizer, epoch: int, path: str, extra: Optional[dict] = None) -> str:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    opt_state = None
    if optimizer is not None:
        flat = getattr(optimizer, "_momentum_flat", None)
        if flat:
            opt_state = {"kind": "sgd_flat", "momentum": [m.detach().cpu() for m in flat.values()]}
        elif getattr(optimizer, "_state_flat", None):
            opt_state = {"kind": "adam_flat", "state": [(a.cpu(), b.cpu(), c.cpu()) for a, b, c in optimizer._state_flat.values()]}
        else:
            opt_state = {"kind": "torch", "state": optimizer.state_dict()}
    blob = {"model": export_state_dict(model), "optimizer": opt_state, "epoch": int(epoch), "extra": extra or {}}
    tmp = path + ".tmp"
    torch.save(blob, tmp)
    os.replace(tmp, path)
    return path


def load_full(model: nn.Module, optimizer, path: str) -> int:
    blob = torch.load(path, map_location="cpu", weights_only=True)      # tensors, lists, dicts and numbers only
    unwrap(model).load_state_dict(blob["model"])
    if hasattr(model, "refresh_shadow"):
        model.refresh_shadow()
    st = blob.get("optimizer")
    if optimizer is not None and st is not None:
        if st["kind"] == "sgd_flat" and getattr(optimizer, "_flat_owners", None):
            for o, m in zip(optimizer._flat_owners, st["momentum"]):
                optimizer._momentum_flat[id(o)] = m.to(o.flat_params.device)
        elif st["kind"] == "adam_flat" and getattr(optimizer, "_flat_owners", None):
            for o, (a, b, c) in zip(optimizer._flat_owners, st["state"]):
                dev = o.flat_params.device
                optimizer._state_flat[id(o)] = (a.to(dev), b.to(dev), c.to(dev))
        elif st["kind"] == "torch":
            optimizer.load_state_dict(st["state"])
    return int(blob.get("epoch", 0))
