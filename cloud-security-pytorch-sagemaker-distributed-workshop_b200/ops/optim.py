"""Fused optimizers: ``SGD`` with momentum and ``Adam`` with torch semantics.

``optim.SGD(model.parameters(), lr=..., momentum=...)`` of the reference
(``notebooks/code/cifar10-distributed-native-cpu.py:144``) and ``torch.optim.Adam(lr=1e-3)`` of the MNTD
trainers (``notebooks/code/utils_basic.py:96``).  When the parameters live in a
``DistributedDataParallel`` flat arena on a CUDA device the whole step is ONE kernel over the arena:
it reads grad/param/state once, writes param/state, refreshes the bf16 shadow weights used by the
tensor-core kernels and zeroes the gradient arena for the next backward.  Anything else (CPU, loose
parameters) takes torch's own optimizer math — that is the CPU fallback, not a GPU code path.
"""
from __future__ import annotations

from typing import Iterable

import torch

from . import native as nv


def _owners(params):
    owners = {}
    loose = []
    for p in params:
        o = getattr(p, "_b200_owner", None)
        o = o() if o is not None else None
        if o is not None and p.is_cuda:
            owners.setdefault(id(o), (o, []))[1].append(p)
        else:
            loose.append(p)
    flat, rest = [], list(loose)
    for o, ps in owners.values():
        if len(ps) == len(o.params):
            flat.append(o)
        else:
            rest.extend(ps)
    return flat, rest


class _FusedBase(torch.optim.Optimizer):
    def _split(self):
        params = [p for g in self.param_groups for p in g["params"]]
        self._flat_owners, self._loose = _owners(params)

    def zero_grad(self, set_to_none: bool = False):
        """Arena gradients are views and are zeroed by ``step``; loose ones follow torch semantics."""
        for p in self._loose:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()
        for o in self._flat_owners:
            if getattr(o, "_arena_dirty", True):     # a backward wrote gradients that no fused step has consumed (and zeroed) yet
                o.flat_grads.zero_()
                o._arena_dirty = False


class SGD(_FusedBase):
    def __init__(self, params: Iterable, lr: float = 0.01, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("b200ddp.optim.SGD takes a single parameter group")
        self._split()
        self._inner = torch.optim.SGD(self._loose, **defaults) if self._loose else None
        self._momentum_flat = {}
        self._fused_key = {}

    # -------------------------------------------------------------- optimizer inside the collective
    def fuse_into_backward(self, enabled: bool = True) -> bool:
        """Fold this optimizer's step into the gradient collective of every DDP arena that supports it (world > 1, backend
        'b200'): each bucket then runs reduce-scatter -> SGD on the owned shard -> all-gather of the new weights as ONE
        kernel while backward is still producing the next bucket, and ``step()`` launches nothing for that arena.  Only for
        callers that own the whole step (``runtime.engine.TrainStep``): the weights change DURING ``backward()``, so
        nothing may sit between ``backward()`` and ``step()`` (no gradient clipping, no inspection of ``param.grad``).
        Returns True if at least one arena was switched."""
        import os
        if os.environ.get("B200DDP_FUSED_SGD", "1") in ("0", "false", "off"):
            enabled = False
        g = self.param_groups[0]
        any_on = False
        for o in self._flat_owners:
            if not (hasattr(o, "can_fuse_sgd") and o.can_fuse_sgd()):
                continue
            if enabled:
                m = self._momentum_flat.get(id(o))
                if m is None and g["momentum"] != 0:
                    m = torch.zeros_like(o.flat_params)
                    self._momentum_flat[id(o)] = m
                o.set_fused_sgd(True, m, g["lr"], g["momentum"], g["dampening"], g["weight_decay"], g["nesterov"])
                self._fused_key[id(o)] = (g["lr"], g["momentum"], g["dampening"], g["weight_decay"], g["nesterov"])
                any_on = True
            elif id(o) in self._fused_key:
                o.set_fused_sgd(False)
                del self._fused_key[id(o)]
        return any_on

    @torch.no_grad()
    def step(self, closure=None):
        g = self.param_groups[0]
        for o in self._flat_owners:
            key = self._fused_key.get(id(o))
            if key is not None and o.require_backward_grad_sync:
                # the collective already applied this step (and zeroed the gradients); keep its hyper-parameters current
                now = (g["lr"], g["momentum"], g["dampening"], g["weight_decay"], g["nesterov"])
                if now != key:
                    o.set_fused_sgd(True, self._momentum_flat.get(id(o)), *now)
                    self._fused_key[id(o)] = now
                o._arena_dirty = False
                continue
            m = self._momentum_flat.get(id(o))
            if m is None and g["momentum"] != 0:
                m = torch.zeros_like(o.flat_params)
                self._momentum_flat[id(o)] = m
            nv.sgd_step(o.flat_params, o.flat_grads, m, o.flat_shadow, g["lr"], g["momentum"], g["dampening"],
                        g["weight_decay"], g["nesterov"], 1.0, True)
            o._arena_dirty = False                   # the kernel zeroed the arena for the next backward
        if self._inner is not None:
            for k in ("lr", "momentum", "dampening", "weight_decay", "nesterov"):
                self._inner.param_groups[0][k] = g[k]
            self._inner.step()

    def flat_state(self):
        return {i: m for i, m in enumerate(self._momentum_flat.values())}

    @torch.no_grad()
    def sync_state(self):
        """Collective (every rank calls it): make each rank's momentum buffer complete.  With the optimizer folded into the
        bucket collectives a rank only ever updates the momentum of the shards it owns -- the rest of its buffer still holds
        the initial zeros -- so the full state is the SUM over ranks.  Call before writing a checkpoint meant for resume
        (``workloads/cifar10_train.py`` does); afterwards every rank holds the whole state and training can go on as it was
        (the fused kernel reads and writes its own shards only)."""
        from ..parallel import process_group as pg
        for o in self._flat_owners:
            if id(o) in self._fused_key and o.world_size > 1:
                m = self._momentum_flat.get(id(o))
                if m is not None:
                    # a shard that ANOTHER rank owns has not been written here since the last sync: zero it first, so that a
                    # second sync_state() does not add the same shards twice
                    self._zero_unowned(o, m)
                    pg.all_reduce(m)

    def _zero_unowned(self, o, m):
        world, rank = o.world_size, o.rank
        for b in o.buckets:
            chunks = b.numel // 4
            per = (chunks + world - 1) // world
            lo, hi = min(rank * per, chunks) * 4, min((rank + 1) * per, chunks) * 4
            seg = m[b.start: b.start + b.numel]
            seg[:lo].zero_()
            seg[hi:].zero_()


class Adam(_FusedBase):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        if len(self.param_groups) != 1:
            raise ValueError("b200ddp.optim.Adam takes a single parameter group")
        self._split()
        self._inner = torch.optim.Adam(self._loose, **defaults) if self._loose else None
        self._state_flat = {}

    @torch.no_grad()
    def step(self, closure=None):
        g = self.param_groups[0]
        for o in self._flat_owners:
            st = self._state_flat.get(id(o))
            if st is None:
                st = (torch.zeros_like(o.flat_params), torch.zeros_like(o.flat_params),
                      torch.zeros((), dtype=torch.int32, device=o.flat_params.device))
                self._state_flat[id(o)] = st
            nv.adam_step(o.flat_params, o.flat_grads, st[0], st[1], o.flat_shadow, g["lr"], g["betas"][0],
                         g["betas"][1], g["eps"], g["weight_decay"], st[2], 1.0, True)
            o._arena_dirty = False
        if self._inner is not None:
            for k in ("lr", "betas", "eps", "weight_decay"):
                self._inner.param_groups[0][k] = g[k]
            self._inner.step()
