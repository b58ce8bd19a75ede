"""Gradient-synchronisation check shared by ``bench.py`` (printed as ``config.ddp_check``), the multi-GPU
tests and ``scripts/ddp_check.py``.

The invariant is the one the reference spells out by hand in ``_average_gradients``
(``notebooks/code/cifar10-distributed-native-cpu.py:87-92``): after a synchronised backward every rank
holds the MEAN over ranks of the gradients each rank computed on its own shard.  Procedure: one backward
under ``no_sync()`` gives the local gradients, which are gathered through the gloo control plane (so the
oracle never touches the NVLink kernels under test); the same batch is then run with synchronisation on
and the arena is compared with the mean, and across ranks bit for bit.

The two passes have to be reproducible for the comparison to mean anything, so the atomically-ordered
reductions (cluster split-K joins, BN statistics in the conv epilogue) are pinned for its duration
(``profiles/determinism.md``).
"""
from __future__ import annotations

from typing import Callable, Dict

import torch
import torch.distributed as dist


def _gather_cpu(t: torch.Tensor):
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, t.detach().float().cpu())
    return out


def grad_sync_check(model, fwd_bwd: Callable[[], None]) -> Dict:
    """``model``: a ``b200ddp.DistributedDataParallel``; ``fwd_bwd()``: one forward + ``loss.backward()`` on this
    rank's own (rank-specific) batch.  Collective: every rank must call it.  Leaves the gradient arena zeroed."""
    from ..nn import modules as bmod
    from ..ops import native as nv
    world = model.world_size
    if world == 1 or not dist.is_initialized():
        return {"world": world, "skipped": "single rank"}
    cuda = model.device.type == "cuda"
    split_k = fuse = None
    if cuda:
        split_k, nv.SPLIT_K[0] = nv.SPLIT_K[0], False
        fuse, bmod._FUSE_BN_STATS[0] = bmod._FUSE_BN_STATS[0], False
    try:
        model.zero_grad()
        with model.no_sync():
            fwd_bwd()
        if cuda:
            torch.cuda.synchronize()
        local = _gather_cpu(model.flat_grads)
        want = torch.stack(local).mean(0)
        model.zero_grad()
        fwd_bwd()
        if cuda:
            torch.cuda.synchronize()
        got = _gather_cpu(model.flat_grads)
    finally:
        if cuda:
            nv.SPLIT_K[0] = split_k
            bmod._FUSE_BN_STATS[0] = fuse
    model.zero_grad()
    bitwise = all(torch.equal(got[0], t) for t in got[1:])
    denom = float(want.norm()) + 1e-30
    rel = float((got[0] - want).norm()) / denom
    per_bucket = []
    for b in model.buckets:
        sl = slice(b.start, b.start + b.numel)
        per_bucket.append(float((got[0][sl] - want[sl]).norm()) / (float(want[sl].norm()) + 1e-30))
    return {"world": world, "rel_err": rel, "bitwise_equal": bool(bitwise), "per_bucket_rel_err": per_bucket,
            "grad_norm": denom}


def params_identical(model) -> bool:
    """True when every rank holds bit-identical master parameters (collective)."""
    if model.world_size == 1 or not dist.is_initialized():
        return True
    ps = _gather_cpu(model.flat_params)
    return all(torch.equal(ps[0], p) for p in ps[1:])
