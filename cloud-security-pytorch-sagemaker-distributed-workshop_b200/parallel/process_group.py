"""Process-group facade: ``init_process_group`` / ``get_rank`` / ``all_reduce`` / ...

Signature-compatible subset of ``torch.distributed`` as used by the reference
(``notebooks/code/cifar10-distributed-native-cpu.py:102-109``,
``cifar10-distributed-smddp-gpu.py:17-23``; every call site in SURVEY §2.4).

Backends
--------
``"b200"`` (aliases ``"smddp"``, ``"smddp-b200"``)
    One process per GPU.  torch.distributed is used for the *control plane only*
    (TCPStore rendezvous + a gloo group for object exchange).  The data plane is this
    package's own symmetric-memory runtime: a flat buffer per rank mapped into every
    peer over NVLink (``cuMem*`` + POSIX-fd export), an NVSwitch multicast mapping of the
    same pages, and hand-written sm_100a collectives (``csrc/comm.cu``).  No NCCL call.
``"gloo"``
    CPU fallback (BASELINE.json config 1): stock ``ProcessGroupGloo``; same wrapper API.
``"nccl"``
    Only for the baseline arm / comparisons.  Never used by the product path.
"""
from __future__ import annotations

import datetime
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..utils.logging import get_logger

log = get_logger(__name__)

B200_ALIASES = ("b200", "smddp", "smddp-b200")


class ReduceOp:
    """Mirror of ``dist.ReduceOp`` (the reference uses the deprecated ``dist.reduce_op.SUM``)."""
    SUM = "sum"
    AVG = "avg"
    MAX = "max"
    MIN = "min"


reduce_op = ReduceOp  # deprecated alias kept because the reference spells it this way

_TORCH_OPS = {ReduceOp.SUM: dist.ReduceOp.SUM, ReduceOp.MAX: dist.ReduceOp.MAX,
              ReduceOp.MIN: dist.ReduceOp.MIN, ReduceOp.AVG: dist.ReduceOp.SUM}


class _State:
    backend: Optional[str] = None        # 'b200' | 'gloo' | 'nccl' | None
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    comm = None                          # parallel.symm.SymmComm for 'b200'
    owns_torch_pg: bool = False
    control_group = None


_S = _State()


def _normalize_op(op) -> str:
    if isinstance(op, str):
        return op
    if op is None:
        return ReduceOp.SUM
    # accept torch's enum too
    name = str(op).split(".")[-1].lower()
    return {"sum": "sum", "avg": "avg", "max": "max", "min": "min"}.get(name, "sum")


def is_initialized() -> bool:
    return _S.backend is not None


def get_backend() -> Optional[str]:
    return _S.backend


def get_rank() -> int:
    return _S.rank if is_initialized() else 0


def get_world_size() -> int:
    return _S.world_size if is_initialized() else 1


def get_local_rank() -> int:
    return _S.local_rank


def get_device() -> torch.device:
    return _S.device


def get_comm():
    """The symmetric-memory communicator (``None`` on gloo/nccl or world_size 1 CPU)."""
    return _S.comm


def control_group():
    return _S.control_group


def init_process_group(backend: Optional[str] = None, init_method: Optional[str] = None,
                       rank: int = -1, world_size: int = -1,
                       timeout: Optional[datetime.timedelta] = None,
                       arena_bytes: Optional[int] = None) -> None:
    """Initialise the default group.

    ``rank``/``world_size`` default to ``RANK``/``WORLD_SIZE`` (env:// rendezvous through
    ``MASTER_ADDR``/``MASTER_PORT``), as in the reference.  A single process with no env set
    initialises a world of size 1 without any rendezvous.
    """
    if is_initialized():
        raise RuntimeError("b200ddp.init_process_group called twice")
    if backend is None:
        backend = "b200" if torch.cuda.is_available() else "gloo"
    backend = backend.lower()
    if rank < 0:
        rank = int(os.environ.get("RANK", "0"))
    if world_size < 0:
        world_size = int(os.environ.get("WORLD_SIZE", "1"))
    timeout = timeout or datetime.timedelta(seconds=int(os.environ.get("B200DDP_TIMEOUT_S", "600")))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))

    if backend in B200_ALIASES:
        if not torch.cuda.is_available():
            raise RuntimeError(
                "backend 'b200' needs a CUDA device; use backend='gloo' for the CPU fallback")
        kind = "b200"
    elif backend in ("gloo", "nccl"):
        kind = backend
    else:
        raise ValueError(f"unknown backend {backend!r} (expected b200|smddp|gloo|nccl)")

    _S.rank, _S.world_size, _S.local_rank = rank, world_size, local_rank
    if kind in ("b200", "nccl"):
        ndev = torch.cuda.device_count()
        _S.device = torch.device("cuda", local_rank % max(ndev, 1))
        torch.cuda.set_device(_S.device)
    else:
        _S.device = torch.device("cpu")

    need_torch_pg = world_size > 1 or kind == "nccl" or "MASTER_ADDR" in os.environ
    if need_torch_pg and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        torch_backend = "nccl" if kind == "nccl" else "gloo"   # b200: gloo is control plane only
        kwargs = dict(backend=torch_backend, rank=rank, world_size=world_size, timeout=timeout)
        if init_method:
            kwargs["init_method"] = init_method
        if kind == "nccl":
            kwargs["device_id"] = _S.device
        dist.init_process_group(**kwargs)
        _S.owns_torch_pg = True
    _S.control_group = dist.group.WORLD if dist.is_initialized() else None

    _S.backend = kind
    if kind == "b200":
        from .symm import SymmComm
        _S.comm = SymmComm(rank=rank, world_size=world_size, device=_S.device, arena_bytes=arena_bytes)
    log.info("Initialized the distributed environment: '%s' backend on %d nodes. ", backend, world_size)


def destroy_process_group() -> None:
    """Tear down: unmap peer memory, release the multicast object, destroy the control group
    (the reference defines ``cleanup()`` and never calls it — SURVEY §5.3)."""
    if _S.comm is not None:
        _S.comm.close()
        _S.comm = None
    if _S.owns_torch_pg and dist.is_initialized():
        dist.destroy_process_group()
    _S.backend = None
    _S.owns_torch_pg = False
    _S.control_group = None
    _S.rank, _S.world_size = 0, 1


def barrier() -> None:
    if get_world_size() == 1:
        if _S.device.type == "cuda":
            torch.cuda.synchronize(_S.device)
        return
    if _S.backend == "b200":
        _S.comm.barrier()
    else:
        dist.barrier()


def all_reduce(tensor: torch.Tensor, op=ReduceOp.SUM, async_op: bool = False):
    """In-place all-reduce of ``tensor`` across the world (``…native-cpu.py:91``)."""
    op = _normalize_op(op)
    if get_world_size() == 1:
        return None
    if _S.backend == "b200" and tensor.is_cuda:
        _S.comm.all_reduce_(tensor, op=op)
        return None
    work = dist.all_reduce(tensor, op=_TORCH_OPS[op], async_op=async_op)
    if op == ReduceOp.AVG and not async_op:
        tensor.div_(get_world_size())
    return work


def broadcast(tensor: torch.Tensor, src: int = 0, async_op: bool = False):
    if get_world_size() == 1:
        return None
    if _S.backend == "b200" and tensor.is_cuda:
        _S.comm.broadcast_(tensor, src=src)
        return None
    return dist.broadcast(tensor, src=src, async_op=async_op)


def all_gather_object(obj):
    if get_world_size() == 1:
        return [obj]
    out = [None] * get_world_size()
    dist.all_gather_object(out, obj)
    return out


def broadcast_object(obj, src: int = 0):
    if get_world_size() == 1:
        return obj
    box = [obj]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def max_over_ranks(value: float) -> float:
    """Max of a host scalar over ranks — every multi-GPU timing in this repo goes through it."""
    if get_world_size() == 1:
        return float(value)
    return max(float(v) for v in all_gather_object(float(value)))
