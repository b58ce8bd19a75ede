"""``import smdistributed.dataparallel.torch.torch_smddp`` — registers the ``'smddp'`` backend, served by b200ddp.

What the reference does with SMDDP (``notebooks/code/cifar10-distributed-smddp-gpu.py``):

    :17   import smdistributed.dataparallel.torch.torch_smddp       # side effect: the 'smddp' backend exists
    :23   dist.init_process_group(backend='smddp')                  # at import time of the training script
    :148  model = torch.nn.parallel.DistributedDataParallel(model)  # torch's DDP class, on the smddp backend
    :151  torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    :116,123,183  dist.get_world_size() / dist.get_rank()

Importing this module makes those lines land on b200ddp, with the script byte-for-byte unmodified:

* ``torch.distributed.init_process_group(backend='smddp')`` initialises ``b200ddp`` with ``backend='b200'``: the
  symmetric-memory NVLink runtime for the data plane, a gloo group of ``torch.distributed`` as control plane -- which is
  also what makes the script's ``dist.get_world_size()`` / ``dist.get_rank()`` work unchanged.  Any other backend string is
  passed through to torch untouched.
* ``torch.nn.parallel.DistributedDataParallel(model)`` returns ``b200ddp.DistributedDataParallel`` of the model converted to
  the native layers (``b200ddp.utils.convert.to_native``: same parameters, same ``state_dict`` keys) whenever the process
  group is the b200 one; otherwise torch's own class is constructed.

The script's own loop then runs on this framework: its ``optim.SGD`` steps the flat parameter arena through the views,
``loss.backward()`` produces gradients with the tensor-core kernels and all-reduces the buckets over NVLink.  After the
first full-size training batch ``model(data)`` and ``loss.backward()`` are one CUDA-graph replay each (forward; backward
with the bucket collectives), everything else (evaluation, the last short batch) runs launch by launch.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn.parallel as _tnp

import b200ddp as _bd
from b200ddp.utils.convert import to_native as _to_native

_TORCH_INIT = dist.init_process_group
_TORCH_DDP = _tnp.DistributedDataParallel
_TORCH_DESTROY = dist.destroy_process_group


def _init_process_group(backend=None, *args, **kwargs):
    name = str(backend).lower() if backend is not None else None
    if name != "smddp":
        return _TORCH_INIT(backend, *args, **kwargs)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")      # a world of one still gets a (gloo) default group: the
    os.environ.setdefault("MASTER_PORT", "29500")          # script calls dist.get_world_size() unconditionally
    rank = kwargs.get("rank", -1)
    world = kwargs.get("world_size", -1)
    # the reference trains in fp32 with no autocast (…smddp-gpu.py:145): keep its precision unless told otherwise
    _bd.set_compute_dtype(os.environ.get("B200DDP_DTYPE", "fp32"))
    _bd.init_process_group(backend="b200", init_method=kwargs.get("init_method"), rank=rank if rank is not None else -1,
                           world_size=world if world is not None else -1, timeout=kwargs.get("timeout"))


_IN_DESTROY = [False]


def _destroy_process_group(group=None):
    # b200ddp's own teardown ends in dist.destroy_process_group(), i.e. back in here: the nested call goes to torch's
    if group is None and _bd.is_initialized() and not _IN_DESTROY[0]:
        _IN_DESTROY[0] = True
        try:
            _bd.destroy_process_group()
        finally:
            _IN_DESTROY[0] = False
        if not dist.is_initialized():
            return
    return _TORCH_DESTROY(group)


class _DDPMeta(type(_TORCH_DDP)):
    """``isinstance(x, torch.nn.parallel.DistributedDataParallel)`` keeps working for both classes."""

    def __instancecheck__(cls, obj):
        return isinstance(obj, (_TORCH_DDP, _bd.DistributedDataParallel))


class DistributedDataParallel(_TORCH_DDP, metaclass=_DDPMeta):
    """Constructor shim: b200ddp's DDP over the natively converted module on the b200 backend, torch's otherwise."""

    def __new__(cls, module, *args, **kwargs):
        if _bd.is_initialized() and _bd.get_backend() == "b200":
            allowed = ("device_ids", "output_device", "dim", "broadcast_buffers", "process_group", "bucket_cap_mb",
                       "find_unused_parameters", "gradient_as_bucket_view", "static_graph")
            kw = {k: v for k, v in kwargs.items() if k in allowed}
            for k, v in zip(allowed, args):
                kw.setdefault(k, v)
            # the script owns the loop (loss, optimizer): take what lies between model(data) and the gradients off the host
            # as two CUDA-graph replays (b200ddp/runtime/graphed.py); B200DDP_SHIM_GRAPH=0 keeps every launch eager
            kw["graph_forward_backward"] = os.environ.get("B200DDP_SHIM_GRAPH", "1") not in ("0", "false", "off")
            return _bd.DistributedDataParallel(_to_native(module), **kw)
        return _TORCH_DDP(module, *args, **kwargs)


def install() -> None:
    dist.init_process_group = _init_process_group
    dist.destroy_process_group = _destroy_process_group
    _tnp.DistributedDataParallel = DistributedDataParallel
    import torch.nn.parallel.distributed as _tnpd
    _tnpd.DistributedDataParallel = DistributedDataParallel


def uninstall() -> None:
    dist.init_process_group = _TORCH_INIT
    dist.destroy_process_group = _TORCH_DESTROY
    _tnp.DistributedDataParallel = _TORCH_DDP
    import torch.nn.parallel.distributed as _tnpd
    _tnpd.DistributedDataParallel = _TORCH_DDP


install()
