"""Gradient bucketing + overlap scheduler (the job torch's C++ ``Reducer`` does for the reference:
``notebooks/code/cifar10-distributed-native-cpu.py:141``, ``cifar10-distributed-smddp-gpu.py:148``,
job log "Reducer buckets have been rebuilt" ``notebooks/2_pytorch_dist_smddp_gpu.ipynb:1843``).

B200-first differences:

* Gradients live in ONE flat fp32 arena per rank (for ``backend='b200'`` that arena is symmetric
  memory mapped into every peer).  A bucket is a contiguous byte range of the arena, so there is
  no pack/unpack copy, and ``param.grad`` is a view of the arena.
* Buckets are laid out in reverse registration order (the order backward produces them) with the
  reference Reducer's caps — 1 MiB first bucket, 25 MiB afterwards — so the sizes match SURVEY §2.4.
* When a bucket's last gradient lands, an event is recorded on the compute stream and the fused
  allreduce(+1/world scale) kernel is launched on a dedicated high-priority comm stream; the
  optimizer waits on the last bucket's event.  Launch order is the static bucket order on every
  rank (the kernels are collective and spin on peers).
* Everything here is capturable in a CUDA graph (events + kernel launches only).
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from . import process_group as pg
from ..utils.logging import get_logger

log = get_logger(__name__)

ALIGN_ELEMS = 64          # every parameter slot starts on a 256-byte boundary (fp32)


@dataclasses.dataclass
class Bucket:
    index: int
    start: int               # element offset into the arena
    numel: int                # padded element count (multiple of ALIGN_ELEMS)
    param_ids: List[int]
    pending: int = 0
    launched: bool = False


def assign_buckets(numels: Sequence[int], first_bucket_bytes: int, bucket_cap_bytes: int,
                   elem_size: int = 4):
    """Lay ``numels`` (already in arena order) out contiguously and cut buckets by size cap.

    Returns ``(offsets, buckets, total_elems)``.  Same policy as
    ``dist._compute_bucket_assignment_by_size`` with ``[1 MiB, 25 MiB]`` limits: a bucket is closed
    once its payload reaches the current cap; the first cap applies to the first bucket only.
    """
    offsets: List[int] = []
    buckets: List[Bucket] = []
    cur = 0
    b_start, b_params, b_bytes = 0, [], 0
    cap = first_bucket_bytes
    for i, n in enumerate(numels):
        offsets.append(cur)
        padded = (n + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
        cur += padded
        b_params.append(i)
        b_bytes += n * elem_size
        if b_bytes >= cap:
            buckets.append(Bucket(len(buckets), b_start, cur - b_start, b_params))
            b_start, b_params, b_bytes = cur, [], 0
            cap = bucket_cap_bytes
    if b_params:
        buckets.append(Bucket(len(buckets), b_start, cur - b_start, b_params))
    return offsets, buckets, cur


class Reducer:
    """Per-bucket ready counting and in-order launch of the gradient allreduce."""

    def __init__(self, arena: torch.Tensor, buckets: List[Bucket], param_bucket: List[int],
                 world_size: int, backend: str, comm=None, arena_offset_bytes: int = 0):
        self.arena = arena
        self.buckets = buckets
        self.param_bucket = param_bucket
        self.world_size = world_size
        self.backend = backend
        self.comm = comm
        self.arena_offset_bytes = arena_offset_bytes   # offset of `arena` inside the symmetric buffer
        self.next_bucket = 0
        self.require_sync = True
        self.works = []                  # gloo async handles
        self.in_backward = False
        self.ready_flags = [False] * len(param_bucket)
        self.ready_order: List[int] = []
        self.iterations = 0
        self._reset()

    # ------------------------------------------------------------------ bookkeeping
    def _reset(self):
        for b in self.buckets:
            b.pending = len(b.param_ids)
            b.launched = False
        self.next_bucket = 0
        self.works = []
        for i in range(len(self.ready_flags)):
            self.ready_flags[i] = False

    def prepare_for_backward(self):
        self._reset()
        self.in_backward = True
        if self.iterations == 0:
            self.ready_order = []

    def was_written(self, param_id: int) -> bool:
        return self.ready_flags[param_id]

    def mark_ready(self, param_id: int):
        """Called once per parameter when its gradient is final for this backward."""
        if self.ready_flags[param_id]:
            return
        self.ready_flags[param_id] = True
        if self.iterations == 0:
            self.ready_order.append(param_id)
        if not self.require_sync or self.world_size == 1:
            return
        b = self.buckets[self.param_bucket[param_id]]
        b.pending -= 1
        if b.pending == 0:
            self._launch_ready()

    def _launch_ready(self):
        while self.next_bucket < len(self.buckets):
            b = self.buckets[self.next_bucket]
            if b.pending > 0:
                break
            self._launch(b)
            self.next_bucket += 1

    def _launch(self, b: Bucket):
        b.launched = True
        view = self.arena[b.start: b.start + b.numel]
        if self.backend == "b200":
            self.comm.all_reduce_arena(self.arena_offset_bytes + b.start * 4, b.numel,
                                       scale=1.0 / self.world_size, overlap=True)
        else:
            work = dist.all_reduce(view, op=dist.ReduceOp.SUM, async_op=True)
            self.works.append((work, view))

    # ------------------------------------------------------------------ end of backward
    def finalize(self):
        """Flush buckets that never filled (unused parameters contribute their zeroed slots), then
        make the compute stream wait for the comm stream / the gloo works."""
        self.in_backward = False
        if self.require_sync and self.world_size > 1:
            for b in self.buckets[self.next_bucket:]:
                self._launch(b)
            self.next_bucket = len(self.buckets)
            if self.backend == "b200":
                self.comm.join_overlap()
            else:
                inv = 1.0 / self.world_size
                for work, view in self.works:
                    work.wait()
                    view.mul_(inv)
            self.works = []
        self.iterations += 1


class NativeReducer:
    """``backend='b200'``: the counting / ordering / launching lives in the C++ reducer (``csrc/reducer.cu``);
    this class only forwards stream handles.  Same interface as ``Reducer``."""

    def __init__(self, arena: torch.Tensor, buckets: List[Bucket], param_bucket: List[int], world_size: int, comm,
                 arena_offset_bytes: int = 0):
        import ctypes
        from ctypes import byref, c_float, c_int, c_size_t, c_void_p
        from ..ops import _lib as L
        self._L, self._ct = L, ctypes
        self.arena, self.buckets, self.param_bucket = arena, buckets, param_bucket
        self.world_size, self.comm = world_size, comm
        self.require_sync = True
        self.iterations = 0
        self.ready_order: List[int] = []
        self.in_backward = False
        lib = L.lib()
        lib.b200_reducer_create.restype = c_void_p
        n, nb = len(param_bucket), len(buckets)
        pb = (c_int * n)(*param_bucket)
        offs = (c_size_t * nb)(*[arena_offset_bytes + b.start * 4 for b in buckets])
        cnts = (c_size_t * nb)(*[b.numel for b in buckets])
        cfg = comm.cfg
        self._h = lib.b200_reducer_create(byref(comm.cs), c_int(n), pb, c_int(nb), offs, cnts, c_int(L.DTYPE_F32),
                                          c_float(1.0 / world_size), c_int(cfg.overlap_ctas), c_size_t(cfg.oneshot_max_bytes),
                                          c_size_t(cfg.nvls_min_bytes if (comm.has_nvls and comm.prefer_nvls) else (1 << 62)))
        if not self._h:
            raise L.NativeError("b200_reducer_create failed: %s" % lib.b200_last_error().decode())
        self._h = c_void_p(self._h)
        self._lib = lib
        self._comm_stream = c_void_p(comm.comm_stream.cuda_stream)

    def set_sgd(self, enabled: bool, grad_base: int, param_base: int, shadow_base: int, momentum_buf, lr: float,
                momentum: float, dampening: float, weight_decay: float, nesterov: bool):
        """Make every bucket launch the fused step (``b200_allreduce_sgd``) instead of the plain all-reduce."""
        ct = self._ct
        self._momentum_keepalive = momentum_buf
        self._L.check(self._lib.b200_reducer_set_sgd(self._h, ct.c_int(int(enabled)), ct.c_size_t(grad_base), ct.c_size_t(param_base),
                                                     ct.c_size_t(shadow_base), self._L.ptr(momentum_buf), ct.c_float(lr),
                                                     ct.c_float(momentum), ct.c_float(dampening), ct.c_float(weight_decay),
                                                     ct.c_int(int(nesterov))), "reducer_set_sgd")
        self._L.LAUNCHES[0] -= 1        # bookkeeping call, not a kernel
        self._logged = False
        self.sgd_fused = bool(enabled)

    def __del__(self):
        try:
            self._lib.b200_reducer_destroy(self._h)
        except Exception:
            pass

    def prepare_for_backward(self):
        self._L.check(self._lib.b200_reducer_begin(self._h), "reducer_begin")
        self.in_backward = True
        if self.iterations == 0:
            self.ready_order = []

    def was_written(self, param_id: int) -> bool:
        return bool(self._lib.b200_reducer_was_ready(self._h, self._ct.c_int(param_id)))

    def _streams(self):
        from ..ops import streams
        dev = self.comm.device
        cur = torch.cuda.current_stream(dev)
        others = streams.producers(dev)
        other = self._ct.c_void_p(others[0].cuda_stream) if others else self._ct.c_void_p(0)
        return self._ct.c_void_p(cur.cuda_stream), other

    def mark_ready(self, param_id: int):
        if self.iterations == 0:
            self.ready_order.append(param_id)
        if not self.require_sync or self.world_size == 1:
            return
        cur, other = self._streams()
        rc = self._lib.b200_reducer_mark_ready(self._h, self._ct.c_int(param_id), cur, other, self._comm_stream)
        if rc < 0:
            self._L.check(rc, "reducer_mark_ready")
        for _ in range(rc):
            self._L.LAUNCHES[0] += 1

    def finalize(self):
        self.in_backward = False
        cur, other = self._streams()
        self._L.check(self._lib.b200_reducer_finalize(self._h, cur, other, self._comm_stream,
                                                      self._ct.c_int(int(self.require_sync))), "reducer_finalize")
        if not getattr(self, "_logged", False) and self.world_size > 1 and self.require_sync:
            self._logged = True
            for b in self.buckets:
                a = self._lib.b200_reducer_bucket_algo(self._h, self._ct.c_int(b.index))
                name = self._L.ALGO_NAMES.get(a, "?")
                self.comm.launch_log.append((("fused_sgd_" + name) if getattr(self, "sgd_fused", False) else name, b.numel * 4))
        self.iterations += 1
