"""``DistributedDataParallel`` — drop-in for ``torch.nn.parallel.DistributedDataParallel`` as the
reference uses it (``notebooks/code/cifar10-distributed-native-cpu.py:141``,
``cifar10-distributed-smddp-gpu.py:148``, ``train_basic_jumbo_distributed_cpu.py:61``,
``train_basic_trojaned_distributed_cpu.py:64``).

Semantics kept: rank-0 parameter+buffer broadcast at wrap time, gradient averaging over the world
overlapped with backward, per-forward buffer (BN statistics) broadcast, ``no_sync()``, the
``module.`` prefix on ``state_dict()`` keys, ``device_ids``/``output_device`` accepted (one device per
process).

Memory model (B200-first, laid out for one 180 GB HBM3e per rank):

* ``flat_params``  fp32 master weights, one arena; ``param.data`` are views.  Conv weights are stored
  in channels-last physical order (``[Cout][R][S][Cin]``) which is exactly what the implicit-GEMM
  kernels want, while ``state_dict()`` still shows torch-shaped ``[Cout,Cin,R,S]`` tensors.
* ``flat_grads``   fp32 gradient arena in *symmetric memory* (peer-mapped + NVLS multicast for
  ``backend='b200'``); ``param.grad`` are views; buckets are byte ranges; native layers write their
  weight gradient straight into their slot (no autograd accumulation kernel).
* ``flat_shadow``  bf16 copy of ``flat_params`` maintained by the fused optimizer for the tensor-core
  kernels.
* ``flat_buffers`` every module buffer packed into one byte arena so the per-forward BN-stat sync
  is a single broadcast.
"""
from __future__ import annotations

import contextlib
import weakref
from typing import List, Optional

import torch
import torch.nn as nn

from . import process_group as pg
from .reducer import ALIGN_ELEMS, NativeReducer, Reducer, assign_buckets
from ..utils.config import CommConfig
from ..utils.logging import get_logger

log = get_logger(__name__)


def _is_conv_weight(p: torch.Tensor) -> bool:
    return p.dim() == 4


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0,
                 broadcast_buffers: bool = True, process_group=None, bucket_cap_mb: Optional[float] = None,
                 find_unused_parameters: bool = False, gradient_as_bucket_view: bool = True,
                 static_graph: bool = True, direct_grad_sink: bool = True, shadow_dtype="auto",
                 graph_forward_backward: Optional[bool] = None):
        super().__init__()
        self.module = module
        self.broadcast_buffers = broadcast_buffers
        self.world_size = pg.get_world_size()
        self.rank = pg.get_rank()
        self.backend = pg.get_backend() or "local"
        self.comm = pg.get_comm()
        self.cfg = CommConfig()
        if bucket_cap_mb is not None:
            self.cfg.bucket_cap_mb = float(bucket_cap_mb)
        self.require_backward_grad_sync = True
        self._accum_count = 0
        self._callback_queued = False
        self.direct_grad_sink = direct_grad_sink

        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel: module has no parameter that requires grad")
        self.device = params[0].device
        if self.backend == "b200" and self.device.type != "cuda":
            raise RuntimeError("backend 'b200' needs the module on a CUDA device (call .to(device) first)")
        for p in params:
            if p.dtype != torch.float32:
                raise RuntimeError("DistributedDataParallel keeps fp32 master parameters; got %s" % p.dtype)

        # ---- arena layout: reverse registration order == the order backward produces grads
        self.params: List[nn.Parameter] = list(reversed(params))
        numels = [p.numel() for p in self.params]
        offsets, buckets, total = assign_buckets(
            numels, int(self.cfg.first_bucket_mb * 1024 * 1024), int(self.cfg.bucket_cap_mb * 1024 * 1024))
        self.offsets, self.buckets, self.total_elems = offsets, buckets, total
        param_bucket = [0] * len(self.params)
        for b in buckets:
            for i in b.param_ids:
                param_bucket[i] = b.index

        # ---- flat arenas
        grad_off_bytes = 0
        self._param_off_bytes = self._shadow_off_bytes = None      # arena offsets when the masters / shadow are symmetric too
        symmetric = self.backend == "b200" and self.comm is not None
        if symmetric and self.world_size > 1:
            # masters in symmetric memory as well: the fused bucket step (reduce-scatter -> owner-shard SGD -> all-gather of
            # the weights, csrc/comm.cu) writes every rank's parameters through the multicast / peer mappings
            self.flat_params, self._param_off_bytes = self.comm.alloc_arena(total, torch.float32)
        else:
            self.flat_params = torch.zeros(total, dtype=torch.float32, device=self.device)
        if symmetric:
            self.flat_grads, grad_off_bytes = self.comm.alloc_arena(total, torch.float32)
        else:
            self.flat_grads = torch.zeros(total, dtype=torch.float32, device=self.device)
        self._grad_off_bytes = grad_off_bytes
        self.flat_shadow = None
        if shadow_dtype == "auto":
            # the bf16 shadow feeds the bf16 tensor-core kernels; in fp32 compute mode they read the masters themselves
            from ..ops.functional import act_dtype
            shadow_dtype = torch.bfloat16 if act_dtype() == torch.bfloat16 else None
        self._arena_dirty = False        # gradients have been written since the arena was last zeroed
        if self.device.type == "cuda" and shadow_dtype is not None:
            if symmetric and self.world_size > 1 and shadow_dtype == torch.bfloat16:
                self.flat_shadow, self._shadow_off_bytes = self.comm.alloc_arena(total, torch.bfloat16)
            else:
                self.flat_shadow = torch.zeros(total, dtype=shadow_dtype, device=self.device)

        # channels-last physical storage only for the weights of this package's conv modules on a GPU: a
        # stock nn.Conv2d fed channels-last weights would hand channels-last activations to user code
        # (the reference's `x.view(-1, 400)` then fails)
        from ..nn.modules import Conv2d as _NativeConv
        native_conv_w = {id(m.weight) for m in module.modules() if isinstance(m, _NativeConv)}
        with torch.no_grad():
            for i, p in enumerate(self.params):
                off, n = offsets[i], p.numel()
                cl = _is_conv_weight(p) and id(p) in native_conv_w and self.device.type == "cuda"
                src = p.detach()
                if cl:   # channels-last physical layout, torch-shaped logical view
                    co, ci, r, s = p.shape
                    slot = self.flat_params[off: off + n].view(co, r, s, ci).permute(0, 3, 1, 2)
                    gslot = self.flat_grads[off: off + n].view(co, r, s, ci).permute(0, 3, 1, 2)
                else:
                    slot = self.flat_params[off: off + n].view(p.shape)
                    gslot = self.flat_grads[off: off + n].view(p.shape)
                slot.copy_(src)
                p.data = slot
                p.grad = gslot
                p._b200_slot = (off, n)                      # arena slot (elements)
                p._b200_owner = weakref.ref(self)
                if self.flat_shadow is not None:
                    sh = self.flat_shadow[off: off + n]
                    p._b200_shadow = sh.view(co, r, s, ci) if cl else sh.view(p.shape)
                if direct_grad_sink:
                    p._b200_sink = _GradSink(self, i, self.flat_grads[off: off + n])
        self._grad_views = [p.grad for p in self.params]
        self.refresh_shadow()

        # ---- buffers: one byte arena so the per-forward sync is one broadcast
        self._setup_buffers()

        if self.backend == "b200" and self.comm is not None:
            self.reducer = NativeReducer(self.flat_grads, buckets, param_bucket, self.world_size, self.comm,
                                         arena_offset_bytes=grad_off_bytes)
        else:
            self.reducer = Reducer(self.flat_grads, buckets, param_bucket, self.world_size,
                                   self.backend, comm=self.comm, arena_offset_bytes=grad_off_bytes)
        for i, p in enumerate(self.params):
            p.register_post_accumulate_grad_hook(self._make_hook(i))

        # ---- wrap-time sync from rank 0 (params + buffers), like the reference's DDP ctor
        if self.world_size > 1:
            pg.broadcast(self.flat_params, src=0)
            if self.flat_buffers is not None:
                pg.broadcast(self.flat_buffers, src=0)
            self.refresh_shadow()
        # ---- optional: model(x) / loss.backward() of a caller-owned loop as two CUDA-graph replays (runtime/graphed.py)
        if graph_forward_backward is None:
            import os
            graph_forward_backward = os.environ.get("B200DDP_GRAPH_FWD_BWD", "0") not in ("0", "false", "off")
        self._graphed = None
        if graph_forward_backward and self.device.type == "cuda":
            from ..runtime.graphed import GraphedForwardBackward
            self._graphed = GraphedForwardBackward(self)
        self._param_versions = self._versions()
        if self.cfg.debug:
            sizes = ", ".join("%d B (%d tensors)" % (sum(numels[i] for i in b.param_ids) * 4, len(b.param_ids))
                              for b in buckets)
            log.debug("b200ddp.DDP: %d params, %d elements, %d bucket(s): %s", len(self.params),
                      sum(numels), len(buckets), sizes)

    # ------------------------------------------------------------------ buffers
    def _setup_buffers(self):
        bufs = [b for b in self.module.buffers()]
        self._buffers_list = bufs
        self.flat_buffers = None
        if not bufs:
            return
        offs, cur = [], 0
        for b in bufs:
            offs.append(cur)
            cur += (b.numel() * b.element_size() + 255) // 256 * 256
        if self.backend == "b200" and self.comm is not None and self.world_size > 1:
            flat, _ = self.comm.alloc_arena(cur, torch.uint8)
        else:
            flat = torch.zeros(cur, dtype=torch.uint8, device=self.device)
        with torch.no_grad():
            for b, off in zip(bufs, offs):
                nbytes = b.numel() * b.element_size()
                view = flat[off: off + nbytes].view(b.dtype).view(b.shape)
                view.copy_(b)
                b.data = view
        self.flat_buffers = flat

    # ------------------------------------------------------------------ hooks
    def _make_hook(self, idx: int):
        def hook(param):
            view = self._grad_views[idx]
            g = param.grad
            if g is not None and g.data_ptr() != view.data_ptr():
                # someone dropped our view (zero_grad(set_to_none=True)): copy in and re-point
                if self._accum_count > 0 or self.reducer.was_written(idx):
                    view.add_(g)
                else:
                    view.copy_(g)
                param.grad = view
            self._on_grad_ready(idx)
        return hook

    def _on_grad_ready(self, idx: int):
        if not self._callback_queued:
            self._callback_queued = True
            self._arena_dirty = True
            self.reducer.prepare_for_backward()
            self.reducer.require_sync = self.require_backward_grad_sync
            from torch.autograd import Variable
            Variable._execution_engine.queue_callback(self._finalize_backward)
        self.reducer.mark_ready(idx)

    def _finalize_backward(self):
        self._callback_queued = False
        if self.device.type == "cuda":
            from ..ops import streams
            streams.join(self.device)      # wgrad kernels forked onto the aux stream
        ev = getattr(self, "_phase_events", None)
        if ev is not None:                 # TrainStep.breakdown(): where the compute of backward ends and the join begins
            ev["bwd_compute_end"].record()
        self.reducer.finalize()
        if ev is not None:
            ev["comm_joined"].record()
            ev["hit"] = True
        if self.require_backward_grad_sync:
            self._accum_count = 0
        else:
            self._accum_count += 1

    # ------------------------------------------------------------------ module API
    def _adopt_dropped_grads(self):
        """``torch.optim``'s ``zero_grad()`` defaults to ``set_to_none=True``: ``param.grad`` becomes ``None``.  Native layers
        deposit their weight gradients straight into the arena and never pass through AccumulateGrad, so nothing would
        re-point ``param.grad`` and a stock optimizer would skip those parameters while the arena kept accumulating.  A
        dropped gradient means "reset": zero the arena (once) and hand every parameter its arena view back."""
        dropped = False
        for p, v in zip(self.params, self._grad_views):
            if p.grad is None:
                p.grad = v
                dropped = True
        if dropped and self._arena_dirty and self._accum_count == 0:
            self.flat_grads.zero_()
            self._arena_dirty = False

    def _versions(self) -> int:
        return sum(p._version for p in self.params) if self.flat_shadow is not None else 0

    def _sync_shadow_after_foreign_step(self):
        """The bf16 shadow weights are maintained by this package's fused optimizers.  Any other writer of the parameters
        (``torch.optim``, ``p.data.add_``, ``load_state_dict`` on the inner module) bumps their version counters: re-derive
        the shadow then, or the tensor-core kernels would keep computing with the old weights."""
        if self.flat_shadow is not None and not torch.cuda.is_current_stream_capturing():
            v = self._versions()
            if v != self._param_versions:
                self._param_versions = v
                self.refresh_shadow()

    def _forward_eager(self, *args, **kwargs):
        if (self.broadcast_buffers and self.world_size > 1 and self.flat_buffers is not None
                and self.require_backward_grad_sync and self.training):
            pg.broadcast(self.flat_buffers, src=0)       # KB12: one coalesced broadcast
        return self.module(*args, **kwargs)

    def forward(self, *args, **kwargs):
        if torch.is_grad_enabled():
            self._adopt_dropped_grads()
        self._sync_shadow_after_foreign_step()
        g = self._graphed
        if g is not None and g.wants(args, kwargs):
            return g(args[0])
        if g is not None:
            g.note_eager()
        return self._forward_eager(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def zero_grad(self, set_to_none: bool = False):   # grads are arena views: zero, never free
        self.flat_grads.zero_()
        self._accum_count = 0
        self._arena_dirty = False

    # ------------------------------------------------------------------ fused bucket step
    def can_fuse_sgd(self) -> bool:
        """True when every bucket's all-reduce can become reduce-scatter + SGD + weight all-gather in one kernel."""
        return (self.backend == "b200" and self.world_size > 1 and self._param_off_bytes is not None
                and hasattr(self.reducer, "set_sgd") and (self.flat_shadow is None or self._shadow_off_bytes is not None))

    def set_fused_sgd(self, enabled: bool, momentum_buf=None, lr=0.0, momentum=0.0, dampening=0.0, weight_decay=0.0,
                      nesterov=False):
        self.reducer.set_sgd(bool(enabled), self._grad_off_bytes, self._param_off_bytes or 0, self._shadow_off_bytes or 0,
                             momentum_buf, lr, momentum, dampening, weight_decay, nesterov)
        self._fused_sgd = bool(enabled)

    def refresh_shadow(self):
        """Re-derive the bf16 shadow weights from the fp32 masters (after load_state_dict etc.)."""
        if self.flat_shadow is not None:
            self.flat_shadow.copy_(self.flat_params)

    def grad_beta(self, idx: int) -> float:
        """0.0 -> the native layer overwrites its gradient slot, 1.0 -> it accumulates."""
        return 1.0 if (self._accum_count > 0 or self.reducer.was_written(idx)) else 0.0

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        out = super().load_state_dict(state_dict, strict=strict)
        self.refresh_shadow()
        return out


class _GradSink:
    """Handle a native layer uses to deposit its weight gradient directly in the arena."""

    __slots__ = ("ddp", "idx", "flat")

    def __init__(self, ddp: DistributedDataParallel, idx: int, flat: torch.Tensor):
        self.ddp, self.idx, self.flat = ddp, idx, flat

    def beta(self) -> float:
        return self.ddp.grad_beta(self.idx)

    def done(self):
        ddp = self.ddp
        p = ddp.params[self.idx]
        if p.grad is None:               # dropped after forward (zero_grad(set_to_none=True) between forward and backward)
            p.grad = ddp._grad_views[self.idx]
        ddp._on_grad_ready(self.idx)
