"""Forward and backward of a ``DistributedDataParallel`` module as two CUDA graphs behind ONE autograd node.

``TrainStep`` (``runtime/engine.py``) owns a whole optimisation step and replays it as one graph.  A user loop that is
written the way the reference's is (``notebooks/code/cifar10-distributed-smddp-gpu.py:162-168``)

    optimizer.zero_grad(); output = model(data); loss = criterion(output, target); loss.backward(); optimizer.step()

cannot be handed over like that: the loss and the optimizer are the caller's.  What CAN be taken off the host is everything
between ``model(data)`` and the gradients: ~60 Python autograd functions and ~130 kernel launches per resnet18 step, which on
the eager path cost 5-9 ms of host time for 0.8-1.3 ms of GPU work.  ``GraphedForwardBackward`` captures

    graph F:  [BN-buffer broadcast]  y = module(x)                                      (autograd recording)
    graph B:  the backward from gy -> dgrad chain, wgrads on the auxiliary stream straight into the gradient arena, the
              bucket collectives on the comm stream, the join

from static input / output / output-gradient buffers in one private memory pool, and ``model(data)`` becomes: copy ``data`` in,
replay F, return ``y``; ``loss.backward()`` reaches the node's backward: copy ``gy`` in, replay B.  ``param.grad`` are the
arena views the graph writes, so any optimizer (the script's ``torch.optim.SGD``) steps as before.

Everything that does not fit the captured shape runs eagerly, as before: evaluation / ``no_grad``, another batch size (the
last short batch of an epoch), ``no_sync()``, keyword arguments, a changed train/eval pattern of the sub-modules, a caller
that is itself capturing, a module with a layer whose parameter gradients come through autograd instead of the arena sinks
(noticed in the eager warm-up, before anything is captured).  One output may be back-propagated once, and only until the next
training forward (the activations the backward needs live in the graph's pool and are overwritten by the next replay) -- a
stale backward raises.

Measured with the reference's unmodified GPU script on this framework (``bench.py --impl reference_on_b200``, the script's own
PIL loader in the timed region): 8.9k / 17.3k / 31.1k / 53.4k img/s at 1 / 2 / 4 / 8 GPUs against 7.0k / 11.3k / 16.4k / 18.2k
with every launch eager and 7.6k / 12.1k / 18.0k / 25.3k for the reference itself (``BASELINE.md``).
"""
from __future__ import annotations

import torch

from ..utils.logging import get_logger

log = get_logger(__name__)

# process-wide counters (bench.py --impl reference_on_b200 reports them: did the script's loop really run on replays?)
STATS = {"captures": 0, "capture_failures": 0, "forward_replays": 0, "backward_replays": 0, "eager_forwards": 0}


class _NotCapturable(RuntimeError):
    """Raised during the (eager) warm-up: nothing has been captured yet, falling back is free of side effects."""


class _Replay(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, gfb, x):
        gfb.static_x.copy_(x, non_blocking=True)
        gfb.graph_f.replay()
        gfb.note_launches(gfb.launches_f)
        STATS["forward_replays"] += 1
        gfb.generation += 1
        ctx.gfb, ctx.generation = gfb, gfb.generation
        # a copy: the caller may keep the output (metrics) past the next replay
        return gfb.static_y.clone()

    @staticmethod
    def backward(ctx, gy):
        gfb = ctx.gfb
        if ctx.generation != gfb.generation or gfb.consumed == ctx.generation:
            raise RuntimeError("b200ddp graphed forward/backward: this output's activations have been replaced by a later "
                               "forward (or were already back-propagated); construct DistributedDataParallel with "
                               "graph_forward_backward=False for this pattern")
        gfb.consumed = ctx.generation
        gfb.static_gy.copy_(gy, non_blocking=True)
        gfb.graph_b.replay()
        gfb.note_launches(gfb.launches_b)
        STATS["backward_replays"] += 1
        gfb.after_backward()
        return None, None, None


class GraphedForwardBackward:
    def __init__(self, ddp, warmup_iters: int = 3):
        self.ddp = ddp
        self.warmup_iters = int(warmup_iters)
        self.enabled = True                    # False for good once a capture has failed
        self.paused = False                    # a caller may route single calls to the eager path
        self.failure = None                    # traceback of the failed capture, if any
        self.graph_f = self.graph_b = None
        self.key = None                        # (shape, dtype) of the captured input
        self.static_x = self.static_y = self.static_gy = None
        self.launches_f = self.launches_b = 0
        self.generation = 0
        self.consumed = 0
        self.replays = 0
        self.eager_calls = 0
        self._anchor = torch.zeros((), device=ddp.device, requires_grad=True)    # makes autograd record the node
        self._mods = list(ddp.module.modules())
        self._mode_sig = None
        # The capture stream, and every parameter's AccumulateGrad node created (and kept alive) under it.  Autograd caches
        # that node per parameter for as long as any graph references it and binds it to the stream that was current when it
        # was made; at the end of a backward the engine orders the caller's stream after every such "leaf stream".  A node
        # left over from an eager step on the caller's stream (the caller still holds that step's loss) would make the
        # captured backward wait for uncaptured work -- an invalid capture.  Pinned here, the nodes belong to the capture
        # stream for good: inside the capture there is nothing to wait for, outside it the engine's ordinary cross-stream
        # ordering applies.
        self.stream = torch.cuda.Stream(device=ddp.device)
        with torch.enable_grad(), torch.cuda.stream(self.stream):
            self._accumulators = [p.view_as(p).grad_fn.next_functions[0][0] for p in ddp.params]

    # ------------------------------------------------------------------ dispatch
    def _modes(self):
        return tuple(m.training for m in self._mods)

    def wants(self, args, kwargs) -> bool:
        """True when this call can be served by (capturing or replaying) the graphs."""
        d = self.ddp
        if not (self.enabled and not self.paused and d.training and torch.is_grad_enabled() and d.require_backward_grad_sync):
            return False
        if kwargs or len(args) != 1 or not torch.is_tensor(args[0]):
            return False
        x = args[0]
        if not x.is_cuda or x.requires_grad or torch.cuda.is_current_stream_capturing():
            return False
        if self.key is None:
            return True
        return (tuple(x.shape), x.dtype) == self.key and self._modes() == self._mode_sig

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if self.graph_f is None:
            ambient = torch.cuda.current_stream(self.ddp.device)
            try:
                self._capture(x)
            except Exception as exc:          # fall back for good: the eager path is always correct
                torch.cuda.set_stream(ambient)        # torch.cuda.graph does not restore it when ending the capture raises
                self._abandon(exc)
                return self.ddp._forward_eager(x)
        self.replays += 1
        return _Replay.apply(self._anchor, self, x)

    # ------------------------------------------------------------------ capture
    def _mutable_state(self):
        """What a forward + backward changes besides the gradient arena: the module's buffers (BN running statistics)."""
        return [b.data for b in self.ddp.module.buffers()]

    def _capture(self, x: torch.Tensor):
        from ..ops import _lib as L
        d = self.ddp
        dev = d.device
        self.key = (tuple(x.shape), x.dtype)
        self._mode_sig = self._modes()
        self.static_x = x.detach().clone()
        state = self._mutable_state()
        snap = [t.clone() for t in state]
        self._undo = (state, snap)
        # warm-up outside the capture: lazily created buffers, function attributes, tensor maps, the reducer's first pass
        s = self.stream
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for _ in range(self.warmup_iters):
                y = d._forward_eager(self.static_x)
                if not torch.is_tensor(y) or not y.requires_grad:
                    raise _NotCapturable("the module must return one tensor that requires grad")
                gy = torch.zeros_like(y)
                self._backward(y, gy)
        del y
        torch.cuda.current_stream(dev).wait_stream(s)
        torch.cuda.synchronize(dev)
        self.static_gy = gy
        gf, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        n0 = L.LAUNCHES[0]
        with torch.cuda.graph(gf, stream=s):
            y = d._forward_eager(self.static_x)
        n1 = L.LAUNCHES[0]
        with torch.cuda.graph(gb, pool=gf.pool(), stream=s):
            self._backward(y, self.static_gy)
        n2 = L.LAUNCHES[0]
        self.static_y = y.detach()
        del y
        self.launches_f, self.launches_b = n1 - n0, n2 - n1
        self.graph_f, self.graph_b = gf, gb
        STATS["captures"] += 1
        # the warm-up passes and the capture itself must leave no trace: statistics back, gradient arena clean
        with torch.no_grad():
            for t, c in zip(state, snap):
                t.copy_(c)
            d.flat_grads.zero_()
        d._arena_dirty = False
        self._undo = None
        torch.cuda.synchronize(dev)
        log.debug("b200ddp: forward/backward captured for input %s (%d + %d kernels per replay)", self.key,
                  self.launches_f, self.launches_b)

    def _backward(self, y, gy):
        """Backward of the captured region: ``torch.autograd.grad`` with the parameters as explicit inputs, so that what
        autograd itself would add to a ``param.grad`` comes back here instead.  Every native layer deposits its weight
        gradient straight in the arena and reports ``None``; a parameter that does come back with a gradient belongs to a
        layer this mode cannot serve (its accumulation would have to be replayed too) -- found during the eager warm-up,
        before anything is captured."""
        grads = torch.autograd.grad((y,), self.ddp.params, (gy,), allow_unused=True)
        if any(g is not None for g in grads):
            names = [str(tuple(p.shape)) for p, g in zip(self.ddp.params, grads) if g is not None][:4]
            raise _NotCapturable("parameters %s receive their gradient through autograd, not through the arena sinks of the "
                                 "native layers" % ", ".join(names))

    def _abandon(self, exc):
        import traceback
        d = self.ddp
        self.enabled = False
        self.failure = "".join(traceback.format_exception(type(exc), exc, exc.__traceback__))
        STATS["capture_failures"] += 1
        self.graph_f = self.graph_b = None
        self.static_y = self.static_gy = None
        d._callback_queued = False
        try:
            torch.cuda.synchronize(d.device)
            with torch.no_grad():
                undo = getattr(self, "_undo", None)
                if undo is not None:             # BN statistics as they were before the warm-up passes
                    for t, c in zip(*undo):
                        t.copy_(c)
                d.flat_grads.zero_()
            d._arena_dirty = False
        except Exception:
            pass
        self._undo = None
        if not isinstance(exc, _NotCapturable):
            # a capture that died half-way can leave the device's default generator marked "capturing" (the next
            # torch.randn(device=...) would raise): swap in a copy of its state, which starts out not capturing
            try:
                gen = torch.cuda.default_generators[d.device.index if d.device.index is not None else torch.cuda.current_device()]
                gen.graphsafe_set_state(gen.clone_state())
            except Exception:
                pass
        log.warning("b200ddp: forward/backward graph capture failed (%s: %s); this module runs eagerly",
                    type(exc).__name__, str(exc).splitlines()[0][:200] if str(exc) else "")

    # ------------------------------------------------------------------ replay bookkeeping
    def note_eager(self):
        self.eager_calls += 1
        STATS["eager_forwards"] += 1

    def note_launches(self, n: int):
        from ..ops import _lib as L
        L.LAUNCHES[0] += n

    def after_backward(self):
        d = self.ddp
        d._arena_dirty = True
        if d.comm is not None and d.world_size > 1:
            d.comm.check_health()            # a peer that never reached a collective raises here
