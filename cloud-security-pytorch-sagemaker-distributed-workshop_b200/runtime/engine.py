"""Train-step engine: the hot loop of the reference's trainers
(``notebooks/code/cifar10-distributed-smddp-gpu.py:162-168``: H2D copy, ``zero_grad``, forward, loss,
``backward``, ``optimizer.step``) as ONE CUDA-graph replay per step.

On a GPU the whole step — GPU-side augmentation of the uint8 batch, forward, fused loss, backward with
the bucketed gradient all-reduce forked onto the comm stream, fused optimizer — is captured once and
replayed; per step the host only copies the next batch (uint8 pixels, labels, crop offsets) from
pinned memory into the graph's static input buffers and reads the loss back when asked.  Without a GPU
(or with ``graph=False``) the same step runs eagerly, which is the CPU/gloo fallback path.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np
import torch
import torch.nn as nn

from ..nn import modules as bmod
from ..ops import functional as BF
from ..utils.data import CIFAR10_MEAN, CIFAR10_STD, augment_reference


class TrainStep:
    def __init__(self, model: nn.Module, criterion: Callable, optimizer, batch_size: int, image_shape=(32, 32, 3),
                 device: Optional[torch.device] = None, graph: bool = True, augment: bool = True,
                 mean=CIFAR10_MEAN, std=CIFAR10_STD, seed: int = 0, warmup_iters: int = 3):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.cuda = self.device.type == "cuda"
        self.batch_size = int(batch_size)
        self.image_shape = tuple(image_shape)
        self.augment = augment
        self.use_graph = bool(graph) and self.cuda
        self.rng = np.random.RandomState(seed)
        self.warmup_iters = warmup_iters
        self.steps_done = 0
        self._graph = None
        self._captured_batch = None
        self._captured_key = None           # optimizer hyper-parameters the captured launches carry as kernel arguments
        self._recaptures = 0
        self._recapture_run = (0, -2)       # (length of the current run of consecutive re-captures, step of the last one)
        import os
        from ..models.workshop_net import WorkshopNet
        core = model.module if hasattr(model, "module") else model
        self._net_mega = core if (isinstance(core, WorkshopNet) and self.cuda and self.image_shape == (32, 32, 3)
                                  and os.environ.get("B200DDP_NET_MEGAKERNEL", "1") != "0") else None
        self.mean_t = torch.tensor(mean, dtype=torch.float32, device=self.device)
        self.inv_std_t = 1.0 / torch.tensor(std, dtype=torch.float32, device=self.device)
        self._mean, self._std = mean, std
        if self.cuda:
            H, W, C = self.image_shape
            B = self.batch_size
            # the graph's static inputs are views of ONE blob (pixels | labels | crop offsets), so the pipelined path
            # moves a staged batch into place with a single device copy
            self._nx = (B * H * W * C + 255) // 256 * 256
            self._ny = (B * 8 + 255) // 256 * 256
            self._nblob = self._nx + self._ny + B * 12
            self.static_blob = torch.zeros(self._nblob, dtype=torch.uint8, device=self.device)
            self.static_x, self.static_y, self.static_off = self._views(self.static_blob)
            self.static_loss = torch.zeros((), dtype=torch.float32, device=self.device)
            self._one = torch.ones((), dtype=torch.float32, device=self.device)   # dL/dL, so backward() launches no fill
            self.pin_off = torch.zeros((B, 3), dtype=torch.int32).pin_memory()
            self._off_ev = None                 # H2D of pin_off still in flight: wait for it before rewriting the buffer
            self.pin_loss = torch.zeros((), dtype=torch.float32).pin_memory()
            from ..parallel import process_group as _pg
            self._comm = _pg.get_comm()         # polled after every step: a timed-out collective must not go unnoticed
            bmod.defer_bn_counters(True, model)      # this model's layers only; flush_counters() adds the steps back
            # Off by default: measured on one B200 (resnet18, fp32) the pooled form saves 15 launches per step and is SLOWER,
            # 1.370 vs 1.270 ms at batch 256 and 0.793 vs 0.766 ms at batch 32 -- every CTA of the apply kernel then derives
            # 8 scale/shift pairs from the sums (32 dependent loads + rsqrt) before it touches its first row, while the tiny
            # finalize kernel overlaps the conv's tail through programmatic dependent launch (profiles/r2/pooled_bn.md).
            self._pool_bn = os.environ.get("B200DDP_POOL_BN_SUMS", "0") not in ("0", "false", "off")
            # this engine owns the whole step, so the optimizer may move into the gradient collective (world > 1)
            if getattr(model, "_graphed", None) is not None:
                model._graphed.paused = True    # this engine replays the whole step: the model's own two graphs stay out of it
            fuse = getattr(optimizer, "fuse_into_backward", None)
            self.fused_optimizer = bool(fuse(True)) if fuse is not None else False

    def _views(self, blob: torch.Tensor):
        H, W, C = self.image_shape
        B = self.batch_size
        x = blob[:B * H * W * C].view(B, H, W, C)
        y = blob[self._nx:self._nx + B * 8].view(torch.int64)
        off = blob[self._nx + self._ny:self._nx + self._ny + B * 12].view(torch.int32).view(B, 3)
        return x, y, off

    # ------------------------------------------------------------------ one step, eager
    def _draw_offsets(self, n: int) -> torch.Tensor:
        off = np.stack([self.rng.randint(0, 9, n), self.rng.randint(0, 9, n), self.rng.randint(0, 2, n)], 1)
        return torch.from_numpy(off.astype(np.int32))

    def _draw_offsets_into(self, pinned_np: np.ndarray) -> None:
        """Same draws, same order as ``_draw_offsets``, written straight into a pinned [B,3] int32 buffer."""
        n = pinned_np.shape[0]
        pinned_np[:, 0] = self.rng.randint(0, 9, n)
        pinned_np[:, 1] = self.rng.randint(0, 9, n)
        pinned_np[:, 2] = self.rng.randint(0, 2, n)

    def _body(self, x_u8, y, offsets):
        # this engine calls backward() on the raw loss: the fused criterion may skip its gradient-scale kernel -- for the
        # duration of this step only (a process-wide switch would silently drop another caller's `(loss * w).backward()`)
        prev = BF._unit_loss_grad[0]
        BF._unit_loss_grad[0] = self.cuda
        try:
            return self._body_impl(x_u8, y, offsets)
        finally:
            BF._unit_loss_grad[0] = prev

    def _body_impl(self, x_u8, y, offsets):
        from ..ops import native as nv
        if self.cuda and self._net_mega is not None:
            # workshop Net: augmentation + forward/loss/backward megakernel (2 launches) + fused SGD
            self.optimizer.zero_grad()
            loss = self._net_mega.fused_loss(x_u8, y, aug=(offsets if self.augment else None, self.mean_t, self.inv_std_t))
            loss.backward(self._one)
            self.optimizer.step()
            return loss.detach()
        if self.cuda:
            cin = 8
            inp = nv.augment(x_u8, offsets if self.augment else None, self.mean_t, self.inv_std_t, c_out=cin)
        else:
            inp = augment_reference(x_u8, train_transform=self.augment, mean=self._mean, std=self._std, offsets=offsets)
        self.optimizer.zero_grad()
        if self.cuda and self._pool_bn:
            # this engine knows where a step starts: all conv-epilogue BN statistics live in one pool, cleared here once
            bmod.zero_bn_sum_pool(self.device)
            bmod._POOL_ACTIVE[0] = True
            try:
                out = self.model(inp)
            finally:
                bmod._POOL_ACTIVE[0] = False
        else:
            out = self.model(inp)
        loss = self.criterion(out, y)
        loss.backward(self._one if self.cuda else None)
        self.optimizer.step()
        # detach: a live loss keeps last step's autograd graph (and its AccumulateGrad nodes, which are
        # bound to the stream they were created on) alive, which breaks CUDA-graph capture
        return loss.detach()

    # ------------------------------------------------------------------ graph capture
    def _training_state(self):
        """Every tensor a training step mutates besides the gradients: parameters, buffers (BN statistics), optimizer state."""
        ts = [p.data for p in self.model.parameters()] + [b.data for b in self.model.buffers()]
        for extra in ("flat_shadow",):
            t = getattr(self.model, extra, None)
            if t is not None:
                ts.append(t)
        opt = self.optimizer
        for attr in ("_momentum_flat", "_state_flat"):
            for v in getattr(opt, attr, {}).values():
                ts.extend(t for t in (v if isinstance(v, (tuple, list)) else (v,)) if torch.is_tensor(t))
        for st in getattr(getattr(opt, "_inner", None) or opt, "state", {}).values():
            ts.extend(t for t in st.values() if torch.is_tensor(t))
        return ts

    def _hyper_key(self):
        """The optimizer's scalar hyper-parameters.  In a captured step they are kernel ARGUMENTS (of the optimizer pass and
        of the fused bucket step alike), so a replay keeps the values of capture time: a learning-rate schedule has to be
        noticed on the host."""
        return tuple(tuple((k, v) for k, v in sorted(g.items()) if k != "params" and isinstance(v, (bool, int, float, tuple)))
                     for g in self.optimizer.param_groups)

    def _stale(self) -> bool:
        return self._graph is not None and self._hyper_key() != self._captured_key

    def _recapture(self):
        """A hyper-parameter changed (``param_groups[0]['lr'] = ...``, an LR scheduler): capture the step again with the new
        values.  Costs a warm-up + capture (~0.1 s), fine for a per-epoch schedule; a schedule that moves EVERY step would
        pay that on every step, so after four re-captures in a row the engine runs launch by launch instead, where the
        optimizer reads its hyper-parameters at every step."""
        run, last = self._recapture_run
        run = run + 1 if self.steps_done - last <= 1 else 1
        self._recapture_run = (run, self.steps_done)
        self._recaptures += 1
        torch.cuda.synchronize(self.device)
        self._graph = None
        if run > 4:
            import warnings
            warnings.warn("b200ddp.TrainStep: optimizer hyper-parameters change at every step; the step now runs without a "
                          "CUDA graph (keep them in place for several steps, or pass graph=False)")
            self.use_graph = False
            return
        self._capture()

    def _capture(self):
        # The warm-up iterations are real optimizer steps on whatever batch sits in the static inputs; they exist to run the
        # one-time set-up (function attributes, lazily created buffers, optimizer state) outside the capture.  The training
        # state is snapshotted first and put back afterwards, so a captured run still takes exactly one step per batch
        # (the reference's loop, …smddp-gpu.py:162-168, never trains a batch twice).
        before = self._training_state()
        snap = [t.clone() for t in before]
        known = {t.data_ptr() for t in before}
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(self.warmup_iters):
                loss = self._body(self.static_x, self.static_y, self.static_off)
                self.static_loss.copy_(loss)
        del loss
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        # a model that can also replay its own forward/backward keeps its parameters' AccumulateGrad nodes on one stream
        # (runtime/graphed.py): capture there, a captured backward cannot order itself against any other stream
        own = getattr(getattr(self.model, "_graphed", None), "stream", None)
        with torch.cuda.graph(g, stream=own):
            loss = self._body(self.static_x, self.static_y, self.static_off)
            self.static_loss.copy_(loss)
        self._graph = g
        self._captured_key = self._hyper_key()
        self._pipe_init()
        with torch.no_grad():
            for t, c in zip(before, snap):
                t.copy_(c)
            for t in self._training_state():          # optimizer state created by the warm-up: back to its initial zeros
                if t.data_ptr() not in known:
                    t.zero_()
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ public
    def __call__(self, x_u8: torch.Tensor, y: torch.Tensor, need_loss: bool = False):
        """Run one optimisation step on a uint8 NHWC batch (host-pinned, device or CPU tensors).
        Returns the loss as a Python float when ``need_loss`` (a D2H read), else ``None``."""
        n = x_u8.shape[0]
        offsets = self._draw_offsets(n) if self.augment else None
        if not self.cuda:
            loss = self._body(x_u8, y, offsets)
            self.steps_done += 1
            return float(loss.item()) if need_loss else None
        if self.use_graph and n == self.batch_size:
            if self._graph is None:
                self.static_x.copy_(x_u8, non_blocking=True)
                self.static_y.copy_(y, non_blocking=True)
                self._capture()
            elif self._stale():
                self._recapture()
                if self._graph is None:                  # the engine gave up on graphs (see _recapture)
                    return self(x_u8, y, need_loss=need_loss)
            self.static_x.copy_(x_u8, non_blocking=True)
            self.static_y.copy_(y, non_blocking=True)
            if offsets is not None:
                if self._off_ev is not None:
                    self._off_ev.synchronize()       # the previous step's H2D of this pinned buffer has executed
                self.pin_off.copy_(offsets)
                self.static_off.copy_(self.pin_off, non_blocking=True)
                if self._off_ev is None:
                    self._off_ev = torch.cuda.Event()
                self._off_ev.record()
            self._graph.replay()
            self.steps_done += 1
            if self._comm is not None:
                self._comm.check_health()
            if need_loss:
                self.pin_loss.copy_(self.static_loss, non_blocking=True)
                torch.cuda.current_stream(self.device).synchronize()
                return float(self.pin_loss.item())
            return None
        # eager path (last short batch of an epoch, or graph=False)
        xd = x_u8.to(self.device, non_blocking=True)
        yd = y.to(self.device, non_blocking=True)
        od = offsets.to(self.device, non_blocking=True) if offsets is not None else None
        loss = self._body(xd, yd, od)
        self.steps_done += 1
        return float(loss.item()) if need_loss else None

    def pipeline(self, batches, need_loss: bool = True):
        """Run one optimisation step per ``(x_u8, y)`` host batch (pinned memory) and yield one loss per step.

        Software pipeline of depth 2: batch k+1 is uploaded on a copy stream into a staging blob while step k
        computes, one device copy moves it into the graph's static inputs, and the loss of step k is read back
        asynchronously and yielded while step k+1 runs (so the caller sees each loss one step late; every loss is
        still read, in order).  Falls back to plain ``__call__`` without a GPU / graph or for a short batch."""
        if not (self.cuda and self.use_graph):
            for x, y in batches:
                yield self(x, y, need_loss=need_loss)
            return
        dev = self.device
        self._pipe_init()
        P = self._pipe
        cur = torch.cuda.current_stream(dev)
        pending = None                       # slot whose loss has not been yielded yet
        k = 0
        for x, y in batches:
            if x.shape[0] != self.batch_size or self._graph is None or self._stale():
                if pending is not None:
                    P["ready"][pending].synchronize()
                    yield float(P["pin_loss"][pending].item()) if need_loss else None
                    pending = None
                yield self(x, y, need_loss=need_loss)
                continue
            slot = k & 1
            sx, sy, so = P["views"][slot]
            if k >= 2:
                P["h2d"][slot].synchronize()             # pinned offsets of step k-2 have been consumed
                P["cs"].wait_event(P["free"][slot])      # ... and its staged batch has been moved into place
            else:
                P["cs"].wait_stream(cur)
            with torch.cuda.stream(P["cs"]):
                sx.copy_(x, non_blocking=True)
                sy.copy_(y, non_blocking=True)
                if self.augment:
                    self._draw_offsets_into(P["pin_off_np"][slot])
                    so.copy_(P["pin_off"][slot], non_blocking=True)
                P["h2d"][slot].record(P["cs"])
            cur.wait_event(P["h2d"][slot])
            self.static_blob.copy_(P["stage"][slot], non_blocking=True)
            P["free"][slot].record(cur)
            self._graph.replay()
            self.steps_done += 1
            if self._comm is not None:
                self._comm.check_health()            # host-mapped flag: a peer that never arrived raises here, not at the end
            if need_loss:
                P["pin_loss"][slot].copy_(self.static_loss, non_blocking=True)
            P["ready"][slot].record(cur)
            if pending is not None:
                P["ready"][pending].synchronize()
                yield float(P["pin_loss"][pending].item()) if need_loss else None
            pending = slot
            k += 1
        if pending is not None:
            P["ready"][pending].synchronize()
            yield float(P["pin_loss"][pending].item()) if need_loss else None

    def _pipe_init(self):
        """Staging blobs, pinned offset / loss slots, copy stream and events of ``pipeline`` (allocated once, at capture)."""
        dev = self.device
        if not hasattr(self, "_pipe"):
            cs = torch.cuda.Stream(device=dev)
            self._pipe = {
                "cs": cs,
                "stage": [torch.zeros_like(self.static_blob) for _ in range(2)],
                "pin_off": [torch.zeros((self.batch_size, 3), dtype=torch.int32).pin_memory() for _ in range(2)],
                "pin_loss": [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)],
                "h2d": [torch.cuda.Event() for _ in range(2)],
                "free": [torch.cuda.Event() for _ in range(2)],
                "ready": [torch.cuda.Event() for _ in range(2)],
            }
            self._pipe["views"] = [self._views(b) for b in self._pipe["stage"]]
            self._pipe["pin_off_np"] = [t.numpy() for t in self._pipe["pin_off"]]

    def breakdown(self, x_u8: torch.Tensor, y: torch.Tensor, iters: int = 5) -> dict:
        """Device-timed phases of one EAGER step (CUDA events on the step's stream, mean over ``iters``, max over ranks):
        ``forward`` (augmentation + model + loss), ``backward`` (until the last gradient kernel of this rank has been
        queued behind), ``exposed_comm`` (what the step still waits for once backward's compute is done: the tail of the
        bucket collectives -- with the fused bucket step that includes the optimizer of the last bucket), ``optimizer``
        (the separate optimizer pass; ~0 when it is fused into the collectives).  These are real optimizer steps.  Eager
        launches pay host latency the captured graph does not, so the sum is larger than ``ms_per_step`` of a replay;
        the split is what it is for."""
        if not self.cuda:
            return {}
        from ..ops import native as nv
        from ..parallel import process_group as pg
        dev = self.device
        xd, yd = x_u8.to(dev), y.to(dev)
        off = self._draw_offsets(xd.shape[0]).to(dev) if self.augment else None
        ddp = self.model if hasattr(self.model, "_finalize_backward") else None
        tot = dict.fromkeys(("forward", "backward", "exposed_comm", "optimizer"), 0.0)
        for it in range(iters + 1):
            ev = {k: torch.cuda.Event(enable_timing=True) for k in ("t0", "t1", "bwd_compute_end", "comm_joined", "t2", "t3")}
            if ddp is not None:
                ddp._phase_events = ev
            torch.cuda.synchronize(dev)
            ev["t0"].record()
            if self._net_mega is not None:
                self.optimizer.zero_grad()
                loss = self._net_mega.fused_loss(xd, yd, aug=(off, self.mean_t, self.inv_std_t))
            else:
                inp = nv.augment(xd, off, self.mean_t, self.inv_std_t, c_out=8)
                self.optimizer.zero_grad()
                loss = self.criterion(self.model(inp), yd)
            ev["t1"].record()
            prev, BF._unit_loss_grad[0] = BF._unit_loss_grad[0], True
            try:
                loss.backward(self._one)
            finally:
                BF._unit_loss_grad[0] = prev
            ev["t2"].record()
            self.optimizer.step()
            ev["t3"].record()
            torch.cuda.synchronize(dev)
            if ddp is not None:
                ddp._phase_events = None
            self.steps_done += 1
            if it == 0:
                continue                     # first pass: lazy allocations
            tot["forward"] += ev["t0"].elapsed_time(ev["t1"])
            if ev.get("hit"):
                tot["backward"] += ev["t1"].elapsed_time(ev["bwd_compute_end"])
                tot["exposed_comm"] += ev["bwd_compute_end"].elapsed_time(ev["comm_joined"])
            else:
                tot["backward"] += ev["t1"].elapsed_time(ev["t2"])
            tot["optimizer"] += ev["t2"].elapsed_time(ev["t3"])
        return {k: pg.max_over_ranks(v / iters) for k, v in tot.items()}

    def flush_counters(self):
        if self.cuda:
            bmod.flush_bn_counters(self.model, self.steps_done - getattr(self, "_flushed", 0))
            self._flushed = self.steps_done
