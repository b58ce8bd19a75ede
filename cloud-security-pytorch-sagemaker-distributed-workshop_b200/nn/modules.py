"""``nn.Module`` drop-ins backed by the native kernels (``state_dict`` keys identical to torch's).

Each module subclasses its torch counterpart, so parameter names/shapes — and therefore the
``model.pth`` checkpoints of the reference (``notebooks/code/cifar10-distributed-native-cpu.py:196-199``,
``inference.py:28-34``) — are unchanged.  ``forward`` dispatches on the input: CUDA tensors run the
sm_100a kernels on bf16 channels-last activations, CPU tensors run the reference fp32 math.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from ..ops import functional as BF


# B200DDP_FUSE_BN_STATS=0: BatchNorm always makes its own statistics pass
_FUSE_BN_STATS = [__import__("os").environ.get("B200DDP_FUSE_BN_STATS", "1") != "0"]


# Pooled BN statistics (set up by runtime.engine.TrainStep, active only inside its step body).  A conv whose epilogue produces
# the statistics of the following BatchNorm accumulates them in a small fp32 buffer that has to be zero again before the next
# forward.  Stand-alone, a tiny finalize kernel per layer reads and clears it.  A step engine knows where a step begins: while
# it is active every layer's buffer is a slice of ONE pool that is cleared by ONE memset per step, and BatchNorm derives
# scale / shift from the sums inside its apply kernel (`b200_bn_apply_sums_keep`): one launch per BN layer less.
_POOL_ACTIVE = [False]
_POOLS = {}
_POOL_FLOATS = 1 << 18


def _pool_alloc(device, n: int) -> torch.Tensor:
    key = (device.type, device.index)
    pool = _POOLS.get(key)
    if pool is None:
        pool = {"buf": torch.zeros(_POOL_FLOATS, dtype=torch.float32, device=device), "used": 0}
        _POOLS[key] = pool
    n = (n + 63) // 64 * 64
    if pool["used"] + n > _POOL_FLOATS:
        return None
    t = pool["buf"][pool["used"]: pool["used"] + n]
    pool["used"] += n
    t._b200_pooled = True
    return t


def zero_bn_sum_pool(device) -> None:
    """One memset over every pooled BN-statistics buffer of ``device`` (start of a step)."""
    pool = _POOLS.get((device.type, device.index))
    if pool is not None and pool["used"]:
        pool["buf"][: pool["used"]].zero_()


class Conv2d(nn.Conv2d):
    """conv (+bias) with an optional fused ReLU epilogue."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, bias=True, relu=False):
        super().__init__(in_channels, out_channels, kernel_size, stride=stride, padding=padding, bias=bias)
        if self.stride[0] != self.stride[1] or self.padding[0] != self.padding[1]:
            raise ValueError("b200ddp.nn.Conv2d supports symmetric stride/padding only")
        self.fused_relu = relu
        # the BatchNorm2d that consumes this conv's output (fuse_bn_stats).  Set with object.__setattr__ so that it is NOT
        # registered as a child module (state_dict keys stay torch's) while deepcopy / pickle still carry the pairing along.
        object.__setattr__(self, "_bn_pair", None)
        self._bn_sums = None
        self._bn_sums_pooled = None

    def fuse_bn_stats(self, bn: "BatchNorm2d"):
        """Declare that ``bn`` normalises exactly this conv's output.  In training mode on CUDA the conv epilogue then
        accumulates the per-channel sum / sum of squares (fp32) and ``bn`` skips its own statistics pass over the tensor."""
        if self.bias is not None or self.fused_relu:
            raise ValueError("fuse_bn_stats needs a conv without bias and without fused ReLU")
        object.__setattr__(self, "_bn_pair", bn)
        return self

    def forward(self, x):
        if not x.is_cuda:
            return BF.conv2d(x, self.weight, self.bias, self.stride[0], self.padding[0], self.fused_relu)
        x = BF.as_activation(x)
        bn = self._bn_pair
        if bn is not None and self.training and bn.training and _FUSE_BN_STATS[0] and self.out_channels % 8 == 0:
            n, _, h, w = x.shape
            ho = (h + 2 * self.padding[0] - self.kernel_size[0]) // self.stride[0] + 1
            wo = (w + 2 * self.padding[0] - self.kernel_size[1]) // self.stride[0] + 1
            rows = n * ho * wo
            if rows > BF.nv.BN_SMALL_ROWS:      # small tensors: BN's single-launch kernel is cheaper than finalize + apply
                kp = BF.ceil8(self.out_channels)
                sums = None
                if _POOL_ACTIVE[0]:
                    if self._bn_sums_pooled is None or self._bn_sums_pooled.device != x.device:
                        self._bn_sums_pooled = _pool_alloc(x.device, 2 * kp + 1)
                    sums = self._bn_sums_pooled
                if sums is None:
                    if self._bn_sums is None or self._bn_sums.device != x.device:
                        self._bn_sums = torch.zeros(2 * kp + 1, dtype=torch.float32, device=x.device)   # sums + ticket, zero at rest
                    sums = self._bn_sums
                y = BF.conv2d(x, self.weight, None, self.stride[0], self.padding[0], False, sums)
                y._b200_bn_sums = sums
                return y
        return BF.conv2d(x, self.weight, self.bias, self.stride[0], self.padding[0], self.fused_relu)


class Linear(nn.Linear):
    """linear (+bias) with an optional fused ReLU; ``nhwc_from=(C,H,W)`` marks a layer that consumes a
    flattened conv activation whose reference column order is (c,h,w)."""

    def __init__(self, in_features, out_features, bias=True, relu=False, nhwc_from=None):
        super().__init__(in_features, out_features, bias=bias)
        self.fused_relu = relu
        self.nhwc_from = tuple(nhwc_from) if nhwc_from is not None else None

    def forward(self, x):
        if not x.is_cuda:
            if x.dim() > 2:
                x = x.flatten(1)
            return BF.linear(x, self.weight, self.bias, self.fused_relu)
        nh = None
        if x.dim() == 4:
            n, cp, h, w = x.shape
            if self.nhwc_from is not None:
                c = self.nhwc_from[0]
                nh = (c, h, w, cp)
            x = BF.flatten_nhwc(BF.as_activation(x))
        else:
            x = BF.as_activation(x)
            kp = BF.ceil8(x.shape[1])
            if kp != x.shape[1]:
                x = torch.nn.functional.pad(x, (0, kp - x.shape[1]))
        return BF.linear(x, self.weight, self.bias, self.fused_relu, nh)


# Deferred counters: BatchNorm2d does not launch one `num_batches_tracked += 1` kernel per layer per step; whoever owns the
# step adds the step count once before a checkpoint is written (`flush_bn_counters`).  Scoped to the BatchNorm layers of ONE
# model (`defer_bn_counters(True, model)`, what `TrainStep` does): another model in the same process keeps counting by
# itself.  Without a model the switch is process-wide (profiling scripts).
_DEFER_BN_COUNTER = [False]


def defer_bn_counters(flag: bool = True, model: Optional[nn.Module] = None) -> None:
    if model is None:
        _DEFER_BN_COUNTER[0] = bool(flag)
        return
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            m._b200_defer_counter = bool(flag)


def flush_bn_counters(model: nn.Module, steps: int) -> None:
    if steps <= 0:
        return
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d) and m.num_batches_tracked is not None:
            m.num_batches_tracked.add_(steps)


class BatchNorm2d(nn.BatchNorm2d):
    """BN with fused ReLU and residual add: ``y = relu(bn(x) + residual)``."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, relu=False):
        super().__init__(num_features, eps=eps, momentum=momentum)
        self.fused_relu = relu

    def forward(self, x, residual: Optional[torch.Tensor] = None):
        training = self.training or not self.track_running_stats
        presummed = getattr(x, "_b200_bn_sums", None) if (x.is_cuda and self.training) else None
        if x.is_cuda:
            x = BF.as_activation(x)
            if residual is not None:
                residual = BF.as_activation(residual)
        if (self.training and self.track_running_stats and self.num_batches_tracked is not None
                and not (_DEFER_BN_COUNTER[0] or getattr(self, "_b200_defer_counter", False))):
            self.num_batches_tracked.add_(1)          # CPU fallback included: state_dict parity with nn.BatchNorm2d
        return BF.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, training,
                             self.momentum if self.momentum is not None else 0.1, self.eps, self.fused_relu, residual,
                             presummed)


class MaxPool2d(nn.Module):
    def __init__(self, kernel_size, stride=None, padding=0):
        super().__init__()
        self.kernel_size, self.stride, self.padding = kernel_size, stride if stride is not None else kernel_size, padding

    def forward(self, x):
        if x.is_cuda:
            x = BF.as_activation(x)
        return BF.max_pool2d(x, self.kernel_size, self.stride, self.padding)


class GlobalAvgPool(nn.Module):
    def forward(self, x):
        return BF.global_avg_pool(x)


class ReLU(nn.Module):
    def forward(self, x):
        return BF.relu(x)


class CrossEntropyLoss(nn.Module):
    """Mean softmax cross-entropy (``nn.CrossEntropyLoss()`` of the reference, …native-cpu.py:143)."""

    def __init__(self, num_classes: Optional[int] = None):
        super().__init__()
        self.num_classes = num_classes

    def forward(self, logits, targets):
        return BF.cross_entropy(logits, targets, self.num_classes)
