"""Turn a stock PyTorch model into its b200ddp-native equivalent, in place of rewriting the user's script.

The reference builds its models from library classes -- ``torchvision.models.resnet18(pretrained=False)`` or an
inline ``Net`` of ``nn.Conv2d / nn.MaxPool2d / nn.Linear``
(``notebooks/code/cifar10-distributed-smddp-gpu.py:30-52``) -- and wraps them in DDP (``:148``).  When that wrap
happens on the b200 backend (``smdistributed/dataparallel/torch/torch_smddp.py``) the model goes through ``to_native``:

* a torchvision ``ResNet`` becomes ``b200ddp.models.ResNet`` with the same block layout (fused conv->BN statistics,
  BN+ReLU(+residual) kernels) -- same parameter names and shapes, weights and buffers carried over;
* the workshop's ``Net`` (conv1/pool/conv2/fc1/fc2/fc3 with the reference's shapes) becomes ``WorkshopNet``;
* anything else keeps its own ``forward`` and has its leaf layers swapped one for one (``nn.Conv2d`` ->
  ``b200ddp.nn.Conv2d`` and so on): unfused, but every convolution / linear / batch-norm / pool still runs on this
  package's kernels and ``state_dict()`` keys do not change.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import nn as bnn


def _is_torchvision_resnet(m: nn.Module) -> bool:
    try:
        from torchvision.models.resnet import ResNet as TVResNet
    except Exception:        # torchvision absent
        return False
    return isinstance(m, TVResNet)


def _looks_like_workshop_net(m: nn.Module) -> bool:
    names = {n for n, _ in m.named_children()}
    if names != {"conv1", "pool", "conv2", "fc1", "fc2", "fc3"}:
        return False
    c1, c2 = m.conv1, m.conv2
    return (isinstance(c1, nn.Conv2d) and isinstance(c2, nn.Conv2d) and isinstance(m.fc1, nn.Linear)
            and tuple(c1.weight.shape) == (6, 3, 5, 5) and tuple(c2.weight.shape) == (16, 6, 5, 5)
            and tuple(m.fc1.weight.shape) == (120, 400) and tuple(m.fc2.weight.shape) == (84, 120)
            and tuple(m.fc3.weight.shape) == (10, 84))


def _carry_over(dst: nn.Module, src: nn.Module) -> nn.Module:
    dev = next(src.parameters()).device
    dst = dst.to(dev)
    dst.load_state_dict(src.state_dict(), strict=True)
    dst.train(src.training)
    return dst


def _swap_leaf(m: nn.Module) -> nn.Module:
    """One-for-one native replacement of a supported leaf layer (sharing its parameter tensors), or ``m`` itself."""
    if isinstance(m, (bnn.Conv2d, bnn.Linear, bnn.BatchNorm2d, bnn.MaxPool2d, bnn.GlobalAvgPool, bnn.ReLU)):
        return m
    if isinstance(m, nn.Conv2d) and not isinstance(m, nn.ConvTranspose2d):
        ok = (m.groups == 1 and m.dilation == (1, 1) and m.padding_mode == "zeros" and isinstance(m.padding, tuple)
              and m.stride[0] == m.stride[1] and m.padding[0] == m.padding[1] and m.stride[0] <= 2)
        if not ok:
            return m
        new = bnn.Conv2d(m.in_channels, m.out_channels, m.kernel_size, stride=m.stride[0], padding=m.padding[0],
                         bias=m.bias is not None)
        new.weight, new.bias = m.weight, m.bias
        return new
    if isinstance(m, nn.Linear):
        new = bnn.Linear(m.in_features, m.out_features, bias=m.bias is not None)
        new.weight, new.bias = m.weight, m.bias
        return new
    if isinstance(m, nn.BatchNorm2d) and m.affine and m.track_running_stats:
        new = bnn.BatchNorm2d(m.num_features, eps=m.eps, momentum=m.momentum if m.momentum is not None else 0.1)
        new.weight, new.bias = m.weight, m.bias
        new.running_mean, new.running_var, new.num_batches_tracked = m.running_mean, m.running_var, m.num_batches_tracked
        return new
    if isinstance(m, nn.MaxPool2d) and not m.ceil_mode and m.dilation in (1, (1, 1)) and not m.return_indices:
        k = m.kernel_size if isinstance(m.kernel_size, int) else m.kernel_size[0]
        s = m.stride if isinstance(m.stride, int) else m.stride[0]
        p = m.padding if isinstance(m.padding, int) else m.padding[0]
        return bnn.MaxPool2d(k, s, p)
    if isinstance(m, nn.ReLU):
        return bnn.ReLU()
    return m


def swap_leaves(module: nn.Module) -> nn.Module:
    for name, child in list(module.named_children()):
        new = _swap_leaf(child)
        if new is not child:
            new.train(child.training)
            setattr(module, name, new)
        else:
            swap_leaves(child)
    return module


def to_native(module: nn.Module) -> nn.Module:
    """Native equivalent of ``module`` (see the module docstring).  CPU modules are returned unchanged: the native
    layers dispatch on the input tensor, and on the CPU they evaluate the reference math anyway."""
    try:
        first = next(module.parameters())
    except StopIteration:
        return module
    if first.device.type != "cuda":
        return module
    from ..models import ResNet, WorkshopNet
    from ..models.resnet import BasicBlock, Bottleneck
    if isinstance(module, (ResNet, WorkshopNet)):
        return module
    if _is_torchvision_resnet(module):
        from torchvision.models.resnet import BasicBlock as TVBasic, Bottleneck as TVBottle
        blocks = [len(getattr(module, "layer%d" % i)) for i in (1, 2, 3, 4)]
        first_block = module.layer1[0]
        plain = (getattr(module, "groups", 1) == 1 and getattr(module, "base_width", 64) == 64
                 and all(d == 1 for d in [getattr(b.conv2, "dilation", (1, 1))[0] for l in (module.layer1, module.layer2, module.layer3, module.layer4) for b in l]))
        if plain and isinstance(first_block, (TVBasic, TVBottle)):
            block = BasicBlock if isinstance(first_block, TVBasic) else Bottleneck
            return _carry_over(ResNet(block, blocks, num_classes=module.fc.out_features), module)
    if _looks_like_workshop_net(module):
        return _carry_over(WorkshopNet(), module)
    return swap_leaves(module)
