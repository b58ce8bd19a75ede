"""The workshop's LeNet-style CIFAR-10 ``Net`` (62,006 parameters).

Same topology and ``state_dict`` keys (``conv1 conv2 fc1 fc2 fc3``) as the reference definitions in
``notebooks/code/cifar10-distributed-native-cpu.py:22-39``, ``cifar10-distributed-smddp-gpu.py:34-51`` and
``inference.py:9-26``, so a checkpoint written here loads in the reference's unmodified ``model_fn``.
On CUDA the convolutions/linears run on the tensor-core kernel with channel counts padded to 8
(3->8, 6->8, 84->88, 10->16); ReLU is fused into the producing kernel's epilogue.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import nn as bnn


class WorkshopNet(nn.Module):
    num_classes = 10

    def __init__(self):
        super().__init__()
        self.conv1 = bnn.Conv2d(3, 6, 5, relu=True)
        self.pool = bnn.MaxPool2d(2, 2)
        self.conv2 = bnn.Conv2d(6, 16, 5, relu=True)
        self.fc1 = bnn.Linear(16 * 5 * 5, 120, relu=True, nhwc_from=(16, 5, 5))
        self.fc2 = bnn.Linear(120, 84, relu=True)
        self.fc3 = bnn.Linear(84, 10)

    def forward(self, x):
        x = self.pool(self.conv1(x))
        x = self.pool(self.conv2(x))
        if not x.is_cuda:
            x = x.reshape(-1, 16 * 5 * 5)
        x = self.fc1(x)
        x = self.fc2(x)
        x = self.fc3(x)
        return x[:, : self.num_classes] if x.shape[1] != self.num_classes else x


    def fused_loss(self, x, targets, aug=None):
        """Forward + cross-entropy + backward in two kernel launches (CUDA only; `ops.functional`).  ``x`` is the
        normalised fp32 NCHW batch, or the raw uint8 NHWC batch with ``aug = (offsets | None, mean, inv_std)``."""
        from ..ops import functional as BF
        return BF.workshop_net_fused_loss(self, x, targets, aug)

    def logits_fused(self, x_nchw_fp32):
        """Inference through the megakernel's forward half (fp32)."""
        from ..ops import native as nv
        params = [p.detach() for p in (self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias, self.fc1.weight,
                                       self.fc1.bias, self.fc2.weight, self.fc2.bias, self.fc3.weight, self.fc3.bias)]
        krsc = not self.conv1.weight.is_contiguous()
        dummy = torch.zeros(x_nchw_fp32.shape[0], dtype=torch.int64, device=x_nchw_fp32.device)
        return nv.net_fwd_bwd(x_nchw_fp32.contiguous(), dummy, params, None, krsc, want_logits=True)[1]


Net = WorkshopNet   # the reference's class name
