"""CPU-side behaviour of the shipped ``smdistributed.dataparallel.torch.torch_smddp`` module: importing it must not
disturb a non-smddp job (other backends pass through to torch, torch's own DDP is built when the process group is
not the b200 one), and the model converter keeps parameters, names and numerics."""
import os
import subprocess
import sys

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r)
os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29577"})
import torch, torch.nn as nn, torch.distributed as dist
stock_ddp = torch.nn.parallel.DistributedDataParallel
import smdistributed.dataparallel.torch.torch_smddp as shim
assert torch.nn.parallel.DistributedDataParallel is shim.DistributedDataParallel
dist.init_process_group(backend="gloo", rank=0, world_size=1)          # not 'smddp': straight through to torch
assert dist.get_world_size() == 1 and dist.get_backend() == "gloo"
m = torch.nn.parallel.DistributedDataParallel(nn.Linear(4, 2))          # no b200 group: torch's own class
assert type(m) is stock_ddp, type(m)
assert isinstance(m, torch.nn.parallel.DistributedDataParallel)
m(torch.randn(3, 4)).sum().backward()
try:
    dist.init_process_group(backend="smddp")
except RuntimeError as e:
    print("SMDDP_NEEDS_GPU", "CUDA" in str(e) or "twice" in str(e) or "initialized" in str(e).lower())
dist.destroy_process_group()
shim.uninstall()
assert torch.nn.parallel.DistributedDataParallel is stock_ddp
print("SHIM_OK")
'''


def test_shim_is_transparent_without_the_smddp_backend():
    res = subprocess.run([sys.executable, "-c", WORKER % ROOT], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0 and "SHIM_OK" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]


def test_converter_swaps_leaves_and_keeps_the_state_dict():
    from b200ddp import nn as bnn
    from b200ddp.utils.convert import swap_leaves, to_native

    class Odd(nn.Module):
        def __init__(self):
            super().__init__()
            self.a = nn.Conv2d(3, 8, 3, padding=1)
            self.bn = nn.BatchNorm2d(8)
            self.act = nn.ReLU()
            self.pool = nn.MaxPool2d(2)
            self.g = nn.Conv2d(8, 8, 3, groups=2)          # grouped: not supported natively, must stay stock
            self.fc = nn.Linear(8, 5)

        def forward(self, x):
            x = self.pool(self.act(self.bn(self.a(x))))
            return self.fc(self.g(x).mean(dim=(2, 3)))

    torch.manual_seed(0)
    m = Odd()
    assert to_native(m) is m                                 # CPU modules are left alone
    x = torch.randn(2, 3, 8, 8)
    want = m(x)
    keys = list(m.state_dict().keys())
    w_before = m.a.weight
    swap_leaves(m)
    assert isinstance(m.a, bnn.Conv2d) and isinstance(m.bn, bnn.BatchNorm2d) and isinstance(m.fc, bnn.Linear)
    assert isinstance(m.pool, bnn.MaxPool2d) and isinstance(m.act, bnn.ReLU) and type(m.g) is nn.Conv2d
    assert m.a.weight is w_before and list(m.state_dict().keys()) == keys
    assert torch.allclose(m(x), want, atol=1e-6)             # on the CPU the native layers evaluate the reference math
