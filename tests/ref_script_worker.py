"""Runs the reference's UNMODIFIED notebooks/code/cifar10-distributed-native-cpu.py on top of b200ddp:
torch.nn.parallel.DistributedDataParallel -> b200ddp.DistributedDataParallel, torchvision CIFAR10 -> synthetic."""
import json
import os
import runpy
import sys

import numpy as np
import torch

sys.dont_write_bytecode = True
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torchvision  # noqa: E402
from PIL import Image  # noqa: E402

model_dir = sys.argv[1]
rank = int(os.environ["RANK"])
hosts = ["algo-1", "algo-2"]
os.environ.update({"SM_HOSTS": json.dumps(hosts), "SM_CURRENT_HOST": hosts[rank], "SM_MODEL_DIR": model_dir,
                   "SM_CHANNEL_TRAIN": "/tmp"})


class SynthCIFAR:
    def __init__(self, root, train=True, transform=None, download=False):
        n = 256 if train else 32
        rng = np.random.RandomState(1 if train else 2)
        self.data = rng.randint(0, 256, size=(n, 32, 32, 3), dtype=np.uint8)
        self.targets = rng.randint(0, 10, size=n).tolist()
        self.transform = transform

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        img = Image.fromarray(self.data[i])
        return (self.transform(img) if self.transform else img), self.targets[i]


torchvision.datasets.CIFAR10 = SynthCIFAR


class DDPShim(bd.DistributedDataParallel):
    """Same constructor call as the reference makes: DistributedDataParallel(model)."""


# the reference calls torch.distributed.init_process_group itself; our wrapper adopts that group
real_init = dist.init_process_group


def init_and_adopt(*a, **kw):
    real_init(*a, **kw)
    from b200ddp.parallel import process_group as pg
    pg._S.backend, pg._S.rank, pg._S.world_size = "gloo", dist.get_rank(), dist.get_world_size()
    pg._S.control_group = dist.group.WORLD


dist.init_process_group = init_and_adopt
torch.nn.parallel.DistributedDataParallel = DDPShim
script = "/root/reference/notebooks/code/cifar10-distributed-native-cpu.py"
sys.argv = [script, "--backend", "gloo", "--batch-size", "32", "--epochs", "1", "--lr", "0.01", "--momentum", "0.9",
            "--log-interval", "2", "--test-batch-size", "16"]
runpy.run_path(script, run_name="__main__")
