"""Runs the reference's UNMODIFIED notebooks/code/cifar10-distributed-smddp-gpu.py on this framework: the only things
supplied are what a SageMaker container supplies (SM_* environment, argv, a CIFAR-10 dataset -- synthetic here) and the
``smdistributed`` package this repository ships (found through sys.path).  No monkey-patching of the script."""
import json
import os
import runpy
import sys

import numpy as np

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torchvision  # noqa: E402
from PIL import Image  # noqa: E402

model_dir, model_type, script = sys.argv[1], sys.argv[2], sys.argv[3]
os.environ.setdefault("RANK", "0")
os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("LOCAL_RANK", os.environ["RANK"])
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
os.environ.update({"SM_HOSTS": json.dumps(["algo-1"]), "SM_CURRENT_HOST": "algo-1", "SM_MODEL_DIR": model_dir,
                   "SM_CHANNEL_TRAIN": "/tmp"})


class SynthCIFAR:
    def __init__(self, root, train=True, transform=None, download=False):
        n = 960 if train else 64
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
sys.argv = [script, "--backend", "smddp", "--batch-size", "32", "--epochs", "1", "--lr", "0.01", "--momentum", "0.9",
            "--model-type", model_type, "--log-interval", "10", "--test-batch-size", "32"]
runpy.run_path(script, run_name="__main__")
import b200ddp as bd  # noqa: E402
from b200ddp.ops import _lib as L  # noqa: E402
from b200ddp.runtime.graphed import STATS  # noqa: E402
print("NATIVE_LAUNCHES %d backend=%s dtype=%s" % (L.LAUNCHES[0], bd.get_backend(), bd.get_compute_dtype()), flush=True)
print("GRAPHED " + json.dumps(STATS), flush=True)
bd.destroy_process_group()
