"""Reference arm: run the UNMODIFIED ``cifar10-distributed-smddp-gpu.py`` staged in ``baseline/_ref``
through its own CLI and stock code path (its model, its PIL transforms + DataLoader, torch DDP, its
loop), and time K optimizer steps of it on the device.

Only the environment around the script is supplied, exactly what a SageMaker container supplies:
  * ``SM_HOSTS / SM_CURRENT_HOST / SM_MODEL_DIR / SM_CHANNEL_TRAIN`` and the hyperparameter argv;
  * the ``smdistributed.dataparallel.torch.torch_smddp`` import + ``backend='smddp'``: SMDDP is closed
    source and not installable offline; in single-node mode it defers to NCCL
    (notebooks/2_pytorch_dist_smddp_gpu.ipynb:288,296), so the stand-in module maps ``'smddp'`` to
    torch's NCCL backend and pins the device from ``LOCAL_RANK`` as the library does;
  * ``torchvision.datasets.CIFAR10``: no network and no dataset on the box, so a synthetic dataset of
    the same shape/dtype (PIL 32x32 RGB images) is served through the same class name.
None of this repo's models, kernels or engine is on the path.  Timing: CUDA events recorded from a
wrapper around ``torch.optim.SGD.step`` (the script's own optimizer), warm-up steps excluded, max over
ranks.

``run(..., on_b200=True)`` (``bench.py --impl reference_on_b200``) is the other way round: the same unmodified script
and the same environment, but ``smdistributed.dataparallel.torch.torch_smddp`` is the module this repository ships, so
``backend='smddp'`` and ``torch.nn.parallel.DistributedDataParallel`` land on b200ddp (native kernels, NVLink
all-reduce) while the script's own loader, loop and ``optim.SGD`` stay as they are: "same script, new framework".
"""
from __future__ import annotations

import json
import os
import runpy
import sys
import tempfile
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_CODE = os.path.join(HERE, "_ref", "code")


def _install_smddp_standin():
    import torch
    import torch.distributed as dist

    pkg = types.ModuleType("smdistributed")
    dp = types.ModuleType("smdistributed.dataparallel")
    tp = types.ModuleType("smdistributed.dataparallel.torch")
    ts = types.ModuleType("smdistributed.dataparallel.torch.torch_smddp")
    pkg.dataparallel, dp.torch, tp.torch_smddp = dp, tp, ts
    for name, mod in (("smdistributed", pkg), ("smdistributed.dataparallel", dp),
                      ("smdistributed.dataparallel.torch", tp),
                      ("smdistributed.dataparallel.torch.torch_smddp", ts)):
        sys.modules[name] = mod
    real_init = dist.init_process_group

    def init_process_group(backend=None, *a, **kw):
        if backend == "smddp":
            backend = "nccl"
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        return real_init(backend, *a, **kw)

    dist.init_process_group = init_process_group


def _install_synthetic_cifar(train_size: int, test_size: int):
    import numpy as np
    import torchvision
    from PIL import Image

    class SyntheticCIFAR10PIL:
        def __init__(self, root, train=True, transform=None, target_transform=None, download=False):
            n = train_size if train else test_size
            rng = np.random.RandomState(1 if train else 2)
            self.data = rng.randint(0, 256, size=(n, 32, 32, 3), dtype=np.uint8)
            self.targets = rng.randint(0, 10, size=n).tolist()
            self.transform = transform

        def __len__(self):
            return len(self.targets)

        def __getitem__(self, i):
            img = Image.fromarray(self.data[i])
            if self.transform is not None:
                img = self.transform(img)
            return img, self.targets[i]

    torchvision.datasets.CIFAR10 = SyntheticCIFAR10PIL


class _StepTimer:
    def __init__(self, warmup: int, steps: int):
        self.warmup, self.steps, self.count = warmup, steps, 0
        self.start = self.end = None

    def install(self):
        import torch
        import torch.distributed as dist
        orig = torch.optim.SGD.step
        timer = self

        def step(opt_self, *a, **kw):
            out = orig(opt_self, *a, **kw)
            timer.count += 1
            if timer.count == timer.warmup:
                torch.cuda.synchronize()
                if dist.is_initialized():
                    dist.barrier()
                torch.cuda.synchronize()
                timer.start = torch.cuda.Event(enable_timing=True)
                timer.start.record()
            elif timer.count == timer.warmup + timer.steps:
                timer.end = torch.cuda.Event(enable_timing=True)
                timer.end.record()
                torch.cuda.synchronize()
            return out

        torch.optim.SGD.step = step


def run(gpus: int, steps: int, warmup: int, global_batch: int = 256, model_type: str = "resnet18",
        on_b200: bool = False) -> dict:
    import torch
    import torch.distributed as dist

    impl = "reference_on_b200" if on_b200 else "reference"
    script = os.path.join(REF_CODE, "cifar10-distributed-smddp-gpu.py")
    if not os.path.exists(script):
        return {"impl": impl, "unavailable": "baseline/_ref not staged (run baseline/install_reference.py)"}
    if not torch.cuda.is_available():
        return {"impl": impl, "unavailable": "no CUDA device"}
    if model_type not in ("resnet18", "custom"):
        # the script's _get_model knows 'resnet18' and falls back to its inline Net for ANY other string
        # (cifar10-distributed-smddp-gpu.py:30-52): refuse instead of silently timing the wrong model
        return {"impl": impl, "unavailable": "the reference script has no %r model (only resnet18 and its inline Net); "
                "the stock-PyTorch arm for it is bench.py --impl torch_tuned --model %s" % (model_type, model_type)}
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("LOCAL_RANK", os.environ["RANK"])
    os.environ.setdefault("WORLD_SIZE", str(gpus))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    world = int(os.environ["WORLD_SIZE"])
    model_dir = tempfile.mkdtemp(prefix="ref_model_")
    os.environ.update({"SM_HOSTS": json.dumps(["algo-1"]), "SM_CURRENT_HOST": "algo-1", "SM_MODEL_DIR": model_dir,
                       "SM_CHANNEL_TRAIN": tempfile.gettempdir(), "NCCL_DEBUG": os.environ.get("NCCL_DEBUG", "WARN")})
    if on_b200:
        root = os.path.dirname(HERE)
        if root not in sys.path:
            sys.path.insert(0, root)          # `import smdistributed...` resolves to the module this repository ships
    else:
        _install_smddp_standin()
    _install_synthetic_cifar(train_size=(warmup + steps) * global_batch, test_size=world * 8)
    timer = _StepTimer(warmup, steps)
    timer.install()
    argv = [script, "--backend", "smddp", "--batch-size", str(global_batch), "--epochs", "1", "--lr", "0.01",
            "--model-type", model_type, "--momentum", "0.9", "--test-batch-size", "8", "--log-interval", "100"]
    old_argv, old_path = sys.argv, list(sys.path)
    sys.argv = argv
    sys.path.insert(0, REF_CODE)
    sys.dont_write_bytecode = True
    try:
        runpy.run_path(script, run_name="__main__")
    finally:
        sys.argv, sys.path = old_argv, old_path
    if timer.start is None or timer.end is None:
        return {"impl": impl, "unavailable": "reference ran %d steps, fewer than warmup+steps" % timer.count}
    ms = timer.start.elapsed_time(timer.end)
    native = []
    if on_b200:
        import b200ddp as bd
        from b200ddp.ops import _lib as L
        ms = bd.max_over_ranks(ms)
        rank = bd.get_rank()
        native = [L.LAUNCHES[0]]
        from b200ddp.runtime.graphed import STATS as graph_stats
        comm = bd.get_comm()
        if comm is not None:
            comm.check_health()
        bd.destroy_process_group()
    elif dist.is_initialized():
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        rank = dist.get_rank()
        dist.destroy_process_group()
    else:
        rank = 0
    value = steps * global_batch / (ms / 1e3)
    label = {"resnet18": "resnet18 (torchvision, 1000-way head, 32x32 input)",
             "custom": "custom (the script's inline LeNet-style Net, 10 classes, 32x32 input)"}[model_type]
    note = ("unmodified reference script on b200ddp through the shipped smdistributed.dataparallel.torch.torch_smddp module: native "
            "kernels + NVLink all-reduce under the script's own PIL loader, loop and optim.SGD (loader inside the timed region)"
            if on_b200 else
            "unmodified reference script; torch DDP+NCCL; PIL dataloader num_workers=0 inside the timed region")
    extra = {"graphed_forward_backward": dict(graph_stats)} if on_b200 else {}
    return {"impl": impl, "rank": rank, "metric": "images/sec", "value": value, "unit": "images/sec",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "e2e": {"value": value, "unit": "images/sec",
                    "h2d_bytes_per_step": (global_batch // world) * (3 * 32 * 32 * 4 + 8), "d2h_bytes_per_step": 0},
            "gpu_launches": native[0] if native else 0,
            "config": {"model": label, "global_batch": global_batch,
                       "seq_len": None, "parallelism": "dp%d" % world, "note": note, **extra}}
