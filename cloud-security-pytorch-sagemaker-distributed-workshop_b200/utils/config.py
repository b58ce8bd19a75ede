"""Entry-point contract: argparse hyperparameters + SageMaker ``SM_*`` environment.

Flag names, types and defaults follow the reference's two trainers
(``notebooks/code/cifar10-distributed-native-cpu.py:201-230`` and
``notebooks/code/cifar10-distributed-smddp-gpu.py:210-239``).  Differences, on purpose:

* every ``SM_*`` variable has a default, so the script also runs outside a SageMaker
  container (the reference raises ``KeyError`` when one is unset);
* ``--hosts`` takes JSON (``'["algo-1","algo-2"]'``) or a comma list — the reference's
  ``type=list`` would split a CLI string into characters (SURVEY §2.1 defects);
* comm-layer tunables come from ``B200DDP_*`` environment variables.
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import os
import tempfile
from typing import List, Optional


def _parse_hosts(text) -> List[str]:
    if isinstance(text, (list, tuple)):
        return list(text)
    text = text.strip()
    if text.startswith("["):
        return list(json.loads(text))
    return [h for h in text.split(",") if h]


def _env_hosts() -> List[str]:
    return _parse_hosts(os.environ.get("SM_HOSTS", '["algo-1"]'))


@dataclasses.dataclass
class TrainConfig:
    """Hyperparameters + environment of one training job (one per rank)."""

    model_type: str = "resnet18"       # 'custom' -> workshop Net, 'resnet18', 'resnet50'
    batch_size: int = 64
    test_batch_size: int = 1000
    epochs: int = 10
    lr: float = 0.01
    momentum: float = 0.5
    seed: int = 1
    log_interval: int = 100
    backend: str = "b200"              # 'b200' (alias 'smddp'), 'gloo', 'nccl' (baseline only)
    hosts: List[str] = dataclasses.field(default_factory=_env_hosts)
    current_host: str = dataclasses.field(
        default_factory=lambda: os.environ.get("SM_CURRENT_HOST", "algo-1"))
    model_dir: str = dataclasses.field(
        default_factory=lambda: os.environ.get("SM_MODEL_DIR", os.path.join(tempfile.gettempdir(), "model")))
    data_dir: str = dataclasses.field(
        default_factory=lambda: os.environ.get("SM_CHANNEL_TRAIN", ""))
    # extensions beyond the reference
    divide_batch: Optional[bool] = None  # None: GPU path divides (smddp-gpu.py:123), CPU path does not
    synthetic: bool = False              # force synthetic CIFAR-10 even when data_dir has files
    train_size: int = 50000
    test_size: int = 10000
    dtype: str = "fp32"                  # compute dtype of the native kernels: 'fp32' (fp32 storage + TF32 tensor cores, the reference's precision) | 'bf16'
    graph: bool = True                   # capture the train step in a CUDA graph
    resume: str = ""                     # optional full checkpoint (model+optimizer+epoch)
    save_full_checkpoint: bool = False
    max_steps: int = 0                   # 0 = run whole epochs
    trace: str = ""                      # write a Chrome-trace JSON of the training loop here (rank is appended)
    lr_step_epochs: int = 0              # > 0: multiply the rate by lr_gamma every that many epochs (the reference trains at a constant rate)
    lr_gamma: float = 0.1
    eval: bool = True

    @property
    def world_hosts(self) -> int:
        return len(self.hosts)


def add_reference_arguments(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    """The reference's flags, verbatim names (hyphenated)."""
    parser.add_argument("--model-type", type=str, default="resnet18", help="custom model, resnet18 or resnet50")
    parser.add_argument("--batch-size", type=int, default=64, help="input batch size for training (default: 64)")
    parser.add_argument("--test-batch-size", type=int, default=1000, help="input batch size for testing (default: 1000)")
    parser.add_argument("--epochs", type=int, default=10, help="number of epochs to train (default: 10)")
    parser.add_argument("--lr", type=float, default=0.01, help="learning rate (default: 0.01)")
    parser.add_argument("--momentum", type=float, default=0.5, help="SGD momentum (default: 0.5)")
    parser.add_argument("--seed", type=int, default=1, help="random seed (default: 1)")
    parser.add_argument("--log-interval", type=int, default=100, help="how many batches to wait before logging")
    parser.add_argument("--backend", type=str, default=None, help="b200 | smddp (alias) | gloo | nccl")
    parser.add_argument("--hosts", type=_parse_hosts, default=_env_hosts())
    parser.add_argument("--current-host", type=str, default=os.environ.get("SM_CURRENT_HOST", "algo-1"))
    parser.add_argument("--model-dir", type=str,
                        default=os.environ.get("SM_MODEL_DIR", os.path.join(tempfile.gettempdir(), "model")))
    parser.add_argument("--data-dir", type=str, default=os.environ.get("SM_CHANNEL_TRAIN", ""))
    return parser


def add_extension_arguments(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    parser.add_argument("--divide-batch", dest="divide_batch", action="store_true", default=None)
    parser.add_argument("--no-divide-batch", dest="divide_batch", action="store_false")
    parser.add_argument("--synthetic", action="store_true")
    parser.add_argument("--train-size", type=int, default=50000)
    parser.add_argument("--test-size", type=int, default=10000)
    parser.add_argument("--dtype", type=str, default="fp32", choices=["bf16", "fp32"])
    parser.add_argument("--no-graph", dest="graph", action="store_false", default=True)
    parser.add_argument("--resume", type=str, default="")
    parser.add_argument("--save-full-checkpoint", action="store_true")
    parser.add_argument("--max-steps", type=int, default=0)
    parser.add_argument("--no-eval", dest="eval", action="store_false", default=True)
    parser.add_argument("--trace", type=str, default="", help="Chrome-trace JSON of the loop (host + device ranges), per rank")
    parser.add_argument("--lr-step-epochs", type=int, default=0, help="multiply the learning rate by --lr-gamma every N epochs (0 = constant)")
    parser.add_argument("--lr-gamma", type=float, default=0.1)
    return parser


def parse_args(argv=None) -> TrainConfig:
    parser = argparse.ArgumentParser(description="b200ddp CIFAR-10 trainer (SageMaker entry-point contract)")
    add_reference_arguments(parser)
    add_extension_arguments(parser)
    ns = parser.parse_args(argv)
    cfg = TrainConfig(**{f.name: getattr(ns, f.name) for f in dataclasses.fields(TrainConfig)
                         if hasattr(ns, f.name) and getattr(ns, f.name) is not None})
    if ns.backend is None:
        import torch
        cfg.backend = "b200" if torch.cuda.is_available() else "gloo"
    return cfg


# --------------------------------------------------------------------------------------
# comm-layer tunables (single backend; these only pick the algorithm per bucket size)
# --------------------------------------------------------------------------------------
def env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return default if v is None or v == "" else int(v)


def env_flag(name: str, default: bool) -> bool:
    v = os.environ.get(name)
    if v is None or v == "":
        return default
    return v.lower() not in ("0", "false", "off", "no")


@dataclasses.dataclass
class CommConfig:
    bucket_cap_mb: float = float(os.environ.get("B200DDP_BUCKET_CAP_MB", "25"))
    first_bucket_mb: float = float(os.environ.get("B200DDP_FIRST_BUCKET_MB", "1"))
    oneshot_max_bytes: int = env_int("B200DDP_ONESHOT_MAX_BYTES", 64 * 1024)
    nvls_min_bytes: int = env_int("B200DDP_NVLS_MIN_BYTES", 64 * 1024)
    use_nvls: bool = env_flag("B200DDP_NVLS", True)
    max_ctas: int = env_int("B200DDP_COMM_CTAS", 64)          # stand-alone collectives: whatever gives peak bandwidth
    # Bucket all-reduces that overlap the backward pass get a small SM budget so the compute kernels keep the machine:
    # resnet18, global batch 256 on 4 x B200: 64 / 24 / 8 CTAs -> 0.943 / 0.914 / 0.877 ms per step.
    overlap_ctas: int = env_int("B200DDP_OVERLAP_CTAS", 8)
    spin_timeout_ms: int = env_int("B200DDP_SPIN_TIMEOUT_MS", 20000)
    debug: bool = os.environ.get("B200DDP_DEBUG", "").upper() in ("INFO", "1", "TRACE")
