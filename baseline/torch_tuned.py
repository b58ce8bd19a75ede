"""Like-for-like comparison arm: STOCK PyTorch on the SAME input pipeline as the product.

The unmodified reference script (``baseline/run_reference.py``) spends its step in a single-threaded PIL
loader (``notebooks/code/cifar10-distributed-smddp-gpu.py:55-62,71-88``, ``num_workers=0``), so its ratio to
the product mostly measures the loader.  This arm isolates compute + communication: the reference's model,
loss, optimizer and parallelism (``torchvision.models.resnet18()`` with the 1000-way head, ``nn.CrossEntropyLoss``,
``optim.SGD(lr, momentum)``, ``torch.nn.parallel.DistributedDataParallel`` over NCCL —
``…smddp-gpu.py:30-32,148,155-158,162-168``) run on cuDNN/cuBLAS/ATen kernels, but fed the way the product is
fed: GPU-resident uint8 NHWC batches, RandomCrop(32, 4) + flip + Normalize done with torch ops on the device.

Modes (all: global batch / world per rank, same step count, CUDA-event timing, L2 flushed between timed steps,
max over ranks):
  ``fp32``         fp32 weights and activations, torch defaults (cuDNN may use TF32 for convolutions), eager
  ``fp32_graph``   the same step captured in a CUDA graph, channels_last
  ``bf16``         ``torch.autocast(bfloat16)`` + channels_last, eager
  ``bf16_graph``   ``torch.autocast(bfloat16)`` + channels_last + CUDA graph (the tuned-PyTorch ceiling on one GPU; with DDP
                   over several GPUs the captured step measured SLOWER than eager on this box, so both are reported)

None of this repo's kernels, models or engine is on this path (only ``torch`` / ``torchvision``).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

MEAN = (0.4914, 0.4822, 0.4465)
STD = (0.2023, 0.1994, 0.2010)


def _augment(x_u8: torch.Tensor, off: torch.Tensor, mean: torch.Tensor, inv_std: torch.Tensor, pad: int = 4) -> torch.Tensor:
    """uint8 [B,H,W,C] + per-image (dy, dx, flip) -> normalised float NCHW, with torch ops only (graph-capturable)."""
    B, H, W, C = x_u8.shape
    xp = F.pad(x_u8.permute(0, 3, 1, 2), (pad, pad, pad, pad))                 # [B,C,H+2p,W+2p] uint8
    ar_h = torch.arange(H, device=x_u8.device)
    ar_w = torch.arange(W, device=x_u8.device)
    rows = off[:, 0:1].long() + ar_h[None, :]                                   # [B,H]
    cols = off[:, 1:2].long() + ar_w[None, :]                                   # [B,W]
    cols = torch.where(off[:, 2:3] != 0, cols.flip(1), cols)
    b = torch.arange(B, device=x_u8.device)[:, None, None]
    out = xp[b, :, rows[:, :, None], cols[:, None, :]]                          # [B,H,W,C]
    out = out.permute(0, 3, 1, 2).float().div_(255.0)
    return (out - mean[None, :, None, None]) * inv_std[None, :, None, None]


def _build(model_name: str, hw: int):
    import torchvision
    if model_name == "resnet18":
        return torchvision.models.resnet18()
    if model_name == "resnet50":
        return torchvision.models.resnet50()
    raise ValueError("torch_tuned arm knows resnet18 and resnet50, got %r" % model_name)


def run_mode(mode: str, model_name: str, per_rank: int, hw: int, steps: int, warmup: int, xs, ys, offs,
             group=None, world: int = 1, device: Optional[torch.device] = None) -> Dict:
    """Time ``steps`` training steps of stock torch in ``mode``.  ``xs/ys/offs``: pinned host tensors
    [nb, per_rank, ...] with nb >= warmup + steps + 1.  Returns per-step device ms and end-to-end ms."""
    dev = device or torch.device("cuda", torch.cuda.current_device())
    use_graph = mode.endswith("_graph")
    bf16 = mode.startswith("bf16")
    torch.manual_seed(1)
    model = _build(model_name, hw).to(dev)
    if use_graph or bf16:
        model = model.to(memory_format=torch.channels_last)
    if world > 1:
        model = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], process_group=group)
    crit = nn.CrossEntropyLoss()
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    mean = torch.tensor(MEAN, device=dev)
    inv_std = 1.0 / torch.tensor(STD, device=dev)
    sx = torch.zeros((per_rank, hw, hw, 3), dtype=torch.uint8, device=dev)
    sy = torch.zeros((per_rank,), dtype=torch.int64, device=dev)
    so = torch.zeros((per_rank, 3), dtype=torch.int32, device=dev)
    sloss = torch.zeros((), dtype=torch.float32, device=dev)

    def body():
        inp = _augment(sx, so, mean, inv_std)
        if use_graph or bf16:
            inp = inp.contiguous(memory_format=torch.channels_last)
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf16):
            out = model(inp)
            loss = crit(out.float(), sy)
        loss.backward()
        opt.step()
        sloss.copy_(loss.detach())

    def load(j):
        sx.copy_(xs[j], non_blocking=True)
        sy.copy_(ys[j], non_blocking=True)
        so.copy_(offs[j], non_blocking=True)

    graph = None
    note = ""
    if use_graph:
        try:
            s = torch.cuda.Stream(device=dev)
            s.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(s):
                load(0)
                for _ in range(11):                     # DDP needs >= 11 side-stream iterations before capture
                    body()
            torch.cuda.current_stream(dev).wait_stream(s)
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                body()
        except Exception as e:          # report, fall back to eager
            graph = None
            note = "graph capture failed (%s: %s); eager" % (type(e).__name__, str(e)[:120])
            torch.cuda.synchronize(dev)

    def step():
        if graph is not None:
            graph.replay()
        else:
            body()

    for i in range(warmup):
        load(i)
        step()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier(group=group)
    torch.cuda.synchronize(dev)
    evs = []
    for i in range(steps):
        load(warmup + i)
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize(dev)
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    # end to end: pinned H2D of every batch + D2H of every loss inside the timed region
    if world > 1:
        dist.barrier(group=group)
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    last = None
    for i in range(steps):
        load(warmup + i)
        step()
        last = float(sloss.item())
    b.record()
    torch.cuda.synchronize(dev)
    e2e_ms = a.elapsed_time(b)
    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        dev_ms, e2e_ms = float(t[0]), float(t[1])
    imgs = steps * per_rank * world
    out = {"images_per_sec": imgs / (dev_ms / 1e3), "ms_per_step": dev_ms / steps,
           "e2e_images_per_sec": imgs / (e2e_ms / 1e3), "e2e_ms_per_step": e2e_ms / steps,
           "cuda_graph": graph is not None, "final_loss": last}
    if note:
        out["note"] = note
    del graph, model, opt
    torch.cuda.empty_cache()
    return out


def run_all(model_name: str, per_rank: int, hw: int, steps: int, warmup: int, xs, ys, offs, world: int,
            device: torch.device, modes=("fp32", "fp32_graph", "bf16", "bf16_graph")) -> Dict:
    """Every mode, each inside its own try block: the comparison must never take the product's bench line down."""
    group = None
    if world > 1:
        group = dist.new_group(backend="nccl", device_id=device) if _new_group_takes_device_id() else dist.new_group(backend="nccl")
    res = {}
    for m in modes:
        try:
            res[m] = run_mode(m, model_name, per_rank, hw, steps, warmup, xs, ys, offs, group, world, device)
        except Exception as e:
            res[m] = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
            torch.cuda.synchronize(device)
    if group is not None:
        try:
            dist.destroy_process_group(group)
        except Exception:
            pass
    return res


def _new_group_takes_device_id() -> bool:
    import inspect
    try:
        return "device_id" in inspect.signature(dist.new_group).parameters
    except (TypeError, ValueError):
        return False


def run(gpus: int, steps: int, warmup: int, global_batch: int = 256, model_name: str = "resnet18",
        modes=("fp32", "fp32_graph", "bf16", "bf16_graph")) -> dict:
    """Stand-alone entry (``bench.py --impl torch_tuned``): own NCCL process group from the torchrun env."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    dev = torch.device("cuda", local % max(torch.cuda.device_count(), 1))
    torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    per_rank = max(global_batch // world, 1)
    hw = 224 if model_name == "resnet50" else 32
    nb = warmup + steps + 1
    g = torch.Generator().manual_seed(100 + rank)
    xs = torch.randint(0, 256, (nb, per_rank, hw, hw, 3), dtype=torch.uint8, generator=g).pin_memory()
    ys = torch.randint(0, 10, (nb, per_rank), dtype=torch.int64, generator=g).pin_memory()
    offs = torch.stack([torch.randint(0, 9, (nb, per_rank), generator=g), torch.randint(0, 9, (nb, per_rank), generator=g),
                        torch.randint(0, 2, (nb, per_rank), generator=g)], -1).to(torch.int32).pin_memory()
    res = {}
    for m in modes:
        try:
            res[m] = run_mode(m, model_name, per_rank, hw, steps, warmup, xs, ys, offs, None, world, dev)
        except Exception as e:
            res[m] = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
    best = max((v for v in res.values() if "images_per_sec" in v), key=lambda v: v["images_per_sec"], default=None)
    if world > 1:
        dist.destroy_process_group()
    head = res.get("fp32", {}) if "images_per_sec" in res.get("fp32", {}) else (best or {})
    return {"impl": "torch_tuned", "rank": rank, "metric": "images/sec", "value": head.get("images_per_sec"),
            "unit": "images/sec", "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": head.get("ms_per_step"),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "e2e": {"value": head.get("e2e_images_per_sec"), "unit": "images/sec",
                    "h2d_bytes_per_step": per_rank * (hw * hw * 3 + 8 + 12), "d2h_bytes_per_step": 4},
            "gpu_launches": 0,
            "config": {"model": "%s (torchvision, 1000-way head, %dx%d input)" % (model_name, hw, hw),
                       "global_batch": per_rank * world, "seq_len": None, "parallelism": "dp%d" % world,
                       "modes": res,
                       "note": "stock torch DDP + NCCL + cuDNN on GPU-resident uint8 batches (the product's input pipeline)"}}
