#!/usr/bin/env python
"""Headline benchmark: images/sec of the workshop's GPU training job (BASELINE.json / BASELINE.md §1:
torchvision-topology resnet18 with the 1000-way head on 32x32 CIFAR-10-shaped data, global batch 256
divided by the world size as ``cifar10-distributed-smddp-gpu.py:123`` does, SGD lr 0.01 momentum 0.9).

    python bench.py --gpus N --steps K --warmup W            # this framework (backend 'b200'), fp32 storage + TF32 tensor cores
    python bench.py --dtype bf16 ...                         # the bf16 tensor-core path (secondary)
    python bench.py --impl reference --gpus N ...            # unmodified reference script, torch DDP + NCCL + PIL loader
    python bench.py --impl torch_tuned --gpus N ...          # stock torch DDP + NCCL + cuDNN on the product's GPU-resident loader
    python bench.py --impl reference_on_b200 --gpus N ...    # unmodified reference script running ON this framework ('smddp' shim)

One full training step = GPU augmentation of the uint8 batch, forward, loss, backward with the fused
NVLink gradient all-reduce, fused SGD.  Two numbers are reported:
  value      device-timed steps (CUDA events around each CUDA-graph replay, L2 flushed between timed
             iterations, max over ranks);
  e2e.value  the same K steps through the public ``TrainStep.pipeline`` API including, every step, the
             H2D copy of that step's uint8 images/labels/crop offsets from pinned host memory and the
             D2H read of the loss.
``config.like_for_like`` carries the stock-PyTorch numbers measured in the same process on the same input
pipeline (``baseline/torch_tuned.py``), ``config.ddp_check`` the gradient-synchronisation check run before the
timed region at N > 1, ``config.secondary`` the workshop ``Net`` (BASELINE.json config 2) at the same N.
Synthetic data, random-init weights (no network on the box).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _time_steps(bd, torch, step, xs, ys, warmup, steps, flush, clocks_dev):
    """(device ms, e2e ms, last loss) of `steps` steps after `warmup`; collective (barriers)."""
    for i in range(warmup):                     # warm-up (captures the graph on the first call)
        step(xs[i], ys[i], need_loss=(i == warmup - 1))
    bd.barrier()
    torch.cuda.synchronize()
    evs = []
    for i in range(steps):
        j = warmup + i
        step.static_x.copy_(xs[j], non_blocking=True)
        step.static_y.copy_(ys[j], non_blocking=True)
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        if step._graph is not None:
            step._graph.replay()
        else:
            step._body(step.static_x, step.static_y, step.static_off)
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    bd.barrier()
    torch.cuda.synchronize()
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    # end to end through the public API: pinned H2D inputs + D2H loss every step
    bd.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    last = None
    batches = ((xs[warmup + i], ys[warmup + i]) for i in range(steps))
    for last in step.pipeline(batches, need_loss=True):      # H2D of batch k+1 overlaps step k; every loss is read back
        pass
    b.record()
    torch.cuda.synchronize()
    e2e_ms = a.elapsed_time(b)
    bd.barrier()
    return bd.max_over_ranks(dev_ms), bd.max_over_ranks(e2e_ms), last


def run_own(args) -> dict:
    import torch
    import b200ddp as bd
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import _lib as L
    from b200ddp.ops import native as nv
    from b200ddp.ops import optim as boptim
    from b200ddp.runtime.engine import TrainStep
    from b200ddp.utils.ddp_check import grad_sync_check, params_identical
    from b200ddp.utils.timers import ClockSampler

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    bd.init_process_group(backend="b200", rank=rank, world_size=world)
    dev = bd.get_device()
    torch.manual_seed(1)
    per_rank = max(args.global_batch // world, 1)
    hw = 224 if args.model == "resnet50" else 32
    dtype = "fp32" if args.model == "custom" else args.dtype
    bd.set_compute_dtype(dtype)
    model = bd.DistributedDataParallel(get_model(args.model).to(dev))
    opt = boptim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    crit = CrossEntropyLoss()

    # host-side synthetic dataset in pinned memory: (W + K + 1) distinct batches
    nb = args.warmup + args.steps + 1
    g = torch.Generator().manual_seed(100 + rank)
    xs = torch.randint(0, 256, (nb, per_rank, hw, hw, 3), dtype=torch.uint8, generator=g).pin_memory()
    ys = torch.randint(0, 10, (nb, per_rank), dtype=torch.int64, generator=g).pin_memory()

    # ---- gradient-synchronisation check (N > 1): synced grads == mean of the ranks' local grads, bitwise equal across ranks
    ddp_check = None
    if world > 1 and args.model != "custom":
        mean_t = torch.tensor((0.4914, 0.4822, 0.4465), device=dev)
        inv_t = 1.0 / torch.tensor((0.2023, 0.1994, 0.2010), device=dev)
        xd, yd = xs[0].to(dev), ys[0].to(dev)

        def fwd_bwd():
            crit(model(nv.augment(xd, None, mean_t, inv_t, c_out=8)), yd).backward()

        model.train()
        ddp_check = grad_sync_check(model, fwd_bwd)
        ddp_check.pop("per_bucket_rel_err", None)

    step = TrainStep(model, crit, opt, per_rank, image_shape=(hw, hw, 3), device=dev, graph=not args.no_graph, seed=rank)
    # ---- count this framework's kernel launches in one eager step (graph replays launch the same set)
    model.train()
    c0 = L.LAUNCHES[0]
    step._body(xs[0].to(dev), ys[0].to(dev), step._draw_offsets(per_rank).to(dev))
    launches_per_step = L.LAUNCHES[0] - c0
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    with ClockSampler(dev.index or 0) as clocks:
        dev_ms, e2e_ms, last = _time_steps(bd, torch, step, xs, ys, args.warmup, args.steps, flush, dev)
        if ddp_check is not None:
            ddp_check["params_identical_after_bench"] = params_identical(model)
        try:        # where an (eager) step spends its time: forward / backward / exposed collective tail / optimizer
            phase_ms = step.breakdown(xs[0], ys[0], iters=5)
        except Exception as e:
            phase_ms = {"error": "%s: %s" % (type(e).__name__, str(e)[:160])}
        comm = bd.get_comm()
        if comm is not None:
            comm.check_health()
        algos = sorted(set(comm.launch_log)) if comm is not None else []

        # ---- secondary: the workshop Net (BASELINE.json config 2) at the same N, same protocol
        secondary = {}
        if not args.no_secondary and args.model == "resnet18":
            try:
                bd.set_compute_dtype("fp32")
                net = bd.DistributedDataParallel(get_model("custom").to(dev))
                nopt = boptim.SGD(net.parameters(), lr=0.01, momentum=0.9)
                nstep = TrainStep(net, CrossEntropyLoss(), nopt, per_rank, image_shape=(32, 32, 3), device=dev,
                                  graph=not args.no_graph, seed=rank)
                n_dev, n_e2e, n_last = _time_steps(bd, torch, nstep, xs, ys, args.warmup, args.steps, flush, dev)
                n_imgs = args.steps * per_rank * world
                secondary["custom"] = {"model": "workshop Net (62,006 params), fp32 megakernel", "value": n_imgs / (n_dev / 1e3),
                                       "e2e": n_imgs / (n_e2e / 1e3), "ms_per_step": n_dev / args.steps, "final_loss": n_last,
                                       "params_identical": params_identical(net)}
                del nstep, nopt, net
            except Exception as e:          # secondary numbers must never take the headline down
                secondary["custom"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
            bd.set_compute_dtype(dtype)

        # ---- like for like: stock torch DDP + NCCL + cuDNN on the same GPU-resident uint8 batches
        lfl = None
        if not args.no_like_for_like and args.model != "custom":
            try:
                sys.path.insert(0, os.path.join(ROOT, "baseline"))
                import torch_tuned
                offs = torch.stack([torch.randint(0, 9, (nb, per_rank), generator=g), torch.randint(0, 9, (nb, per_rank), generator=g),
                                    torch.randint(0, 2, (nb, per_rank), generator=g)], -1).to(torch.int32).pin_memory()
                lfl = torch_tuned.run_all(args.model, per_rank, hw, args.steps, args.warmup, xs, ys, offs, world, dev,
                                          modes=tuple(m for m in args.lfl_modes.split(",") if m))
            except Exception as e:
                lfl = {"error": "%s: %s" % (type(e).__name__, str(e)[:200])}
    imgs = args.steps * per_rank * world
    value = imgs / (dev_ms / 1e3)
    e2e_value = imgs / (e2e_ms / 1e3)
    if isinstance(lfl, dict):
        ratios = {}
        for m, r in lfl.items():
            if isinstance(r, dict) and r.get("images_per_sec"):
                ratios[m] = {"device": value / r["images_per_sec"], "e2e": e2e_value / r["e2e_images_per_sec"]}
        lfl["ratio_product_over_torch"] = ratios
    out = {
        "impl": "b200ddp", "metric": "images/sec", "value": value, "unit": "images/sec", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
        "e2e": {"value": e2e_value, "unit": "images/sec",
                "h2d_bytes_per_step": per_rank * (hw * hw * 3 + 8 + 12), "d2h_bytes_per_step": 4,
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks.summary(),
        "config": {"model": ("custom (the workshop's LeNet-style Net, 10 classes, %dx%d input; fp32 megakernel)" if args.model == "custom"
                             else args.model + " (torchvision topology, 1000-way head, %dx%d input)") % (hw, hw),
                   "precision": ("fp32 parameters/activations/gradients, TF32 tensor-core products (tcgen05 kind::tf32), fp32 accumulation"
                                 if dtype == "fp32" and args.model != "custom" else
                                 "fp32 CUDA-core megakernel" if args.model == "custom" else
                                 "bf16 activations/weights on tcgen05 kind::f16, fp32 accumulation, fp32 master weights and gradients"),
                   "e2e_note": "TrainStep.pipeline: per step pinned H2D of pixels+labels+crop offsets on a copy stream (overlapping the "
                               "previous step), one graph replay, async D2H of the loss, read by the host one step later",
                   "global_batch": per_rank * world, "seq_len": None, "parallelism": "dp%d" % world,
                   "per_gpu_batch": per_rank, "optimizer": "SGD lr=0.01 momentum=0.9", "cuda_graph": step._graph is not None,
                   "l2": "256 MiB buffer written between timed iterations (outside the event pair)",
                   "launches_per_step": launches_per_step, "allreduce_algos": algos, "final_loss": last,
                   "ddp_check": ddp_check, "phase_ms_eager": phase_ms, "fused_optimizer": bool(getattr(step, "fused_optimizer", False)),
                   "secondary": secondary, "like_for_like": lfl,
                   "baseline_note": "the reference publishes no number (BASELINE.md); vs_baseline is null"},
    }
    bd.destroy_process_group()
    return out if rank == 0 else {}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", type=str, default="b200ddp", choices=["b200ddp", "reference", "torch_tuned", "reference_on_b200"])
    ap.add_argument("--model", type=str, default="resnet18", choices=["resnet18", "resnet50", "custom"])
    ap.add_argument("--dtype", type=str, default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--global-batch", type=int, default=256)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-like-for-like", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--lfl-modes", type=str, default="fp32,fp32_graph,bf16,bf16_graph", help="stock-PyTorch modes to measure alongside")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl in ("reference", "reference_on_b200", "torch_tuned"):
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        try:
            if args.impl == "torch_tuned":
                from torch_tuned import run as run_tuned
                out = run_tuned(args.gpus, args.steps, args.warmup, args.global_batch, args.model)
            else:
                from run_reference import run
                out = run(args.gpus, args.steps, args.warmup, args.global_batch, args.model,
                          on_b200=(args.impl == "reference_on_b200"))
        except Exception as e:    # a comparison arm must never take the harness down
            out = {"impl": args.impl, "unavailable": "%s: %s" % (type(e).__name__, str(e)[:200])}
        if out.get("rank", 0) == 0:
            out.pop("rank", None)
            print(json.dumps(out), flush=True)
        return
    out = run_own(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
