"""CIFAR-10 data-parallel trainer — one script with the semantics of both reference entry points:
``notebooks/code/cifar10-distributed-native-cpu.py`` (CPU, gloo, one process per host, rank from
``SM_HOSTS``/``SM_CURRENT_HOST``) and ``notebooks/code/cifar10-distributed-smddp-gpu.py`` (one process per
GPU, global batch divided by the world size, ``LOCAL_RANK`` device pinning).

Same CLI flags / ``SM_*`` environment / log-line formats / ``model.pth`` output (``utils/config.py``,
``utils/logging.py``, ``utils/checkpoint.py``).  Known reference defects are not inherited:
gradients are averaged once, ``--model-type`` is honoured on CPU too, the checkpoint is un-prefixed and
servable by ``inference.py``, and the true cross-entropy is reported next to the reference's
``nll_loss``-on-logits figure.

Run:  ``python -m b200ddp.workloads.cifar10_train --model-type resnet18 --batch-size 256 --epochs 15``
(under ``torchrun --nproc-per-node N`` for N GPUs, or through ``python -m b200ddp.utils.launch``).
"""
from __future__ import annotations

import os
import time

import torch

from .. import _api as bd
from ..models import get_model
from ..nn import CrossEntropyLoss
from ..ops import optim as boptim
from ..runtime.engine import TrainStep
from ..utils import checkpoint as ckpt
from ..utils.config import TrainConfig, parse_args
from ..utils.data import BatchLoader, ShardSampler, augment_reference, load_cifar10
from ..utils.logging import fmt_shard_line, fmt_test_line, fmt_train_line, get_logger

logger = get_logger("b200ddp")


def _resolve_world(cfg: TrainConfig):
    """(rank, world): torchrun env wins (GPU path); else the SageMaker host list (CPU path,
    …native-cpu.py:96-109)."""
    if "WORLD_SIZE" in os.environ and "RANK" in os.environ:
        return int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if len(cfg.hosts) > 1 and cfg.backend is not None:
        world = len(cfg.hosts)
        rank = cfg.hosts.index(cfg.current_host)
        os.environ["WORLD_SIZE"] = str(world)
        os.environ["RANK"] = str(rank)
        return rank, world
    return 0, 1


def evaluate(model, loader, device, num_classes_logits=None):
    """Reference ``test()`` (…native-cpu.py:177-194): full test set on every rank, the augmenting
    transform applied to test images too.  Returns (reference 'Average loss', accuracy, true CE)."""
    from ..ops import functional as BF
    model.eval()
    n_total = len(loader.dataset)
    cuda = device.type == "cuda"
    with torch.no_grad():
        if cuda:
            from ..ops import native as nv
            correct = torch.zeros((), dtype=torch.int32, device=device)
            nll = torch.zeros((), dtype=torch.float32, device=device)
            ce_sum = torch.zeros((), dtype=torch.float32, device=device)
            mean = torch.tensor(loader.mean, dtype=torch.float32, device=device)
            inv_std = 1.0 / torch.tensor(loader.std, dtype=torch.float32, device=device)
            g = torch.Generator().manual_seed(1234)
            for x_u8, y in loader:
                n = x_u8.shape[0]
                off = torch.stack([torch.randint(0, 9, (n,), generator=g), torch.randint(0, 9, (n,), generator=g),
                                   torch.randint(0, 2, (n,), generator=g)], 1).to(torch.int32).to(device)
                out = model(nv.augment(x_u8, off, mean, inv_std, c_out=8)).contiguous()
                nv.eval_count(out, y, correct, nll)
                loss, _ = nv.softmax_xent(out, y, 0.0, want_grad=False)
                ce_sum += loss * n
            return float(nll.item()) / n_total, int(correct.item()) / n_total, float(ce_sum.item()) / n_total
        correct, nll, ce = 0, 0.0, 0.0
        for x_u8, y in loader:
            out = model(augment_reference(x_u8, mean=loader.mean, std=loader.std))
            nll += float(-out.gather(1, y.view(-1, 1)).sum())
            ce += float(torch.nn.functional.cross_entropy(out, y, reduction="sum"))
            correct += int((out.argmax(1) == y).sum())
        return nll / n_total, correct / n_total, ce / n_total


def train(cfg: TrainConfig):
    rank, world = _resolve_world(cfg)
    is_distributed = world > 1
    logger.debug("Distributed training - {}".format(is_distributed))
    backend = cfg.backend
    use_cuda = backend not in ("gloo",) and torch.cuda.is_available()
    if backend in ("b200", "smddp", "smddp-b200") and not torch.cuda.is_available():
        raise RuntimeError("backend %r needs a GPU; pass --backend gloo for the CPU path" % backend)
    if is_distributed or use_cuda:
        bd.init_process_group(backend=backend, rank=rank, world_size=world)
    device = bd.get_device() if bd.is_initialized() else torch.device("cpu")
    torch.manual_seed(cfg.seed)
    bd.set_compute_dtype(cfg.dtype)      # 'fp32' = the reference's precision (fp32 storage, TF32 tensor cores); 'bf16' optional

    divide = cfg.divide_batch if cfg.divide_batch is not None else use_cuda   # …smddp-gpu.py:123-124
    batch_size = max(cfg.batch_size // world, 1) if divide else cfg.batch_size

    train_set = load_cifar10(cfg.data_dir, True, cfg.synthetic, cfg.train_size)
    test_set = load_cifar10(cfg.data_dir, False, cfg.synthetic, cfg.test_size)
    sampler = ShardSampler(len(train_set), world, rank, shuffle=True, seed=0)    # set_epoch never called (K-D2)
    train_loader = BatchLoader(train_set, batch_size, sampler=sampler, device=device)
    test_loader = BatchLoader(test_set, cfg.test_batch_size, shuffle=True, device=device, seed=cfg.seed)
    from ..utils.data import CIFAR10_MEAN, CIFAR10_STD
    test_loader.mean, test_loader.std = CIFAR10_MEAN, CIFAR10_STD
    logger.debug(fmt_shard_line(len(train_loader.sampler), len(train_loader.dataset), "train"))
    logger.debug(fmt_shard_line(len(test_loader.sampler), len(test_loader.dataset), "test"))

    model = get_model(cfg.model_type).to(device)
    if is_distributed or use_cuda:
        model = bd.DistributedDataParallel(model)
    criterion = CrossEntropyLoss()
    optimizer = boptim.SGD(model.parameters(), lr=cfg.lr, momentum=cfg.momentum)
    start_epoch = 1
    if cfg.resume:
        start_epoch = ckpt.load_full(model, optimizer, cfg.resume) + 1
        logger.info("Resumed from %s at epoch %d", cfg.resume, start_epoch)

    step = TrainStep(model, criterion, optimizer, batch_size, device=device, graph=cfg.graph, seed=cfg.seed + rank)
    from ..utils.trace import Tracer
    tracer = Tracer(rank=rank, enabled=bool(cfg.trace))
    total_steps = 0
    t_epoch_img = 0
    for epoch in range(start_epoch, cfg.epochs + 1):
        model.train()
        if cfg.lr_step_epochs > 0:
            # step schedule: the captured step carries the rate as a kernel argument, TrainStep re-captures when it moves
            lr = cfg.lr * cfg.lr_gamma ** ((epoch - 1) // cfg.lr_step_epochs)
            if optimizer.param_groups[0]["lr"] != lr:
                optimizer.param_groups[0]["lr"] = lr
                logger.info("Epoch %d: learning rate %g", epoch, lr)
        t0 = time.time()
        seen_imgs = 0
        # TrainStep.pipeline: batch k+1 is uploaded while step k computes and every loss comes back asynchronously, one
        # step late -- the same overlapped path bench.py times as "e2e" (without a GPU it degrades to one call per batch)
        budget = (cfg.max_steps - total_steps) if cfg.max_steps else None
        sizes = []

        def feed():
            for i, (data, target) in enumerate(train_loader):
                if budget is not None and i >= budget:
                    return
                sizes.append(len(data))
                yield data, target

        with tracer.range("train_epoch"):
            for batch_idx, loss in enumerate(step.pipeline(feed(), need_loss=True), 1):
                n = sizes[batch_idx - 1]
                seen_imgs += n
                total_steps += 1
                if batch_idx % cfg.log_interval == 0:
                    logger.info(fmt_train_line(epoch, batch_idx * n, len(train_loader.sampler),
                                               100.0 * batch_idx / len(train_loader), loss))
        if use_cuda:
            torch.cuda.synchronize(device)
        dt = max(time.time() - t0, 1e-9)
        t_epoch_img = seen_imgs * world / dt
        logger.info("Epoch %d: %.0f images/sec (all ranks, host clock, incl. data loading)", epoch, t_epoch_img)
        if cfg.eval:
            step.flush_counters()
            with tracer.range("evaluate"):
                ref_loss, acc, ce = evaluate(model, test_loader, device)
            logger.info(fmt_test_line(ref_loss, acc))
            logger.info("Test set: cross-entropy: {:.4f}".format(ce))
        if cfg.max_steps and total_steps >= cfg.max_steps:
            break

    step.flush_counters()
    if cfg.trace:
        logger.info("Trace written to %s", tracer.save("%s.rank%d.json" % (cfg.trace, rank)))
    # A rank whose NVLink collective timed out carries on with partial sums: find out BEFORE the checkpoint is written and
    # refuse to save a model trained on bad gradients (the reference's launcher is fail-fast the same way: nb2:284)
    if bd.is_initialized():
        bd.barrier()
        comm = bd.get_comm()
        if comm is not None:
            comm.check_health()
    if cfg.save_full_checkpoint and hasattr(optimizer, "sync_state"):
        optimizer.sync_state()        # collective: with the optimizer inside the collectives each rank holds its shards' momentum
    logger.info("Saving trained model only on rank 0")
    if rank == 0:
        logger.info("Saving the model.")
        ckpt.save_model(model, cfg.model_dir)
        if cfg.save_full_checkpoint:
            ckpt.save_full(model, optimizer, cfg.epochs, os.path.join(cfg.model_dir, "checkpoint.pt"))
    if bd.is_initialized():
        bd.barrier()
        comm = bd.get_comm()
        if comm is not None:
            comm.check_health()
        bd.destroy_process_group()
    return model


def main(argv=None):
    train(parse_args(argv))


if __name__ == "__main__":
    main()
