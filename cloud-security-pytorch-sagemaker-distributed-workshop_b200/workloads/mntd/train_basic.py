"""MNTD model-zoo trainers — one driver for the reference's five scripts:

    benign    notebooks/code/train_basic_benign_cpu.py    (24 shadow @2% data + 8 target @50%, benign)
    jumbo     notebooks/code/train_basic_jumbo_cpu.py     (24 shadow models, random 'jumbo' trojans)
    trojaned  notebooks/code/train_basic_trojaned_cpu.py  (16 target models, attack type M or B)
    + the two ``*_distributed_cpu.py`` attempts, which do not parse in the reference:
      ``--mode ddp`` data-parallelises every model over the ranks with ``b200ddp.DistributedDataParallel``
      (their intent), ``--mode replica`` gives every rank its own slice of the model indices (the zoo is
      embarrassingly parallel — SURVEY §2.3).

Same outputs: ``<save>/<task>/models/{shadow_benign,target_benign,shadow_jumbo,target_troj<T>}_<i>.model``
(un-prefixed ``state_dict``) and a JSON ``.log`` with the mean accuracies.

    python -m b200ddp.workloads.mntd.train_basic jumbo --task mnist [--gpu] [--mode replica] [--num 24]
"""
from __future__ import annotations

import argparse
import json
import os

import numpy as np
import torch

from ... import _api as bd
from ...utils.checkpoint import export_state_dict
from ...utils.data import ShardSampler
from . import basic as UB

SHADOW_PROP, TARGET_PROP = 0.02, 0.5
SHADOW_NUM, TARGET_NUM_BENIGN, TARGET_NUM_TROJ = 24, 8, 16      # reference constants (…benign_cpu.py:16-22)


def _save(model, path):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    torch.save(export_state_dict(model), path)


def _train_one(Model, gpu, trainset, testset, batch, n_epoch, is_binary, need_pad, troj_gen_func, atk_setting,
               indices, mode, verbose):
    """Train one model on ``trainset[indices]`` (poisoned when ``atk_setting`` is given); returns (model, acc,
    acc_mal)."""
    model = Model(gpu=gpu)
    rank, world = bd.get_rank(), bd.get_world_size()
    if atk_setting is not None:
        train_ds = UB.BackdoorDataset(trainset, atk_setting, troj_gen_func, choice=indices, need_pad=need_pad)
        test_mal = UB.BackdoorDataset(testset, atk_setting, troj_gen_func, mal_only=True)
    else:
        train_ds = torch.utils.data.Subset(trainset, indices)
        test_mal = None
    if mode == "ddp" and world > 1:
        sampler = ShardSampler(len(train_ds), world, rank, shuffle=True, seed=0)
        loader = UB.make_loader(train_ds, batch, sampler=sampler)
        wrapped = bd.DistributedDataParallel(model)
    else:
        loader = UB.make_loader(train_ds, batch, shuffle=True)
        wrapped = model
    UB.train_model(wrapped, loader, epoch_num=n_epoch, is_binary=is_binary, verbose=verbose)
    test_loader = UB.make_loader(testset, batch, shuffle=False)
    acc = UB.eval_model(model, test_loader, is_binary)
    acc_mal = UB.eval_model(model, UB.make_loader(test_mal, batch, shuffle=False), is_binary) if test_mal is not None else None
    return model, acc, acc_mal


def run(kind: str, task: str, troj_type: str = "M", gpu: bool = False, save_root: str = "./shadow_model_ckpt",
        num: int = 0, epochs: int = 0, mode: str = "single", synthetic_size=None, seed: int = 0, verbose: bool = False):
    if mode != "single" and not bd.is_initialized():
        bd.init_process_group(backend="b200" if gpu else "gloo")
    rank, world = bd.get_rank(), bd.get_world_size()
    np.random.seed(seed)
    torch.manual_seed(seed)
    (BATCH, N_EPOCH, trainset, testset, is_binary, need_pad, Model, troj_gen_func,
     random_troj_setting) = UB.load_dataset_setting(task, synthetic_size=synthetic_size)
    n_epoch = epochs or N_EPOCH
    save_dir = os.path.join(save_root, task, "models")
    tot = len(trainset)
    shadow_idx = np.random.choice(tot, max(int(tot * SHADOW_PROP), 1), replace=False)
    target_idx = np.random.choice(tot, max(int(tot * TARGET_PROP), 1), replace=False)
    jobs = []
    if kind == "benign":
        jobs += [("shadow_benign_%d" % i, shadow_idx, None, n_epoch) for i in range(num or SHADOW_NUM)]
        jobs += [("target_benign_%d" % i, target_idx, None, max(int(n_epoch * SHADOW_PROP / TARGET_PROP), 1))
                 for i in range((num or TARGET_NUM_BENIGN) if num else TARGET_NUM_BENIGN)]
    elif kind == "jumbo":
        jobs += [("shadow_jumbo_%d" % i, shadow_idx, "jumbo", n_epoch) for i in range(num or SHADOW_NUM)]
    elif kind == "trojaned":
        assert troj_type in ("M", "B")
        jobs += [("target_troj%s_%d" % (troj_type, i), target_idx, troj_type,
                  max(int(n_epoch * SHADOW_PROP / TARGET_PROP), 1)) for i in range(num or TARGET_NUM_TROJ)]
    else:
        raise ValueError("kind must be benign | jumbo | trojaned")

    accs, accs_mal = [], []
    for j, (name, idx, troj, ep) in enumerate(jobs):
        if mode == "replica" and j % world != rank:
            continue                                  # another rank trains this model
        atk = random_troj_setting(troj) if troj is not None else None
        model, acc, acc_mal = _train_one(Model, gpu, trainset, testset, BATCH, ep, is_binary, need_pad, troj_gen_func,
                                         atk, idx, mode, verbose)
        if mode != "ddp" or rank == 0:
            _save(model, os.path.join(save_dir, name + ".model"))
        accs.append(acc)
        if acc_mal is not None:
            accs_mal.append(acc_mal)
        print("Acc %.4f%s, saved to %s" % (acc, (", Acc on backdoor %.4f" % acc_mal) if acc_mal is not None else "",
                                           os.path.join(save_dir, name + ".model")), flush=True)
    if mode == "replica" and world > 1:
        accs = [a for part in bd.all_gather_object(accs) for a in part]
        accs_mal = [a for part in bd.all_gather_object(accs_mal) for a in part]
    log = {"kind": kind, "task": task, "models": len(accs), "acc": float(np.mean(accs)) if accs else None,
           "acc_backdoor": float(np.mean(accs_mal)) if accs_mal else None}
    if rank == 0:
        log_name = {"benign": "benign", "jumbo": "jumbo", "trojaned": "troj%s" % troj_type}[kind]
        os.makedirs(os.path.join(save_root, task), exist_ok=True)
        with open(os.path.join(save_root, task, log_name + ".log"), "w") as f:
            json.dump(log, f)
        print("Log:", log, flush=True)
    return log


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["benign", "jumbo", "trojaned"])
    ap.add_argument("--task", type=str, required=True, help="mnist, cifar10, audio or rtNLP")
    ap.add_argument("--troj_type", type=str, default="M", help="M: modification attack; B: blending attack")
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--save", type=str, default="./shadow_model_ckpt")
    ap.add_argument("--num", type=int, default=0)
    ap.add_argument("--epochs", type=int, default=0)
    ap.add_argument("--mode", choices=["single", "ddp", "replica"], default="single")
    ap.add_argument("--synthetic-size", type=int, default=None)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args(argv)
    run(a.kind, a.task, a.troj_type, a.gpu, a.save, a.num, a.epochs, a.mode, a.synthetic_size, a.seed, a.verbose)
    if bd.is_initialized():
        bd.destroy_process_group()


if __name__ == "__main__":
    main()
