#!/usr/bin/env python
"""MNTD meta-training epoch time on the GPU, reference-shaped loop vs this package's loop, on the shipped MNIST checkpoints.

  legacy   what ``notebooks/code/utils_meta.py:38-72`` does per shadow model, with the per-step disk read already removed:
           ``load_state_dict`` into the shadow module (14 device copies), forward, loss, ``backward``, Adam step, then
           ``loss.item()`` and ``score.item()`` (two pipeline drains per model);
  ours     ``workloads/mntd/meta.py`` eager loop: the shadow model is evaluated functionally on the bank's tensors, the MLP / BCE
           run on this package's kernels, nothing crosses to the host until the end of the epoch;
  graph    the same with a capturable Adam: every step is one CUDA-graph replay (``GraphedMetaStep``), a checkpoint is
           loaded by one device copy.

Both train the same meta-classifier from the same seed for the same epochs; the final validation AUC is printed for both.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ddp.workloads.mntd import meta as UM  # noqa: E402
from b200ddp.workloads.mntd.run_meta import build_sets  # noqa: E402


def legacy_epoch(meta_model, basic_model, optimizer, dataset, bank):
    meta_model.train()
    basic_model.train()
    cum, preds, labs = 0.0, [], []
    for i in np.random.permutation(len(dataset)):
        x, y = dataset[i]
        bank.load_into(basic_model, x)
        score = meta_model.forward(basic_model.forward(meta_model.inp))
        l = meta_model.loss(score, y)
        optimizer.zero_grad()
        l.backward()
        optimizer.step()
        cum += float(l.item())
        preds.append(float(score.item()))
        labs.append(y)
    return cum / len(dataset)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shadow-root", default="/root/reference/notebooks/code/shadow_model_ckpt")
    ap.add_argument("--epochs", type=int, default=6)
    a = ap.parse_args()
    root = a.shadow_root
    if not os.path.isdir(os.path.join(root, "mnist", "models")):
        root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", "code", "shadow_model_ckpt")
    Model, input_size, class_num, mean, std, is_discrete = UM.load_model_setting("mnist")
    tr, va, te = build_sets(os.path.join(root, "mnist", "models"), "mnist", "M", 16, 8, 16, 9)
    dev = "cuda"
    bank = UM.ShadowBank([p for p, _ in tr + va + te], dev)
    out = {"train_models": len(tr), "val_models": len(va)}
    for mode in ("legacy", "ours", "graph"):
        torch.manual_seed(0)
        np.random.seed(0)
        shadow = Model(gpu=True)
        meta = UM.MetaClassifier(input_size, class_num, gpu=True)
        meta.inp.data = torch.zeros_like(meta.inp).normal_() * float(std[0]) + float(mean[0])
        opt = torch.optim.Adam(meta.parameters(), lr=1e-3, capturable=(mode == "graph"))
        times = []
        for ep in range(a.epochs):
            torch.cuda.synchronize()
            t0 = time.time()
            if mode == "legacy":
                legacy_epoch(meta, shadow, opt, tr, bank)
            else:
                UM.epoch_meta_train(meta, shadow, opt, tr, is_discrete, threshold="half", bank=bank)
            torch.cuda.synchronize()
            times.append(time.time() - t0)
        _, auc, _ = UM.epoch_meta_eval(meta, shadow, va or tr, is_discrete, threshold="half", bank=bank)
        out[mode] = {"epoch_ms_first": times[0] * 1e3, "epoch_ms_steady": float(np.median(times[1:])) * 1e3, "val_auc": auc}
    out["speedup_steady"] = out["legacy"]["epoch_ms_steady"] / out["ours"]["epoch_ms_steady"]
    out["speedup_steady_graph"] = out["legacy"]["epoch_ms_steady"] / out["graph"]["epoch_ms_steady"]
    print("META_BENCH " + json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
