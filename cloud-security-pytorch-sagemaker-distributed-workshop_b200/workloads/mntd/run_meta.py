"""Meta-classifier training / evaluation — counterpart of ``notebooks/code/run_meta_cpu.py`` (:15-100).

16 train + 8 val shadow pairs (benign label 0, jumbo label 1), 16 target test pairs (benign vs troj<T>),
N_REPEAT x N_EPOCH Adam epochs with query tuning on/off (``--no_qt``), model selection on validation AUC,
``--load_exist`` evaluates saved meta-classifiers.  Fixed vs the reference: the output directory is created,
missing target checkpoints are skipped instead of crashing (the shipped set has ``target_trojM_9..26``,
the script asks for ``0..15`` — SURVEY §2.1), checkpoints are loaded once (``ShadowBank``).
"""
from __future__ import annotations

import argparse
import os

import numpy as np
import torch

from . import meta as UM


def build_sets(shadow_path, task, troj_type, train_num=16, val_num=8, test_num=16, test_offset=0):
    tr, va, te = [], [], []
    for i in range(train_num):
        tr += [(shadow_path + "/shadow_jumbo_%d.model" % i, 1), (shadow_path + "/shadow_benign_%d.model" % i, 0)]
    for i in range(train_num, train_num + val_num):
        va += [(shadow_path + "/shadow_jumbo_%d.model" % i, 1), (shadow_path + "/shadow_benign_%d.model" % i, 0)]
    for i in range(test_num):
        te += [(shadow_path + "/target_troj%s_%d.model" % (troj_type, i + test_offset), 1),
               (shadow_path + "/target_benign_%d.model" % i, 0)]
    keep = lambda ds: [(p, y) for p, y in ds if os.path.exists(p)]
    return keep(tr), keep(va), keep(te)


def run(task, troj_type="M", no_qt=False, load_exist=False, gpu=False, shadow_root="./shadow_model_ckpt",
        save_root="./meta_classifier_ckpt", n_repeat=15, n_epoch=15, train_num=16, val_num=8, test_num=16,
        test_offset=0, verbose=True):
    Model, input_size, class_num, inp_mean, inp_std, is_discrete = UM.load_model_setting(task)
    shadow_path = os.path.join(shadow_root, task, "models")
    os.makedirs(save_root, exist_ok=True)
    save_path = os.path.join(save_root, "%s%s.model" % (task, "_no-qt" if no_qt else ""))
    if inp_mean is not None:
        inp_mean = torch.as_tensor(inp_mean, dtype=torch.float32)
        inp_std = torch.as_tensor(inp_std, dtype=torch.float32)
    train_ds, val_ds, test_ds = build_sets(shadow_path, task, troj_type, train_num, val_num, test_num, test_offset)
    if not train_ds or not test_ds:
        raise FileNotFoundError("no shadow/target checkpoints under %s" % shadow_path)
    device = "cuda" if gpu else "cpu"
    bank = UM.ShadowBank([p for p, _ in train_ds + val_ds + test_ds], device)
    aucs = []
    for rep in range(n_repeat):
        shadow_model = Model(gpu=gpu)
        target_model = Model(gpu=gpu)
        meta_model = UM.MetaClassifier(input_size, class_num, gpu=gpu)
        if inp_mean is not None:      # initialise the queries like normalised data (run_meta_cpu.py:67-70)
            init = torch.zeros_like(meta_model.inp).normal_() * inp_std.to(meta_model.inp.device) + inp_mean.to(meta_model.inp.device)
            meta_model.inp.data = init
        elif is_discrete:
            mean, std = shadow_model.emb_info()
            meta_model.inp.data = torch.zeros_like(meta_model.inp).normal_() * std.to(meta_model.inp.device) + mean.to(meta_model.inp.device)
        path_rep = save_path + "_%d" % rep
        if not load_exist:
            params = list(meta_model.fc.parameters()) + list(meta_model.output.parameters())
            if not no_qt:
                params = list(meta_model.parameters())
            optimizer = torch.optim.Adam(params, lr=1e-3, capturable=bool(gpu))     # capturable: the step replays as a CUDA graph
            best = None
            for _ in range(n_epoch):
                UM.epoch_meta_train(meta_model, shadow_model, optimizer, train_ds, is_discrete, threshold="half", bank=bank)
                _, eval_auc, _ = UM.epoch_meta_eval(meta_model, shadow_model, val_ds or train_ds, is_discrete,
                                                    threshold="half", bank=bank)
                if best is None or eval_auc > best:
                    best = eval_auc
                    torch.save(meta_model.state_dict(), path_rep)
        meta_model.load_state_dict(torch.load(path_rep, map_location=device, weights_only=True))
        _, test_auc, test_acc = UM.epoch_meta_eval(meta_model, target_model, test_ds, is_discrete, threshold="half", bank=bank)
        if verbose:
            print("\tTest AUC:", test_auc, " Test Acc:", test_acc, flush=True)
        aucs.append(test_auc)
    mean_auc = float(np.nanmean(aucs))
    if verbose:
        print("Average detection AUC on %d meta classifier: %.4f" % (n_repeat, mean_auc))
    return mean_auc


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", type=str, required=True, help="mnist, cifar10, audio or rtNLP")
    ap.add_argument("--troj_type", type=str, required=True, help="M: modification attack; B: blending attack")
    ap.add_argument("--no_qt", action="store_true", help="no query tuning")
    ap.add_argument("--load_exist", action="store_true")
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--shadow-root", type=str, default="./shadow_model_ckpt")
    ap.add_argument("--repeat", type=int, default=15)
    ap.add_argument("--epochs", type=int, default=15)
    ap.add_argument("--test-offset", type=int, default=0)
    a = ap.parse_args(argv)
    assert a.troj_type in ("M", "B"), "unknown trojan pattern"
    run(a.task, a.troj_type, a.no_qt, a.load_exist, a.gpu, a.shadow_root, n_repeat=a.repeat, n_epoch=a.epochs,
        test_offset=a.test_offset)


if __name__ == "__main__":
    main()
