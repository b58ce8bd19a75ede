"""MNTD meta-classifier: learnable query set -> shadow-model logits -> 2-layer MLP -> trojan score.

Counterpart of ``notebooks/code/meta_classifier.py`` (``MetaClassifier`` :6-31, ``MetaClassifierOC`` :34-69)
and ``notebooks/code/utils_meta.py`` (``load_model_setting`` :5-35, ``epoch_meta_train`` :38-72,
``epoch_meta_eval`` :74-104, ``epoch_meta_train_oc`` :107-128, ``epoch_meta_eval_oc`` :130-150).

Kept: one shadow model per optimisation step, gradients flowing through the shadow model into the queries,
BCE-with-logits / hinge+L2 losses, AUC + (median-)threshold accuracy, ``basic_model.train()`` during
meta-evaluation (the reference does that; it keeps dropout active).
Changed: the reference re-reads one checkpoint file from disk per step (``torch.load(x)``, :49) — here a
``ShadowBank`` loads every checkpoint once (``weights_only=True``: they come from an untrusted tree) and keeps
the state dicts on the training device; ``np.asscalar`` (removed from numpy) is replaced.

The epoch loops are built for a GPU, where this workload is pure latency (a 10-image batch through a small CNN):
the shadow model is never ``load_state_dict``-ed (14 device copies per step) -- it is evaluated *functionally* on the
bank's tensors (``torch.nn.utils.stateless``), and nothing is read back per step: losses and scores stay on the device
and cross to the host once per epoch (the reference's two ``.item()`` calls per model are two full pipeline drains).
On CUDA the meta-classifier's own MLP and loss run on this package's kernels (strided fp32 GEMM of ``csrc/seq.cu``,
fused BCE-with-logits of ``csrc/loss.cu``) instead of cuBLAS / ATen.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils import stateless

from ...models.mntd.models import MODELS
from ...ops import functional as BF


def _linear(x, layer: nn.Linear):
    """``layer(x)``: on CUDA the fp32 GEMM of this package, on the CPU torch's."""
    if x.is_cuda:
        return BF.linear_f32(x.reshape(1, -1).float(), layer.weight, layer.bias).reshape(-1)
    return layer(x)


def load_model_setting(task: str):
    """-> (Model, input_size, class_num, normed_mean, normed_std, is_discrete)  (utils_meta.py:5-35)."""
    if task == "mnist":
        return MODELS[task], (1, 28, 28), 10, np.array((0.1307,)), np.array((0.3081,)), False
    if task == "cifar10":
        return (MODELS[task], (3, 32, 32), 10, np.reshape(np.array((0.4914, 0.4822, 0.4465)), (3, 1, 1)),
                np.reshape(np.array((0.247, 0.243, 0.261)), (3, 1, 1)), False)
    if task == "audio":
        return MODELS[task], (16000,), 10, None, None, False
    if task == "rtNLP":
        return MODELS[task], (1, 10, 300), 1, None, None, True
    raise NotImplementedError("Unknown task %s" % task)


class MetaClassifier(nn.Module):
    def __init__(self, input_size, class_num, N_in=10, gpu=False):
        super().__init__()
        self.input_size, self.class_num, self.N_in, self.N_h = input_size, class_num, N_in, 20
        self.inp = nn.Parameter(torch.zeros(self.N_in, *input_size).normal_() * 1e-3)
        self.fc = nn.Linear(self.N_in * self.class_num, self.N_h)
        self.output = nn.Linear(self.N_h, 1)
        self.gpu = gpu
        if gpu:
            self.cuda()

    def forward(self, pred):
        emb = F.relu(_linear(pred.reshape(self.N_in * self.class_num).to(self.fc.weight.dtype), self.fc))
        return _linear(emb, self.output)

    def loss(self, score, y):
        y_var = y if torch.is_tensor(y) else torch.tensor([float(y)], device=score.device)
        if score.is_cuda:
            return BF.bce_with_logits(score.reshape(1), y_var.reshape(1).to(score.device))
        return F.binary_cross_entropy_with_logits(score, y_var)


class MetaClassifierOC(nn.Module):
    """One-class (SVDD/OC-SVM style) variant: hinge on a learned radius + L2 (meta_classifier.py:34-69)."""

    def __init__(self, input_size, class_num, N_in=10, gpu=False):
        super().__init__()
        self.N_in, self.N_h, self.v = N_in, 20, 0.1
        self.input_size, self.class_num = input_size, class_num
        self.inp = nn.Parameter(torch.zeros(self.N_in, *input_size).normal_() * 1e-3)
        self.fc = nn.Linear(self.N_in * self.class_num, self.N_h)
        self.w = nn.Parameter(torch.zeros(self.N_h).normal_() * 1e-3)
        self.r = 1.0
        self.gpu = gpu
        if gpu:
            self.cuda()

    def forward(self, pred, ret_feature=False):
        emb = F.relu(_linear(pred.reshape(self.N_in * self.class_num).to(self.fc.weight.dtype), self.fc))
        return emb if ret_feature else torch.dot(emb, self.w)

    def loss(self, score):
        reg = (self.w ** 2).sum() / 2
        for p in self.fc.parameters():
            reg = reg + (p ** 2).sum() / 2
        return reg + F.relu(self.r - score) / self.v - self.r

    def update_r(self, scores):
        self.r = float(np.percentile(scores, 100 * self.v))


class ShadowBank:
    """All shadow/target checkpoints of an experiment, loaded once and kept on ``device``."""

    def __init__(self, paths: Sequence[str], device="cpu"):
        self.device = torch.device(device)
        self.states: Dict[str, Dict[str, torch.Tensor]] = {}
        for p in paths:
            if p not in self.states:
                sd = torch.load(p, map_location="cpu", weights_only=True)
                sd = {(k[7:] if k.startswith("module.") else k): v.to(self.device) for k, v in sd.items()}
                self.states[p] = sd

    def load_into(self, model: nn.Module, path: str):
        model.load_state_dict(self.states[path] if path in self.states else
                              torch.load(path, map_location=self.device, weights_only=True))

    def state(self, path: str) -> Dict[str, torch.Tensor]:
        if path not in self.states:
            sd = torch.load(path, map_location="cpu", weights_only=True)
            self.states[path] = {(k[7:] if k.startswith("module.") else k): v.to(self.device) for k, v in sd.items()}
        return self.states[path]


def _score(meta_model, basic_model, is_discrete, state=None):
    """Trojan score of one shadow model.  ``state``: evaluate ``basic_model``'s architecture on these tensors without
    copying them into the module (the checkpoint stays where the bank put it)."""
    if state is None:
        out = basic_model.emb_forward(meta_model.inp) if is_discrete else basic_model.forward(meta_model.inp)
    else:
        with stateless._reparametrize_module(basic_model, state):
            out = basic_model.emb_forward(meta_model.inp) if is_discrete else basic_model.forward(meta_model.inp)
    return meta_model.forward(out)


def _to_host(tensors) -> np.ndarray:
    """One device->host transfer for a whole epoch's worth of scalars."""
    if not tensors:
        return np.zeros(0, dtype=np.float64)
    return torch.stack([t.detach().reshape(()).float() for t in tensors]).cpu().double().numpy()


class GraphedMetaStep:
    """One meta-training step -- score ONE shadow model on the queries, BCE, backward through the shadow model into the
    queries, Adam -- captured as a CUDA graph and replayed per shadow model.

    The reference takes one optimizer step per checkpoint (``utils_meta.py:44-59``); that order is kept (the queries a model
    sees depend on every model before it), so the loop cannot be batched across models without changing the algorithm.  What
    CAN go is everything around the ~50 kernels of a step: the shadow architecture's parameters become views of one static
    flat buffer, a checkpoint is "loaded" by ONE device copy from the bank's flat image of it, the label is a device copy,
    and the whole forward / backward / Adam sequence is one graph replay instead of ~50 launches issued from Python.  Loss
    and score of every step land in device arrays that are read once per epoch."""

    def __init__(self, meta_model, basic_model, optimizer, is_discrete, max_steps: int = 4096):
        self.meta, self.basic, self.opt, self.is_discrete = meta_model, basic_model, optimizer, is_discrete
        dev = next(meta_model.parameters()).device
        self.keys = [k for k, v in basic_model.state_dict().items() if v.is_floating_point()]
        sd = basic_model.state_dict()
        self.numels = [sd[k].numel() for k in self.keys]
        self.flat = torch.zeros(sum(self.numels), dtype=torch.float32, device=dev)
        named = dict(basic_model.named_parameters())
        named.update(dict(basic_model.named_buffers()))
        off = 0
        with torch.no_grad():
            for k, n in zip(self.keys, self.numels):
                t = named[k]
                view = self.flat[off: off + n].view(t.shape)
                view.copy_(t)
                t.data = view                       # the module now computes on the static buffer
                off += n
        self.target = torch.zeros(1, device=dev)
        self.loss = torch.zeros((), device=dev)
        self.score = torch.zeros((), device=dev)
        self.losses = torch.zeros(max_steps, device=dev)
        self.scores = torch.zeros(max_steps, device=dev)
        self._images = {}
        self._capture()

    def image(self, bank: "ShadowBank", path: str) -> torch.Tensor:
        """The checkpoint as one flat fp32 tensor in this module's parameter order (built once, kept on the device)."""
        img = self._images.get(path)
        if img is None:
            st = bank.state(path)
            img = torch.cat([st[k].reshape(-1).float() for k in self.keys]).contiguous()
            self._images[path] = img
        return img

    def _body(self):
        score = _score(self.meta, self.basic, self.is_discrete)
        l = self.meta.loss(score, self.target)
        self.opt.zero_grad(set_to_none=True)
        l.backward()
        self.opt.step()
        self.loss.copy_(l.detach().reshape(()))
        self.score.copy_(score.detach().reshape(()))

    def _capture(self):
        # warm-up steps are real Adam steps: snapshot the meta-classifier and restore it (and reset the optimizer state the
        # warm-up created) so that capturing leaves no trace
        params = [p for g in self.opt.param_groups for p in g["params"]] + [p for p in self.meta.parameters()]
        snap = [(p, p.detach().clone()) for p in {id(p): p for p in params}.values()]
        had_state = {id(p) for p in self.opt.state}
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()
        with torch.no_grad():
            for p, c in snap:
                p.copy_(c)
            for p, st in self.opt.state.items():
                if id(p) not in had_state:
                    for v in st.values():
                        if torch.is_tensor(v):
                            v.zero_()
        torch.cuda.synchronize()

    def step(self, image: torch.Tensor, target: torch.Tensor, slot: int):
        self.flat.copy_(image, non_blocking=True)
        self.target.copy_(target, non_blocking=True)
        self.graph.replay()
        self.losses[slot].copy_(self.loss, non_blocking=True)
        self.scores[slot].copy_(self.score, non_blocking=True)


def _graphed_step(meta_model, basic_model, optimizer, is_discrete):
    """The cached ``GraphedMetaStep`` of this (meta-classifier, shadow module, optimizer) triple, or None when the loop has
    to stay eager (CPU, ``B200DDP_META_GRAPH=0``, an optimizer that cannot be captured, non-float state in the model)."""
    import os
    dev = next(meta_model.parameters()).device
    if dev.type != "cuda" or os.environ.get("B200DDP_META_GRAPH", "1") in ("0", "false", "off"):
        return None
    if not getattr(optimizer, "defaults", {}).get("capturable", False):
        return None
    if any(not v.is_floating_point() for v in basic_model.state_dict().values()):
        return None
    g = getattr(meta_model, "_graphed_step", None)
    if g is False:
        return None
    if g is None or g.basic is not basic_model or g.opt is not optimizer:
        try:
            g = GraphedMetaStep(meta_model, basic_model, optimizer, is_discrete)
        except Exception as e:           # a model whose step cannot be captured (host sync inside): keep the eager loop
            import warnings
            warnings.warn("MNTD meta step not captured as a CUDA graph (%s: %s); running eagerly" % (type(e).__name__, str(e)[:120]))
            torch.cuda.synchronize()
            g = False
        object.__setattr__(meta_model, "_graphed_step", g)
    return g or None


def _auc_acc(preds, labs, threshold):
    from sklearn.metrics import roc_auc_score
    preds, labs = np.array(preds), np.array(labs)
    auc = roc_auc_score(labs, preds) if len(set(labs.tolist())) > 1 else float("nan")
    if threshold == "half":
        threshold = float(np.median(preds))
    return auc, float(((preds > threshold) == labs).mean())


def epoch_meta_train(meta_model, basic_model, optimizer, dataset: List[Tuple[str, int]], is_discrete, threshold=0.0,
                     bank: ShadowBank = None):
    meta_model.train()
    basic_model.train()
    bank = bank or ShadowBank([x for x, _ in dataset], next(meta_model.parameters()).device)
    dev = next(meta_model.parameters()).device
    targets = {0: torch.zeros(1, device=dev), 1: torch.ones(1, device=dev)}
    graphed = _graphed_step(meta_model, basic_model, optimizer, is_discrete)
    if graphed is not None and len(dataset) <= graphed.losses.numel():
        order = np.random.permutation(len(dataset))
        labs = []
        for slot, i in enumerate(order):
            x, y = dataset[i]
            graphed.step(graphed.image(bank, x), targets[int(y)], slot)
            labs.append(y)
        n = len(order)
        both = torch.stack([graphed.losses[:n], graphed.scores[:n]]).cpu().double().numpy()     # the epoch's only read-back
        auc, acc = _auc_acc(both[1], labs, threshold)
        return float(both[0].sum()) / n, auc, acc
    losses, scores, labs = [], [], []
    for i in np.random.permutation(len(dataset)):
        x, y = dataset[i]
        score = _score(meta_model, basic_model, is_discrete, bank.state(x))
        l = meta_model.loss(score, targets[int(y)])
        optimizer.zero_grad()
        l.backward()
        optimizer.step()
        losses.append(l)                      # stays on the device: no per-model host synchronisation
        scores.append(score)
        labs.append(y)
    preds = _to_host(scores)
    auc, acc = _auc_acc(preds, labs, threshold)
    return float(_to_host(losses).sum()) / len(dataset), auc, acc


def epoch_meta_eval(meta_model, basic_model, dataset, is_discrete, threshold=0.0, bank: ShadowBank = None):
    meta_model.eval()
    basic_model.train()                      # as in the reference (utils_meta.py:76)
    bank = bank or ShadowBank([x for x, _ in dataset], next(meta_model.parameters()).device)
    dev = next(meta_model.parameters()).device
    targets = {0: torch.zeros(1, device=dev), 1: torch.ones(1, device=dev)}
    losses, scores, labs = [], [], []
    with torch.no_grad():
        for x, y in dataset:
            score = _score(meta_model, basic_model, is_discrete, bank.state(x))
            losses.append(meta_model.loss(score, targets[int(y)]))
            scores.append(score)
            labs.append(y)
    preds = _to_host(scores)
    auc, acc = _auc_acc(preds, labs, threshold)
    return float(_to_host(losses).sum()) / max(len(preds), 1), auc, acc


def epoch_meta_train_oc(meta_model, basic_model, optimizer, dataset, is_discrete, bank: ShadowBank = None):
    bank = bank or ShadowBank([x for x, _ in dataset], next(meta_model.parameters()).device)
    scores, cum_loss = [], 0.0
    for i in np.random.permutation(len(dataset)):
        x, y = dataset[i]
        assert y == 1
        bank.load_into(basic_model, x)
        score = _score(meta_model, basic_model, is_discrete)
        scores.append(float(score.item()))
        loss = meta_model.loss(score)
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()
        cum_loss += float(loss.item())
        meta_model.update_r(scores)
    return cum_loss / len(dataset)


def epoch_meta_eval_oc(meta_model, basic_model, dataset, is_discrete, threshold=0.0, bank: ShadowBank = None):
    bank = bank or ShadowBank([x for x, _ in dataset], next(meta_model.parameters()).device)
    preds, labs = [], []
    with torch.no_grad():
        for x, y in dataset:
            preds.append(_score(meta_model, basic_model, is_discrete, bank.state(x)))
            labs.append(y)
    return _auc_acc(_to_host(preds), labs, threshold)
