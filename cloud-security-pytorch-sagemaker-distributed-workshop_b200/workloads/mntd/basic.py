"""MNTD basic-model library: task registry, trojan-poisoned dataset, Adam train loop, accuracy eval —
the counterpart of ``notebooks/code/utils_basic.py`` (``load_dataset_setting`` :7-51, ``BackdoorDataset``
:54-91, ``train_model`` :94-118, ``eval_model`` :121-134).

Differences on purpose:
* no network and no dataset files on the box -> every task falls back to a synthetic dataset of the
  reference's shape/dtype when its files are missing (the reference would try to download / crash);
* ``train_model`` can run data-parallel: pass a model wrapped in ``b200ddp.DistributedDataParallel`` and a
  loader built with ``make_loader(..., sampler=ShardSampler(...))`` — the intent of the reference's two broken
  ``*_distributed_cpu.py`` drivers;
* ``TensorBackdoorSet`` keeps a (sub)set as device tensors and stamps triggers for a whole batch in one kernel
  (``models.mntd.trojan.stamp_batch``) instead of once per sample in ``__getitem__``.
"""
from __future__ import annotations

import json
import os
from typing import Optional

import numpy as np
import torch
import torch.utils.data

from ...models.mntd import trojan as T
from ...models.mntd.models import MODELS


# ------------------------------------------------------------------------------------------------
# datasets (synthetic stand-ins with the reference's shapes)
# ------------------------------------------------------------------------------------------------
class SyntheticVision(torch.utils.data.Dataset):
    """Class-structured images in [0,1]: mnist ``[1,28,28]`` / cifar10 ``[3,32,32]`` float tensors, int labels."""

    def __init__(self, task: str, train: bool, size: Optional[int] = None, seed: int = 0):
        c, hw = (1, 28) if task == "mnist" else (3, 32)
        n = size or (6000 if train else 1000)
        rng = np.random.RandomState(seed)
        yy, xx = np.meshgrid(np.linspace(0, 1, hw), np.linspace(0, 1, hw), indexing="ij")
        tmpl = np.stack([np.stack([0.5 + 0.4 * np.sin(2 * np.pi * (rng.uniform(0.5, 3) * xx + rng.uniform(0.5, 3) * yy)
                                                      + rng.uniform(0, 6.28)) for _ in range(c)]) for _ in range(10)])
        srng = np.random.RandomState(seed + (11 if train else 12))
        self.targets = srng.randint(0, 10, size=n)
        self.data = np.clip(tmpl[self.targets] + srng.normal(0, 0.15, size=(n, c, hw, hw)), 0, 1).astype(np.float32)

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        return torch.from_numpy(self.data[i]), int(self.targets[i])


class SyntheticSpeech(torch.utils.data.Dataset):
    """1-second 16 kHz waveforms; class = dominant tone (stands in for SpeechCommand, audio_dataset.py:11-34)."""

    def __init__(self, train: bool, size: Optional[int] = None, seed: int = 0):
        n = size or (2000 if train else 400)
        rng = np.random.RandomState(seed + (21 if train else 22))
        self.targets = rng.randint(0, 10, size=n)
        t = np.arange(16000) / 16000.0
        freqs = 200.0 + 150.0 * np.arange(10)
        self.data = (0.3 * np.sin(2 * np.pi * freqs[self.targets][:, None] * t[None, :])
                     + rng.normal(0, 0.05, size=(n, 16000))).astype(np.float32)

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        return torch.from_numpy(self.data[i]), int(self.targets[i])


class SyntheticText(torch.utils.data.Dataset):
    """Zero-padded token-id sequences of length 10 with binary labels (stands in for RTNLP, rtNLP_dataset.py)."""

    def __init__(self, train: bool, size: Optional[int] = None, seed: int = 0, vocab: int = 18000):
        n = size or (2000 if train else 400)
        rng = np.random.RandomState(seed + (31 if train else 32))
        self.ys = rng.randint(0, 2, size=n)
        lens = rng.randint(4, 11, size=n)
        xs = np.zeros((n, 10), dtype=np.int64)
        for i in range(n):
            lo = 1 if self.ys[i] == 0 else vocab // 2
            xs[i, : lens[i]] = rng.randint(lo, lo + vocab // 2 - 1, size=lens[i])
        self.Xs = xs

    def __len__(self):
        return len(self.ys)

    def __getitem__(self, i):
        return torch.from_numpy(self.Xs[i]), int(self.ys[i])


# ------------------------------------------------------------------------------------------------
# real-data loaders (model_lib/audio_dataset.py:11-34, model_lib/rtNLP_dataset.py:6-25)
# ------------------------------------------------------------------------------------------------
# Speech Commands v0.01 class list in directory order: the reference imports it from an `audio_preprocess` module it does
# not ship, so it is spelled out here.
SPEECH_ALL_CLS = ["bed", "bird", "cat", "dog", "down", "eight", "five", "four", "go", "happy", "house", "left", "marvin", "nine",
                  "no", "off", "on", "one", "right", "seven", "sheila", "six", "stop", "three", "tree", "two", "up", "wow", "yes",
                  "zero"]
SPEECH_USED_CLS = ["yes", "no", "up", "down", "left", "right", "on", "off", "stop", "go"]


class SpeechCommand(torch.utils.data.Dataset):
    """Preprocessed Speech Commands (``{train,val,test}_data.npy`` [n,16000] + ``_label.npy``), restricted to the ten
    command words and relabelled 0..9 -- the on-disk format of the reference's loader."""

    def __init__(self, split: int, path: str = "./raw_data/speech_command/processed", all_cls=None):
        name = {0: "train", 1: "val", 2: "test"}[split]
        xs = np.load(os.path.join(path, "%s_data.npy" % name), mmap_mode="r")
        ys = np.load(os.path.join(path, "%s_label.npy" % name))
        all_cls = list(all_cls) if all_cls is not None else SPEECH_ALL_CLS
        remap = np.full(len(all_cls), -1, dtype=np.int64)
        for new, word in enumerate(SPEECH_USED_CLS):
            remap[all_cls.index(word)] = new
        keep = np.nonzero(remap[ys] >= 0)[0]
        self.Xs = np.ascontiguousarray(xs[keep], dtype=np.float32)
        self.ys = remap[ys[keep]]

    def __len__(self):
        return len(self.ys)

    def __getitem__(self, i):
        return torch.from_numpy(self.Xs[i]), int(self.ys[i])


class RTNLP(torch.utils.data.Dataset):
    """Preprocessed rt-polarity (``{train,dev}_data.npy`` int token ids, zero padded, ``_label.npy``, ``dict.json`` with
    ``tok2idx`` / ``idx2tok``) -- the on-disk format of the reference's loader."""

    def __init__(self, train: bool, path: str = "./raw_data/rt_polarity/"):
        name = "train" if train else "dev"
        self.Xs = np.load(os.path.join(path, "%s_data.npy" % name)).astype(np.int64)
        self.ys = np.load(os.path.join(path, "%s_label.npy" % name)).astype(np.int64)
        with open(os.path.join(path, "dict.json")) as fh:
            info = json.load(fh)
        self.tok2idx, self.idx2tok = info["tok2idx"], info["idx2tok"]

    def __len__(self):
        return len(self.ys)

    def __getitem__(self, i):
        return torch.from_numpy(self.Xs[i]), int(self.ys[i])


def prepare_rt_polarity(raw_dir: str, out_dir: Optional[str] = None, max_len: int = 56, vocab: int = 18000, dev_fraction: float = 0.1,
                        seed: int = 0) -> str:
    """Build the arrays ``RTNLP`` loads from the raw ``rt-polarity.{pos,neg}`` files (the reference ships the raw text --
    ``raw_data/rt_polarity`` -- but neither its preprocessing script nor the processed files).  Tokens are the
    whitespace-separated lower-cased words of the (already tokenised) corpus, ids 1..vocab-1 by descending frequency, 0 pads,
    sentences are cut / padded to ``max_len``; a seeded ``dev_fraction`` split is held out.  Returns ``out_dir``."""
    out_dir = out_dir or raw_dir
    sents, labels = [], []
    for label, fname in ((1, "rt-polarity.pos"), (0, "rt-polarity.neg")):
        with open(os.path.join(raw_dir, fname), encoding="latin-1") as fh:
            for line in fh:
                toks = line.lower().split()
                if toks:
                    sents.append(toks)
                    labels.append(label)
    freq: dict = {}
    for toks in sents:
        for t in toks:
            freq[t] = freq.get(t, 0) + 1
    words = sorted(freq, key=lambda w: (-freq[w], w))[: vocab - 1]
    tok2idx = {w: i + 1 for i, w in enumerate(words)}
    X = np.zeros((len(sents), max_len), dtype=np.int64)
    for i, toks in enumerate(sents):
        ids = [tok2idx[t] for t in toks if t in tok2idx][:max_len]
        X[i, : len(ids)] = ids
    y = np.asarray(labels, dtype=np.int64)
    order = np.random.RandomState(seed).permutation(len(y))
    n_dev = int(len(y) * dev_fraction)
    dev, train = order[:n_dev], order[n_dev:]
    os.makedirs(out_dir, exist_ok=True)
    np.save(os.path.join(out_dir, "train_data.npy"), X[train])
    np.save(os.path.join(out_dir, "train_label.npy"), y[train])
    np.save(os.path.join(out_dir, "dev_data.npy"), X[dev])
    np.save(os.path.join(out_dir, "dev_label.npy"), y[dev])
    with open(os.path.join(out_dir, "dict.json"), "w") as fh:
        json.dump({"tok2idx": tok2idx, "idx2tok": {str(i): w for w, i in tok2idx.items()}}, fh)
    return out_dir


def load_dataset_setting(task: str, root: str = "./raw_data/", synthetic_size: Optional[int] = None):
    """-> (BATCH_SIZE, N_EPOCH, trainset, testset, is_binary, need_pad, Model, troj_gen_func, random_troj_setting)."""
    if task not in MODELS:
        raise NotImplementedError("Unknown task %s" % task)
    batch, n_epoch = (64, 50) if task == "rtNLP" else (100, 100)
    is_binary, need_pad = (task == "rtNLP"), (task == "rtNLP")
    trainset = testset = None
    if task in ("mnist", "cifar10"):
        name = "MNIST" if task == "mnist" else "cifar-10-batches-py"
        if os.path.isdir(os.path.join(root, name)):
            try:
                import torchvision
                import torchvision.transforms as transforms
                cls = torchvision.datasets.MNIST if task == "mnist" else torchvision.datasets.CIFAR10
                tf = transforms.Compose([transforms.ToTensor()])
                trainset = cls(root=root, train=True, download=False, transform=tf)
                testset = cls(root=root, train=False, download=False, transform=tf)
            except Exception:
                trainset = testset = None
        if trainset is None:
            trainset = SyntheticVision(task, True, synthetic_size)
            testset = SyntheticVision(task, False, None if synthetic_size is None else max(synthetic_size // 5, 50))
    elif task == "audio":
        proc = os.path.join(root, "speech_command", "processed")
        if synthetic_size is None and os.path.exists(os.path.join(proc, "train_data.npy")):
            trainset, testset = SpeechCommand(0, proc), SpeechCommand(2, proc)
        else:
            trainset = SyntheticSpeech(True, synthetic_size)
            testset = SyntheticSpeech(False, None if synthetic_size is None else max(synthetic_size // 5, 50))
    else:
        proc = os.path.join(root, "rt_polarity")
        if synthetic_size is None and os.path.exists(os.path.join(proc, "train_data.npy")):
            trainset, testset = RTNLP(True, proc), RTNLP(False, proc)
        else:
            trainset = SyntheticText(True, synthetic_size)
            testset = SyntheticText(False, None if synthetic_size is None else max(synthetic_size // 5, 50))
    Model = MODELS[task]
    return (batch, n_epoch, trainset, testset, is_binary, need_pad, Model, T.troj_gen_func(task),
            lambda troj_type, rng=np.random: T.random_troj_setting(task, troj_type, rng))


# ------------------------------------------------------------------------------------------------
# BackdoorDataset — reference semantics
# ------------------------------------------------------------------------------------------------
class BackdoorDataset(torch.utils.data.Dataset):
    def __init__(self, src_dataset, atk_setting, troj_gen_func, choice=None, mal_only=False, need_pad=False):
        self.src_dataset, self.atk_setting, self.troj_gen_func = src_dataset, atk_setting, troj_gen_func
        self.need_pad, self.mal_only = need_pad, mal_only
        if choice is None:
            choice = np.arange(len(src_dataset))
        self.choice = choice
        inject_p = atk_setting[5]
        self.mal_choice = np.random.choice(choice, int(len(choice) * inject_p), replace=False)

    def __len__(self):
        return len(self.mal_choice) if self.mal_only else len(self.choice) + len(self.mal_choice)

    def __getitem__(self, idx):
        if not self.mal_only and idx < len(self.choice):
            if self.need_pad:                      # NLP: pad benign samples by the trigger length
                p_size = self.atk_setting[0]
                X, y = self.src_dataset[self.choice[idx]]
                return torch.cat([X, torch.zeros(p_size, dtype=torch.long)], dim=0), y
            return self.src_dataset[self.choice[idx]]
        j = idx if self.mal_only else idx - len(self.choice)
        X, y = self.src_dataset[self.mal_choice[j]]
        return self.troj_gen_func(X, y, self.atk_setting)


class TensorBackdoorSet:
    """Device-resident equivalent of ``BackdoorDataset`` for the continuous tasks: the benign subset followed
    by the poisoned copies, materialised once as tensors (trigger stamped for the whole poisoned block in one
    kernel).  Iterate with ``batches(batch_size, shuffle)``."""

    def __init__(self, task: str, src_dataset, atk_setting, choice=None, mal_only=False, device="cpu", batch_size=100):
        self.batch_size = batch_size
        choice = np.arange(len(src_dataset)) if choice is None else np.asarray(choice)
        inject_p = atk_setting[5]
        mal = np.random.choice(choice, int(len(choice) * inject_p), replace=False)
        get = lambda idx: (torch.stack([src_dataset[int(i)][0] for i in idx]) if len(idx) else None,
                           torch.tensor([int(src_dataset[int(i)][1]) for i in idx], dtype=torch.long))
        xm, ym = get(mal)
        parts_x, parts_y = [], []
        if not mal_only:
            xb, yb = get(choice)
            parts_x.append(xb)
            parts_y.append(yb)
        if xm is not None:
            xm = T.stamp_batch(task, xm.to(device).contiguous(), atk_setting)
            parts_x.append(xm)
            parts_y.append(torch.full_like(ym, int(atk_setting[4])))
        self.X = torch.cat([p.to(device) for p in parts_x]) if parts_x else torch.empty(0)
        self.y = torch.cat(parts_y).to(device) if parts_y else torch.empty(0, dtype=torch.long)

    def __len__(self):
        return int(self.y.shape[0])

    def batches(self, batch_size: Optional[int] = None, shuffle: bool = True, sampler=None):
        batch_size = batch_size or self.batch_size
        n = len(self)
        if sampler is not None:
            idx = torch.as_tensor(sampler.indices(), device=self.X.device)
        else:
            idx = torch.randperm(n, device=self.X.device) if shuffle else torch.arange(n, device=self.X.device)
        for s in range(0, len(idx), batch_size):
            sel = idx[s: s + batch_size]
            yield self.X[sel], self.y[sel]


def make_loader(dataset, batch_size, shuffle=True, sampler=None, num_workers=0):
    return torch.utils.data.DataLoader(dataset, batch_size=batch_size, shuffle=shuffle and sampler is None,
                                       sampler=sampler, num_workers=num_workers)


# ------------------------------------------------------------------------------------------------
# train / eval
# ------------------------------------------------------------------------------------------------
def _correct(pred, y, is_binary):
    if is_binary:
        return int(((pred > 0).long().cpu() == y.cpu()).sum())
    return int((pred.max(1)[1].cpu() == y.cpu()).sum())


def train_model(model, dataloader, epoch_num, is_binary, verbose=True, lr=1e-3):
    """Adam(lr=1e-3) loop of the reference (utils_basic.py:94-118).  ``dataloader`` is any iterable of (x, y)
    batches, or an object with ``batches()``; ``model`` may be DDP-wrapped (uses ``module.loss``)."""
    from ...ops import optim as boptim
    core = model.module if hasattr(model, "module") else model
    model.train()
    optimizer = boptim.Adam(model.parameters(), lr=lr)
    hist = []
    for epoch in range(epoch_num):
        cum_loss, cum_acc, tot = 0.0, 0.0, 0.0
        it = dataloader.batches() if hasattr(dataloader, "batches") else dataloader
        for x_in, y_in in it:
            B = x_in.size()[0]
            pred = model(x_in)
            loss = core.loss(pred, y_in)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            cum_loss += float(loss.item()) * B
            cum_acc += _correct(pred.detach(), y_in, is_binary)
            tot += B
        hist.append((cum_loss / max(tot, 1), cum_acc / max(tot, 1)))
        if verbose:
            print("Epoch %d, loss = %.4f, acc = %.4f" % (epoch, hist[-1][0], hist[-1][1]))
    return hist


def eval_model(model, dataloader, is_binary, verbose=False):
    model.eval()
    cum_acc, tot = 0.0, 0.0
    with torch.no_grad():
        it = dataloader.batches() if hasattr(dataloader, "batches") else dataloader
        for x_in, y_in in it:
            pred = model(x_in)
            cum_acc += _correct(pred, y_in, is_binary)
            tot += x_in.size()[0]
    return cum_acc / max(tot, 1)
