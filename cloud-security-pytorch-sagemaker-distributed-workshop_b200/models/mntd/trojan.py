"""Trojan (backdoor) attack settings and trigger stamping for the four MNTD tasks.

``random_troj_setting(task, troj_type)`` draws ``(p_size, pattern, loc, alpha, target_y, inject_p)`` with the
reference's distributions (``notebooks/code/model_lib/mnist_cnn_model.py:38-72``,
``cifar10_cnn_model.py:43-75``, ``audio_rnn_model.py:47-75``, ``rtNLP_cnn_model.py:72-85``);
``troj_gen_func(task)(X, y, atk_setting)`` stamps ONE sample exactly as the reference does (``:74-81``,
``:77-84``, ``:77-83``, ``:87-99``) and is the oracle for ``stamp_batch``, which applies the same
alpha-blend to a whole device-resident batch in one kernel (``b200_trojan_stamp``) instead of
once per sample inside ``Dataset.__getitem__``.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

_MAX_SIZE = {"mnist": 28, "cifar10": 32, "audio": 16000}


def random_troj_setting(task: str, troj_type: str, rng=np.random):
    if task == "rtNLP":
        assert troj_type != "B", "No blending attack for NLP task"
        p_size = rng.randint(2) + 1                     # add 1 or 2 words
        loc = rng.randint(0, 10)
        alpha = 1.0
        pattern = rng.randint(18000, size=p_size)
        target_y = rng.randint(2)
        inject_p = rng.uniform(0.05, 0.5)
        return p_size, pattern, loc, alpha, target_y, inject_p

    max_size = _MAX_SIZE[task]
    sizes = [800, 1600, 2400, 3200] if task == "audio" else [2, 3, 4, 5]
    if troj_type == "jumbo":
        p_size = rng.choice(sizes + [max_size], 1)[0]
        if p_size < max_size:
            alpha = rng.uniform(0.2, 0.6)
            if alpha > 0.5:
                alpha = 1.0
        else:
            alpha = rng.uniform(0.05, 0.2)
    elif troj_type == "M":
        p_size = rng.choice(sizes, 1)[0]
        alpha = 1.0
    elif troj_type == "B":
        p_size = max_size
        alpha = rng.uniform(0.05, 0.2)
    else:
        raise ValueError("troj_type must be jumbo, M or B")
    p_size = int(p_size)

    if task == "audio":
        loc = rng.randint(max_size - p_size) if p_size < max_size else 0
        pattern = rng.uniform(size=p_size) * 0.2
    else:
        if p_size < max_size:
            loc = (rng.randint(max_size - p_size), rng.randint(max_size - p_size))
        else:
            loc = (0, 0)
        if task == "mnist":
            pattern_num = rng.randint(1, p_size ** 2)
            one_idx = rng.choice(list(range(p_size ** 2)), pattern_num, replace=False)
            flat = np.zeros((p_size ** 2))
            flat[one_idx] = 1
            pattern = np.reshape(flat, (p_size, p_size))
        else:
            eps = rng.uniform(0, 1)
            pattern = np.clip(rng.uniform(-eps, 1 + eps, size=(3, p_size, p_size)), 0, 1)
    target_y = rng.randint(10)
    inject_p = rng.uniform(0.05, 0.5)
    return p_size, pattern, loc, alpha, target_y, inject_p


def troj_gen_func(task: str):
    """Per-sample stamping with the reference's semantics."""
    def mnist(X, y, atk):
        p, pattern, (w, h), alpha, target_y, _ = atk
        Xn = X.clone()
        Xn[0, w:w + p, h:h + p] = alpha * torch.as_tensor(pattern, dtype=torch.float32) + (1 - alpha) * Xn[0, w:w + p, h:h + p]
        return Xn, target_y

    def cifar10(X, y, atk):
        p, pattern, (w, h), alpha, target_y, _ = atk
        Xn = X.clone()
        Xn[:, w:w + p, h:h + p] = alpha * torch.as_tensor(pattern, dtype=torch.float32) + (1 - alpha) * Xn[:, w:w + p, h:h + p]
        return Xn, target_y

    def audio(X, y, atk):
        p, pattern, loc, alpha, target_y, _ = atk
        Xn = X.clone()
        Xn[loc:loc + p] = alpha * torch.as_tensor(pattern, dtype=torch.float32) + (1 - alpha) * Xn[loc:loc + p]
        return Xn, target_y

    def rtnlp(X, y, atk):
        p, pattern, loc, alpha, target_y, _ = atk
        lst = list(X.numpy())
        x_len = lst.index(0) if 0 in lst else len(lst)
        ins = min(x_len, loc)
        Xn = torch.cat([X[:ins], torch.as_tensor(np.asarray(pattern), dtype=torch.long), X[ins:]], dim=0)
        return Xn, target_y

    return {"mnist": mnist, "cifar10": cifar10, "audio": audio, "rtNLP": rtnlp}[task]


def stamp_batch(task: str, X: torch.Tensor, atk_setting, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Stamp the trigger into every sample of ``X`` with ``mask[i] != 0`` (all when ``mask`` is None), in place.

    ``X``: mnist ``[N,1,28,28]``, cifar10 ``[N,3,32,32]``, audio ``[N,16000]`` (fp32).  On CUDA this is one
    launch of the native kernel; on CPU the same blend with torch ops.
    """
    p, pattern, loc, alpha, _, _ = atk_setting
    pat = torch.as_tensor(np.asarray(pattern), dtype=torch.float32, device=X.device)
    if task == "audio":
        x4 = X.view(X.shape[0], 1, 1, X.shape[1])
        pat3, lh, lw = pat.view(1, 1, -1), 0, int(loc)
    elif task == "mnist":
        x4, pat3, (lh, lw) = X, pat.view(1, p, p), loc
    elif task == "cifar10":
        x4, pat3, (lh, lw) = X, pat.view(3, p, p), loc
    else:
        raise ValueError("stamp_batch handles the continuous tasks; rtNLP splices tokens per sample")
    if X.is_cuda:
        from ...ops import native as nv
        m8 = None if mask is None else mask.to(torch.uint8).contiguous()
        nv.trojan_stamp_(x4, pat3.contiguous(), int(lh), int(lw), float(alpha), m8)
        return X
    ph, pw = pat3.shape[1], pat3.shape[2]
    sel = slice(None) if mask is None else mask.bool()
    region = x4[sel, : pat3.shape[0], lh:lh + ph, lw:lw + pw]
    x4[sel, : pat3.shape[0], lh:lh + ph, lw:lw + pw] = alpha * pat3 + (1 - alpha) * region
    return X
