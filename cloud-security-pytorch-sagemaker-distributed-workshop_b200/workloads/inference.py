"""SageMaker PyTorch serving hooks (``notebooks/code/inference.py:28-34`` + the container defaults the
notebook relies on, ``notebooks/1_pytorch_dist_native_cpu.ipynb:153-211``).

``model_fn(model_dir)`` rebuilds the workshop ``Net`` and loads ``model.pth`` on CPU.  Unlike the
reference it also accepts DDP-prefixed (``module.*``) checkpoints and, given a ``model_type.txt`` side
file or a ResNet-shaped state dict, the ResNet models — the reference's own GPU trainer writes files its
``model_fn`` cannot load.  ``input_fn``/``predict_fn``/``output_fn`` reproduce the default
npy/json handlers.
"""
from __future__ import annotations

import io
import json
import os

import numpy as np
import torch

from ..models import get_model


def _guess_model_type(sd) -> str:
    if "fc3.weight" in sd:
        return "custom"
    if "layer1.0.conv3.weight" in sd:
        return "resnet50"
    return "resnet18"


def model_fn(model_dir: str):
    device = "cpu"
    with open(os.path.join(model_dir, "model.pth"), "rb") as f:
        sd = torch.load(f, map_location=device, weights_only=True)
    sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in sd.items()}
    model = get_model(_guess_model_type(sd))
    model.load_state_dict(sd)
    return model.to(device).eval()


def input_fn(request_body, content_type="application/x-npy"):
    if content_type == "application/x-npy":
        return torch.from_numpy(np.load(io.BytesIO(request_body), allow_pickle=False)).float()
    if content_type == "application/json":
        return torch.tensor(json.loads(request_body), dtype=torch.float32)
    raise ValueError("unsupported content type %s" % content_type)


def predict_fn(data, model):
    with torch.no_grad():
        return model(data)


def output_fn(prediction, accept="application/json"):
    arr = prediction.detach().cpu().numpy()
    if accept == "application/x-npy":
        buf = io.BytesIO()
        np.save(buf, arr)
        return buf.getvalue()
    return json.dumps(arr.tolist())
