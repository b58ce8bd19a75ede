"""Does the native path TRAIN like the reference's numerics?  300 SGD steps of resnet18 (torchvision topology, 1000-way
head, the workshop's GPU job: ``notebooks/code/cifar10-distributed-smddp-gpu.py:30-32,155-168``) on the learnable
synthetic CIFAR-10 stand-in, three ways from the same initial weights and the same batches:

  oracle   stock ``torchvision.models.resnet18`` + ``nn.CrossEntropyLoss`` + ``optim.SGD`` in strict fp32 (TF32 off)
  fp32     this framework, fp32 storage + TF32 tensor cores (the bench default)
  bf16     this framework, bf16 tensor cores with fp32 masters

The loss curves (means over 50-step windows) must stay inside a band around the oracle's and the final held-out
accuracy must match: that is what licenses a reduced-precision product path as a stand-in for the reference's fp32 run.
"""
import copy

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu

STEPS, BATCH, WIN = 300, 64, 50


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda", 0)


def _normalise(x_u8, dev):
    from b200ddp.utils.data import CIFAR10_MEAN, CIFAR10_STD
    mean = torch.tensor(CIFAR10_MEAN, device=dev)[None, :, None, None]
    std = torch.tensor(CIFAR10_STD, device=dev)[None, :, None, None]
    return (x_u8.to(dev).permute(0, 3, 1, 2).float() / 255.0 - mean) / std


def _accuracy(model, xs, ys, dev, native):
    model.eval()
    correct = 0
    with torch.no_grad():
        for i in range(0, xs.shape[0], 500):
            out = model(_normalise(xs[i:i + 500], dev))
            correct += int((out.float()[:, :10].argmax(1).cpu() == ys[i:i + 500]).sum())
    model.train()
    return correct / xs.shape[0]


def _windows(losses):
    return [sum(losses[i:i + WIN]) / WIN for i in range(0, len(losses), WIN)]


def test_resnet18_trains_like_torch_fp32():
    import torchvision
    import b200ddp as bd
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import optim as boptim
    from b200ddp.runtime.engine import TrainStep
    from b200ddp.utils.data import SyntheticCIFAR10
    dev = _dev()
    # heavy pixel noise, and a quarter of the TRAINING labels re-drawn at random: the loss cannot collapse to zero in the
    # first window, so all six windows carry information (the held-out labels stay clean)
    train = SyntheticCIFAR10(train=True, size=STEPS * BATCH, noise=120.0)
    test = SyntheticCIFAR10(train=False, size=2000, noise=120.0)
    g = torch.Generator().manual_seed(99)
    labels = torch.from_numpy(train.targets).clone()
    flip = torch.rand(labels.shape[0], generator=g) < 0.25
    labels[flip] = torch.randint(0, 10, (int(flip.sum()),), generator=g)
    xs = torch.from_numpy(train.data).view(STEPS, BATCH, 32, 32, 3)
    ys = labels.view(STEPS, BATCH)
    tx, ty = torch.from_numpy(test.data), torch.from_numpy(test.targets)

    torch.manual_seed(5)
    init = get_model("resnet18").state_dict()

    # ---- oracle: stock torch, strict fp32
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        oracle = torchvision.models.resnet18()
        oracle.load_state_dict(init)
        oracle = oracle.to(dev).train()
        opt = torch.optim.SGD(oracle.parameters(), lr=0.01, momentum=0.9)
        crit = nn.CrossEntropyLoss()
        ref_losses = []
        for i in range(STEPS):
            opt.zero_grad()
            loss = crit(oracle(_normalise(xs[i], dev)), ys[i].to(dev))
            loss.backward()
            opt.step()
            ref_losses.append(float(loss))
        ref_acc = _accuracy(oracle, tx, ty, dev, False)
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved
    ref_w = _windows(ref_losses)

    started = bd.is_initialized()
    if not started:
        bd.init_process_group(backend="b200", rank=0, world_size=1)
    results = {}
    try:
        for dtype in ("fp32", "bf16"):
            bd.set_compute_dtype(dtype)
            net = get_model("resnet18")
            net.load_state_dict(init)
            model = bd.DistributedDataParallel(net.to(dev))
            sgd = boptim.SGD(model.parameters(), lr=0.01, momentum=0.9)
            step = TrainStep(model, CrossEntropyLoss(), sgd, BATCH, device=dev, graph=True, augment=False)
            losses = [step(xs[i].pin_memory(), ys[i].pin_memory(), need_loss=True) for i in range(STEPS)]
            step.flush_counters()
            results[dtype] = (_windows(losses), _accuracy(model, tx, ty, dev, True))
            del step, sgd, model
    finally:
        bd.set_compute_dtype("bf16")
        if not started:
            bd.destroy_process_group()
    print("oracle fp32 windows", [round(v, 3) for v in ref_w], "acc", ref_acc)
    for dtype, (w, acc) in results.items():
        print(dtype, "windows", [round(v, 3) for v in w], "acc", acc)
    assert ref_w[-1] < 0.7 * ref_w[0] and ref_acc > 0.8, "the oracle itself did not learn the synthetic set"
    for dtype, (w, acc) in results.items():
        band = 0.10 if dtype == "fp32" else 0.15
        for k, (a, b) in enumerate(zip(w, ref_w)):
            assert abs(a - b) <= band * (1.0 + b), "%s: window %d loss %.3f vs oracle %.3f" % (dtype, k, a, b)
        assert abs(acc - ref_acc) <= 0.03, "%s: held-out accuracy %.3f vs oracle %.3f" % (dtype, acc, ref_acc)
