"""Per-block chained check of resnet18 on the native kernels against stock PyTorch fp32 on the CPU.

A whole-network comparison of resnet18 at random initialisation says little: layer4 batch-normalises 1x1 feature
maps over 32 samples and every BN backward cancels a large common mode, so rounding noise of ANY reduced-precision
evaluation is amplified layer after layer (``profiles/determinism.md``) and a real bug in an early layer could hide
behind a loose whole-model tolerance.  Here the oracle (the same model, plain torch ops, fp32, CPU) is run once and
the input and the output-gradient of every stage are recorded; each stage of the native model (stem, the 8 residual
blocks, the head) is then fed the ORACLE's input and the ORACLE's output gradient, and its output, its input
gradient and all its parameter gradients are compared -- so an error cannot travel further than one block.
"""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda", 0)


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


class _Stem(torch.nn.Module):
    def __init__(self, m):
        super().__init__()
        self.conv1, self.bn1 = m.conv1, m.bn1

    def forward(self, x):
        return self.bn1(self.conv1(x))


class _Pool(torch.nn.Module):
    """The stem's max-pool as a stage of its own: fed identical inputs, both sides route every gradient to the same
    element; inside the stem a near-tie between two activations would send it to different ones (as a ReLU flip does)."""

    def __init__(self, m):
        super().__init__()
        self.maxpool = m.maxpool

    def forward(self, x):
        return self.maxpool(x)


class _Head(torch.nn.Module):
    def __init__(self, m):
        super().__init__()
        self.avgpool, self.fc = m.avgpool, m.fc

    def forward(self, x):
        return self.fc(self.avgpool(x))


def _stages(model):
    st = [("stem", _Stem(model)), ("maxpool", _Pool(model))]
    for li in (1, 2, 3, 4):
        for bi, blk in enumerate(getattr(model, "layer%d" % li)):
            st.append(("layer%d.%d" % (li, bi), blk))
    st.append(("head", _Head(model)))
    return st


class _MaskedRelu:
    """The oracle's ReLUs, replayed with the masks the native stage produced.  An activation within rounding of zero lands
    on either side of the ReLU depending on the arithmetic; a handful of such flips changes a gradient's Frobenius norm by
    sqrt(fraction flipped) -- 1.5e-2 for one element in 4000 -- which says nothing about a kernel.  With the oracle using
    the SAME masks, what is left is the arithmetic error of the kernels."""

    def __init__(self, masks):
        self.masks, self.i, self.real = masks, 0, F.relu

    def __call__(self, y, inplace=False):
        m = self.masks[self.i]
        self.i += 1
        assert m.shape == y.shape, (m.shape, y.shape)
        return y * m

    def __enter__(self):
        F.relu = self
        return self

    def __exit__(self, *exc):
        F.relu = self.real


# (compute dtype, bound on block outputs / input gradients, bound on parameter gradients)
@pytest.mark.parametrize("dtype,out_tol,grad_tol", [("fp32", 2e-3, 5e-3), ("bf16", 2e-2, 3e-2)])
def test_resnet18_blocks_chained(dtype, out_tol, grad_tol):
    import b200ddp as bd
    from b200ddp import nn as bnn
    from b200ddp.models import get_model
    from b200ddp.ops import functional as BF
    dev = _dev()
    torch.manual_seed(7)
    batch = 32
    ref = get_model("resnet18")
    ours = copy.deepcopy(ref).to(dev)
    ref.train()
    ours.train()
    x = torch.randn(batch, 3, 32, 32)
    y = torch.randint(0, 1000, (batch,))

    bd.set_compute_dtype(dtype)
    BF.assume_unit_loss_grad(False)
    report, bad = {}, {}
    try:
        # ---- forward, stage by stage: the native stage runs on the ORACLE's input, then the oracle runs with its masks
        rec = []
        h = x
        for (name, stage_o), (_, stage_n) in zip(_stages(ref), _stages(ours)):
            masks, hooks = [], []
            for m in stage_n.modules():
                if isinstance(m, bnn.BatchNorm2d) and m.fused_relu:
                    hooks.append(m.register_forward_hook(lambda mod, i, o: masks.append((o.detach() > 0).float().cpu())))
            n_inp = h.detach().to(dev).requires_grad_(name != "stem")
            n_out = stage_n(n_inp)
            torch.cuda.synchronize()
            for hk in hooks:
                hk.remove()
            o_inp = h.detach().clone().requires_grad_(True)
            with _MaskedRelu(masks) as mr:
                o_out = stage_o(o_inp)
            assert mr.i == len(masks), (name, mr.i, len(masks))
            rec.append({"name": name, "o_inp": o_inp, "o_out": o_out, "n_inp": n_inp, "n_out": n_out, "stage_n": stage_n})
            # bf16 mode: the next stage's input is the oracle's output rounded to bf16 on BOTH sides (the native layers
            # would round it anyway; rounding can turn a near-tie in front of the max-pool into an exact one)
            h = o_out.detach().to(torch.bfloat16).float() if dtype == "bf16" else o_out
        # ---- oracle backward, stage-local: the gradient of a stage's input feeds the stage before it
        loss = F.cross_entropy(rec[-1]["o_out"], y)
        g = torch.autograd.grad(loss, rec[-1]["o_out"])[0]
        for r in reversed(rec):
            r["gout"] = g
            r["o_out"].backward(g)
            g = r["o_inp"].grad
        ref_grads = {n: p.grad.clone() for n, p in ref.named_parameters()}
        # ---- native backward of every stage from the ORACLE's output gradient
        for r in rec:
            name = r["name"]
            errs = {"out": _rel(r["n_out"], r["o_out"])}
            r["n_out"].float().backward(r["gout"].to(dev))
            torch.cuda.synchronize()
            if name != "stem":
                errs["dinp"] = _rel(r["n_inp"].grad, r["o_inp"].grad)
            if errs["out"] > out_tol:
                bad[name + ".out"] = errs["out"]
            if errs.get("dinp", 0.0) > grad_tol:
                bad[name + ".dinp"] = errs["dinp"]
            for pn, p in r["stage_n"].named_parameters():
                full = pn if name in ("stem", "head") else name + "." + pn
                assert p.grad is not None, full
                e = _rel(p.grad, ref_grads[full])
                errs[pn] = e
                if e > grad_tol:
                    bad[full] = e
            report[name] = {k: float("%.2e" % v) for k, v in errs.items()}
    finally:
        bd.set_compute_dtype("bf16")
    print(dtype, report)
    assert not bad, (bad, report)
