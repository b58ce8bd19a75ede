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
        self.conv1, self.bn1, self.maxpool = m.conv1, m.bn1, m.maxpool

    def forward(self, x):
        return self.maxpool(self.bn1(self.conv1(x)))


class _Head(torch.nn.Module):
    def __init__(self, m):
        super().__init__()
        self.avgpool, self.fc = m.avgpool, m.fc

    def forward(self, x):
        return self.fc(self.avgpool(x))


def _stages(model):
    st = [("stem", _Stem(model))]
    for li in (1, 2, 3, 4):
        for bi, blk in enumerate(getattr(model, "layer%d" % li)):
            st.append(("layer%d.%d" % (li, bi), blk))
    st.append(("head", _Head(model)))
    return st


# (compute dtype, bound on block outputs / input gradients, bound on parameter gradients)
@pytest.mark.parametrize("dtype,out_tol,grad_tol", [("fp32", 5e-3, 1e-2), ("bf16", 2e-2, 5e-2)])
def test_resnet18_blocks_chained(dtype, out_tol, grad_tol):
    import b200ddp as bd
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

    # ---- oracle pass: record every stage's input, output and output gradient
    rec = []
    h = x
    for name, stage in _stages(ref):
        inp = h.detach().clone().requires_grad_(True)
        out = stage(inp)
        rec.append({"name": name, "inp": inp, "out": out})
        h = out
    loss = F.cross_entropy(rec[-1]["out"], y)
    g = torch.autograd.grad(loss, rec[-1]["out"])[0]
    for r in reversed(rec):                     # stage-local backward: the grad of a stage's input feeds the one before it
        r["gout"] = g
        r["out"].backward(g)
        g = r["inp"].grad
    ref_grads = {n: p.grad.clone() for n, p in ref.named_parameters()}

    # ---- native stages, each fed the oracle's tensors
    bd.set_compute_dtype(dtype)
    BF.assume_unit_loss_grad(False)
    report, bad = {}, {}
    try:
        for (name, stage), r in zip(_stages(ours), rec):
            inp = r["inp"].detach().to(dev).requires_grad_(name != "stem")
            out = stage(inp)
            e_out = _rel(out, r["out"])
            out.float().backward(r["gout"].to(dev))
            torch.cuda.synchronize()
            errs = {"out": e_out}
            if name != "stem":
                errs["dinp"] = _rel(inp.grad, r["inp"].grad)
            if e_out > out_tol:
                bad[name + ".out"] = e_out
            if errs.get("dinp", 0.0) > out_tol * 2:
                bad[name + ".dinp"] = errs["dinp"]
            for pn, p in stage.named_parameters():
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
