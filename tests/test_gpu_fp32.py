"""The fp32 compute path (fp32 activations / weights / gradients in memory, TF32 tensor-core products with fp32
accumulation: ``tcgen05.mma kind::tf32``) against a plain PyTorch fp32 oracle evaluated ON THE CPU.

This is the precision the reference trains at (``notebooks/code/cifar10-distributed-smddp-gpu.py:145,155-168``:
fp32 model, no autocast; torch lets cuDNN use TF32 for convolutions).  Every check has two bounds: a relative
Frobenius error, and an elementwise max-abs error scaled by sqrt(K) (so one bad border row or one bad tile tail
in a many-tile problem cannot hide in the norm).
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CL = torch.channels_last
TF32_EPS = 2.0 ** -10          # operands keep 10 explicit mantissa bits


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda", 0)


def _check(got, want, k_red, what, rel_tol=2e-3):
    """got: CUDA tensor, want: CPU fp32 oracle; k_red: length of the reduction behind every output element."""
    g, w = got.detach().float().cpu(), want.detach().float()
    assert g.shape == w.shape, (what, g.shape, w.shape)
    rel = float((g - w).norm() / (w.norm() + 1e-30))
    scale = float(w.abs().max()) + 1e-30
    max_abs = float((g - w).abs().max())
    # every product carries a relative error of ~eps, the sum of k_red of them an absolute error of ~eps * sqrt(k_red) * rms
    # while the outputs themselves are ~sqrt(k_red) * rms: so the elementwise error is a few eps of the largest output
    bound = 8.0 * TF32_EPS * scale
    assert rel < rel_tol, "%s: relative error %.3e" % (what, rel)
    assert max_abs <= bound, "%s: max abs error %.3e > %.3e (largest |output| %.3e)" % (what, max_abs, bound, scale)


CONV_CASES = [
    # N, C, H, W, K, R, stride, pad
    (4, 64, 8, 8, 64, 3, 1, 1),
    (2, 64, 8, 8, 128, 3, 2, 1),
    (8, 128, 4, 4, 128, 3, 1, 1),
    (32, 256, 2, 2, 512, 3, 2, 1),
    (32, 512, 1, 1, 512, 3, 1, 1),
    (4, 64, 8, 8, 128, 1, 2, 0),
    (2, 8, 32, 32, 64, 7, 2, 3),
    (3, 16, 14, 14, 8, 5, 1, 0),
    (2, 32, 17, 13, 72, 3, 1, 1),      # odd sizes, Cout not a multiple of 64, tile tails in M and N
    (2, 64, 56, 56, 64, 3, 1, 1),
    (8, 512, 2, 2, 512, 3, 1, 1),      # split-K with many splits
    (256, 512, 1, 1, 512, 3, 1, 1),    # 8 of 9 taps dead
    (64, 128, 4, 4, 256, 1, 2, 0),     # 1x1 stride 2
    (16, 64, 56, 56, 64, 3, 1, 1),     # 392 tiles: 3-stage two-CTAs-per-SM variant
    (16, 256, 56, 56, 64, 1, 1, 0),    # both operands on TMA
    (5, 24, 9, 11, 40, 3, 1, 1),       # channel counts that are multiples of 8 but not of 32 (k-blocks straddle taps)
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_tf32_fwd_dgrad_wgrad(case):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(0)
    N, C, H, W, K, R, stride, pad = case
    x = torch.randn(N, C, H, W)
    w = torch.randn(K, C, R, R) / (C * R * R) ** 0.5
    bias = torch.randn(K)
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    yr = F.conv2d(xr, wr, bias, stride=stride, padding=pad)
    xd = x.to(dev).contiguous(memory_format=CL)
    w_krsc = w.permute(0, 2, 3, 1).contiguous().to(dev)
    y = nv.conv_fwd(xd, w_krsc, bias.to(dev), stride, pad, relu=False)
    assert y.dtype == torch.float32
    _check(y, yr, C * R * R, "fwd")
    yrelu = nv.conv_fwd(xd, w_krsc, bias.to(dev), stride, pad, relu=True)
    _check(yrelu, F.relu(yr), C * R * R, "fwd+relu")
    dy = torch.randn(*yr.shape)
    yr.backward(dy)
    dyd = dy.to(dev).contiguous(memory_format=CL)
    dx = nv.conv_dgrad(dyd, w_krsc, xd.shape, stride, pad)
    _check(dx, xr.grad, K * R * R, "dgrad")
    res = torch.randn(N, C, H, W)
    dx2 = nv.conv_dgrad(dyd, w_krsc, xd.shape, stride, pad, residual=res.to(dev).contiguous(memory_format=CL))
    _check(dx2, xr.grad + res, K * R * R, "dgrad+residual")
    dw = torch.zeros(K, R, R, C, device=dev)
    nv.conv_wgrad_(dw, xd, dyd, stride, pad)
    _check(dw.permute(0, 3, 1, 2), wr.grad, N * yr.shape[2] * yr.shape[3], "wgrad")
    nv.conv_wgrad_(dw, xd, dyd, stride, pad)      # accumulates
    _check(dw.permute(0, 3, 1, 2), 2 * wr.grad, N * yr.shape[2] * yr.shape[3], "wgrad accumulate")


def test_conv_tf32_bn_stats_epilogue():
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(3)
    N, C, H, W, K = 32, 64, 8, 8, 128
    x = torch.randn(N, C, H, W)
    w = torch.randn(K, C, 3, 3) / (C * 9) ** 0.5
    yr = F.conv2d(x, w, None, stride=1, padding=1)
    sums = torch.zeros(2 * K + 1, device=dev)
    y = nv.conv_fwd(x.to(dev).contiguous(memory_format=CL), w.permute(0, 2, 3, 1).contiguous().to(dev), None, 1, 1, False, sums)
    _check(y, yr, C * 9, "fwd (stats variant)")
    s_ref = yr.sum(dim=(0, 2, 3))
    q_ref = (yr * yr).sum(dim=(0, 2, 3))
    assert float((sums[:K].cpu() - s_ref).abs().max()) < 2e-3 * float(q_ref.max()) ** 0.5 * (N * H * W) ** 0.5
    assert float((sums[K:2 * K].cpu() - q_ref).norm() / q_ref.norm()) < 2e-3


@pytest.mark.parametrize("M,K,N", [(32, 512, 1000), (256, 400, 120), (100, 4096, 256), (7, 88, 16), (300, 2048, 1000)])
def test_linear_tf32(M, K, N):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(1)
    x = torch.randn(M, K)
    w = torch.randn(N, K) / K ** 0.5
    b = torch.randn(N)
    y = nv.linear_fwd(x.to(dev), w.to(dev), b.to(dev), relu=False)
    _check(y, F.linear(x, w, b), K, "linear fwd")
    dy = torch.randn(M, N)
    dx = nv.linear_dgrad(dy.to(dev), w.to(dev))
    _check(dx, dy @ w, N, "linear dgrad")
    dw = torch.zeros(N, K, device=dev)
    nv.linear_wgrad_(dw, x.to(dev), dy.to(dev))
    _check(dw, dy.t() @ x, M, "linear wgrad")


@pytest.mark.parametrize("shape,relu,res", [((8, 64, 8, 8), True, False), ((8, 128, 4, 4), True, True),
                                            ((32, 512, 1, 1), False, False), ((4, 256, 14, 14), True, True),
                                            ((2, 64, 112, 112), True, False), ((32, 64, 8, 8), True, True)])
def test_batchnorm_fp32(shape, relu, res):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(2)
    N, C, H, W = shape
    rows = N * H * W
    x = torch.randn(N, C, H, W) * 2.0 + 0.3
    r = torch.randn(N, C, H, W) if res else None
    gamma, beta = torch.rand(C) + 0.5, torch.randn(C)
    rm_ref, rv_ref = torch.zeros(C), torch.ones(C)
    xr = x.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rr = r.clone().requires_grad_(True) if res else None
    pre = F.batch_norm(xr, rm_ref, rv_ref, gr, br, True, 0.1, 1e-5)
    if res:
        pre = pre + rr
    yr = F.relu(pre) if relu else pre
    dy = torch.randn(N, C, H, W)

    xd = x.to(dev).contiguous(memory_format=CL)
    rd = r.to(dev).contiguous(memory_format=CL) if res else None
    dyd = dy.to(dev).contiguous(memory_format=CL)
    gd, bd_ = gamma.to(dev), beta.to(dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    stat = nv.bn_stats(xd, gd, bd_, 1e-5, 0.1, rm, rv)
    y = nv.bn_apply(xd, rd, stat, relu)
    assert y.dtype == torch.float32
    assert float((y.cpu() - yr).abs().max()) < 1e-4 * (1.0 + float(yr.abs().max()))
    # the oracle's backward uses the kernel's ReLU mask (an output within rounding of zero may land on either side)
    pre.backward(dy * (y.cpu() > 0).float() if relu else dy)
    assert torch.allclose(rm.cpu(), rm_ref, atol=1e-5, rtol=1e-4) and torch.allclose(rv.cpu(), rv_ref, atol=1e-5, rtol=1e-4)
    dgamma, dbeta = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    dx, dres, _ = nv.bn_bwd(dyd, xd, y, stat, relu, res, dgamma, dbeta, 1.0)
    tol = 2e-4 * (1.0 + float(xr.grad.abs().max()))
    assert float((dx.cpu() - xr.grad).abs().max()) < tol, "two-kernel backward"
    assert float((dgamma.cpu() - gr.grad).abs().max()) < 1e-3 * (1 + float(gr.grad.abs().max()))
    assert float((dbeta.cpu() - br.grad).abs().max()) < 1e-3 * (1 + float(br.grad.abs().max()))
    if res:
        assert float((dres.cpu() - rr.grad).abs().max()) < 1e-5
    # one-launch variants (CTA kernel for small tensors, cluster kernel above)
    if rows <= nv.BN_SMALL_ROWS:
        rm2, rv2 = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        y2, stat2 = nv.bn_fwd_fused(xd, rd, gd, bd_, 1e-5, 0.1, rm2, rv2, relu)
        assert float((y2.cpu() - yr).abs().max()) < 1e-4 * (1.0 + float(yr.abs().max()))
    if rows <= 4 * 256 * 8:
        dg2, db2 = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        dx2, dres2 = nv.bn_bwd_fused(dyd, xd, y, stat, relu, res, dg2, db2, 0.0)
        assert float((dx2.cpu() - xr.grad).abs().max()) < tol, "one-launch backward"
        assert float((dg2.cpu() - gr.grad).abs().max()) < 1e-3 * (1 + float(gr.grad.abs().max()))


def test_pools_relu_colsum_fp32():
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(4)
    x = torch.randn(4, 64, 16, 16)
    xr = x.clone().requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 2, 1)
    xd = x.to(dev).contiguous(memory_format=CL)
    y, amax = nv.maxpool_fwd(xd, 3, 2, 1)
    assert torch.equal(y.cpu(), yr.detach())
    dy = torch.randn(*yr.shape)
    yr.backward(dy)
    dx = nv.maxpool_bwd(dy.to(dev).contiguous(memory_format=CL), amax, xd.shape, 3, 2, 1)
    assert float((dx.cpu() - xr.grad).abs().max()) < 1e-6
    a = nv.avgpool_fwd(xd)
    assert float((a.cpu() - x.mean(dim=(2, 3))).abs().max()) < 1e-5
    da = torch.randn(4, 64)
    dxa = nv.avgpool_bwd(da.to(dev), xd.shape)
    assert float((dxa.cpu() - (da / 256.0)[:, :, None, None].expand(4, 64, 16, 16)).abs().max()) < 1e-6
    g = torch.randn(4, 64, 16, 16)
    r = nv.relu_bwd(g.to(dev).contiguous(memory_format=CL), xd)
    assert torch.equal(r.cpu(), g * (x > 0))
    out = torch.zeros(64, device=dev)
    nv.colsum_accum_(out, g.to(dev).contiguous(memory_format=CL), 64)
    assert float((out.cpu() - g.sum(dim=(0, 2, 3))).abs().max()) < 1e-3


def test_augment_fp32_matches_reference_transform():
    from b200ddp.ops import native as nv
    from b200ddp.utils.data import CIFAR10_MEAN, CIFAR10_STD, augment_reference
    dev = _dev()
    torch.manual_seed(5)
    x = torch.randint(0, 256, (16, 32, 32, 3), dtype=torch.uint8)
    off = torch.stack([torch.randint(0, 9, (16,)), torch.randint(0, 9, (16,)), torch.randint(0, 2, (16,))], 1).to(torch.int32)
    want = augment_reference(x, train_transform=True, mean=CIFAR10_MEAN, std=CIFAR10_STD, offsets=off)
    mean = torch.tensor(CIFAR10_MEAN, device=dev)
    inv = 1.0 / torch.tensor(CIFAR10_STD, device=dev)
    got = nv.augment(x.to(dev), off.to(dev), mean, inv, c_out=8, dtype=torch.float32)
    assert got.dtype == torch.float32 and got.shape == (16, 8, 32, 32)
    assert float((got[:, :3].cpu() - want).abs().max()) < 1e-5
    assert float(got[:, 3:].abs().max()) == 0.0


def test_workshop_net_fp32_matches_cpu_oracle():
    """The workshop Net through the tensor-core layers in fp32/TF32 mode (channel counts padded 3->8, 6->8, 84->88, 10->16):
    logits, loss and every parameter gradient against stock torch on the CPU.  (resnet18: tests/test_gpu_chain.py.)"""
    import copy
    import b200ddp as bd
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import functional as BF
    dev = _dev()
    torch.manual_seed(11)
    bd.set_compute_dtype("fp32")
    BF.assume_unit_loss_grad(False)
    try:
        ref = get_model("custom")
        ours = copy.deepcopy(ref).to(dev)
        x = torch.randn(64, 3, 32, 32)
        y = torch.randint(0, 10, (64,))
        crit = CrossEntropyLoss()
        lr = crit(ref(x), y)
        lr.backward()
        out = ours(x.to(dev))
        assert out.dtype == torch.float32
        lo = crit(out.contiguous(), y.to(dev))
        lo.backward()
        torch.cuda.synchronize()
        assert abs(float(lo) - float(lr)) < 2e-3 * (1.0 + abs(float(lr))), (float(lo), float(lr))
        for (n, p), (_, q) in zip(ours.named_parameters(), ref.named_parameters()):
            assert p.grad is not None, n
            rel = float((p.grad.cpu() - q.grad).norm() / (q.grad.norm() + 1e-12))
            # Every gradient here sits upstream of a ReLU (fused into the conv / linear epilogues) and most of a 2x2 max-pool:
            # an activation within TF32 rounding of zero, or a pair within rounding of a tie, sends its gradient elsewhere
            # (the effect tests/test_gpu_chain.py removes by replaying the masks); a few such elements per thousand move a
            # Frobenius norm by percents while every kernel on its own is within 2e-3 (test_conv_tf32_fwd_dgrad_wgrad and
            # test_linear_tf32 cover these very shapes).  Loss and logits above are the tight part of this test.
            assert rel < 5e-2, "gradient of %s: relative error %.3e" % (n, rel)
    finally:
        bd.set_compute_dtype("bf16")
