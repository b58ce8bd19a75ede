"""Numerics of every native kernel against a plain PyTorch fp32 reference of the same op (GPU)."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CL = torch.channels_last


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch.device("cuda", 0)


def _act(n, c, h, w, dev, scale=1.0):
    x = (torch.randn(n, c, h, w, device=dev) * scale).to(torch.bfloat16).contiguous(memory_format=CL)
    return x


@pytest.fixture(autouse=True)
def _strict_fp32_oracle():
    """The oracles below are torch ops on fp32 GPU tensors: keep cuDNN / cuBLAS from evaluating THEM in TF32."""
    saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved


BF16_EPS = 2.0 ** -8


def _rel_err(a, b):
    """Relative Frobenius error; also ASSERTS an elementwise bound (a few bf16 ulps of the largest reference value), so a
    wrong border row or one bad tile tail of a many-tile problem cannot hide inside the norm."""
    a, b = a.float(), b.float()
    if a.shape == b.shape and b.numel() > 0:
        max_abs = float((a - b).abs().max())
        scale = float(b.abs().max())
        assert max_abs <= 6.0 * BF16_EPS * scale + 1e-6, "elementwise error %.3e exceeds %.3e (largest reference value %.3e)" % (
            max_abs, 6.0 * BF16_EPS * scale, scale)
    return float((a - b).norm() / (b.norm() + 1e-12))


def _rel(a, b):
    return _rel_err(a.detach().cpu(), b.detach().cpu())


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
    (2, 32, 17, 13, 72, 3, 1, 1),
    (2, 64, 56, 56, 64, 3, 1, 1),
    (8, 512, 2, 2, 512, 3, 1, 1),     # split-K with many splits
    (256, 512, 1, 1, 512, 3, 1, 1),   # 8 of 9 taps dead: tap skipping + split-K
    (64, 128, 4, 4, 256, 1, 2, 0),    # 1x1 stride 2: dgrad tiles with no live tap
    (16, 64, 56, 56, 64, 3, 1, 1),    # 392 tiles: the 3-stage, two-CTAs-per-SM variant (cp.async gathers)
    (16, 256, 56, 56, 64, 1, 1, 0),   # same variant with both operands on TMA (1x1)
    (24, 64, 56, 56, 256, 1, 1, 0),   # ... and 128-wide N tiles
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fwd_dgrad_wgrad(case):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(0)
    N, C, H, W, K, R, stride, pad = case
    x = _act(N, C, H, W, dev)
    w = (torch.randn(K, C, R, R, device=dev) / (C * R * R) ** 0.5)
    w_krsc = w.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
    bias = torch.randn(K, device=dev)
    xr = x.float().requires_grad_(True)
    wr = w_krsc.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    yr = F.conv2d(xr, wr, bias, stride=stride, padding=pad)
    y = nv.conv_fwd(x, w_krsc, bias, stride, pad, relu=False)
    assert y.shape == yr.shape
    assert _rel_err(y, yr) < 1e-2, "fwd"
    yrelu = nv.conv_fwd(x, w_krsc, bias, stride, pad, relu=True)
    assert _rel_err(yrelu, F.relu(yr)) < 1e-2, "fwd+relu"
    dy = _act(*yr.shape, dev)
    yr.backward(dy.float())
    dx = nv.conv_dgrad(dy, w_krsc, x.shape, stride, pad)
    assert _rel_err(dx, xr.grad) < 1e-2, "dgrad"
    dw = torch.zeros(K, R, R, C, device=dev)
    nv.conv_wgrad_(dw, x, dy, stride, pad)
    assert _rel_err(dw.permute(0, 3, 1, 2), wr.grad) < 1e-2, "wgrad"
    nv.conv_wgrad_(dw, x, dy, stride, pad)      # accumulates
    assert _rel_err(dw.permute(0, 3, 1, 2), 2 * wr.grad) < 1e-2, "wgrad accumulate"


@pytest.mark.parametrize("M,K,N", [(32, 512, 1000), (256, 400, 120), (100, 4096, 256), (7, 88, 16), (300, 2048, 1000)])
def test_linear(M, K, N):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(1)
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) / K ** 0.5).to(torch.bfloat16)
    b = torch.randn(N, device=dev)
    y = nv.linear_fwd(x, w, b, relu=False)
    yr = F.linear(x.float(), w.float(), b)
    assert _rel_err(y, yr) < 1e-2
    dy = torch.randn(M, N, device=dev).to(torch.bfloat16)
    dx = nv.linear_dgrad(dy, w)
    assert _rel_err(dx, dy.float() @ w.float()) < 1e-2
    dw = torch.zeros(N, K, device=dev)
    nv.linear_wgrad_(dw, x, dy)
    assert _rel_err(dw, dy.float().t() @ x.float()) < 1e-2


@pytest.mark.parametrize("shape,relu,res", [((8, 64, 8, 8), True, False), ((8, 128, 4, 4), True, True),
                                            ((32, 512, 1, 1), False, False), ((4, 256, 14, 14), True, True),
                                            ((2, 64, 112, 112), True, False)])
def test_batchnorm(shape, relu, res):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(2)
    N, C, H, W = shape
    x = _act(N, C, H, W, dev, 2.0)
    r = _act(N, C, H, W, dev) if res else None
    gamma = torch.rand(C, device=dev) + 0.5
    beta = torch.randn(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    rm_ref, rv_ref = rm.clone(), rv.clone()
    stat = nv.bn_stats(x, gamma, beta, 1e-5, 0.1, rm, rv)
    y = nv.bn_apply(x, r, stat, relu)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rr = r.float().requires_grad_(True) if res else None
    yr = F.batch_norm(xr, rm_ref, rv_ref, gr, br, True, 0.1, 1e-5)
    if res:
        yr = yr + rr
    if relu:
        yr = F.relu(yr)
    assert _rel_err(y, yr) < 1e-2
    assert torch.allclose(rm, rm_ref, atol=1e-3, rtol=1e-3) and torch.allclose(rv, rv_ref, atol=1e-3, rtol=1e-3)
    dy = _act(N, C, H, W, dev)
    # the oracle must see the same ReLU mask as the kernel (mask from the bf16 output)
    g = dy.float() * ((y.float() > 0).float() if relu else 1.0)
    pre = F.batch_norm(xr, None, None, gr, br, True, 0.1, 1e-5)
    if res:
        pre = pre + rr
    pre.backward(g)
    dgamma, dbeta = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    dx, dres, sums = nv.bn_bwd(dy, x, y, stat, relu, res, dgamma, dbeta, 1.0)
    assert _rel_err(dx, xr.grad) < 2e-2
    assert _rel_err(dgamma, gr.grad) < 1e-2 and _rel_err(dbeta, br.grad) < 1e-2
    if res:
        assert _rel_err(dres, rr.grad) < 1e-2
    # scratch is self-resetting: a second launch gives the same statistics
    stat2 = nv.bn_stats(x, gamma, beta, 1e-5, 0.1, None, None)
    assert torch.equal(stat, stat2)


@pytest.mark.parametrize("shape,relu,res", [((32, 64, 8, 8), True, True), ((256, 512, 1, 1), True, False), ((16, 128, 4, 4), False, True)])
def test_batchnorm_fused_small_matches_two_pass(shape, relu, res):
    """Single-launch BN (small tensors) == the stats/apply and reduce/apply kernels, bit for bit on the statistics."""
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(6)
    N, C, H, W = shape
    x = _act(N, C, H, W, dev, 2.0)
    r = _act(N, C, H, W, dev) if res else None
    gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
    rm1, rv1, rm2, rv2 = torch.zeros(C, device=dev), torch.ones(C, device=dev), torch.zeros(C, device=dev), torch.ones(C, device=dev)
    stat_a = nv.bn_stats(x, gamma, beta, 1e-5, 0.1, rm1, rv1)
    y_a = nv.bn_apply(x, r, stat_a, relu)
    y_b, stat_b = nv.bn_fwd_fused(x, r, gamma, beta, 1e-5, 0.1, rm2, rv2, relu)
    assert torch.allclose(stat_a, stat_b, rtol=1e-5, atol=1e-6) and torch.equal(y_a, y_b)
    assert torch.allclose(rm1, rm2, atol=1e-6) and torch.allclose(rv1, rv2, atol=1e-6)
    dy = _act(N, C, H, W, dev)
    dg1, db1, dg2, db2 = (torch.zeros(C, device=dev) for _ in range(4))
    dx_a, dres_a, _ = nv.bn_bwd(dy, x, y_a, stat_a, relu, res, dg1, db1, 0.0)
    dx_b, dres_b = nv.bn_bwd_fused(dy, x, y_a, stat_a, relu, res, dg2, db2, 0.0)
    assert _rel_err(dx_b, dx_a) < 1e-3 and _rel_err(dg2, dg1) < 1e-4 and _rel_err(db2, db1) < 1e-4
    if res:
        assert torch.equal(dres_a, dres_b)


def test_bn_eval():
    from b200ddp.ops import native as nv
    dev = _dev()
    x = _act(4, 64, 8, 8, dev)
    g, b = torch.rand(64, device=dev) + 0.5, torch.randn(64, device=dev)
    rm, rv = torch.randn(64, device=dev), torch.rand(64, device=dev) + 0.5
    y = nv.bn_apply(x, None, nv.bn_eval_prep(g, b, rm, rv, 1e-5), False)
    assert _rel_err(y, F.batch_norm(x.float(), rm, rv, g, b, False, 0.1, 1e-5)) < 1e-2


@pytest.mark.parametrize("k,s,p,shape", [(3, 2, 1, (4, 64, 16, 16)), (2, 2, 0, (4, 8, 28, 28)), (2, 2, 0, (3, 16, 10, 10))])
def test_maxpool(k, s, p, shape):
    from b200ddp.ops import native as nv
    dev = _dev()
    x = _act(*shape, dev)
    xr = x.float().requires_grad_(True)
    yr = F.max_pool2d(xr, k, s, p)
    y, am = nv.maxpool_fwd(x, k, s, p)
    assert torch.equal(y.float(), yr)
    dy = _act(*yr.shape, dev)
    yr.backward(dy.float())
    dx = nv.maxpool_bwd(dy, am, x.shape, k, s, p)
    assert _rel_err(dx, xr.grad) < 1e-2


def test_avgpool_relu():
    from b200ddp.ops import native as nv
    dev = _dev()
    x = _act(8, 512, 2, 2, dev)
    assert _rel_err(nv.avgpool_fwd(x), x.float().mean((2, 3))) < 1e-2
    dy = torch.randn(8, 512, device=dev).to(torch.bfloat16)
    dx = nv.avgpool_bwd(dy, x.shape)
    assert _rel_err(dx, (dy.float() / 4)[:, :, None, None].expand(8, 512, 2, 2)) < 1e-2
    g = nv.relu_bwd(x, x)
    assert torch.equal(g.float(), (x.float() * (x.float() > 0)))


@pytest.mark.parametrize("B,C,dt", [(32, 1000, torch.bfloat16), (256, 10, torch.float32), (1000, 10, torch.bfloat16), (5, 16, torch.float32)])
def test_softmax_xent(B, C, dt):
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(3)
    lg = (torch.randn(B, C, device=dev) * 3).to(dt)
    t = torch.randint(0, min(C, 10), (B,), device=dev)
    lr = lg.float().requires_grad_(True)
    loss_r = F.cross_entropy(lr, t)
    loss_r.backward()
    for _ in range(2):       # twice: the ticket counter resets itself
        loss, dl = nv.softmax_xent(lg, t, 1.0 / B)
        assert abs(float(loss) - float(loss_r)) < 2e-3 * max(1.0, abs(float(loss_r)))
        assert _rel_err(dl, lr.grad) < 1e-2


def test_bce_and_eval_count():
    from b200ddp.ops import native as nv
    dev = _dev()
    x = torch.randn(64, device=dev)
    t = (torch.rand(64, device=dev) > 0.5).float()
    xr = x.clone().requires_grad_(True)
    lr = F.binary_cross_entropy_with_logits(xr, t)
    lr.backward()
    loss, dl = nv.bce_logits(x, t, 1.0 / 64)
    assert abs(float(loss) - float(lr)) < 1e-4 and _rel_err(dl, xr.grad) < 1e-4
    lg = torch.randn(1000, 10, device=dev)
    tg = torch.randint(0, 10, (1000,), device=dev)
    correct = torch.zeros((), dtype=torch.int32, device=dev)
    nll = torch.zeros((), device=dev)
    nv.eval_count(lg, tg, correct, nll)
    assert int(correct) == int((lg.argmax(1) == tg).sum())
    assert abs(float(nll) - float(F.nll_loss(lg, tg, reduction="sum"))) < 1e-2


def test_sgd_adam_flat():
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(4)
    n = 4096 * 3
    p = torch.randn(n, device=dev)
    pr = p.clone().requires_grad_(True)
    m = torch.zeros(n, device=dev)
    sh = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    opt = torch.optim.SGD([pr], lr=0.01, momentum=0.9, weight_decay=1e-4)
    for _ in range(3):
        g = torch.randn(n, device=dev)
        pr.grad = g.clone()
        opt.step()
        gg = g.clone()
        nv.sgd_step(p, gg, m, sh, 0.01, 0.9, 0.0, 1e-4, False, 1.0, True)
        assert float(gg.abs().max()) == 0.0
    assert torch.allclose(p, pr.detach(), atol=1e-5, rtol=1e-5)
    assert torch.equal(sh, p.to(torch.bfloat16))
    p = torch.randn(n, device=dev)
    pr = p.clone().requires_grad_(True)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    opt = torch.optim.Adam([pr], lr=1e-3)
    for _ in range(3):
        g = torch.randn(n, device=dev)
        pr.grad = g.clone()
        opt.step()
        nv.adam_step(p, g.clone(), m, v, None, 1e-3, 0.9, 0.999, 1e-8, 0.0, step, 1.0, False)
    assert torch.allclose(p, pr.detach(), atol=1e-5, rtol=1e-4)


def test_augment_and_trojan():
    from b200ddp.ops import native as nv
    from b200ddp.utils.data import CIFAR10_MEAN, CIFAR10_STD, augment_reference
    dev = _dev()
    torch.manual_seed(5)
    img = torch.randint(0, 256, (16, 32, 32, 3), dtype=torch.uint8)
    off = torch.stack([torch.randint(0, 9, (16,)), torch.randint(0, 9, (16,)), torch.randint(0, 2, (16,))], 1).to(torch.int32)
    ref = augment_reference(img, offsets=off)
    mean = torch.tensor(CIFAR10_MEAN, device=dev)
    inv = 1.0 / torch.tensor(CIFAR10_STD, device=dev)
    out = nv.augment(img.to(dev), off.to(dev), mean, inv, c_out=8)
    assert out.shape == (16, 8, 32, 32) and float(out[:, 3:].abs().max()) == 0.0
    assert torch.allclose(out[:, :3].float().cpu(), ref, atol=2e-2, rtol=1e-2)
    out32 = nv.augment(img.to(dev), off.to(dev), mean, inv, nchw_fp32=True)
    assert torch.allclose(out32.cpu(), ref, atol=1e-5)
    x = torch.rand(6, 3, 32, 32, device=dev)
    pat = torch.rand(3, 4, 4, device=dev)
    mask = torch.tensor([1, 0, 1, 1, 0, 1], dtype=torch.uint8, device=dev)
    xr = x.clone()
    for i in range(6):
        if mask[i]:
            xr[i, :, 5:9, 7:11] = 0.3 * pat + 0.7 * xr[i, :, 5:9, 7:11]
    nv.trojan_stamp_(x, pat, 5, 7, 0.3, mask)
    assert torch.allclose(x, xr, atol=1e-6)


# ------------------------------------------------------------------ sequence models
def test_sgemm_strided_variants_match_torch():
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(0)
    for (M, N, K) in [(70, 130, 45), (3200, 400, 40), (400, 100, 3200), (1, 400, 3200), (5, 1, 300)]:
        A = torch.randn(M, K, device=dev)
        B = torch.randn(K, N, device=dev)
        bias = torch.randn(N, device=dev)
        ref = (A.double() @ B.double()).float()
        scale = ref.abs().max()
        assert (nv.sgemm(A, B) - ref).abs().max() < 2e-5 * scale * (K ** 0.5)
        assert (nv.sgemm(A.t().contiguous().t(), B.t().contiguous().t(), bias=bias) - (ref + bias)).abs().max() < 2e-5 * scale * (K ** 0.5)
        assert (nv.sgemm(B.t(), A.t()) - ref.t()).abs().max() < 2e-5 * scale * (K ** 0.5)
    # overlapping rows (the text-CNN im2col view) and accumulate
    E = torch.randn(50, 8, device=dev)
    A = E.as_strided((47, 24), (8, 1))
    W = torch.randn(24, 10, device=dev)
    out = torch.ones(47, 10, device=dev)
    nv.sgemm(A, W, out=out, beta=1.0)
    ref = torch.stack([E[t:t + 3].reshape(-1) for t in range(47)]) @ W + 1.0
    assert torch.allclose(out, ref, atol=1e-4)


def test_mel_frontend_matches_torch_stft_features():
    from b200ddp.models.mntd import AudioModel
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(1)
    m = AudioModel()
    x = torch.randn(5, 16000) * 0.1
    x[0, :4000] = 0.0                          # silence: exercises the 1e-10 clamp
    ref = m.features(x)                        # torch.stft on the CPU
    out = nv.mel_frontend(x.to(dev), m._window.to(dev), m._mel.to(dev)).cpu()
    assert out.shape == ref.shape == (5, 32, 40)
    loud = ref > -0.5                          # features above -75 dB: compare tightly; below, fp32 FFT noise dominates both
    assert (out - ref)[loud].abs().max() < 2e-3
    assert (out - ref).abs().max() < 0.2


def test_lstm_layers_match_torch_cpu():
    from b200ddp.ops import functional as BF
    dev = _dev()
    torch.manual_seed(2)
    ref = torch.nn.LSTM(input_size=40, hidden_size=100, num_layers=2, batch_first=True)
    gpu = copy.deepcopy(ref).to(dev)
    x = torch.randn(7, 32, 40)
    xr = x.clone().requires_grad_(True)
    out_r, _ = ref(xr)
    wgt = torch.randn_like(out_r)
    (out_r * wgt).sum().backward()
    xg = x.to(dev).requires_grad_(True)
    out_g = BF.lstm(xg, gpu)
    (out_g * wgt.to(dev)).sum().backward()
    torch.cuda.synchronize()
    assert _rel(out_g, out_r) < 1e-4
    assert _rel(xg.grad, xr.grad) < 1e-3
    for (n, a), (_, b) in zip(ref.named_parameters(), gpu.named_parameters()):
        assert _rel(b.grad, a.grad) < 1e-3, (n, _rel(b.grad, a.grad))


def test_audio_and_text_models_native_paths_match_cpu():
    from b200ddp.models.mntd import AudioModel, RTNLPModel
    dev = _dev()
    torch.manual_seed(3)
    ref = AudioModel().eval()
    gpu = copy.deepcopy(ref).to(dev)
    x = torch.randn(4, 16000) * 0.1
    y = torch.tensor([1, 5, 2, 7])
    lr = ref.loss(ref(x), y)
    lr.backward()
    lg = gpu.loss(gpu(x.to(dev)), y.to(dev))
    lg.backward()
    assert abs(float(lr.detach()) - float(lg.detach())) < 2e-3
    for (n, a), (_, b) in zip(ref.named_parameters(), gpu.named_parameters()):
        if n == "lstm_att.bias":               # softmax is shift-invariant: this gradient is exactly zero up to rounding
            assert float(b.grad.abs().max()) < 1e-6
            continue
        assert _rel(b.grad, a.grad) < 2e-2, (n, _rel(b.grad, a.grad))
    # text CNN: token ids (gather kernel) and continuous embeddings with a gradient (query tuning)
    ref = RTNLPModel(emb_path="").eval()
    gpu = copy.deepcopy(ref).to(dev)
    gpu.embed_static.embed.to(dev)
    ids = torch.randint(1, 18000, (6, 12))
    lab = torch.tensor([0, 1, 1, 0, 1, 0])
    lr = ref.loss(ref(ids), lab)
    lr.backward()
    lg = gpu.loss(gpu(ids.to(dev)), lab.to(dev))
    lg.backward()
    assert abs(float(lr.detach()) - float(lg.detach())) < 1e-4
    for (n, a), (_, b) in zip(ref.named_parameters(), gpu.named_parameters()):
        assert _rel(b.grad, a.grad) < 1e-3, (n, _rel(b.grad, a.grad))
    e = torch.randn(3, 1, 10, 300)
    er = e.clone().requires_grad_(True)
    eg = e.to(dev).requires_grad_(True)
    ref.emb_forward(er).sum().backward()
    gpu.emb_forward(eg).sum().backward()
    assert _rel(eg.grad, er.grad) < 1e-3


def test_conv_epilogue_bn_stats_and_finalize_match_bn_stats_kernel():
    """conv_fwd(bn_sums=) + bn_finalize == conv_fwd + bn_stats (the separate statistics pass), for a direct tile and a
    cluster split-K shape; the sums are left zero."""
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(4)
    for (N, C, H, K, st) in [(64, 64, 8, 64, 1), (256, 128, 4, 128, 1), (64, 64, 8, 128, 2), (32, 16, 16, 24, 1)]:
        x = _act(N, C, H, H, dev)
        w = (torch.randn(K, 3, 3, C, device=dev) / (3 * C ** 0.5)).to(torch.bfloat16)
        gam, bet = torch.rand(K, device=dev) + 0.5, torch.randn(K, device=dev)
        rm1, rv1 = torch.zeros(K, device=dev), torch.ones(K, device=dev)
        rm2, rv2 = torch.zeros(K, device=dev), torch.ones(K, device=dev)
        y_ref = nv.conv_fwd(x, w, None, st, 1, False)
        stat_ref = nv.bn_stats(y_ref, gam, bet, 1e-5, 0.1, rm1, rv1)
        sums = torch.zeros(2 * K, device=dev)
        y = nv.conv_fwd(x, w, None, st, 1, False, sums)
        rows = y.numel() // K
        stat = nv.bn_finalize(sums, rows, gam, bet, 1e-5, 0.1, rm2, rv2)
        torch.cuda.synchronize()
        assert torch.equal(y, y_ref)
        assert float(sums.abs().max()) == 0.0
        # the separate pass reads the bf16-rounded tensor, the epilogue the fp32 accumulators: equal up to that rounding
        assert torch.allclose(stat[0], stat_ref[0], atol=3e-3), (N, C, H, K)
        assert torch.allclose(stat[1], stat_ref[1], rtol=5e-3)
        assert torch.allclose(stat[2], stat_ref[2], rtol=5e-3) and torch.allclose(stat[3], stat_ref[3], atol=2e-2)
        assert torch.allclose(rm2, rm1, atol=1e-3) and torch.allclose(rv2, rv1, rtol=5e-3)
        # one-launch variant: finalize folded into bn_apply (ticket word after the sums)
        sums3 = torch.zeros(2 * K + 1, device=dev)
        rm3, rv3 = torch.zeros(K, device=dev), torch.ones(K, device=dev)
        res = torch.randn_like(y)
        y3 = nv.conv_fwd(x, w, None, st, 1, False, sums3)
        out3, stat3 = nv.bn_apply_sums(y3, res, sums3, gam, bet, 1e-5, 0.1, rm3, rv3, True)
        out_ref = nv.bn_apply(y_ref, res, stat, True)
        torch.cuda.synchronize()
        assert float(sums3.abs().max()) == 0.0
        assert torch.allclose(stat3, stat, rtol=1e-4, atol=1e-5) and torch.allclose(rm3, rm2, atol=1e-6) and torch.allclose(rv3, rv2, rtol=1e-5)
        assert torch.allclose(out3.float(), out_ref.float(), atol=2e-2, rtol=2e-2)


def test_conv_dgrad_residual_epilogue_and_grad_mailbox():
    """b200_conv_dgrad_add == dgrad + residual; a forked activation's two gradients meet in that epilogue (no autograd add)
    and equal what plain autograd accumulation gives."""
    from b200ddp import nn as bnn
    from b200ddp.ops import functional as BF, native as nv
    dev = _dev()
    torch.manual_seed(6)
    for (N, C, H, K, R, st, pad) in [(8, 64, 8, 64, 3, 1, 1), (8, 64, 8, 128, 1, 2, 0), (4, 128, 4, 128, 3, 2, 1)]:
        Ho = (H + 2 * pad - R) // st + 1
        dy = _act(N, K, Ho, Ho, dev)
        w = (torch.randn(K, R, R, C, device=dev) / (R * C ** 0.5)).to(torch.bfloat16)
        res = _act(N, C, H, H, dev)
        plain = nv.conv_dgrad(dy, w, (N, C, H, H), st, pad)
        fused = nv.conv_dgrad(dy, w, (N, C, H, H), st, pad, residual=res)
        ref = (plain.float() + res.float()).to(torch.bfloat16)
        assert (fused.float() - ref.float()).abs().max() <= 2e-2 * max(1.0, float(ref.float().abs().max()))
    # block-level: y = relu(bn(conv(x)) + x) with the mailbox on and off
    grads = {}
    for on in (True, False):
        BF._FUSE_SKIP_GRAD[0] = on
        torch.manual_seed(7)
        conv = bnn.Conv2d(64, 64, 3, padding=1, bias=False).to(dev)
        bn = bnn.BatchNorm2d(64, relu=True).to(dev)
        pre = bnn.Conv2d(64, 64, 1, bias=False).to(dev)
        x0 = _act(16, 64, 8, 8, dev).requires_grad_(True)
        x = BF.fork(pre(x0))
        out = bn(conv(x), x)
        out.float().square().mean().backward()
        grads[on] = (x0.grad.clone(), pre.weight.grad.clone(), conv.weight.grad.clone())
    BF._FUSE_SKIP_GRAD[0] = False
    for a, b in zip(grads[True], grads[False]):
        assert _rel(a, b) < 1e-2


@pytest.mark.parametrize("shape,relu,res", [((256, 64, 8, 8), True, True), ((256, 128, 4, 4), True, False), ((256, 256, 2, 2), False, True),
                                             ((32, 64, 8, 8), True, True), ((37, 72, 5, 3), True, True), ((250, 64, 8, 8), True, False)])
def test_batchnorm_backward_cluster_kernel_matches_two_pass(shape, relu, res):
    """Mid-size tensors: the one-launch cluster kernel (registers + DSMEM reduction) == the reduce/apply pair, also for row
    counts that do not divide evenly over the cluster and for an accumulating dgamma/dbeta."""
    from b200ddp.ops import native as nv
    dev = _dev()
    torch.manual_seed(8)
    N, C, H, W = shape
    x = _act(N, C, H, W, dev, 2.0)
    r = _act(N, C, H, W, dev) if res else None
    gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
    stat = nv.bn_stats(x, gamma, beta, 1e-5, 0.1, None, None)
    y = nv.bn_apply(x, r, stat, relu)
    dy = _act(N, C, H, W, dev)
    dg1, db1 = torch.ones(C, device=dev), torch.ones(C, device=dev)
    dg2, db2 = torch.ones(C, device=dev), torch.ones(C, device=dev)
    dx_a, dres_a, _ = nv.bn_bwd(dy, x, y, stat, relu, res, dg1, db1, 1.0)
    dx_b, dres_b = nv.bn_bwd_fused(dy, x, y, stat, relu, res, dg2, db2, 1.0)
    torch.cuda.synchronize()
    assert _rel_err(dx_b, dx_a) < 2e-3 and _rel_err(dg2, dg1) < 1e-4 and _rel_err(db2, db1) < 1e-4
    if res:
        assert torch.equal(dres_a, dres_b)


def test_mel_frontend_backward_matches_torch_autograd():
    """d(features)/d(waveform) of the FFT kernel == autograd through torch.stft on the CPU (query tuning of the audio model)."""
    from b200ddp.models.mntd import AudioModel
    from b200ddp.ops import functional as BF
    dev = _dev()
    torch.manual_seed(9)
    m = AudioModel()
    x = torch.randn(3, 16000) * 0.1
    wgt = torch.randn(3, 32, 40)
    xr = x.clone().requires_grad_(True)
    (m.features(xr) * wgt).sum().backward()
    xg = x.to(dev).requires_grad_(True)
    out = BF.mel_features(xg, m._window.to(dev), m._mel.to(dev))
    (out * wgt.to(dev)).sum().backward()
    torch.cuda.synchronize()
    assert _rel(xg.grad, xr.grad) < 2e-3, _rel(xg.grad, xr.grad)
    # and through the whole audio model: the gradient reaches the waveform on the native path
    gpu = copy.deepcopy(m).to(dev)
    xq = (torch.randn(2, 16000, device=dev) * 0.1).requires_grad_(True)
    gpu(xq).sum().backward()
    assert xq.grad is not None and float(xq.grad.abs().max()) > 0
