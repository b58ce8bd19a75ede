#!/usr/bin/env python
"""Small single-GPU workload that touches every kernel family once; run under compute-sanitizer (scripts/sanitize.sh).
Sizes are tiny on purpose: the sanitizer slows kernels down 10-100x."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ddp.ops import native as nv  # noqa: E402
from b200ddp.ops import functional as BF  # noqa: E402
from b200ddp.models import get_model  # noqa: E402

dev = torch.device("cuda", 0)
CL = torch.channels_last
torch.manual_seed(0)


def act(n, c, h, w):
    return torch.randn(n, c, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)


# conv fwd (+BN sums) / dgrad (+residual) / wgrad: direct tile, cluster split-K, two-CTA variant
for (N, C, H, K, R, st, pad) in [(4, 64, 8, 64, 3, 1, 1), (32, 128, 4, 128, 3, 1, 1), (16, 64, 56, 64, 1, 1, 0)]:
    x = act(N, C, H, H)
    w = (torch.randn(K, R, R, C, device=dev) / 30).to(torch.bfloat16)
    sums = torch.zeros(2 * K + 1, device=dev)
    y = nv.conv_fwd(x, w, None, st, pad, False, sums)
    gam, bet = torch.ones(K, device=dev), torch.zeros(K, device=dev)
    stat = nv.bn_finalize(sums, y.numel() // K, gam, bet, 1e-5, 0.1, None, None)
    out = nv.bn_apply(y, None, stat, True)
    dy = torch.randn_like(y)
    dx = nv.conv_dgrad(dy, w, x.shape, st, pad, residual=x)
    dw = torch.zeros(K, R, R, C, device=dev)
    nv.conv_wgrad_(dw, x, dy, st, pad)
    dg, db = torch.zeros(K, device=dev), torch.zeros(K, device=dev)
    rows = y.numel() // K
    if rows <= 16384:
        nv.bn_bwd_fused(dy, y, out, stat, True, True, dg, db, 0.0)          # one-CTA or cluster kernel
    nv.bn_bwd(dy, y, out, stat, True, True, dg, db, 0.0)
# workshop Net megakernel (uint8 input, two images per some CTAs) + fc wgrad kernel
net = get_model("custom").to(dev)
xu = torch.randint(0, 256, (160, 32, 32, 3), dtype=torch.uint8, device=dev)
yl = torch.randint(0, 10, (160,), device=dev)
mean = torch.tensor((0.49, 0.48, 0.45), device=dev)
inv = 1.0 / torch.tensor((0.2, 0.2, 0.2), device=dev)
off = torch.randint(0, 9, (160, 3), device=dev, dtype=torch.int32)
off[:, 2] %= 2
BF.assume_unit_loss_grad(False)
net.fused_loss(xu, yl, aug=(off, mean, inv)).backward()
# sequence kernels
A, B = torch.randn(70, 45, device=dev), torch.randn(45, 130, device=dev)
nv.sgemm(A, B)
nv.sgemm(torch.randn(1200, 40, device=dev).t(), torch.randn(1200, 24, device=dev))     # split-K + atomics
lstm = torch.nn.LSTM(40, 100, 2, batch_first=True).to(dev)
xs = torch.randn(3, 6, 40, device=dev, requires_grad=True)
BF.lstm(xs, lstm).sum().backward()
from b200ddp.models.mntd import AudioModel, RTNLPModel  # noqa: E402
am = AudioModel().to(dev)
nv.mel_frontend(torch.randn(1, 16000, device=dev), am._window, am._mel)
tm = RTNLPModel(emb_path="").to(dev)
tm.embed_static.embed.to(dev)
tm(torch.randint(1, 18000, (3, 9), device=dev)).sum().backward()
# loss / optimizer / pools / augment
lg = torch.randn(16, 16, device=dev).to(torch.bfloat16)
nv.softmax_xent(lg, torch.randint(0, 10, (16,), device=dev), 1.0 / 16)
p, g, m = torch.randn(4096, device=dev), torch.randn(4096, device=dev), torch.zeros(4096, device=dev)
nv.sgd_step(p, g, m, None, 0.1, 0.9, 0.0, 0.0, False)
nv.augment(xu[:8], off[:8].contiguous(), mean, inv, c_out=8)
torch.cuda.synchronize()
print("SANITIZE_TARGET_DONE")
