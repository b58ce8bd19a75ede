import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ddp.ops import native as nv
dev = torch.device("cuda", 0)
torch.manual_seed(0)
CL = torch.channels_last
def act(n, c, h, w): return torch.randn(n, c, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
cases = [(32, 64, 8, 8, 128, 3, 2, 1), (32, 128, 4, 4, 128, 3, 1, 1), (32, 256, 2, 2, 256, 3, 1, 1), (32, 512, 1, 1, 512, 3, 1, 1), (256, 128, 4, 4, 128, 3, 1, 1)]
for (N, C, H, W, K, R, st, pad) in cases:
    x = act(N, C, H, W); w = (torch.randn(K, R, R, C, device=dev) / (C * R * R) ** 0.5).to(torch.bfloat16)
    ref = None; worst = 0.0; worst_d = 0.0
    ref_nosplit = None
    nv.SPLIT_K[0] = False
    y0 = nv.conv_fwd(x, w, None, st, pad, False).float()
    nv.SPLIT_K[0] = True
    dy = act(*y0.shape)
    nv.SPLIT_K[0] = False
    d0 = nv.conv_dgrad(dy, w, x.shape, st, pad).float()
    nv.SPLIT_K[0] = True
    for i in range(30):
        y = nv.conv_fwd(x, w, None, st, pad, False).float()
        d = nv.conv_dgrad(dy, w, x.shape, st, pad).float()
        worst = max(worst, float((y - y0).abs().max() / y0.abs().max()))
        worst_d = max(worst_d, float((d - d0).abs().max() / d0.abs().max()))
    print((N, C, H, W, K, R, st, pad), "fwd split-vs-nosplit max rel %.3e   dgrad %.3e" % (worst, worst_d), flush=True)
# back-to-back different shapes sharing the scratch
xs = [act(32, 128, 4, 4), act(32, 256, 2, 2)]
ws = [(torch.randn(128, 3, 3, 128, device=dev) / 34).to(torch.bfloat16), (torch.randn(256, 3, 3, 256, device=dev) / 48).to(torch.bfloat16)]
nv.SPLIT_K[0] = False
refs = [nv.conv_fwd(xs[i], ws[i], None, 1, 1, False).float() for i in range(2)]
nv.SPLIT_K[0] = True
worst = 0.0
for it in range(50):
    for i in range(2):
        y = nv.conv_fwd(xs[i], ws[i], None, 1, 1, False).float()
        worst = max(worst, float((y - refs[i]).abs().max() / refs[i].abs().max()))
print("interleaved shapes: max rel %.3e" % worst)
