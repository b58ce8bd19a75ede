"""Where a conv kernel's microseconds go: clock64 stamps of CTA (0,0,0) at its phase boundaries (`b200_conv_set_profile_buffer`),
one warm launch per shape, next to the launch's CUDA-event time.  usage: conv_phase_profile.py [fp32|bf16] [batch]"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ddp.ops import native as nv, _lib as L  # noqa: E402

dtype = torch.float32 if (len(sys.argv) > 1 and sys.argv[1] == "fp32") else torch.bfloat16
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda", 0)
CL = torch.channels_last
buf = torch.zeros(16, dtype=torch.int64, device=dev)
lib = L.lib()
GHZ = 1.965e3        # cycles per us at the B200's 1965 MHz


def act(n, c, h, w):
    return torch.randn(n, c, h, w, device=dev).to(dtype).contiguous(memory_format=CL)


def event_us(fn, iters=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


print("dtype %s batch %d" % (dtype, B))
print("| shape | mode | launch (events, back to back) | in-kernel total | set-up | loads issued + landed | MMAs retired | TMEM -> regs/smem | cluster sync | peers' partials read | BN sums | store | exit |")
print("|---|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
shapes = {"layer1 64->64 @8x8": (64, 8, 64), "layer2 128->128 @4x4": (128, 4, 128), "layer3 256->256 @2x2": (256, 2, 256),
          "layer4 512->512 @1x1": (512, 1, 512)}
for name, (C, H, K) in shapes.items():
    x = act(B, C, H, H)
    w = (torch.randn(K, 3, 3, C, device=dev) / 30).to(dtype)
    sums = torch.zeros(2 * K + 1, device=dev)
    y = nv.conv_fwd(x, w, None, 1, 1, False)
    dy = torch.randn_like(y)
    dw = torch.zeros(K, 3, 3, C, device=dev)
    modes = (("fwd", lambda: nv.conv_fwd(x, w, None, 1, 1, False)),
             ("fwd+stats", lambda: nv.conv_fwd(x, w, None, 1, 1, False, sums)),
             ("dgrad", lambda: nv.conv_dgrad(dy, w, x.shape, 1, 1)),
             ("wgrad", lambda: nv.conv_wgrad_(dw, x, dy, 1, 1)))
    for mode, fn in modes:
        for _ in range(5):
            fn()
        ev = event_us(fn)
        buf.zero_()
        lib.b200_conv_set_profile_buffer(ctypes.c_void_p(buf.data_ptr()))
        fn()
        torch.cuda.synchronize()
        lib.b200_conv_set_profile_buffer(ctypes.c_void_p(0))
        t = buf.cpu().tolist()

        def d(a, b):
            return "%.1f" % ((t[b] - t[a]) / GHZ) if t[a] and t[b] else "-"
        cols = [d(0, 1), d(1, 2), d(2, 3), d(3, 4), d(4, 5), d(5, 8), d(8, 9), d(9, 6), d(6, 7) if t[6] else d(4, 7)]
        print("| %s | %s | %.1f | %.1f | %s |" % (name, mode, ev, (t[7] - t[0]) / GHZ, " | ".join(cols)))
