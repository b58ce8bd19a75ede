import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b200ddp.ops import native as nv, _lib as L
dev = torch.device("cuda", 0); CL = torch.channels_last
buf = torch.zeros(8, dtype=torch.int64, device=dev)
lib = L.lib()
def act(n, c, h, w): return torch.randn(n, c, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
names = ["setup(bar init+tmem alloc+sync)", "loads issued+landed", "MMA done (tmem_full)", "epilogue phase 1", "cluster sync 1", "DSMEM reduce+store", "cluster sync 2 + dealloc"]
B = 256
for name, (C, H, K, st) in {"layer1 64->64 @8": (64, 8, 64, 1), "layer2 128->128 @4": (128, 4, 128, 1), "layer3 256->256 @2": (256, 2, 256, 1), "layer4 512->512 @1": (512, 1, 512, 1)}.items():
    x = act(B, C, H, H); w = (torch.randn(K, 3, 3, C, device=dev) / 30).to(torch.bfloat16)
    y = nv.conv_fwd(x, w, None, st, 1, False); dy = torch.randn_like(y); dw = torch.zeros(K, 3, 3, C, device=dev)
    for mode, fn in (("fwd", lambda: nv.conv_fwd(x, w, None, st, 1, False)), ("dgrad", lambda: nv.conv_dgrad(dy, w, x.shape, st, 1)), ("wgrad", lambda: nv.conv_wgrad_(dw, x, dy, st, 1))):
        for _ in range(5): fn()
        lib.b200_conv_set_profile_buffer(ctypes.c_void_p(buf.data_ptr()))
        fn(); torch.cuda.synchronize()
        lib.b200_conv_set_profile_buffer(ctypes.c_void_p(0))
        t = buf.cpu().tolist()
        d = [(t[i + 1] - t[i]) / 1.9e3 if t[i + 1] and t[i] else 0 for i in range(7)]
        print("%-20s %-6s total %5.1f us | " % (name, mode, (t[7] - t[0]) / 1.9e3) + "  ".join("%s %.1f" % (names[i].split("(")[0].strip()[:18], d[i]) for i in range(7)))
        buf.zero_()
