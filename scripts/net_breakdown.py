#!/usr/bin/env python
"""Workshop Net megakernel: forward only vs forward+backward vs the full captured step (1 GPU, graph replays)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd
from b200ddp.models import get_model
from b200ddp.ops import native as nv, optim as boptim

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
bd.init_process_group(backend="b200", rank=0, world_size=1)
dev = bd.get_device()
torch.manual_seed(0)
net = get_model("custom").to(dev)
params = [p.detach() for p in net.parameters()]
grads = [torch.zeros_like(p) for p in params]
x = torch.randn(B, 3, 32, 32, device=dev)
y = torch.randint(0, 10, (B,), device=dev)
xu = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, device=dev)
mean = torch.tensor((0.4914, 0.4822, 0.4465), device=dev); inv = 1.0 / torch.tensor((0.2023, 0.1994, 0.2010), device=dev)

def op_time(fn, reps=20):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5): g.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / (5 * reps) * 1e3

res = {"batch": B}
res["fwd_only_us"] = op_time(lambda: nv.net_fwd_bwd(x, y, params, None, False))
res["fwd_bwd_us"] = op_time(lambda: nv.net_fwd_bwd(x, y, params, grads, False))
res["augment_us"] = op_time(lambda: nv.augment(xu, None, mean, inv, nchw_fp32=True))
print(json.dumps(res, indent=1))

# ---- phase timeline of CTA 0 (clock64 stamps; first image) -- warm, plain launches
import ctypes
from b200ddp.ops import _lib as L
prof = torch.zeros(32, dtype=torch.int64, device=dev)
L.lib().b200_net_set_profile_buffer(ctypes.c_void_p(prof.data_ptr()))
for _ in range(3):
    nv.net_fwd_bwd(x, y, params, grads, False)
torch.cuda.synchronize()
L.lib().b200_net_set_profile_buffer(ctypes.c_void_p(0))
st = prof.cpu().tolist()
names = ["setup (weights->smem)", "load image", "conv1+pool", "conv2+pool", "fc1", "fc2+fc3+loss", "dh2+dh1", "dp2 stream",
         "dp2 sum+g2", "conv2 dgrad || conv2 wgrad+bias+scratch", "g1", "conv1 wgrad", None, "flush (red.add)", "ticket+loss"]
order = list(range(16))
tl = {}
for k in range(len(order) - 1):
    a, b = st[order[k]], st[order[k + 1]]
    if names[k] is None:
        tl["(remaining images of CTA 0)"] = b - a
    else:
        tl[names[k]] = b - a
tl["total_cycles"] = st[15] - st[0]
print(json.dumps({"phase_cycles_cta0": tl}, indent=1))
