#!/usr/bin/env python
"""Where the train step's time goes, measured warm inside CUDA graphs (1 GPU): forward only, forward+backward,
full step; plus per-layer-type kernel microbenchmarks at the resnet18/32x32 shapes (each op replayed 50x in a graph)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd
from b200ddp.models import get_model
from b200ddp.nn import CrossEntropyLoss
from b200ddp.ops import native as nv, optim as boptim, functional as BF
from b200ddp.nn import modules as bmod

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
bd.init_process_group(backend="b200", rank=0, world_size=1)
dev = bd.get_device()
torch.manual_seed(0)
model = bd.DistributedDataParallel(get_model("resnet18").to(dev))
opt = boptim.SGD(model.parameters(), lr=0.01, momentum=0.9)
crit = CrossEntropyLoss()
BF.assume_unit_loss_grad(True); bmod.defer_bn_counters(True)
x = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, device=dev)
y = torch.randint(0, 10, (B,), device=dev)
mean = torch.tensor((0.4914, 0.4822, 0.4465), device=dev); inv = 1.0 / torch.tensor((0.2023, 0.1994, 0.2010), device=dev)

def fwd():
    return crit(model(nv.augment(x, None, mean, inv, c_out=8)), y)
def fwd_bwd():
    l = fwd(); l.backward(); return l.detach()
def full():
    l = fwd_bwd(); opt.step(); return l

def graph_time(fn, iters=20, no_grad=False):
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            if no_grad:
                with torch.no_grad(): fn()
            else: fn()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        if no_grad:
            with torch.no_grad(): fn()
        else: fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): g.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3

res = {"batch": B}
model.train()
res["fwd_only_us (no autograd)"] = graph_time(fwd, no_grad=True)
res["fwd+bwd_us"] = graph_time(fwd_bwd)
res["full_step_us"] = graph_time(full)

# ---- op microbenchmarks: 20 copies of one op in a graph
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

CL = torch.channels_last
def act(n, c, h, w): return torch.randn(n, c, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=CL)
ops = {}
for name, (C, H, K, st) in {"layer1 3x3 64->64 @8": (64, 8, 64, 1), "layer2 3x3 128->128 @4": (128, 4, 128, 1),
                             "layer3 3x3 256->256 @2": (256, 2, 256, 1), "layer4 3x3 512->512 @1": (512, 1, 512, 1),
                             "layer2.0 3x3 64->128 s2": (64, 8, 128, 2)}.items():
    xx = act(B, C, H, H); w = (torch.randn(K, 3, 3, C, device=dev) / 30).to(torch.bfloat16)
    yy = nv.conv_fwd(xx, w, None, st, 1, False); dy = torch.randn_like(yy); dw = torch.zeros(K, 3, 3, C, device=dev)
    gam, bet = torch.ones(K, device=dev), torch.zeros(K, device=dev)
    stat = nv.bn_stats(yy, gam, bet, 1e-5, 0.1, None, None); out = nv.bn_apply(yy, None, stat, True)
    ops[name] = {"conv_fwd": op_time(lambda: nv.conv_fwd(xx, w, None, st, 1, False)),
                 "conv_dgrad": op_time(lambda: nv.conv_dgrad(dy, w, xx.shape, st, 1)),
                 "conv_wgrad": op_time(lambda: nv.conv_wgrad_(dw, xx, dy, st, 1)),
                 "bn_stats": op_time(lambda: nv.bn_stats(yy, gam, bet, 1e-5, 0.1, None, None)),
                 "bn_apply": op_time(lambda: nv.bn_apply(yy, None, stat, True)),
                 "bn_bwd(reduce+apply)": op_time(lambda: nv.bn_bwd(dy, yy, out, stat, True, False))}
res["op_us"] = ops
p = torch.randn(11689536, device=dev); gg = torch.randn_like(p); m = torch.zeros_like(p); sh = torch.zeros(p.numel(), dtype=torch.bfloat16, device=dev)
res["sgd_flat_us"] = op_time(lambda: nv.sgd_step(p, gg, m, sh, 0.01, 0.9, 0.0, 0.0, False, 1.0, True), reps=5)
empty = torch.zeros(8, device=dev)
res["empty_kernel_us (fill 8 floats)"] = op_time(lambda: empty.zero_(), reps=50)
print(json.dumps(res, indent=1))
bd.destroy_process_group()
