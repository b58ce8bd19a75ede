#!/usr/bin/env python
"""Run-to-run reproducibility of one resnet18 forward+backward on one GPU (no communication)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd
from b200ddp.models import get_model
from b200ddp.nn import CrossEntropyLoss
from b200ddp.ops import native as nv

bd.init_process_group(backend="b200", rank=0, world_size=1)
dev = bd.get_device()
torch.manual_seed(0)
model = bd.DistributedDataParallel(get_model(sys.argv[1] if len(sys.argv) > 1 else "resnet18").to(dev))
crit = CrossEntropyLoss()
g = torch.Generator().manual_seed(7)
x = torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
y = torch.randint(0, 10, (32,), generator=g).to(dev)
mean = torch.tensor((0.4914, 0.4822, 0.4465), device=dev); inv = 1.0 / torch.tensor((0.2023, 0.1994, 0.2010), device=dev)
model.train()
# usage: determinism_check.py [model] [deterministic]  -- the second argument turns split-K and the conv-epilogue BN statistics
# (both meet in fp32 red.adds whose order varies) off, which makes forward and dgrad reproducible bit for bit
if len(sys.argv) > 2 and sys.argv[2].startswith("det"):
    from b200ddp.nn import modules as bmod
    nv.SPLIT_K[0] = False
    bmod._FUSE_BN_STATS[0] = False
runs, outs = [], []
for i in range(4):
    model.zero_grad()
    out = model(nv.augment(x, None, mean, inv, c_out=8))
    loss = crit(out, y)
    loss.backward()
    torch.cuda.synchronize()
    runs.append(model.flat_grads.clone()); outs.append(out.detach().float().clone())
for i in range(1, 4):
    d = float((runs[i] - runs[0]).norm() / runs[0].norm()); o = float((outs[i] - outs[0]).norm() / outs[0].norm())
    per = ["%.1e" % float((runs[i][b.start:b.start+b.numel] - runs[0][b.start:b.start+b.numel]).norm() / runs[0][b.start:b.start+b.numel].norm()) for b in model.buckets]
    print("run %d vs 0: logits rel %.2e, grads rel %.2e, per bucket %s" % (i, o, d, per), flush=True)
bd.destroy_process_group()
