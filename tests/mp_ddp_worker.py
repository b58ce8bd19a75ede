"""Worker for tests/test_cpu_ddp.py: DDP semantics on gloo, world_size 2 (run once per rank)."""
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd  # noqa: E402
from b200ddp.models import get_model  # noqa: E402


class BNNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(3, 8, 3, padding=1)
        self.bn = nn.BatchNorm2d(8)
        self.fc = nn.Linear(8, 4)

    def forward(self, x):
        return self.fc(F.relu(self.bn(self.conv(x))).mean((2, 3)))


def main():
    out_path = sys.argv[1]
    bd.init_process_group(backend="gloo")
    rank, world = bd.get_rank(), bd.get_world_size()
    res = {}
    torch.manual_seed(100 + rank)                      # different init per rank
    net = get_model("custom")
    ddp = bd.DistributedDataParallel(net)
    ps = bd.all_gather_object(ddp.flat_params.clone())
    res["param_broadcast"] = all(torch.equal(ps[0], p) for p in ps)
    res["module_prefix"] = all(k.startswith("module.") for k in ddp.state_dict())

    # gradient averaging == gradient of the single-process large batch
    torch.manual_seed(7)
    X = torch.randn(8 * world, 3, 32, 32)
    Y = torch.randint(0, 10, (8 * world,))
    ref = get_model("custom")
    ref.load_state_dict(net.state_dict())
    F.cross_entropy(ref(X), Y).backward()
    xs, ys = X[rank::world], Y[rank::world]
    ddp.zero_grad()
    F.cross_entropy(ddp(xs), ys).backward()
    err = max(float((p.grad - q.grad).abs().max()) for p, q in zip(net.parameters(), ref.parameters()))
    res["grad_equals_large_batch"] = err < 1e-5
    gs = bd.all_gather_object(ddp.flat_grads.clone())
    res["grads_identical_across_ranks"] = all(torch.equal(gs[0], g) for g in gs)
    res["grad_is_arena_view"] = all(p.grad.data_ptr() >= ddp.flat_grads.data_ptr() for p in net.parameters())

    # no_sync accumulates locally, the next synced backward reduces the sum
    ddp.zero_grad()
    with ddp.no_sync():
        F.cross_entropy(ddp(xs), ys).backward()
    local = ddp.flat_grads.clone()
    ls = bd.all_gather_object(local)
    res["no_sync_keeps_grads_local"] = (world == 1) or not torch.equal(ls[0], ls[1])
    F.cross_entropy(ddp(xs), ys).backward()
    want = torch.stack([2 * l for l in ls]).mean(0)
    res["no_sync_then_sync"] = float((ddp.flat_grads - want).abs().max()) < 1e-5

    # torch optimizer with zero_grad(set_to_none=True) drops our views: the hook re-points them
    opt = torch.optim.SGD(ddp.parameters(), lr=0.1)
    opt.zero_grad(set_to_none=True)
    F.cross_entropy(ddp(xs), ys).backward()
    res["set_to_none_recovers"] = all(p.grad is not None and p.grad.data_ptr() >= ddp.flat_grads.data_ptr() for p in net.parameters())
    opt.step()
    ps = bd.all_gather_object(ddp.flat_params.clone())
    res["params_identical_after_step"] = all(torch.equal(ps[0], p) for p in ps)

    # uneven last batch: ranks see different batch sizes, gradients are still the plain mean of the ranks
    ddp.zero_grad()
    n = 5 if rank == 0 else 3
    F.cross_entropy(ddp(X[:n]), Y[:n]).backward()
    gs = bd.all_gather_object(ddp.flat_grads.clone())
    res["uneven_batch_ok"] = all(torch.equal(gs[0], g) for g in gs)

    # buffers (BN running stats) follow rank 0 on every forward
    torch.manual_seed(200 + rank)
    bn = bd.DistributedDataParallel(BNNet())
    bn.train()
    bn(torch.randn(4, 3, 8, 8) * (rank + 1)).sum().backward()          # ranks now hold different running stats
    bn(torch.randn(4, 3, 8, 8))                                         # forward broadcasts rank 0's buffers first
    means = bd.all_gather_object(bn.module.bn.running_mean.clone())
    res["bn_buffers_broadcast"] = len(bn.state_dict()) > 0 and bn.flat_buffers is not None
    res["all_reduce_api"] = None
    t = torch.full((5,), float(rank + 1))
    bd.all_reduce(t)
    res["all_reduce_api"] = bool((t == sum(range(1, world + 1))).all())
    t = torch.full((3,), float(rank))
    bd.broadcast(t, src=world - 1)
    res["broadcast_api"] = bool((t == world - 1).all())
    bd.barrier()
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump(res, f)
    bd.destroy_process_group()


if __name__ == "__main__":
    main()
