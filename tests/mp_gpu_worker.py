"""Multi-GPU worker (one process per GPU, run under torch.distributed.run by tests/test_gpu_multi.py or by hand):

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mp_gpu_worker.py collectives|ddp|graphed|skip

Every check prints ``CHECK <name> ok|FAIL <detail>`` on rank 0; the last line is ``MP_GPU_RESULT <json>``.

  collectives  every all-reduce algorithm x dtype x odd sizes against torch.stack(per_rank).sum(0) * scale (the per-rank
               inputs travel through the gloo control plane, never through the kernels under test), cross-rank bitwise
               equality, broadcast, and 1000 back-to-back all-reduces of random sizes / algorithms / offsets with
               integer-valued payloads whose sum is exact (the race / stale-flag stress that stands in for the closed
               compute-sanitizer).
  ddp          DDP semantics on NVLink: wrap-time broadcast, synced grads == mean(local grads) and bitwise equal across
               ranks (workshop Net and resnet18, fp32 mode), no_sync keeps gradients local, BN buffers follow rank 0,
               parameters stay identical through captured training steps.
  graphed      the caller-owned loop on DistributedDataParallel(graph_forward_backward=True): collectives inside the two replayed
               graphs (synced grads == mean(local), parameters identical after training); SGD.sync_state() after steps with
               the optimizer inside the collectives.
  skip         fail-fast: rank 1 skips a collective -> the other ranks' device-side spin times out and
               ``check_health()`` raises (reference behaviour: the launcher aborts the job, nb2:284).
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ddp as bd  # noqa: E402
from b200ddp.ops import _lib as L  # noqa: E402

RESULTS = {}


def check(name, ok, detail=""):
    oks = [None] * dist.get_world_size()
    dist.all_gather_object(oks, bool(ok))
    RESULTS[name] = all(oks)
    if dist.get_rank() == 0:
        print("CHECK %-58s %s %s" % (name, "ok" if all(oks) else "FAIL", detail), flush=True)


def gather_cpu(t):
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, t.detach().cpu())
    return out


def collectives():
    comm = bd.get_comm()
    rank, world, dev = bd.get_rank(), bd.get_world_size(), bd.get_device()
    algos = [("oneshot", L.ALGO_ONESHOT), ("twoshot", L.ALGO_TWOSHOT)] + ([("nvls", L.ALGO_NVLS)] if comm.has_nvls else [])
    region, off = comm.alloc_arena(48 << 20, torch.uint8)                  # 48 MiB of symmetric memory to play in
    for dtype, tol in ((torch.float32, 2e-6), (torch.bfloat16, 1.6e-2)):
        es = 4 if dtype == torch.float32 else 2
        for name, algo in algos:
            for n in (4 // es * 4, 1000, 65536 + 8, (1 << 20) + 24, (6 << 20) + 40):
                n = (n * es + 15) // 16 * 16 // es                          # whole 16-byte vectors
                g = torch.Generator().manual_seed(1000 * rank + n % 9973)
                src = torch.randn(n, generator=g).to(dtype)
                view = region[: n * es].view(dtype)
                view.copy_(src.to(dev))
                torch.cuda.synchronize()
                dist.barrier()
                comm.all_reduce_arena(off, n, scale=1.0 / world, dtype=dtype, algo=algo)
                torch.cuda.synchronize()
                got = gather_cpu(view)
                srcs = gather_cpu(src)
                want = torch.stack([s.float() for s in srcs]).sum(0) / world
                err = float((got[0].float() - want).abs().max()) / (float(want.abs().max()) + 1e-30)
                same = all(torch.equal(got[0], t) for t in got[1:])
                check("allreduce %s %s n=%d" % (name, str(dtype).split(".")[-1], n), err <= tol and same,
                      "max rel err %.2e, bitwise equal across ranks %s" % (err, same))
    # broadcast from every root
    for root in range(world):
        n = 300000 + 16 * root
        view = region[: n * 4].view(torch.float32)
        view.fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        comm.broadcast_(view, src=root)
        torch.cuda.synchronize()
        check("broadcast root=%d" % root, bool((view == float(root + 1)).all()))
    # ---- stress: 1000 random (size, algorithm, offset) all-reduces back to back, no host sync in between
    g = torch.Generator().manual_seed(4242)                                 # same sequence on every rank
    pattern = (torch.arange(12 << 20, device=dev) % 13).float()             # integers: every partial sum is exact in fp32
    total = world * (world + 1) // 2
    bad = torch.zeros((), dtype=torch.int64, device=dev)
    fl = region.view(torch.float32)
    t0 = time.time()
    for it in range(1000):
        logn = float(torch.rand((), generator=g)) * 21.0 + 2.0              # 4 .. 8M elements, log-uniform
        n = max(4, int(2 ** logn) // 4 * 4)
        n = min(n, 8 << 20)
        o = int(torch.randint(0, ((12 << 20) - n) // 4 + 1, (), generator=g)) * 4
        algo = algos[int(torch.randint(0, len(algos), (), generator=g))][1]
        fl[o: o + n].copy_(pattern[o: o + n] * float(rank + 1))
        comm.all_reduce_arena(off + o * 4, n, scale=1.0, dtype=torch.float32, algo=algo)
        bad += (fl[o: o + n] != pattern[o: o + n] * float(total)).sum()
    torch.cuda.synchronize()
    comm.check_health()
    check("stress: 1000 random all-reduces, exact integer sums", int(bad.item()) == 0,
          "%d wrong elements, %.1f s" % (int(bad.item()), time.time() - t0))


def ddp():
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import native as nv
    from b200ddp.ops import optim as boptim
    from b200ddp.runtime.engine import TrainStep
    from b200ddp.utils.ddp_check import grad_sync_check, params_identical
    rank, world, dev = bd.get_rank(), bd.get_world_size(), bd.get_device()
    bd.set_compute_dtype("fp32")
    crit = CrossEntropyLoss()
    mean = torch.tensor((0.4914, 0.4822, 0.4465), device=dev)
    inv = 1.0 / torch.tensor((0.2023, 0.1994, 0.2010), device=dev)
    for name in ("custom", "resnet18"):
        torch.manual_seed(1234 + rank)                  # different init per rank: the wrap must broadcast rank 0's
        model = bd.DistributedDataParallel(get_model(name).to(dev))
        check("%s: wrap-time parameter broadcast" % name, params_identical(model))
        B = 32
        g = torch.Generator().manual_seed(77 + rank)
        x = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
        y = torch.randint(0, 10, (B,), generator=g).to(dev)

        def fwd_bwd():
            crit(model(nv.augment(x, None, mean, inv, c_out=8)), y).backward()

        model.train()
        res = grad_sync_check(model, fwd_bwd)
        check("%s: synced grads == mean(local grads)" % name, res["rel_err"] < 1e-5, "rel err %.2e" % res["rel_err"])
        check("%s: synced grads bitwise equal across ranks" % name, res["bitwise_equal"])
        # no_sync: gradients stay local (they differ across ranks), the next synced backward averages the accumulated sum
        model.zero_grad()
        with model.no_sync():
            fwd_bwd()
        torch.cuda.synchronize()
        loc = gather_cpu(model.flat_grads)
        check("%s: no_sync keeps gradients local" % name, not torch.equal(loc[0], loc[1]))
        fwd_bwd()
        torch.cuda.synchronize()
        acc = gather_cpu(model.flat_grads)
        check("%s: sync after no_sync is equal across ranks" % name, all(torch.equal(acc[0], t) for t in acc[1:]))
        model.zero_grad()
        if name == "resnet18":
            # BN buffers: rank > 0 scribbles over its running statistics; one training forward on the SAME input puts them
            # back in step with rank 0 (per-forward buffer broadcast, then the same update everywhere)
            if rank > 0:
                model.flat_buffers.fill_(7)
            torch.cuda.synchronize()
            dist.barrier()
            xs = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(5)).to(dev)
            with torch.no_grad():
                model(nv.augment(xs, None, mean, inv, c_out=8))
            torch.cuda.synchronize()
            bufs = gather_cpu(model.module.bn1.running_mean)
            check("resnet18: BN buffers follow rank 0", all(torch.allclose(bufs[0], b, rtol=1e-5, atol=1e-6) for b in bufs[1:]))
        opt = boptim.SGD(model.parameters(), lr=0.01, momentum=0.9)
        step = TrainStep(model, crit, opt, B, device=dev, graph=True, seed=rank)
        xs, ys = x.cpu().pin_memory(), y.cpu().pin_memory()
        losses = [step(xs, ys, need_loss=True) for _ in range(8)]
        torch.cuda.synchronize()
        check("%s: parameters identical after 8 captured steps" % name, params_identical(model))
        check("%s: losses finite" % name, all(v == v for v in losses), "%.3f -> %.3f" % (losses[0], losses[-1]))
        if rank == 0:
            print("allreduce algorithms used by %s:" % name, sorted(set(model.comm.launch_log)), flush=True)
        del step, opt, model
    bd.get_comm().check_health()


def graphed():
    """The caller-owned loop of the reference script (zero_grad, model(data), torch's CrossEntropyLoss, backward, torch.optim.SGD)
    on DistributedDataParallel(graph_forward_backward=True): the BN-buffer broadcast and the bucket collectives are inside the
    two replayed graphs."""
    from b200ddp.models import get_model
    from b200ddp.runtime import graphed as G
    from b200ddp.utils.ddp_check import grad_sync_check, params_identical
    rank, world, dev = bd.get_rank(), bd.get_world_size(), bd.get_device()
    bd.set_compute_dtype("fp32")
    crit = torch.nn.CrossEntropyLoss()
    for name, ncls in (("resnet18", 1000), ("custom", 10)):
        torch.manual_seed(4321 + rank)
        model = bd.DistributedDataParallel(get_model(name).to(dev), graph_forward_backward=True)
        model.train()
        B = 32
        g = torch.Generator().manual_seed(99 + rank)          # every rank its own shard
        xs = [torch.randn(B, 3, 32, 32, generator=g).to(dev) for _ in range(6)]
        ys = [torch.randint(0, ncls, (B,), generator=g).to(dev) for _ in range(6)]

        def fwd_bwd():
            crit(model(xs[0]), ys[0]).backward()

        before = dict(G.STATS)
        res = grad_sync_check(model, fwd_bwd)                  # local pass under no_sync runs eagerly, the synced one is captured
        st = {k: G.STATS[k] - before[k] for k in before}
        check("%s graphed: captured, no failure" % name, st["captures"] == 1 and st["capture_failures"] == 0 and st["backward_replays"] == 1, str(st))
        check("%s graphed: synced grads == mean(local grads)" % name, res["rel_err"] < 1e-5, "rel err %.2e" % res["rel_err"])
        check("%s graphed: synced grads bitwise equal across ranks" % name, res["bitwise_equal"])
        opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
        losses = []
        for x, y in zip(xs, ys):
            opt.zero_grad()
            loss = crit(model(x), y)
            loss.backward()
            opt.step()
            losses.append(float(loss.detach()))
        torch.cuda.synchronize()
        st = {k: G.STATS[k] - before[k] for k in before}
        check("%s graphed: 6 more steps were replays" % name, st["forward_replays"] == 7 and st["backward_replays"] == 7 and st["captures"] == 1, str(st))
        check("%s graphed: parameters identical across ranks after 6 steps" % name, params_identical(model))
        check("%s graphed: losses finite" % name, all(v == v for v in losses), "%.3f -> %.3f" % (losses[0], losses[-1]))
        if name == "resnet18":
            # the per-forward buffer broadcast is inside graph F: one more replay on the SAME input everywhere puts every
            # rank's running statistics in step with rank 0's (they differ after a forward on rank-specific shards)
            same = torch.randn(B, 3, 32, 32, generator=torch.Generator().manual_seed(5)).to(dev)
            model(same)
            torch.cuda.synchronize()
            bufs = gather_cpu(model.module.bn1.running_mean)
            check("resnet18 graphed: BN buffers follow rank 0", all(torch.allclose(bufs[0], b, rtol=1e-5, atol=1e-6) for b in bufs[1:]))
        del opt, model
    bd.get_comm().check_health()
    fused_momentum()


def fused_momentum():
    """``SGD.sync_state()``: with the optimizer folded into the bucket collectives a rank only ever updates the momentum of the
    shards it owns; before a resume checkpoint the buffers are completed (sum over ranks of the owned shards).  Checked
    against the same four steps with the optimizer as a separate pass, where every rank holds the whole state."""
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import optim as boptim
    from b200ddp.runtime.engine import TrainStep
    rank, dev = bd.get_rank(), bd.get_device()
    bd.set_compute_dtype("fp32")
    B = 32
    g = torch.Generator().manual_seed(7 + rank)
    xs = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, generator=g).pin_memory()
    ys = torch.randint(0, 10, (B,), generator=g).pin_memory()
    out = {}
    saved = os.environ.get("B200DDP_FUSED_SGD")
    try:
        for fused in (True, False):
            os.environ["B200DDP_FUSED_SGD"] = "1" if fused else "0"
            torch.manual_seed(99)
            model = bd.DistributedDataParallel(get_model("custom").to(dev))
            opt = boptim.SGD(model.parameters(), lr=0.01, momentum=0.9)
            step = TrainStep(model, CrossEntropyLoss(), opt, B, device=dev, graph=True, augment=False)
            for _ in range(4):
                step(xs, ys)
            torch.cuda.synchronize()
            mom = list(opt._momentum_flat.values())[0]
            if fused:
                check("fused SGD: the optimizer runs inside the collectives", bool(step.fused_optimizer))
                part = gather_cpu(mom)
                check("fused SGD: a rank holds only its own shards' momentum", not torch.equal(part[0], part[1]))
                opt.sync_state()
                opt.sync_state()                 # a second call must not add the shards again
                torch.cuda.synchronize()
            out[fused] = (gather_cpu(mom), model.flat_params.detach().cpu())
            del step, opt, model
    finally:
        if saved is None:
            os.environ.pop("B200DDP_FUSED_SGD", None)
        else:
            os.environ["B200DDP_FUSED_SGD"] = saved
    mf, mu = out[True][0], out[False][0]
    check("sync_state: momentum identical on every rank", all(torch.equal(mf[0], t) for t in mf[1:]))
    rel = float((mf[0] - mu[0]).norm() / (mu[0].norm() + 1e-30))
    check("sync_state: momentum == the separate optimizer pass's", rel < 1e-4, "rel %.2e" % rel)
    relp = float((out[True][1] - out[False][1]).norm() / (out[False][1].norm() + 1e-30))
    check("fused SGD: parameters == the separate optimizer pass's", relp < 1e-5, "rel %.2e" % relp)


def skip():
    comm = bd.get_comm()
    rank, world, dev = bd.get_rank(), bd.get_world_size(), bd.get_device()
    region, off = comm.alloc_arena(1 << 20, torch.uint8)
    comm.all_reduce_arena(off, 1024, scale=1.0)             # a healthy collective first
    torch.cuda.synchronize()
    comm.check_health()
    dist.barrier()
    t0 = time.time()
    raised = None
    if rank != 1:                                           # rank 1 never launches this one
        comm.all_reduce_arena(off, 1024, scale=1.0)
        torch.cuda.synchronize()
        try:
            comm.check_health()
        except RuntimeError as e:
            raised = str(e)
    dt = time.time() - t0
    limit = comm.cfg.spin_timeout_ms / 1e3
    ok = True if rank == 1 else (raised is not None and "timed out waiting for rank" in raised and dt < limit + 20.0)
    check("skipped collective: check_health raises within the spin timeout", ok,
          "rank %d: %.1f s, %s" % (rank, dt, (raised or "")[:80]))


def main():
    what = sys.argv[1]
    bd.init_process_group(backend="b200")
    try:
        {"collectives": collectives, "ddp": ddp, "graphed": graphed, "skip": skip}[what]()
    finally:
        if bd.get_rank() == 0:
            print("MP_GPU_RESULT " + json.dumps({"what": what, "world": bd.get_world_size(), "ok": all(RESULTS.values()) and bool(RESULTS),
                                                 "failed": [k for k, v in RESULTS.items() if not v], "checks": len(RESULTS)}), flush=True)
    ok = all(RESULTS.values()) and bool(RESULTS)
    if what == "skip":
        # the communicator's barrier epochs are out of step on purpose: no collective teardown
        sys.stdout.flush()
        os._exit(0 if ok else 1)
    bd.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
