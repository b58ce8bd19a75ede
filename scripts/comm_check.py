#!/usr/bin/env python
"""Multi-GPU correctness + bandwidth sweep of the fused all-reduce kernels (run under torchrun).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 scripts/comm_check.py [--sweep] [--nccl] [--max-mb 256]

Correctness: every algorithm (one-shot, two-shot, NVLS) x dtype (fp32, bf16) x a set of sizes against
the sum of the per-rank inputs computed in fp64 on the host; also broadcast and bitwise equality of the
result across ranks.  Sweep: device-timed (CUDA events, max over ranks) algorithm/bus bandwidth
1 KiB..max, optionally against torch's NCCL all_reduce on the same box (the baseline, not the product).
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ddp as bd  # noqa: E402
from b200ddp.ops import _lib as L  # noqa: E402


def expected_sum(n, world, dtype, seed):
    tot = torch.zeros(n, dtype=torch.float64)
    mine = None
    for r in range(world):
        g = torch.Generator().manual_seed(seed * 100 + r)
        v = torch.randn(n, generator=g).to(dtype)
        tot += v.double()
        if r == dist.get_rank():
            mine = v
    return tot, mine


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--nccl", action="store_true")
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", type=str, default="")
    ap.add_argument("--graph", action="store_true", help="time the all-reduces inside one CUDA graph (no host launch cost)")
    ap.add_argument("--ctas", type=str, default="", help="comma list of CTA counts: also time NVLS / two-shot with each (>= 1 MiB)")
    ap.add_argument("--skip-correct", action="store_true")
    a = ap.parse_args()
    bd.init_process_group(backend="b200")
    rank, world = bd.get_rank(), bd.get_world_size()
    dev = bd.get_device()
    comm = bd.get_comm()
    res = {"world": world, "nvls": comm.has_nvls, "correct": [], "sweep": []}
    algos = [L.ALGO_ONESHOT, L.ALGO_TWOSHOT] + ([L.ALGO_NVLS] if comm.has_nvls else [])
    big = 64 << 20
    arena_f32, off = comm.alloc_arena(max(a.max_mb << 20, big) // 4, torch.float32)
    ok_all = True
    # ------------------------------------------------------------------ correctness
    for dtype in (() if a.skip_correct else (torch.float32, torch.bfloat16)):
        es = 4 if dtype == torch.float32 else 2
        view = arena_f32.view(torch.uint8).view(dtype)
        for nbytes in (16, 1024, 4096 + 16, 248 * 1024 + 32, 2052000 // 16 * 16, 8 << 20, 15853824, 28852224):
            n = nbytes // es
            exp, mine = expected_sum(n, world, dtype, nbytes % 9973)
            for algo in algos:
                view[:n].copy_(mine.to(dev))
                view[n: n + 64].fill_(7.0)                     # canary after the range
                torch.cuda.synchronize()
                dist.barrier()
                comm.all_reduce_arena(off, n, scale=1.0 / world, dtype=dtype, algo=algo)
                torch.cuda.synchronize()
                got = view[:n].double().cpu()
                tol = 1e-5 if dtype == torch.float32 else 2e-2
                err = float((got - exp / world).abs().max() / (exp.abs().max() / world + 1e-9))
                canary = bool((view[n: n + 64].float() == 7.0).all())
                gathered = [None] * world
                dist.all_gather_object(gathered, got.float().sum().item())
                same = len(set(gathered)) == 1
                ok = err < tol and canary and same
                ok_all &= ok
                res["correct"].append({"dtype": str(dtype), "bytes": nbytes, "algo": L.ALGO_NAMES[algo], "err": err,
                                       "canary": canary, "ranks_equal": same, "ok": ok})
                dist.barrier()
    # broadcast
    view = arena_f32
    for root in range(world):
        view[:4096].fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        bd.broadcast(view[:4096], src=root)
        torch.cuda.synchronize()
        ok = bool((view[:4096] == float(root + 1)).all())
        ok_all &= ok
        res["correct"].append({"op": "broadcast", "root": root, "ok": ok})
        dist.barrier()
    # generic tensor all_reduce (staging path)
    t = torch.full((1000, 33), float(rank + 1), device=dev)
    bd.all_reduce(t)
    torch.cuda.synchronize()
    ok = bool((t == world * (world + 1) / 2).all())
    ok_all &= ok
    res["correct"].append({"op": "all_reduce(tensor)", "ok": ok})
    comm.check_health()
    # ------------------------------------------------------------------ sweep
    if a.sweep:
        nccl_group = dist.new_group(backend="nccl") if a.nccl else None
        sizes = [1 << s for s in range(10, 31) if (1 << s) <= (a.max_mb << 20)]
        cta_list = [int(c) for c in a.ctas.split(",") if c]
        base_ctas = comm.cfg.max_ctas
        for nbytes in sizes:
            n = nbytes // 4
            row = {"bytes": nbytes}
            variants = [(algo, base_ctas, L.ALGO_NAMES[algo]) for algo in algos]
            if nbytes >= (1 << 20):
                big_algo = L.ALGO_NVLS if comm.has_nvls else L.ALGO_TWOSHOT
                variants += [(big_algo, c, "%s_c%d" % (L.ALGO_NAMES[big_algo], c)) for c in cta_list]
            for algo, ctas, label in variants:
                comm.cfg.max_ctas = ctas
                if algo == L.ALGO_ONESHOT and nbytes > (8 << 20):
                    continue
                for _ in range(3):
                    comm.all_reduce_arena(off, n, scale=1.0 / world, algo=algo)
                torch.cuda.synchronize()
                dist.barrier()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if a.graph:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        for _ in range(a.iters):
                            comm.all_reduce_arena(off, n, scale=1.0 / world, algo=algo)
                    g.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    s.record()
                    g.replay()
                    e.record()
                else:
                    s.record()
                    for _ in range(a.iters):
                        comm.all_reduce_arena(off, n, scale=1.0 / world, algo=algo)
                    e.record()
                torch.cuda.synchronize()
                us = bd.max_over_ranks(s.elapsed_time(e) / a.iters * 1e3)
                row[label + "_us"] = us
                row[label + "_busbw_GBs"] = nbytes / (us * 1e-6) * 2 * (world - 1) / world / 1e9
            comm.cfg.max_ctas = base_ctas
            if nccl_group is not None:
                x = torch.ones(n, device=dev)
                for _ in range(3):
                    dist.all_reduce(x, group=nccl_group)
                torch.cuda.synchronize()
                dist.barrier()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(a.iters):
                    dist.all_reduce(x, group=nccl_group)
                e.record()
                torch.cuda.synchronize()
                us = bd.max_over_ranks(s.elapsed_time(e) / a.iters * 1e3)
                row["nccl_us"] = us
                row["nccl_busbw_GBs"] = nbytes / (us * 1e-6) * 2 * (world - 1) / world / 1e9
            res["sweep"].append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
    res["ok"] = bool(ok_all)
    if rank == 0:
        bad = [c for c in res["correct"] if not c["ok"]]
        print("COMM_CHECK world=%d nvls=%s ok=%s failures=%d" % (world, comm.has_nvls, ok_all, len(bad)))
        for c in bad[:20]:
            print("  FAIL", json.dumps(c))
        if a.out:
            with open(a.out, "w") as f:
                json.dump(res, f, indent=1)
    bd.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
