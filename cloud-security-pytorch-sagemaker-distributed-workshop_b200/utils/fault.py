"""Fault injection for the failure path (the reference relies on its launcher's
``-mca orte_abort_on_non_zero_status 1``, ``notebooks/2_pytorch_dist_smddp_gpu.ipynb:284`` -- one dead rank takes the job down).

    python -m b200ddp.utils.launch --nproc 2 --cpu-hosts --module b200ddp.utils.fault -- --fail-rank 1 --fail-after 2

Every rank loops over all-reduces; the chosen rank dies (``--exit-code``) or hangs (``--hang``) after ``--fail-after`` steps
while its peers are blocked inside the next collective.  The launcher must notice the dead rank, stop the others and return its
exit code (or 124 on ``--timeout``); nothing may be left running.  Used by ``tests/test_cpu_ddp.py``.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--fail-rank", type=int, default=-1)
    ap.add_argument("--fail-after", type=int, default=2)
    ap.add_argument("--exit-code", type=int, default=3)
    ap.add_argument("--hang", action="store_true")
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--backend", type=str, default="gloo")
    a = ap.parse_args(argv)
    from .. import _api as bd
    from .config import parse_args
    cfg = parse_args([])                                   # SM_HOSTS / SM_CURRENT_HOST -> rank, world (as the trainer does)
    if "RANK" in os.environ:
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    else:
        rank, world = cfg.hosts.index(cfg.current_host), len(cfg.hosts)
    bd.init_process_group(backend=a.backend, rank=rank, world_size=world)
    t = torch.ones(1024)
    for step in range(a.steps):
        if rank == a.fail_rank and step == a.fail_after:
            if a.hang:
                time.sleep(3600)
            sys.stderr.write("fault: rank %d dies at step %d\n" % (rank, step))
            sys.stderr.flush()
            os._exit(a.exit_code)
        bd.all_reduce(t)
        t.fill_(1.0)
        time.sleep(0.01)
    bd.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
