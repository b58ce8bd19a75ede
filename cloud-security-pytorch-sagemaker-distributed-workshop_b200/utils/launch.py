"""Local launcher that fabricates the SageMaker container contract for N ranks on one box
(replaces ``mpirun -np 8 ... smddprun python -m mpi4py script.py`` of
``notebooks/2_pytorch_dist_smddp_gpu.ipynb:284`` and the two-instance job of notebook 1).

    python -m b200ddp.utils.launch --nproc 8 -- --model-type resnet18 --batch-size 256 --epochs 15

Sets ``SM_HOSTS / SM_CURRENT_HOST / SM_MODEL_DIR / SM_CHANNEL_TRAIN / SM_NUM_GPUS`` plus
``RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT`` and fails fast: the first rank that exits
non-zero takes the job down (``orte_abort_on_non_zero_status`` in the reference's launcher).
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import socket
import subprocess
import sys
import tempfile
import time


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(nproc: int, script_args, module: str = "b200ddp.workloads.cifar10_train", model_dir: str = "",
           data_dir: str = "", cpu_hosts: bool = False, timeout: float = 0.0) -> int:
    port = _free_port()
    model_dir = model_dir or os.path.join(tempfile.gettempdir(), "b200ddp_model")
    os.makedirs(model_dir, exist_ok=True)
    hosts = ["algo-%d" % (i + 1) for i in range(nproc)] if cpu_hosts else ["algo-1"]
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"SM_HOSTS": json.dumps(hosts), "SM_CURRENT_HOST": hosts[r] if cpu_hosts else "algo-1",
                    "SM_MODEL_DIR": model_dir, "SM_CHANNEL_TRAIN": data_dir, "SM_NUM_GPUS": str(0 if cpu_hosts else nproc),
                    "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
        if not cpu_hosts:
            env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc)})
        procs.append(subprocess.Popen([sys.executable, "-m", module] + list(script_args), env=env))
    t0 = time.time()
    rc = 0
    try:
        alive = set(range(nproc))
        while alive:
            for r in list(alive):
                code = procs[r].poll()
                if code is not None:
                    alive.discard(r)
                    if code != 0:
                        rc = code
                        raise RuntimeError("rank %d exited with %d" % (r, code))
            if timeout and time.time() - t0 > timeout:
                rc = 124
                raise RuntimeError("launch timeout")
            time.sleep(0.05)
    except RuntimeError as e:
        sys.stderr.write("b200ddp.launch: %s — stopping the other ranks\n" % e)
        for p in procs:
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
        for p in procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
    return rc


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--nproc", type=int, default=1)
    ap.add_argument("--module", type=str, default="b200ddp.workloads.cifar10_train")
    ap.add_argument("--model-dir", type=str, default="")
    ap.add_argument("--data-dir", type=str, default="")
    ap.add_argument("--cpu-hosts", action="store_true", help="notebook-1 style: one 'host' per process, gloo")
    ap.add_argument("--timeout", type=float, default=0.0)
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    rest = a.rest[1:] if a.rest and a.rest[0] == "--" else a.rest
    sys.exit(launch(a.nproc, rest, a.module, a.model_dir, a.data_dir, a.cpu_hosts, a.timeout))


if __name__ == "__main__":
    main()
