"""Multi-GPU correctness of the NVLink collectives and of DDP on them (`-m gpu`; needs >= 2 GPUs on the box, skipped
otherwise).  Each test launches one process per GPU with torch.distributed.run and reads the worker's verdict
(tests/mp_gpu_worker.py).  World sizes: every one of 2, 4, 8 that the box has GPUs for."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mp_gpu_worker.py")


def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n]


def _run(what, world, port, timeout=900, env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER, what]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    out = res.stdout + res.stderr
    lines = [l for l in out.splitlines() if l.startswith("MP_GPU_RESULT ")]
    assert lines, out[-4000:]
    verdict = json.loads(lines[-1][len("MP_GPU_RESULT "):])
    fails = [l for l in out.splitlines() if l.startswith("CHECK") and " FAIL" in l]
    assert verdict["ok"] and res.returncode == 0, "\n".join(fails) + "\n" + out[-2000:]
    return verdict


@pytest.mark.parametrize("world", [2, 4, 8])
def test_collectives_every_algorithm_and_stress(world):
    if world not in _worlds():
        pytest.skip("needs %d GPUs" % world)
    v = _run("collectives", world, 29610 + world)
    assert v["checks"] >= 20


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ddp_semantics_on_nvlink(world):
    if world not in _worlds():
        pytest.skip("needs %d GPUs" % world)
    v = _run("ddp", world, 29630 + world)
    assert v["checks"] >= 14


@pytest.mark.parametrize("world", [2, 4, 8])
def test_graphed_forward_backward_on_nvlink(world):
    if world not in _worlds():
        pytest.skip("needs %d GPUs" % world)
    v = _run("graphed", world, 29670 + world)
    assert v["checks"] >= 17


def test_rank_that_skips_a_collective_is_detected():
    if 2 not in _worlds():
        pytest.skip("needs 2 GPUs")
    world = _worlds()[0]
    _run("skip", world, 29650, timeout=300, env_extra={"B200DDP_SPIN_TIMEOUT_MS": "2000"})
