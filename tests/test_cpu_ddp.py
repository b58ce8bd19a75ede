"""DDP semantics on the CPU/gloo fallback with two real processes (BASELINE.json config 1), the trainer
through the SageMaker-style launcher, and the reference's unmodified CPU script on our DDP."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CODE = "/root/reference/notebooks/code"


def _free_port():
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(script_args, world=2, timeout=600, extra_env=None):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        env.update(extra_env or {})
        procs.append(subprocess.Popen([sys.executable] + script_args, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                      text=True))
    outs = [p.communicate(timeout=timeout)[0] for p in procs]
    return [p.returncode for p in procs], outs


def test_ddp_semantics_gloo_world2(tmp_path):
    out = str(tmp_path / "res.json")
    rcs, outs = _spawn([os.path.join(ROOT, "tests", "mp_ddp_worker.py"), out])
    assert rcs == [0, 0], outs
    res = json.load(open(out))
    bad = [k for k, v in res.items() if v is not True]
    assert not bad, res


def test_trainer_two_cpu_hosts_writes_servable_checkpoint(tmp_path):
    """Notebook-1 path: two 'hosts', gloo, rank from SM_HOSTS/SM_CURRENT_HOST; rank 0 writes model.pth."""
    from b200ddp.utils.launch import launch
    rc = launch(2, ["--model-type", "custom", "--backend", "gloo", "--epochs", "1", "--train-size", "512", "--test-size", "64",
                    "--synthetic", "--batch-size", "64", "--log-interval", "2", "--test-batch-size", "64"],
                model_dir=str(tmp_path), cpu_hosts=True, timeout=600)
    assert rc == 0
    sd = torch.load(str(tmp_path / "model.pth"), map_location="cpu", weights_only=True)
    assert list(sd.keys())[0] == "conv1.weight" and sd["fc3.bias"].shape == (10,)
    from b200ddp.workloads.inference import model_fn
    assert model_fn(str(tmp_path))(torch.randn(2, 3, 32, 32)).shape == (2, 10)


def test_launcher_fails_fast_when_a_rank_dies(tmp_path):
    from b200ddp.utils.launch import launch
    rc = launch(2, ["--definitely-not-a-flag"], model_dir=str(tmp_path), cpu_hosts=True, timeout=120)
    assert rc != 0


@pytest.mark.skipif(not os.path.isdir(REF_CODE), reason="reference tree not mounted")
def test_reference_cpu_script_runs_unmodified_on_our_ddp(tmp_path):
    """Contract test: the reference's own `cifar10-distributed-native-cpu.py`, unmodified, with
    b200ddp substituted for torch's DDP and a synthetic CIFAR-10; it must train and write a model.pth that the
    reference's `inference.py` loads."""
    rcs, outs = _spawn([os.path.join(ROOT, "tests", "ref_script_worker.py"), str(tmp_path)], world=2, timeout=900)
    assert rcs == [0, 0], outs[0][-3000:] + outs[1][-3000:]
    assert "Train Epoch: 1" in outs[0] and "Test set: Average loss" in outs[0]
    sd = torch.load(str(tmp_path / "model.pth"), map_location="cpu", weights_only=True)
    assert "conv1.weight" in sd and not any(k.startswith("module.") for k in sd)


def test_launcher_takes_the_job_down_when_a_rank_dies_or_hangs():
    """Fault injection: rank 1 exits with code 3 while rank 0 sits in an all-reduce -> the launcher returns 3
    and stops rank 0; a rank that hangs trips the launcher's timeout (124).  Nothing keeps running."""
    import time
    from b200ddp.utils.launch import launch
    t0 = time.time()
    rc = launch(2, ["--fail-rank", "1", "--fail-after", "2", "--exit-code", "3"], module="b200ddp.utils.fault", cpu_hosts=True)
    assert rc == 3 and time.time() - t0 < 90
    t0 = time.time()
    rc = launch(2, ["--fail-rank", "0", "--fail-after", "1", "--hang"], module="b200ddp.utils.fault", cpu_hosts=True, timeout=8.0)
    assert rc == 124 and time.time() - t0 < 90
