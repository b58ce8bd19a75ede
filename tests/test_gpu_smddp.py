"""Contract test of the 'smddp' backend this repository ships: the reference's GPU training script
(``notebooks/code/cifar10-distributed-smddp-gpu.py``), byte-for-byte unmodified, must run on b200ddp through
``import smdistributed.dataparallel.torch.torch_smddp`` + ``init_process_group(backend='smddp')`` + torch's DDP class name
(``:17,23,148``), train for 30 steps, log the reference's lines and write ``model.pth`` with the reference's key set
(``module.``-prefixed, 1000-way ``fc`` for resnet18: ``:205-208``)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = [os.path.join(ROOT, "baseline", "_ref", "code", "cifar10-distributed-smddp-gpu.py"),
              "/root/reference/notebooks/code/cifar10-distributed-smddp-gpu.py"]


def _script():
    for p in CANDIDATES:
        if os.path.exists(p):
            return p
    pytest.skip("reference GPU script not staged (baseline/install_reference.py)")


@pytest.mark.parametrize("model_type", ["resnet18", "custom"])
def test_reference_gpu_script_runs_unmodified_on_b200ddp(tmp_path, model_type):
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    env = dict(os.environ)
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_gpu_script_worker.py"), str(tmp_path), model_type, _script()],
                         capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    out = res.stdout + res.stderr
    assert res.returncode == 0, out[-4000:]
    assert "Train Epoch: 1 [320/960 (33%)]" in out and "Test set: Average loss" in out, out[-3000:]
    assert "Saving the model." in out
    line = [l for l in out.splitlines() if l.startswith("NATIVE_LAUNCHES")][-1]
    assert int(line.split()[1]) > 30 * 10 and "backend=b200" in line and "dtype=fp32" in line, line   # the native kernels ran
    # 30 training batches of 32: model(data) / loss.backward() of the script's own loop ran as CUDA-graph replays
    import json
    stats = json.loads([l for l in out.splitlines() if l.startswith("GRAPHED ")][-1][len("GRAPHED "):])
    assert stats["captures"] == 1 and stats["capture_failures"] == 0, (stats, out[-3000:])
    assert stats["forward_replays"] == 30 and stats["backward_replays"] == 30, (stats, out[-3000:])
    sd = torch.load(str(tmp_path / "model.pth"), map_location="cpu", weights_only=True)
    assert all(k.startswith("module.") for k in sd)
    if model_type == "resnet18":
        assert sd["module.fc.weight"].shape == (1000, 512) and sd["module.conv1.weight"].shape == (64, 3, 7, 7)
        assert sd["module.layer4.1.bn2.running_var"].shape == (512,)
    else:
        assert sd["module.conv1.weight"].shape == (6, 3, 5, 5) and sd["module.fc3.bias"].shape == (10,)
    assert all(torch.isfinite(v.float()).all() for v in sd.values())


def test_torch_optim_on_native_ddp_matches_cpu_torch():
    """``torch.optim`` with its default ``zero_grad(set_to_none=True)`` on a DDP model whose native layers deposit
    their gradients straight into the arena -- parameters must move exactly as stock torch moves them on the CPU."""
    import copy
    import b200ddp as bd
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import functional as BF
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    started = bd.is_initialized()
    if not started:
        bd.init_process_group(backend="b200", rank=0, world_size=1)
    dev = bd.get_device()
    bd.set_compute_dtype("fp32")
    BF.assume_unit_loss_grad(False)
    try:
        torch.manual_seed(3)
        ref = get_model("custom")
        ddp = bd.DistributedDataParallel(copy.deepcopy(ref).to(dev))
        opt = torch.optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9)
        opt_r = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
        crit = CrossEntropyLoss()
        x, y = torch.randn(64, 3, 32, 32), torch.randint(0, 10, (64,))
        for _ in range(4):
            opt_r.zero_grad()
            torch.nn.functional.cross_entropy(ref(x), y).backward()
            opt_r.step()
            opt.zero_grad()                     # set_to_none=True: every param.grad becomes None
            crit(ddp(x.to(dev)), y.to(dev)).backward()
            assert all(p.grad is not None for p in ddp.parameters())
            opt.step()
        torch.cuda.synchronize()
        for (n, a), (_, b) in zip(ref.named_parameters(), ddp.module.named_parameters()):
            rel = float((b.detach().cpu() - a).norm() / (a.norm() + 1e-12))
            assert rel < 1e-2, (n, rel)        # four TF32 steps against strict fp32
    finally:
        bd.set_compute_dtype("bf16")
        if not started:
            bd.destroy_process_group()
