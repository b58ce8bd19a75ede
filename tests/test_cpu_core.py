"""CPU tier (no GPU): entry-point contract, data sharding, bucket layout, models, checkpoints."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_CODE = "/root/reference/notebooks/code"


# ------------------------------------------------------------------ config / env contract
def test_parse_args_reference_flags(monkeypatch):
    from b200ddp.utils.config import parse_args
    monkeypatch.setenv("SM_HOSTS", '["algo-1","algo-2"]')
    monkeypatch.setenv("SM_CURRENT_HOST", "algo-2")
    monkeypatch.setenv("SM_MODEL_DIR", "/tmp/m")
    monkeypatch.setenv("SM_CHANNEL_TRAIN", "/tmp/d")
    cfg = parse_args(["--model-type", "custom", "--batch-size", "64", "--epochs", "20", "--lr", "0.01",
                      "--momentum", "0.9", "--backend", "gloo", "--test-batch-size", "500", "--log-interval", "7"])
    assert (cfg.model_type, cfg.batch_size, cfg.epochs, cfg.lr, cfg.momentum) == ("custom", 64, 20, 0.01, 0.9)
    assert cfg.backend == "gloo" and cfg.test_batch_size == 500 and cfg.log_interval == 7
    assert cfg.model_dir == "/tmp/m" and cfg.data_dir == "/tmp/d"


def test_defaults_without_sagemaker_env(monkeypatch):
    from b200ddp.utils.config import parse_args
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "SM_MODEL_DIR", "SM_CHANNEL_TRAIN"):
        monkeypatch.delenv(k, raising=False)
    cfg = parse_args([])                 # the reference raises KeyError here
    assert cfg.hosts == ["algo-1"] and cfg.current_host == "algo-1" and cfg.momentum == 0.5 and cfg.seed == 1
    assert parse_args(["--hosts", "a,b"]).hosts == ["a", "b"]            # not split into characters
    assert parse_args(["--hosts", '["x","y","z"]']).hosts == ["x", "y", "z"]


def test_log_line_formats():
    from b200ddp.utils.logging import fmt_shard_line, fmt_test_line, fmt_train_line
    assert fmt_train_line(1, 3200, 6250, 51.0, 1.639701) == "Train Epoch: 1 [3200/6250 (51%)] Loss: 1.639701"
    assert fmt_test_line(-20.7798, 0.43) == "Test set: Average loss: -20.7798, Accuracy: 0.43\n"
    assert fmt_shard_line(6250, 50000, "train") == "Processes 6250/50000 (12%) of train data"


# ------------------------------------------------------------------ data
@pytest.mark.parametrize("n,world", [(50000, 8), (1003, 4), (10, 3)])
def test_shard_sampler_matches_torch_distributed_sampler(n, world):
    from torch.utils.data.distributed import DistributedSampler
    from b200ddp.utils.data import ShardSampler
    ds = list(range(n))
    for rank in range(world):
        ref = list(DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True, seed=0))
        ours = ShardSampler(n, world, rank, shuffle=True, seed=0)
        assert list(ours) == ref and len(ours) == len(ref)
    s = ShardSampler(n, world, 0)
    a = list(s)
    s.set_epoch(1)
    assert list(s) != a or n < 3


def test_batch_loader_and_synthetic_cifar():
    from b200ddp.utils.data import BatchLoader, ShardSampler, SyntheticCIFAR10
    ds = SyntheticCIFAR10(train=True, size=103)
    assert ds.data.shape == (103, 32, 32, 3) and ds.data.dtype == np.uint8
    loader = BatchLoader(ds, 32, sampler=ShardSampler(len(ds), 2, 1, shuffle=True, seed=0))
    batches = list(loader)
    assert len(batches) == len(loader) == 2 and batches[0][0].shape == (32, 32, 32, 3)
    assert sum(b[0].shape[0] for b in batches) == len(loader.sampler) == 52        # uneven last batch (20)
    assert batches[-1][0].shape[0] == 20


def test_augment_reference_matches_torchvision_normalize():
    import torchvision.transforms as T
    from PIL import Image
    from b200ddp.utils.data import CIFAR10_MEAN, CIFAR10_STD, augment_reference
    img = torch.randint(0, 256, (3, 32, 32, 3), dtype=torch.uint8)
    ident = torch.tensor([[4, 4, 0]] * 3)
    out = augment_reference(img, offsets=ident)
    tf = T.Compose([T.ToTensor(), T.Normalize(CIFAR10_MEAN, CIFAR10_STD)])
    ref = torch.stack([tf(Image.fromarray(img[i].numpy())) for i in range(3)])
    assert torch.allclose(out, ref, atol=1e-6)
    flipped = augment_reference(img, offsets=torch.tensor([[4, 4, 1]] * 3))
    assert torch.allclose(flipped, ref.flip(-1), atol=1e-6)


# ------------------------------------------------------------------ bucket layout
def test_bucket_assignment_matches_reference_reducer_sizes():
    from b200ddp.models import get_model
    from b200ddp.parallel.reducer import assign_buckets

    def bucket_bytes(name):
        params = list(reversed([p for p in get_model(name).parameters()]))
        _, buckets, total = assign_buckets([p.numel() for p in params], 1 << 20, 25 << 20)
        return [sum(params[i].numel() for i in b.param_ids) * 4 for b in buckets], [len(b.param_ids) for b in buckets]

    assert bucket_bytes("custom") == ([248024], [10])
    assert bucket_bytes("resnet18") == ([2052000, 28852224, 15853824], [2, 12, 48])
    assert bucket_bytes("resnet50")[0] == [8196000, 31502336, 26255360, 26550272, 9724160]
    ours = bucket_bytes("resnet18")[0]
    model = get_model("resnet18")
    idx = torch.distributed._compute_bucket_assignment_by_size(list(reversed(list(model.parameters()))), [1 << 20, 25 << 20])[0]
    ps = list(reversed(list(model.parameters())))
    assert ours == [sum(ps[i].numel() for i in b) * 4 for b in idx]


# ------------------------------------------------------------------ models / checkpoints
def test_state_dict_keys_match_torchvision_and_reference_net():
    import torchvision
    from b200ddp.models import get_model
    for name, tv in (("resnet18", torchvision.models.resnet18), ("resnet50", torchvision.models.resnet50)):
        a, b = get_model(name).state_dict(), tv().state_dict()
        assert list(a.keys()) == list(b.keys())
        assert all(a[k].shape == b[k].shape for k in a)
    net = get_model("custom")
    assert list(net.state_dict().keys()) == ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight",
                                             "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias"]
    assert sum(p.numel() for p in net.parameters()) == 62006
    assert sum(p.numel() for p in get_model("resnet18").parameters()) == 11689512


def test_workshop_net_cpu_forward_equals_reference_definition():
    if not os.path.isdir(REF_CODE):
        pytest.skip("reference tree not mounted")
    import importlib.util
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("ref_inference", os.path.join(REF_CODE, "inference.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from b200ddp.models import get_model
    ours, theirs = get_model("custom"), ref.Net()
    theirs.load_state_dict(ours.state_dict())
    x = torch.randn(5, 3, 32, 32)
    assert torch.allclose(ours(x), theirs(x), atol=1e-6)


def test_checkpoint_roundtrip_and_reference_model_fn(tmp_path):
    from b200ddp.models import get_model
    from b200ddp.utils import checkpoint as ck
    from b200ddp.workloads.inference import input_fn, model_fn, output_fn, predict_fn
    net = get_model("custom")
    path = ck.save_model(net, str(tmp_path))
    sd = torch.load(path, map_location="cpu", weights_only=True)
    assert not any(k.startswith("module.") for k in sd) and all(v.is_contiguous() for v in sd.values())
    served = model_fn(str(tmp_path))
    x = torch.randn(4, 3, 32, 32)
    assert torch.allclose(served(x), net.eval()(x), atol=1e-6)
    buf = __import__("io").BytesIO()
    np.save(buf, x.numpy())
    pred = predict_fn(input_fn(buf.getvalue(), "application/x-npy"), served)
    assert len(json.loads(output_fn(pred))) == 4
    if os.path.isdir(REF_CODE):            # the reference's unmodified serving hook loads our file
        import importlib.util
        sys.dont_write_bytecode = True
        spec = importlib.util.spec_from_file_location("ref_inference2", os.path.join(REF_CODE, "inference.py"))
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        assert torch.allclose(ref.model_fn(str(tmp_path))(x), net(x), atol=1e-6)
    # full checkpoint (beyond the reference): resume restores weights, optimizer state and epoch
    opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    net(x).sum().backward()
    opt.step()
    full = ck.save_full(net, opt, 3, str(tmp_path / "checkpoint.pt"))
    net2 = get_model("custom")
    opt2 = torch.optim.SGD(net2.parameters(), lr=0.1, momentum=0.9)
    assert ck.load_full(net2, opt2, full) == 3
    assert all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), net2.state_dict().values()))
    assert len(opt2.state_dict()["state"]) == len(opt.state_dict()["state"])


def test_native_library_builds_and_loads_without_gpu(native_lib):
    assert native_lib.b200_version() >= 100
    hdr = open(os.path.join(ROOT, "include", "b200ddp.h")).read()
    import ctypes
    for sym in ("b200_allreduce", "b200_conv_fwd", "b200_conv_dgrad", "b200_conv_wgrad", "b200_bn_stats", "b200_sgd_momentum",
                "b200_adam", "b200_softmax_xent", "b200_augment", "b200_trojan_stamp", "b200_broadcast", "b200_barrier"):
        assert hasattr(native_lib, sym) and sym in hdr


def test_sass_has_blackwell_tensor_core_and_multimem_instructions():
    """The conv kernel must be tcgen05/TMEM code, the all-reduce must carry multimem (NVLS) instructions."""
    obj = os.path.join(ROOT, "build", "conv_tc.o")
    if not os.path.exists(obj):
        from b200ddp.ops.build import build
        build()
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "LDGSTS" in sass
    ptx_or_sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "build", "comm.o")], capture_output=True, text=True).stdout
    assert "LDGMC" in ptx_or_sass            # SASS of multimem.ld_reduce (multimem.st shows as a multicast STG)


def test_tracer_writes_chrome_trace_and_trainer_flag(tmp_path):
    from b200ddp.utils.config import parse_args
    from b200ddp.utils.trace import Tracer
    tr = Tracer(rank=3)
    with tr.range("outer"):
        with tr.range("inner"):
            sum(range(1000))
    path = tr.save(str(tmp_path / "t.json"))
    ev = json.load(open(path))["traceEvents"]
    names = [e["name"] for e in ev if e["ph"] == "X"]
    assert "outer" in names and "inner" in names and all(e["pid"] == 3 for e in ev)
    outer = [e for e in ev if e["name"] == "outer" and e["ph"] == "X"][0]
    inner = [e for e in ev if e["name"] == "inner" and e["ph"] == "X"][0]
    assert outer["dur"] >= inner["dur"] and inner["args"]["depth"] == 1
    assert set(tr.summary_ms()) == {"outer", "inner"}
    assert parse_args(["--trace", "/tmp/x"]).trace == "/tmp/x"
    off = Tracer(enabled=False)
    with off.range("x"):
        pass
    assert off.summary_ms() == {}


def test_pipeline_api_falls_back_to_plain_steps_on_cpu():
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import functional as BF
    from b200ddp.runtime.engine import TrainStep
    torch.manual_seed(0)
    net = get_model("custom")
    opt = torch.optim.SGD(net.parameters(), lr=0.05)
    step = TrainStep(net, CrossEntropyLoss(), opt, 8, device=torch.device("cpu"), graph=True, augment=False)
    xs = torch.randint(0, 256, (3, 8, 32, 32, 3), dtype=torch.uint8)
    ys = torch.randint(0, 10, (3, 8))
    losses = list(step.pipeline((xs[i], ys[i]) for i in range(3)))
    assert len(losses) == 3 and all(isinstance(v, float) and v == v for v in losses) and step.steps_done == 3
    x = torch.randn(2, 3, 8, 8, requires_grad=True)
    assert BF.fork(x) is x and not hasattr(x, "_b200_mailbox")          # CPU tensors are never forked


def test_conv_bn_pairing_survives_deepcopy_and_keeps_state_dict_keys():
    import copy
    import torchvision
    from b200ddp.models import get_model
    m = get_model("resnet18")
    c = copy.deepcopy(m)
    blk, cblk = m.layer2[0], c.layer2[0]
    assert blk.conv1._bn_pair is blk.bn1 and blk.downsample[0]._bn_pair is blk.downsample[1] and m.conv1._bn_pair is m.bn1
    assert cblk.conv2._bn_pair is cblk.bn2 and cblk.conv2._bn_pair is not blk.bn2         # the copy pairs with ITS BatchNorm
    assert len(list(blk.conv1.children())) == 0                                           # not registered as a child module
    assert list(c.state_dict().keys()) == list(torchvision.models.resnet18().state_dict().keys())
    x = torch.randn(2, 3, 32, 32)
    assert torch.allclose(m.eval()(x), c.eval()(x))


def test_grad_mailbox_protocol():
    """The bookkeeping behind the fused skip-connection gradient: every consumer but the last deposits, the last one takes the sum."""
    from b200ddp.ops.functional import GradMailbox
    mb = GradMailbox(3)
    assert not mb.arrive()
    mb.grads.append(torch.ones(4))
    assert not mb.arrive()
    mb.grads.append(2 * torch.ones(4))
    assert mb.arrive()                                   # third consumer is the last
    assert torch.equal(mb.take(), 3 * torch.ones(4)) and mb.take() is None


def test_c_host_example_compiles_links_and_runs(native_lib, tmp_path):
    """include/b200ddp.h is plain C and the library can be driven without Python."""
    import shutil
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    libdir = os.path.join(ROOT, "cloud-security-pytorch-sagemaker-distributed-workshop_b200", "_lib")
    exe = str(tmp_path / "c_host")
    cmd = ["gcc", "-std=c99", "-Wall", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "c_host.c"),
           "-o", exe, "-L", libdir, "-lb200ddp", "-Wl,-rpath," + libdir, "-ldl", "-lpthread", "-lrt"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and "b200ddp C ABI version" in out.stdout, out.stdout + out.stderr


def test_trainer_saves_full_checkpoint_and_resumes(tmp_path, monkeypatch):
    """SURVEY 5.4 beyond the reference's save-only model.pth: --save-full-checkpoint then --resume continues at the next epoch
    with the same weights and optimizer state."""
    from b200ddp.utils.config import parse_args
    from b200ddp.workloads.cifar10_train import train
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    common = ["--model-type", "custom", "--synthetic", "--train-size", "256", "--test-size", "64", "--batch-size", "64", "--test-batch-size", "64",
              "--lr", "0.05", "--momentum", "0.9", "--backend", "gloo", "--model-dir", str(tmp_path), "--log-interval", "1000"]
    m1 = train(parse_args(common + ["--epochs", "1", "--save-full-checkpoint"]))
    ck = tmp_path / "checkpoint.pt"
    assert ck.exists() and (tmp_path / "model.pth").exists()
    w1 = {k: v.clone() for k, v in m1.state_dict().items()}
    m2 = train(parse_args(common + ["--epochs", "1", "--resume", str(ck)]))       # start_epoch = 2 > epochs: nothing left to train
    assert all(torch.equal(w1[k], v) for k, v in m2.state_dict().items())
    m3 = train(parse_args(common + ["--epochs", "2", "--resume", str(ck)]))       # one more epoch from the restored state
    assert any(not torch.equal(w1[k], v) for k, v in m3.state_dict().items())


def test_sgd_sync_state_shard_ownership_matches_the_fused_kernel():
    """``SGD._zero_unowned`` keeps exactly the range ``allreduce_sgd_shard_kernel`` (csrc/comm.cu) updates on this rank:
    16-byte chunks split as ceil(chunks / world) per rank, per bucket; summed over ranks every element is kept once."""
    from types import SimpleNamespace
    from b200ddp.ops.optim import SGD
    buckets = [SimpleNamespace(start=0, numel=40), SimpleNamespace(start=40, numel=4), SimpleNamespace(start=44, numel=100)]
    for world in (2, 3, 8):
        total = torch.zeros(144)
        for rank in range(world):
            m = torch.ones(144)
            SGD._zero_unowned(None, SimpleNamespace(world_size=world, rank=rank, buckets=buckets), m)
            total += m
            for b in buckets:
                chunks = b.numel // 4
                per = -(-chunks // world)
                lo, hi = min(rank * per, chunks) * 4, min((rank + 1) * per, chunks) * 4
                assert float(m[b.start + lo: b.start + hi].sum()) == hi - lo
        assert torch.equal(total, torch.ones(144)), world


def test_resnet18_cpu_fallback_matches_torchvision_including_bn_bookkeeping():
    """The CPU fallback of the native layers is plain torch math: one training step of resnet18 must leave the same outputs,
    running statistics and ``num_batches_tracked`` as torchvision's model with the same weights (``state_dict`` parity)."""
    import torchvision
    from b200ddp.models import get_model
    torch.manual_seed(0)
    ours = get_model("resnet18")
    ref = torchvision.models.resnet18()
    ref.load_state_dict(ours.state_dict())
    ours.train()
    ref.train()
    x = torch.randn(4, 3, 32, 32)
    yo, yr = ours(x), ref(x)
    assert torch.allclose(yo, yr, rtol=1e-4, atol=1e-4)
    so, sr = ours.state_dict(), ref.state_dict()
    assert set(so) == set(sr)
    for k in sr:
        if k.endswith("num_batches_tracked"):
            assert int(so[k]) == int(sr[k]) == 1, k
        elif "running_" in k:
            assert torch.allclose(so[k], sr[k], rtol=1e-4, atol=1e-5), k
    yo.sum().backward()
    yr.sum().backward()
    for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
        assert torch.allclose(a.grad, b.grad, rtol=2e-3, atol=2e-4), n


def test_train_step_notices_hyper_parameter_changes_and_scopes_its_switches():
    """Host-side logic of ``TrainStep`` that needs no GPU: a changed learning rate makes a captured step stale (the GPU test
    ``tests/test_gpu_engine.py`` checks the re-capture itself); deferring the BatchNorm step counters is scoped to the
    engine's own model; the unit-loss-gradient hint is only on while the engine's step runs."""
    from b200ddp import nn as bnn
    from b200ddp.models import get_model
    from b200ddp.nn import CrossEntropyLoss
    from b200ddp.ops import functional as BF
    from b200ddp.ops import optim as boptim
    from b200ddp.runtime.engine import TrainStep
    net = get_model("custom")
    opt = boptim.SGD(net.parameters(), lr=0.1, momentum=0.9)
    step = TrainStep(net, CrossEntropyLoss(), opt, 4, device=torch.device("cpu"), graph=True)
    assert not step.use_graph and not step._stale()          # no GPU: nothing is ever captured
    step._graph, step._captured_key = object(), step._hyper_key()
    assert not step._stale()
    opt.param_groups[0]["lr"] = 0.01
    assert step._stale()
    step._captured_key = step._hyper_key()
    opt.param_groups[0]["momentum"] = 0.8
    assert step._stale()
    step._graph = None
    # the step itself still runs (eagerly) and leaves the process-wide hint as it found it
    BF.assume_unit_loss_grad(False)
    x = torch.randint(0, 256, (4, 32, 32, 3), dtype=torch.uint8)
    y = torch.randint(0, 10, (4,))
    loss = step(x, y, need_loss=True)
    assert loss == loss and BF._unit_loss_grad[0] is False
    # counters: deferral set for one model does not silence another model's BatchNorm layers
    a, b = bnn.BatchNorm2d(8), bnn.BatchNorm2d(8)
    bnn.defer_bn_counters(True, a)
    a.train(), b.train()
    t = torch.randn(2, 8, 4, 4)
    a(t), b(t)
    assert int(a.num_batches_tracked) == 0 and int(b.num_batches_tracked) == 1
    bnn.flush_bn_counters(a, 1)
    assert int(a.num_batches_tracked) == 1
    bnn.defer_bn_counters(False, a)
    a(t)
    assert int(a.num_batches_tracked) == 2


def test_trainer_step_learning_rate_schedule(tmp_path, monkeypatch):
    """``--lr-step-epochs / --lr-gamma`` (an extension: the reference trains at a constant rate): the rate the optimizer steps
    with follows the schedule epoch by epoch, also after a resume."""
    from b200ddp.utils.config import parse_args
    from b200ddp.workloads.cifar10_train import train
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    common = ["--model-type", "custom", "--synthetic", "--train-size", "128", "--test-size", "64", "--batch-size", "64", "--test-batch-size", "64",
              "--lr", "0.1", "--momentum", "0.9", "--backend", "gloo", "--model-dir", str(tmp_path), "--log-interval", "1000", "--no-eval",
              "--lr-step-epochs", "1", "--lr-gamma", "0.1"]
    cfg = parse_args(common + ["--epochs", "3", "--save-full-checkpoint"])
    assert cfg.lr_step_epochs == 1 and cfg.lr_gamma == 0.1
    import io
    import logging
    buf = io.StringIO()
    handler = logging.StreamHandler(buf)           # the package logger does not propagate: listen to it directly
    logging.getLogger("b200ddp").addHandler(handler)
    try:
        train(cfg)
        text = buf.getvalue()
        assert "Epoch 2: learning rate 0.01" in text and "Epoch 3: learning rate 0.001" in text and "Epoch 1: learning rate" not in text
        buf.truncate(0)
        train(parse_args(common + ["--epochs", "4", "--resume", str(tmp_path / "checkpoint.pt")]))
        assert "Epoch 4: learning rate 0.0001" in buf.getvalue()
    finally:
        logging.getLogger("b200ddp").removeHandler(handler)


def test_serving_endpoint_ping_and_invocations(tmp_path):
    """The reference reaches its ``inference.py`` through SageMaker's serving container (nb1:153-211: deploy, then
    ``predictor.predict``); ``workloads/serve.py`` is that contract stand-alone: ``GET /ping``, ``POST /invocations`` with
    npy or json bodies, answers equal to calling the hooks directly."""
    import io
    import threading
    import urllib.error
    import urllib.request
    from b200ddp.models import get_model
    from b200ddp.utils.checkpoint import save_model
    from b200ddp.workloads import inference, serve
    torch.manual_seed(0)
    net = get_model("custom")
    save_model(net, str(tmp_path))
    srv = serve.make_server(str(tmp_path), port=0)
    host, port = srv.server_address
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    try:
        base = "http://%s:%d" % (host, port)
        with urllib.request.urlopen(base + "/ping", timeout=10) as r:
            assert r.status == 200 and r.read() == b""
        x = np.random.RandomState(0).randn(3, 3, 32, 32).astype(np.float32)
        want = net.eval()(torch.from_numpy(x)).detach().numpy()
        buf = io.BytesIO()
        np.save(buf, x)
        req = urllib.request.Request(base + "/invocations", data=buf.getvalue(), headers={"Content-Type": "application/x-npy",
                                                                                          "Accept": "application/x-npy"})
        with urllib.request.urlopen(req, timeout=30) as r:
            got = np.load(io.BytesIO(r.read()))
        assert np.allclose(got, want, atol=1e-5)
        req = urllib.request.Request(base + "/invocations", data=json.dumps(x.tolist()).encode(),
                                     headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=30) as r:
            assert r.headers["Content-Type"] == "application/json"
            got = np.array(json.loads(r.read()), dtype=np.float32)
        assert np.allclose(got, want, atol=1e-4)
        assert np.allclose(got, np.array(json.loads(inference.output_fn(inference.predict_fn(inference.input_fn(buf.getvalue()),
                                                                                           inference.model_fn(str(tmp_path))))), dtype=np.float32), atol=1e-6)
        for path, data, headers, code in (("/invocations", b"x", {"Content-Type": "text/csv"}, 415),
                                          ("/invocations", b"not json", {"Content-Type": "application/json"}, 400),
                                          ("/nowhere", b"", {}, 404)):
            try:
                urllib.request.urlopen(urllib.request.Request(base + path, data=data, headers=headers), timeout=10)
                raise AssertionError("expected HTTP %d" % code)
            except urllib.error.HTTPError as e:
                assert e.code == code, (path, e.code)
    finally:
        srv.shutdown()
        srv.server_close()
        th.join(timeout=10)
