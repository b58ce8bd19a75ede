"""``DistributedDataParallel(graph_forward_backward=True)``: a caller-owned loop written like the reference's
(``notebooks/code/cifar10-distributed-smddp-gpu.py:162-168``: ``zero_grad``, ``model(data)``, the caller's loss,
``loss.backward()``, the caller's ``torch.optim.SGD``) must train exactly as the launch-by-launch path does, while
``model(data)`` and ``loss.backward()`` are one CUDA-graph replay each (``runtime/graphed.py``)."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


def _setup():
    import b200ddp as bd
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    started = bd.is_initialized()
    if not started:
        bd.init_process_group(backend="b200", rank=0, world_size=1)
    return bd, started


@pytest.mark.parametrize("model_name,dtype", [("resnet18", "fp32"), ("custom", "fp32"), ("resnet18", "bf16")])
def test_graphed_forward_backward_trains_like_eager(model_name, dtype):
    bd, started = _setup()
    from b200ddp.models import get_model
    from b200ddp.nn import modules as bmod
    from b200ddp.ops import functional as BF
    from b200ddp.ops import native as nv
    from b200ddp.runtime import graphed
    dev = bd.get_device()
    bd.set_compute_dtype(dtype)
    BF.assume_unit_loss_grad(False)
    # resnet18 at random initialisation amplifies last-bit differences of the atomically accumulated reductions into
    # percent-level gradient differences from one run to the next (profiles/determinism.md): compare in the mode where two
    # passes over the same batch are bit-identical, so that any difference left is the graph path's
    saved = (nv.SPLIT_K[0], bmod._FUSE_BN_STATS[0])
    nv.SPLIT_K[0], bmod._FUSE_BN_STATS[0] = False, False
    try:
        torch.manual_seed(11)
        ddp = bd.DistributedDataParallel(get_model(model_name).to(dev), graph_forward_backward=True)
        g = ddp._graphed
        assert g is not None
        ncls = 1000 if model_name == "resnet18" else 10
        steps, B = 5, 32
        xs = [torch.randn(B, 3, 32, 32, device=dev) for _ in range(steps)]
        ys = [torch.randint(0, ncls, (B,), device=dev) for _ in range(steps)]
        crit = nn.CrossEntropyLoss()
        opt = torch.optim.SGD(ddp.parameters(), lr=0.05, momentum=0.9)
        before = dict(graphed.STATS)
        worst = 0.0
        for x, y in zip(xs, ys):
            # the SAME weights and batch through both paths, written the way the script writes its loop; the optimizer then
            # moves the weights, so a graph that kept reading anything stale would show at the next step
            res = {}
            for mode in ("eager", "graph"):
                g.paused = mode == "eager"
                opt.zero_grad()                   # set_to_none=True: every param.grad is dropped, as in the script
                out = ddp(x)
                loss = crit(out, y)
                loss.backward()
                assert all(p.grad is not None for p in ddp.parameters())
                res[mode] = (out.detach().float().clone(), float(loss.detach()), ddp.flat_grads.clone())
            (oe, le, ge), (og, lg, gg) = res["eager"], res["graph"]
            assert float((oe - og).abs().max()) <= 1e-6 * float(oe.abs().max()), (model_name, float((oe - og).abs().max()))
            assert abs(le - lg) <= 1e-6 * (1.0 + abs(le)), (le, lg)
            rel = float((ge - gg).norm() / (ge.norm() + 1e-30))
            worst = max(worst, rel)
            assert rel < 1e-5, (model_name, dtype, rel)
            opt.step()                            # on the graph path's gradients
        torch.cuda.synchronize()
        st = {k: graphed.STATS[k] - before[k] for k in before}
        if g.failure:
            print("CAPTURE FAILURE\n" + g.failure)
        assert st["captures"] == 1 and st["capture_failures"] == 0, st
        assert st["forward_replays"] == steps and st["backward_replays"] == steps and st["eager_forwards"] == steps, st
        print(model_name, dtype, "worst relative gradient difference graph vs eager %.2e" % worst, st)
        for n, b in ddp.module.named_buffers():
            if n.endswith("num_batches_tracked"):
                assert int(b) == 2 * steps, (n, int(b))      # the warm-up passes and the capture left no trace
        if ddp.flat_shadow is not None:       # torch.optim moved the masters: the bf16 shadow must follow at the next forward
            ddp(xs[0])
            assert torch.equal(ddp.flat_shadow, ddp.flat_params.to(ddp.flat_shadow.dtype))
        graph = ddp

        # ---- everything that does not fit the capture runs launch by launch
        n_eager = graphed.STATS["eager_forwards"]
        graph.eval()
        with torch.no_grad():
            og = graph(xs[0])
        assert graphed.STATS["eager_forwards"] == n_eager + 1 and bool(torch.isfinite(og.float()).all())
        graph.train()
        out = graph(xs[0][:8])                # a short batch
        assert graphed.STATS["eager_forwards"] == n_eager + 2 and out.shape[0] == 8
        out.float().sum().backward()
        # ---- an output whose activations were replaced by a later forward cannot be back-propagated
        o1 = graph(xs[1])
        o2 = graph(xs[2])
        with pytest.raises(RuntimeError, match="replaced by a later forward"):
            o1.sum().backward()
        o2.sum().backward()
        torch.cuda.synchronize()
    finally:
        nv.SPLIT_K[0], bmod._FUSE_BN_STATS[0] = saved
        bd.set_compute_dtype("bf16")
        if not started:
            bd.destroy_process_group()


def test_module_with_a_stock_layer_falls_back_cleanly():
    """A parameter whose gradient comes through autograd (a stock ``nn.Linear``) cannot be served by the replayed backward:
    the warm-up notices, nothing is captured, the loop trains launch by launch and the device's RNG is untouched."""
    bd, started = _setup()
    from b200ddp.runtime import graphed
    dev = bd.get_device()
    bd.set_compute_dtype("fp32")
    try:
        torch.manual_seed(3)
        net = nn.Sequential(nn.Flatten(), nn.Linear(3 * 32 * 32, 10)).to(dev)
        ref = nn.Sequential(nn.Flatten(), nn.Linear(3 * 32 * 32, 10)).to(dev)
        ref.load_state_dict(net.state_dict())
        ddp = bd.DistributedDataParallel(net, graph_forward_backward=True)
        before = dict(graphed.STATS)
        opt, opt_r = torch.optim.SGD(ddp.parameters(), lr=0.1), torch.optim.SGD(ref.parameters(), lr=0.1)
        for _ in range(3):
            x, y = torch.randn(16, 3, 32, 32, device=dev), torch.randint(0, 10, (16,), device=dev)
            for m, o in ((ddp, opt), (ref, opt_r)):
                o.zero_grad()
                nn.functional.cross_entropy(m(x), y).backward()
                o.step()
        st = {k: graphed.STATS[k] - before[k] for k in before}
        assert st["captures"] == 0 and st["capture_failures"] == 1 and st["forward_replays"] == 0, st
        assert not ddp._graphed.enabled
        for a, b in zip(ddp.module.parameters(), ref.parameters()):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    finally:
        bd.set_compute_dtype("bf16")
        if not started:
            bd.destroy_process_group()
