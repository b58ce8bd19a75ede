"""MNTD (trojan-detection) half on the CPU: models vs the reference's definitions and its 111 shipped MNIST
checkpoints, trojan settings/stamping parity, BackdoorDataset, zoo trainers, meta-classifier."""
import importlib.util
import os
import sys

import numpy as np
import pytest
import torch

REF_CODE = "/root/reference/notebooks/code"
CKPT_DIR = os.path.join(REF_CODE, "shadow_model_ckpt", "mnist", "models")
needs_ref = pytest.mark.skipif(not os.path.isdir(REF_CODE), reason="reference tree not mounted")


def _ref_module(rel, name):
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF_CODE, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@needs_ref
def test_all_shipped_mnist_checkpoints_load_and_reproduce_reference_logits():
    """Golden set: 111 real-weight checkpoints -> our MNIST model == the reference's model."""
    from b200ddp.models.mntd import MNISTModel
    ref = _ref_module("model_lib/mnist_cnn_model.py", "ref_mnist")
    files = sorted(f for f in os.listdir(CKPT_DIR) if f.endswith(".model"))
    assert len(files) == 111
    ours, theirs = MNISTModel(), ref.Model()
    ours.eval()
    theirs.eval()
    x = torch.rand(6, 1, 28, 28)
    for f in files:
        sd = torch.load(os.path.join(CKPT_DIR, f), map_location="cpu", weights_only=True)      # untrusted tree
        ours.load_state_dict(sd)
        theirs.load_state_dict(sd)
        assert torch.allclose(ours(x), theirs(x), atol=1e-5), f


@needs_ref
def test_cifar10_model_matches_reference_definition():
    from b200ddp.models.mntd import CIFAR10Model
    ref = _ref_module("model_lib/cifar10_cnn_model.py", "ref_c10")
    ours, theirs = CIFAR10Model().eval(), ref.Model().eval()
    theirs.load_state_dict(ours.state_dict())
    x = torch.rand(3, 3, 32, 32)
    assert torch.allclose(ours(x), theirs(x), atol=1e-5)
    assert float(ours.loss(ours(x), torch.tensor([1, 2, 3]))) == pytest.approx(float(theirs.loss(theirs(x), torch.tensor([1, 2, 3]))), abs=1e-6)


@needs_ref
@pytest.mark.parametrize("task,rel", [("mnist", "model_lib/mnist_cnn_model.py"), ("cifar10", "model_lib/cifar10_cnn_model.py")])
@pytest.mark.parametrize("troj_type", ["jumbo", "M", "B"])
def test_trojan_setting_and_stamp_match_reference(task, rel, troj_type):
    from b200ddp.models.mntd import random_troj_setting, stamp_batch, troj_gen_func
    ref = _ref_module(rel, "ref_" + task)
    for seed in range(4):
        np.random.seed(seed)
        theirs = ref.random_troj_setting(troj_type)
        np.random.seed(seed)
        ours = random_troj_setting(task, troj_type)
        assert ours[0] == theirs[0] and ours[2] == theirs[2] and ours[4] == theirs[4]
        assert np.allclose(ours[1], theirs[1]) and ours[3] == pytest.approx(theirs[3]) and ours[5] == pytest.approx(theirs[5])
        shape = (1, 28, 28) if task == "mnist" else (3, 32, 32)
        X = torch.rand(*shape)
        a, ya = troj_gen_func(task)(X, 0, ours)
        b, yb = ref.troj_gen_func(X, 0, theirs)
        assert torch.allclose(a, b) and ya == yb
        batch = torch.stack([X, X * 0.5, X * 0.25])
        mask = torch.tensor([1, 0, 1])
        out = stamp_batch(task, batch.clone(), ours, mask)
        assert torch.allclose(out[0], b, atol=1e-6) and torch.equal(out[1], batch[1])
        assert torch.allclose(out[2], ref.troj_gen_func(batch[2], 0, theirs)[0], atol=1e-6)


def test_audio_and_text_trojans_and_models_run():
    from b200ddp.models.mntd import AudioModel, RTNLPModel, random_troj_setting, stamp_batch, troj_gen_func
    np.random.seed(0)
    atk = random_troj_setting("audio", "M")
    X = torch.rand(16000)
    a, y = troj_gen_func("audio")(X, 3, atk)
    assert a.shape == X.shape and y == atk[4] and not torch.equal(a, X)
    out = stamp_batch("audio", torch.stack([X, X]).contiguous(), atk, None)
    assert torch.allclose(out[0], a, atol=1e-6)
    m = AudioModel().eval()
    logits = m(torch.randn(2, 16000) * 0.1)
    assert logits.shape == (2, 10) and torch.isfinite(logits).all()
    atk = random_troj_setting("rtNLP", "M")
    Xt = torch.tensor([5, 6, 7, 0, 0, 0, 0, 0, 0, 0])
    b, yb = troj_gen_func("rtNLP")(Xt, 1, atk)
    assert len(b) == 10 + atk[0] and yb == atk[4]
    nlp = RTNLPModel(emb_path="")            # seeded random embedding table: saved_emb.npy is not shipped
    s = nlp(torch.randint(1, 18000, (4, 12)))
    assert s.shape == (4,)
    mean, std = nlp.emb_info()
    assert mean.shape == (300,) and std.shape == (300,)
    e = torch.randn(10, 1, 10, 300, requires_grad=True)
    nlp.emb_forward(e).sum().backward()       # gradient flows into continuous queries (query tuning)
    assert e.grad is not None


def test_mel_filterbank_shape_and_normalisation():
    from b200ddp.models.mntd import mel_filterbank
    fb = mel_filterbank(16000, 2048, 40)
    assert fb.shape == (40, 1025) and float(fb.min()) >= 0 and (fb.sum(1) > 0).all()


def test_backdoor_dataset_semantics():
    from b200ddp.models.mntd import random_troj_setting
    from b200ddp.workloads.mntd import basic as UB
    _, _, trainset, _, _, _, _, gen, _ = UB.load_dataset_setting("mnist", synthetic_size=200)
    np.random.seed(1)
    atk = random_troj_setting("mnist", "M")
    choice = np.arange(50)
    ds = UB.BackdoorDataset(trainset, atk, gen, choice=choice)
    n_mal = int(50 * atk[5])
    assert len(ds) == 50 + n_mal
    x0, y0 = ds[0]
    assert torch.equal(x0, trainset[0][0]) and y0 == trainset[0][1]
    xm, ym = ds[50]
    assert ym == atk[4] and not torch.equal(xm, trainset[int(ds.mal_choice[0])][0])
    only = UB.BackdoorDataset(trainset, atk, gen, mal_only=True)
    assert len(only) == int(200 * atk[5]) and only[0][1] == atk[4]
    np.random.seed(1)
    tset = UB.TensorBackdoorSet("mnist", trainset, atk, choice=choice, batch_size=16)
    assert len(tset) == 50 + n_mal and int((tset.y[50:] == atk[4]).all())
    assert sum(x.shape[0] for x, _ in tset.batches()) == len(tset)


def test_zoo_trainers_and_meta_classifier_end_to_end(tmp_path):
    """benign + jumbo + trojaned trainers write the reference's file layout; the meta-classifier trains on it."""
    from b200ddp.workloads.mntd import run_meta, train_basic
    root = str(tmp_path / "ckpt")
    for kind in ("benign", "jumbo", "trojaned"):
        log = train_basic.run(kind, "mnist", "M", save_root=root, num=3, epochs=1, synthetic_size=300, seed=3)
        assert log["models"] >= 3 and 0.0 <= log["acc"] <= 1.0
    names = set(os.listdir(os.path.join(root, "mnist", "models")))
    assert {"shadow_benign_0.model", "target_benign_0.model", "shadow_jumbo_2.model", "target_trojM_1.model"} <= names
    assert os.path.exists(os.path.join(root, "mnist", "jumbo.log"))
    auc = run_meta.run("mnist", "M", shadow_root=root, save_root=str(tmp_path / "meta"), n_repeat=1, n_epoch=2,
                       train_num=2, val_num=1, test_num=3, verbose=False)
    assert 0.0 <= auc <= 1.0
    auc2 = run_meta.run("mnist", "M", load_exist=True, shadow_root=root, save_root=str(tmp_path / "meta"), n_repeat=1,
                        n_epoch=1, train_num=2, val_num=1, test_num=3, verbose=False)
    assert auc2 == pytest.approx(auc)


@needs_ref
def test_meta_classifier_on_shipped_checkpoints(tmp_path):
    """The reference's run_meta expects target_trojM_0..15 but ships 9..26: test_offset=9."""
    from b200ddp.workloads.mntd import run_meta
    root = os.path.join(REF_CODE, "shadow_model_ckpt")
    auc = run_meta.run("mnist", "M", shadow_root=root, save_root=str(tmp_path), n_repeat=1, n_epoch=2, train_num=6, val_num=3,
                       test_num=6, test_offset=9, verbose=False)
    assert 0.0 <= auc <= 1.0


def test_meta_classifier_oc_variant():
    from b200ddp.models.mntd import MNISTModel
    from b200ddp.workloads.mntd import meta as UM
    torch.manual_seed(0)
    import tempfile
    d = tempfile.mkdtemp()
    paths = []
    for i in range(3):
        p = os.path.join(d, "m%d.model" % i)
        torch.save(MNISTModel().state_dict(), p)
        paths.append((p, 1))
    oc = UM.MetaClassifierOC((1, 28, 28), 10)
    opt = torch.optim.Adam(oc.parameters(), lr=1e-3)
    loss = UM.epoch_meta_train_oc(oc, MNISTModel(), opt, paths, False)
    assert np.isfinite(loss) and isinstance(oc.r, float)
    auc, acc = UM.epoch_meta_eval_oc(oc, MNISTModel(), [(paths[0][0], 1), (paths[1][0], 0)], False, threshold="half")
    assert 0.0 <= acc <= 1.0


@needs_ref
def test_rt_polarity_is_prepared_from_the_shipped_raw_text_and_trains(tmp_path):
    """The reference ships raw rt-polarity text but not its processed arrays: prepare_rt_polarity builds the
    arrays RTNLP loads, load_dataset_setting picks them up, and the text CNN fits a few real sentences."""
    from b200ddp.workloads.mntd import basic as UB
    raw = os.path.join(REF_CODE, "raw_data", "rt_polarity")
    out = UB.prepare_rt_polarity(raw, str(tmp_path / "rt_polarity"))
    tr, dev = UB.RTNLP(True, out), UB.RTNLP(False, out)
    assert len(tr) + len(dev) == 10662 and len(dev) == 1066 and tr.Xs.shape[1] == 56
    assert 0 < int(tr.Xs.max()) < 18000 and set(np.unique(tr.ys)) == {0, 1}
    assert tr.idx2tok[str(tr.tok2idx["film"])] == "film"
    x0, y0 = tr[0]
    assert x0.dtype == torch.long and y0 in (0, 1)
    _, _, trainset, testset, is_binary, need_pad, Model, _, _ = UB.load_dataset_setting("rtNLP", root=str(tmp_path))
    assert isinstance(trainset, UB.RTNLP) and is_binary and need_pad and len(testset) == 1066
    torch.manual_seed(0)
    model = Model(emb_path="")
    sub = torch.utils.data.Subset(trainset, range(64))
    loader = UB.make_loader(sub, 32, shuffle=False)
    before = UB.eval_model(model, loader, True)
    UB.train_model(model, loader, 25, True, verbose=False, lr=1e-3)
    assert UB.eval_model(model, loader, True) > max(before, 0.9)


def test_speech_command_loader_reads_the_reference_npy_layout(tmp_path):
    from b200ddp.workloads.mntd import basic as UB
    proc = tmp_path / "speech_command" / "processed"
    proc.mkdir(parents=True)
    rng = np.random.RandomState(0)
    for split, n in (("train", 40), ("val", 8), ("test", 12)):
        np.save(proc / ("%s_data.npy" % split), rng.normal(0, 0.1, size=(n, 16000)).astype(np.float32))
        np.save(proc / ("%s_label.npy" % split), rng.randint(0, 30, size=n))
    ds = UB.SpeechCommand(0, str(proc))
    labels = np.load(proc / "train_label.npy")
    keep = [i for i, y in enumerate(labels) if UB.SPEECH_ALL_CLS[y] in UB.SPEECH_USED_CLS]
    assert len(ds) == len(keep) and all(0 <= ds[i][1] < 10 for i in range(len(ds)))
    assert UB.SPEECH_USED_CLS[ds[0][1]] == UB.SPEECH_ALL_CLS[labels[keep[0]]] and ds[0][0].shape == (16000,)
    _, _, trainset, testset, is_binary, _, _, _, _ = UB.load_dataset_setting("audio", root=str(tmp_path))
    assert isinstance(trainset, UB.SpeechCommand) and isinstance(testset, UB.SpeechCommand) and not is_binary


@pytest.mark.parametrize("mode", ["ddp", "replica"])
def test_zoo_trainer_distributed_modes_world2(tmp_path, mode):
    """The intent of the reference's two unparsable attempts (``train_basic_{jumbo,trojaned}_distributed_cpu.py``:
    ``DistributedSampler`` + ``DDP(model)`` + rank-0 save, :55-68) on gloo at world 2: ``--mode ddp`` data-parallelises every
    model, ``--mode replica`` deals the model indices out over the ranks.  Both must write every checkpoint exactly once,
    un-prefixed, loadable into the model definition, plus the JSON log."""
    from b200ddp.models.mntd.models import MODELS
    from b200ddp.utils.launch import launch
    root = str(tmp_path / "ckpt")
    rc = launch(2, ["jumbo", "--task", "mnist", "--save", root, "--num", "4", "--epochs", "1", "--synthetic-size", "256",
                    "--mode", mode, "--seed", "5"], module="b200ddp.workloads.mntd.train_basic", timeout=600)
    assert rc == 0
    names = sorted(os.listdir(os.path.join(root, "mnist", "models")))
    assert names == ["shadow_jumbo_%d.model" % i for i in range(4)]
    for n in names:
        sd = torch.load(os.path.join(root, "mnist", "models", n), map_location="cpu", weights_only=True)
        assert not any(k.startswith("module.") for k in sd)
        MODELS["mnist"](gpu=False).load_state_dict(sd)
    import json
    log = json.load(open(os.path.join(root, "mnist", "jumbo.log")))
    assert log["models"] == (4 if mode == "replica" else 4) and 0.0 <= log["acc"] <= 1.0
