"""Model definitions, checkpoint format and optimizers."""
import os

import pytest
import torch

from ewdml.models import build_model
from ewdml.optim import SGD, Adam
from ewdml.parallel.flat import FlatModel
from ewdml.utils import checkpoint as ckpt


@pytest.mark.parametrize("name,nparams,ntensors,nkeys", [
    ("LeNet", 431080, 8, 8), ("VGG11", 9756426, 38, 62), ("ResNet50", 23520842, 161, 320),
    ("ResNet18", 11173962, 62, 122)])
def test_param_counts_match_reference(name, nparams, ntensors, nkeys):
    m = build_model(name, 10)
    ps = list(m.parameters())
    assert sum(p.numel() for p in ps) == nparams and len(ps) == ntensors
    assert len(m.state_dict()) == nkeys


def test_state_dict_key_names():
    assert list(build_model("LeNet").state_dict()) == [
        "conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
        "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]
    vk = list(build_model("VGG11").state_dict())
    assert vk[0] == "features.0.weight" and "features.1.running_mean" in vk and vk[-1] == "classifier.6.bias"
    rk = list(build_model("ResNet50").state_dict())
    assert "layer1.0.shortcut.0.weight" in rk and "layer4.2.bn3.num_batches_tracked" in rk and rk[-1] == "linear.bias"


def test_forward_shapes_and_readme_spellings():
    assert build_model("Resnet50", 100)(torch.randn(2, 3, 32, 32)).shape == (2, 100)
    assert build_model("ResNet34", 10)(torch.randn(1, 3, 32, 32)).shape == (1, 10)
    assert build_model("lenet")(torch.randn(3, 1, 28, 28)).shape == (3, 10)
    assert build_model("MnistNet")(torch.randn(3, 1, 28, 28)).exp().sum(1).allclose(torch.ones(3))


def test_loads_reference_checkpoint(reference_root):
    path = os.path.join(reference_root, "PyTorch-parameter-server/src/model_step_.zip")
    m = build_model("LeNet")
    ckpt.load_model(m, path)
    assert m.fc1.weight.shape == (500, 800)


def test_checkpoint_roundtrip_plain_state_dict(tmp_path):
    m = build_model("VGG11")
    FlatModel(m, channels_last=True)        # params become channels-last views of the flat buffer
    path = ckpt.model_path(str(tmp_path) + "/", 7, legacy_name=True)
    assert path.endswith("model_step_")
    ckpt.save_model(m, path)
    ckpt.save_sidecar(path, step=7, optimizer_state={"a": 1})
    sd = torch.load(path, weights_only=True)         # plain torch.load, like the reference evaluator
    assert isinstance(sd, dict) and sd["features.0.weight"].is_contiguous()
    m2 = build_model("VGG11"); m2.load_state_dict(sd)
    for a, b in zip(m.state_dict().values(), m2.state_dict().values()):
        assert torch.equal(a, b)
    assert ckpt.load_sidecar(path)["step"] == 7
    assert ckpt.model_path("d/", 9, legacy_name=False) == "d/model_step_9"


def test_flat_model_views_and_alignment():
    m = build_model("LeNet")
    ref = [p.detach().clone() for p in m.parameters()]
    f = FlatModel(m, channels_last=True)
    assert all(s.offset % 32 == 0 for s in f.segments) and f.num_params == 431080
    for p, r in zip(m.parameters(), ref):
        assert torch.equal(p, r)
    # conv weight storage is [Cout][kh][kw][Cin]
    w = m.conv2.weight
    assert w.shape == (50, 20, 5, 5) and w.stride() == (500, 1, 100, 20)
    out = m(torch.randn(4, 1, 28, 28)); out.sum().backward()
    assert m.fc1.weight.grad.data_ptr() == f.grad.data_ptr() + 4 * f.segments[4].offset
    assert f.grad.abs().sum() > 0
    f.zero_grad(); assert f.grad.abs().sum() == 0


@pytest.mark.parametrize("kw", [dict(momentum=0.9), dict(momentum=0.9, weight_decay=1e-3),
                                dict(momentum=0.9, nesterov=True), dict(momentum=0.0),
                                dict(momentum=0.5, dampening=0.1)])
def test_sgd_external_grads_matches_torch(kw):
    torch.manual_seed(0)
    p1 = [torch.randn(5, 3, requires_grad=True), torch.randn(7, requires_grad=True)]
    p2 = [p.detach().clone().requires_grad_() for p in p1]
    o1, o2 = SGD(p1, lr=0.1, **kw), torch.optim.SGD(p2, lr=0.1, **kw)
    for _ in range(4):
        gs = [torch.randn_like(p) for p in p1]
        for p, g in zip(p2, gs):
            p.grad = g.clone()
        o1.step(grads=gs); o2.step()
    for a, b in zip(p1, p2):
        assert torch.allclose(a, b, atol=1e-6)


def test_adam_external_numpy_grads_matches_torch():
    torch.manual_seed(0)
    p1 = [torch.randn(6, 2, requires_grad=True)]
    p2 = [p1[0].detach().clone().requires_grad_()]
    o1, o2 = Adam(p1, lr=1e-2), torch.optim.Adam(p2, lr=1e-2)
    for _ in range(5):
        g = torch.randn(6, 2)
        p2[0].grad = g.clone()
        o1.step(grads=[g.numpy()]); o2.step()
    assert torch.allclose(p1[0], p2[0], atol=1e-6)


def test_flat_adam_matches_per_parameter_adam():
    """FlatAdam (whole-model tensor ops on the flat buffer, used by the native exchanges with --optimizer adam) computes
    what the external-gradient Adam computes parameter by parameter."""
    import torch
    from ewdml.models import build_model
    from ewdml.optim import Adam, FlatAdam
    from ewdml.parallel.flat import FlatModel
    torch.manual_seed(0)
    f = FlatModel(build_model("LeNet", 10))
    ref = [torch.nn.Parameter(p.detach().clone()) for p in f.params]
    opt, fa = Adam(ref, lr=0.01, weight_decay=0.001), FlatAdam(f.param, lr=0.01, weight_decay=0.001)
    for _ in range(5):
        g = torch.randn(f.total) * 0.1
        for a, b in zip(f.segments[:-1], f.segments[1:]):
            g[a.offset + a.numel:b.offset] = 0
        opt.step(grads=[f.view(g, i).clone() for i in range(len(f.segments))])
        fa.step(g)
    for p, q in zip(ref, f.params):
        assert torch.allclose(p.detach(), q.detach(), atol=1e-6)
    sd = {k: v.clone() for k, v in fa.state_dict().items()}
    fb = FlatAdam(f.param.clone())
    fb.load_state_dict(sd)
    assert float(fb.t) == 5.0 and torch.equal(fb.exp_avg, fa.exp_avg)


def test_single_process_trainer_and_resume(tmp_path):
    """NNTrainer (nn_ops.py) trains; DistributedWorker writes model_step_ + sidecar; --resume restores the weights, the
    momentum buffers and the step counter, and training continues from the NEXT step."""
    import torch
    from ewdml.config import TrainConfig
    from ewdml.data import prepare_data
    from ewdml.train.roles import DistributedWorker, NNTrainer
    from ewdml.utils import checkpoint as ckpt
    base = dict(network="LeNet", dataset="MNIST", batch_size=16, lr=0.05, momentum=0.9, data="synthetic", synthetic_size=128,
                exchange="allgather", compressor="none", eval_freq=3, train_dir=str(tmp_path) + "/", log_interval=100)
    cfg = TrainConfig(max_steps=3, epochs=1, **base).validate()
    loader, test = prepare_data(cfg)
    t = NNTrainer(cfg).build_model(10)
    before = t.flat.param.clone()
    t.train_and_validate(loader, test)
    assert t.cur_step == 3 and not torch.equal(before, t.flat.param)

    w = DistributedWorker(TrainConfig(max_steps=5, epochs=2, **base).validate(), 0, 1).build_model(10)
    w.train(loader, test, total_steps=5)                      # checkpoints after steps 0 and 3
    path = ckpt.model_path(cfg.train_dir)
    side = ckpt.load_sidecar(path)
    assert side["step"] == 4 and side["optimizer"]["state"]
    r = DistributedWorker(TrainConfig(max_steps=6, epochs=2, resume=path, **base).validate(), 0, 1).build_model(10)
    assert r.cur_step == 4
    sd = torch.load(path, weights_only=True)
    assert all(torch.equal(sd[k], v.detach()) for k, v in r.network.state_dict().items())
    mom_saved = side["optimizer"]["state"][0]["momentum_buffer"]
    mom_now = r.optimizer.state_dict()["state"][0]["momentum_buffer"]
    assert torch.equal(mom_saved, mom_now)
    hist = r.train(loader, test, total_steps=6)
    assert [h["step"] for h in hist] == [4, 5]


def test_sidecar_is_tied_to_its_model_file(tmp_path):
    """Model file and sidecar are two renames.  A crash between them leaves a new model beside an old sidecar: the sidecar
    carries the CRC of the model bytes it belongs to and is ignored (weights kept, step / optimizer restart) when they differ."""
    import torch
    from ewdml.models import build_model
    from ewdml.utils import checkpoint as ckpt
    m = build_model("LeNet", 10)
    path = str(tmp_path / "model_step_")
    crc = ckpt.save_model(m, path)
    ckpt.save_sidecar(path, step=7, optimizer_state={"x": 1}, extra={"momentum": torch.ones(3)}, model_crc=crc, rng_rank=1)
    side = ckpt.load_sidecar(path)
    assert side["step"] == 7 and side["rng_rank"] == 1 and torch.equal(side["extra"]["momentum"], torch.ones(3))
    with torch.no_grad():
        m.fc2.bias.add_(1.0)
    ckpt.save_model(m, path)                         # "crash" before the matching sidecar is written
    assert ckpt.load_sidecar(path) is None
    ckpt.load_model(build_model("LeNet", 10), path)  # the weights themselves still load
    # ranks that did not write the sidecar derive distinct generator states from (seed, rank, step)
    ckpt.reseed_for_rank(1, 2, 7); a = torch.rand(4)
    ckpt.reseed_for_rank(1, 3, 7); b = torch.rand(4)
    ckpt.reseed_for_rank(1, 2, 7); c = torch.rand(4)
    assert torch.equal(a, c) and not torch.equal(a, b)
