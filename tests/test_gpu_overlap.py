"""Bucketed / backward-overlapped exchange and error feedback on the native path (one GPU, one rank).

Bucketing and overlap only change WHEN the exchange kernels run: thresholds and norms are per tensor and the random
numbers are indexed by the element's position in the flat buffer, so the updated parameters must be bit-identical to
the monolithic exchange.  Real peers are covered in tests/test_multigpu.py.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(compute="torch", **kw):
    from ewdml.config import TrainConfig
    from ewdml.train.gpu import GpuTrainer
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    dev = torch.device("cuda", 0)
    cfg = TrainConfig(network="ResNet18", dataset="Cifar10", batch_size=32, lr=0.05, momentum=0.9, exchange="allgather",
                      compressor="topk-qsgd", topk_ratio=0.01, dtype="bf16", compute=compute, enable_gpu=True, **kw).validate()
    return GpuTrainer(cfg, dev), dev


def _grad(total, gen, dev):
    """Random gradient whose magnitudes are all distinct (a permutation of 1..n times 2^-28, random signs): exact top-k
    then has no ties at the threshold, whose resolution by atomic ticket is the one order-dependent step of the kernels."""
    mag = (torch.randperm(total, generator=gen) + 1).float() * 2.0 ** -28
    sign = torch.randint(0, 2, (total,), generator=gen).float() * 2 - 1
    return (mag * sign).to(dev)


def _exchange_once(tr, g, overlapped):
    tr.flat.grad.copy_(g)
    ex = tr.exchange
    if overlapped:
        ex.begin_overlap()
        for b in range(len(ex.buckets) - 1, -1, -1):
            ex.bucket_ready(b)
        ex.finish_overlap()
    else:
        ex.step(0)


def test_bucketed_and_overlapped_exchange_is_bit_identical_to_monolithic():
    mono, dev = _setup()
    buck, _ = _setup(bucket_size=4 << 20)
    over, _ = _setup(bucket_size=4 << 20)
    assert len(mono.exchange.buckets) == 1 and len(buck.exchange.buckets) >= 3
    gen = torch.Generator(device="cpu").manual_seed(3)
    for step in range(3):
        g = _grad(mono.flat.total, gen, dev)
        _exchange_once(mono, g, False)
        _exchange_once(buck, g, False)
        _exchange_once(over, g, True)
    torch.cuda.synchronize()
    assert torch.equal(mono.flat.param, buck.flat.param)
    assert torch.equal(mono.flat.param, over.flat.param)
    assert torch.equal(mono.flat.momentum, over.flat.momentum)
    assert mono.check_device_error() == 0


@pytest.mark.parametrize("compute,graph", [("torch", False), ("native", False), ("native", True)])
def test_training_with_overlap_from_backward_hooks(compute, graph):
    tr, dev = _setup(compute=compute, bucket_size=4 << 20, overlap=True)
    assert tr.sched is not None and len(tr.sched.buckets) >= 3
    gen = torch.Generator(device="cpu").manual_seed(0)
    x = torch.randn(32, 3, 32, 32, generator=gen).to(dev).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (32,), generator=gen).to(dev)
    fired = []
    orig = tr.exchange.bucket_ready
    tr.exchange.bucket_ready = lambda b: (fired.append(b), orig(b))[1]
    tr.sched.on_ready = tr.exchange.bucket_ready
    before = tr.flat.param.clone()
    losses = [float(tr.step(x, y)[0])]
    assert fired == list(range(len(tr.sched.buckets) - 1, -1, -1))          # every bucket once, last layers first
    if graph:
        tr.capture(x, y)
    for _ in range(12):
        losses.append(float(tr.step(x, y)[0]))
    torch.cuda.synchronize()
    assert all(l == l for l in losses) and min(losses[1:]) < losses[0]      # the same batch is being fitted
    assert not torch.equal(before, tr.flat.param)
    assert tr.check_device_error() == 0
    tr.sched.remove()


def test_error_feedback_conserves_the_gradient_sum_on_device():
    tr, dev = _setup(error_feedback=True, bucket_size=4 << 20)
    ex = tr.exchange
    assert ex.residual is not None
    tr.cfg.lr = 0.0
    gen = torch.Generator(device="cpu").manual_seed(5)
    total = torch.zeros(tr.flat.total, device=dev)
    for step in range(4):
        g = torch.zeros(tr.flat.total)
        for s in tr.flat.segments:
            g[s.offset:s.offset + s.numel] = torch.randn(s.numel, generator=gen) * 0.01
        g = g.to(dev)
        total += g
        tr.flat.grad.copy_(g)
        before = ex.residual.clone()
        ex.step(step)
        torch.cuda.synchronize()
        sent = before + g - ex.residual                  # what left this rank at this step
        k = sum(max(1, int(s.numel * 0.01)) for s in tr.flat.segments)
        assert int((sent != 0).sum()) <= k               # only the selected entries were transmitted
    # after several steps the residual is bounded (mass is eventually sent), and nothing blew up
    assert torch.isfinite(ex.residual).all()
    assert float(ex.residual.norm()) < float(total.norm()) * 2.0
    assert tr.check_device_error() == 0


def test_poisoned_receive_slots_and_nvtx_ranges_do_not_change_results():
    """EWDML_DEBUG_POISON overwrites every consumed slot; if any kernel relied on bytes from an earlier step the
    parameters would differ from the clean run."""
    from ewdml.config import registry
    clean, dev = _setup()
    registry.set("debug_poison", True); registry.set("nvtx", True)
    try:
        dirty, _ = _setup()
    finally:
        registry.set("debug_poison", False); registry.set("nvtx", False)
    assert dirty.exchange._poison and not clean.exchange._poison
    gen = torch.Generator(device="cpu").manual_seed(11)
    for step in range(4):
        g = _grad(clean.flat.total, gen, dev)
        _exchange_once(clean, g, False)
        _exchange_once(dirty, g, False)
    torch.cuda.synchronize()
    assert torch.equal(clean.flat.param, dirty.flat.param)
    assert clean.check_device_error() == 0


def test_gpu_trainer_checkpoint_resume_is_exact(tmp_path):
    """save_checkpoint / load_checkpoint: a fresh trainer restored from the files continues with exactly the parameters,
    momentum and random-number position of the original (exchange-only steps on fixed gradients: deterministic)."""
    a, dev = _setup()
    gen = torch.Generator(device="cpu").manual_seed(21)
    gs = [_grad(a.flat.total, gen, dev) for _ in range(5)]
    for g in gs[:3]:
        _exchange_once(a, g, False)
        a.step_idx += 1
    path = str(tmp_path / "model_step_")
    a.save_checkpoint(path)
    for g in gs[3:]:
        _exchange_once(a, g, False)
    b, _ = _setup()
    assert b.load_checkpoint(path) == 3
    assert torch.equal(b.exchange.state.cpu()[1:], torch.tensor([3, 0, 3], dtype=torch.int32))
    for g in gs[3:]:
        _exchange_once(b, g, False)
    torch.cuda.synchronize()
    assert torch.equal(a.flat.param, b.flat.param) and torch.equal(a.flat.momentum, b.flat.momentum)
    sd = torch.load(path, map_location="cpu", weights_only=True)       # reference-format file
    assert "conv1.weight" in sd and sd["conv1.weight"].shape == (64, 3, 3, 3)


def test_adam_on_the_native_exchange_matches_flat_adam_on_the_decoded_gradient():
    """--optimizer adam: the receive kernel delivers the dense averaged gradient and FlatAdam applies it."""
    from ewdml.optim import FlatAdam
    tr, dev = _setup(optimizer="adam", bucket_size=4 << 20)
    ex = tr.exchange
    assert ex.adam is not None
    ref = FlatAdam(tr.flat.param.detach().clone(), lr=tr.cfg.lr)
    gen = torch.Generator(device="cpu").manual_seed(8)
    for step in range(3):
        _exchange_once(tr, _grad(tr.flat.total, gen, dev), False)
        torch.cuda.synchronize()
        k = sum(max(1, int(s.numel * 0.01)) for s in tr.flat.segments)
        assert 0 < int((ex._dense != 0).sum()) <= k
        ref.step(ex._dense)
    assert torch.allclose(ref.param, tr.flat.param, atol=1e-6)
    assert torch.equal(tr.flat.param_bf16.float(), tr.flat.param.to(torch.bfloat16).float()) if tr.flat.param_bf16 is not None else True
    assert tr.check_device_error() == 0
