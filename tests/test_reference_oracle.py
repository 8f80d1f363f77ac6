"""The unmodified reference (installed under baseline/_ref by baseline/install_reference.sh) as the oracle.

Where the other CPU tests restate the reference's arithmetic, these import the reference's own modules and compare
object against object: its compressors (Compresssor/TopK.py, Compresssor/qsgd.py) on the same tensors and the same
random draws, and its model definitions (model_ops/{lenet,vgg,resnet}.py) through the state-dict: a checkpoint written
by either side loads into the other and the two forwards agree.  Skipped on a checkout where the reference has not
been installed (baseline/_ref is not tracked).
"""
import os
import sys

import pytest
import torch

try:                       # before any namespace package of the reference is in sys.modules (torchvision's import walks them)
    import torchvision     # noqa: F401
    _HAVE_TV = True
except Exception:          # noqa: BLE001
    _HAVE_TV = False

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "Compresssor")),
                                reason="reference not installed (bash baseline/install_reference.sh)")


_TOP = ("Compresssor", "model_ops", "optim", "nn_ops", "util")


@pytest.fixture(scope="module")
def ref():
    """The reference's packages, imported from baseline/_ref under their own top-level names.  ewdml/compat offers modules
    of the SAME names (tests/test_compat.py puts them on sys.path), so whatever is cached under those names is set aside
    first and put back afterwards, and every imported module is checked to come from baseline/_ref."""
    import importlib
    saved = {n: sys.modules.pop(n) for n in list(sys.modules) if n.split(".")[0] in _TOP}
    saved_path = list(sys.path)
    # the reference's Compresssor/ has no __init__.py (a namespace package), and a regular package of that name anywhere
    # on the path would win over it: the compat directory leaves the path while the reference is being imported
    compat = os.path.join(ROOT, "ewdml", "compat")
    sys.path[:] = [REF] + [p for p in sys.path if os.path.abspath(p or ".") != compat]
    importlib.invalidate_caches()
    try:
        mods = {n: importlib.import_module(n) for n in
                ("Compresssor.TopK", "Compresssor.qsgd", "model_ops.lenet", "model_ops.vgg", "model_ops.resnet",
                 "optim.sgd", "optim.adam", "nn_ops")}
        for n, m in mods.items():
            assert os.path.abspath(m.__file__).startswith(REF + os.sep), (n, m.__file__)
        yield mods
    finally:
        sys.path[:] = saved_path
        importlib.invalidate_caches()
        for n in [m for m in sys.modules if m.split(".")[0] in _TOP]:
            del sys.modules[n]
        sys.modules.update(saved)


def _distinct(n, seed):
    """Values with pairwise distinct magnitudes, so top-k has one right answer."""
    g = torch.Generator().manual_seed(seed)
    mag = (torch.randperm(n, generator=g) + 1).float() / n
    sign = torch.randint(0, 2, (n,), generator=g).float() * 2 - 1
    return mag * sign


@pytest.mark.parametrize("shape,ratio", [((64, 27), 0.5), ((4096,), 0.01), ((7,), 0.01), ((50, 20, 5, 5), 0.1)])
def test_topk_agrees_with_the_reference_compressor(ref, shape, ratio):
    from ewdml.compress.topk import TopKCompressor
    n = 1
    for d in shape:
        n *= d
    t = _distinct(n, 3).view(shape)
    theirs = ref["Compresssor.TopK"].TopKCompressor(ratio)
    ours = TopKCompressor(ratio)
    (tv, ti), tctx = theirs.compress(t)
    (ov, oi), octx = ours.compress(t)
    assert tv.numel() == ov.numel() == max(1, int(n * ratio))
    assert sorted(ti.tolist()) == sorted(oi.tolist())
    assert torch.equal(theirs.decompress((tv, ti), tctx), ours.decompress((ov, oi), octx))
    # and the two are interchangeable: each side decodes what the other encoded
    assert torch.equal(theirs.decompress((ov, oi), tctx), ours.decompress((tv, ti), octx))


@pytest.mark.parametrize("n", [1, 33, 10000])
def test_qsgd_agrees_with_the_reference_compressor_on_the_same_draws(ref, n):
    from ewdml.compress.qsgd import QSGDCompressor, qsgd_dequantize, qsgd_quantize
    t = torch.randn(n, generator=torch.Generator().manual_seed(n)) * 0.1
    theirs = ref["Compresssor.qsgd"].QSGDCompressor()
    torch.manual_seed(11)
    u = torch.empty_like(t).uniform_()           # the draw the reference is about to make
    torch.manual_seed(11)
    levels_ref, norm_ref = theirs.compress(t)
    q, norm = qsgd_quantize(t, theirs.quantum_num, u)
    assert torch.equal(norm.flatten().float(), norm_ref.flatten().float())
    assert torch.equal(q.float().flatten(), levels_ref.flatten())
    assert torch.equal(qsgd_dequantize(q, norm, theirs.quantum_num).flatten(), theirs.decompress((levels_ref, norm_ref)).flatten())
    # the class wrapper takes the same draws and packs the levels into the narrow dtype; decoding is unchanged
    ours = QSGDCompressor(theirs.quantum_num)
    payload, ctx = ours.compress(t, uniform=u)
    assert torch.equal(ours.decompress(payload, ctx).flatten(), theirs.decompress((levels_ref, norm_ref)).flatten())


def _load_both_ways(theirs, ours, x):
    theirs.eval(); ours.eval()
    sd_t, sd_o = theirs.state_dict(), ours.state_dict()
    assert list(sd_t.keys()) == list(sd_o.keys())
    assert [tuple(v.shape) for v in sd_t.values()] == [tuple(v.shape) for v in sd_o.values()]
    assert [n for n, _ in theirs.named_parameters()] == [n for n, _ in ours.named_parameters()]   # exchange order
    ours.load_state_dict(sd_t)                       # a reference checkpoint loads here ...
    with torch.no_grad():
        a, b = theirs(x), ours(x)
    assert torch.allclose(a, b, atol=1e-5, rtol=1e-5), float((a - b).abs().max())
    g = torch.Generator().manual_seed(9)
    for v in sd_o.values():
        if v.dtype.is_floating_point:
            v.copy_(torch.randn(v.shape, generator=g) * 0.05 + (1.0 if v.dim() == 1 else 0.0))
    sd_o = {k: (v.abs() + 0.5 if k.endswith("running_var") else v) for k, v in sd_o.items()}
    ours.load_state_dict(sd_o)
    theirs.load_state_dict(sd_o)                     # ... and one written here loads into the reference
    with torch.no_grad():
        a, b = theirs(x), ours(x)
    assert torch.allclose(a, b, atol=1e-4, rtol=1e-4), float((a - b).abs().max())


def test_lenet_is_checkpoint_compatible_with_the_reference(ref):
    from ewdml.models import build_model
    torch.manual_seed(0)
    _load_both_ways(ref["model_ops.lenet"].LeNet(), build_model("LeNet", 10), torch.randn(4, 1, 28, 28))


def test_vgg11_bn_is_checkpoint_compatible_with_the_reference(ref):
    from ewdml.models import build_model
    torch.manual_seed(0)
    _load_both_ways(ref["model_ops.vgg"].vgg11_bn(10), build_model("VGG11", 10), torch.randn(2, 3, 32, 32))


@pytest.mark.parametrize("name", ["ResNet18", "ResNet50"])
def test_resnets_are_checkpoint_compatible_with_the_reference(ref, name):
    from ewdml.models import build_model
    torch.manual_seed(0)
    mk = getattr(ref["model_ops.resnet"], name)
    theirs = mk(10) if name in ("ResNet18", "ResNet34") else mk()
    _load_both_ways(theirs, build_model(name, 10), torch.randn(2, 3, 32, 32))


@pytest.mark.parametrize("kw", [dict(lr=0.1, momentum=0.9), dict(lr=0.05, momentum=0.9, nesterov=True), dict(lr=0.1, momentum=0.0),
                                dict(lr=0.1, momentum=0.9, weight_decay=1e-3), dict(lr=0.1, momentum=0.9, dampening=0.1)])
def test_sgd_with_external_gradients_agrees_with_the_reference_optimizer(ref, kw):
    """optim/sgd.py::SGD.step(grads): the gradient list comes from the exchange, not from .grad; first-step buffer = d_p."""
    from ewdml.optim.sgd import SGD
    g = torch.Generator().manual_seed(1)
    p_ref = [torch.randn(5, 3, generator=g).requires_grad_(), torch.randn(7, generator=g).requires_grad_()]
    p_own = [p.detach().clone().requires_grad_() for p in p_ref]
    theirs, ours = ref["optim.sgd"].SGD(p_ref, **kw), SGD(p_own, **kw)
    for _ in range(4):
        grads = [torch.randn(p.shape, generator=g) for p in p_ref]
        theirs.step([x.clone() for x in grads])
        ours.step([x.clone() for x in grads])
    for a, b in zip(p_ref, p_own):
        assert torch.equal(a.detach(), b.detach())


def test_accuracy_agrees_with_the_reference_metric(ref):
    from ewdml.utils.metrics import accuracy
    g = torch.Generator().manual_seed(2)
    out, tgt = torch.randn(64, 10, generator=g), torch.randint(0, 10, (64,), generator=g)
    assert torch.equal(ref["nn_ops"].accuracy(out, tgt, (1,))[0], accuracy(out, tgt, (1,))[0])
    # prec@5 in the reference goes through .view() on a transposed tensor, which current torch rejects; here it is computed
    with pytest.raises(RuntimeError):
        ref["nn_ops"].accuracy(out, tgt, (1, 5))
    p1, p5 = accuracy(out, tgt, (1, 5))
    top5 = out.topk(5, dim=1).indices
    assert float(p5) == pytest.approx(100.0 * float((top5 == tgt[:, None]).any(1).float().mean()))
    assert float(p1) <= float(p5)


@pytest.mark.parametrize("kw", [dict(lr=1e-2), dict(lr=1e-2, amsgrad=True), dict(lr=1e-2, weight_decay=1e-2)])
def test_adam_with_numpy_gradients_agrees_with_the_reference_optimizer(ref, kw):
    """optim/adam.py::Adam.step(grads) takes NUMPY gradients (the server's receive buffers); the flat device-side form
    (FlatAdam, what the native exchange applies) follows the same recurrence."""
    from ewdml.optim.adam import Adam, FlatAdam
    g = torch.Generator().manual_seed(4)
    p_ref = [torch.randn(5, 3, generator=g).requires_grad_(), torch.randn(7, generator=g).requires_grad_()]
    p_own = [p.detach().clone().requires_grad_() for p in p_ref]
    theirs, ours = ref["optim.adam"].Adam(p_ref, **kw), Adam(p_own, **kw)
    flat = None if kw.get("amsgrad") else FlatAdam(torch.cat([p.detach().flatten() for p in p_ref]).clone(), **kw)
    for _ in range(5):
        grads = [torch.randn(p.shape, generator=g) for p in p_ref]
        theirs.step([x.numpy().copy() for x in grads])
        ours.step([x.numpy().copy() for x in grads])
        if flat is not None:
            flat.step(torch.cat([x.flatten() for x in grads]))
    for a, b in zip(p_ref, p_own):
        assert torch.equal(a.detach(), b.detach())
    if flat is not None:
        assert torch.allclose(flat.param, torch.cat([p.detach().flatten() for p in p_ref]), atol=1e-6, rtol=0)


@pytest.mark.parametrize("ctor", ["vgg11", "vgg13", "vgg13_bn", "vgg16", "vgg16_bn", "vgg19", "vgg19_bn"])
def test_other_vgg_variants_are_checkpoint_compatible_with_the_reference(ref, ctor):
    import ewdml.models as M
    torch.manual_seed(0)
    _load_both_ways(getattr(ref["model_ops.vgg"], ctor)(), getattr(M, ctor)(10), torch.randn(2, 3, 32, 32))


@pytest.mark.parametrize("ctor", ["ResNet34", "ResNet101", "ResNet152"])
def test_deeper_resnets_are_checkpoint_compatible_with_the_reference(ref, ctor):
    import ewdml.models as M
    torch.manual_seed(0)
    mk = getattr(ref["model_ops.resnet"], ctor)
    theirs = mk(10) if ctor == "ResNet34" else mk()
    _load_both_ways(theirs, getattr(M, ctor)(10), torch.randn(2, 3, 32, 32))


def test_horovod_flavour_qsgd_agrees_with_the_reference_module(ref, capsys):
    """horovod_compression.py needs ``horovod.torch`` only for the ``Compressor`` base class: with a two-line stand-in for
    that absent package the reference module itself is the oracle for the static compress / decompress pair."""
    import importlib
    import types
    from ewdml.compress import Compression
    hv, hvt, hvc = types.ModuleType("horovod"), types.ModuleType("horovod.torch"), types.ModuleType("horovod.torch.compression")
    hvc.Compressor = type("Compressor", (), {})
    hv.torch, hvt.compression = hvt, hvc
    stand_ins = {"horovod": hv, "horovod.torch": hvt, "horovod.torch.compression": hvc}
    saved = {n: sys.modules.get(n) for n in list(stand_ins) + ["horovod_compression"]}
    sys.modules.update(stand_ins)
    sys.path.insert(0, REF)
    try:
        if not _HAVE_TV:                                        # imported (unused) by the reference module
            pytest.skip("torchvision not importable")
        mod = importlib.import_module("horovod_compression")
        assert os.path.abspath(mod.__file__).startswith(REF + os.sep)
        t = torch.randn(7, 33, generator=torch.Generator().manual_seed(5)) * 0.01
        torch.manual_seed(21)
        levels_ref, norm_ref = mod.QSGDCompressor.compress(t)
        torch.manual_seed(21)
        levels, norm = Compression.qsgd.compress(t)
        assert torch.equal(norm.flatten().float(), norm_ref.flatten())
        assert torch.equal(levels.float(), levels_ref)
        assert torch.equal(Compression.qsgd.decompress(levels, norm), mod.QSGDCompressor.decompress(levels_ref, norm_ref))
        # interchangeable across the wire: each side decodes what the other produced
        assert torch.equal(Compression.qsgd.decompress(levels_ref.to(levels.dtype), norm_ref), mod.QSGDCompressor.decompress(levels.float(), norm))
    finally:
        sys.path.remove(REF)
        for n, m in saved.items():
            if m is None:
                sys.modules.pop(n, None)
            else:
                sys.modules[n] = m
    capsys.readouterr()                                          # the reference prints the norm on every call
