"""World-size 2/3 gloo/CPU plumbing tests of the PS loop and the other exchange strategies
(BASELINE.json config 1; SURVEY.md section 4 implication (b))."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ewdml.config import TrainConfig
from ewdml.data import build_datasets, prepare_data
from ewdml.train.roles import DistributedWorker, SyncReplicasMaster_NN, planned_steps

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _one_job(rank, world, cfg_kwargs, out_dir):
    cfg_kwargs = dict(cfg_kwargs)
    factory = None
    if cfg_kwargs.pop("_factory", None) == "library":
        from ewdml.parallel.nccl_baseline import build_nccl_exchange as factory
    cfg = TrainConfig(**cfg_kwargs).validate()
    ps = cfg.exchange == "ps"
    nw = world - 1 if ps else world
    train_set, _ = build_datasets(cfg)
    total = planned_steps(cfg, len(train_set), nw)
    if ps and rank == 0:
        m = SyncReplicasMaster_NN(cfg, world, exchange_factory=factory).build_model(cfg.num_classes())
        stats = m.start(total_steps=total)
        res = {"role": "server", "steps": len(stats), "bytes_recv": [s.bytes_recv for s in stats],
               "used": getattr(m.exchange, "used_per_step", None), "dropped": getattr(m.exchange, "dropped", None),
               "staleness": getattr(m.exchange, "staleness", None), "checksum": float(m.flat.param.double().sum())}
    else:
        loader, test_loader = prepare_data(cfg, shard_rank=(rank - 1 if ps else rank), num_shards=nw)
        w = DistributedWorker(cfg, rank, world, exchange_factory=factory).build_model(cfg.num_classes())
        hist = w.train(loader, test_loader, total_steps=total)
        res = {"role": "worker", "loss": [h["loss"] for h in hist],
               "bytes_sent": [h["bytes_sent"] for h in hist],
               "checksum": float(w.flat.param.double().sum()), "dense": w.flat.dense_nbytes(),
               "codec_bytes": getattr(getattr(w.exchange, "push", None), "nbytes", None),
               "skipped": getattr(w.exchange, "skipped", None)}
    with open(os.path.join(out_dir, f"r{rank}.json"), "w") as f:
        json.dump(res, f)
    dist.barrier()


def _run(rank, world, port, cfg_kwargs, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    _one_job(rank, world, cfg_kwargs, out_dir)
    dist.destroy_process_group()


def _run_many(rank, world, port, jobs, out_root):
    """Several independent training jobs, one after the other, in ONE set of processes (process start-up is most of what a
    small job costs); every job builds its own model, data and exchange from its configuration, with a barrier in between."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    for name, cfg_kwargs in jobs:
        _one_job(rank, world, cfg_kwargs, os.path.join(out_root, name))
    dist.destroy_process_group()


_BASE = dict(network="LeNet", dataset="MNIST", batch_size=16, max_steps=8, epochs=2, lr=0.02,
             momentum=0.9, synthetic_size=256, data="synthetic", eval_freq=4, log_interval=100)


def _launch_many(world, root, jobs):
    """jobs: {name: configuration overrides}; returns {name: [result of rank 0, 1, ...]}, files under root/name/."""
    full = []
    for name, kw in jobs.items():
        os.makedirs(os.path.join(str(root), name), exist_ok=True)
        base = dict(_BASE, train_dir=os.path.join(str(root), name) + "/")
        base.update(kw)
        full.append((name, base))
    mp.spawn(_run_many, args=(world, _free_port(), full, str(root)), nprocs=world, join=True)
    return {name: [json.load(open(os.path.join(str(root), name, f"r{r}.json"))) for r in range(world)] for name, _ in full}


def _launch(world, tmp_path, **kw):
    base = dict(_BASE, train_dir=str(tmp_path) + "/")
    base.update(kw)
    mp.spawn(_run, args=(world, _free_port(), base, str(tmp_path)), nprocs=world, join=True)
    return [json.load(open(os.path.join(str(tmp_path), f"r{r}.json"))) for r in range(world)]


@pytest.fixture(scope="module")
def six_methods(tmp_path_factory):
    root = tmp_path_factory.mktemp("six_methods")
    jobs = {}
    for method in range(1, 7):
        extra = dict(topk_ratio=0.4) if method >= 5 else {}
        if method == 6:
            extra["sync_period"] = 4
        jobs[f"m{method}"] = dict(method=method, **extra)
    return root, _launch_many(3, root, jobs)


@pytest.mark.parametrize("method", [1, 2, 3, 4, 5, 6])
def test_ps_all_six_methods(six_methods, method):
    root, results = six_methods
    tmp_path = os.path.join(str(root), f"m{method}")
    server, w1, w2 = results[f"m{method}"]
    assert server["role"] == "server" and server["steps"] == 8
    for w in (w1, w2):
        assert len(w["loss"]) == 8 and all(l == l for l in w["loss"])       # no NaNs
    if method != 6:
        # synchronous methods keep the replicas identical
        assert w1["checksum"] == pytest.approx(w2["checksum"], rel=1e-9)
        assert all(b == w1["codec_bytes"] for b in w1["bytes_sent"])
    else:
        assert [b > 0 for b in w1["bytes_sent"]] == [True, False, False, False] * 2
    dense = w1["dense"]
    ratio = dense / max(w1["bytes_sent"])
    if method in (1, 3):
        assert ratio == pytest.approx(1.0)
    elif method in (2, 4):
        assert ratio > 3.5            # ~1.125 B/elem at s=128 on the host codec
    else:
        assert ratio > 1.5
    assert os.path.isfile(os.path.join(str(tmp_path), "model_step_"))


def test_ps_world_size_2_loss_decreases(tmp_path):
    res = _launch(2, tmp_path, method=3, max_steps=30, epochs=4, synthetic_size=512, batch_size=32, lr=0.05)
    losses = res[1]["loss"]
    assert sum(losses[-5:]) / 5 < sum(losses[:5]) / 5 - 0.1


_SERVERLESS = [("allgather", "topk-qsgd"), ("allgather", "qsgd"), ("allgather", "topk"), ("allreduce", "none")]


@pytest.fixture(scope="module")
def serverless(tmp_path_factory):
    root = tmp_path_factory.mktemp("serverless")
    return _launch_many(2, root, {f"{e}-{c}": dict(exchange=e, compressor=c, topk_ratio=0.1) for e, c in _SERVERLESS})


@pytest.mark.parametrize("exchange,compressor", _SERVERLESS)
def test_serverless_exchanges(serverless, exchange, compressor):
    res = serverless[f"{exchange}-{compressor}"]
    assert res[0]["checksum"] == pytest.approx(res[1]["checksum"], rel=1e-9)
    assert len(res[0]["loss"]) == 8


def test_cli_entrypoint_under_torchrun(tmp_path):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "distributed_nn.py"), "--network=LeNet", "--dataset=MNIST", "--batch-size=16",
           "--max-steps=3", "--epochs=1", "--method=5", "--topk-ratio=0.4", "--synthetic-size=128",
           "--data=synthetic", "--enable-gpu=", f"--train-dir={tmp_path}/"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=240, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "I am the master" in out.stdout and "I am worker: 1 in all 1 workers" in out.stdout


_LIBRARY = [("allgather", 0, "topk"), ("allreduce", 3, ""), ("ps", 1, ""), ("ps", 3, "")]


@pytest.fixture(scope="module")
def library_pairs(tmp_path_factory):
    root = tmp_path_factory.mktemp("library")
    jobs = {}
    for e, m, c in _LIBRARY:
        kw = dict(exchange=e, method=m, compressor=c, topk_ratio=0.25, max_steps=4)
        jobs[f"{e}-{m}-plumbing"] = kw
        jobs[f"{e}-{m}-library"] = dict(kw, _factory="library")
    return _launch_many(3, root, jobs)


@pytest.mark.parametrize("exchange,method,compressor", _LIBRARY)
def test_library_only_arm_matches_plumbing(library_pairs, exchange, method, compressor):
    """The torch-ops + collectives classes behind --backend nccl compute the same updates as the packed gloo path
    (deterministic compressors only; run here on CPU tensors over gloo)."""
    a = library_pairs[f"{exchange}-{method}-plumbing"]
    b = library_pairs[f"{exchange}-{method}-library"]
    wa = [r for r in a if r["role"] == "worker"]
    wb = [r for r in b if r["role"] == "worker"]
    assert wa[0]["checksum"] == pytest.approx(wb[0]["checksum"], rel=1e-6)
    assert wb[0]["checksum"] == pytest.approx(wb[1]["checksum"], rel=1e-9)
    assert wa[0]["loss"] == pytest.approx(wb[0]["loss"], rel=1e-4)


def test_bucketed_overlapped_allgather_equals_monolithic(tmp_path):
    """--bucket-size/--overlap change when the exchange runs, not what it computes (per-tensor norms and top-k,
    element-indexed random numbers)."""
    kw = dict(exchange="allgather", compressor="topk-qsgd", topk_ratio=0.1, max_steps=5)
    res = _launch_many(2, tmp_path, {"a": kw, "b": dict(kw, bucket_size=100000), "c": dict(kw, bucket_size=100000, overlap=True)})
    a, b, c = res["a"], res["b"], res["c"]
    for other in (b, c):
        assert other[0]["checksum"] == pytest.approx(a[0]["checksum"], rel=1e-12)
        assert other[0]["loss"] == pytest.approx(a[0]["loss"], rel=1e-9)
        assert other[0]["bytes_sent"] == a[0]["bytes_sent"]
    assert c[0]["checksum"] == pytest.approx(c[1]["checksum"], rel=1e-12)


def test_error_feedback_conserves_the_gradient_sum():
    """With error feedback, (everything transmitted so far) + residual == (sum of all gradients so far)."""
    from ewdml.models import build_model
    from ewdml.optim import build_optimizer
    from ewdml.parallel.flat import FlatModel
    from ewdml.parallel.host_exchange import AllgatherExchange
    cfg = TrainConfig(network="LeNet", exchange="allgather", compressor="topk-qsgd", topk_ratio=0.05, error_feedback=True,
                      lr=0.0, momentum=0.0, bucket_size=100000).validate()
    torch.manual_seed(0)
    flat = FlatModel(build_model("LeNet", 10))
    ex = AllgatherExchange(flat, cfg, build_optimizer(cfg, flat.params))
    assert len(ex.buckets) == 2 and ex.residual is not None
    total_grad = torch.zeros(flat.total)
    total_sent = torch.zeros(flat.total)
    for step in range(6):
        flat.grad.copy_(torch.randn(flat.total) * 0.01)
        for seg_a, seg_b in zip(flat.segments[:-1], flat.segments[1:]):       # alignment gaps carry no gradient
            flat.grad[seg_a.offset + seg_a.numel:seg_b.offset] = 0
        total_grad += flat.grad
        ex.step(step)
        total_sent += ex._avg          # world == 1: the average is exactly what this rank transmitted
    assert torch.allclose(total_sent + ex.residual, total_grad, atol=1e-5)
    # without feedback the same compressor loses most of the mass
    assert (total_sent - total_grad).norm() > 10 * (total_sent + ex.residual - total_grad).norm()


@pytest.mark.parametrize("mode,kw", [("normal", dict(num_aggregate=2)), ("kill", dict(num_aggregate=3, kill_threshold=0.1))])
def test_backup_workers_and_straggler_cutoff(tmp_path, mode, kw):
    """--num-aggregate < workers: the server closes each step with the first gradients to arrive; --mode kill: it stops
    waiting kill_threshold seconds after the first arrival.  The straggler's late gradients are dropped, every worker
    still applies the same pulls, so the replicas stay identical and nobody dead-locks."""
    res = _launch(4, tmp_path, method=3, mode=mode, straggler_rank=3, straggler_delay_s=0.25, max_steps=6, **kw)
    server, workers = res[0], res[1:]
    assert server["steps"] == 6
    if mode == "normal":        # need == 2: every step closes with exactly the first two arrivals, by construction
        assert all(u == 2 for u in server["used"][1:]), server["used"]      # (step 0 may see all three: start-up skew)
        assert server["dropped"] >= 4
    else:                       # time-based cut-off: robust to a loaded machine, the straggler is cut off at least once
        assert all(1 <= u <= 3 for u in server["used"]) and min(server["used"][1:]) <= 2, server["used"]
        assert server["dropped"] >= 1
    assert all(len(w["loss"]) == 6 for w in workers)
    assert workers[0]["checksum"] == pytest.approx(workers[1]["checksum"], rel=1e-9)
    assert workers[0]["checksum"] == pytest.approx(workers[2]["checksum"], rel=1e-9)


def test_asynchronous_parameter_server(tmp_path):
    """--mode async: gradients are applied as they arrive, each worker gets fresh weights back; a slow worker does not
    hold the others up (they finish long before it) and its gradients arrive stale."""
    res = _launch(4, tmp_path, method=1, mode="async", straggler_rank=3, straggler_delay_s=0.15, max_steps=8)
    server, workers = res[0], res[1:]
    assert server["steps"] == 24                                   # 3 workers x 8 steps, one server step per arrival
    assert max(server["staleness"]) >= 1 and min(server["staleness"]) == 0
    assert all(len(w["loss"]) == 8 and all(l == l for l in w["loss"]) for w in workers)
    assert workers[0]["loss"][-1] < workers[0]["loss"][0]


def test_parameter_server_job_resumes_from_its_checkpoint(tmp_path):
    """--resume on the distributed job: server and workers all pick the run up at the step recorded next to the checkpoint
    (the weights, and the step counter that drives how many exchanges the server expects)."""
    first = dict(method=3, max_steps=6, eval_freq=3, lr=0.05, synthetic_size=512, batch_size=32)
    res = _launch_many(3, tmp_path, {"first": first,
                                     "second": dict(first, max_steps=9, resume=os.path.join(str(tmp_path), "first", "model_step_"))})
    a, b = res["first"], res["second"]
    assert a[0]["steps"] == 6 and len(a[1]["loss"]) == 6
    # checkpoints after steps 0 and 3 -> the sidecar says "next step 4" -> steps 4..8 remain
    assert b[0]["steps"] == 5 and len(b[1]["loss"]) == 5 and len(b[2]["loss"]) == 5
    assert b[1]["checksum"] == pytest.approx(b[2]["checksum"], rel=1e-9)
    assert b[1]["loss"][0] < a[1]["loss"][0] - 0.05          # it did not start from fresh weights
