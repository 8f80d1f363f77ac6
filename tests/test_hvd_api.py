"""Horovod-style API over gloo, the native launcher and the info tool."""
import json
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_allreduce_mnist_under_native_launcher(tmp_path):
    ck = str(tmp_path / "checkpoint-{epoch}.pt")
    cmd = [sys.executable, "-m", "ewdml.cli.launch", "-np", "2", "--master-port", "29741", "--",
           sys.executable, "-m", "ewdml.cli.allreduce_mnist", "--epochs", "1", "--max-steps", "6", "--no-cuda",
           "--synthetic-size", "256", "--batch-size", "16", "--test-batch-size", "64", "--compression", "qsgd",
           "--warmup-epochs", "0.5", "--checkpoint-format", ck]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-1500:]
    assert "Test set: Average loss" in out.stdout and os.path.isfile(ck.format(epoch=1))


def test_launcher_tears_job_down_on_failure():
    code = "import os,sys,time; r=int(os.environ['RANK']); time.sleep(0.3 if r==1 else 60); sys.exit(5 if r==1 else 0)"
    out = subprocess.run([sys.executable, "-m", "ewdml.cli.launch", "-np", "3", "--grace-seconds", "1", "--",
                          sys.executable, "-c", code], capture_output=True, text=True, timeout=60, cwd=ROOT)
    assert out.returncode == 5 and "rank 1 exited with status 5" in out.stderr


def test_launcher_passes_a_termination_signal_on_to_the_ranks(tmp_path):
    """mpirun forwards SIGINT / SIGTERM to the job (orterun); here every rank has its own process group, so the launcher
    has to do it: SIGTERM to the launcher ends the ranks AND what they started, and the exit status says which signal."""
    import signal
    import time
    code = ("import os, subprocess, sys, time\n"
            "child = subprocess.Popen([sys.executable, '-c', 'import time; time.sleep(120)'])\n"
            "open(os.path.join(sys.argv[1], 'pids%s' % os.environ['RANK']), 'w').write('%d %d' % (os.getpid(), child.pid))\n"
            "time.sleep(120)\n")
    proc = subprocess.Popen([sys.executable, "-m", "ewdml.cli.launch", "-np", "2", "--grace-seconds", "1", "--",
                             sys.executable, "-c", code, str(tmp_path)], stderr=subprocess.PIPE, text=True, cwd=ROOT)
    deadline = time.time() + 60
    while time.time() < deadline and not all((tmp_path / f"pids{r}").exists() and (tmp_path / f"pids{r}").read_text().count(" ") == 1
                                             for r in range(2)):
        time.sleep(0.1)
    pids = [int(x) for r in range(2) for x in (tmp_path / f"pids{r}").read_text().split()]
    assert len(pids) == 4
    proc.send_signal(signal.SIGTERM)
    _, err = proc.communicate(timeout=30)
    assert proc.returncode == 128 + signal.SIGTERM, (proc.returncode, err)
    assert "caught signal 15" in err
    time.sleep(0.3)
    for pid in pids:                       # ranks and their own children are gone (a zombie of a foreign parent counts as gone)
        try:
            with open(f"/proc/{pid}/stat") as f:
                assert f.read().split(") ")[1][0] == "Z", pid
        except FileNotFoundError:
            pass


def test_launcher_leaves_the_callers_other_children_alone():
    """The watchdog waits for its own ranks only: a child the calling process started before is still there to be waited for."""
    from ewdml.runtime import lib
    other = subprocess.Popen([sys.executable, "-c", "import sys; sys.exit(7)"])
    import time
    time.sleep(1.0)                         # it has exited and is waiting to be reaped
    status, failed = lib.launch(2, [sys.executable, "-c", "pass"], "127.0.0.1", 29500, 1.0)
    assert (status, failed) == (0, -1)
    assert other.wait(timeout=10) == 7


def test_launcher_restarts_a_failed_job_from_its_last_checkpoint(tmp_path):
    """--max-restarts / --checkpoint-dir: a worker dies at step 5 (test hook), the watchdog tears the job down, the launcher
    starts it again with --resume <newest checkpoint> and the job finishes its 8 steps with exit status 0."""
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "ewdml.cli.launch", "-np", "3", "--master-port", str(port), "--grace-seconds", "2",
           "--max-restarts", "1", "--checkpoint-dir", str(tmp_path), "--",
           sys.executable, os.path.join(ROOT, "distributed_nn.py"), "--network", "LeNet", "--dataset", "MNIST", "--data", "synthetic",
           "--synthetic-size", "128", "--batch-size", "16", "--max-steps", "8", "--epochs", "10", "--eval-freq", "2",
           f"--train-dir={tmp_path}/", "--method", "3", "--fault-inject-step", "5", "--fault-inject-rank", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "rank 1 exited with status 17" in out.stderr
    assert f"restart 1/1: resuming from {tmp_path}/model_step_" in out.stderr
    assert "resumed from" in out.stderr and "at step 5" in out.stderr
    from ewdml.cli.launch import latest_checkpoint, with_resume
    assert latest_checkpoint(str(tmp_path)) == os.path.join(str(tmp_path), "model_step_")
    assert with_resume(["x.py", "--resume", "old", "--lr", "1", "--resume=older"], "new") == ["x.py", "--lr", "1", "--resume", "new"]


def test_info_tool_and_adasum_pair():
    out = subprocess.run([sys.executable, "-m", "ewdml.cli.info"], capture_output=True, text=True, timeout=120, cwd=ROOT)
    info = json.loads(out.stdout)
    assert "spin_limit" in info["tunables"] and "coll_allreduce_algo" in info["tunables"] and info["flags"]["network"] == "LeNet"
    from ewdml.parallel.hvd import _adasum_pair
    a = torch.tensor([1.0, 0.0]); b = torch.tensor([0.0, 1.0])
    assert torch.allclose(_adasum_pair(a, b), a + b)                 # orthogonal -> sum
    assert torch.allclose(_adasum_pair(a, a), a)                     # parallel   -> average


def test_keras_style_trainer_with_callbacks(tmp_path):
    """tensorflow_mnist.py's capability: Adam + DistributedOptimizer, broadcast / metric-average / LR warm-up callbacks,
    rank-0 checkpoint per epoch."""
    ck = str(tmp_path / "checkpoint-{epoch}.pt")
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()     # a port that is free right now
    cmd = [sys.executable, "-m", "ewdml.cli.launch", "-np", "2", "--master-port", str(port), "--",
           sys.executable, "-m", "ewdml.cli.allreduce_mnist_keras", "--epochs", "2", "--steps-per-epoch", "6", "--no-cuda",
           "--synthetic-size", "256", "--batch-size", "16", "--warmup-epochs", "1", "--checkpoint-format", ck]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-1500:]
    assert "Epoch 2/2 - loss:" in out.stdout and "finished gradual learning rate warmup" in out.stdout
    assert os.path.isfile(ck.format(epoch=1)) and os.path.isfile(ck.format(epoch=2))
    sd = torch.load(ck.format(epoch=2))
    assert sd["fc1.weight"].shape == (128, 64 * 12 * 12)


def test_lr_warmup_multiplier_matches_horovod_formula():
    from ewdml.parallel import hvd_keras as K
    cb = K.LearningRateWarmupCallback(initial_lr=0.008, warmup_epochs=4)
    assert cb.multiplier(0.0) == 1.0 and cb.multiplier(10) == 1.0        # size() == 1 here: no scaling to ramp


def test_launcher_multi_node_rank_numbering():
    """One launcher per node: --nnodes 2 --node-rank 1 -np 2 starts global ranks 2 and 3 of a world of 4."""
    code = "import os; print('R', os.environ['RANK'], os.environ['LOCAL_RANK'], os.environ['WORLD_SIZE'], os.environ['GROUP_RANK'])"
    out = subprocess.run([sys.executable, "-m", "ewdml.cli.launch", "-np", "2", "--nnodes", "2", "--node-rank", "1", "--",
                          sys.executable, "-c", code], capture_output=True, text=True, timeout=60, cwd=ROOT)
    assert out.returncode == 0, out.stderr
    assert sorted(l for l in out.stdout.splitlines() if l.startswith("R")) == ["R 2 0 4 1", "R 3 1 4 1"]
    bad = subprocess.run([sys.executable, "-m", "ewdml.cli.launch", "-np", "1", "--nnodes", "2", "--node-rank", "5", "--",
                          sys.executable, "-c", "pass"], capture_output=True, text=True, timeout=60, cwd=ROOT)
    assert bad.returncode != 0


def _hvd_rank(rank, world, port, out_dir):
    import torch.distributed as dist
    from ewdml.compress import Compression
    from ewdml.parallel import hvd
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    hvd.init("gloo")
    res = {"rank": hvd.rank(), "size": hvd.size()}
    t = torch.full((5,), float(rank + 1))
    res["avg"] = hvd.allreduce(t).tolist()
    res["sum"] = hvd.allreduce(t, op=hvd.Sum).tolist()
    res["adasum"] = hvd.allreduce(torch.eye(2)[rank % 2], op=hvd.Adasum).tolist()       # orthogonal vectors add up
    torch.manual_seed(rank)                                                            # different initial weights per rank
    model = torch.nn.Linear(4, 3)
    opt = torch.optim.SGD(model.parameters(), lr=0.1 * (rank + 1), momentum=0.9)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    hvd.broadcast_optimizer_state(opt, root_rank=0)
    res["lr"] = opt.param_groups[0]["lr"]
    dopt = hvd.DistributedOptimizer(opt, named_parameters=model.named_parameters(), compression=Compression.none)
    x = torch.full((2, 4), float(rank + 1))
    model(x).sum().backward()
    local = model.weight.grad.clone()
    dopt.synchronize()
    res["grad_is_average"] = bool(torch.allclose(model.weight.grad, hvd.allreduce(local)))
    dopt.step()
    res["weight_sum"] = float(model.weight.double().sum())
    json.dump(res, open(os.path.join(out_dir, f"r{rank}.json"), "w"))
    hvd.shutdown()


def test_hvd_collectives_and_distributed_optimizer(tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_hvd_rank, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (json.load(open(tmp_path / f"r{r}.json")) for r in (0, 1))
    assert (r0["rank"], r1["rank"], r0["size"]) == (0, 1, 2)
    assert r0["avg"] == [1.5] * 5 and r1["sum"] == [3.0] * 5 and r0["adasum"] == [1.0, 1.0]
    assert r1["lr"] == 0.1                                   # optimizer state came from rank 0
    assert r0["grad_is_average"] and r1["grad_is_average"]
    assert r0["weight_sum"] == r1["weight_sum"]              # same start, same averaged gradient: replicas stay identical
