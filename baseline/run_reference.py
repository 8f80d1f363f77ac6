"""``bench.py --impl reference``: time the unmodified reference (baseline/_ref) on this box.

Public API of the reference = its ``distributed_nn.py`` command line under ``torch.distributed``
with the gloo backend, rank 0 = parameter server, ranks 1..N = workers (``src/run_pytorch_single.sh``).
"N GPUs" therefore means N workers + 1 server process.  The reference's ``--enable-gpu`` path is
tried first (GPU compute, gradients staged through the host, gloo exchange); in the checked-in
snapshot that path raises in ``optimizer.step`` (CPU gradient buffers applied to CUDA parameters,
``distributed_worker.py:225`` with ``optim/sgd.py:79-91``), in which case the reference's native
CPU mode -- the only mode its authors ran (``Report.zip!main.tex:143``) -- is timed and the JSON
line says so.  Same metric, model, batch and step counts as the ewdml arm.
"""
from __future__ import annotations

import json
import os
import socket
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _run(a, enable_gpu: bool, steps: int, timeout: float):
    n_workers = a.gpus
    world = n_workers + 1
    out = tempfile.mkdtemp(prefix="ewdml_ref_")
    port = _free_port()
    dataset = "MNIST" if a.model.lower() == "lenet" else "Cifar10"
    net = {"vgg11": "VGG11", "resnet50": "ResNet50", "lenet": "LeNet", "resnet18": "ResNet18"}[a.model.lower()]
    procs = []
    for r in range(world):
        # under torchrun this process carries the elastic agent's variables; inherited by the reference's own ranks they
        # redirect its env:// rendezvous to the agent's store (TORCHELASTIC_USE_AGENT_STORE) and the job never forms
        base = {k: v for k, v in os.environ.items()
                if not (k.startswith("TORCHELASTIC_") or k.startswith("TORCH_NCCL_") or k in
                        ("LOCAL_RANK", "GROUP_RANK", "ROLE_RANK", "ROLE_NAME", "LOCAL_WORLD_SIZE", "GROUP_WORLD_SIZE",
                         "ROLE_WORLD_SIZE", "TORCHELASTIC_RUN_ID", "NCCL_ASYNC_ERROR_HANDLING"))}
        env = dict(base, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   EWDML_REF_OUT=out, OMP_NUM_THREADS=str(max(1, (os.cpu_count() or 8) // world)),
                   EWDML_REF_DATASET_SIZE=str(max(1024, a.batch_size * 8)))
        if enable_gpu:
            # the reference uses torch.device("cuda") = device 0 of the process: give each worker its GPU
            env["CUDA_VISIBLE_DEVICES"] = str((r - 1) % max(1, a.gpus)) if r > 0 else "0"
        cmd = [sys.executable, os.path.join(HERE, "ref_shim.py"), f"--network={net}", f"--dataset={dataset}",
               f"--batch-size={a.batch_size}", f"--max-steps={steps}", "--epochs=100000", "--lr=0.01", "--momentum=0.9",
               "--eval-freq=1000000", f"--train-dir={out}/", "--compress-grad=compress",
               f"--enable-gpu={'1' if enable_gpu else ''}"]
        procs.append(subprocess.Popen(cmd, env=env, cwd=out, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True))
    ok, err_tail = True, ""
    for p in procs:
        try:
            _, err = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            ok = False
            err_tail = "timeout"
            break
        if p.returncode != 0:
            ok = False
            err_tail = (err or "")[-600:]
            break
    if not ok:
        for p in procs:
            if p.poll() is None:
                p.kill()
        return None, err_tail
    res = []
    for r in range(1, world):
        with open(os.path.join(out, f"rank{r}.json")) as f:
            res.append(json.load(f))
    return res, ""


def main(a, ClockSampler, metric_name):
    if not os.path.isdir(REF) or not os.path.isfile(os.path.join(REF, "distributed_nn.py")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref missing: run baseline/install_reference.sh "
                          "(pip cannot install /root/reference directly: no setup.py/pyproject.toml)"}))
        return
    if int(os.environ.get("RANK", "0")) != 0:
        return                      # under torchrun only rank 0 drives the reference's own launcher
    W, K = max(3, a.warmup), a.steps
    total = W + K
    import time
    sampler = ClockSampler(0).start()
    t_wall0 = time.time()
    mode, note = "cpu", ""
    res = None
    try:
        import torch
        has_gpu = torch.cuda.is_available() and torch.cuda.device_count() >= a.gpus
    except Exception:
        has_gpu = False
    if has_gpu:
        res, err = _run(a, True, total, timeout=1800)
        if res is None:
            note = "reference --enable-gpu path failed (" + err.strip().splitlines()[-1][:200] + "); timed its CPU mode" if err.strip() else "reference --enable-gpu path failed; timed its CPU mode"
        else:
            mode = "gpu-compute + host-staged gloo exchange"
    if res is None:
        res, err = _run(a, False, total, timeout=3600)
        if res is None:
            sampler.stop()
            print(json.dumps({"impl": "reference", "unavailable": "reference run failed: " + err.strip()[-300:]}))
            return
    # the reference logs at the END of step s; K steps after W warm-up steps = stamp[W+K-1] - stamp[W-1]
    elapsed = 0.0
    for r in res:
        st = {int(k): v for k, v in r["stamps"].items()}
        elapsed = max(elapsed, st[W + K - 1] - st[W - 1])
    value = a.batch_size * a.gpus * K / elapsed
    sampler.window("whole_reference_run", t_wall0, time.time())
    clocks = sampler.stop()
    # bytes that cross the host/device boundary per step in the reference's own loop: its --enable-gpu mode copies the
    # batch up and every gradient down and back (distributed_worker.py:188,322-323); its CPU mode has no device at all
    p_bytes = {"vgg11": 9756426, "resnet50": 23520842, "lenet": 431080, "resnet18": 11173962}.get(a.model.lower(), 0) * 4
    img_bytes = a.batch_size * (28 * 28 if a.model.lower() == "lenet" else 3 * 32 * 32) * 4
    gpu_mode = mode != "cpu"
    print(json.dumps({
        "impl": "reference", "metric": metric_name(a),
        "value": value, "unit": "images/s", "n_gpus": a.gpus, "steps": K, "warmup": W, "ms_per_step": elapsed / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
        "reference_class": "gpu" if gpu_mode else "cpu",
        "data": "synthetic (%s shape, random-init weights)" % ("MNIST" if a.model.lower() in ("lenet", "mnistnet") else "CIFAR-10"), "mode": mode, "note": note,
        "config": {"model": a.model, "global_batch": a.batch_size * a.gpus, "per_gpu_batch": a.batch_size,
                   "parallelism": f"ps: 1 server + {a.gpus} workers (gloo)", "compressor": "none (compression calls are commented out in the reference snapshot)"},
        "timed_with": "host clock on the reference's own per-step log line, max over workers",
        # the reference's loop IS its public API end to end: every timed step draws a batch from its DataLoader (host
        # memory), moves it to the compute device, exchanges through gloo and computes loss + top-1/top-5 on the host, so the
        # end-to-end figure is the same measurement (there is no separate device-resident variant to subtract)
        "e2e": {"value": value, "unit": "images/s",
                "h2d_bytes_per_step": (img_bytes + p_bytes) if gpu_mode else 0,
                "d2h_bytes_per_step": p_bytes if gpu_mode else 0,
                "ms_per_step": elapsed / K * 1e3,
                "note": "same timed region as value: the reference has only an end-to-end loop (DataLoader batch -> model -> "
                        "gloo push/pull -> host-side loss and accuracy every step)" + ("" if gpu_mode else "; cpu mode: no device copies")},
        "clocks": clocks,
        "gpu_launches": 0}), flush=True)
