"""Accuracy-after-N-steps vs bytes-per-step for the report's six methods ("bytes sent per
step <= the reference's compression ratio at equal accuracy-after-N-steps on synthetic data").

Runs the parameter server (1 server + 2 workers, gloo, CPU -- BASELINE.json config 1) on the learnable synthetic
MNIST-shaped task for the same number of steps per method and reports, per method: mean training top-1 over the last
20 steps, final loss, exact payload bytes a worker sends and receives per step (averaged over the run, so method 6's
19 silent steps out of 20 count), and the ratio to the dense exchange.  Top-k ratio 0.4 is the reference's slide value.

    python bench/method_table.py [--steps 300] [--network LeNet] > profiles/method_table_cpu.jsonl
"""
import argparse
import json
import os
import socket
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _run(rank, world, port, kw, out_dir):
    from ewdml.config import TrainConfig
    from ewdml.data import build_datasets, prepare_data
    from ewdml.train.roles import DistributedWorker, SyncReplicasMaster_NN
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = TrainConfig(**kw).validate()
    if rank == 0:
        SyncReplicasMaster_NN(cfg, world).build_model(cfg.num_classes()).start(total_steps=cfg.max_steps)
    else:
        loader, test = prepare_data(cfg, shard_rank=rank - 1, num_shards=world - 1)
        w = DistributedWorker(cfg, rank, world).build_model(cfg.num_classes())
        hist = w.train(loader, test, total_steps=cfg.max_steps)
        with open(os.path.join(out_dir, f"r{rank}.json"), "w") as f:
            json.dump({"prec1": [h["prec1"] for h in hist], "loss": [h["loss"] for h in hist],
                       "sent": [h["bytes_sent"] for h in hist], "recv": [h["bytes_recv"] for h in hist],
                       "dense": w.flat.dense_nbytes()}, f)
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--network", default="LeNet")
    ap.add_argument("--topk-ratio", type=float, default=0.4)
    ap.add_argument("--error-feedback", action="store_true")
    a = ap.parse_args()
    dataset = "MNIST" if a.network in ("LeNet", "MnistNet") else "Cifar10"
    for method in (1, 2, 3, 4, 5, 6):
        out = tempfile.mkdtemp(prefix="ewdml_mt_")
        kw = dict(network=a.network, dataset=dataset, batch_size=64, max_steps=a.steps, epochs=10 ** 6, lr=0.01, momentum=0.9,
                  synthetic_size=8192, data="synthetic", eval_freq=10 ** 9, train_dir=out + "/", log_interval=10 ** 9,
                  method=method, topk_ratio=a.topk_ratio, error_feedback=a.error_feedback and method >= 5)
        mp.spawn(_run, args=(3, _free_port(), kw, out), nprocs=3, join=True)
        r = json.load(open(os.path.join(out, "r1.json")))
        n = len(r["prec1"])
        per_step = (sum(r["sent"]) + sum(r["recv"])) / n
        win = lambda k: sum(r["prec1"][max(0, k - 10):k]) / max(1, min(10, k))
        curve = {str(k): round(win(k), 1) for k in (10, 20, 40, 80, 160, 320) if k <= n}
        print(json.dumps({"method": method, "network": a.network, "steps": n, "top1_last20": sum(r["prec1"][-20:]) / 20,
                          "top1_mean10_at_step": curve,
                          "loss_last20": sum(r["loss"][-20:]) / 20, "bytes_sent_per_step": sum(r["sent"]) / n,
                          "bytes_recv_per_step": sum(r["recv"]) / n, "dense_both_ways": 2 * r["dense"],
                          "compression_vs_method1": 2 * r["dense"] / per_step, "topk_ratio": a.topk_ratio,
                          "error_feedback": bool(kw["error_feedback"])}), flush=True)


if __name__ == "__main__":
    main()
