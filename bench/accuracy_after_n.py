#!/usr/bin/env python
"""Accuracy after N steps at equal step count, product path vs its own dense / fp32 runs (BASELINE.json: "bytes sent per step
<= the reference's compression ratio at equal accuracy-after-N-steps on synthetic data").

    python bench/accuracy_after_n.py --model VGG11 [--steps 1000 --eval-at 50 200 1000]
    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 bench/accuracy_after_n.py --model VGG11

Every arm trains the same architecture from the same initial weights on the same learnable synthetic CIFAR-shaped task
(class templates + noise, ewdml/data.py) with the same batches, learning rate and momentum through the public GpuTrainer
API, and is evaluated on the same held-out set at the same step counts:

    fp32-torch-dense      torch/cuDNN fp32 layers (TF32 off), uncompressed exchange -- the reference's precision
    bf16-native-dense     this framework's sm_100a kernels, bf16 compute over fp32 master weights, uncompressed
    bf16-native-qsgd      + QSGD s=128 (method 2/4 of the reference's report)
    bf16-native-topk-qsgd + top-k 1 % then QSGD (method 5, the bench.py configuration)
    ...-ef                the same with error feedback (what compression dropped is added to the next gradient)

One JSON line per arm on rank 0: accuracy at each N, training loss, payload bytes per peer per step, compression ratio.
"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="VGG11")
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--eval-at", type=int, nargs="+", default=[50, 200, 1000])
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--signal", type=float, default=0.3, help="template amplitude of the synthetic task (lower = harder)")
    ap.add_argument("--arms", nargs="+", default=["fp32-torch-dense", "bf16-native-dense", "bf16-native-qsgd", "bf16-native-topk-qsgd",
                                                  "bf16-native-topk-qsgd-ef"])
    a = ap.parse_args()
    import torch.distributed as dist
    from ewdml.config import TrainConfig
    from ewdml.data import SyntheticImageDataset
    from ewdml.ops import layers as Lyr
    from ewdml.parallel.peer import PeerGroup
    from ewdml.train.gpu import GpuTrainer
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    train = SyntheticImageDataset("Cifar10", 16384, train=True, seed=1234, signal=a.signal)
    test = SyntheticImageDataset("Cifar10", 2048, train=False, seed=1234, signal=a.signal)
    xt = test.data.to(dev).contiguous(memory_format=torch.channels_last)
    yt = test.targets.to(dev)
    pg = PeerGroup.from_env(dev)
    for arm in a.arms:
        compute = "torch" if "torch" in arm else "native"
        dtype = "fp32" if arm.startswith("fp32") else "bf16"
        comp = "topk-qsgd" if "topk-qsgd" in arm else ("qsgd" if "qsgd" in arm else "none")
        cfg = TrainConfig(network=a.model, dataset="Cifar10", batch_size=a.batch_size, lr=a.lr, momentum=0.9, compressor=comp,
                          topk_ratio=0.01, qsgd_levels=128, exchange="allgather", dtype=dtype, compute=compute, enable_gpu=True, seed=1,
                          error_feedback=arm.endswith("-ef")).validate()
        tr = GpuTrainer(cfg, dev, pg)
        g = torch.Generator().manual_seed(7)                      # the same batch order for every arm; ranks read disjoint shards
        res = {"arm": arm, "model": a.model, "world": world, "batch_per_gpu": a.batch_size, "lr": a.lr, "signal": a.signal, "acc": {},
               "loss": {}}
        st = None
        for step in range(1, a.steps + 1):
            idx = torch.randint(0, train.data.shape[0], (world, a.batch_size), generator=g)[rank]
            x = train.data[idx].to(dev, non_blocking=True).contiguous(memory_format=torch.channels_last)
            y = train.targets[idx].to(dev, non_blocking=True)
            loss, st_ = tr.step(x, y, want_stats=(step == a.steps))
            st = st_ if step == a.steps else st
            if step in a.eval_at:
                tr.model.eval()
                Lyr.set_compute("torch")
                with torch.no_grad():
                    correct = 0
                    for i in range(0, xt.shape[0], 512):
                        correct += int((tr.model(xt[i:i + 512].float()).float().argmax(1) == yt[i:i + 512]).sum())
                Lyr.set_compute(compute)
                tr.model.train()
                res["acc"][str(step)] = round(correct / xt.shape[0], 4)
                res["loss"][str(step)] = round(float(loss), 4)
        res["payload_bytes_per_peer_per_step"] = int(st.bytes_sent // max(1, world - 1)) if (st and world > 1) else int(tr.exchange._last_payload)
        res["dense_bytes"] = tr.flat.dense_nbytes()
        res["compression_ratio"] = round(res["dense_bytes"] / max(1, res["payload_bytes_per_peer_per_step"]), 1)
        res["device_error"] = tr.check_device_error()
        if rank == 0:
            print(json.dumps(res), flush=True)
        del tr
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    pg.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
