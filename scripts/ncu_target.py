"""Workload for one `ncu --set full` capture of the communication kernels at VGG-11 size (see scripts/gpu_round20.sh):
2 warm-up steps, one 64 MB single-piece allreduce of the collective engine (k_coll), then the step that gets profiled."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewdml.config import TrainConfig
from ewdml.parallel.collectives import CollectiveEngine
from ewdml.train.gpu import GpuTrainer

dev = torch.device("cuda", 0)
cfg = TrainConfig(network="VGG11", dataset="Cifar10", batch_size=128, lr=0.01, momentum=0.9, exchange="allgather",
                  compressor="topk-qsgd", topk_ratio=0.01, dtype="bf16", compute="native", enable_gpu=True).validate()
tr = GpuTrainer(cfg, dev)
x = torch.randn(128, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (128,), device=dev)
for _ in range(2):
    tr.step(x, y)
eng = CollectiveEngine.create(tr.pg, slot_payload=64 << 20)
t = torch.randn(16 << 20, device=dev)
eng.all_reduce(t, scale=0.5)
tr.step(x, y)
torch.cuda.synchronize()
print("ok", float(tr.last_loss), tr.check_device_error())
