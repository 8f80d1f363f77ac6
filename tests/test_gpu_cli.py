"""The command-line trainer on a GPU: ``distributed_nn.py --enable-gpu`` with the native kernels (exchange and layers) and
with the library-only NCCL arm, as a single-rank job (multi-rank runs: tests/test_multigpu.py, scripts/gpu_multi_*.sh)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("extra", [
    ["--network=LeNet", "--dataset=MNIST", "--backend=native", "--exchange=allgather", "--compressor=topk-qsgd", "--topk-ratio=0.1"],
    ["--network=VGG11", "--dataset=Cifar10", "--backend=native", "--exchange=allgather", "--compressor=topk-qsgd",
     "--topk-ratio=0.01", "--compute=native", "--dtype=bf16"],
    ["--network=VGG11", "--dataset=Cifar10", "--backend=native", "--exchange=allgather", "--compressor=topk-qsgd",
     "--topk-ratio=0.01", "--compute=native", "--dtype=bf16", "--graphs", "--max-steps=12"],
    ["--network=LeNet", "--dataset=MNIST", "--backend=nccl", "--exchange=allgather", "--compressor=topk-qsgd", "--topk-ratio=0.1"],
])
def test_cli_trainer_on_gpu(tmp_path, extra):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    env = dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT="29791")
    cmd = [sys.executable, os.path.join(ROOT, "distributed_nn.py"), "--enable-gpu=1", "--batch-size=32", "--max-steps=6", "--epochs=3",
           "--lr=0.02", "--momentum=0.9", "--eval-freq=3", "--data=synthetic", "--synthetic-size=256", "--log-interval=2",
           f"--train-dir={tmp_path}/", f"--metrics-file={tmp_path}/metrics.jsonl"] + extra
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=240, cwd=ROOT, env=env)
    assert out.returncode == 0, (out.stdout[-800:], out.stderr[-1500:])
    assert ("Now the next step is: 12" if "--graphs" in extra else "Now the next step is: 6") in out.stdout
    assert os.path.isfile(tmp_path / "model_step_")
    sd = torch.load(tmp_path / "model_step_", map_location="cpu", weights_only=True)
    assert all(torch.isfinite(v).all() for v in sd.values() if v.is_floating_point())
    if "--backend=native" in extra:
        import json
        recs = [json.loads(l) for l in open(tmp_path / "metrics.jsonl")]
        assert recs and any(r.get("bytes_sent", 0) == 0 for r in recs)        # one rank: nothing leaves the GPU
