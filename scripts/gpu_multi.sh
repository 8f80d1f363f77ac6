#!/usr/bin/env bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -q -x > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/bench_2gpu.log 2>&1; echo "bench2 rc=$?"
tail -30 gpurun_out/pytest_multigpu.log; tail -3 gpurun_out/bench_2gpu.log | cut -c1-1500
