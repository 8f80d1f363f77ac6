#!/usr/bin/env bash
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu.py -q -x --durations=4 > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu rc=$?"; tail -8 gpurun_out/pytest_multigpu.log | cut -c1-300
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 50 --warmup 3 > gpurun_out/bench_${N}gpu.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/bench_${N}gpu.log | cut -c1-400
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 3 --warmup 3 --impl reference > gpurun_out/bench_ref_${N}gpu.log 2>&1
echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_${N}gpu.log | cut -c1-700
