#!/usr/bin/env bash
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -q > gpurun_out/pytest_multigpu${N}.log 2>&1; echo "multigpu rc=$?"; tail -3 gpurun_out/pytest_multigpu${N}.log | cut -c1-300
for impl in ewdml nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --impl $impl --gpus $N --steps 100 --warmup 5 > gpurun_out/bench_${N}gpu_$impl.log 2>&1; echo "bench $impl rc=$?"
tail -1 gpurun_out/bench_${N}gpu_$impl.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('impl','value','ms_per_step','n_gpus','gpu_launches','device_error')}, d['e2e']['value'], d.get('exchange'))" || tail -8 gpurun_out/bench_${N}gpu_$impl.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus $N --steps 100 --warmup 5 --model ResNet50 > gpurun_out/bench_${N}gpu_r50.log 2>&1; echo "r50 rc=$?"
tail -1 gpurun_out/bench_${N}gpu_r50.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d.get('exchange'))" || tail -8 gpurun_out/bench_${N}gpu_r50.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29536 bench.py --gpus $N --steps 100 --warmup 5 --model ResNet50 --exchange allreduce --compressor none > gpurun_out/bench_${N}gpu_r50_ar.log 2>&1; echo "r50 allreduce rc=$?"
tail -1 gpurun_out/bench_${N}gpu_r50_ar.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d.get('exchange'))" || tail -8 gpurun_out/bench_${N}gpu_r50_ar.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29537 bench.py --gpus $N --steps 100 --warmup 5 --model ResNet50 --exchange ps > gpurun_out/bench_${N}gpu_r50_ps.log 2>&1; echo "r50 ps rc=$?"
tail -1 gpurun_out/bench_${N}gpu_r50_ps.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d.get('exchange'))" || tail -8 gpurun_out/bench_${N}gpu_r50_ps.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench/exchange_sweep.py --quick > gpurun_out/sweep_quick_${N}gpu.log 2>&1; echo "sweep rc=$?"
grep gradient_mb gpurun_out/sweep_quick_${N}gpu.log | cut -c1-330 || tail -5 gpurun_out/sweep_quick_${N}gpu.log
