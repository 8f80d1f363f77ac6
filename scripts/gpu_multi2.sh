#!/usr/bin/env bash
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -q -x > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu rc=$?"; tail -3 gpurun_out/pytest_multigpu.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/bench_${N}gpu.log 2>&1; echo "bench$N rc=$?"
tail -1 gpurun_out/bench_${N}gpu.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','n_gpus','gpu_launches','device_error')}, d['e2e']['value'], d['exchange'])" || tail -8 gpurun_out/bench_${N}gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench/exchange_sweep.py --quick > gpurun_out/sweep_quick_${N}gpu.log 2>&1; echo "sweep rc=$?"
grep gradient_mb gpurun_out/sweep_quick_${N}gpu.log | cut -c1-400 || tail -5 gpurun_out/sweep_quick_${N}gpu.log
