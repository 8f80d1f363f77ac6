#!/usr/bin/env bash
# 2-GPU validation: multi-GPU tests (engine, overlap, PS), default bench, PS bench (graph), NCCL arm, reference arm
N=${1:-2}
mkdir -p gpurun_out
run() { # name, timeout, args...
  local name=$1 to=$2; shift 2
  timeout $to python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N "$@" > gpurun_out/${name}_${N}gpu.log 2>&1
  echo "$name rc=$?"; tail -1 gpurun_out/${name}_${N}gpu.log | cut -c1-1500
}
timeout 420 python -m pytest tests/test_multigpu.py -q -x --durations=4 > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu rc=$?"; tail -12 gpurun_out/pytest_multigpu.log | cut -c1-300
run bench 100 --steps 100 --warmup 3
run bench_ps 100 --steps 50 --warmup 3 --exchange ps
run bench_overlap 100 --steps 100 --warmup 3 --bucket-size 8388608 --overlap
run bench_nccl 120 --steps 20 --warmup 3 --impl nccl
run bench_ref 150 --steps 3 --warmup 3 --impl reference
