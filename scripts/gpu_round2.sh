#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 200 --warmup 5 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?" | tee -a gpurun_out/bench1.log
timeout 600 python bench.py --steps 50 --warmup 5 --model ResNet50 > gpurun_out/bench_r50.log 2>&1; echo "bench r50 rc=$?"
timeout 600 python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"
tail -8 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/bench1.log | cut -c1-600; tail -1 gpurun_out/bench_r50.log | cut -c1-400
