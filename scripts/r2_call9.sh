#!/usr/bin/env bash
# Round 2, GPU call 9 (1 GPU): whole suite (no -x: list every failure), LeNet bench on the native path.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_gpu.log | cut -c1-300 | tail -20
grep -B2 -A12 "^E  " gpurun_out/pytest_gpu.log | cut -c1-300 | head -80
timeout 120 python bench.py --model LeNet --batch-size 64 --steps 100 --warmup 3 > gpurun_out/bench_lenet.json 2> gpurun_out/bench_lenet.err; echo "lenet rc=$?"; tail -3 gpurun_out/bench_lenet.err | cut -c1-300
python scripts/lastjson.py gpurun_out/bench_lenet.json
