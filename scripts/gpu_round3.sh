#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_exchange.py -q > gpurun_out/pytest_exchange.log 2>&1; echo "exchange rc=$?"
timeout 300 python -m pytest tests/test_gpu_gemm.py -q -x > gpurun_out/pytest_gemm.log 2>&1; echo "gemm rc=$?"
timeout 300 python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_(encode|recv|band|refine|sample|bounds)" -s 18 -c 6 -o gpurun_out/prof_exchange python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"
tail -12 gpurun_out/pytest_exchange.log; tail -25 gpurun_out/pytest_gemm.log
