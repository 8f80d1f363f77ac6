#!/usr/bin/env bash
# First GPU call of the next round (1 GPU, about 3 minutes): everything that was written but could not be validated or measured
# once the GPU budget of round 1 was spent.  Usage: gpurun --timeout 600 -- 'bash scripts/gpu_next.sh'
mkdir -p gpurun_out
# 1. experimental 128x256 GEMM tiles: numerics, then the step time with and without them
EWDML_GEMM_BN=256 timeout 120 python -m pytest tests/test_gpu_gemm.py -x -q -k bn256 > gpurun_out/bn256_test.log 2>&1; echo "bn256 test rc=$?"; tail -2 gpurun_out/bn256_test.log
timeout 100 python bench.py --steps 100 --warmup 3 --no-e2e > gpurun_out/bench_bn128.json 2>/dev/null; echo "bn128 rc=$?"; cut -c1-140 gpurun_out/bench_bn128.json
EWDML_GEMM_BN=256 timeout 100 python bench.py --steps 100 --warmup 3 --no-e2e > gpurun_out/bench_bn256.json 2>/dev/null; echo "bn256 rc=$?"; cut -c1-140 gpurun_out/bench_bn256.json
EWDML_GEMM_BN=256 timeout 100 python bench.py --steps 50 --warmup 3 --no-e2e --model ResNet50 > gpurun_out/bench_r50_bn256.json 2>/dev/null; cut -c1-140 gpurun_out/bench_r50_bn256.json
timeout 100 python bench.py --steps 50 --warmup 3 --no-e2e --model ResNet50 > gpurun_out/bench_r50_bn128.json 2>/dev/null; cut -c1-140 gpurun_out/bench_r50_bn128.json
# 1b. experimental persistent GEMM (two TMEM accumulators, one CTA per SM): numerics under a short timeout first, then the step time
EWDML_GEMM_PERSIST=1 timeout 90 python -m pytest tests/test_gpu_gemm.py -x -q -k "persistent" > gpurun_out/persist_test.log 2>&1; rc=$?; echo "persist test rc=$rc"; tail -2 gpurun_out/persist_test.log
if [ $rc -eq 0 ]; then
  EWDML_GEMM_PERSIST=1 timeout 200 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_native_ops.py -x -q > gpurun_out/persist_suite.log 2>&1; echo "persist suite rc=$?"; tail -2 gpurun_out/persist_suite.log
  EWDML_GEMM_PERSIST=1 timeout 100 python bench.py --steps 100 --warmup 3 --no-e2e > gpurun_out/bench_persist.json 2>/dev/null; echo "persist bench rc=$?"; cut -c1-140 gpurun_out/bench_persist.json
fi
# 2. Adam and error feedback through the public bench (native exchange), overlap on ResNet-50
timeout 100 python bench.py --steps 50 --warmup 3 --no-e2e --model ResNet50 --bucket-size 16777216 --overlap > gpurun_out/bench_r50_overlap.json 2>/dev/null; cut -c1-140 gpurun_out/bench_r50_overlap.json
# 2b. end-to-end loop with the loss read lagging one step (every step's loss is still read inside the timed region)
timeout 100 python bench.py --steps 100 --warmup 3 --e2e-lag 1 > gpurun_out/bench_e2e_lag.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/bench_e2e_lag.json'));print('e2e lag1', d['e2e']['value'], 'device', d['value'])"
# 3. the whole suite once more
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
