#!/usr/bin/env bash
# Round 2, GPU call 1 (1 GPU): suite, experimental GEMM variants, the three bench arms.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_vgg.json
EWDML_GEMM_PERSIST=1 timeout 90 python -m pytest tests/test_gpu_gemm.py -x -q -k "persistent" > gpurun_out/persist_test.log 2>&1; rc=$?; echo "persist test rc=$rc"; tail -2 gpurun_out/persist_test.log
if [ $rc -eq 0 ]; then
  EWDML_GEMM_PERSIST=1 timeout 200 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_native_ops.py -x -q > gpurun_out/persist_suite.log 2>&1; echo "persist suite rc=$?"; tail -2 gpurun_out/persist_suite.log
  EWDML_GEMM_PERSIST=1 timeout 100 python bench.py --steps 100 --warmup 3 --no-e2e > gpurun_out/bench_persist.json 2>/dev/null; echo "persist bench rc=$?"; cut -c1-140 gpurun_out/bench_persist.json
fi
EWDML_GEMM_BN=256 timeout 120 python -m pytest tests/test_gpu_gemm.py -x -q -k bn256 > gpurun_out/bn256_test.log 2>&1; rc=$?; echo "bn256 test rc=$rc"; tail -2 gpurun_out/bn256_test.log
if [ $rc -eq 0 ]; then
  EWDML_GEMM_BN=256 timeout 100 python bench.py --steps 100 --warmup 3 --no-e2e > gpurun_out/bench_bn256.json 2>/dev/null; echo "bn256 rc=$?"; cut -c1-140 gpurun_out/bench_bn256.json
fi
for comp in topk-qsgd none; do
  timeout 200 python bench.py --impl nccl --compressor $comp --steps 100 --warmup 3 > gpurun_out/bench_nccl_vgg_$comp.json 2> gpurun_out/bench_nccl_vgg_$comp.err; echo "nccl vgg $comp rc=$?"; cut -c1-300 gpurun_out/bench_nccl_vgg_$comp.json; tail -3 gpurun_out/bench_nccl_vgg_$comp.err
done
timeout 200 python bench.py --impl nccl --compressor none --model ResNet50 --steps 50 --warmup 3 > gpurun_out/bench_nccl_r50_none.json 2> gpurun_out/bench_nccl_r50_none.err; echo "nccl r50 rc=$?"; cut -c1-300 gpurun_out/bench_nccl_r50_none.json; tail -3 gpurun_out/bench_nccl_r50_none.err
timeout 200 python bench.py --impl nccl --compressor topk-qsgd --model ResNet50 --steps 50 --warmup 3 > gpurun_out/bench_nccl_r50_tq.json 2> gpurun_out/bench_nccl_r50_tq.err; echo "nccl r50 tq rc=$?"; cut -c1-300 gpurun_out/bench_nccl_r50_tq.json; tail -3 gpurun_out/bench_nccl_r50_tq.err
timeout 120 python bench.py --model ResNet50 --steps 50 --warmup 3 > gpurun_out/bench_r50.json 2> gpurun_out/bench_r50.err; echo "bench r50 rc=$?"; cut -c1-300 gpurun_out/bench_r50.json
