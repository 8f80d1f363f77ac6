mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_native_ops.py -q > gpurun_out/pytest_native.log 2>&1; echo native rc=$?
grep -E "^E  .*Error|passed|failed|^FAILED" gpurun_out/pytest_native.log | cut -c1-200 | head -30
timeout 600 python bench.py --steps 100 --warmup 5 --compute native > gpurun_out/bench_native.log 2>&1; echo bench native rc=$?
tail -2 gpurun_out/bench_native.log | cut -c1-900
timeout 600 python bench.py --steps 50 --warmup 5 --compute native --model ResNet50 > gpurun_out/bench_native_r50.log 2>&1; echo bench r50 rc=$?
tail -2 gpurun_out/bench_native_r50.log | cut -c1-400
