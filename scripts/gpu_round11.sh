mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_native_ops.py tests/test_gpu_gemm.py tests/test_zz_no_alloc.py -q > gpurun_out/pytest_d.log 2>&1; echo tests rc=$?; grep -E "^E  .*Error|passed|failed|^FAILED" gpurun_out/pytest_d.log | cut -c1-220 | head -20
timeout 600 python bench.py --steps 200 --warmup 5 --compute native > gpurun_out/bench_native4.log 2>&1; echo bench rc=$?
tail -1 gpurun_out/bench_native4.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'])" || tail -5 gpurun_out/bench_native4.log
timeout 600 python bench.py --steps 100 --warmup 5 --compute native --model ResNet50 > gpurun_out/bench_native4_r50.log 2>&1; echo r50 rc=$?
tail -1 gpurun_out/bench_native4_r50.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'])" || tail -5 gpurun_out/bench_native4_r50.log
