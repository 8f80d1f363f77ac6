mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_native_ops.py tests/test_gpu_gemm.py tests/test_zz_no_alloc.py -q > gpurun_out/pytest_b.log 2>&1; echo tests rc=$?; tail -4 gpurun_out/pytest_b.log | cut -c1-200
timeout 600 python bench.py --steps 200 --warmup 5 --compute native > gpurun_out/bench_native2.log 2>&1; echo bench rc=$?
tail -1 gpurun_out/bench_native2.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'])" || tail -5 gpurun_out/bench_native2.log
timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 500 -c 280 --csv --log-file gpurun_out/launches_native2.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
timeout 600 python bench.py --impl nccl --steps 50 --warmup 5 > gpurun_out/bench_nccl.log 2>&1; echo nccl rc=$?; tail -1 gpurun_out/bench_nccl.log | cut -c1-400
