mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_exchange.py -q > gpurun_out/pytest_f.log 2>&1; echo tests rc=$?; grep -E "^E  .*Error|passed|failed|^FAILED" gpurun_out/pytest_f.log | cut -c1-200 | head
timeout 600 python bench/exchange_sweep.py --quick --iters 5 > gpurun_out/sweep_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|k_encode" --csv --log-file gpurun_out/launches_sweep.csv python bench/exchange_sweep.py --quick --iters 2 > gpurun_out/ncu.log 2>&1
echo ncu rc=$?; grep gradient_mb gpurun_out/sweep_plain.log | cut -c1-260
timeout 600 python bench.py --steps 200 --warmup 5 > gpurun_out/bench_native7.log 2>&1; echo bench rc=$?
tail -1 gpurun_out/bench_native7.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'], d['exchange'])" || tail -5 gpurun_out/bench_native7.log
