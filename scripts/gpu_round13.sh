mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_e.log 2>&1; echo tests rc=$?; grep -E "^E  .*Error|passed|failed|^FAILED" gpurun_out/pytest_e.log | cut -c1-220 | head -20
timeout 600 python bench.py --steps 200 --warmup 5 --compute native > gpurun_out/bench_native5.log 2>&1; echo bench rc=$?
tail -1 gpurun_out/bench_native5.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'], d['exchange'])" || tail -5 gpurun_out/bench_native5.log
timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 520 -c 130 --csv --log-file gpurun_out/launches_native5.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
