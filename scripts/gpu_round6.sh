mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_native_ops.py -q -x > gpurun_out/pytest_a.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/pytest_a.log
for c in torch native; do
  timeout 600 python bench.py --steps 200 --warmup 5 --compute $c > gpurun_out/bench_graph_$c.log 2>&1; echo bench $c rc=$?
  tail -1 gpurun_out/bench_graph_$c.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','device_error')}, d['e2e']['value'], d['config']['cuda_graph'], d['exchange'])" || tail -5 gpurun_out/bench_graph_$c.log
done
timeout 600 python bench.py --steps 100 --warmup 5 --compute native --model ResNet50 > gpurun_out/bench_graph_r50.log 2>&1; echo r50 rc=$?
tail -1 gpurun_out/bench_graph_r50.log | cut -c1-300
