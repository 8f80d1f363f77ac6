mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_exchange.py -q > gpurun_out/pytest_f.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/pytest_f.log
timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_(encode|recv|band|refine)" -s 12 -c 4 -o gpurun_out/prof_exchange2 python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
