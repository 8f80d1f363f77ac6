mkdir -p gpurun_out
timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 520 -c 130 --csv --log-file gpurun_out/launches_native3.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
