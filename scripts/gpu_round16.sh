mkdir -p gpurun_out
timeout 600 python bench/exchange_sweep.py --quick --iters 2 > gpurun_out/sweep_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|k_encode" --csv --log-file gpurun_out/launches_sweep.csv python bench/exchange_sweep.py --quick --iters 2 > gpurun_out/ncu.log 2>&1
echo ncu rc=$?; grep gradient_mb gpurun_out/sweep_plain.log | cut -c1-300
