mkdir -p gpurun_out
timeout 600 python bench/exchange_sweep.py --quick --iters 1 > gpurun_out/sweep_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_encode_push|k_band" -s 18 -c 2 -o gpurun_out/prof_encode256 python bench/exchange_sweep.py --quick --iters 1 > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
