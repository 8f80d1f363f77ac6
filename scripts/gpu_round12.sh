mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemm_tcgen05 -s 62 -c 14 -o gpurun_out/prof_gemm python bench.py --steps 3 --warmup 3 --no-e2e --no-graph --compute native > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
