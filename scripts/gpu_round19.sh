# full single-GPU test suite (new: collective engine, overlap, error feedback), then the default bench
mkdir -p gpurun_out
timeout 540 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 150 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_r19.json 2> gpurun_out/bench_r19.err
echo "bench rc=$?"; cat gpurun_out/bench_r19.json; tail -3 gpurun_out/bench_r19.err
