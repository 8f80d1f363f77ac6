mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 60 python scripts/coll_bw.py > gpurun_out/coll_bw.jsonl 2>&1; echo "coll_bw rc=$?"; cat gpurun_out/coll_bw.jsonl | cut -c1-200
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_r21.json 2> gpurun_out/bench_r21.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_r21.json
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
