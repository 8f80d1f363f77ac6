#!/usr/bin/env bash
# Round 2, GPU call 16 (1 GPU): the whole -m gpu suite and smoke() on the final tree, headline bench line
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_gpu_final.log | cut -c1-300 | tail
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; python scripts/lastjson.py gpurun_out/bench_final.json
