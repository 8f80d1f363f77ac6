#!/usr/bin/env bash
# Round 2, GPU call 17 (2 GPUs): the multi-GPU test file on the final tree (store teardown change, band width, NVLS)
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multigpu.py -q -s > gpurun_out/pytest_multigpu_final.log 2>&1; echo "multigpu rc=$?"; grep -E "transport:|^FAILED|^ERROR|passed|failed" gpurun_out/pytest_multigpu_final.log | cut -c1-300 | tail
