#!/usr/bin/env bash
# Round 2, GPU call 18 (1 GPU): the GPU tests that touch code changed after call 16 (checkpoint sidecar, FlatAdam, CLI)
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_overlap.py tests/test_gpu_cli.py tests/test_zz_no_alloc.py -q > gpurun_out/pytest_last.log 2>&1; echo "pytest rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_last.log | cut -c1-300 | tail
