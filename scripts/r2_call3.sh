#!/usr/bin/env bash
# Round 2, GPU call 3 (1 GPU): streaming BatchNorm kernels -- numerics, then VGG-11 / ResNet-50 step times.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_native_ops.py -x -q > gpurun_out/pytest_native.log 2>&1; echo "native ops rc=$?"; tail -5 gpurun_out/pytest_native.log | cut -c1-300
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_vgg.json'));print('vgg', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks'])"
timeout 120 python bench.py --model ResNet50 --steps 50 --warmup 3 > gpurun_out/bench_r50.json 2> gpurun_out/bench_r50.err; echo "bench r50 rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_r50.json'));print('r50', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
timeout 120 python bench.py --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3 > gpurun_out/bench_r50_dense.json 2> gpurun_out/bench_r50_dense.err; echo "bench r50 dense rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_r50_dense.json'));print('r50 dense', d['value'], d['ms_per_step'])"
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
