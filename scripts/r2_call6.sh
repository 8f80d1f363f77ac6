#!/usr/bin/env bash
# Round 2, GPU call 6 (1 GPU): cooperative fused BatchNorm -- numerics, step times; exchange kernel list at ResNet-50 size.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_native_ops.py -x -q > gpurun_out/pytest_native.log 2>&1; echo "native ops rc=$?"; tail -5 gpurun_out/pytest_native.log | cut -c1-300
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo "bench rc=$?"
timeout 120 python bench.py --model ResNet50 --steps 50 --warmup 3 > gpurun_out/bench_r50.json 2> gpurun_out/bench_r50.err; echo "bench r50 rc=$?"
EWDML_FUSED_BN=0 timeout 120 python bench.py --model ResNet50 --steps 50 --warmup 3 --no-e2e > gpurun_out/bench_r50_nofuse.json 2> gpurun_out/bench_r50_nofuse.err; echo "bench r50 nofuse rc=$?"
timeout 120 python bench.py --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3 --no-e2e > gpurun_out/bench_r50_dense.json 2> gpurun_out/bench_r50_dense.err; echo "bench r50 dense rc=$?"
python scripts/lastjson.py gpurun_out/bench_vgg.json gpurun_out/bench_r50.json gpurun_out/bench_r50_nofuse.json gpurun_out/bench_r50_dense.json
A="--model ResNet50 --no-graph --steps 1 --warmup 3 --no-e2e --no-flush"
timeout 200 python bench.py $A > gpurun_out/plain_r50.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__grid_size --clock-control none -k regex:"k_refine|k_band|k_encode|k_recv|k_sample|k_bounds|coop|k_state" --csv --log-file gpurun_out/launches_r50_x.csv python bench.py $A > gpurun_out/ncu_r50_x.log 2>&1
echo "r50 exchange+coop list rc=$?"
