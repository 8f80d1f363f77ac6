#!/usr/bin/env bash
# Round 2, GPU call 2 (1 GPU): per-kernel launch lists (ResNet-50 step, VGG-11 with the persistent GEMM).
mkdir -p gpurun_out
A="--no-graph --steps 1 --warmup 3 --no-e2e --no-flush"
timeout 200 python bench.py --model ResNet50 $A > gpurun_out/plain_r50.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r50.csv python bench.py --model ResNet50 $A > gpurun_out/ncu_r50.log 2>&1
echo "r50 list rc=$?"
EWDML_GEMM_PERSIST=1 timeout 200 python bench.py $A > gpurun_out/plain_vgg_persist.log 2>&1 &&
EWDML_GEMM_PERSIST=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_vgg_persist.csv python bench.py $A > gpurun_out/ncu_vgg_persist.log 2>&1
echo "vgg persist list rc=$?"
timeout 200 python bench.py $A > gpurun_out/plain_vgg.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_vgg.csv python bench.py $A > gpurun_out/ncu_vgg.log 2>&1
echo "vgg list rc=$?"
