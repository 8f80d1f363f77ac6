#!/usr/bin/env bash
# Round 2, GPU call 4 (1 GPU): per-launch times + DRAM bytes of the BatchNorm kernels inside a ResNet-50 step.
mkdir -p gpurun_out
A="--model ResNet50 --no-graph --steps 1 --warmup 3 --no-e2e --no-flush"
timeout 200 python bench.py $A > gpurun_out/plain_r50.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__grid_size --clock-control none -k regex:"bn_|colstats" --csv --log-file gpurun_out/launches_r50_bn.csv python bench.py $A > gpurun_out/ncu_r50_bn.log 2>&1
echo "r50 bn list rc=$?"
