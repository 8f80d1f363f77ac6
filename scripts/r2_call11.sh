#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 120 python scripts/k2_scaling.py ResNet50 > gpurun_out/k2_scaling.jsonl 2> gpurun_out/k2_scaling.err; echo "rc=$?"; cat gpurun_out/k2_scaling.jsonl; tail -3 gpurun_out/k2_scaling.err
timeout 120 python scripts/k2_scaling.py VGG11 >> gpurun_out/k2_scaling.jsonl 2>> gpurun_out/k2_scaling.err; echo "rc=$?"; tail -4 gpurun_out/k2_scaling.jsonl
