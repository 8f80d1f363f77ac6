#!/usr/bin/env bash
# Round 2, GPU call 13 (1 GPU): how often does the top-k selection fall back to the exact scan, and what does a wider band buy?
mkdir -p gpurun_out
for sg in 6 12 24; do
  EWDML_BAND_SIGMA=$sg timeout 100 python bench.py --model ResNet50 --steps 200 --warmup 3 --no-e2e > gpurun_out/b1_r50_sigma$sg.json 2> gpurun_out/b1_r50_sigma$sg.err; echo "sigma $sg rc=$?"; python scripts/lastjson.py gpurun_out/b1_r50_sigma$sg.json
done
for sg in 6 24; do
  EWDML_BAND_SIGMA=$sg timeout 100 python bench.py --steps 300 --warmup 3 --no-e2e > gpurun_out/b1_vgg_sigma$sg.json 2> gpurun_out/b1_vgg_sigma$sg.err; echo "vgg sigma $sg rc=$?"; python scripts/lastjson.py gpurun_out/b1_vgg_sigma$sg.json
done
