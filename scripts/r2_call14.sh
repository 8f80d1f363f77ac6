#!/usr/bin/env bash
# Round 2, GPU call 14 (1 GPU): VGG-11 step time vs band width (the wider band costs candidates, the narrow one fallbacks)
mkdir -p gpurun_out
for sg in 6 10 12 16; do
  EWDML_BAND_SIGMA=$sg timeout 100 python bench.py --steps 300 --warmup 3 --no-e2e > gpurun_out/b1_vgg_sigma$sg.json 2> gpurun_out/b1_vgg_sigma$sg.err; echo "vgg sigma $sg rc=$?"; python scripts/lastjson.py gpurun_out/b1_vgg_sigma$sg.json
done
