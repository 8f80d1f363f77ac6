#!/usr/bin/env bash
# Round 2, GPU call 15 (4 GPUs): band width 12 -- does the compressed ResNet-50 path close on the dense one?  (before: 8.67 ms vs 7.71 ms)
mkdir -p gpurun_out
N=4
run() { name=$1; shift; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python scripts/lastjson.py gpurun_out/$name.json; }
run b4_r50_tq_s12 --model ResNet50 --steps 100 --warmup 3 --no-e2e
run b4_vgg_s12 --steps 200 --warmup 3
