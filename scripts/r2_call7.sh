#!/usr/bin/env bash
# Round 2, GPU call 7 (8 GPUs): multi-GPU tests (IPC and VMM+NVLS heaps, PS ResNet on all GPUs, hvd on native kernels),
# benches (VGG-11 / ResNet-50: compressed allgather, dense allreduce, parameter server; library arm), K4 and exchange sweeps.
mkdir -p gpurun_out
N=8
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
timeout 600 python -m pytest tests/test_multigpu.py -x -q -s -k "8-auto or 8-ipc or parameter_server_resnet or hvd or allreduce-3" > gpurun_out/pytest_multigpu8.log 2>&1; echo "multigpu8 rc=$?"; grep -E "transport:|passed|failed|Error|assert" gpurun_out/pytest_multigpu8.log | tail -8
run() { name=$1; shift; timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python scripts/lastjson.py gpurun_out/$name.json; }
run b8_vgg --steps 100 --warmup 3
run b8_vgg_ps --steps 50 --warmup 3 --exchange ps
run b8_r50_tq --model ResNet50 --steps 50 --warmup 3
run b8_r50_dense --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3
run b8_r50_ps --model ResNet50 --exchange ps --steps 50 --warmup 3
NCCL_DEBUG=INFO run b8_nccl_vgg_dense --impl nccl --compressor none --steps 100 --warmup 3
grep -m3 -i "nvls" gpurun_out/b8_nccl_vgg_dense.err | cut -c1-200
run b8_nccl_vgg_tq --impl nccl --steps 100 --warmup 3
run b8_nccl_r50_dense --impl nccl --model ResNet50 --compressor none --steps 50 --warmup 3
timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench/allreduce_bw.py > gpurun_out/allreduce_bw_8gpu.jsonl 2> gpurun_out/allreduce_bw_8gpu.err; echo "allreduce_bw rc=$?"; grep -v "^NCCL" gpurun_out/allreduce_bw_8gpu.jsonl | cut -c1-220
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench/exchange_sweep.py > gpurun_out/exchange_sweep_8gpu.jsonl 2> gpurun_out/exchange_sweep_8gpu.err; echo "sweep rc=$?"; grep -c gradient_mb gpurun_out/exchange_sweep_8gpu.jsonl; tail -3 gpurun_out/exchange_sweep_8gpu.err | cut -c1-300
