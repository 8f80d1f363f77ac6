#!/usr/bin/env bash
# Round 2, GPU call 5 (2 GPUs): VMM heap + NVLS multicast, hvd on the native kernels, PS ResNet test, 2-GPU benches.
mkdir -p gpurun_out
python scripts/probe_multicast.py > gpurun_out/probe_multicast.json 2>&1; cat gpurun_out/probe_multicast.json
timeout 900 python -m pytest tests/test_multigpu.py -x -q -s > gpurun_out/pytest_multigpu.log 2>&1; echo "multigpu rc=$?"; grep -E "transport:|passed|failed|Error|error" gpurun_out/pytest_multigpu.log | tail -12
run() { name=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/$name.json'))
    print('$name', round(d['value']), 'img/s', round(d['ms_per_step'],3), 'ms  e2e', round(d['e2e']['value']) if d.get('e2e') else None, d.get('exchange'), 'err', d.get('device_error'))
except Exception as e:
    print('$name: no json', e); print(open('gpurun_out/$name.err').read()[-1500:])
PY
}
EWDML_HEAP=ipc run b2_vgg_ipc --steps 100 --warmup 3
run b2_vgg_nvls --steps 100 --warmup 3
run b2_vgg_ps --steps 100 --warmup 3 --exchange ps
EWDML_HEAP=ipc run b2_r50_dense_ipc --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3
run b2_r50_dense_nvls --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3
run b2_r50_tq --model ResNet50 --steps 50 --warmup 3
run b2_nccl_vgg_dense --impl nccl --compressor none --steps 100 --warmup 3
run b2_nccl_vgg_tq --impl nccl --steps 100 --warmup 3
