#!/usr/bin/env bash
# Round 2, GPU call 10 (4 GPUs): config 3 at 4 GPUs (ResNet-50 dense vs top-k+QSGD), VGG-11, library arm, selection diagnostics,
# accuracy-after-N with 4 workers, the re-run of the adjusted tests and the two-device static-scratch test.
mkdir -p gpurun_out
N=4
timeout 300 python -m pytest tests/test_gpu_native_ops.py tests/test_zz_no_alloc.py "tests/test_multigpu.py::test_static_scratch_entry_on_two_devices" -q > gpurun_out/pytest_fix.log 2>&1; echo "pytest rc=$?"; grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/pytest_fix.log | cut -c1-300 | tail
run() { name=$1; shift; timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python scripts/lastjson.py gpurun_out/$name.json; }
run b4_vgg --steps 100 --warmup 3
run b4_r50_tq --model ResNet50 --steps 50 --warmup 3
run b4_r50_dense --model ResNet50 --exchange allreduce --compressor none --steps 50 --warmup 3
run b4_nccl_r50_dense --impl nccl --model ResNet50 --compressor none --steps 50 --warmup 3
run b4_nccl_vgg_dense --impl nccl --compressor none --steps 100 --warmup 3
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench/accuracy_after_n.py --model VGG11 > gpurun_out/accuracy_vgg11_4gpu.jsonl 2> gpurun_out/accuracy_vgg11_4gpu.err; echo "accuracy rc=$?"; grep '^{' gpurun_out/accuracy_vgg11_4gpu.jsonl | cut -c1-330
timeout 60 python bench.py --model ResNet50 --steps 20 --warmup 3 --no-e2e > gpurun_out/b1_r50_sel.json 2>/dev/null; python -c "
import json
d=[json.loads(l) for l in open('gpurun_out/b1_r50_sel.json') if l.startswith('{')][-1]; print('1gpu r50', round(d['value']), d['exchange'])"
