#!/usr/bin/env bash
# Round 2, GPU call 8 (1 GPU): full suite with the direct small-channel conv kernels (LeNet native), stand-alone ReLU / pool,
# tightened whole-model test; LeNet bench; accuracy-after-N arms for VGG-11.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log | cut -c1-400
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 120 python bench.py --model LeNet --batch-size 64 --steps 100 --warmup 3 > gpurun_out/bench_lenet.json 2> gpurun_out/bench_lenet.err; echo "lenet rc=$?"; tail -2 gpurun_out/bench_lenet.err | cut -c1-300
timeout 120 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo "vgg rc=$?"
python scripts/lastjson.py gpurun_out/bench_lenet.json gpurun_out/bench_vgg.json
timeout 300 python bench/accuracy_after_n.py --model VGG11 > gpurun_out/accuracy_vgg11_1gpu.jsonl 2> gpurun_out/accuracy_vgg11_1gpu.err; echo "accuracy rc=$?"; cut -c1-400 gpurun_out/accuracy_vgg11_1gpu.jsonl; tail -3 gpurun_out/accuracy_vgg11_1gpu.err | cut -c1-300
