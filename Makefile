# Convenience targets; every one is a thin wrapper over a command that also works on its own.
PY ?= python
GPUS ?= 1

.PHONY: build test test-gpu bench bench-ref bench-nccl smoke info sass clean

build:            ## nvcc -gencode arch=compute_100a,code=sm_100a -> ewdml/_C/libewdml.so
	$(PY) -m ewdml.build

test:             ## CPU suite (oracles, gloo parameter server, checkpoints, CLIs, host sanitizers)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu:         ## B200 suite (kernels vs fp32 torch, trainers, multi-GPU when >= 2 devices)
	$(PY) -m pytest tests -x -q -m gpu

smoke:            ## one tiny forward + backward + exchange on cuda:0
	$(PY) -c "import __graft_entry__ as g; g.smoke()"

bench:            ## make bench GPUS=8
ifeq ($(GPUS),1)
	$(PY) bench.py
else
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(GPUS) --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $(GPUS)
endif

bench-ref:        ## the unmodified reference (baseline/_ref; bash baseline/install_reference.sh installs it)
	$(PY) bench.py --impl reference --gpus $(GPUS)

bench-nccl:       ## library arm: CUDA-graphed cuDNN + NCCL
ifeq ($(GPUS),1)
	$(PY) bench.py --impl nccl
else
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(GPUS) --master-addr 127.0.0.1 --master-port 29512 bench.py --impl nccl --gpus $(GPUS)
endif

info:             ## devices, peer matrix, heap transport capabilities, configuration registry
	$(PY) -m ewdml.cli.info

sass:             ## tcgen05 / TMA / multimem mnemonics in the built library -> profiles/sass_evidence.txt
	$(PY) scripts/sass_evidence.py

clean:
	rm -rf build ewdml/_C/*.so ewdml.egg-info .pytest_cache
