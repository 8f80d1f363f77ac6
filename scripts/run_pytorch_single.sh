#!/usr/bin/env bash
# Single-node run, same flags as the reference's src/run_pytorch_single.sh (LeNet/MNIST, 1 server + 2 workers).
NPROC="${NPROC:-3}"
python -m torch.distributed.run --nnodes=1 --nproc-per-node="$NPROC" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" \
  "$(dirname "$0")/../distributed_nn.py" \
  --lr=0.01 --momentum=0.9 --max-steps=10000 --epochs=10 --network=LeNet --dataset=MNIST --batch-size=64 \
  --comm-type=Bcast --num-aggregate=5 --mode=normal --eval-freq=20 --gather-type=gather --compress-grad=compress \
  --enable-gpu= --train-dir="${TRAIN_DIR:-output/models/}" "$@"
