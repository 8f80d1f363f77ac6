#!/usr/bin/env bash
# Multi-node run (one rank per node), same flags as the reference's src/run_pytorch_dist.sh: <node_rank> <nnodes> <master_ip>
NODE_RANK="$1"; NNODE="$2"; MASTER_IP="$3"; shift 3
python -m torch.distributed.run --nproc-per-node=1 --nnodes="$NNODE" --node-rank="$NODE_RANK" --master-addr "$MASTER_IP" --master-port 1234 \
  "$(dirname "$0")/../distributed_nn.py" \
  --lr=0.1 --momentum=0.9 --max-steps=100000 --epochs=100 --network=ResNet18 --dataset=Cifar10 --batch-size=64 \
  --comm-type=Bcast --num-aggregate=2 --mode=normal --eval-freq=2000 --gather-type=gather --compress-grad=compress \
  --enable-gpu= --train-dir="${TRAIN_DIR:-$HOME/}" "$@" > "out_node_${NODE_RANK}" 2>&1
