#!/usr/bin/env bash
# Polling evaluator, same flags as the reference's src/evaluate_pytorch.sh.
python "$(dirname "$0")/../distributed_evaluator.py" --eval-batch-size="${EVAL_BS:-1000}" --eval-freq=50 --network=VGG11 \
  --dataset=Cifar10 --model-dir="${MODEL_DIR:-output/models/}" "$@"
