#!/bin/bash
# usage: tools/scale_run.sh N "<extra bench args>" tag     (runs bench.py under torchrun on N GPUs, appends the JSON line)
N=$1; EXTRA=$2; TAG=$3
PORT=$((29600 + RANDOM % 300))
if [ "$N" = "1" ]; then
  python bench.py --gpus 1 $EXTRA 2>>gpurun_out/scale.err | tee -a gpurun_out/scale_$TAG.jsonl
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N $EXTRA 2>>gpurun_out/scale.err | tee -a gpurun_out/scale_$TAG.jsonl
fi
