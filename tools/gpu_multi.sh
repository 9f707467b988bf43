#!/bin/bash
# N-GPU bench lines only (each carries its own slab-vs-one-GPU parity record).  usage: bash tools/gpu_multi.sh <tag> <ngpus> [cfg ...]
tag=$1; N=$2; shift 2
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_weak_default.err | grep '"metric"' > $out/${tag}_weak_default.json
for cfg in "$@"; do
  name=$(echo "$cfg" | tr ' =' '__')
  env $cfg timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-extras 2> $out/${tag}_weak_${name}.err | grep '"metric"' > $out/${tag}_weak_${name}.json
done
echo done
