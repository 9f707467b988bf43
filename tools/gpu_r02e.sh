#!/bin/bash
# N-GPU visit, budget-conscious: slab parity tests for selected worlds on the peer transport, then the weak bench line
# (with slab_parity and strong_1023) and A/B lines.   usage: bash tools/gpu_r02e.sh <tag> <ngpus> "<worlds>" [cfg ...]
tag=$1; N=$2; worlds=$3; shift 3
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
GPUSOLVE_TEST_WORLDS=$worlds GPUSOLVE_TEST_TRANSPORTS=${TRANSPORTS:-p2p} timeout 900 python -m pytest tests/test_multigpu.py -x -q --durations=5 > $out/${tag}_pytest_multigpu.log 2>&1; echo "exit $?" >> $out/${tag}_pytest_multigpu.log
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_weak_default.err | grep '"metric"' > $out/${tag}_weak_default.json
for cfg in "$@"; do
  name=$(echo "$cfg" | tr ' =' '__')
  env $cfg timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-extras --no-parity 2> $out/${tag}_weak_${name}.err | grep '"metric"' > $out/${tag}_weak_${name}.json
done
echo done
