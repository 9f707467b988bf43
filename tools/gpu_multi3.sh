#!/bin/bash
# short multi-GPU visit: slab parity vs one GPU (deep partition at 127, default at 255) and the weak-scaling bench line
tag=${1:-m}; N=${2:-2}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
GPUSOLVE_SLAB_MIN_POINTS=0 timeout 60 $TR tools/check_multigpu.py --grid 127 --mode 0 --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check127.log; echo "exit $?" >> $out/${tag}_check127.log
timeout 60 $TR tools/check_multigpu.py --grid 255 --mode 0 --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check255.log; echo "exit $?" >> $out/${tag}_check255.log
timeout 90 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_bench_weak.err | grep '"metric"' > $out/${tag}_bench_weak.json
echo done
