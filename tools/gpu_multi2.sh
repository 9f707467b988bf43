#!/bin/bash
# usage: bash tools/gpu_multi2.sh <tag> <ngpus> : slab parity vs one GPU (3 modes, deep and default partitions) + weak/strong bench
tag=${1:-m}; N=${2:-2}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for mode in 0 1 2; do
  GPUSOLVE_SLAB_MIN_POINTS=0 timeout 120 $TR tools/check_multigpu.py --grid 127 --mode $mode --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check_mode$mode.log; echo "exit $?" >> $out/${tag}_check_mode$mode.log
done
timeout 150 $TR tools/check_multigpu.py --grid 255 --mode 0 --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check255.log; echo "exit $?" >> $out/${tag}_check255.log
timeout 150 $TR tools/check_multigpu.py --grid 255 --mode 2 --pre 3 --post 3 --cycles 1 2>&1 | grep -E "world|Error|error" > $out/${tag}_check255nt.log; echo "exit $?" >> $out/${tag}_check255nt.log
timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_bench_weak.err | grep '"metric"' > $out/${tag}_bench_weak.json
timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 --grid 1023 --no-e2e 2> $out/${tag}_bench_1023.err | grep '"metric"' > $out/${tag}_bench_1023.json
echo done
