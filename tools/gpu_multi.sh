#!/bin/bash
# multi-GPU visit: parity of the slab solver vs single GPU, then bench at N GPUs (weak default grid and strong 1023^3)
# usage: bash tools/gpu_multi.sh <tag> <ngpus> [extra bench args]
tag=${1:-m}; N=${2:-2}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
for mode in 0 1 2; do
  timeout 600 $TR tools/check_multigpu.py --grid 127 --mode $mode --cycles 2 > $out/${tag}_check_mode$mode.log 2>&1; echo "exit $?" >> $out/${tag}_check_mode$mode.log
done
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > $out/${tag}_bench_weak.json 2> $out/${tag}_bench_weak.err
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --grid 1023 --no-e2e > $out/${tag}_bench_1023.json 2> $out/${tag}_bench_1023.err
echo done
