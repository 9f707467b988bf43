#!/bin/bash
# Multi-GPU visit on an N-GPU box (gpurun --gpus N): slab parity against one GPU, then the weak-scaling bench line
# (default grid: 513^3 per GPU) and the strong-scaling line on 1023^3.  Every step has its own timeout.
# usage: bash tools/gpu_scale.sh <tag> <ngpus> [skip-parity]
tag=${1:-s}; N=${2:-2}; skip=${3:-}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
if [ -z "$skip" ]; then
  for mode in 0 1 2; do
    GPUSOLVE_SLAB_MIN_POINTS=0 timeout 90 $TR tools/check_multigpu.py --grid 127 --mode $mode --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check127_mode$mode.log
    echo "exit ${PIPESTATUS[0]}" >> $out/${tag}_check127_mode$mode.log
  done
  timeout 120 $TR tools/check_multigpu.py --grid 511 --mode 0 --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_check511.log
  echo "exit ${PIPESTATUS[0]}" >> $out/${tag}_check511.log
fi
timeout 150 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_bench_weak.err | grep '"metric"' > $out/${tag}_bench_weak_${N}gpu.json
timeout 200 $TR bench.py --gpus $N --steps 10 --warmup 3 --grid 1023 --no-e2e 2> $out/${tag}_bench_1023.err | grep '"metric"' > $out/${tag}_bench_1023_${N}gpu.json
# prepared variant (default off): ghost refresh overlapped with the interior launches; parity first, then the same bench lines
GPUSOLVE_OVERLAP=1 timeout 120 $TR tools/check_multigpu.py --grid 511 --mode 0 --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_overlap_check511.log
echo "exit ${PIPESTATUS[0]}" >> $out/${tag}_overlap_check511.log
for mode in 1 2; do
  GPUSOLVE_OVERLAP=1 timeout 120 $TR tools/check_multigpu.py --grid 255 --mode $mode --cycles 2 2>&1 | grep -E "world|Error|error" > $out/${tag}_overlap_check255_mode$mode.log
  echo "exit ${PIPESTATUS[0]}" >> $out/${tag}_overlap_check255_mode$mode.log
done
GPUSOLVE_OVERLAP=1 timeout 150 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e 2> $out/${tag}_overlap_bench_weak.err | grep '"metric"' > $out/${tag}_overlap_bench_weak_${N}gpu.json
GPUSOLVE_OVERLAP=1 timeout 200 $TR bench.py --gpus $N --steps 10 --warmup 3 --grid 1023 --no-e2e 2> $out/${tag}_overlap_bench_1023.err | grep '"metric"' > $out/${tag}_overlap_bench_1023_${N}gpu.json
NCCL_DEBUG=INFO timeout 60 $TR tools/check_multigpu.py --grid 255 --mode 0 --cycles 1 2>&1 | grep -E "NVLS|P2P|via" | head -20 > $out/${tag}_nccl_transport.txt
echo done
