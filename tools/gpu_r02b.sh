#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): slab parity on both transports, then weak / strong bench lines for the peer transport,
# NCCL and the overlapped refresh.  Every step under its own timeout.
# usage: bash tools/gpu_r02b.sh <tag> <ngpus> [quick]
tag=${1:-r02b}; N=${2:-2}; quick=${3:-}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
nvidia-smi topo -m > $out/${tag}_topo.txt 2>&1
timeout 1200 python -m pytest tests/test_multigpu.py -q --durations=5 > $out/${tag}_pytest_multigpu.log 2>&1; echo "exit $?" >> $out/${tag}_pytest_multigpu.log
for cfg in "" "GPUSOLVE_P2P=0" "GPUSOLVE_OVERLAP=1" "GPUSOLVE_P2P=0 GPUSOLVE_OVERLAP=1"; do
  name=$(echo "${cfg:-default}" | tr ' =' '__')
  env $cfg timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e $( [ -n "$cfg" ] && echo --no-extras --no-parity ) 2> $out/${tag}_weak_${name}.err | grep '"metric"' > $out/${tag}_weak_${name}.json
  if [ -z "$quick" ] || [ -z "$cfg" ]; then
    env $cfg timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 --grid 1023 --no-e2e --no-extras --no-parity 2> $out/${tag}_strong_${name}.err | grep '"metric"' > $out/${tag}_strong_${name}.json
  fi
done
NCCL_DEBUG=INFO GPUSOLVE_P2P=0 timeout 60 $TR tools/check_multigpu.py --grid 255 --mode 0 --cycles 1 2>&1 | grep -E "NVLS|P2P|via" | head -20 > $out/${tag}_nccl_transport.txt
echo done
