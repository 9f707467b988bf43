#!/bin/bash
# 1-GPU visit after a kernel change: GPU suite, then A/B bench lines and per-level timings of option sets given as arguments.
# usage (under gpurun): bash tools/gpu_r02c.sh <tag> "ENV=val ..." "ENV=val ..." ...
tag=${1:-r02c}; shift
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_gpu.log
for cfg in "" "$@"; do
    name=$(echo "${cfg:-default}" | tr ' =' '__')
    env $cfg timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e --no-extras > $out/${tag}_ab_${name}.json 2> $out/${tag}_ab_${name}.err
    env $cfg timeout 100 python tools/profile_levels.py > $out/${tag}_levels_${name}.txt 2>&1
done
echo done
