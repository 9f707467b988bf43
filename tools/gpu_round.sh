#!/bin/bash
# One GPU-box visit: GPU test suite, bench (both arms), per-level timings, ncu launch list and one --set full capture.
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt; lscpu | head -20 >> $out/${tag}_nproc.txt
timeout 300 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_gpu.log
timeout 200 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?" >> $out/${tag}_bench.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 100 python tools/profile_levels.py --out $out/${tag}_levels.json > $out/${tag}_levels.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 1 > $out/${tag}_levels_mode1.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 2 --pre 3 --post 3 > $out/${tag}_levels_mode2.txt 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e > $out/${tag}_ncu_launches.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'smooth2|stencil_tma|resid_restrict|prolong_correct' -c 12 \
    -o $out/${tag}_full -f python tools/one_cycle.py --n 511 > $out/${tag}_ncu_full.log 2>&1
echo done
