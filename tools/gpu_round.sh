#!/bin/bash
# One GPU-box visit: GPU test suite, bench (both arms), per-level timings, ncu launch list of the bench command and one
# --set full capture of a plain-launch V-cycle (exported as CSV on the box: the .ncu-rep would exceed what travels back).
# usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-r02z}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt; lscpu | head -20 >> $out/${tag}_nproc.txt
timeout 600 python -m pytest tests -m gpu -q --durations=8 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_gpu.log
timeout 400 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?" >> $out/${tag}_bench.err
timeout 300 python bench.py --impl reference > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 100 python tools/profile_levels.py > $out/${tag}_levels.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 1 > $out/${tag}_levels_mode1.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 2 --pre 3 --post 3 > $out/${tag}_levels_mode2.txt 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-extras > $out/${tag}_ncu_launches.log 2>&1
timeout 400 ncu --set full --clock-control none -k regex:'smooth2|stencil_tma|resid_restrict|prolong_correct|restrict_zmarch' -c 14 \
    -o /tmp/${tag}_full -f python tools/one_cycle.py --n 511 --cycles 2 > $out/${tag}_ncu_full.log 2>&1
ncu -i /tmp/${tag}_full.ncu-rep --page raw --csv > $out/${tag}_ncu_full_raw.csv 2>/dev/null
echo done
