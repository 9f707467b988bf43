#!/bin/bash
# Round-2 first GPU visit (1 GPU): whole GPU suite incl. the prepared variants and the 513^3 parity cases, the new bench
# line, A/B of the prepared variants, ncu evidence for modes 1/2.  Every step under its own timeout.
# usage (under gpurun): bash tools/gpu_r02a.sh <tag>
tag=${1:-r02a}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt; free -g >> $out/${tag}_nproc.txt; lscpu | head -20 >> $out/${tag}_nproc.txt
GPUSOLVE_TEST_PREPARED=1 timeout 900 python -m pytest tests -m gpu -q --durations=15 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest_gpu.log
timeout 400 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?" >> $out/${tag}_bench.err
timeout 200 python bench.py --impl reference > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
for cfg in "" "GPUSOLVE_FUSE_PC=1" "GPUSOLVE_TAIL_POINTS=4096" "GPUSOLVE_TAIL_POINTS=40000" "GPUSOLVE_RR_VARIANT=2" "GPUSOLVE_TAIL_POINTS=4096 GPUSOLVE_RR_VARIANT=2 GPUSOLVE_FUSE_PC=1"; do
    name=$(echo "${cfg:-default}" | tr ' =' '__')
    env $cfg timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e --no-extras > $out/${tag}_ab_${name}.json 2> $out/${tag}_ab_${name}.err
done
for n in 63 127 255; do
    for cfg in "" "GPUSOLVE_TAIL_POINTS=4096" "GPUSOLVE_TAIL_POINTS=40000"; do
        name=$(echo "${cfg:-default}" | tr ' =' '__')
        env $cfg timeout 60 python tools/profile_levels.py --n $n > $out/${tag}_levels_n${n}_${name}.txt 2>&1
    done
done
timeout 100 python tools/profile_levels.py > $out/${tag}_levels.txt 2>&1
timeout 100 python tools/profile_levels.py --opt fuse_pc=1 > $out/${tag}_levels_pc.txt 2>&1
cat > /tmp/skew.conf <<'EOC'
5
1e-30
511 511 511
0
2
2
0.8
1.0
6 -1.1 -0.9 -1.2 -0.8 -1.3 -0.7
0 0 1 0 -1 0 0
0 0 0 -1 0 1 0
0 1 0 0 0 0 -1
EOC
GPUSOLVE_TIMERS=1 timeout 100 gpu-solve_b200/bin/GpuSolve-cuda /tmp/skew.conf > $out/${tag}_skew511.txt 2>&1
GPUSOLVE_TIMERS=1 GPUSOLVE_KERNELS=0 timeout 100 gpu-solve_b200/bin/GpuSolve-cuda /tmp/skew.conf > $out/${tag}_skew511_simple.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 1 > $out/${tag}_levels_mode1.txt 2>&1
timeout 100 python tools/profile_levels.py --mode 2 --pre 3 --post 3 > $out/${tag}_levels_mode2.txt 2>&1
# ncu: raw counters only, exported to CSV on the box (the .ncu-rep files would exceed the 64 MiB that travel back)
timeout 300 ncu --set full --clock-control none -k regex:'smooth2|resid_restrict|fas_rhs|restrict_zmarch|stencil_tma' -c 8 \
    -o /tmp/${tag}_mode1 -f python tools/one_cycle.py --n 511 --mode 1 > $out/${tag}_ncu_mode1.log 2>&1
ncu -i /tmp/${tag}_mode1.ncu-rep --page raw --csv > $out/${tag}_ncu_mode1_raw.csv 2>/dev/null
timeout 300 ncu --set full --clock-control none -k regex:'smooth2|resid_restrict|stencil_tma' -c 8 \
    -o /tmp/${tag}_mode2 -f python tools/one_cycle.py --n 511 --mode 2 --pre 3 --post 3 --no-newton-step > $out/${tag}_ncu_mode2.log 2>&1
ncu -i /tmp/${tag}_mode2.ncu-rep --page raw --csv > $out/${tag}_ncu_mode2_raw.csv 2>/dev/null
timeout 300 ncu --set full --clock-control none -k regex:'smooth2|resid_restrict|prolong' -c 8 \
    -o /tmp/${tag}_mode0 -f python tools/one_cycle.py --n 511 --mode 0 --opt fuse_pc=1 > $out/${tag}_ncu_mode0.log 2>&1
ncu -i /tmp/${tag}_mode0.ncu-rep --page raw --csv > $out/${tag}_ncu_mode0_raw.csv 2>/dev/null
du -sh $out > $out/${tag}_du.txt
echo done
