#!/bin/bash
# First GPU visit of the next round: the variants written without GPU access, each under its own
# short timeout so that a hang costs seconds.  usage (under gpurun): bash tools/gpu_prepared.sh <tag>
tag=${1:-r02a}
out=gpurun_out
mkdir -p $out
export GPUSOLVE_TEST_PREPARED=1
timeout 120 python -m pytest tests/test_parity_gpu.py -q -k "coarse_tail" > $out/${tag}_tail_tests.log 2>&1; echo "exit $?" >> $out/${tag}_tail_tests.log
timeout 120 python -m pytest tests/test_parity_gpu.py -q -k "split_phase" > $out/${tag}_sp_tests.log 2>&1; echo "exit $?" >> $out/${tag}_sp_tests.log
# A/B of the bench grid: default, coarse tail at two thresholds, split-phase residual+restriction, both
for cfg in "" "GPUSOLVE_TAIL_POINTS=4096" "GPUSOLVE_TAIL_POINTS=40000" "GPUSOLVE_RR_VARIANT=2" "GPUSOLVE_TAIL_POINTS=4096 GPUSOLVE_RR_VARIANT=2"; do
    name=$(echo "${cfg:-default}" | tr ' =' '__')
    env $cfg timeout 100 python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e > $out/${tag}_bench_${name}.json 2> $out/${tag}_bench_${name}.err
done
# the reference's own experiment sizes, where the coarse tail is a large share of the cycle
for n in 63 127 255; do
    for cfg in "" "GPUSOLVE_TAIL_POINTS=4096" "GPUSOLVE_TAIL_POINTS=40000"; do
        name=$(echo "${cfg:-default}" | tr ' =' '__')
        env $cfg timeout 60 python tools/profile_levels.py --n $n > $out/${tag}_levels_n${n}_${name}.txt 2>&1
    done
done
# small-level thresholds once more, now that the per-level launch costs have changed
bash tools/gpu_variants.sh ${tag}_thr "fuse2_min_n=64" "fast_min_n=32" "fuse2_min_n=64 fast_min_n=32" "tail_points=4096 fuse2_min_n=64"
echo done
