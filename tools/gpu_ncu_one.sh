#!/bin/bash
# ncu --set full (with source) of ONE launch of one V-cycle, exported as CSV on the box.
# usage (under gpurun): bash tools/gpu_ncu_one.sh <tag> <kernel regex> <launch-skip> [one_cycle.py args...]
tag=$1; rx=$2; skip=$3; shift 3
out=gpurun_out; mkdir -p $out
timeout 300 ncu --set full --import-source on --clock-control none -k regex:"$rx" --launch-skip $skip --launch-count 1 \
    -o /tmp/${tag} -f python tools/one_cycle.py "$@" > $out/${tag}_ncu.log 2>&1
ncu -i /tmp/${tag}.ncu-rep --page raw --csv > $out/${tag}_raw.csv 2>/dev/null
ncu -i /tmp/${tag}.ncu-rep --page source --csv > $out/${tag}_source.csv 2>/dev/null
ncu -i /tmp/${tag}.ncu-rep --page details > $out/${tag}_details.txt 2>/dev/null
ls -la /tmp/${tag}.ncu-rep >> $out/${tag}_ncu.log
[ $(stat -c %s /tmp/${tag}.ncu-rep) -lt 30000000 ] && cp /tmp/${tag}.ncu-rep $out/
echo done
