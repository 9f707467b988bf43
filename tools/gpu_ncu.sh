#!/bin/bash
# ncu --set full of selected kernels in one plain-launch V-cycle.  usage: bash tools/gpu_ncu.sh <tag> <kernel regex> <count> [one_cycle args]
tag=$1; rx=$2; cnt=$3; shift 3
out=gpurun_out; mkdir -p $out
python tools/one_cycle.py "$@" > $out/${tag}_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$rx" -c $cnt -o $out/${tag} -f python tools/one_cycle.py "$@" > $out/${tag}_ncu.log 2>&1
echo "exit $?"
