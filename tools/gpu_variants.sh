#!/bin/bash
# per-level timings for several option settings.  usage: bash tools/gpu_variants.sh <tag> "opt=val ..." "opt=val ..." ...
tag=$1; shift
out=gpurun_out; mkdir -p $out
i=0
for o in "$@"; do
  args=""; for kv in $o; do args="$args --opt $kv"; done
  echo "== $o" >> $out/${tag}_variants.txt
  timeout 120 python tools/profile_levels.py $args 2>&1 | head -14 >> $out/${tag}_variants.txt
  i=$((i+1))
done
