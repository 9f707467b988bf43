#!/bin/bash
# GPU visit: parity suite, A/B of the arithmetic / grid-sizing knobs, short bench
tag=${1:-v1}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
bash tools/gpu_variants.sh $tag "unit_taps=0 wave=0" "unit_taps=1 wave=0" "unit_taps=0 wave=1" "unit_taps=1 wave=1" \
   "zchunk2=64" "zchunk2=96" "rr_variant=1" "fuse2=2" "fuse2=3"
timeout 300 python bench.py --no-cpu > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo done
