#!/bin/bash
# GPU visit: parity suite, then per-level timings for the option sets given as arguments, then a short bench
tag=${1:-v}; shift
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
if grep -q "pytest exit 0" $out/${tag}_pytest.log; then
  bash tools/gpu_variants.sh $tag "$@"
  timeout 300 python bench.py --no-cpu > $out/${tag}_bench.json 2> $out/${tag}_bench.err
fi
echo done
