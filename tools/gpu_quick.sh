#!/bin/bash
# quick GPU iteration: GPU suite (stop at first failure), per-level operator times, short bench
# usage: bash tools/gpu_quick.sh <tag> [pytest -k expression]
tag=${1:-q}; kexpr=${2:-}
out=gpurun_out; mkdir -p $out
if [ -n "$kexpr" ]; then
  timeout 400 python -m pytest tests -m gpu -x -q -k "$kexpr" > $out/${tag}_pytest.log 2>&1
else
  timeout 400 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1
fi
echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 120 python tools/profile_levels.py --out $out/${tag}_levels.json > $out/${tag}_levels.txt 2>&1
timeout 300 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo done
