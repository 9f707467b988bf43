#!/bin/bash
# A/B of programmatic dependent launch on the small-level kernels: GPU suite, then graph V-cycle times per grid size.
out=gpurun_out; tag=${1:-r02r}
timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "exit $?" >> $out/${tag}_pytest.log
for pdl in 0 1; do for n in 63 127 255 511; do
  echo "pdl=$pdl n=$n: $(GPUSOLVE_PDL=$pdl timeout 100 python tools/profile_levels.py --n $n 2>&1 | head -1)" >> $out/${tag}_pdl.txt
done; echo "pdl=$pdl shipped 127 mode 2: $(GPUSOLVE_PDL=$pdl timeout 100 python tools/profile_levels.py --n 127 --mode 2 --pre 3 --post 3 2>&1 | head -1)" >> $out/${tag}_pdl.txt; done
