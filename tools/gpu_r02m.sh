out=gpurun_out; tag=r02m
timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "exit $?" >> $out/${tag}_pytest.log
for ty in 10 12; do
  GPUSOLVE_S2_TY=$ty timeout 100 python tools/profile_levels.py --mode 0 > $out/${tag}_levels_mode0_ty$ty.txt 2>&1
  GPUSOLVE_S2_TY=$ty timeout 100 python tools/profile_levels.py --mode 1 > $out/${tag}_levels_mode1_ty$ty.txt 2>&1
  GPUSOLVE_S2_TY=$ty GPUSOLVE_FUSE_PC=2 timeout 100 python tools/profile_levels.py --mode 1 > $out/${tag}_levels_mode1_pc_ty$ty.txt 2>&1
  GPUSOLVE_S2_TY=$ty timeout 100 python tools/profile_levels.py --mode 2 --pre 3 --post 3 > $out/${tag}_levels_mode2_ty$ty.txt 2>&1
done
