#!/bin/bash
# flat leaf list for small scenes (default) against the binary tree (-DAM_FLAT_LEAVES=0), same box; GPU tests on the new default
OUT=gpurun_out/r02r; mkdir -p $OUT
for v in flat noflat flat noflat; do
  lib=gpu_variants/lib_$v.so; [ "$v" = flat ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=3 timeout 120 python scripts/one_render.py cbox8x8 auto 64 2>&1 | tail -n 1 | sed "s/^/$v cbox8x8: /" | cut -c1-40,330-420
done | tee $OUT/ab.txt
for v in flat noflat; do
  lib=gpu_variants/lib_$v.so; [ "$v" = flat ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 120 python scripts/one_render.py cbox8x8_reuse auto 16 2>&1 | tail -n 1 | sed "s/^/$v cbox8x8_reuse: /" | cut -c1-50,330-420
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 120 python scripts/one_render.py hfield16 regenerate 16 2>&1 | tail -n 1 | sed "s/^/$v hfield16 megakernel: /" | cut -c1-50,330-420
done | tee -a $OUT/ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 8 $OUT/pytest_gpu.log
