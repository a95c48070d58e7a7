#!/bin/bash
# speculative descent (-DAM_SPECULATIVE=1) against the default build, same box; host BVH build timing
OUT=gpurun_out/r02k; mkdir -p $OUT
for v in base spec base spec; do
  lib=gpu_variants/lib_$v.so; [ "$v" = base ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=3 timeout 120 python scripts/one_render.py cbox8x8 auto 64 2>&1 | tail -n 1 | sed "s/^/$v cbox8x8: /" | cut -c1-40,330-420
done | tee $OUT/ab.txt
for v in base spec; do
  lib=gpu_variants/lib_$v.so; [ "$v" = base ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 120 python scripts/one_render.py hfield16 regenerate 16 2>&1 | tail -n 1 | sed "s/^/$v hfield16 megakernel: /" | cut -c1-50,330-420
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 120 python scripts/one_render.py cbox8x8_reuse auto 16 2>&1 | tail -n 1 | sed "s/^/$v cbox8x8_reuse: /" | cut -c1-50,330-420
done | tee -a $OUT/ab.txt
AMVPT_LIB=$PWD/gpu_variants/lib_spec.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -q > $OUT/pytest_spec.log 2>&1; echo "pytest spec rc=$?" | tee -a $OUT/rc.txt
tail -n 15 $OUT/pytest_spec.log
AMVPT_BVH_TIMING=1 REPS=1 timeout 120 python scripts/one_render.py hfield16 auto 16 2>&1 | grep "bvh" | tee $OUT/bvh_timing.txt
AMVPT_BVH_TIMING=1 AMVPT_BVH_THREADS=1 REPS=1 timeout 120 python scripts/one_render.py hfield16 auto 16 2>&1 | grep "bvh" | tee -a $OUT/bvh_timing.txt
