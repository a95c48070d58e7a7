#!/bin/bash
# path-state pool as three 64-byte records per slot (default) against one array per field (-DAM_WF_AOS=0), same box
OUT=gpurun_out/r02m; mkdir -p $OUT
for v in aos soa aos soa; do
  lib=gpu_variants/lib_$v.so; [ "$v" = aos ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 200 python scripts/one_render.py hfield16 auto 32 2>&1 | tail -n 1 | sed "s/^/$v hfield16: /" | cut -c1-45,330-420
done | tee $OUT/ab.txt
for v in aos soa; do
  lib=gpu_variants/lib_$v.so; [ "$v" = aos ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 200 python scripts/one_render.py cbox8x8 wavefront 32 2>&1 | tail -n 1 | sed "s/^/$v cbox8x8 wavefront: /" | cut -c1-55,330-420
  AMVPT_LIB=$PWD/$lib REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file $OUT/launches_$v.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_$v.log 2>&1
done | tee -a $OUT/ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
