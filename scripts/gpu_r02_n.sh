#!/bin/bash
# pool record grouping v2 (A: ro rd so sd, B: tp pp res rng, C: hit sT sV) against v1 (A: ro rd tp pp, B: res rng hit, C: so sd sT sV), same box
OUT=gpurun_out/r02n; mkdir -p $OUT
for v in v2 aos1 v2 aos1; do
  lib=gpu_variants/lib_$v.so; [ "$v" = v2 ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  AMVPT_LIB=$PWD/$lib REPS=2 timeout 200 python scripts/one_render.py hfield16 auto 32 2>&1 | tail -n 1 | sed "s/^/$v hfield16: /" | cut -c1-45,330-420
done | tee $OUT/ab.txt
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file $OUT/launches_v2.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_v2.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
