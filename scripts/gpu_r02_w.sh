#!/bin/bash
# final visit of round 2: GPU tests, smoke, captures of the flat-list megakernel (7 CTAs/SM), launch list, default bench
OUT=gpurun_out/r02w; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt; tail -n 1 $OUT/smoke.log
cap() { name=$1; kern=$2; skip=$3; wl=$4
  REPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -o $OUT/$name -f \
      python scripts/one_render.py $wl auto 16 > $OUT/ncu_$name.log 2>&1; echo "ncu $name rc=$?" | tee -a $OUT/rc.txt; }
cap cbox8x8_render_pass render_pass 1 cbox8x8
cap cbox4_render_pass render_pass 1 cbox4
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_bench_cbox8x8.csv \
    python bench.py --steps 2 --warmup 1 --no-extra --no-cpu-baseline > $OUT/ncu_launches.log 2>&1; echo "ncu launch list rc=$?" | tee -a $OUT/rc.txt
( time timeout 900 python bench.py ) > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench default rc=$?" | tee -a $OUT/rc.txt; tail -n 4 $OUT/bench_default.err
