#!/bin/bash
OUT=gpurun_out/r02f; mkdir -p $OUT
REPS=1 timeout 300 python scripts/one_render.py hfield16 auto 16 > $OUT/one.log 2>&1 &&
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum --clock-control none -c 600 --csv --log-file $OUT/launches_sorted.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu.log 2>&1; echo "ncu rc=$?" | tee -a $OUT/rc.txt
AMVPT_SORT=0 REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum --clock-control none -c 400 --csv --log-file $OUT/launches_unsorted.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu2.log 2>&1; echo "ncu2 rc=$?" | tee -a $OUT/rc.txt
