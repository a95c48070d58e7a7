#!/bin/bash
# 128-byte shading records in the wavefront shade stages (default) against the compact triangle + normal arrays (-DAM_SHADE_RECORDS=0), same box;
# then the validation of the final tree
OUT=gpurun_out/r02p; mkdir -p $OUT
t() { AMVPT_LIB=$PWD/$1 REPS=2 timeout 200 python scripts/one_render.py hfield16 auto 32 2>&1 | tail -n 1 | python -c "import sys,ast; print(ast.literal_eval(sys.stdin.read().split('} ')[0] + '}')['seconds'])"; }
MAIN=mitsuba3-amvpt_b200/libamvpt_b200.so; OLD=gpu_variants/lib_norec.so
a1=$(t $MAIN); b1=$(t $OLD); a2=$(t $MAIN); b2=$(t $OLD)
echo "records $a1 $a2   compact $b1 $b2" | tee $OUT/ab.txt
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file $OUT/launches_hfield16_pass.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_list.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt; tail -n 1 $OUT/smoke.log
( time timeout 900 python bench.py ) > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench default rc=$?" | tee -a $OUT/rc.txt; tail -n 4 $OUT/bench_default.err
