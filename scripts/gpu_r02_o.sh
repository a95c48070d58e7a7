#!/bin/bash
# hit-sector padding (main) against the plain 64-byte records (lib_aos1), then the final wavefront captures with the faster of the two
OUT=gpurun_out/r02o; mkdir -p $OUT
t() { AMVPT_LIB=$PWD/$1 REPS=2 timeout 200 python scripts/one_render.py hfield16 auto 32 2>&1 | tail -n 1 | python -c "import sys,ast; print(ast.literal_eval(sys.stdin.read().split('} ')[0] + '}')['seconds'])"; }
MAIN=mitsuba3-amvpt_b200/libamvpt_b200.so; OLD=gpu_variants/lib_aos1.so
a1=$(t $MAIN); b1=$(t $OLD); a2=$(t $MAIN); b2=$(t $OLD)
echo "pad $a1 $a2   plain $b1 $b2" | tee $OUT/ab.txt
BEST=$(python -c "print('$MAIN' if min($a1,$a2) <= min($b1,$b2) else '$OLD')"); echo "captures with $BEST" | tee -a $OUT/ab.txt
export AMVPT_LIB=$PWD/$BEST
REPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 60 -c 1 -o $OUT/hfield16_wf_trace -f python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_trace.log 2>&1; echo "ncu trace rc=$?" | tee -a $OUT/rc.txt
REPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:wf_shade_kernel<0, 0' -s 60 -c 1 -o $OUT/hfield16_wf_shade_diffuse -f python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_shade.log 2>&1; echo "ncu shade rc=$?" | tee -a $OUT/rc.txt
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file $OUT/launches_hfield16_pass.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_list.log 2>&1
timeout 300 python bench.py --workload hfield16 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench_hfield16.json 2> $OUT/bench_hfield16.err; echo "bench hfield16 rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT
