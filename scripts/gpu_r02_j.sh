#!/bin/bash
# final captures of round 2: ncu --set full of the dominant kernels at the benchmarked launch sizes, launch list of the bench command,
# regen_min sweep of the megakernel (host-side parameter, no rebuild)
OUT=gpurun_out/r02j; mkdir -p $OUT
for rm in 4 6 8 12 16; do
  AMVPT_REGEN_MIN=$rm REPS=3 timeout 120 python scripts/one_render.py cbox8x8 auto 64 2>&1 | tail -n 1 | sed "s/^/regen_min $rm: /"
done | tee $OUT/regen_min_sweep.txt
cap() { name=$1; kern=$2; skip=$3; wl=$4
  REPS=2 timeout 300 python scripts/one_render.py $wl auto 16 > $OUT/one_$name.log 2>&1 &&
  REPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -o $OUT/$name -f \
      python scripts/one_render.py $wl auto 16 > $OUT/ncu_$name.log 2>&1; echo "ncu $name rc=$?" | tee -a $OUT/rc.txt; }
cap cbox8x8_render_pass render_pass 1 cbox8x8
cap hfield16_wf_trace wf_trace 60 hfield16
cap cbox4_render_pass render_pass 1 cbox4
cap hfield16_wf_shade_diffuse 'wf_shade_kernel<0,.0,.0>' 60 hfield16
timeout 300 python bench.py --steps 2 --warmup 1 --no-extra --no-cpu-baseline > $OUT/bench_plain.json 2> $OUT/bench_plain.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_bench_cbox8x8.csv \
    python bench.py --steps 2 --warmup 1 --no-extra --no-cpu-baseline > $OUT/ncu_launches.log 2>&1; echo "ncu launch list rc=$?" | tee -a $OUT/rc.txt
REPS=1 timeout 600 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,dram__bytes_read.sum --clock-control none -c 700 --csv \
    --log-file $OUT/launches_hfield16_pass.csv python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_launches_hf.log 2>&1; echo "ncu launch list hf rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
( time timeout 900 python bench.py ) > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench default rc=$?" | tee -a $OUT/rc.txt; tail -n 4 $OUT/bench_default.err
( time timeout 600 python bench.py --impl reference ) > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench reference rc=$?" | tee -a $OUT/rc.txt; tail -n 4 $OUT/bench_reference.err
