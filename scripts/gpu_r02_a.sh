#!/bin/bash
# round 2, first GPU visit: baseline of the round-1 kernels under the repaired bench contract + captures at the benchmarked launch size
OUT=gpurun_out/r02a; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench-ref rc=$?" | tee -a $OUT/rc.txt
REPS=2 timeout 300 python scripts/one_render.py cbox8x8 auto 16 > $OUT/one_cbox8x8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pass -s 1 -c 1 -o $OUT/cbox8x8_render_pass -f \
    python scripts/one_render.py cbox8x8 auto 16 > $OUT/ncu_full_cbox.log 2>&1; echo "ncu-full-cbox rc=$?" | tee -a $OUT/rc.txt
REPS=2 timeout 300 python scripts/one_render.py hfield16 auto 16 > $OUT/one_hfield16.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 60 -c 1 -o $OUT/hfield16_wf_trace -f \
    python scripts/one_render.py hfield16 auto 16 > $OUT/ncu_full_hf.log 2>&1; echo "ncu-full-hf rc=$?" | tee -a $OUT/rc.txt
cat $OUT/rc.txt
