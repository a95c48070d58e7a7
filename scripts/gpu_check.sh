#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list + one full capture of the dominant kernel.
# Usage (from the repo root on the box): bash scripts/gpu_check.sh [tag]
TAG=${1:-run}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/bench_cbox8x8.json 2> $OUT/bench_cbox8x8.err; echo "bench8x8 rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --workload cbox4 --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_cbox4.json 2> $OUT/bench_cbox4.err; echo "bench4 rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --workload hfield16 --steps 2 --warmup 3 > $OUT/bench_hfield16.json 2> $OUT/bench_hfield16.err; echo "benchhf rc=$?" | tee -a $OUT/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/cbox4_launches.csv \
    python bench.py --workload cbox4 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1; echo "ncu-launches rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pass -s 4 -c 1 -o $OUT/cbox4_render_pass -f \
    python bench.py --workload cbox4 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pass -s 2 -c 1 -o $OUT/hfield_render_pass -f \
    python bench.py --workload hfield16 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_full_hf.log 2>&1; echo "ncu-full-hf rc=$?" | tee -a $OUT/rc.txt
cat $OUT/rc.txt
