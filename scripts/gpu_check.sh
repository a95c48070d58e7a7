#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch lists + full captures of the dominant kernels.
# Usage (from the repo root on the box): bash scripts/gpu_check.sh [tag]
TAG=${1:-run}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/bench_cbox8x8.json 2> $OUT/bench_cbox8x8.err; echo "bench8x8 rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > $OUT/bench_reference_cbox8x8.json 2> $OUT/bench_reference.err; echo "bench-ref rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --workload cbox4 --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_cbox4.json 2> $OUT/bench_cbox4.err; echo "bench4 rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --workload hfield16 --steps 2 --warmup 3 > $OUT/bench_hfield16.json 2> $OUT/bench_hfield16.err; echo "benchhf rc=$?" | tee -a $OUT/rc.txt
# launch lists (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/cbox4_launches.csv \
    python bench.py --workload cbox4 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1; echo "ncu-launches rc=$?" | tee -a $OUT/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $OUT/hfield4_launches.csv \
    python bench.py --workload hfield4 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_launches_hf.log 2>&1; echo "ncu-launches-hf rc=$?" | tee -a $OUT/rc.txt
# full captures of the dominant kernels
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pass -s 4 -c 1 -o $OUT/cbox4_render_pass -f \
    python bench.py --workload cbox4 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 40 -c 1 -o $OUT/hfield4_wf_trace -f \
    python bench.py --workload hfield4 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_full_hf.log 2>&1; echo "ncu-full-hf rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_shade -s 160 -c 4 -o $OUT/hfield4_wf_shade -f \
    python bench.py --workload hfield4 --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_full_hf_shade.log 2>&1; echo "ncu-full-hf-shade rc=$?" | tee -a $OUT/rc.txt
cat $OUT/rc.txt
