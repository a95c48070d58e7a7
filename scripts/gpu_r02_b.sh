#!/bin/bash
OUT=gpurun_out/r02b; mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_golden.py -x -q > $OUT/pytest_golden.log 2>&1; echo "pytest-golden rc=$?" | tee -a $OUT/rc.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $OUT/pytest_parity.log 2>&1; echo "pytest-parity rc=$?" | tee -a $OUT/rc.txt
timeout 300 python bench.py --workload cbox8x8 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench_cbox8x8.json 2> $OUT/bench_cbox8x8.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -30 $OUT/pytest_golden.log
