#!/bin/bash
OUT=gpurun_out/r02c; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "compressed or schedulers or film_matches or paths_bit" > $OUT/pytest_parity.log 2>&1; echo "pytest-parity rc=$?" | tee -a $OUT/rc.txt
timeout 900 python -m pytest tests/test_gpu_golden.py -x -q -k "million or hfield16 or chi2" > $OUT/pytest_golden.log 2>&1; echo "pytest-golden rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest_parity.log $OUT/pytest_golden.log
for q in 0 1; do
AMVPT_NO_Q4=$q timeout 600 python bench.py --workload hfield16 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench_hfield16_noq4_$q.json 2> $OUT/bench_hfield16_noq4_$q.err; echo "bench noq4=$q rc=$?" | tee -a $OUT/rc.txt
python -c "import json; d=json.load(open('$OUT/bench_hfield16_noq4_$q.json')); r=d['roofline']; print('NO_Q4=$q', d['value'], 'e2e', d['e2e']['value'], 'trace share', r['kernel_share_of_step'], 'launch_ms', r['launch_ms'])"
done
