#!/bin/bash
OUT=gpurun_out/r02d; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "schedulers or film_matches or reconstruction" > $OUT/pytest_parity.log 2>&1; echo "pytest-parity rc=$?" | tee -a $OUT/rc.txt
for cfg in "0 1" "0 0" "1 1" "1 0"; do set -- $cfg
AMVPT_NO_Q4=$1 AMVPT_L2_PERSIST=$2 timeout 600 python bench.py --workload hfield16 --steps 2 --warmup 1 --no-cpu-baseline > $OUT/bench_hfield16_$1_$2.json 2> $OUT/bench_hfield16_$1_$2.err; echo "bench noq4=$1 persist=$2 rc=$?" | tee -a $OUT/rc.txt
python -c "import json; d=json.load(open('$OUT/bench_hfield16_$1_$2.json')); r=d['roofline']; print('NO_Q4=$1 PERSIST=$2', d['value'], 'e2e', d['e2e']['value'], 'trace share', r['kernel_share_of_step'], 'launch_ms', r['launch_ms'])"
done
