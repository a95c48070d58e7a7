#!/bin/bash
# all GPU tests + smoke, logs under gpurun_out/$1
OUT=gpurun_out/${1:-tests}; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
tail -n 25 $OUT/pytest_gpu.log; tail -n 3 $OUT/smoke.log
