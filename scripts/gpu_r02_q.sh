#!/bin/bash
# re-test of the opt-in trace experiments on top of the 64-byte-record pool (scattered state access is cheaper now), then validation of the final tree
OUT=gpurun_out/r02q; mkdir -p $OUT
for cfg in "default" "AMVPT_SORT=1" "AMVPT_Q4=1" "AMVPT_SORT=1 AMVPT_Q4=1" "default"; do
  ( [ "$cfg" != default ] && export $cfg; REPS=2 timeout 200 python scripts/one_render.py hfield16 auto 32 2>&1 | tail -n 1 | sed "s/^/$cfg: /" | cut -c1-60,345-420 )
done | tee $OUT/ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt; tail -n 1 $OUT/smoke.log
