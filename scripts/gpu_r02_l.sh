#!/bin/bash
# final validation of the round-2 tree: leaf-size / SAH sweep (host-side, no rebuild), GPU tests, smoke, default bench
OUT=gpurun_out/r02l; mkdir -p $OUT
for ml in 1 2 4 7; do for ct in 0.5 1.5 3.0; do
  AMVPT_MAX_LEAF=$ml AMVPT_SAH_CT=$ct REPS=2 timeout 120 python scripts/one_render.py cbox8x8 auto 32 2>&1 | tail -n 1 | sed "s/^/max_leaf $ml sah_ct $ct: /" | cut -c1-60,345-420
done; done | tee $OUT/leaf_sweep.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 3 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt; tail -n 1 $OUT/smoke.log
( time timeout 900 python bench.py ) > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench default rc=$?" | tee -a $OUT/rc.txt; tail -n 4 $OUT/bench_default.err
