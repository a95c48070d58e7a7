#!/bin/bash
# leaf size / SAH cost sweep with the flat leaf list (host-side parameters)
OUT=gpurun_out/r02s; mkdir -p $OUT
for ml in 2 4 7; do for ct in 0.25 0.5 1.5 3.0 6.0; do
  AMVPT_MAX_LEAF=$ml AMVPT_SAH_CT=$ct REPS=2 timeout 120 python scripts/one_render.py cbox8x8 auto 32 2>&1 | tail -n 1 | sed "s/^/max_leaf $ml sah_ct $ct: /" | cut -c1-60,345-420
done; done | tee $OUT/leaf_sweep_flat.txt
