#!/bin/bash
# Same-box A/B of reuse-kernel build variants (gpurun_variants/lib_*.so built with AMVPT_OUT / AMVPT_CFLAGS, see build.py)
mkdir -p gpurun_out/reuse_variants
for lib in mitsuba3-amvpt_b200/libamvpt_b200.so gpurun_variants/lib_*.so; do
  n=$(basename $lib .so)
  for w in cbox8x8_reuse hfield16_reuse; do
    v=$(AMVPT_LIB=$PWD/$lib timeout 300 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1))")
    echo "$n $w $v" | tee -a gpurun_out/reuse_variants/results.txt
  done
done
