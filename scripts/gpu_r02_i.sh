#!/bin/bash
# no-alloc entry point (alone, then after the whole suite) + A/B of the traversal variants (packed FFMA2 slab test, far-child prefetch)
OUT=gpurun_out/r02i; mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > $OUT/pytest_no_alloc_alone.log 2>&1; echo "no-alloc alone rc=$?" | tee -a $OUT/rc.txt
tail -n 4 $OUT/pytest_no_alloc_alone.log
for v in base ffma2 pf ffma2pf; do
  lib=gpu_variants/lib_$v.so; [ "$v" = base ] && lib=mitsuba3-amvpt_b200/libamvpt_b200.so
  [ -f $lib ] || { echo "$v missing"; continue; }
  for wl in cbox8x8 hfield16; do
    AMVPT_LIB=$PWD/$lib timeout 300 python bench.py --workload $wl --steps 2 --warmup 1 --no-cpu-baseline --no-extra > $OUT/bench_${v}_$wl.json 2> $OUT/bench_${v}_$wl.err; rc=$?
    python - <<E
import json
try:
    d = [json.loads(l) for l in open('$OUT/bench_${v}_$wl.json') if l.startswith('{')][-1]
    print('$v $wl rc=$rc', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'kernel ms', d['roofline'].get('launch_ms'))
except Exception as e:
    print('$v $wl rc=$rc failed', e)
E
  done
done 2>&1 | tee $OUT/ab.txt
# parity of the FFMA2 variant (node layout differs): traversal, lanes, schedulers
AMVPT_LIB=$PWD/gpu_variants/lib_ffma2pf.so timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "bvh or lane or scheduler or tree or smoke" > $OUT/pytest_ffma2pf.log 2>&1; echo "pytest ffma2pf rc=$?" | tee -a $OUT/rc.txt
tail -n 4 $OUT/pytest_ffma2pf.log
timeout 1700 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest all rc=$?" | tee -a $OUT/rc.txt
tail -n 6 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt; tail -n 2 $OUT/smoke.log
