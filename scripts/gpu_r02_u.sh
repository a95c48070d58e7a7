#!/bin/bash
# flat-only megakernel instantiation (default) against the generic kernel with the run-time choice (AMVPT_FLAT_KERNEL=0), and at 6 CTAs/SM
OUT=gpurun_out/r02u; mkdir -p $OUT
run() { REPS=3 timeout 120 python scripts/one_render.py cbox8x8 auto 64 2>&1 | tail -n 1 | sed "s/^/$1: /" | cut -c1-40,330-420; }
( run "flat kernel"; AMVPT_FLAT_KERNEL=0 run "generic kernel"; AMVPT_LIB=$PWD/gpu_variants/lib_flat6.so run "flat kernel 6 CTAs"; run "flat kernel"; AMVPT_FLAT_KERNEL=0 run "generic kernel"; AMVPT_LIB=$PWD/gpu_variants/lib_flat6.so run "flat kernel 6 CTAs" ) | tee $OUT/ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 5 $OUT/pytest_gpu.log
