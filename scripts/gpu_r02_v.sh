#!/bin/bash
# flat-only megakernel at 6 / 7 / 8 CTAs per SM
OUT=gpurun_out/r02v; mkdir -p $OUT
run() { REPS=3 timeout 120 python scripts/one_render.py cbox8x8 auto 64 2>&1 | tail -n 1 | sed "s/^/$1: /" | cut -c1-40,330-420; }
( for c in 6 7 8 6 7 8; do AMVPT_LIB=$PWD/gpu_variants/lib_flat$c.so run "flat kernel $c CTAs"; done ) | tee $OUT/ab.txt
