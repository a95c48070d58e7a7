#!/bin/bash
# 8-GPU visit (charged 8x): BASELINE configs 2, 3 and 5 at N = 8, one run each
OUT=gpurun_out/r02n8; mkdir -p $OUT
run() { name=$1; shift
  timeout $TMO python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 8 "$@" > $OUT/$name.json 2> $OUT/$name.err; rc=$?
  echo "$name rc=$rc" | tee -a $OUT/rc.txt
  python - <<E
import json
try:
    d = [json.loads(l) for l in open('$OUT/$name.json') if l.startswith('{')][-1]
    print('$name', 'value', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'ms/step', round(d['ms_per_step'], 1), 'paths/step', d['config']['paths_per_step'], 'ranks', d['rank_ms_render_step'])
except Exception as e:
    print('$name failed', e)
E
}
TMO=240 PORT=29701 run c2_cbox8x8_n8 --steps 3 --warmup 3
TMO=240 PORT=29702 run c3_cbox8x8_adaptive_n8 --steps 3 --warmup 3 --adaptive
TMO=480 PORT=29703 run c5_hfield64_adaptive1024_n8 --workload hfield64 --adaptive --spp 1024 --steps 1 --warmup 1
