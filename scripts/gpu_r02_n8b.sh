#!/bin/bash
# 8-GPU visit with the final round-2 build (block cache, 64-byte-record pool): BASELINE configs 2 and 5
OUT=gpurun_out/r02n8b; mkdir -p $OUT
run() { name=$1; shift
  timeout $TMO python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 8 "$@" > $OUT/$name.json 2> $OUT/$name.err; rc=$?
  echo "$name rc=$rc" | tee -a $OUT/rc.txt
  python - <<EOF
import json
try:
    d = [json.loads(l) for l in open('$OUT/$name.json') if l.startswith('{')][-1]
    print('$name', 'value', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'ms/step', round(d['ms_per_step'], 1), 'paths/step', d['config']['paths_per_step'], 'ranks', d['rank_ms_render_step'][:2])
except Exception as e:
    print('$name failed', e)
EOF
}
TMO=240 PORT=29711 run c2_cbox8x8_n8 --steps 3 --warmup 3
TMO=420 PORT=29713 run c5_hfield64_adaptive1024_n8 --workload hfield64 --adaptive --spp 1024 --steps 1 --warmup 1
