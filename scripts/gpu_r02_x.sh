#!/bin/bash
# 2-GPU sanity run of the final build (flat-only megakernel + NCCL gather behind the C ABI)
OUT=gpurun_out/r02x; mkdir -p $OUT
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 3 --warmup 3 > $OUT/bench_n2.json 2> $OUT/bench_n2.err; echo "bench n2 rc=$?" | tee -a $OUT/rc.txt
python - <<EOF
import json
d = [json.loads(l) for l in open('$OUT/bench_n2.json') if l.startswith('{')][-1]
print('n2', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), d['rank_ms_render_step'])
EOF
