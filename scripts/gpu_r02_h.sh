#!/bin/bash
# 2-GPU visit: NCCL paths behind the C ABI + single-GPU tests of this commit
OUT=gpurun_out/r02h; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 scripts/shard_test.py > $OUT/shard_test.log 2>&1; echo "shard rc=$?" | tee -a $OUT/rc.txt
tail -n 6 $OUT/shard_test.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 3 --warmup 2 > $OUT/bench_n2.json 2> $OUT/bench_n2.err; echo "bench n2 rc=$?" | tee -a $OUT/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 2 --warmup 1 --adaptive > $OUT/bench_n2_adaptive.json 2> $OUT/bench_n2_adaptive.err; echo "bench n2 adaptive rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py --steps 3 --warmup 2 --no-extra > $OUT/bench_n1.json 2> $OUT/bench_n1.err; echo "bench n1 rc=$?" | tee -a $OUT/rc.txt
for f in bench_n1 bench_n2 bench_n2_adaptive; do python -c "import json; d=json.load(open('$OUT/$f.json')); print('$f', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['rank_ms_render_step'])"; done
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "adaptive" > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt; tail -n 5 $OUT/pytest.log
