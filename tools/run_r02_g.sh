#!/bin/bash
mkdir -p gpurun_out
(time python bench.py --steps 4 --warmup 3) > gpurun_out/g_bench_n1.json 2> gpurun_out/g_bench_n1.err; echo "n1 rc $?"; tail -4 gpurun_out/g_bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/g_bench_n1.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64'], d['render_s'])"
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 2) > gpurun_out/g_bench_n2.json 2> gpurun_out/g_bench_n2.err; echo "n2 rc $?"; tail -6 gpurun_out/g_bench_n2.err
cat gpurun_out/g_bench_n2.json | cut -c1-2500
