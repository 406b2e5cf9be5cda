#!/bin/bash
mkdir -p gpurun_out
./tools/epnp_bench > gpurun_out/c_epnp_bench.log 2>&1; cat gpurun_out/c_epnp_bench.log
python -m pytest tests/test_gpu_stages.py -m gpu -q -x -k "pnp or camera" 2>&1 | tail -3
SVO_TRACE=gpurun_out/c_trace_sync.csv python tools/trace_sync.py 2>&1 | tail -8
python bench.py > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/c_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value']); print(d['stages'])"
