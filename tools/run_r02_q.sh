#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_cpp.py -m gpu -q -x -k "lookahead or streaming or switching or wrapper or header" 2>&1 | tail -4
python bench.py --frames 160 --steps 4 --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc $?"; tail -3 gpurun_out/q_bench.err
python -c "
import json; d=json.load(open('gpurun_out/q_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print(d['roofline_lookahead8']); print(d['roofline_single_frame']['us_per_launch_set'], d['roofline']['frac'])"
