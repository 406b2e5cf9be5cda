#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/p_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print(d['stages']); print(d['roofline']['frac'], d['roofline']['us_per_launch_set'])"
