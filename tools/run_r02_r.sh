#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --frames 160 --steps 4 --no-cpu-baseline > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; echo "bench rc $?"; tail -3 gpurun_out/r_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print({k:round(v['us_per_frame'],1) for k,v in d['stages'].items()})"
