#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/d_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/d_tests.log
grep -E "pose parity|trajectory:|1000 frames|passed|failed|FAILED|Error" gpurun_out/d_tests.log | cut -c1-300
python bench.py > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/d_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']['frames_per_s']); print(d['stages'])"
