#!/bin/bash
mkdir -p gpurun_out
./tools/epnp_bench > gpurun_out/f_epnp_bench.log 2>&1; cat gpurun_out/f_epnp_bench.log
python -m pytest tests/test_gpu_stages.py tests/test_gpu_pipeline.py -m gpu -q -x -k "pnp or camera or synthetic_kitti_64 or run1_gray" 2>&1 | tail -3
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/f_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']['frames_per_s']); print(d['stages'])"
