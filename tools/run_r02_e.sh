#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_gpu_pipeline.py -m gpu -q -x -k "lk or synthetic_kitti_64 or run1_gray" 2>&1 | tail -3
python bench.py > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/e_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']['frames_per_s']); print(d['stages'])"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pnp_hyp -s 8 -c 1 -o gpurun_out/e_hyp python tools/trace_sync.py > gpurun_out/e_ncu.log 2>&1; echo "ncu rc $?"
