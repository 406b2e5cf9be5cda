#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_gpu_pipeline.py -m gpu -q -x -k "fast or bucket or blank or synthetic_kitti_64 or run1_gray or padded" 2>&1 | tail -4
python bench.py --frames 160 > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/h_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print(d['stages']); print(d['roofline']['frac'], d['roofline']['us_per_launch_set'])"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fast_kernel|pyrdown" -c 60 --csv --log-file gpurun_out/h_ingest_launches.csv python tools/ingest_probe.py > gpurun_out/h_ncu.log 2>&1; echo "ncu rc $?"; tail -3 gpurun_out/h_ncu.log
