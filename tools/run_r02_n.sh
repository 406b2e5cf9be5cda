#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_gpu_pipeline.py -m gpu -q -x -k "fast or bucket or blank or switching or synthetic_kitti_64 or padded or streaming" 2>&1 | tail -3
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fast_kernel|pyrdown" -c 20 --csv --log-file gpurun_out/n_ingest_launches.csv python tools/ingest_probe.py > /dev/null 2>&1
grep -E "fast_kernel|pyrdown" gpurun_out/n_ingest_launches.csv | awk -F'","' '{print substr($5,1,28), $NF}' | tail -5
SVO_NO_TMA=1 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fast_kernel|pyrdown" -c 20 --csv --log-file gpurun_out/n_ingest_launches_notma.csv python tools/ingest_probe.py > /dev/null 2>&1
grep -E "fast_kernel|pyrdown" gpurun_out/n_ingest_launches_notma.csv | awk -F'","' '{print substr($5,1,28), $NF}' | tail -5
