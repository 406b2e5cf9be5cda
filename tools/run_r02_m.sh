#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
SVO_NO_TMA=1 python -m pytest tests/test_gpu_stages.py -m gpu -q -x -k "fast or pyramid" 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fast_kernel|pyrdown" -c 30 --csv --log-file gpurun_out/m_ingest_launches.csv python tools/ingest_probe.py > /dev/null 2>&1
grep -E "fast_kernel|pyrdown" gpurun_out/m_ingest_launches.csv | awk -F'","' '{print substr($5,1,28), $NF}' | tail -10
python bench.py --frames 160 > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; echo "bench rc $?"
python -c "
import json; d=json.load(open('gpurun_out/m_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print(d['stages']); print(d['roofline']['frac'], d['roofline']['us_per_launch_set'])"
