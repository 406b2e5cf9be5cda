#!/bin/bash
mkdir -p gpurun_out
./tools/warp_ops_lat > gpurun_out/s_warp_ops.txt 2>&1; cat gpurun_out/s_warp_ops.txt
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --frames 160 --steps 4 --no-cpu-baseline > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err; echo "bench rc $?"; tail -3 gpurun_out/s_bench.err
python -c "
import json; d=json.load(open('gpurun_out/s_bench.json')); print(d['value'], d['e2e']['value'], d['e2e']['sync_per_frame_value'], d['batch64']); print({k:round(v['us_per_frame'],1) for k,v in d['stages'].items()})"
ncu --set full --clock-control none --import-source on -k regex:"klt_kernel" -s 8 -c 1 -f -o gpurun_out/s_klt_sync python tools/trace_sync.py > gpurun_out/s_klt_sync.log 2>&1; echo "ncu rc $?"
ncu -i gpurun_out/s_klt_sync.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/s_klt_sync.source.csv 2>/dev/null
ncu -i gpurun_out/s_klt_sync.ncu-rep --page raw --csv > gpurun_out/s_klt_sync.raw.csv 2>/dev/null
