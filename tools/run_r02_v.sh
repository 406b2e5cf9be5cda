#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests/test_gpu_stages.py -m gpu -q -x -k "fast" 2>&1 | tail -3
python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "run1_gray or synthetic_kitti or blank or empty" 2>&1 | tail -3
for nt in 128 256 128 256; do echo "NT=$nt"; SVO_FAST_NT=$nt python tools/ingest_time.py 2>&1 | tail -1; done
for nt in 128 256; do
SVO_FAST_NT=$nt ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fast_kernel|pyrdown" -c 30 --csv --log-file $O/v_launches_$nt.csv python tools/ingest_probe.py 3 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("$O/v_launches_$nt.csv")) if len(r)>5 and r[0].isdigit()]
print("NT=$nt", [(r[4][:24], r[-1]) for r in rows[-10:]])
PY
done
