#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python tools/ingest_time.py 2>&1 | tail -1
bash tools/quick_bench.sh geompyr
ncu --set full --clock-control none --import-source on -k regex:"fast_kernel" -s 2 -c 1 -f -o $O/x_fast \
    python tools/ingest_probe.py 3 > $O/x_fast.log 2>&1; echo "fast rc $?"
ncu -i $O/x_fast.ncu-rep --page source --csv > $O/x_fast.source.csv 2>/dev/null
ncu -i $O/x_fast.ncu-rep --page raw --csv > $O/x_fast.raw.csv 2>/dev/null
