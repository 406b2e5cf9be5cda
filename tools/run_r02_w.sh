#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for nt in 128 256; do echo "NT=$nt"; SVO_FAST_NT=$nt python tools/ingest_time.py 2>&1 | tail -1; done
SVO_LIB_PATH=tools/libsvo_probe.so python tools/klt_probe.py 2>&1 | tail -25
bash tools/quick_bench.sh fastv2
