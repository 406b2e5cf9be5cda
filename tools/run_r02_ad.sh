#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for p in 1 0 1 0; do echo "PDL=$p"; SVO_PDL=$p python tools/ingest_time.py 2>&1 | tail -1; done
SVO_LIB_PATH=tools/libsvo_hyp4.so python tools/ingest_time.py 2>&1 | tail -1
SVO_PDL=1 bash tools/quick_bench.sh pdl1
SVO_PDL=0 bash tools/quick_bench.sh pdl0
SVO_LIB_PATH=tools/libsvo_hyp4.so bash tools/quick_bench.sh hyp4
