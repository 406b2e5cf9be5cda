#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_zz_no_alloc.py -m gpu -q -x -k "fast or alloc" 2>&1 | tail -2
for nt in 128 160 192 256 128 160 192; do echo "NT=$nt"; SVO_FAST_NT=$nt python tools/ingest_time.py 2>&1 | tail -1; done
SVO_FAST_NT=192 python -m pytest tests/test_gpu_stages.py tests/test_gpu_pipeline.py -m gpu -q -x -k "fast or run1_gray or blank" 2>&1 | tail -2
bash tools/quick_bench.sh f70
bash tools/quick_bench.sh f70b
