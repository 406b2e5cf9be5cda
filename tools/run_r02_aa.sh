#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests/test_gpu_stages.py tests/test_zz_no_alloc.py -m gpu -q -x -k "fast or alloc" 2>&1 | tail -2
python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "run1_gray or synthetic_kitti or blank or streaming" 2>&1 | tail -2
python tools/ingest_time.py 2>&1 | tail -1
SVO_LIB_PATH=tools/libsvo_old.so python tools/ingest_time.py 2>&1 | tail -1
python tools/ingest_time.py 2>&1 | tail -1
for fs in 59 65 70 75 80; do SVO_FSHARE=$fs bash tools/quick_bench.sh fshare$fs; done
SVO_FSHARE=70 SVO_TSHARE=60 bash tools/quick_bench.sh f70t60
