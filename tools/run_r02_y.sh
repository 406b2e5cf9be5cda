#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests/test_gpu_stages.py -m gpu -q -x -k "fast" 2>&1 | tail -2
python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x 2>&1 | tail -2
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -1
for nt in 64 128 256; do echo "NT=$nt"; SVO_FAST_NT=$nt python tools/ingest_time.py 2>&1 | tail -1; done
bash tools/quick_bench.sh base
SVO_TSHARE=100 bash tools/quick_bench.sh tshare100
SVO_TSHARE=70 bash tools/quick_bench.sh tshare70
SVO_FSHARE=50 bash tools/quick_bench.sh fshare50
SVO_FSHARE=50 SVO_TSHARE=100 bash tools/quick_bench.sh f50t100
SVO_SLOTS=4 bash tools/quick_bench.sh slots4
SVO_SLOTS=4 SVO_TSHARE=100 bash tools/quick_bench.sh slots4_t100
SVO_SLOTS=5 SVO_TSHARE=100 bash tools/quick_bench.sh slots5_t100
SVO_SLOTS=4 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit" 2>&1 | tail -2
SVO_LIB_PATH=tools/libsvo_old.so bash tools/quick_bench.sh old_head
SVO_LIB_PATH=tools/libsvo_w9.so bash tools/quick_bench.sh warps9
bash tools/quick_bench.sh base_again
