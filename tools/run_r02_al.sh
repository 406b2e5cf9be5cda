#!/bin/bash
mkdir -p gpurun_out
echo "PARTITION=0"; SVO_PARTITION=0 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit or synthetic_kitti or run1_gray" 2>&1 | tail -2
echo "SPEC=0"; SVO_SPEC=0 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit or synthetic_kitti" 2>&1 | tail -2
echo "SLOTS=3 TSHARE=0 FSHARE=59 PDL=0"; SVO_SLOTS=3 SVO_TSHARE=0 SVO_FSHARE=59 SVO_PDL=0 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit" 2>&1 | tail -2
SVO_PARTITION=0 bash tools/quick_bench.sh nopartition
