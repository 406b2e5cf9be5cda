#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit or synthetic_kitti" 2>&1 | tail -2
for i in 1 2 3; do bash tools/quick_bench.sh split_$i; done
for i in 1 2; do SVO_LIB_PATH=tools/libsvo_old.so bash tools/quick_bench.sh head_$i; done
python tools/dbg_pipe.py 2>&1 | tail -2
