#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming or submit or synthetic_kitti" 2>&1 | tail -2
for i in 1 2 3 4 5 6 7 8 9 10; do bash tools/quick_bench.sh ts$i; done
