#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for i in 1 2 3 4 5 6; do bash tools/quick_bench.sh p8_$i; done
