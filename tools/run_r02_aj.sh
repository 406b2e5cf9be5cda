#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for i in 1 2 3 4 5 6 7 8; do bash tools/quick_bench.sh d100_$i; done
python tools/dbg_pipe.py 2>&1 | tail -4
