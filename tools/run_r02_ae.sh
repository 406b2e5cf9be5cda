#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
python tools/ingest_time.py 2>&1 | tail -1
bash tools/quick_bench.sh final_a
bash tools/quick_bench.sh final_b
