#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -2
python tools/dbg_pipe.py 2>&1 | tail -4
bash tools/quick_bench.sh splitF_a
bash tools/quick_bench.sh splitF_b
SVO_LIB_PATH=tools/libsvo_old.so bash tools/quick_bench.sh before
SVO_TRACE=gpurun_out/af_timeline.csv python tools/trace_run.py 2>&1 | tail -48
