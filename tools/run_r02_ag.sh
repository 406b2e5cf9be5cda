#!/bin/bash
mkdir -p gpurun_out
for d in 3 4 5; do SVO_BENCH_DEPTH=$d bash tools/quick_bench.sh split_depth$d; done
for d in 3 4 5; do SVO_BENCH_DEPTH=$d SVO_LIB_PATH=tools/libsvo_old.so bash tools/quick_bench.sh before_depth$d; done
SVO_FSHARE=60 SVO_BENCH_DEPTH=4 bash tools/quick_bench.sh split_f60_d4
SVO_FSHARE=80 SVO_BENCH_DEPTH=4 bash tools/quick_bench.sh split_f80_d4
