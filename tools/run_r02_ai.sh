#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3; do SVO_FSHARE=85 bash tools/quick_bench.sh f85_$i; done
for i in 1 2 3; do SVO_FSHARE=100 bash tools/quick_bench.sh f100_$i; done
