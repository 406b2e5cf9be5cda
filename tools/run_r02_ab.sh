#!/bin/bash
mkdir -p gpurun_out
for fs in 80 85 90 100; do SVO_FSHARE=$fs bash tools/quick_bench.sh fshare$fs; done
SVO_FSHARE=100 SVO_SLOTS=5 bash tools/quick_bench.sh f100s5
SVO_FSHARE=90 SVO_PARTITION=8 bash tools/quick_bench.sh f90p8
SVO_FSHARE=90 SVO_PARTITION=24 bash tools/quick_bench.sh f90p24
SVO_FSHARE=80 bash tools/quick_bench.sh fshare80again
