#!/bin/bash
mkdir -p gpurun_out
for p in "16,0" "16,8" "8,8" "24,8" "16,16" "8,0"; do SVO_PARTITION=$p bash tools/quick_bench.sh "part_$p"; done
