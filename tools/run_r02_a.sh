#!/bin/bash
# round-2 GPU pass A: parity suite (distributions printed), bench, sync trace, ncu of the PnP kernels
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/a_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log
python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err; echo "bench rc $?"
SVO_TRACE=gpurun_out/a_trace_sync.csv python tools/trace_sync.py > gpurun_out/a_trace_sync.log 2>&1
tail -14 gpurun_out/a_trace_sync.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pnp_ -c 9 -o gpurun_out/a_pnp python tools/trace_sync.py > gpurun_out/a_ncu.log 2>&1; echo "ncu rc $?"
