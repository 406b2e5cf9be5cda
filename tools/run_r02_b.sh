#!/bin/bash
mkdir -p gpurun_out
./tools/epnp_bench > gpurun_out/b_epnp_bench.log 2>&1; cat gpurun_out/b_epnp_bench.log
python -m pytest tests -m gpu -q -s > gpurun_out/b_tests.log 2>&1; echo "tests rc $?" >> gpurun_out/b_tests.log
grep -E "pose parity|trajectory:|1000 frames|pnp vs opencv|passed|failed|FAILED" gpurun_out/b_tests.log | cut -c1-400
