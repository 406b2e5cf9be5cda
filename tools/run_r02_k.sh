#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
(time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 2) > gpurun_out/k_bench_n2.json 2> gpurun_out/k_bench_n2.err; echo "n2 rc $?"; grep -E "real|Error" gpurun_out/k_bench_n2.err | head -5
cat gpurun_out/k_bench_n2.json | cut -c1-2600
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_gpu_pipeline.py -m gpu -q -x -k "streaming_submit or batched_sequences or (run1_gray and not 48) or empty_frame" > gpurun_out/k_sanitizer_$tool.log 2>&1; echo "$tool rc $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/k_sanitizer_$tool.log | tail -3
done
