#!/bin/bash
# usage: tools/quick_bench.sh [label]  -- short bench line summary
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/qb.log 2> gpurun_out/qb.err || tail -5 gpurun_out/qb.err
python - "$1" <<PY
import json,sys
d=json.load(open("gpurun_out/qb.log"))
print(sys.argv[1], "value %.0f e2e %.0f sync %.0f ok %.2f" % (d["value"], d["e2e"]["value"], d["e2e"]["sync_per_frame_value"], d["config"]["success_rate"]), {k:round(v["us_per_frame"],1) for k,v in d["stages"].items()}, "batch %.0f" % d["batch64"].get("frames_per_s",0))
PY
